// oracle/stubs/hbt/src/phase/PerCpuMetaDataConsumer.h -- TEST INFRASTRUCTURE ONLY.
//
// hbt/src/mon/MonData.h:11 includes hbt/src/phase/PerCpuMetaDataConsumer.h, a module that is
// absent from the reference tree.  MonData only stores a shared_ptr<phase::MetaDataContainer>,
// clones it and prints it (MonData.h:304-323,409-413); the binner (MonData.h:528-607) never
// looks at it.  This stand-in supplies exactly that much so that the REAL MonData / CuMonData /
// TagStackIdBinner compile; the harness always passes a null metadata pointer.
#pragma once
#include <memory>
#include <ostream>

namespace facebook::hbt::phase {
struct MetaDataContainer {
  std::unique_ptr<MetaDataContainer> clone() const { return std::make_unique<MetaDataContainer>(); }
};
inline std::ostream& operator<<(std::ostream& os, const MetaDataContainer&) { return os << "<MetaDataContainer>"; }
}  // namespace facebook::hbt::phase
