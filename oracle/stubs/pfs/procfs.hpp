// oracle/stubs/pfs/procfs.hpp -- TEST INFRASTRUCTURE ONLY.
//
// hbt/src/common/ProcFsHelpers.h:8 includes the third-party `pfs` /proc reader, a
// library with its own sources that the thread-switch decode does not depend on: it is used
// only by readProcFsCmdLine (ProcFsHelpers.h:16-24) to fill ThreadInfo::cmdline.  The harness
// puts this directory on the include path instead of building that library; every lookup
// throws std::system_error, which readProcFsCmdLine turns into std::nullopt.  The Event
// stream and the virtual-id assignment (what the parity tests compare) are unaffected.
#pragma once
#include <string>
#include <system_error>
#include <vector>
#include <fmt/format.h>
#include <fmt/ranges.h>

namespace pfs {
struct mem_region {};  // named in a declaration of hbt/src/mon/MonData.h:666 only
struct task {
  std::vector<std::string> get_cmdline() const {
    throw std::system_error(std::make_error_code(std::errc::no_such_file_or_directory));
  }
};
struct procfs {
  task get_task(int) const { return task{}; }
};
}  // namespace pfs
