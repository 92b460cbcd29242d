// oracle/ref_bin_harness.cpp -- drives the UNMODIFIED reference binner.  TEST INFRASTRUCTURE ONLY
// (see oracle/hbtslices_api.h).
//
// Real reference code exercised (hbt/src/mon/MonData.h, compiled with HBT_ENABLE_TRACING):
//   * CuMonData::stackIdFreqs        :236-256     * MonData::stackIdFreqs / numCountIntervals :334-398
//   * binmap::BinMapper::markBinKey  :497-523     * binmap::TagStackIdBinner (constructor)    :528-607
//   * binmap::IntervalBinMatrix      :425-468     * CountData::write / numIntervals            :69-92
// Stand-ins for what the tree or the image lacks (oracle/stubs/): hbt/src/phase (absent module;
// MonData only stores a pointer to it), pfs (third-party /proc reader), boost::intrusive.
// Glue: MonData.h:587 formats a Slice with fmt, which the vendored fmt (v11) only does through an
// explicit formatter; the harness declares fmt::ostream_formatter for it (the reference's
// operator<<, Slicer.h:111-116, does the printing).
#include <vector>  // hbt/src/common/Defs.h uses std::vector without including it

#define HBT_ENABLE_TRACING 1
#include "hbt/src/common/Defs.h"
#undef HBT_LOG_WARNING
#undef HBT_LOG_ERROR
#undef HBT_LOG_INFO
struct HbtrefNullLog {
  template <class T>
  HbtrefNullLog& operator<<(const T&) { return *this; }
  HbtrefNullLog& operator<<(std::ostream& (*)(std::ostream&)) { return *this; }
};
#define HBT_LOG_WARNING() HbtrefNullLog()
#define HBT_LOG_ERROR() HbtrefNullLog()
#define HBT_LOG_INFO() HbtrefNullLog()
#undef HBT_LOG_WARNING_IF
#define HBT_LOG_WARNING_IF(c) HbtrefNullLog()

#include "hbt/src/tagstack/Slicer.h"
#include <fmt/ostream.h>
template <>
struct fmt::formatter<facebook::hbt::tagstack::slice::Slice> : fmt::ostream_formatter {};

#include <errno.h>
#include <string.h>

#include <map>

#include "hbt/src/mon/MonData.h"
#include "hbtslices_api.h"

using namespace facebook::hbt;
using tagstack::slice::Slice;

extern "C" int hbtref_bin(const uint32_t* counts, const uint64_t* int_begin, const hbtslices_slice* slices,
                          uint32_t n_units, uint64_t min_distinct_intervals, uint64_t min_duration,
                          hbtslices_freq* freqs, uint64_t freq_cap, uint64_t* n_freqs, uint64_t* durations,
                          uint64_t dur_cap, uint64_t* num_rows, uint64_t* num_bins) {
  try {
    mon::MonData::PerCuMonData per_cu;
    uint64_t s = 0;
    for (uint32_t u = 0; u < n_units; ++u) {
      auto cd = std::make_shared<mon::CountData>(std::vector<std::string>{"event"});
      auto pis = std::make_shared<tagstack::PerIntervalSlices>();
      for (uint64_t k = int_begin[u]; k < int_begin[u + 1]; ++k) {
        const uint64_t row[3] = {k, 1, 0};  // the binner only counts the rows
        cd->write(row);
        tagstack::slice::Slices v;
        for (uint32_t i = 0; i < counts[k]; ++i, ++s)
          v.push_back(Slice::makeSlice(slices[s].tstamp, slices[s].duration, slices[s].stack_id,
                                       static_cast<Slice::TransitionType>(slices[s].swin_type),
                                       static_cast<Slice::TransitionType>(slices[s].swout_type)));
        pis->push_back(std::move(v));
      }
      per_cu[static_cast<tagstack::CompUnitId>(u)] = std::make_shared<mon::CuMonData>(cd, pis);
    }
    mon::MonData md("short", "long", TimeStampNsConverter{}, "key", std::move(per_cu), nullptr);
    mon::binmap::TagStackIdBinner binner(md, std::nullopt, min_distinct_intervals, min_duration);
    auto fr = md.stackIdFreqs(std::nullopt);
    std::map<uint64_t, mon::SliceFreq> sorted(fr.begin(), fr.end());
    *n_freqs = sorted.size();
    *num_rows = binner.durations->kNumIntervals;
    *num_bins = binner.durations->kNumBids;
    if (sorted.size() > freq_cap || binner.durations->getSize() > dur_cap) return -ENOSPC;
    uint64_t i = 0;
    for (const auto& [sid, f] : sorted) {
      hbtslices_freq& o = freqs[i++];
      o.stack_id = sid;
      o.duration = f.duration;
      o.num_obs = f.num_obs;
      o.num_intervals = f.num_intervals;
      o.bid = binner.sidToBid(sid);
      o.pad = 0;
    }
    memcpy(durations, binner.durations->getDataPtr(), binner.durations->getSize() * sizeof(uint64_t));
    return 0;
  } catch (...) {
    return -EINVAL;
  }
}
