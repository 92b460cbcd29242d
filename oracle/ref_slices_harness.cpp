// oracle/ref_slices_harness.cpp -- drives the UNMODIFIED reference tagstack classes.
// TEST INFRASTRUCTURE ONLY (see oracle/hbtslices_api.h).
//
// Built by oracle/Makefile into oracle/_ref/libhbtref_slices.so together with
// hbt/src/tagstack/IntervalSlicer.cpp compiled from where it lies under /root/reference.
// Real reference code exercised here:
//   * tagstack::IntervalSlicer::{write, makeIntervalIntersectingSlices}
//       hbt/src/tagstack/IntervalSlicer.h:115-135, IntervalSlicer.cpp:52-188
//     on its own ringbuffer (hbt/src/ringbuffer/{RingBuffer,Producer,Consumer}.h)
//   * tagstack::slice::Slicer<TEventsStream, TSlices>::processNext and everything it calls
//       hbt/src/tagstack/Slicer.h:303-869, TagStack.h, Event.h
// Harness glue (the reference has no standalone form of it):
//   * the row loop of TraceMonitor::releaseCountData_ (hbt/src/mon/TraceMonitor.cpp:52-96;
//     that file needs hbt/src/phase, which is absent from the tree), restated around the real
//     makeIntervalIntersectingSlices;
//   * an events stream over an array (prepareNext/complete as Stream.h:26-36 specifies them)
//     and a slice sink with the producer's write().
#include <vector>  // hbt/src/common/Defs.h uses std::vector without including it

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

#include <errno.h>
#include <string.h>

#include <map>
#include <memory>

#include "hbt/src/tagstack/IntervalSlicer.h"
#include "hbt/src/tagstack/Slicer.h"
#include "hbtslices_api.h"

using namespace facebook::hbt;
using tagstack::slice::Slice;

static_assert(sizeof(Slice) == sizeof(hbtslices_slice));
static_assert(offsetof(Slice, stack_id) == offsetof(hbtslices_slice, stack_id));
static_assert(offsetof(Slice, swin_type) == offsetof(hbtslices_slice, swin_type));
static_assert(offsetof(Slice, swout_type) == offsetof(hbtslices_slice, swout_type));
static_assert(sizeof(tagstack::Event) == sizeof(hbtslices_event));
static_assert(offsetof(tagstack::Event, comp_unit_id) == offsetof(hbtslices_event, comp_unit_id));
static_assert(offsetof(tagstack::Event, tag) == offsetof(hbtslices_event, tag));

namespace {

Slice toRef(const hbtslices_slice& s) {
  return Slice::makeSlice(s.tstamp, s.duration, s.stack_id,
                          static_cast<Slice::TransitionType>(s.swin_type),
                          static_cast<Slice::TransitionType>(s.swout_type));
}

hbtslices_slice fromRef(const Slice& s) {
  hbtslices_slice o;
  memset(&o, 0, sizeof(o));
  o.tstamp = s.tstamp;
  o.duration = s.duration;
  o.stack_id = s.stack_id;
  o.swin_type = static_cast<uint8_t>(s.swin_type);
  o.swout_type = static_cast<uint8_t>(s.swout_type);
  return o;
}

std::unique_ptr<tagstack::IntervalSlicer> fill(const hbtslices_slice* slices, uint64_t ns) {
  auto is = std::make_unique<tagstack::IntervalSlicer>(ns + 2, 0);
  for (uint64_t i = 0; i < ns; ++i) {
    ssize_t ret = is->write(toRef(slices[i]));
    if (ret < 0) return nullptr;
  }
  return is;
}

struct ArrayStream {
  const tagstack::Event* ev;
  uint64_t n, pos = 0;
  const tagstack::Event* prepareNext(TimeStamp stop_ts) {
    if (pos >= n) return nullptr;
    return ev[pos].tstamp <= stop_ts ? &ev[pos] : nullptr;
  }
  void complete() { ++pos; }
};

struct Sink {
  std::vector<Slice> v;
  ssize_t write(const Slice& s) {
    v.push_back(s);
    return sizeof(Slice);
  }
};

}  // namespace

extern "C" int hbtref_islice_make(const hbtslices_slice* slices, uint64_t ns, uint64_t offset,
                                  uint64_t tstamp, uint64_t duration, uint32_t flags,
                                  hbtslices_slice* out, uint64_t cap, uint64_t* n_out,
                                  uint64_t* consumed) {
  *n_out = 0;
  *consumed = 0;
  try {
    auto is = fill(slices, ns);
    if (!is) return -ENOMEM;
    auto [res, n] = is->makeIntervalIntersectingSlices(offset, tstamp, duration,
                                                       flags & HBTSLICES_TRIM,
                                                       flags & HBTSLICES_ABORT_ON_MISSING);
    if (res == nullptr) return 1;
    if (res->size() > cap) return -ENOSPC;
    for (size_t i = 0; i < res->size(); ++i) out[i] = fromRef((*res)[i]);
    *n_out = res->size();
    *consumed = n;
    return 0;
  } catch (...) {
    return -EINVAL;
  }
}

extern "C" int hbtref_islice_release(const hbtslices_slice* slices, uint64_t ns,
                                     const uint64_t* rows, uint64_t stride, uint64_t n_rows,
                                     uint64_t stop_ts, uint32_t flags, uint64_t* mark,
                                     uint32_t* counts, hbtslices_slice* out, uint64_t cap,
                                     uint64_t* n_out, uint64_t* n_intervals,
                                     uint64_t* err_interval) {
  *n_out = 0;
  *n_intervals = 0;
  *err_interval = ~0ull;
  std::unique_ptr<tagstack::IntervalSlicer> is;
  try {
    is = fill(slices, ns);
  } catch (...) {
    return -ENOMEM;
  }
  if (!is) return -ENOMEM;
  uint64_t k = 0;
  try {
    for (; k < n_rows; ++k) {
      const uint64_t* p = rows + k * stride;
      TimeStamp tstamp = p[0], duration = p[1];
      if (tstamp + duration >= stop_ts) break;  // TraceMonitor.cpp:63-70
      auto [res, n] = is->makeIntervalIntersectingSlices(*mark, tstamp, duration,
                                                         flags & HBTSLICES_TRIM,
                                                         flags & HBTSLICES_ABORT_ON_MISSING);
      if (res == nullptr) {
        counts[k] = 0;  // pislices->push_back({}), :82-85
      } else {
        if (*n_out + res->size() > cap) return -ENOSPC;
        for (const auto& s : *res) out[(*n_out)++] = fromRef(s);
        counts[k] = static_cast<uint32_t>(res->size());
      }
      *mark += n;  // :89
    }
  } catch (...) {
    *err_interval = k;
    *n_intervals = k;
    return -EINVAL;
  }
  *n_intervals = k;
  return 0;
}

extern "C" int hbtref_slicer_run(const hbtslices_event* events, uint64_t n_events,
                                 uint32_t n_units, hbtslices_slice* out, uint64_t cap_per_unit,
                                 uint64_t* n_slices, hbtslices_slicer_stats* stats,
                                 hbtslices_stack* stacks, uint64_t stack_cap, uint64_t* n_stacks,
                                 uint64_t* last_ts) {
  using TSlicer = tagstack::slice::Slicer<ArrayStream, Sink>;
  auto stream = std::make_shared<ArrayStream>();
  stream->ev = reinterpret_cast<const tagstack::Event*>(events);
  stream->n = n_events;
  auto grouped = std::make_shared<TSlicer::TGroupedSlices>();
  for (uint32_t u = 0; u < n_units; ++u) grouped->push_back(std::make_shared<Sink>());
  int rc = 0;
  try {
    TSlicer slicer{stream, grouped};
    while (stream->pos < n_events) {
      unsigned n = slicer.processNext(1u << 20, kInvalidTimeStamp);
      if (n == 0 && stream->pos < n_events) break;
    }
    for (uint32_t u = 0; u < n_units; ++u) {
      const auto& v = (*grouped)[u]->v;
      n_slices[u] = v.size();
      if (v.size() > cap_per_unit) {
        rc = -ENOSPC;
        continue;
      }
      for (size_t i = 0; i < v.size(); ++i) out[u * cap_per_unit + i] = fromRef(v[i]);
    }
    auto st = slicer.getStats();
    stats->num_processed = st.num_processed;
    stats->num_error_start = st.num_error_start;
    stats->num_error_end = st.num_error_end;
    stats->num_error_out_sync = st.num_error_out_sync;
    stats->num_dropped = st.num_dropped;
    stats->num_out_of_order = st.num_out_of_order;
    stats->num_slices_completed = st.num_slices_completed;
    stats->num_logic_errors = st.num_logic_errors;
    *last_ts = slicer.getLastTimeStamp();
    auto dump = slicer.getTagStackStatsDump();
    *n_stacks = dump->size();
    uint64_t i = 0;
    for (const auto& [sid, ss] : *dump) {
      if (i >= stack_cap) {
        rc = -ENOSPC;
        break;
      }
      hbtslices_stack& o = stacks[i++];
      memset(&o, 0, sizeof(o));
      o.stack_id = ss.stack_id;
      o.parent_stack_id = ss.parent_stack_id;
      // tags at or above num_set_tags are indeterminate in the reference (TagStack.h:103-112)
      for (int t = 0; t < 6; ++t) o.tags[t] = t < ss.stack.num_set_tags ? ss.stack.tags[t] : 0xffffffffu;
      o.num_set_tags = ss.stack.num_set_tags;
      o.known_num_levels = ss.stack.known_num_levels;
      o.cum_active_time = ss.cum_active_time;
      o.num_swins = ss.num_swins;
      o.num_swouts = ss.num_swouts;
    }
  } catch (...) {
    return -EINVAL;
  }
  return rc;
}
