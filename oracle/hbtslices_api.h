/*
 * oracle/hbtslices_api.h -- CPU checkers for the tagstack slice path (SURVEY.md section 8(f)
 * row f3): events -> slices (Slicer) and slices x count intervals -> per-interval slices
 * (IntervalSlicer), plus the per-interval duration matrix of row f1 (TagStackIdBinner).
 * TEST INFRASTRUCTURE ONLY, like hbtref_api.h.
 *
 *   hbtref_*     oracle/_ref/libhbtref_slices.so: the REAL reference classes
 *                  tagstack::IntervalSlicer::makeIntervalIntersectingSlices
 *                      hbt/src/tagstack/IntervalSlicer.cpp:52-188 (compiled from where it lies)
 *                  tagstack::slice::Slicer<>::processNext    hbt/src/tagstack/Slicer.h:303-869
 *                driven the way TraceMonitor::releaseCountData_ drives them
 *                (hbt/src/mon/TraceMonitor.cpp:25-103).  boost::intrusive is absent from the
 *                image: oracle/stubs/boost/intrusive/list.hpp stands in for it.
 *   hbtoracle_*  oracle/libhbtoracle.so: plain-C restatement (hbt_slices_oracle.c); its
 *                slicer_run covers level-0 thread events only (-ENOTSUP for phase events
 *                and other levels, which need the general stack dictionary)
 */
#ifndef HBTSLICES_API_H
#define HBTSLICES_API_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* tagstack::slice::Slice, hbt/src/tagstack/Slicer.h:30-49 (same size and field order) */
typedef struct {
  uint64_t tstamp;
  uint64_t duration;
  uint64_t stack_id;
  uint8_t swin_type;  /* Slice::TransitionType: 0 NA, 1 Analysis, 2 ThreadPreempted, 3 ThreadYield, 4 PhaseChange */
  uint8_t swout_type;
  uint8_t pad[6];
} hbtslices_slice;

/* tagstack::Event, hbt/src/tagstack/Event.h:28-56 */
typedef struct {
  uint64_t tstamp;
  uint8_t type;
  uint8_t level;
  uint16_t comp_unit_id;
  uint32_t tag;
} hbtslices_event;

/* SlicerStats, hbt/src/tagstack/Slicer.h:276-285 */
typedef struct {
  uint64_t num_processed, num_error_start, num_error_end, num_error_out_sync, num_dropped,
      num_out_of_order, num_slices_completed, num_logic_errors;
} hbtslices_slicer_stats;

/* one entry of getTagStackStatsDump(), Slicer.h:160-193,478-485 (ordered by stack_id) */
typedef struct {
  uint64_t stack_id, parent_stack_id;
  uint32_t tags[6];
  uint8_t num_set_tags, known_num_levels, pad[6];
  uint64_t cum_active_time, num_swins, num_swouts;
} hbtslices_stack;

/* One stack id of MonData::stackIdFreqs (SliceFreq, hbt/src/mon/MonData.h:31-51) with the bin
 * TagStackIdBinner gave it (MonData.h:528-607; bin 0 = ignored_bid). */
typedef struct {
  uint64_t stack_id, duration, num_obs, num_intervals;
  uint32_t bid, pad;
} hbtslices_freq;

#define HBTSLICES_TRIM 1u
#define HBTSLICES_ABORT_ON_MISSING 2u

#define HBTSLICES_DECL(prefix)                                                                   \
  /* One makeIntervalIntersectingSlices call on a slicer holding slices[0..ns).  Returns 1    */ \
  /* when the reference returns {nullptr, 0}, 0 otherwise (*n_out slices in out, *consumed),  */ \
  /* -EINVAL when the reference throws.                                                        */ \
  int prefix##_islice_make(const hbtslices_slice* slices, uint64_t ns, uint64_t offset,          \
      uint64_t tstamp, uint64_t duration, uint32_t flags, hbtslices_slice* out, uint64_t cap,    \
      uint64_t* n_out, uint64_t* consumed);                                                      \
  /* The loop of TraceMonitor::releaseCountData_ (TraceMonitor.cpp:52-96) for one comp unit:  */ \
  /* rows[k*stride+0..1] = {tstamp, duration}; stops in front of the first row with           */ \
  /* tstamp + duration >= stop_ts; counts[k] = slices of interval k (0 = the empty entry      */ \
  /* pushed for a nullptr result); *mark advances by the consumed slices.  Returns 0, or      */ \
  /* -EINVAL with *err_interval = the row on which the reference throws, -ENOSPC when out is  */ \
  /* too small.                                                                                */ \
  int prefix##_islice_release(const hbtslices_slice* slices, uint64_t ns, const uint64_t* rows,  \
      uint64_t stride, uint64_t n_rows, uint64_t stop_ts, uint32_t flags, uint64_t* mark,        \
      uint32_t* counts, hbtslices_slice* out, uint64_t cap, uint64_t* n_out,                     \
      uint64_t* n_intervals, uint64_t* err_interval);                                            \
  /* Slicer::processNext over events already in stream order (Event::happenedBefore, i.e.     */ \
  /* what the Combinator tree delivers) until they run out.  Slices of comp unit u are        */ \
  /* appended to out + u * cap_per_unit (n_slices[u] of them).  stacks: the stats dump in     */ \
  /* stack_id order.  Returns 0, -EINVAL when the reference throws, -ENOSPC on overflow.      */ \
  int prefix##_slicer_run(const hbtslices_event* events, uint64_t n_events, uint32_t n_units,    \
      hbtslices_slice* out, uint64_t cap_per_unit, uint64_t* n_slices,                           \
      hbtslices_slicer_stats* stats, hbtslices_stack* stacks, uint64_t stack_cap,                \
      uint64_t* n_stacks, uint64_t* last_ts);

/* TagStackIdBinner over a MonData whose comp unit u holds n_int = int_begin[u+1] - int_begin[u]
 * count intervals with counts[int_begin[u] + k] slices each (slices: all of them, unit after
 * unit, interval after interval).  freqs come out in ascending stack_id order; durations is the
 * IntervalBinMatrix, num_rows x num_bins, row-major.  The REAL binner numbers the bins in the
 * iteration order of a std::unordered_map (implementation-defined); the restatement numbers
 * them in ascending stack_id order -- compare through freqs[].bid.  Returns 0, -ENOSPC. */
#define HBTSLICES_BIN_DECL(prefix)                                                               \
  int prefix##_bin(const uint32_t* counts, const uint64_t* int_begin, const hbtslices_slice* slices, \
      uint32_t n_units, uint64_t min_distinct_intervals, uint64_t min_duration,                  \
      hbtslices_freq* freqs, uint64_t freq_cap, uint64_t* n_freqs, uint64_t* durations,          \
      uint64_t dur_cap, uint64_t* num_rows, uint64_t* num_bins);

HBTSLICES_DECL(hbtref)
HBTSLICES_DECL(hbtoracle)
HBTSLICES_BIN_DECL(hbtref)
HBTSLICES_BIN_DECL(hbtoracle)

#ifdef __cplusplus
}
#endif
#endif
