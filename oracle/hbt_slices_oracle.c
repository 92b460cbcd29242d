/*
 * oracle/hbt_slices_oracle.c -- plain-C restatement of the interval/slice intersection of
 * hbt's tagstack.  TEST INFRASTRUCTURE ONLY: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline leg may call it.
 *
 * Pinned against the real reference (oracle/_ref/libhbtref_slices.so, built from
 * hbt/src/tagstack/IntervalSlicer.cpp) and against the hand-built vectors of
 * hbt/src/tagstack/tests/IntervalSlicerTest.cpp by tests/test_slices_oracle.py.
 *
 * All time arithmetic is wrapping u64, as TimeStamp = uint64_t in the reference
 * (hbt/src/common/System.h:105).
 */
#include <errno.h>
#include <stdint.h>
#include <string.h>

#include "hbtslices_api.h"

#define TT_ANALYSIS 1 /* Slice::TransitionType::Analysis, Slicer.h:36 */

static inline uint64_t s_end(const hbtslices_slice* s) { return s->tstamp + s->duration; }

/*
 * IntervalSlicer::makeIntervalIntersectingSlices, IntervalSlicer.cpp:52-188.
 * The ring buffer transaction (startTx / seekInTx(offset) / readInTx / cancelTx) is a cursor
 * into slices[offset..ns): nextSlice returns NULL at the end (-ENODATA, :36-44).
 * A seek beyond the end is clamped to the end (hbt/src/ringbuffer/Consumer.h:273-286).
 */
int hbtoracle_islice_make(const hbtslices_slice* slices, uint64_t ns, uint64_t offset,
                          uint64_t tstamp, uint64_t duration, uint32_t flags,
                          hbtslices_slice* out, uint64_t cap, uint64_t* n_out,
                          uint64_t* consumed) {
  const int trim = (flags & HBTSLICES_TRIM) != 0;
  const int abort_missing = (flags & HBTSLICES_ABORT_ON_MISSING) != 0;
  const uint64_t tstamp_end = tstamp + duration; /* :62 */
  *n_out = 0;
  *consumed = 0;
  if (duration == 0) return -EINVAL; /* HBT_ARG_CHECK_GT(duration, 0), :64 */
  if (offset > ns) offset = ns;      /* seekInTx clamps to [tail, head], Consumer.h:281 */

  uint64_t cur = offset; /* index of s; cur == ns <=> s == nullptr */
  uint64_t num_consumed = 0;
  /* :76-82 move to the first slice that ends at or after the interval's start */
  while (cur < ns && s_end(&slices[cur]) < tstamp) {
    if (slices[cur].duration == 0) return -EINVAL; /* :77-78 */
    ++cur;
    ++num_consumed;
  }
  if (cur >= ns) return 1;                                          /* :85-89 */
  if (abort_missing && slices[cur].tstamp > tstamp) return 1;       /* :91-95 */

  uint64_t slices_duration = 0;
  uint64_t last_end = slices[cur].tstamp; /* :99 */
  uint64_t n = 0;
  /* :103-122 */
  while (cur < ns && slices[cur].tstamp < tstamp_end) {
    const hbtslices_slice* s = &slices[cur];
    if (abort_missing && last_end != s->tstamp) return 1; /* hole, :106-110 */
    if (!trim && s_end(s) > tstamp_end) break;            /* :113-115 */
    slices_duration += s->duration;
    if (n >= cap) return -ENOSPC;
    out[n++] = *s;
    last_end = s_end(s);
    ++cur;
    ++num_consumed;
  }
  if (n == 0) return 1; /* :128-133 */

  hbtslices_slice* last = &out[n - 1];
  const uint64_t last_slice_end = s_end(last); /* :137, before any trimming */
  if (abort_missing && last_slice_end < tstamp_end) return 1; /* :139-143 */

  if (trim) { /* :146-159 */
    hbtslices_slice* f = &out[0];
    if (f->tstamp < tstamp) {
      const uint64_t overlap = s_end(f) - tstamp;
      const uint64_t diff = f->duration - overlap;
      f->duration = overlap;
      f->tstamp = tstamp;
      f->swin_type = TT_ANALYSIS;
      slices_duration -= diff;
    }
  }
  if (last_slice_end > tstamp_end) --num_consumed; /* :161-164 */
  if (trim && last_slice_end > tstamp_end) {       /* :167-173 */
    const uint64_t delta = last_slice_end - tstamp_end;
    last->duration -= delta;
    last->swout_type = TT_ANALYSIS;
    slices_duration -= delta;
  }
  if (slices_duration > duration) return -EINVAL;                          /* :175-177 */
  if (trim && abort_missing && slices_duration != duration) return -EINVAL; /* :178-183 */
  *n_out = n;
  *consumed = num_consumed;
  return 0;
}

/* The row loop of TraceMonitor::releaseCountData_, hbt/src/mon/TraceMonitor.cpp:52-96. */
int hbtoracle_islice_release(const hbtslices_slice* slices, uint64_t ns, const uint64_t* rows,
                             uint64_t stride, uint64_t n_rows, uint64_t stop_ts, uint32_t flags,
                             uint64_t* mark, uint32_t* counts, hbtslices_slice* out,
                             uint64_t cap, uint64_t* n_out, uint64_t* n_intervals,
                             uint64_t* err_interval) {
  uint64_t k, total = 0;
  *n_out = 0;
  *n_intervals = 0;
  *err_interval = ~0ull;
  for (k = 0; k < n_rows; ++k) {
    const uint64_t tstamp = rows[k * stride], duration = rows[k * stride + 1];
    uint64_t n = 0, consumed = 0;
    int rc;
    if (tstamp + duration >= stop_ts) break; /* :63-70 */
    rc = hbtoracle_islice_make(slices, ns, *mark, tstamp, duration, flags, out + total,
                               cap - total, &n, &consumed);
    if (rc == -ENOSPC) return rc;
    if (rc < 0) {
      *err_interval = k;
      *n_intervals = k;
      *n_out = total;
      return rc;
    }
    counts[k] = (uint32_t)n; /* 0 for the nullptr result, :82-85 */
    total += n;
    *mark += consumed; /* :89 */
  }
  *n_intervals = k;
  *n_out = total;
  return 0;
}
