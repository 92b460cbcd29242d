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
#include <stdlib.h>
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

/* ---- row f1: MonData::stackIdFreqs + binmap::TagStackIdBinner (hbt/src/mon/MonData.h) --------
 *
 * Freqs (CuMonData::stackIdFreqs :236-256, summed over the units :382-398): per stack id the
 * total duration, the number of slices, and the number of intervals it appears in (once per
 * interval and unit).  Binner (:539-607): bin 0 is the ignored bin; a stack id with
 * num_intervals >= min_distinct_intervals and duration >= min_duration gets a bin of its own
 * -- here in ascending stack_id order (the reference walks a std::unordered_map).  The matrix
 * has one row per count interval of all units; every unit restarts at row 0 (`unsigned i = 0`
 * sits inside the loop over the units, :563), skips intervals without slices, adds each
 * slice's duration to [i][bid] unless it is 0 (:581-590), and moves to the next row only if
 * the interval's durations added up to something (:592-597).
 */
static int freq_cmp(const void* a, const void* b) {
  const uint64_t x = ((const hbtslices_freq*)a)->stack_id, y = ((const hbtslices_freq*)b)->stack_id;
  return x < y ? -1 : x > y;
}

static hbtslices_freq* freq_find(hbtslices_freq* f, uint64_t n, uint64_t sid) {
  uint64_t lo = 0, hi = n;
  while (lo < hi) {
    const uint64_t mid = (lo + hi) / 2;
    if (f[mid].stack_id < sid) lo = mid + 1;
    else hi = mid;
  }
  return (lo < n && f[lo].stack_id == sid) ? &f[lo] : 0;
}

int hbtoracle_bin(const uint32_t* counts, const uint64_t* int_begin, const hbtslices_slice* slices,
                  uint32_t n_units, uint64_t min_distinct_intervals, uint64_t min_duration,
                  hbtslices_freq* freqs, uint64_t freq_cap, uint64_t* n_freqs, uint64_t* durations,
                  uint64_t dur_cap, uint64_t* num_rows, uint64_t* num_bins) {
  const uint64_t n_int = int_begin[n_units] - int_begin[0];
  uint64_t n_slices = 0, n = 0, k, s, i;
  uint32_t u, next_bid = 1;
  for (k = int_begin[0]; k < int_begin[n_units]; ++k) n_slices += counts[k];
  /* distinct stack ids, ascending: collect, sort, unique */
  for (s = 0; s < n_slices; ++s) {
    if (n == freq_cap) { /* compact what we have before giving up */
      uint64_t m = 0;
      qsort(freqs, n, sizeof(*freqs), freq_cmp);
      for (i = 0; i < n; ++i)
        if (m == 0 || freqs[m - 1].stack_id != freqs[i].stack_id) freqs[m++] = freqs[i];
      n = m;
      if (n == freq_cap) return -ENOSPC;
    }
    memset(&freqs[n], 0, sizeof(*freqs));
    freqs[n++].stack_id = slices[s].stack_id;
  }
  qsort(freqs, n, sizeof(*freqs), freq_cmp);
  {
    uint64_t m = 0;
    for (i = 0; i < n; ++i)
      if (m == 0 || freqs[m - 1].stack_id != freqs[i].stack_id) freqs[m++] = freqs[i];
    n = m;
  }
  /* stackIdFreqs */
  s = 0;
  for (k = int_begin[0]; k < int_begin[n_units]; ++k) {
    for (i = 0; i < counts[k]; ++i) {
      hbtslices_freq* f = freq_find(freqs, n, slices[s + i].stack_id);
      uint64_t j;
      int seen = 0;
      ++f->num_obs;
      f->duration += slices[s + i].duration;
      for (j = 0; j < i && !seen; ++j) seen = slices[s + j].stack_id == slices[s + i].stack_id;
      if (!seen) ++f->num_intervals;
    }
    s += counts[k];
  }
  for (i = 0; i < n; ++i)
    freqs[i].bid = (freqs[i].num_intervals >= min_distinct_intervals && freqs[i].duration >= min_duration)
                       ? next_bid++ : 0;
  *n_freqs = n;
  *num_rows = n_int;
  *num_bins = next_bid;
  if (n_int * next_bid > dur_cap) return -ENOSPC;
  memset(durations, 0, n_int * next_bid * sizeof(uint64_t));
  s = 0;
  for (u = 0; u < n_units; ++u) {
    uint64_t row = 0;
    for (k = int_begin[u]; k < int_begin[u + 1]; ++k) {
      uint64_t row_dur = 0;
      for (i = 0; i < counts[k]; ++i) {
        const hbtslices_slice* sl = &slices[s + i];
        if (sl->duration == 0) continue;
        row_dur += sl->duration;
        durations[row * next_bid + freq_find(freqs, n, sl->stack_id)->bid] += sl->duration;
      }
      s += counts[k];
      if (row_dur) ++row;
    }
  }
  return 0;
}
