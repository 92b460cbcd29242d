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

/* ---- row f3, Slicer side: tagstack::slice::Slicer<>::processNext (hbt/src/tagstack/Slicer.h) ----
 *
 * Restated for the events the thread-switch decode produces: every event at level 0
 * (ThreadCreation / ThreadDestruction / SwitchIn / SwitchOut* / WriteErrors*).  Phase events
 * (Start / End) and switches at other levels return -ENOTSUP: their stacks need the general
 * dictionary, which this restatement does not carry (the real Slicer in oracle/_ref does).
 * With only level 0 in play the stack dictionary (Slicer.h:313) collapses to
 *   E[k]      the empty stack {} with known_num_levels = k         (k = 0, 1)
 *   T[v][k]   the one-tag stack {v} with known_num_levels = k
 * and the dormant list of tag v (Slicer.h:308-311) can hold T[v][0] and T[v][1]; since
 * Stack::contains demands equal known_num_levels (TagStack.h:122-126) at most one of them matches
 * a lookup, so the list order never decides anything.
 */
#define SL_INVALID (~0ull)
#define SL_NA 0xffffffffu
enum { EV_START = 0, EV_END, EV_CREATE, EV_DESTROY, EV_IN, EV_OUT_PREEMPT, EV_OUT_YIELD, EV_WE_START, EV_WE_END };
enum { TT_PREEMPTED = 2, TT_YIELD = 3 };

typedef struct {
  uint8_t exists, linked, nset, known;
  uint32_t tag;
  uint64_t id, parent, cum, swins, swouts;
} sl_state;

typedef struct {
  sl_state* st; /* active state of the group, NULL if none */
  uint64_t swin_ts;
  uint8_t swin_type;
} sl_active;

typedef struct {
  sl_state E[2];
  sl_state* T; /* [n_tags][2] */
  uint64_t n_tags, next_id;
  sl_active* act;
  uint32_t n_units;
} sl_ctx;

static int sl_aliased(const sl_state* s) { /* Slicer::isAliased, :655-665 (kHandleTID0 = true) */
  return s->nset == 0 || (s->nset == 1 && s->tag == 0) || (s->nset == 1 && s->tag == SL_NA);
}

static void sl_create(sl_ctx* c, sl_state* s) { /* createStackState_, :560-584 */
  uint64_t parent = SL_INVALID;
  if (s->nset > 0) {
    sl_state* p = &c->E[s->known];
    if (!p->exists) sl_create(c, p);
    parent = p->id;
  }
  if (s->exists) { /* emplace fails, but next_stack_id_++ was evaluated */
    ++c->next_id;
    return;
  }
  s->exists = 1;
  s->id = c->next_id++;
  s->parent = parent;
}

static void sl_activate(sl_ctx* c, uint32_t g, sl_state* s, uint64_t ts, uint8_t type) { /* :515-526, 231-241 */
  c->act[g].st = s;
  c->act[g].swin_ts = ts;
  c->act[g].swin_type = type;
  ++s->swins;
}

static void sl_activate_exact(sl_ctx* c, uint32_t g, sl_state* s, uint64_t ts, uint8_t type) { /* :586-609 */
  if (!s->exists) sl_create(c, s);
  else if (s->linked) s->linked = 0;
  sl_activate(c, g, s, ts, type);
}

static void sl_reset(sl_ctx* c) { /* resetState_, :636-652 */
  uint64_t i;
  uint32_t g;
  for (g = 0; g < c->n_units; ++g) c->act[g].st = 0;
  c->E[0].linked = c->E[1].linked = 0;
  for (i = 0; i < 2 * c->n_tags; ++i) c->T[i].linked = 0;
}

static int stack_cmp(const void* a, const void* b) {
  const uint64_t x = ((const hbtslices_stack*)a)->stack_id, y = ((const hbtslices_stack*)b)->stack_id;
  return x < y ? -1 : x > y;
}

int hbtoracle_slicer_run(const hbtslices_event* events, uint64_t n_events, uint32_t n_units,
                         hbtslices_slice* out, uint64_t cap_per_unit, uint64_t* n_slices,
                         hbtslices_slicer_stats* stats, hbtslices_stack* stacks, uint64_t stack_cap,
                         uint64_t* n_stacks, uint64_t* last_ts_out) {
  sl_ctx c;
  uint64_t i, last_ts = 0, max_tag = 0, n_all;
  uint32_t num_active_write_errors = 0, g;
  int rc = 0;
  memset(&c, 0, sizeof(c));
  memset(stats, 0, sizeof(*stats));
  *n_stacks = 0;
  for (i = 0; i < n_events; ++i)
    if (events[i].tag != SL_NA && events[i].tag > max_tag) max_tag = events[i].tag;
  if (max_tag >= (1u << 26)) return -ENOTSUP;
  c.n_tags = max_tag + 1;
  c.n_units = n_units;
  c.T = (sl_state*)calloc(2 * c.n_tags, sizeof(sl_state));
  c.act = (sl_active*)calloc(n_units ? n_units : 1, sizeof(sl_active));
  if (!c.T || !c.act) {
    free(c.T);
    free(c.act);
    return -ENOMEM;
  }
  c.E[1].known = 1;
  for (i = 0; i < c.n_tags; ++i) {
    c.T[2 * i].nset = c.T[2 * i + 1].nset = 1;
    c.T[2 * i].tag = c.T[2 * i + 1].tag = (uint32_t)i;
    c.T[2 * i + 1].known = 1;
  }
  for (g = 0; g < n_units; ++g) n_slices[g] = 0;

  for (i = 0; i < n_events && rc == 0; ++i) {
    const hbtslices_event* ev = &events[i];
    const uint64_t ts = ev->tstamp;
    sl_active* a;
    uint8_t sw, last_known, last_nset;
    uint32_t last_tag0;
    g = ev->comp_unit_id;
    if (g >= n_units) { rc = -EINVAL; break; }                 /* :351-352 */
    if (ts < last_ts) {                                         /* :358-370; the event is then processed */
      ++stats->num_out_of_order;
      sl_reset(&c);
      last_ts = ts;
    }
    if (ev->type == EV_WE_END) {                                /* handleWriteErrors_, :668-703 */
      last_ts = ts; ++stats->num_processed; ++stats->num_error_end; --num_active_write_errors;
      continue;
    }
    if (ev->type == EV_WE_START) {
      last_ts = ts; ++stats->num_processed; ++stats->num_error_start; sl_reset(&c); ++num_active_write_errors;
      continue;
    }
    if (num_active_write_errors) {
      last_ts = ts; ++stats->num_processed; ++stats->num_dropped;
      continue;
    }
    if (ev->type > EV_OUT_YIELD) { rc = -EINVAL; break; }      /* :384-385 */
    if (ev->type == EV_START || ev->type == EV_END) { rc = -ENOTSUP; break; }
    if (ev->type == EV_CREATE || ev->type == EV_DESTROY) {      /* handleThreadLifetime_, :705-745 */
      if (ev->level != 0) {
        last_ts = ts; ++stats->num_processed; ++stats->num_dropped; ++stats->num_logic_errors;
        continue;
      }
      if (ev->type == EV_CREATE) {
        sl_state* s;
        if (ev->tag == SL_NA) { rc = -ENOTSUP; break; }
        s = &c.T[2 * (uint64_t)ev->tag + 1];
        sl_create(&c, s);
        s->linked = 1; /* dormant_states_[tag].push_front */
      }
      last_ts = ts; ++stats->num_processed;
      continue;
    }
    if (ev->level != 0) { rc = -ENOTSUP; break; }
    sw = ev->type == EV_OUT_PREEMPT ? TT_PREEMPTED : TT_YIELD; /* eventToTransitionType_, :789-805 */
    a = &c.act[g];
    last_known = 0;
    last_nset = ev->type == EV_IN ? 0 : 1;                      /* get_last_stack, :405-412 */
    last_tag0 = SL_NA;
    if (a->st) {                                                /* handleLastState_, :747-787 */
      sl_state* s = a->st;
      uint64_t dur;
      if (!sl_aliased(s)) s->linked = 1;
      if (ts < a->swin_ts) { rc = -EINVAL; break; }             /* HBT_ARG_CHECK_GE, :543 */
      dur = ts - a->swin_ts;                                    /* completeSlice_, :529-558 */
      if (dur > 0) {
        if (n_slices[g] < cap_per_unit) {
          hbtslices_slice* o = &out[(uint64_t)g * cap_per_unit + n_slices[g]];
          memset(o, 0, sizeof(*o));
          o->tstamp = a->swin_ts;
          o->duration = dur;
          o->stack_id = s->id;
          o->swin_type = a->swin_type;
          o->swout_type = sw;
        } else {
          rc = -ENOSPC;
        }
        ++n_slices[g];
      }
      s->cum += dur;
      ++s->swouts;
      ++stats->num_slices_completed;
      last_known = s->known;
      last_nset = s->nset;
      last_tag0 = s->nset ? s->tag : SL_NA;
      a->st = 0;
    }
    if (ev->type != EV_IN) {
      /* handleEventOutPreConditionError_ (:808-837) needs num_set_tags <= level = 0: never */
      sl_activate_exact(&c, g, &c.E[1], ts, sw);                /* copyTruncated(0, true), :422-428 */
    } else {
      sl_state* s;
      if (last_nset > 0 && last_tag0 != SL_NA) {                /* handleEventInPreConditionError_, :839-868 */
        last_ts = ts; sl_reset(&c); ++stats->num_processed;
        continue;
      }
      if (ev->tag == SL_NA) { rc = -ENOTSUP; break; }
      s = &c.T[2 * (uint64_t)ev->tag + last_known];            /* bottom_stack = {tag}, known of the last stack */
      if (sl_aliased(s)) {
        sl_activate_exact(&c, g, s, ts, sw);                    /* :443-445 */
      } else if (s->exists && s->linked) {                      /* activateStackStateThatContains_, :611-626 */
        s->linked = 0;
        sl_activate(&c, g, s, ts, sw);
      } else {
        sl_activate_exact(&c, g, &c.T[2 * (uint64_t)ev->tag], ts, sw); /* {tag} with unknown levels, :632-633 */
      }
    }
    last_ts = ts;
    ++stats->num_processed;
  }
  *last_ts_out = last_ts;
  /* getTagStackStatsDump, :478-485: every state, by stack id */
  n_all = 0;
  if (rc == 0 || rc == -ENOSPC) {
    uint64_t k;
    for (k = 0; k < 2 + 2 * c.n_tags; ++k) {
      const sl_state* s = k < 2 ? &c.E[k] : &c.T[k - 2];
      if (!s->exists) continue;
      if (n_all < stack_cap) {
        hbtslices_stack* o = &stacks[n_all];
        int t;
        memset(o, 0, sizeof(*o));
        o->stack_id = s->id;
        o->parent_stack_id = s->parent;
        for (t = 0; t < 6; ++t) o->tags[t] = SL_NA;
        if (s->nset) o->tags[0] = s->tag;
        o->num_set_tags = s->nset;
        o->known_num_levels = s->known;
        o->cum_active_time = s->cum;
        o->num_swins = s->swins;
        o->num_swouts = s->swouts;
      } else {
        rc = -ENOSPC;
      }
      ++n_all;
    }
    *n_stacks = n_all;
    if (n_all <= stack_cap) qsort(stacks, n_all, sizeof(*stacks), stack_cmp); /* ids are unique */
  }
  free(c.T);
  free(c.act);
  return rc;
}
