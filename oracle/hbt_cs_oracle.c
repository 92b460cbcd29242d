/*
 * oracle/hbt_cs_oracle.c -- plain-C restatement of hbt's ContextSwitch-mode decode
 * (SURVEY.md 8(f) row f2).  TEST INFRASTRUCTURE ONLY: see oracle/hbtcs_api.h.
 * Pinned against the real reference (oracle/_ref/libhbtref.so, hbtref_cs_run) by
 * tests/test_cs_oracle.py.
 *
 * Follows, record by record:
 *   PerfEventsGroup<CpuThreadSwitchGenerator, mode::ContextSwitch>::consume
 *       hbt/src/perf_event/PerfEventsGroup.h:1394-1589 (dispatch, counters)
 *   CpuThreadSwitchGenerator::handleRecord{Lost,Comm,Exit,Fork,SwitchCpuWide}, pendingRecordLost_
 *       hbt/src/perf_event/PerCpuThreadSwitchGenerator.h:219-446
 *   PerThreadStats::{createThreadInfo,selectNextVid,registerFork,registerExit}  :52-135
 *   tagstack::Event factories  hbt/src/tagstack/Event.h:65-176
 *
 * Fences (the reference misbehaves; same rules as the Sampling-mode oracle): a record with
 * size 0, size % 8 != 0, running past the stream end, or too short for the fields its handler
 * reads stops the walk with -EPERM.  Inputs on which the reference's noexcept handlers would
 * throw (std::terminate) are out of contract and not reproduced: COMM/EXIT for tid 0
 * (info.at(0)), a second EXIT of an active vid, a FORK of an already forked active tid, a COMM
 * at or before its thread's fork time (:262-265, :125-131, :322-325).
 * The output ring is assumed large enough (no -ENOSPC paths).
 */
#include <errno.h>
#include <stdlib.h>
#include <string.h>

#include "hbtcs_api.h"

enum { R_LOST = 2, R_COMM = 3, R_EXIT = 4, R_THROTTLE = 5, R_UNTHROTTLE = 6, R_FORK = 7, R_SWITCH_CPU_WIDE = 15 };
enum { EV_CREATE = 2, EV_DESTROY = 3, EV_IN = 4, EV_OUT_PREEMPT = 5, EV_OUT_YIELD = 6, EV_WE_START = 7, EV_WE_END = 8 };
#define MISC_SWITCH_OUT (1u << 13)
#define MISC_SWITCH_OUT_PREEMPT (1u << 14)
#define K_NA 0xFFFFFFFFu
#define K_INVALID_LEVEL 255
#define K_INVALID_TS (~0ull)

typedef struct {
  uint64_t ts;
  uint8_t is_exec, len;
  char name[16];
} name_pt;

typedef struct {
  int32_t ppid, ptid, pid, tid;
  uint64_t fork_ts, end_ts;
  name_pt* names;
  uint32_t n_names, cap_names;
} thread_info;

/* PerThreadStats (:52-135): info by vid, active_tids as an open-addressing map tid -> vid */
typedef struct {
  uint32_t next_vid;
  thread_info* info; /* info[vid - 1] */
  uint32_t cap_info;
  int32_t* keys; /* -1 empty, -2 deleted */
  uint32_t* vals;
  uint32_t cap, used;
} stats_t;

static uint32_t hash_tid(int32_t t) {
  uint32_t x = (uint32_t)t * 0x9E3779B1u;
  return x ^ (x >> 15);
}

static void map_clear(stats_t* st) {
  for (uint32_t i = 0; i < st->cap; ++i) st->keys[i] = -1;
  st->used = 0;
}

static void map_put_raw(int32_t* keys, uint32_t* vals, uint32_t cap, int32_t k, uint32_t v) {
  uint32_t i = hash_tid(k) & (cap - 1);
  while (keys[i] >= 0) i = (i + 1) & (cap - 1);
  keys[i] = k;
  vals[i] = v;
}

static void map_grow(stats_t* st) {
  uint32_t ncap = st->cap ? st->cap * 2 : 1024;
  int32_t* nk = (int32_t*)malloc(sizeof(int32_t) * ncap);
  uint32_t* nv = (uint32_t*)malloc(sizeof(uint32_t) * ncap);
  for (uint32_t i = 0; i < ncap; ++i) nk[i] = -1;
  uint32_t live = 0;
  for (uint32_t i = 0; i < st->cap; ++i)
    if (st->keys[i] >= 0) {
      map_put_raw(nk, nv, ncap, st->keys[i], st->vals[i]);
      ++live;
    }
  free(st->keys);
  free(st->vals);
  st->keys = nk;
  st->vals = nv;
  st->cap = ncap;
  st->used = live;
}

static int map_find(const stats_t* st, int32_t k, uint32_t* slot) {
  if (!st->cap) return 0;
  uint32_t i = hash_tid(k) & (st->cap - 1);
  while (st->keys[i] != -1) {
    if (st->keys[i] == k) {
      *slot = i;
      return 1;
    }
    i = (i + 1) & (st->cap - 1);
  }
  return 0;
}

static void map_put(stats_t* st, int32_t k, uint32_t v) {
  if ((st->used + 1) * 2 > st->cap) map_grow(st);
  map_put_raw(st->keys, st->vals, st->cap, k, v);
  st->used += 1; /* deleted slots stay counted until the next grow */
}

static void map_erase(stats_t* st, int32_t k) {
  uint32_t s;
  if (map_find(st, k, &s)) st->keys[s] = -2;
}

/* selectNextVid (:96-108): tid 0 is never remapped */
static uint32_t select_vid(const stats_t* st, int32_t tid, int* is_new) {
  uint32_t s;
  *is_new = 0;
  if (tid == 0) return 0;
  if (map_find(st, tid, &s)) return st->vals[s];
  *is_new = 1;
  return st->next_vid;
}

/* createThreadInfo (:60-91); cmdline / proc comm are host lookups and are left out */
static thread_info* create_thread(stats_t* st, int32_t ppid, int32_t ptid, int32_t pid, int32_t tid) {
  uint32_t vid = st->next_vid++;
  if (vid > st->cap_info) {
    uint32_t nc = st->cap_info ? st->cap_info * 2 : 1024;
    st->info = (thread_info*)realloc(st->info, sizeof(thread_info) * nc);
    st->cap_info = nc;
  }
  thread_info* t = &st->info[vid - 1];
  memset(t, 0, sizeof(*t));
  t->ppid = ppid;
  t->ptid = ptid;
  t->pid = pid;
  t->tid = tid;
  t->fork_ts = K_INVALID_TS;
  t->end_ts = K_INVALID_TS;
  map_put(st, tid, vid);
  return t;
}

/* Series::emplace (utils/ValueTimeSeries.h:46-52): append, sort by tstamp if out of order */
static void names_emplace(thread_info* t, uint64_t ts, const uint8_t* name, uint32_t len, int is_exec) {
  if (t->n_names == t->cap_names) {
    t->cap_names = t->cap_names ? t->cap_names * 2 : 4;
    t->names = (name_pt*)realloc(t->names, sizeof(name_pt) * t->cap_names);
  }
  uint32_t i = t->n_names++;
  while (i > 0 && t->names[i - 1].ts > ts) {
    t->names[i] = t->names[i - 1];
    --i;
  }
  name_pt* p = &t->names[i];
  memset(p, 0, sizeof(*p));
  p->ts = ts;
  p->is_exec = (uint8_t)is_exec;
  p->len = (uint8_t)(len > 16 ? 16 : len);
  memcpy(p->name, name, p->len);
}

static uint32_t rd32(const uint8_t* p) {
  uint32_t v;
  memcpy(&v, p, 4);
  return v;
}
static uint64_t rd64(const uint8_t* p) {
  uint64_t v;
  memcpy(&v, p, 8);
  return v;
}

typedef struct {
  hbtcs_event* out;
  uint64_t cap, n;
  uint16_t cu;
  uint8_t level;
} sink_t;

static void emit(sink_t* k, uint64_t ts, uint8_t type, uint8_t level, uint32_t tag) {
  if (k->out && k->n < k->cap) {
    hbtcs_event* e = &k->out[k->n];
    e->tstamp = ts;
    e->type = type;
    e->level = level;
    e->comp_unit_id = k->cu;
    e->tag = tag;
  }
  k->n += 1;
}

/* pendingRecordLost_ (:424-435) */
static void pending_lost(sink_t* k, hbtcs_state* s, uint64_t ts) {
  if (!s->in_write_errors) return;
  emit(k, ts, EV_WE_END, K_INVALID_LEVEL, K_NA);
  s->last_tstamp = ts;
  s->in_write_errors = 0;
}

int hbtoracle_cs_run(uint32_t n_streams, const uint8_t* const* bytes, const uint64_t* lens,
    uint32_t tid_level, uint32_t batch, const hbtcs_state* in, hbtcs_state* out,
    hbtcs_event* events, const uint64_t* ev_off, const uint64_t* ev_cap, uint64_t* n_events,
    hbtref_counters* ctrs, hbtcs_thread* threads, uint64_t thread_cap, uint64_t* n_threads,
    hbtcs_name* names, uint64_t name_cap, uint64_t* n_names) {
  (void)batch; /* how many records one consume() call takes changes nothing observable */
  stats_t st;
  memset(&st, 0, sizeof(st));
  st.next_vid = 1;
  for (uint32_t s = 0; s < n_streams; ++s) {
    const uint8_t* b = bytes[s];
    const uint64_t len = lens[s];
    hbtcs_state state = {0, 0, 0}; /* last_tstamp_ = numeric_limits<TimeStamp>::min() (:419) */
    if (in) state = in[s];
    sink_t k = {events ? events + ev_off[s] : NULL, events ? ev_cap[s] : 0, 0, (uint16_t)s, (uint8_t)tid_level};
    hbtref_counters c;
    memset(&c, 0, sizeof(c));
    uint64_t pos = 0;
    while (pos < len) {
      if (pos + 8 > len) {
        c.err = -EPERM;
        break;
      }
      const uint64_t h = rd64(b + pos);
      const uint32_t type = (uint32_t)h, misc = (uint32_t)(h >> 32) & 0xFFFFu, size = (uint32_t)(h >> 48);
      if (size < 8 || (size & 7) || pos + size > len) {
        c.err = -EPERM;
        break;
      }
      const uint8_t* r = b + pos;
      int bad = 0;
      switch (type) {
        case R_LOST: { /* :219-253, consume :1440-1450 */
          if (size < 40) { bad = 1; break; }
          const uint64_t ts = rd64(r + 32);
          pending_lost(&k, &state, ts);
          emit(&k, state.last_tstamp + 1, EV_WE_START, K_INVALID_LEVEL, K_NA);
          state.last_tstamp = ts;
          state.in_write_errors = 1;
          map_clear(&st);
          c.lost += rd64(r + 16);
        } break;
        case R_COMM: { /* :255-292, consume :1451-1464; is_exec = misc | COMM_EXEC is always true */
          if (size < 32) { bad = 1; break; }
          const uint32_t name_len = size - 32;
          const uint64_t ts = rd64(r + size - 8);
          const int32_t pid = (int32_t)rd32(r + 8), tid = (int32_t)rd32(r + 12);
          pending_lost(&k, &state, ts);
          int is_new;
          uint32_t vid = select_vid(&st, tid, &is_new);
          if (is_new) create_thread(&st, -1, -1, pid, tid);
          if (vid >= 1 && vid < st.next_vid) names_emplace(&st.info[vid - 1], ts, r + 16, name_len, 1);
          state.last_tstamp = ts;
          c.comm += 1;
        } break;
        case R_EXIT:   /* :294-328 */
        case R_FORK: { /* :330-372 */
          if (size < 48) { bad = 1; break; }
          const int32_t pid = (int32_t)rd32(r + 8), ppid = (int32_t)rd32(r + 12);
          const int32_t tid = (int32_t)rd32(r + 16), ptid = (int32_t)rd32(r + 20);
          const uint64_t ts = rd64(r + 24);
          pending_lost(&k, &state, ts);
          int is_new;
          uint32_t vid = select_vid(&st, tid, &is_new);
          if (type == R_EXIT) {
            emit(&k, ts, EV_DESTROY, k.level, vid);
            if (is_new) create_thread(&st, ppid, ptid, pid, tid);
            if (vid >= 1) { /* registerExit :122-134 */
              st.info[vid - 1].end_ts = ts;
              map_erase(&st, tid);
            }
            c.exit += 1;
          } else {
            emit(&k, ts, EV_CREATE, k.level, vid);
            if (!is_new) {
              if (vid >= 1) st.info[vid - 1].fork_ts = ts;
            } else { /* registerFork :116-120 */
              create_thread(&st, ppid, ptid, pid, tid)->fork_ts = ts;
            }
            c.fork += 1;
          }
          state.last_tstamp = ts;
        } break;
        case R_SWITCH_CPU_WIDE: { /* :374-414 */
          if (size < 32) { bad = 1; break; }
          const int32_t pid = (int32_t)rd32(r + 16), tid = (int32_t)rd32(r + 20);
          const uint64_t ts = rd64(r + 24);
          pending_lost(&k, &state, ts);
          if (misc & MISC_SWITCH_OUT) {
            emit(&k, ts, (misc & MISC_SWITCH_OUT_PREEMPT) ? EV_OUT_PREEMPT : EV_OUT_YIELD, k.level, K_NA);
          } else {
            int is_new;
            uint32_t vid = select_vid(&st, tid, &is_new);
            emit(&k, ts, EV_IN, k.level, vid);
            if (is_new) create_thread(&st, -1, -1, pid, tid);
          }
          state.last_tstamp = ts;
          c.switch_cpu_wide += 1;
        } break;
        case R_THROTTLE:
          c.throttle += 1;
          break;
        case R_UNTHROTTLE:
          c.unthrottle += 1;
          break;
        default: /* SAMPLE, READ, AUX, ITRACE_START are void in this mode; unknown types are skipped */
          break;
      }
      if (bad) {
        c.err = -EPERM;
        break;
      }
      c.num_records += 1;
      pos += size;
    }
    c.num_rows = k.n;
    c.err_offset = pos;
    c.last_tstamp = state.last_tstamp;
    if (ctrs) ctrs[s] = c;
    if (n_events) n_events[s] = k.n;
    if (out) out[s] = state;
  }
  uint64_t nt = 0, nn = 0;
  for (uint32_t vid = 1; vid < st.next_vid; ++vid) {
    const thread_info* t = &st.info[vid - 1];
    for (uint32_t i = 0; i < t->n_names; ++i) {
      if (names && nn < name_cap) {
        hbtcs_name* o = &names[nn];
        memset(o, 0, sizeof(*o));
        o->vid = vid;
        o->is_exec = t->names[i].is_exec;
        o->len = t->names[i].len;
        o->tstamp = t->names[i].ts;
        memcpy(o->name, t->names[i].name, 16);
      }
      ++nn;
    }
    if (threads && nt < thread_cap) {
      hbtcs_thread* o = &threads[nt];
      o->vid = vid;
      o->ppid = t->ppid;
      o->ptid = t->ptid;
      o->pid = t->pid;
      o->tid = t->tid;
      o->n_names = t->n_names;
      o->fork_ts = t->fork_ts;
      o->end_ts = t->end_ts;
    }
    ++nt;
    free(t->names);
  }
  if (n_threads) *n_threads = nt;
  if (n_names) *n_names = nn;
  free(st.info);
  free(st.keys);
  free(st.vals);
  return 0;
}
