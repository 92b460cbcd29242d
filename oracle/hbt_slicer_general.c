/* hbt_slicer_general.c -- TEST INFRASTRUCTURE ONLY (the product never links or loads this).
 *
 * tagstack::slice::Slicer<>::processNext (hbt/src/tagstack/Slicer.h:336-459) restated in plain C
 * for ANY event stream: switch events at every level, phase events (Start / End), the general
 * stack dictionary (Stack -> TagStackState, :313), the dormant lists by bottom tag in
 * most-recently-used order (:308-311), resets.  hbt_slices_oracle.c holds the collapsed form for
 * level-0 streams; this one is the checker of the GPU's general path.  Pinned against the real
 * class in oracle/_ref by tests/test_slicer_oracle.py (the vectors of
 * hbt/src/tagstack/tests/SlicerTest.cpp:63,150,213,323,432 and random phase streams).
 *
 * Where the reference reads indeterminate memory -- Stack::copyTruncated / copyFillUpTo leave the
 * tags above num_set_tags uninitialised (TagStack.h:103-118) and a later copyFillUpTo at a higher
 * level copies them -- this restatement has kNA there; the tests stay off such input.
 * The dormant lists follow the stand-in for boost::intrusive in oracle/stubs (push_front of a
 * linked element moves it to the front; clear() unlinks everything).
 */
#include <errno.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "hbtslices_api.h"

#define G_INVALID (~0ull)
#define G_NA 0xffffffffu
#define G_LEVELS 6
enum { EV_START = 0, EV_END, EV_CREATE, EV_DESTROY, EV_IN, EV_OUT_PREEMPT, EV_OUT_YIELD, EV_WE_START, EV_WE_END };
enum { TT_PREEMPTED = 2, TT_YIELD = 3, TT_PHASE = 4 };

typedef struct {
  uint32_t tags[G_LEVELS];
  uint8_t nset, known;
} g_stack;

typedef struct {
  g_stack st;
  uint64_t parent, cum, swins, swouts;
  int64_t dprev, dnext; /* dormant list of the bottom tag; -1: none */
  uint8_t linked;
} g_state; /* index in the array = stack id (ids are handed out in creation order, :509) */

typedef struct {
  g_state* s;
  uint64_t n, cap;     /* states; n = next_stack_id_ minus the ids burnt by failed emplaces */
  uint64_t* id_of;     /* stack id of state k (ids can be burnt, so id != index) */
  uint64_t next_id;
  int64_t* hslot;      /* open addressing: state index or -1 */
  uint64_t hcap;
  /* dormant heads: tag -> first state (most recently used) */
  uint32_t* dkey;
  int64_t* dhead;
  uint8_t* dused;
  uint64_t dcap;
  /* per unit */
  int64_t* active;
  uint64_t* swin_ts;
  uint8_t* swin_type;
  uint32_t n_units;
} g_ctx;

static g_stack stack_na(uint8_t nset, uint8_t known) { /* Stack(Level, bool), TagStack.h:92-96 */
  g_stack s;
  int i;
  for (i = 0; i < G_LEVELS; ++i) s.tags[i] = G_NA;
  s.nset = nset;
  s.known = known;
  return s;
}
static g_stack copy_truncated(const g_stack* a, uint8_t n, uint8_t known) { /* :104-112 */
  g_stack s = stack_na(n, known);
  int i;
  for (i = 0; i < n && i < G_LEVELS; ++i) s.tags[i] = a->tags[i];
  return s;
}
static g_stack copy_fill_up_to(const g_stack* a, uint8_t idx, uint32_t tag, uint8_t known) { /* :114-120 */
  g_stack s = stack_na((uint8_t)(idx + 1), known);
  int i;
  for (i = 0; i < idx; ++i) s.tags[i] = a->tags[i];
  s.tags[idx] = tag;
  return s;
}
static int stack_eq(const g_stack* a, const g_stack* b) { /* operator==, :139-149 */
  return a->known == b->known && a->nset == b->nset && memcmp(a->tags, b->tags, 4u * a->nset) == 0;
}
static int stack_contains(const g_stack* self, const g_stack* s, uint8_t num_levels) { /* :122-137 */
  if (self->known != s->known) return 0;
  if (num_levels == 0) return self->nset == 0;
  if (self->nset < num_levels) return 0;
  return memcmp(s->tags, self->tags, 4u * num_levels) == 0;
}
static int stack_has_nas(const g_stack* s) {
  int i;
  for (i = 0; i < s->nset; ++i)
    if (s->tags[i] == G_NA) return 1;
  return 0;
}
static int is_aliased(const g_stack* s) { /* Slicer::isAliased, :655-665 (kHandleTID0 = true) */
  return (s->nset == 1 && s->tags[0] == 0) || s->nset == 0 || stack_has_nas(s);
}
static uint64_t stack_hash(const g_stack* s) {
  uint64_t h = s->known ? 1469598103934665603ull : 1099511628211ull;
  int i;
  h = (h ^ s->nset) * 1099511628211ull;
  for (i = 0; i < s->nset; ++i) h = (h ^ s->tags[i]) * 1099511628211ull;
  return h ^ (h >> 29);
}

static int64_t find_state(const g_ctx* c, const g_stack* s) {
  uint64_t k = stack_hash(s) & (c->hcap - 1);
  for (;;) {
    const int64_t i = c->hslot[k];
    if (i < 0) return -1;
    if (stack_eq(&c->s[i].st, s)) return i;
    k = (k + 1) & (c->hcap - 1);
  }
}

/* createStackState_, :560-584: parents first; a stack that already exists only burns an id */
static int64_t create_state(g_ctx* c, const g_stack* s) {
  uint64_t parent = G_INVALID;
  int64_t i;
  if (s->nset > 0) {
    const g_stack p = copy_truncated(s, (uint8_t)(s->nset - 1), s->known);
    int64_t pi = find_state(c, &p);
    if (pi < 0) pi = create_state(c, &p);
    if (pi < 0) return pi;
    parent = c->id_of[pi];
  }
  i = find_state(c, s);
  if (i >= 0) { /* emplace fails, but next_stack_id_++ was evaluated */
    ++c->next_id;
    return i;
  }
  if (c->n == c->cap) return -ENOSPC;
  i = (int64_t)c->n++;
  memset(&c->s[i], 0, sizeof(g_state));
  c->s[i].st = *s;
  c->s[i].parent = parent;
  c->s[i].dprev = c->s[i].dnext = -1;
  c->id_of[i] = c->next_id++;
  {
    uint64_t k = stack_hash(s) & (c->hcap - 1);
    while (c->hslot[k] >= 0) k = (k + 1) & (c->hcap - 1);
    c->hslot[k] = i;
  }
  return i;
}

static uint64_t dslot(g_ctx* c, uint32_t tag) { /* slot of the list of `tag` (created on demand) */
  uint64_t k = ((uint64_t)tag * 0x9E3779B97F4A7C15ull >> 20) & (c->dcap - 1);
  while (c->dused[k] && c->dkey[k] != tag) k = (k + 1) & (c->dcap - 1);
  if (!c->dused[k]) {
    c->dused[k] = 1;
    c->dkey[k] = tag;
    c->dhead[k] = -1;
  }
  return k;
}
static void dormant_unlink(g_ctx* c, int64_t i) {
  g_state* s = &c->s[i];
  if (!s->linked) return;
  if (s->dprev >= 0) c->s[s->dprev].dnext = s->dnext;
  else c->dhead[dslot(c, s->st.tags[0])] = s->dnext;
  if (s->dnext >= 0) c->s[s->dnext].dprev = s->dprev;
  s->dprev = s->dnext = -1;
  s->linked = 0;
}
static void dormant_push_front(g_ctx* c, int64_t i) {
  uint64_t k;
  dormant_unlink(c, i);
  k = dslot(c, c->s[i].st.tags[0]);
  c->s[i].dnext = c->dhead[k];
  c->s[i].dprev = -1;
  if (c->dhead[k] >= 0) c->s[c->dhead[k]].dprev = i;
  c->dhead[k] = i;
  c->s[i].linked = 1;
}

static void reset_state(g_ctx* c) { /* resetState_, :636-652 */
  uint64_t i;
  uint32_t g;
  for (g = 0; g < c->n_units; ++g) {
    c->active[g] = -1;
    c->swin_ts[g] = G_INVALID; /* restart() of every state */
  }
  for (i = 0; i < c->n; ++i) {
    c->s[i].linked = 0;
    c->s[i].dprev = c->s[i].dnext = -1;
  }
  memset(c->dused, 0, c->dcap);
}

static void activate(g_ctx* c, uint32_t g, int64_t i, uint64_t ts, uint8_t type) { /* :515-526 */
  c->active[g] = i;
  c->swin_ts[g] = ts;
  c->swin_type[g] = type;
  ++c->s[i].swins;
}
static int activate_exact(g_ctx* c, uint32_t g, const g_stack* st, uint64_t ts, uint8_t type) { /* :586-609 */
  int64_t i = find_state(c, st);
  if (i < 0) {
    i = create_state(c, st);
    if (i < 0) return (int)i;
  } else if (c->s[i].linked) {
    dormant_unlink(c, i);
  }
  activate(c, g, i, ts, type);
  return 0;
}
static int activate_that_contains(g_ctx* c, uint32_t g, const g_stack* st, uint8_t num_levels, uint64_t ts,
                                  uint8_t type) { /* :611-634 */
  int64_t i = c->dhead[dslot(c, st->tags[0])];
  g_stack unknown;
  for (; i >= 0; i = c->s[i].dnext) {
    if (stack_contains(&c->s[i].st, st, num_levels)) {
      dormant_unlink(c, i);
      activate(c, g, i, ts, type);
      return 0;
    }
  }
  unknown = *st;
  unknown.known = 0;
  return activate_exact(c, g, &unknown, ts, type);
}

static int stack_cmp(const void* a, const void* b) {
  const uint64_t x = ((const hbtslices_stack*)a)->stack_id, y = ((const hbtslices_stack*)b)->stack_id;
  return x < y ? -1 : x > y;
}

int hbtoracle_slicer_general_run(const hbtslices_event* events, uint64_t n_events, uint32_t n_units,
                                 hbtslices_slice* out, uint64_t cap_per_unit, uint64_t* n_slices,
                                 hbtslices_slicer_stats* stats, hbtslices_stack* stacks, uint64_t stack_cap,
                                 uint64_t* n_stacks, uint64_t* last_ts_out) {
  g_ctx c;
  uint64_t i, last_ts = 0;
  uint32_t num_active_write_errors = 0, g;
  int rc = 0;
  memset(&c, 0, sizeof(c));
  memset(stats, 0, sizeof(*stats));
  *n_stacks = 0;
  c.n_units = n_units;
  c.cap = 7 * n_events + 8;
  c.hcap = 16;
  while (c.hcap < 4 * c.cap) c.hcap <<= 1;
  c.dcap = 16;
  while (c.dcap < 4 * (n_events + 4)) c.dcap <<= 1;
  c.s = (g_state*)calloc(c.cap, sizeof(g_state));
  c.id_of = (uint64_t*)calloc(c.cap, sizeof(uint64_t));
  c.hslot = (int64_t*)malloc(c.hcap * sizeof(int64_t));
  c.dkey = (uint32_t*)calloc(c.dcap, sizeof(uint32_t));
  c.dhead = (int64_t*)calloc(c.dcap, sizeof(int64_t));
  c.dused = (uint8_t*)calloc(c.dcap, 1);
  c.active = (int64_t*)malloc((n_units ? n_units : 1) * sizeof(int64_t));
  c.swin_ts = (uint64_t*)malloc((n_units ? n_units : 1) * sizeof(uint64_t));
  c.swin_type = (uint8_t*)calloc(n_units ? n_units : 1, 1);
  if (!c.s || !c.id_of || !c.hslot || !c.dkey || !c.dhead || !c.dused || !c.active || !c.swin_ts || !c.swin_type) {
    rc = -ENOMEM;
    goto done;
  }
  for (i = 0; i < c.hcap; ++i) c.hslot[i] = -1;
  for (g = 0; g < n_units; ++g) {
    c.active[g] = -1;
    c.swin_ts[g] = G_INVALID;
    n_slices[g] = 0;
  }

  for (i = 0; i < n_events && rc == 0; ++i) {
    const hbtslices_event* ev = &events[i];
    const uint64_t ts = ev->tstamp;
    const int is_out = ev->type == EV_END || ev->type == EV_OUT_PREEMPT || ev->type == EV_OUT_YIELD;
    const int is_phase = ev->type == EV_START || ev->type == EV_END;
    uint8_t sw;
    g_stack last_stack;
    int64_t la;
    g = ev->comp_unit_id;
    if (g >= n_units) { rc = -EINVAL; break; }                 /* :351-352 */
    if (ts < last_ts) {                                         /* :358-370; the event is then processed */
      ++stats->num_out_of_order;
      reset_state(&c);
      last_ts = ts;
    }
    if (ev->type == EV_WE_END) {                                /* handleWriteErrors_, :668-703 */
      last_ts = ts; ++stats->num_processed; ++stats->num_error_end; --num_active_write_errors;
      continue;
    }
    if (ev->type == EV_WE_START) {
      last_ts = ts; ++stats->num_processed; ++stats->num_error_start; reset_state(&c); ++num_active_write_errors;
      continue;
    }
    if (num_active_write_errors) {
      last_ts = ts; ++stats->num_processed; ++stats->num_dropped;
      continue;
    }
    if (ev->type > EV_OUT_YIELD) { rc = -EINVAL; break; }      /* :384-385 */
    if (ev->type == EV_CREATE || ev->type == EV_DESTROY) {      /* handleThreadLifetime_, :705-745 */
      if (ev->level != 0) {
        last_ts = ts; ++stats->num_processed; ++stats->num_dropped; ++stats->num_logic_errors;
        continue;
      }
      if (ev->type == EV_CREATE) {
        g_stack s = stack_na(1, 1);
        int64_t k;
        s.tags[0] = ev->tag;
        k = create_state(&c, &s);
        if (k < 0) { rc = (int)k; break; }
        dormant_push_front(&c, k);
      }
      last_ts = ts; ++stats->num_processed;
      continue;
    }
    if (ev->level >= G_LEVELS) { rc = -EINVAL; break; }         /* HBT_DCHECK_LT, :431 */
    sw = is_phase ? TT_PHASE : (ev->type == EV_OUT_PREEMPT ? TT_PREEMPTED : TT_YIELD); /* :789-805 */
    la = c.active[g];
    if (la >= 0) {                                              /* handleLastState_, :747-787 */
      g_state* s = &c.s[la];
      uint64_t dur;
      if (!is_aliased(&s->st)) dormant_push_front(&c, la);
      if (ts < c.swin_ts[g]) { rc = -EINVAL; break; }           /* HBT_ARG_CHECK_GE, :543 */
      dur = ts - c.swin_ts[g];                                  /* completeSlice_, :529-558 */
      if (dur > 0) {
        if (n_slices[g] < cap_per_unit) {
          hbtslices_slice* o = &out[(uint64_t)g * cap_per_unit + n_slices[g]];
          memset(o, 0, sizeof(*o));
          o->tstamp = c.swin_ts[g];
          o->duration = dur;
          o->stack_id = c.id_of[la];
          o->swin_type = c.swin_type[g];
          o->swout_type = sw;
        } else {
          rc = -ENOSPC;
        }
        ++n_slices[g];
      }
      s->cum += dur;
      ++s->swouts;
      ++stats->num_slices_completed;
      c.swin_ts[g] = G_INVALID;
      c.active[g] = -1;
      last_stack = s->st;
    } else {
      last_stack = stack_na((uint8_t)(ev->level + (is_out ? 1 : 0)), (uint8_t)is_phase); /* get_last_stack, :405-412 */
      if (last_stack.nset >= G_LEVELS) { rc = -EINVAL; break; }
    }
    if (is_out) {
      g_stack ns;
      if (last_stack.known && last_stack.nset > 0 && last_stack.nset <= ev->level) { /* :808-837 */
        last_ts = ts; ++stats->num_processed; ++stats->num_error_out_sync; reset_state(&c);
        continue;
      }
      if (la >= 0 && ev->level > last_stack.nset) { rc = -ENOTSUP; break; } /* fence: copyTruncated would copy
                                                                    tags the reference never set */
      ns = copy_truncated(&last_stack, ev->level, 1);           /* :422-428 */
      rc = activate_exact(&c, g, &ns, ts, sw);
    } else {
      g_stack bottom;
      uint8_t known;
      if (last_stack.nset > ev->level && last_stack.tags[ev->level] != G_NA) { /* :839-868 */
        last_ts = ts; reset_state(&c); ++stats->num_processed;
        continue;
      }
      if (la >= 0 && ev->level > last_stack.nset) { rc = -ENOTSUP; break; } /* fence: copyFillUpTo would copy
                                                                    tags the reference never set (see the header) */
      known = (uint8_t)(last_stack.known | is_phase);           /* :433 */
      bottom = copy_fill_up_to(&last_stack, ev->level, ev->tag, known);
      if (is_aliased(&bottom) || is_phase) rc = activate_exact(&c, g, &bottom, ts, sw);
      else rc = activate_that_contains(&c, g, &bottom, (uint8_t)(ev->level + 1), ts, sw);
    }
    if (rc) break;
    last_ts = ts;
    ++stats->num_processed;
  }
  *last_ts_out = last_ts;
  if (rc == 0 || rc == -ENOSPC) { /* getTagStackStatsDump, :478-485: every state, by stack id */
    uint64_t k;
    *n_stacks = c.n;
    for (k = 0; k < c.n; ++k) {
      hbtslices_stack* o;
      int t;
      if (k >= stack_cap) { rc = -ENOSPC; break; }
      o = &stacks[k];
      memset(o, 0, sizeof(*o));
      o->stack_id = c.id_of[k];
      o->parent_stack_id = c.s[k].parent;
      for (t = 0; t < G_LEVELS; ++t) o->tags[t] = t < c.s[k].st.nset ? c.s[k].st.tags[t] : G_NA;
      o->num_set_tags = c.s[k].st.nset;
      o->known_num_levels = c.s[k].st.known;
      o->cum_active_time = c.s[k].cum;
      o->num_swins = c.s[k].swins;
      o->num_swouts = c.s[k].swouts;
    }
    if (c.n <= stack_cap) qsort(stacks, c.n, sizeof(*stacks), stack_cmp);
  }
done:
  free(c.s); free(c.id_of); free(c.hslot); free(c.dkey); free(c.dhead); free(c.dused);
  free(c.active); free(c.swin_ts); free(c.swin_type);
  return rc;
}
