// hbtgpu_slicer.cu -- tagstack events -> slices on the GPU (Slicer side).
//
// Reference: tagstack::slice::Slicer<>::processNext (hbt/src/tagstack/Slicer.h:336-459) and what
// it calls: one state machine over the time-ordered merge of all comp units' events
// (Event::happenedBefore, tagstack/Event.h:217-223), a dictionary Stack -> TagStackState whose
// ids follow the order of creation (:560-584), dormant lists per bottom tag (:308-311), resets.
//
// Scope of this form: the events the thread-switch decode emits -- everything at level 0
// (ThreadCreation / ThreadDestruction / SwitchIn / SwitchOut* / WriteErrors*), every unit's
// events in time order.  Then the dictionary is E[k] (the empty stack) and T[v][k] (the stack
// {v}), k = known_num_levels, and:
//   * a switch-out on unit g closes the slice of g's active state and activates E[1];
//   * a switch-in of v closes it too and activates T[v][1] if g's previous event was a
//     switch-out (known levels) and T[v][1] is dormant, else T[v][0]; v = 0 takes T[0][k];
//   * WriteErrorsStart resets (active states, dormant flags); events between it and its End
//     are dropped; a switch-in on a unit whose active state is a thread closes that slice, resets
//     everything and activates nothing (:839-868).  Whether a switch-in is such depends on the
//     resets before it, so these are settled in merge order: link, take the earliest, repeat;
//   * ids: one per createStackState_ call in merge order, parents first; a ThreadCreation of
//     an existing stack still takes an id.
// Anything else -- phase events, other levels, a unit whose events go back in time, a
// WriteErrorsEnd without its Start -- is refused with -ENOTSUP; nothing is computed on the CPU.
// The state between two windows stays on the device (HBTGPU_SLICER_CONTINUE): a unit's active
// state travels as one extra event in front of the unit's events.
//
// Steps: (1) stable radix sort of the events by tstamp (units arrive one after the other, so
// ties keep comp-unit order = the merge order); (2) scans in merge order: write-error depth ->
// dropped events, reset epochs; (3) a max-scan in unit-major order gives every switch event its
// predecessor on the same unit (valid within an epoch): slice bounds, `known`; (4) per thread
// id v the list {creation, switch-in, the switch-out that parks it} sorted by (v, merge
// position), walked by one thread per v: which state a switch-in takes, who creates what;
// (5) creation counts -> exclusive sum -> stack ids; (6) slices, compacted where the duration
// is 0; statistics by atomics.  cub's radix sort and scans are library code; the rest is here.
#include <cuda_runtime.h>
#include <errno.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <string>
#include <vector>

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_select.cuh>
#include <cub/iterator/counting_input_iterator.cuh>
#include <cub/device/device_scan.cuh>

#include "hbtgpu.h"

namespace {

constexpr uint32_t EV_START = 0, EV_END = 1, EV_CREATE = 2, EV_DESTROY = 3, EV_IN = 4, EV_OUT_PREEMPT = 5,
                   EV_OUT_YIELD = 6, EV_WE_START = 7, EV_WE_END = 8;
constexpr uint8_t TT_PREEMPTED = 2, TT_YIELD = 3;  // Slice::TransitionType, Slicer.h:31-43
constexpr uint32_t NA = 0xffffffffu;
constexpr uint32_t NOIDX = 0xffffffffu;
constexpr uint64_t NONE64 = ~0ull;
// The state a unit was left in by the previous window travels as one extra event in front of the
// unit's events (level VIRT_K0 / VIRT_K1: the state has unknown / known levels; type DESTROY = none).
constexpr uint8_t VIRT_K0 = 0xee, VIRT_K1 = 0xef;
__device__ __forceinline__ bool is_virtual(const hbtgpu_cs_event& e) { return e.level >= VIRT_K0 && e.level <= VIRT_K1; }

enum : uint32_t {  // why a window is outside the parallel form
  BAD_TYPE = 1, BAD_LEVEL = 2, BAD_ORDER = 4, BAD_UNIT = 8, BAD_TAG = 16, BAD_DOUBLE_IN = 32, BAD_WE_END = 64
};
// per-event creation flags (step 4)
enum : uint8_t { CF_T1 = 1, CF_BURN = 2, CF_T0 = 4, CF_PICK_T1 = 8, CF_KNOWN = 16, CF_PREV = 32 };
enum : uint32_t { K_CREATE = 0, K_IN = 1, K_PARK = 2, K_CARRIED = 3 };

// small device-wide numbers
struct Globals {
  unsigned int bad;
  unsigned int e_first[2];  // merge position where E[k] is first needed
  unsigned int a_first[2];  // ... T[0][k] (tid 0)
  unsigned long long n_dropped, n_we_start, n_we_end, n_completed, n_logic;
  unsigned int max_tag;
  unsigned long long max_ts;  // largest sort key (time since min_ts, + 1)
  unsigned long long min_ts;  // earliest event of the window (every unit's first one is its earliest)
  int depth0;                 // write-error depth the window starts with
  unsigned long long last_ts0;  // last timestamp of the previous window
  unsigned int have;          // bit k: E[k] exists already, bit 2 + k: T[0][k]
  unsigned int cand;          // earliest switch-in on a unit whose active state is a thread (merge position)
};

struct MaxOp {
  __device__ uint32_t operator()(uint32_t a, uint32_t b) const { return a > b ? a : b; }
};

__device__ __forceinline__ bool is_switch(uint32_t t) { return t >= EV_IN && t <= EV_OUT_YIELD; }

// (1) sort keys, input checks
__global__ void k_sl_prepare(const hbtgpu_cs_event* __restrict__ ev, uint32_t n, const uint32_t* __restrict__ unit_of,
                             const uint64_t* __restrict__ ub, uint64_t* __restrict__ key, uint32_t* __restrict__ key32,
                             uint32_t* __restrict__ val, Globals* g) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const hbtgpu_cs_event e = ev[i];
  const bool virt = is_virtual(e);
  // Sort key: time since the window's earliest event, + 1; carried states (0) sort in front.
  // Absolute kernel times have 50 bits, a window's span about 32: fewer radix passes, and
  // mostly over 4-byte keys (key32 is what is sorted when every key fits)
  const uint64_t k = virt ? 0 : e.tstamp - g->min_ts + 1;
  key[i] = k;
  key32[i] = static_cast<uint32_t>(k);
  val[i] = i;
  uint32_t bad = 0;
  const uint32_t u = unit_of[i];
  if (virt && i != ub[u]) bad |= BAD_LEVEL;  // levels 0xee / 0xef mark the carried state in slot ub[u] only
  if (e.comp_unit_id != u) bad |= BAD_UNIT;
  if (!virt) {
    if (e.type > EV_WE_END || e.type == EV_START || e.type == EV_END) bad |= BAD_TYPE;
    // (a thread lifetime event at another level is dropped with a logic error, as in the reference)
    if (e.type >= EV_IN && e.type <= EV_OUT_YIELD && e.level != 0) bad |= BAD_LEVEL;
    if ((e.type == EV_CREATE || e.type == EV_IN) && e.tag == NA) bad |= BAD_TAG;
    // slot ub[u] holds the carried state; the unit's events must not go back in time, nor behind
    // the previous window's last event
    if (i > ub[u] + 1 ? ev[i - 1].tstamp > e.tstamp : e.tstamp < g->last_ts0) bad |= BAD_ORDER;
  }
  // (threads past the end have left: aggregate over the lanes that are here)
  const unsigned m = __activemask();
  const uint32_t wbad = __reduce_or_sync(m, bad);
  const uint32_t wmax = __reduce_max_sync(m, ((e.type == EV_CREATE || e.type == EV_IN) && e.tag != NA) ? e.tag : 0u);
  uint64_t wts = k;  // (a unit's last event would do, but the check above may fail)
  for (int o = 16; o; o >>= 1) {
    const uint64_t other = __shfl_xor_sync(m, wts, o);
    if ((m >> ((threadIdx.x & 31) ^ o)) & 1u) wts = max(wts, other);
  }
  if ((threadIdx.x & 31) == __ffs(m) - 1) {
    // (plain reads as a filter: most warps have nothing new to report)
    if (wbad & ~g->bad) atomicOr(&g->bad, wbad);
    if (wmax > g->max_tag) atomicMax(&g->max_tag, wmax);
    if (wts > g->max_ts) atomicMax(&g->max_ts, static_cast<unsigned long long>(wts));
  }
}

__global__ void k_sl_unit_of(const uint64_t* __restrict__ ub, uint32_t n_units, uint32_t* __restrict__ unit_of,
                             const hbtgpu_cs_event* __restrict__ ev, Globals* g) {
  const uint32_t u = blockIdx.x;
  for (uint64_t i = ub[u] + threadIdx.x; i < ub[u + 1]; i += blockDim.x) unit_of[i] = u;
  // (parallel form only: slot ub[u] holds the carried state, the unit's first event follows)
  if (g != nullptr && threadIdx.x == 0 && ub[u] + 1 < ub[u + 1])
    atomicMin(&g->min_ts, static_cast<unsigned long long>(ev[ub[u] + 1].tstamp));
}

// (2) merge order: rank, write-error deltas
__global__ void k_sl_gather(const hbtgpu_cs_event* __restrict__ ev, const uint32_t* __restrict__ order, uint32_t n,
                            uint32_t* __restrict__ rank, int32_t* __restrict__ delta, uint32_t* __restrict__ start) {
  const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n) return;
  const uint32_t i = order[p];
  rank[i] = p;
  const uint32_t t = ev[i].type;
  delta[p] = t == EV_WE_START ? 1 : (t == EV_WE_END ? -1 : 0);
  start[p] = t == EV_WE_START;
}

// (3a) effective events in unit-major order; input of the predecessor scan
__global__ void k_sl_effective(const hbtgpu_cs_event* __restrict__ ev, uint32_t n, const uint32_t* __restrict__ rank,
                               const int32_t* __restrict__ depth_incl, const int32_t* __restrict__ delta,
                               uint8_t* __restrict__ eff, uint32_t* __restrict__ scan_in, Globals* g) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t t = ev[i].type, p = rank[i];
  const bool virt = is_virtual(ev[i]);
  const int32_t before = g->depth0 + depth_incl[p] - delta[p];  // num_active_write_errors_ as the event finds it
  const bool we_end = t == EV_WE_END, we_start = t == EV_WE_START;
  const bool dropped = !virt && !we_end && !we_start && before > 0;  // handleWriteErrors_, Slicer.h:691-701
  // handleThreadLifetime_, :711-720: a lifetime event that is not at level 0 is dropped as a logic error
  const bool logic = !virt && !dropped && (t == EV_CREATE || t == EV_DESTROY) && ev[i].level != 0;
  const bool e = virt ? is_switch(t) : (!we_end && !we_start && !dropped && !logic);
  const unsigned m = __activemask();
  const unsigned b_end = __ballot_sync(m, we_end), b_start = __ballot_sync(m, we_start), b_drop = __ballot_sync(m, dropped);
  const unsigned b_bad = __ballot_sync(m, we_end && before <= 0);  // the reference's unsigned counter would wrap
  const unsigned b_logic = __ballot_sync(m, logic);
  if ((threadIdx.x & 31) == __ffs(m) - 1) {
    if (b_end) atomicAdd(&g->n_we_end, static_cast<unsigned long long>(__popc(b_end)));
    if (b_start) atomicAdd(&g->n_we_start, static_cast<unsigned long long>(__popc(b_start)));
    if (b_drop) atomicAdd(&g->n_dropped, static_cast<unsigned long long>(__popc(b_drop)));
    if (b_bad) atomicOr(&g->bad, BAD_WE_END);
    if (b_logic) atomicAdd(&g->n_logic, static_cast<unsigned long long>(__popc(b_logic)));
  }
  eff[i] = e;
  scan_in[i] = (e && is_switch(t)) ? i + 1 : 0;
}

// (3b) predecessor / successor on the same unit, `known`, per-thread entry counts
__global__ void k_sl_links(const hbtgpu_cs_event* __restrict__ ev, uint32_t n, const uint32_t* __restrict__ unit_of,
                           const uint64_t* __restrict__ ub, const uint32_t* __restrict__ rank,
                           const uint32_t* __restrict__ epoch, const uint8_t* __restrict__ eff,
                           const uint8_t* __restrict__ rin, const uint32_t* __restrict__ prev_scan,
                           uint32_t* __restrict__ prev, uint32_t* __restrict__ next, uint8_t* __restrict__ cf,
                           uint32_t* __restrict__ own_epoch, Globals* g) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  own_epoch[i] = epoch[rank[i]];  // by event, for the per-thread walks (one look-up instead of two)
  uint32_t pv = NOIDX;
  uint8_t f = 0;
  const uint32_t t = ev[i].type;
  if (eff[i] && is_switch(t)) {
    const uint32_t raw = prev_scan[i];  // last effective switch event before i, + 1
    if (raw != 0 && raw - 1 >= ub[unit_of[i]]) {
      const uint32_t r = raw - 1;
      next[r] = i;  // (each event is the predecessor of at most one)
      // no reset in between; a switch-in that reset everything (rin) belongs to the epoch it ended
      // and activated nothing
      if (!rin[r] && epoch[rank[r]] == epoch[rank[i]] - (rin[i] ? 1u : 0u)) {
        pv = r;
        f |= CF_PREV;
        if (t == EV_IN) {
          // handleEventInPreConditionError_, :839-868: the slice of the thread is closed, then
          // everything is reset and this switch-in activates nothing.  Only the earliest one is
          // certain (later ones may disappear with that reset): the host marks it and repeats.
          if (ev[r].type == EV_IN && !rin[i]) atomicMin(&g->cand, rank[i]);
          f |= CF_KNOWN;  // the state left behind is E[1]: known_num_levels
        }
      }
    }
  }
  if (is_virtual(ev[i]) && ev[i].level == VIRT_K1) f |= ev[i].tag == 0 ? CF_KNOWN : CF_PICK_T1;
  prev[i] = pv;
  cf[i] = f;
}

__global__ void k_sl_entry_counts(const hbtgpu_cs_event* __restrict__ ev, uint32_t n, const uint32_t* __restrict__ rank,
                                  const uint32_t* __restrict__ epoch, const uint8_t* __restrict__ eff,
                                  const uint8_t* __restrict__ rin, const uint32_t* __restrict__ next,
                                  uint32_t* __restrict__ ecnt) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t c = 0;
  if (eff[i] && ev[i].tag != 0 && !rin[i]) {
    const uint32_t t = ev[i].type;
    if (t == EV_CREATE) c = 1;
    else if (t == EV_IN) {
      const uint32_t nx = next[i];
      c = 1 + ((nx != NOIDX && epoch[rank[nx]] - (rin[nx] ? 1u : 0u) == epoch[rank[i]]) ? 1u : 0u);
    }
  }
  ecnt[i] = c;
}

__global__ void k_sl_entries(const hbtgpu_cs_event* __restrict__ ev, uint32_t n, const uint32_t* __restrict__ rank,
                             const uint32_t* __restrict__ next, const uint32_t* __restrict__ ecnt,
                             const uint32_t* __restrict__ eoff, const uint8_t* __restrict__ cf, int pos_bits,
                             uint64_t* __restrict__ ekey, uint32_t* __restrict__ eval) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n || ecnt[i] == 0) return;
  const uint64_t tag = ev[i].tag;
  const uint32_t o = eoff[i];
  const uint32_t kind = ev[i].type == EV_CREATE ? K_CREATE : (is_virtual(ev[i]) ? K_CARRIED : K_IN);
  // key: thread id | merge position | `known` of a switch-in -- everything the walk needs of the
  // event travels with the entry, so that the walk reads its list and nothing else
  ekey[o] = (tag << (pos_bits + 1)) | (static_cast<uint64_t>(rank[i]) << 1) | ((cf[i] & CF_KNOWN) ? 1u : 0u);
  eval[o] = (i << 2) | kind;  // (fewer than 2^30 events per call)
  if (ecnt[i] == 2) {
    ekey[o + 1] = (tag << (pos_bits + 1)) | (static_cast<uint64_t>(rank[next[i]]) << 1);
    eval[o + 1] = (i << 2) | K_PARK;
  }
}

// (4) one thread per thread id: walk its events in merge order.  The steps of a walk depend on
// each other, so a step must not wait for memory: it reads its list (consecutive entries), finds
// the reset epoch of an entry by its merge position in the short list of reset positions, keeps
// the one switch-in it may have to come back to in registers, and sends its flags off with
// atomics that return nothing.  (Reading own_epoch[i] and cf[i] of a random event in every step
// made the longest list, times a memory latency per entry, the duration of the kernel.)
__device__ __forceinline__ void cf_or(uint8_t* cf, uint32_t i, uint32_t flag) {
  atomicOr(reinterpret_cast<unsigned int*>(cf + (i & ~3u)), flag << (8 * (i & 3u)));
}
// (the first entry of every thread id's list, for cub::DeviceSelect::If over the entry indices: the
// walkers are then neighbours in their warps -- one thread per ENTRY that returns unless it is a
// first one left one lane in fifty working)
struct IsListHead {
  const uint64_t* ekey;
  int sh;
  __device__ bool operator()(uint32_t e) const { return e == 0 || (ekey[e] >> sh) != (ekey[e - 1] >> sh); }
};
__global__ void k_sl_walk(const uint64_t* __restrict__ ekey, const uint32_t* __restrict__ eval, uint32_t n_entries, int pos_bits,
                          const uint32_t* __restrict__ heads, const uint32_t* __restrict__ n_heads,
                          const uint32_t* __restrict__ rpos /* merge positions of the resets, ascending */,
                          const uint32_t* __restrict__ n_rpos, const uint32_t* __restrict__ own_epoch, uint8_t* cf,
                          uint32_t* __restrict__ t_create /* [n_tags][2]: event that created T[v][k] */,
                          uint8_t* __restrict__ pt_flags /* [n_tags][2]: bit 0 exists, bit 1 dormant */,
                          uint32_t* __restrict__ last_epoch /* [n_tags] */) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= *n_heads) return;
  const uint32_t e0 = heads[j];
  const int sh = pos_bits + 1;
  const uint32_t v = static_cast<uint32_t>(ekey[e0] >> sh);
  bool ex1 = pt_flags[2 * v + 1] & 1, ln1 = pt_flags[2 * v + 1] & 2, ex0 = pt_flags[2 * v] & 1, ln0 = pt_flags[2 * v] & 2;
  uint32_t cur_epoch = 0;  // what was carried over is valid until the window's first reset
  const uint32_t R = *n_rpos;
  uint32_t r = 0;          // resets at merge positions <= the current entry's
  uint32_t in_i = NOIDX, in_ep = 0;  // the activation a parking entry most likely belongs to
  bool in_pick = false;
  const uint64_t pos_mask = (1ull << pos_bits) - 1;
  for (uint32_t e = e0; e < n_entries; ++e) {
    const uint64_t key = ekey[e];
    if (static_cast<uint32_t>(key >> sh) != v) break;
    const uint32_t kind = eval[e] & 3, i = eval[e] >> 2;
    uint32_t ep;  // (a parking switch-out is in its switch-in's epoch)
    if (kind == K_PARK) {
      ep = i == in_i ? in_ep : own_epoch[i];
    } else {
      const uint32_t pos = static_cast<uint32_t>((key >> 1) & pos_mask);
      if (r < R && rpos[r] <= pos) {  // gallop, then bisect: positions only move forward
        uint32_t step = 1, lo = r;
        while (lo + step < R && rpos[lo + step] <= pos) {
          lo += step;
          step <<= 1;
        }
        uint32_t hi = lo + step < R ? lo + step : R;  // rpos[lo] <= pos, rpos[hi] > pos (or hi == R)
        while (hi - lo > 1) {
          const uint32_t mid = lo + (hi - lo) / 2;
          if (rpos[mid] <= pos) lo = mid;
          else hi = mid;
        }
        r = hi;
      }
      ep = r;
    }
    if (ep != cur_epoch) {               // resetState_, :636-652: dormant lists cleared
      ln1 = ln0 = false;
      cur_epoch = ep;
    }
    if (kind == K_CREATE) {  // handleThreadLifetime_, :722-732
      if (!ex1) {
        ex1 = true;
        cf_or(cf, i, CF_T1);
        t_create[2 * v + 1] = i;
      } else {
        cf_or(cf, i, CF_BURN);  // emplace fails, next_stack_id_++ was evaluated
      }
      ln1 = true;
    } else if (kind == K_IN) {
      const bool known = key & 1;
      in_i = i;
      in_ep = ep;
      if (known && ex1 && ln1) {  // activateStackStateThatContains_, :611-626
        ln1 = false;
        in_pick = true;
        cf_or(cf, i, CF_PICK_T1);
      } else {                    // {v} with unknown levels, :632-633 / exact match :586-609
        in_pick = false;
        if (!ex0) {
          ex0 = true;
          cf_or(cf, i, CF_T0);
          t_create[2 * v] = i;
        }
        ln0 = false;
      }
    } else if (kind == K_PARK) {  // the next switch event on that unit parks the state: handleLastState_, :761-768
      // (not the latest activation: the thread runs on two units at once -- read what that one took)
      const bool pick = i == in_i ? in_pick : (*reinterpret_cast<volatile uint8_t*>(cf + i) & CF_PICK_T1) != 0;
      if (pick) ln1 = true;
      else ln0 = true;
    } else {  // K_CARRIED: the state was activated in an earlier window; it is neither created nor unlinked here
      in_i = i;
      in_ep = ep;
      in_pick = (cf[i] & CF_PICK_T1) != 0;
    }
  }
  pt_flags[2 * v + 1] = (ex1 ? 1 : 0) | (ln1 ? 2 : 0);
  pt_flags[2 * v] = (ex0 ? 1 : 0) | (ln0 ? 2 : 0);
  last_epoch[v] = cur_epoch;
}

// (5a) where E[k] and T[0][k] are first needed
__global__ void k_sl_firsts(const hbtgpu_cs_event* __restrict__ ev, uint32_t n, const uint32_t* __restrict__ rank,
                            const uint8_t* __restrict__ eff, const uint8_t* __restrict__ rin,
                            const uint8_t* __restrict__ cf, Globals* g) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t e0 = NOIDX, e1 = NOIDX, a0 = NOIDX, a1 = NOIDX;
  if (i < n && eff[i] && !is_virtual(ev[i]) && !rin[i]) {
    const uint32_t t = ev[i].type, p = rank[i];
    const uint8_t f = cf[i];
    if (t == EV_OUT_PREEMPT || t == EV_OUT_YIELD) e1 = p;
    else if (t == EV_CREATE) {
      e1 = p;  // parent of {v} known (exists already when this one only burns an id)
      if (ev[i].tag == 0) a1 = p;
    } else if (t == EV_IN) {
      if (ev[i].tag == 0) {
        if (f & CF_KNOWN) a1 = e1 = p;
        else a0 = e0 = p;
      } else if (f & CF_T0) {
        e0 = p;
      }
    }
  }
  const unsigned have = g->have;
  if (have & 1) e0 = NOIDX;
  if (have & 2) e1 = NOIDX;
  if (have & 4) a0 = NOIDX;
  if (have & 8) a1 = NOIDX;
  e0 = __reduce_min_sync(0xffffffffu, e0);
  e1 = __reduce_min_sync(0xffffffffu, e1);
  a0 = __reduce_min_sync(0xffffffffu, a0);
  a1 = __reduce_min_sync(0xffffffffu, a1);
  if ((threadIdx.x & 31) == 0) {
    if (e0 < g->e_first[0]) atomicMin(&g->e_first[0], e0);
    if (e1 < g->e_first[1]) atomicMin(&g->e_first[1], e1);
    if (a0 < g->a_first[0]) atomicMin(&g->a_first[0], a0);
    if (a1 < g->a_first[1]) atomicMin(&g->a_first[1], a1);
  }
}

// (5b) ids taken by the event at merge position p
__global__ void k_sl_id_counts(const hbtgpu_cs_event* __restrict__ ev, const uint32_t* __restrict__ order, uint32_t n,
                               const uint8_t* __restrict__ eff, const uint8_t* __restrict__ rin,
                               const uint8_t* __restrict__ cf, const Globals* g, uint32_t* __restrict__ idcnt) {
  const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n) return;
  const uint32_t i = order[p];
  uint32_t c = 0;
  if (eff[i] && !is_virtual(ev[i]) && !rin[i]) {
    const uint32_t t = ev[i].type;
    const uint8_t f = cf[i];
    if (t == EV_OUT_PREEMPT || t == EV_OUT_YIELD) c = p == g->e_first[1];
    else if (t == EV_CREATE) {
      if (ev[i].tag == 0) c = 1 + (p == g->a_first[1] && p == g->e_first[1]);  // create (with parent) or burn
      else c = ((f & (CF_T1 | CF_BURN)) ? 1u : 0u) + ((f & CF_T1) && p == g->e_first[1]);
    } else if (t == EV_IN) {
      if (ev[i].tag == 0) {
        const uint32_t k = (f & CF_KNOWN) ? 1 : 0;
        c = (p == g->a_first[k]) + (p == g->a_first[k] && p == g->e_first[k]);
      } else if (f & CF_T0) {
        c = 1 + (p == g->e_first[0]);
      }
    }
  }
  idcnt[p] = c;
}

// The id of the state event i activated (i: an effective switch event)
struct Ids {
  const uint32_t* pt_id;  // [n_tags][2]: id of T[v][k]
  uint64_t e_id[2], a_id[2];
  __device__ __forceinline__ uint64_t thread_state(uint32_t v, uint32_t k) const { return pt_id[2 * v + k]; }
};

// ids of the thread states this window created: base of the creating event (+ 1 when the same
// event created the parent E[k] first)
__global__ void k_sl_store_ids(uint32_t n_tags, const uint32_t* __restrict__ t_create, const uint32_t* __restrict__ rank,
                               const uint32_t* __restrict__ idbase, uint32_t id0, const Globals* g,
                               uint32_t* __restrict__ pt_id) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;  // 2 v + k
  if (j >= 2 * n_tags || t_create[j] == NOIDX) return;
  const uint32_t p = rank[t_create[j]];
  pt_id[j] = id0 + idbase[p] + (p == g->e_first[j & 1] ? 1u : 0u);
}

// state slots for the statistics: 0,1 = E[k]; 2,3 = T[0][k]; 4 + 2v + k = T[v][k]
__device__ __forceinline__ uint32_t state_slot(const hbtgpu_cs_event& e, uint8_t f) {
  if (e.type != EV_IN) return 1;
  if (e.tag == 0) return 2 + ((f & CF_KNOWN) ? 1 : 0);
  return 4 + 2 * e.tag + ((f & CF_PICK_T1) ? 1 : 0);
}
__device__ __forceinline__ uint64_t state_id(const Ids& ids, const hbtgpu_cs_event& e, uint8_t f) {
  if (e.type != EV_IN) return ids.e_id[1];
  if (e.tag == 0) return ids.a_id[(f & CF_KNOWN) ? 1 : 0];
  return ids.thread_state(e.tag, (f & CF_PICK_T1) ? 1 : 0);
}

// (6a) which switch events close a slice with a duration
__global__ void k_sl_slice_flags(const hbtgpu_cs_event* __restrict__ ev, uint32_t n, const uint8_t* __restrict__ eff,
                                 const uint8_t* __restrict__ rin, const uint32_t* __restrict__ prev, const uint8_t* __restrict__ cf,
                                 uint32_t* __restrict__ sflag, unsigned long long* __restrict__ st_cum,
                                 unsigned long long* __restrict__ st_in, unsigned long long* __restrict__ st_out,
                                 Globals* g) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t fl = 0;
  uint32_t in_slot = NOIDX, out_slot = NOIDX;
  uint64_t dur = 0;
  if (i < n && eff[i] && is_switch(ev[i].type) && !is_virtual(ev[i])) {
    if (!rin[i]) in_slot = state_slot(ev[i], cf[i]);  // markSwitchIn, :231-241
    const uint32_t r = prev[i];
    if (r != NOIDX) {
      dur = ev[i].tstamp - ev[r].tstamp;
      out_slot = state_slot(ev[r], cf[r]);  // markSwitchOut, :243-252
      fl = dur != 0;                        // completeSlice_ drops slices of no duration, :545-554
    }
  }
  if (i < n) sflag[i] = fl;
  // The empty stack and tid 0 take half of all updates: those four slots are summed per warp, then
  // per block in shared memory, and reach global memory once per block; the thread states (spread
  // over the table) take one atomic per event.
  __shared__ unsigned long long sh_acc[13];  // cum[4], in[4], out[4], completed
  if (threadIdx.x < 13) sh_acc[threadIdx.x] = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31;
#pragma unroll
  for (uint32_t s = 0; s < 4; ++s) {
    const unsigned b_in = __ballot_sync(0xffffffffu, in_slot == s), b_out = __ballot_sync(0xffffffffu, out_slot == s);
    if (b_out) {
      uint64_t d = out_slot == s ? dur : 0;
      for (int o = 16; o; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
      if (lane == 0) {
        if (d) atomicAdd(&sh_acc[s], static_cast<unsigned long long>(d));
        atomicAdd(&sh_acc[8 + s], static_cast<unsigned long long>(__popc(b_out)));
      }
    }
    if (b_in && lane == 0) atomicAdd(&sh_acc[4 + s], static_cast<unsigned long long>(__popc(b_in)));
  }
  if (in_slot != NOIDX && in_slot >= 4) atomicAdd(&st_in[in_slot], 1ull);
  if (out_slot != NOIDX && out_slot >= 4) {
    if (dur) atomicAdd(&st_cum[out_slot], static_cast<unsigned long long>(dur));
    atomicAdd(&st_out[out_slot], 1ull);
  }
  const unsigned b_done = __ballot_sync(0xffffffffu, out_slot != NOIDX);
  if (b_done && lane == 0) atomicAdd(&sh_acc[12], static_cast<unsigned long long>(__popc(b_done)));
  __syncthreads();
  if (threadIdx.x < 13 && sh_acc[threadIdx.x]) {
    const int j = threadIdx.x;
    unsigned long long* dst = j < 4 ? &st_cum[j] : (j < 8 ? &st_in[j - 4] : (j < 12 ? &st_out[j - 8] : &g->n_completed));
    atomicAdd(dst, sh_acc[j]);
  }
}

// (6b) the slices, unit after unit (the exclusive sum of the flags is their place)
__global__ void k_sl_emit(const hbtgpu_cs_event* __restrict__ ev, uint32_t n, const uint32_t* __restrict__ prev,
                          const uint8_t* __restrict__ cf, const uint32_t* __restrict__ sflag,
                          const uint32_t* __restrict__ soff, Ids ids, hbtgpu_slice* __restrict__ out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n || !sflag[i]) return;
  const uint32_t r = prev[i];
  const hbtgpu_cs_event a = ev[r], b = ev[i];
  hbtgpu_slice s;
  memset(&s, 0, sizeof(s));
  s.tstamp = a.tstamp;
  s.duration = b.tstamp - a.tstamp;
  s.stack_id = state_id(ids, a, cf[r]);
  s.swin_type = a.type == EV_OUT_PREEMPT ? TT_PREEMPTED : TT_YIELD;  // eventToTransitionType_, :789-805
  s.swout_type = b.type == EV_OUT_PREEMPT ? TT_PREEMPTED : TT_YIELD;
  out[soff[i]] = s;
}

__global__ void k_sl_unit_counts(const uint64_t* __restrict__ ub, uint32_t n_units, uint32_t n,
                                 const uint32_t* __restrict__ soff, const uint32_t* __restrict__ sflag,
                                 uint64_t* __restrict__ slice_begin) {
  const uint32_t u = blockIdx.x * blockDim.x + threadIdx.x;
  if (u > n_units) return;
  const uint64_t b = u < n_units ? ub[u] : n;
  slice_begin[u] = b < n ? soff[b] : (n ? soff[n - 1] + sflag[n - 1] : 0);
}

// (7) the stack table: one entry per state that exists, ordered by id (two passes around a scan)
__global__ void k_sl_stacks(uint32_t n_tags, Ids ids, const uint8_t* __restrict__ pt_flags,
                            const unsigned long long* __restrict__ st_cum,
                            const unsigned long long* __restrict__ st_in, const unsigned long long* __restrict__ st_out,
                            uint32_t* __restrict__ id_flag, const uint32_t* __restrict__ id_pos,
                            hbtgpu_stack_stats* __restrict__ out) {
  const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
  if (slot >= 4 + 2 * n_tags) return;
  const uint32_t k = slot & 1;
  uint64_t id, parent = NONE64;
  uint32_t tag = NA, nset = 0;
  if (slot < 2) {
    if (ids.e_id[k] == NONE64) return;
    id = ids.e_id[k];
  } else if (slot < 4) {
    if (ids.a_id[k] == NONE64) return;
    id = ids.a_id[k];
    parent = ids.e_id[k];
    tag = 0;
    nset = 1;
  } else {
    const uint32_t v = (slot - 4) >> 1;
    if (v == 0 || !(pt_flags[2 * v + k] & 1)) return;
    id = ids.thread_state(v, k);
    parent = ids.e_id[k];
    tag = v;
    nset = 1;
  }
  if (id_pos == nullptr) {  // first pass: which ids belong to a state (creations of existing stacks burn ids)
    id_flag[id] = 1;
    return;
  }
  hbtgpu_stack_stats s;
  memset(&s, 0, sizeof(s));
  s.stack_id = id;
  s.parent_stack_id = parent;
  for (int t = 0; t < 6; ++t) s.tags[t] = NA;
  if (nset) s.tags[0] = tag;
  s.num_set_tags = static_cast<uint8_t>(nset);
  s.known_num_levels = static_cast<uint8_t>(k);
  s.cum_active_time = st_cum[slot];
  s.num_swins = st_in[slot];
  s.num_swouts = st_out[slot];
  out[id_pos[id]] = s;  // in stack_id order
}

// what every unit is left with: the state its last effective switch event activated, unless a
// reset came after it
struct Carried {
  uint64_t tstamp;
  uint32_t tag;
  uint8_t type, level, none, pad;
};
__global__ void k_sl_final_active(const hbtgpu_cs_event* __restrict__ ev, const uint64_t* __restrict__ ub,
                                  uint32_t n_units, const uint32_t* __restrict__ rank,
                                  const uint32_t* __restrict__ epoch, const uint8_t* __restrict__ eff,
                                  const uint8_t* __restrict__ rin, const uint8_t* __restrict__ cf, uint32_t final_epoch,
                                  Carried* __restrict__ out) {
  const uint32_t u = blockIdx.x * blockDim.x + threadIdx.x;
  if (u >= n_units) return;
  Carried c{0, 0, EV_DESTROY, VIRT_K0, 1, 0};
  for (uint64_t i = ub[u + 1]; i > ub[u]; --i) {
    const hbtgpu_cs_event e = ev[i - 1];
    if (!eff[i - 1] || !is_switch(e.type)) continue;
    if (epoch[rank[i - 1]] == final_epoch && !rin[i - 1]) {
      const uint8_t f = cf[i - 1];
      c.tstamp = e.tstamp;
      c.tag = e.tag;
      c.type = e.type;
      // E[1] after a switch-out; T[v][1] / T[0][1] when the switch-in took the known-levels state
      c.level = (e.type != EV_IN || (e.tag == 0 ? (f & CF_KNOWN) : (f & CF_PICK_T1))) ? VIRT_K1 : VIRT_K0;
      c.none = 0;
    }
    break;
  }
  out[u] = c;
}
// a reset clears every dormant list, also of threads that had no event after it
__global__ void k_sl_clear_dormant(uint32_t n_tags, uint32_t final_epoch, const uint32_t* __restrict__ last_epoch,
                                   uint8_t* __restrict__ pt_flags) {
  const uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= n_tags || last_epoch[v] == final_epoch) return;
  pt_flags[2 * v] &= 1;
  pt_flags[2 * v + 1] &= 1;
}
__global__ void k_sl_put_carried(hbtgpu_cs_event* __restrict__ ev, const uint64_t* __restrict__ ub, uint32_t n_units,
                                 const Carried* __restrict__ carried) {
  const uint32_t u = blockIdx.x * blockDim.x + threadIdx.x;
  if (u >= n_units) return;
  const Carried c = carried[u];
  hbtgpu_cs_event e;
  e.tstamp = c.tstamp;
  e.type = c.none ? EV_DESTROY : c.type;
  e.level = c.none ? VIRT_K0 : c.level;
  e.comp_unit_id = static_cast<uint16_t>(u);
  e.tag = c.tag;
  ev[ub[u]] = e;
}

}  // namespace

#include "hbtgpu_slicer_general.inc"

struct hbtgpu_slicer_ctx {
  int device = 0;
  std::string err;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  float last_ms = 0.f;
  char* arena = nullptr;   // per-run scratch sized by the number of events
  size_t arena_cap = 0, arena_used = 0;
  char* arena2 = nullptr;  // per-run scratch sized by the number of thread ids / stack ids
  size_t arena2_cap = 0;
  // the Slicer's state between windows (HBTGPU_SLICER_CONTINUE)
  uint8_t* p_flags = nullptr;            // [p_tags][2]: bit 0 exists, bit 1 dormant
  uint32_t* p_id = nullptr;              // [p_tags][2]
  unsigned long long *p_cum = nullptr, *p_in = nullptr, *p_out = nullptr;  // [4 + 2 p_tags]
  uint32_t p_tags = 0;
  Carried* d_carried = nullptr;
  uint32_t carried_units = 0;
  uint64_t e_id[2] = {NONE64, NONE64}, a_id[2] = {NONE64, NONE64};
  uint64_t next_id = 0, last_ts = 0;
  int depth = 0;
  hbtgpu_slicer_stats total{};
  bool have_state = false;
  // the general path (any level, phase events): its dictionary, and whether this Slicer runs on it
  GenHost gen;
  bool general_mode = false;
  // results of the last run
  hbtgpu_slice* d_slices = nullptr;
  uint64_t* d_slice_begin = nullptr;
  std::vector<uint64_t> slice_begin;
  hbtgpu_stack_stats* d_stacks = nullptr;
  uint64_t n_stacks = 0;
  bool have = false;
};

namespace {

int sfail(hbtgpu_slicer_ctx* c, int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  if (c) c->err = buf;
  return code;
}

#define SK(call)                                                                                    \
  do {                                                                                              \
    cudaError_t e_ = (call);                                                                        \
    if (e_ != cudaSuccess)                                                                          \
      return sfail(c, -EIO, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
  } while (0)

template <class T>
T* carve(hbtgpu_slicer_ctx* c, size_t n) {
  const size_t bytes = (n * sizeof(T) + 255) & ~static_cast<size_t>(255);
  if (c->arena_used + bytes > c->arena_cap) return nullptr;
  T* p = reinterpret_cast<T*>(c->arena + c->arena_used);
  c->arena_used += bytes;
  return p;
}

inline unsigned blocks(size_t n, unsigned t = 256) { return static_cast<unsigned>((n + t - 1) / t); }

// a persistent array grown to n elements: old contents kept, the rest zero
template <class T>
int regrow(hbtgpu_slicer_ctx* c, T** p, size_t old_n, size_t new_n) {
  T* q = nullptr;
  SK(cudaMalloc(&q, std::max<size_t>(new_n, 1) * sizeof(T)));
  SK(cudaMemsetAsync(q, 0, std::max<size_t>(new_n, 1) * sizeof(T), c->stream));
  if (*p && old_n) SK(cudaMemcpyAsync(q, *p, old_n * sizeof(T), cudaMemcpyDeviceToDevice, c->stream));
  SK(cudaStreamSynchronize(c->stream));
  cudaFree(*p);
  *p = q;
  return 0;
}

int ensure_tags(hbtgpu_slicer_ctx* c, uint32_t n_tags) {
  if (n_tags <= c->p_tags) return 0;
  const uint32_t nt = std::max(n_tags, c->p_tags + c->p_tags / 2 + 16);
  if (int rc = regrow(c, &c->p_flags, 2 * static_cast<size_t>(c->p_tags), 2 * static_cast<size_t>(nt))) return rc;
  if (int rc = regrow(c, &c->p_id, 2 * static_cast<size_t>(c->p_tags), 2 * static_cast<size_t>(nt))) return rc;
  const size_t os = c->p_tags ? 4 + 2 * static_cast<size_t>(c->p_tags) : 0, ns = 4 + 2 * static_cast<size_t>(nt);
  if (int rc = regrow(c, &c->p_cum, os, ns)) return rc;
  if (int rc = regrow(c, &c->p_in, os, ns)) return rc;
  if (int rc = regrow(c, &c->p_out, os, ns)) return rc;
  c->p_tags = nt;
  return 0;
}

int reset_state(hbtgpu_slicer_ctx* c, uint32_t n_units) {
  cudaStream_t st = c->stream;
  if (c->p_tags) {
    SK(cudaMemsetAsync(c->p_flags, 0, 2 * static_cast<size_t>(c->p_tags), st));
    SK(cudaMemsetAsync(c->p_id, 0, 2 * static_cast<size_t>(c->p_tags) * 4, st));
    const size_t ns = 4 + 2 * static_cast<size_t>(c->p_tags);
    SK(cudaMemsetAsync(c->p_cum, 0, ns * 8, st));
    SK(cudaMemsetAsync(c->p_in, 0, ns * 8, st));
    SK(cudaMemsetAsync(c->p_out, 0, ns * 8, st));
  }
  if (c->carried_units != n_units) {
    cudaFree(c->d_carried);
    c->d_carried = nullptr;
    SK(cudaMalloc(&c->d_carried, n_units * sizeof(Carried)));
    c->carried_units = n_units;
  }
  std::vector<Carried> none(n_units, Carried{0, 0, static_cast<uint8_t>(EV_DESTROY), VIRT_K0, 1, 0});
  SK(cudaMemcpyAsync(c->d_carried, none.data(), n_units * sizeof(Carried), cudaMemcpyHostToDevice, st));
  SK(cudaStreamSynchronize(st));
  c->e_id[0] = c->e_id[1] = c->a_id[0] = c->a_id[1] = NONE64;
  c->next_id = 0;
  c->last_ts = 0;
  c->depth = 0;
  memset(&c->total, 0, sizeof(c->total));
  c->have_state = true;
  return 0;
}

}  // namespace

namespace {

template <class T>
int gen_grow(hbtgpu_slicer_ctx* c, T** p, size_t* cap, size_t need) {
  if (need <= *cap && *p) return 0;
  const size_t ncap = std::max(need, *cap * 2);
  T* fresh = nullptr;
  SK(cudaMalloc(reinterpret_cast<void**>(&fresh), ncap * sizeof(T)));
  if (*p) cudaFree(*p);
  *p = fresh;
  *cap = ncap;
  return 0;
}

void gen_free(GenHost& g) {
  for (void* p : {static_cast<void*>(g.h.s), static_cast<void*>(g.h.hslot), static_cast<void*>(g.h.dkey),
                  static_cast<void*>(g.h.dhead), static_cast<void*>(g.h.depoch), static_cast<void*>(g.h.active),
                  static_cast<void*>(g.h.swin_ts), static_cast<void*>(g.h.swin_type), static_cast<void*>(g.d),
                  static_cast<void*>(g.stacks)})
    if (p) cudaFree(p);
  g = GenHost{};
}

// room for `need_states` stacks: the state array grows by copy, the tables are rebuilt
int gen_reserve(hbtgpu_slicer_ctx* c, uint32_t n_units, uint64_t need_states) {
  GenHost& g = c->gen;
  cudaStream_t st = c->stream;
  if (need_states >= 0x40000000ull) return sfail(c, -EOVERFLOW, "more than 2^30 stacks");
  if (!g.d) SK(cudaMalloc(reinterpret_cast<void**>(&g.d), sizeof(GenDev)));
  if (g.h.n_units != n_units) {
    for (void* p : {static_cast<void*>(g.h.active), static_cast<void*>(g.h.swin_ts), static_cast<void*>(g.h.swin_type)})
      if (p) cudaFree(p);
    SK(cudaMalloc(reinterpret_cast<void**>(&g.h.active), n_units * sizeof(int32_t)));
    SK(cudaMalloc(reinterpret_cast<void**>(&g.h.swin_ts), n_units * sizeof(uint64_t)));
    SK(cudaMalloc(reinterpret_cast<void**>(&g.h.swin_type), n_units));
    SK(cudaMemsetAsync(g.h.active, 0xff, n_units * sizeof(int32_t), st));
    SK(cudaMemsetAsync(g.h.swin_ts, 0xff, n_units * sizeof(uint64_t), st));
    SK(cudaMemsetAsync(g.h.swin_type, 0, n_units, st));
    g.h.n_units = n_units;
  }
  if (need_states > g.h.cap) {
    uint32_t cap = 1024;
    while (cap < need_states) cap <<= 1;
    GState* ns = nullptr;
    SK(cudaMalloc(reinterpret_cast<void**>(&ns), static_cast<size_t>(cap) * sizeof(GState)));
    if (g.h.n) SK(cudaMemcpyAsync(ns, g.h.s, static_cast<size_t>(g.h.n) * sizeof(GState), cudaMemcpyDeviceToDevice, st));
    SK(cudaStreamSynchronize(st));
    for (void* p : {static_cast<void*>(g.h.s), static_cast<void*>(g.h.hslot), static_cast<void*>(g.h.dkey),
                    static_cast<void*>(g.h.dhead), static_cast<void*>(g.h.depoch)})
      if (p) cudaFree(p);
    g.h.s = ns;
    g.h.cap = cap;
    g.h.hcap = cap * 4;
    SK(cudaMalloc(reinterpret_cast<void**>(&g.h.hslot), static_cast<size_t>(g.h.hcap) * 4));
    SK(cudaMalloc(reinterpret_cast<void**>(&g.h.dkey), static_cast<size_t>(g.h.hcap) * 4));
    SK(cudaMalloc(reinterpret_cast<void**>(&g.h.dhead), static_cast<size_t>(g.h.hcap) * 4));
    SK(cudaMalloc(reinterpret_cast<void**>(&g.h.depoch), static_cast<size_t>(g.h.hcap) * 4));
    SK(cudaMemsetAsync(g.h.hslot, 0xff, static_cast<size_t>(g.h.hcap) * 4, st));
    SK(cudaMemsetAsync(g.h.depoch, 0, static_cast<size_t>(g.h.hcap) * 4, st));
    if (g.h.n) {
      k_gen_rehash<<<blocks(g.h.n), 256, 0, st>>>(g.h.s, g.h.n, g.h.hslot, g.h.hcap);
      k_gen_reheads<<<blocks(g.h.n), 256, 0, st>>>(g.h.s, g.h.n, g.h.epoch, g.h.dkey, g.h.dhead, g.h.depoch, g.h.hcap);
      SK(cudaGetLastError());
    }
  }
  return 0;
}

// an empty Slicer on the general path
int gen_reset(hbtgpu_slicer_ctx* c, uint32_t n_units) {
  GenHost& g = c->gen;
  cudaStream_t st = c->stream;
  if (g.h.n_units != n_units && g.h.active) {  // gen_reserve makes them anew
    cudaFree(g.h.active); cudaFree(g.h.swin_ts); cudaFree(g.h.swin_type);
    g.h.active = nullptr; g.h.swin_ts = nullptr; g.h.swin_type = nullptr;
    g.h.n_units = 0;
  }
  g.h.n = 0;
  g.h.next_id = 0;
  g.h.last_ts = 0;
  g.h.we_depth = 0;
  g.h.epoch = 1;
  memset(&g.h.stats, 0, sizeof(g.h.stats));
  g.h.rc = 0;
  g.h.rc_at = 0;
  if (g.h.hcap) {
    SK(cudaMemsetAsync(g.h.hslot, 0xff, static_cast<size_t>(g.h.hcap) * 4, st));
    SK(cudaMemsetAsync(g.h.depoch, 0, static_cast<size_t>(g.h.hcap) * 4, st));
  }
  if (g.h.active) {
    SK(cudaMemsetAsync(g.h.active, 0xff, n_units * sizeof(int32_t), st));
    SK(cudaMemsetAsync(g.h.swin_ts, 0xff, n_units * sizeof(uint64_t), st));
  }
  return 0;
}

// hbtgpu_slicer_run on the general path (hbtgpu_slicer_general.inc)
int run_general(hbtgpu_slicer_ctx* c, const hbtgpu_cs_event* events, const uint64_t* event_begin, uint32_t n_units,
                uint32_t flags, hbtgpu_slicer_stats* stats, uint64_t* n_slices, uint64_t* n_stacks, uint64_t* last_ts) {
  GenHost& g = c->gen;
  cudaStream_t st = c->stream;
  const uint64_t base = event_begin[0];
  const uint64_t N64 = event_begin[n_units] - base;
  const uint32_t N = static_cast<uint32_t>(N64);
  const bool fresh = !(flags & HBTGPU_SLICER_CONTINUE) || !c->have_state || !c->general_mode || g.h.n_units != n_units;
  if (fresh)
    if (int rc = gen_reset(c, n_units)) return rc;
  c->have_state = false;  // until this window is through: an error leaves no state to continue from
  c->general_mode = true;
  // most events find their stack in the dictionary: start with room for a fraction, grow when the walk asks
  if (int rc = gen_reserve(c, n_units, static_cast<uint64_t>(g.h.n) + std::min<uint64_t>(7 * N64, 4096 + N64 / 8) + 8)) return rc;

  std::vector<uint64_t> ub(n_units + 1);
  for (uint32_t u = 0; u <= n_units; ++u) ub[u] = event_begin[u] - base;
  size_t tmp_bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, static_cast<uint64_t*>(nullptr), static_cast<uint64_t*>(nullptr),
                                  static_cast<uint32_t*>(nullptr), static_cast<uint32_t*>(nullptr), static_cast<int>(N), 0, 64, st);
  tmp_bytes += 256;
  const size_t n1 = static_cast<size_t>(N) + 1;
  size_t need = 0;
  auto add = [&](size_t n, size_t sz) { need += (n * sz + 255) & ~static_cast<size_t>(255); };
  add(n1, 16); add(n_units + 1, 8); add(n1, 4);        // events, unit begins, unit_of
  add(n1, 8); add(n1, 8); add(n1, 4); add(n1, 4);      // key, key_s, val, order
  add(n1, 32); add(n1, 32);                            // slices by unit slot, slices packed
  add(n_units + 2, 8); add(n_units + 2, 8);            // slices per unit, slice_begin
  add(1, sizeof(Globals)); add(tmp_bytes, 1);
  if (c->arena_cap < need) {
    cudaFree(c->arena);
    c->arena = nullptr;
    c->arena_cap = 0;
    const size_t cap = need + need / 4 + (1 << 20);
    SK(cudaMalloc(&c->arena, cap));
    c->arena_cap = cap;
  }
  c->arena_used = 0;
  auto* d_ev = carve<hbtgpu_cs_event>(c, n1);
  auto* d_ub = carve<uint64_t>(c, n_units + 1);
  auto* d_unit_of = carve<uint32_t>(c, n1);
  auto* d_key = carve<uint64_t>(c, n1);
  auto* d_key_s = carve<uint64_t>(c, n1);
  auto* d_val = carve<uint32_t>(c, n1);
  auto* d_order = carve<uint32_t>(c, n1);
  auto* d_raw = carve<hbtgpu_slice>(c, n1);
  auto* d_slices = carve<hbtgpu_slice>(c, n1);
  auto* d_nout = carve<uint64_t>(c, n_units + 2);
  auto* d_slice_begin = carve<uint64_t>(c, n_units + 2);
  auto* d_g = carve<Globals>(c, 1);
  auto* d_tmp = carve<char>(c, tmp_bytes);
  if (!d_tmp) return sfail(c, -EIO, "internal: arena too small");

  SK(cudaMemcpyAsync(d_ub, ub.data(), (n_units + 1) * sizeof(uint64_t), cudaMemcpyHostToDevice, st));
  const cudaMemcpyKind kind = (flags & HBTGPU_SLICER_DEVICE_INPUT) ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
  if (N) SK(cudaMemcpyAsync(d_ev, events + base, static_cast<size_t>(N) * sizeof(hbtgpu_cs_event), kind, st));
  SK(cudaMemsetAsync(d_g, 0, sizeof(Globals), st));
  SK(cudaMemsetAsync(d_nout, 0, (n_units + 2) * sizeof(uint64_t), st));
  SK(cudaEventRecord(c->ev0, st));
  Globals gh;
  memset(&gh, 0, sizeof(gh));
  if (N) {
    k_sl_unit_of<<<n_units, 256, 0, st>>>(d_ub, n_units, d_unit_of, nullptr, nullptr);
    k_gen_prepare<<<blocks(N), 256, 0, st>>>(d_ev, N, d_unit_of, d_ub, g.h.last_ts, d_key, d_val, d_g);
    SK(cudaGetLastError());
    SK(cudaMemcpyAsync(&gh, d_g, sizeof(gh), cudaMemcpyDeviceToHost, st));
    SK(cudaStreamSynchronize(st));
    if (gh.bad) {
      const unsigned b = gh.bad;
      return sfail(c, (b & BAD_ORDER) ? -ENOTSUP : -EINVAL, "events the Slicer cannot take:%s%s%s",
                   (b & BAD_TYPE) ? " unknown event types;" : "",
                   (b & BAD_ORDER) ? " a unit's events go back in time;" : "",
                   (b & BAD_UNIT) ? " comp_unit_id differs from the unit the event was filed under;" : "");
    }
    int ts_bits = 1;
    while (ts_bits < 64 && (gh.max_ts >> ts_bits) != 0) ++ts_bits;
    size_t t = tmp_bytes;
    cub::DeviceRadixSort::SortPairs(d_tmp, t, d_key, d_key_s, d_val, d_order, static_cast<int>(N), 0, ts_bits, st);
  }
  for (uint32_t p0 = 0;;) {
    SK(cudaMemcpyAsync(g.d, &g.h, sizeof(GenDev), cudaMemcpyHostToDevice, st));
    k_gen_walk<<<1, 32, 0, st>>>(g.d, d_ev, d_order, p0, N, d_ub, d_raw, d_nout);
    SK(cudaGetLastError());
    SK(cudaMemcpyAsync(&g.h, g.d, sizeof(GenDev), cudaMemcpyDeviceToHost, st));
    SK(cudaStreamSynchronize(st));
    if (g.h.rc != -ENOSPC) break;
    p0 = g.h.rc_at;
    g.h.rc = 0;
    if (int rc = gen_reserve(c, n_units, 2 * static_cast<uint64_t>(g.h.cap))) return rc;
  }
  if (g.h.rc) {
    const int rc = g.h.rc;
    hbtgpu_cs_event bad_ev;
    uint32_t idx = 0;
    memset(&bad_ev, 0, sizeof(bad_ev));
    if (g.h.rc_at < N) {
      SK(cudaMemcpy(&idx, d_order + g.h.rc_at, 4, cudaMemcpyDeviceToHost));
      SK(cudaMemcpy(&bad_ev, d_ev + idx, sizeof(bad_ev), cudaMemcpyDeviceToHost));
    }
    return sfail(c, rc, "%s at event %u of the merged stream (tstamp %llu, type %u, level %u, unit %u)",
                 rc == -ENOTSUP ? "an event above the active stack's height: the reference reads tags it never set there (TagStack.h:103-118)"
                                : "the Slicer's argument checks fail (level >= 6, or a slice that would end before it began)",
                 g.h.rc_at, static_cast<unsigned long long>(bad_ev.tstamp), static_cast<unsigned>(bad_ev.type),
                 static_cast<unsigned>(bad_ev.level), static_cast<unsigned>(bad_ev.comp_unit_id));
  }
  // slices: unit after unit, densely
  std::vector<uint64_t> nout(n_units);
  SK(cudaMemcpyAsync(nout.data(), d_nout, n_units * sizeof(uint64_t), cudaMemcpyDeviceToHost, st));
  SK(cudaStreamSynchronize(st));
  c->slice_begin.assign(n_units + 1, 0);
  for (uint32_t u = 0; u < n_units; ++u) c->slice_begin[u + 1] = c->slice_begin[u] + nout[u];
  SK(cudaMemcpyAsync(d_slice_begin, c->slice_begin.data(), (n_units + 1) * sizeof(uint64_t), cudaMemcpyHostToDevice, st));
  k_gen_pack<<<n_units, 256, 0, st>>>(d_raw, d_ub, d_slice_begin, d_slices);
  if (int rc = gen_grow(c, &g.stacks, &g.cap_stacks, std::max<size_t>(g.h.n, 1))) return rc;
  if (g.h.n) k_gen_stacks<<<blocks(g.h.n), 256, 0, st>>>(g.h.s, g.h.n, g.stacks);
  SK(cudaGetLastError());
  SK(cudaEventRecord(c->ev1, st));
  SK(cudaStreamSynchronize(st));
  SK(cudaEventElapsedTime(&c->last_ms, c->ev0, c->ev1));
  if (stats) *stats = g.h.stats;
  if (last_ts) *last_ts = g.h.last_ts;
  if (n_slices)
    for (uint32_t u = 0; u < n_units; ++u) n_slices[u] = nout[u];
  c->n_stacks = g.h.n;
  c->d_stacks = g.stacks;
  if (n_stacks) *n_stacks = c->n_stacks;
  c->d_slices = d_slices;
  c->d_slice_begin = d_slice_begin;
  c->have_state = true;
  c->have = true;
  return 0;
}

}  // namespace

extern "C" {

hbtgpu_slicer_ctx* hbtgpu_slicer_create(int device) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || device < 0 || device >= n) return nullptr;  // no CPU fallback
  if (cudaSetDevice(device) != cudaSuccess) return nullptr;
  auto* c = new hbtgpu_slicer_ctx();
  c->device = device;
  if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess ||
      cudaEventCreate(&c->ev0) != cudaSuccess || cudaEventCreate(&c->ev1) != cudaSuccess) {
    hbtgpu_slicer_destroy(c);
    return nullptr;
  }
  return c;
}

void hbtgpu_slicer_destroy(hbtgpu_slicer_ctx* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  cudaFree(c->arena);
  cudaFree(c->arena2);
  cudaFree(c->p_flags); cudaFree(c->p_id); cudaFree(c->p_cum); cudaFree(c->p_in); cudaFree(c->p_out);
  cudaFree(c->d_carried);
  gen_free(c->gen);
  if (c->ev0) cudaEventDestroy(c->ev0);
  if (c->ev1) cudaEventDestroy(c->ev1);
  if (c->stream) cudaStreamDestroy(c->stream);
  delete c;
}

const char* hbtgpu_slicer_last_error(hbtgpu_slicer_ctx* c) { return c ? c->err.c_str() : "no context"; }
float hbtgpu_slicer_last_ms(const hbtgpu_slicer_ctx* c) { return c ? c->last_ms : 0.f; }

int hbtgpu_slicer_run(hbtgpu_slicer_ctx* c, const hbtgpu_cs_event* events, const uint64_t* event_begin,
                      uint32_t n_units, uint32_t flags, hbtgpu_slicer_stats* stats, uint64_t* n_slices,
                      uint64_t* n_stacks, uint64_t* last_ts) {
  if (!c) return -EINVAL;
  c->have = false;
  if (!event_begin || n_units == 0 || n_units > 65535) return sfail(c, -EINVAL, "null argument or bad unit count");
  SK(cudaSetDevice(c->device));
  const uint64_t base = event_begin[0];
  for (uint32_t u = 0; u < n_units; ++u)
    if (event_begin[u + 1] < event_begin[u]) return sfail(c, -EINVAL, "event_begin must not decrease");
  const uint64_t Nreal = event_begin[n_units] - base;
  if (Nreal + n_units >= 0x3fffffffull) return sfail(c, -EINVAL, "more than 2^30 events in one call");
  if (Nreal && !events) return sfail(c, -EINVAL, "null events");
  const bool continues = (flags & HBTGPU_SLICER_CONTINUE) && c->have_state;
  if ((flags & HBTGPU_SLICER_GENERAL) || (continues && c->general_mode)) {
    if (continues && !c->general_mode)
      return sfail(c, -ENOTSUP, "this Slicer runs the parallel form: HBTGPU_SLICER_GENERAL must be given from its first window on");
    return run_general(c, events, event_begin, n_units, flags, stats, n_slices, n_stacks, last_ts);
  }
  if (!continues || c->carried_units != n_units)
    if (int rc = reset_state(c, n_units)) return rc;
  c->general_mode = false;
  // every unit gets one slot in front of its events for the state the previous window left it in
  const uint32_t N = static_cast<uint32_t>(Nreal + n_units);
  std::vector<uint64_t> ub(n_units + 1);
  for (uint32_t u = 0; u <= n_units; ++u) ub[u] = event_begin[u] - base + u;

  cudaStream_t st = c->stream;
  size_t tmp_sort = 0, tmp_sort2 = 0, tmp_scan = 0, t2 = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, tmp_sort, static_cast<uint64_t*>(nullptr), static_cast<uint64_t*>(nullptr),
                                  static_cast<uint32_t*>(nullptr), static_cast<uint32_t*>(nullptr), static_cast<int>(N), 0, 64, st);
  cub::DeviceRadixSort::SortPairs(nullptr, tmp_sort2, static_cast<uint64_t*>(nullptr), static_cast<uint64_t*>(nullptr),
                                  static_cast<uint32_t*>(nullptr), static_cast<uint32_t*>(nullptr), static_cast<int>(2 * static_cast<uint64_t>(N)), 0, 64, st);
  cub::DeviceRadixSort::SortPairs(nullptr, t2, static_cast<uint32_t*>(nullptr), static_cast<uint32_t*>(nullptr),
                                  static_cast<uint32_t*>(nullptr), static_cast<uint32_t*>(nullptr), static_cast<int>(N), 0, 32, st);
  tmp_sort = std::max(tmp_sort, t2);
  cub::DeviceScan::InclusiveSum(nullptr, tmp_scan, static_cast<int32_t*>(nullptr), static_cast<int32_t*>(nullptr), static_cast<int>(N), st);
  cub::DeviceScan::ExclusiveScan(nullptr, t2, static_cast<uint32_t*>(nullptr), static_cast<uint32_t*>(nullptr), MaxOp(), 0u, static_cast<int>(N), st);
  tmp_scan = std::max(tmp_scan, t2);
  cub::DeviceScan::ExclusiveSum(nullptr, t2, static_cast<uint32_t*>(nullptr), static_cast<uint32_t*>(nullptr),
                                static_cast<int>(c->next_id + N + 9), st);  // events, or all stack ids so far
  tmp_scan = std::max(tmp_scan, t2);
  cub::DeviceSelect::Flagged(nullptr, t2, cub::CountingInputIterator<uint32_t>(0), static_cast<uint32_t*>(nullptr),
                             static_cast<uint32_t*>(nullptr), static_cast<uint32_t*>(nullptr), static_cast<int>(N), st);
  tmp_scan = std::max(tmp_scan, t2);
  cub::DeviceSelect::If(nullptr, t2, cub::CountingInputIterator<uint32_t>(0), static_cast<uint32_t*>(nullptr),
                        static_cast<uint32_t*>(nullptr), static_cast<int>(2 * static_cast<uint64_t>(N)), IsListHead{nullptr, 0}, st);
  tmp_scan = std::max(tmp_scan, t2);
  const size_t tmp_bytes = std::max(std::max(tmp_sort, tmp_sort2), tmp_scan) + 256;
  const size_t n1 = static_cast<size_t>(N) + 1;
  size_t need = 0;
  auto add = [&](size_t n, size_t sz) { need += (n * sz + 255) & ~static_cast<size_t>(255); };
  add(n1, 16); add(n_units + 1, 8); add(n1, 4);                 // events, unit begins, unit_of
  add(n1, 8); add(n1, 8); add(n1, 4); add(n1, 4);               // key, key_s, val, order
  add(n1, 4); add(n1, 4); add(n1, 4); add(n1, 4); add(n1, 4);   // rank, delta, depth, start, epoch
  add(n1, 1); add(n1, 4); add(n1, 4); add(n1, 4); add(n1, 4);   // eff, scan_in, prev_scan, prev, next
  add(n1, 1); add(n1, 4); add(n1, 4); add(n1, 1); add(n1, 4);   // cf, ecnt, eoff, rin, own_epoch
  add(2 * n1, 8); add(2 * n1, 4); add(2 * n1, 8); add(2 * n1, 4);  // entries + sorted
  add(n1, 4); add(n1, 4); add(n1, 4); add(n1, 4);               // idcnt, idbase, sflag, soff
  add(n1, 32); add(n_units + 2, 8);                             // slices, slice_begin
  add(n1, 4); add(64, 4); add(n1, 4);                           // reset positions, their number (and the lists'), list heads
  add(n1, 4); add(n1, 4);                                       // 4-byte sort keys
  add(1, sizeof(Globals)); add(tmp_bytes, 1);
  if (c->arena_cap < need) {
    cudaFree(c->arena);
    c->arena = nullptr;
    c->arena_cap = 0;
    const size_t cap = need + need / 4 + (1 << 20);
    SK(cudaMalloc(&c->arena, cap));
    c->arena_cap = cap;
  }
  c->arena_used = 0;
  auto* d_ev = carve<hbtgpu_cs_event>(c, n1);
  auto* d_ub = carve<uint64_t>(c, n_units + 1);
  auto* d_unit_of = carve<uint32_t>(c, n1);
  auto* d_key = carve<uint64_t>(c, n1);
  auto* d_key_s = carve<uint64_t>(c, n1);
  auto* d_val = carve<uint32_t>(c, n1);
  auto* d_order = carve<uint32_t>(c, n1);
  auto* d_rank = carve<uint32_t>(c, n1);
  auto* d_delta = carve<int32_t>(c, n1);
  auto* d_depth = carve<int32_t>(c, n1);
  auto* d_start = carve<uint32_t>(c, n1);
  auto* d_epoch = carve<uint32_t>(c, n1);
  auto* d_eff = carve<uint8_t>(c, n1);
  auto* d_scan_in = carve<uint32_t>(c, n1);
  auto* d_prev_scan = carve<uint32_t>(c, n1);
  auto* d_prev = carve<uint32_t>(c, n1);
  auto* d_next = carve<uint32_t>(c, n1);
  auto* d_cf = carve<uint8_t>(c, n1);
  auto* d_ecnt = carve<uint32_t>(c, n1);
  auto* d_eoff = carve<uint32_t>(c, n1);
  auto* d_rin = carve<uint8_t>(c, n1);
  auto* d_own_epoch = carve<uint32_t>(c, n1);
  auto* d_ekey = carve<uint64_t>(c, 2 * n1);
  auto* d_eval = carve<uint32_t>(c, 2 * n1);
  auto* d_ekey_s = carve<uint64_t>(c, 2 * n1);
  auto* d_eval_s = carve<uint32_t>(c, 2 * n1);
  auto* d_idcnt = carve<uint32_t>(c, n1);
  auto* d_idbase = carve<uint32_t>(c, n1);
  auto* d_sflag = carve<uint32_t>(c, n1);
  auto* d_soff = carve<uint32_t>(c, n1);
  auto* d_slices = carve<hbtgpu_slice>(c, n1);
  auto* d_slice_begin = carve<uint64_t>(c, n_units + 2);
  auto* d_rpos = carve<uint32_t>(c, n1);
  auto* d_nrpos = carve<uint32_t>(c, 64);
  auto* d_heads = carve<uint32_t>(c, n1);
  auto* d_key32 = carve<uint32_t>(c, n1);
  auto* d_key32_s = carve<uint32_t>(c, n1);
  auto* d_g = carve<Globals>(c, 1);
  auto* d_tmp = carve<char>(c, tmp_bytes);
  if (!d_tmp) return sfail(c, -EIO, "internal: arena too small");

  // stage: [carried state, the unit's events] for every unit
  SK(cudaMemcpyAsync(d_ub, ub.data(), (n_units + 1) * sizeof(uint64_t), cudaMemcpyHostToDevice, st));
  k_sl_put_carried<<<blocks(n_units), 256, 0, st>>>(d_ev, d_ub, n_units, c->d_carried);
  const cudaMemcpyKind kind = (flags & HBTGPU_SLICER_DEVICE_INPUT) ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
  for (uint32_t u = 0; u < n_units; ++u) {
    const uint64_t n = event_begin[u + 1] - event_begin[u];
    if (n) SK(cudaMemcpyAsync(d_ev + ub[u] + 1, events + event_begin[u], n * sizeof(hbtgpu_cs_event), kind, st));
  }
  const hbtgpu_cs_event* ev = d_ev;
  Globals g0;
  memset(&g0, 0, sizeof(g0));
  g0.e_first[0] = g0.e_first[1] = g0.a_first[0] = g0.a_first[1] = NOIDX;
  g0.cand = NOIDX;
  g0.min_ts = ~0ull;
  g0.depth0 = c->depth;
  g0.last_ts0 = c->last_ts;
  g0.have = (c->e_id[0] != NONE64 ? 1u : 0u) | (c->e_id[1] != NONE64 ? 2u : 0u) | (c->a_id[0] != NONE64 ? 4u : 0u) |
            (c->a_id[1] != NONE64 ? 8u : 0u);
  SK(cudaMemcpyAsync(d_g, &g0, sizeof(g0), cudaMemcpyHostToDevice, st));
  SK(cudaEventRecord(c->ev0, st));

  Globals gh = g0;
  size_t t = tmp_bytes;
  k_sl_unit_of<<<n_units, 256, 0, st>>>(d_ub, n_units, d_unit_of, ev, d_g);
  k_sl_prepare<<<blocks(N), 256, 0, st>>>(ev, N, d_unit_of, d_ub, d_key, d_key32, d_val, d_g);
  SK(cudaMemcpyAsync(&gh, d_g, sizeof(gh), cudaMemcpyDeviceToHost, st));
  SK(cudaStreamSynchronize(st));
  int ts_bits = 1;  // radix passes only over the bits the keys use
  while (ts_bits < 64 && (gh.max_ts >> ts_bits) != 0) ++ts_bits;
  if (ts_bits <= 32)
    cub::DeviceRadixSort::SortPairs(d_tmp, t, d_key32, d_key32_s, d_val, d_order, static_cast<int>(N), 0, ts_bits, st);
  else
    cub::DeviceRadixSort::SortPairs(d_tmp, t, d_key, d_key_s, d_val, d_order, static_cast<int>(N), 0, ts_bits, st);
  k_sl_gather<<<blocks(N), 256, 0, st>>>(ev, d_order, N, d_rank, d_delta, d_start);
  t = tmp_bytes;
  cub::DeviceScan::InclusiveSum(d_tmp, t, d_delta, d_depth, static_cast<int>(N), st);
  t = tmp_bytes;
  cub::DeviceScan::InclusiveSum(d_tmp, t, d_start, d_epoch, static_cast<int>(N), st);
  k_sl_effective<<<blocks(N), 256, 0, st>>>(ev, N, d_rank, d_depth, d_delta, d_eff, d_scan_in, d_g);
  t = tmp_bytes;
  cub::DeviceScan::ExclusiveScan(d_tmp, t, d_scan_in, d_prev_scan, MaxOp(), 0u, static_cast<int>(N), st);
  SK(cudaMemsetAsync(d_rin, 0, static_cast<size_t>(N), st));
  // Switch-ins on a unit whose active state is a thread reset everything (Slicer.h:839-868).  Whether
  // one is such depends on the resets before it, so they are settled in merge order: link, take the
  // earliest candidate, make it a reset, link again -- until none is left (rarely more than a few).
  for (int round = 0;; ++round) {
    SK(cudaMemsetAsync(d_next, 0xff, static_cast<size_t>(N) * sizeof(uint32_t), st));
    k_sl_links<<<blocks(N), 256, 0, st>>>(ev, N, d_unit_of, d_ub, d_rank, d_epoch, d_eff, d_rin, d_prev_scan, d_prev,
                                          d_next, d_cf, d_own_epoch, d_g);
    SK(cudaGetLastError());
    SK(cudaMemcpyAsync(&gh, d_g, sizeof(gh), cudaMemcpyDeviceToHost, st));
    SK(cudaStreamSynchronize(st));
    if (gh.bad || gh.cand == NOIDX) break;
    if (round >= 4096) return sfail(c, -ENOTSUP, "more than 4096 switch-ins on units whose active state is a thread in one window");
    uint32_t idx = 0;
    const uint32_t one = 1, none = NOIDX;
    SK(cudaMemcpyAsync(&idx, d_order + gh.cand, 4, cudaMemcpyDeviceToHost, st));
    SK(cudaStreamSynchronize(st));
    SK(cudaMemsetAsync(d_rin + idx, 1, 1, st));
    SK(cudaMemcpyAsync(d_start + gh.cand, &one, 4, cudaMemcpyHostToDevice, st));
    SK(cudaMemcpyAsync(&d_g->cand, &none, 4, cudaMemcpyHostToDevice, st));
    t = tmp_bytes;
    cub::DeviceScan::InclusiveSum(d_tmp, t, d_start, d_epoch, static_cast<int>(N), st);
  }
  if (gh.bad) {
    const unsigned b = gh.bad;
    // phases, or switches above level 0: the general path has an answer, if this Slicer may still change over
    constexpr unsigned general_has_it = BAD_TYPE | BAD_LEVEL | BAD_WE_END;
    if (!(b & ~general_has_it)) {
      if (!continues) return run_general(c, events, event_begin, n_units, flags, stats, n_slices, n_stacks, last_ts);
      c->have_state = true;  // (nothing of the window was applied)
      return sfail(c, -ENOTSUP, "phase events, events above level 0 or a WriteErrorsEnd without its Start in a later "
                                "window of a Slicer that runs the parallel form: give HBTGPU_SLICER_GENERAL from the "
                                "first window on");
    }
    return sfail(c, (b & BAD_UNIT) ? -EINVAL : -ENOTSUP, "events outside the parallel form:%s%s%s%s%s%s%s",
                 (b & BAD_TYPE) ? " phase or unknown event types;" : "", (b & BAD_LEVEL) ? " events above level 0;" : "",
                 (b & BAD_ORDER) ? " a unit's events go back in time;" : "",
                 (b & BAD_UNIT) ? " comp_unit_id differs from the unit the event was filed under;" : "",
                 (b & BAD_TAG) ? " switch-in or creation without a tag;" : "",
                 (b & BAD_DOUBLE_IN) ? " (internal);" : "",
                 (b & BAD_WE_END) ? " WriteErrorsEnd without its Start;" : "");
  }
  if (gh.max_tag >= (1u << 27)) return sfail(c, -ENOTSUP, "thread ids up to %u: the per-thread tables would not fit", gh.max_tag);
  if (int rc = ensure_tags(c, gh.max_tag + 1)) return rc;
  const uint32_t n_tags = c->p_tags;
  const size_t n_slots = 4 + 2 * static_cast<size_t>(n_tags);

  // per-run scratch by thread id / stack id
  const uint64_t ids_bound = c->next_id + N + 8;
  size_t need2 = 0;
  auto add2 = [&](size_t n, size_t sz) { need2 += (n * sz + 255) & ~static_cast<size_t>(255); };
  add2(2 * static_cast<size_t>(n_tags), 4); add2(n_tags, 4); add2(ids_bound, 4); add2(ids_bound, 4);
  add2(n_slots, sizeof(hbtgpu_stack_stats));
  if (need2 > c->arena2_cap) {
    cudaFree(c->arena2);
    c->arena2 = nullptr;
    c->arena2_cap = 0;
    SK(cudaMalloc(&c->arena2, need2 + need2 / 4 + 4096));
    c->arena2_cap = need2 + need2 / 4 + 4096;
  }
  char* a2 = c->arena2;
  auto carve2 = [&](size_t n, size_t sz) {
    char* q = a2;
    a2 += (n * sz + 255) & ~static_cast<size_t>(255);
    return q;
  };
  auto* d_tcreate = reinterpret_cast<uint32_t*>(carve2(2 * static_cast<size_t>(n_tags), 4));
  auto* d_last_epoch = reinterpret_cast<uint32_t*>(carve2(n_tags, 4));
  auto* d_idflag = reinterpret_cast<uint32_t*>(carve2(ids_bound, 4));
  auto* d_idpos = reinterpret_cast<uint32_t*>(carve2(ids_bound, 4));
  auto* d_stacks = reinterpret_cast<hbtgpu_stack_stats*>(carve2(n_slots, sizeof(hbtgpu_stack_stats)));
  SK(cudaMemsetAsync(d_tcreate, 0xff, 2 * static_cast<size_t>(n_tags) * sizeof(uint32_t), st));
  SK(cudaMemsetAsync(d_last_epoch, 0, static_cast<size_t>(n_tags) * sizeof(uint32_t), st));

  k_sl_entry_counts<<<blocks(N), 256, 0, st>>>(ev, N, d_rank, d_epoch, d_eff, d_rin, d_next, d_ecnt);
  t = tmp_bytes;
  cub::DeviceScan::ExclusiveSum(d_tmp, t, d_ecnt, d_eoff, static_cast<int>(N), st);
  uint32_t last_off = 0, last_cnt = 0;
  SK(cudaMemcpyAsync(&last_off, d_eoff + (N - 1), 4, cudaMemcpyDeviceToHost, st));
  SK(cudaMemcpyAsync(&last_cnt, d_ecnt + (N - 1), 4, cudaMemcpyDeviceToHost, st));
  SK(cudaStreamSynchronize(st));
  const uint32_t n_entries = last_off + last_cnt;
  if (n_entries) {
    int pos_bits = 1, tag_bits = 1;  // key = tag << pos_bits | merge position
    while ((static_cast<uint64_t>(N) >> pos_bits) != 0) ++pos_bits;
    while ((static_cast<uint64_t>(gh.max_tag) >> tag_bits) != 0) ++tag_bits;
    k_sl_entries<<<blocks(N), 256, 0, st>>>(ev, N, d_rank, d_next, d_ecnt, d_eoff, d_cf, pos_bits, d_ekey, d_eval);
    t = tmp_bytes;
    cub::DeviceRadixSort::SortPairs(d_tmp, t, d_ekey, d_ekey_s, d_eval, d_eval_s, static_cast<int>(n_entries), 0,
                                    pos_bits + 1 + tag_bits, st);
    // the merge positions where a reset happens (few): the walks find an entry's epoch there
    t = tmp_bytes;
    cub::DeviceSelect::Flagged(d_tmp, t, cub::CountingInputIterator<uint32_t>(0), d_start, d_rpos, d_nrpos, static_cast<int>(N), st);
    t = tmp_bytes;
    cub::DeviceSelect::If(d_tmp, t, cub::CountingInputIterator<uint32_t>(0), d_heads, d_nrpos + 1, static_cast<int>(n_entries),
                          IsListHead{d_ekey_s, pos_bits + 1}, st);
    const uint32_t max_lists = static_cast<uint32_t>(std::min<uint64_t>(n_entries, static_cast<uint64_t>(gh.max_tag) + 1));
    k_sl_walk<<<blocks(max_lists, 128), 128, 0, st>>>(d_ekey_s, d_eval_s, n_entries, pos_bits, d_heads, d_nrpos + 1, d_rpos,
                                                      d_nrpos, d_own_epoch, d_cf, d_tcreate, c->p_flags, d_last_epoch);
  }
  k_sl_firsts<<<blocks(N), 256, 0, st>>>(ev, N, d_rank, d_eff, d_rin, d_cf, d_g);
  k_sl_id_counts<<<blocks(N), 256, 0, st>>>(ev, d_order, N, d_eff, d_rin, d_cf, d_g, d_idcnt);
  t = tmp_bytes;
  cub::DeviceScan::ExclusiveSum(d_tmp, t, d_idcnt, d_idbase, static_cast<int>(N), st);
  SK(cudaGetLastError());
  SK(cudaMemcpyAsync(&gh, d_g, sizeof(gh), cudaMemcpyDeviceToHost, st));
  // ids of the shared states created in this window (E[k] is the parent: first when one event makes both),
  // the number of ids the window took, where it ends
  uint32_t pos_list[4] = {0, 0, 0, 0}, base_at[4] = {0, 0, 0, 0}, last_base = 0, last_idcnt = 0, final_epoch = 0, last_idx = 0;
  int32_t depth_last = 0;
  SK(cudaStreamSynchronize(st));
  pos_list[0] = gh.e_first[0]; pos_list[1] = gh.e_first[1]; pos_list[2] = gh.a_first[0]; pos_list[3] = gh.a_first[1];
  for (int j = 0; j < 4; ++j)
    if (pos_list[j] != NOIDX) SK(cudaMemcpyAsync(&base_at[j], d_idbase + pos_list[j], 4, cudaMemcpyDeviceToHost, st));
  SK(cudaMemcpyAsync(&last_base, d_idbase + (N - 1), 4, cudaMemcpyDeviceToHost, st));
  SK(cudaMemcpyAsync(&last_idcnt, d_idcnt + (N - 1), 4, cudaMemcpyDeviceToHost, st));
  SK(cudaMemcpyAsync(&final_epoch, d_epoch + (N - 1), 4, cudaMemcpyDeviceToHost, st));
  SK(cudaMemcpyAsync(&depth_last, d_depth + (N - 1), 4, cudaMemcpyDeviceToHost, st));
  SK(cudaMemcpyAsync(&last_idx, d_order + (N - 1), 4, cudaMemcpyDeviceToHost, st));
  SK(cudaStreamSynchronize(st));
  const uint32_t id0 = static_cast<uint32_t>(c->next_id);
  for (int k = 0; k < 2; ++k) {
    if (gh.e_first[k] != NOIDX) c->e_id[k] = static_cast<uint64_t>(id0) + base_at[k];
    if (gh.a_first[k] != NOIDX)
      c->a_id[k] = static_cast<uint64_t>(id0) + base_at[2 + k] + (gh.a_first[k] == gh.e_first[k] ? 1 : 0);
  }
  c->next_id += static_cast<uint64_t>(last_base) + last_idcnt;
  if (c->next_id >= 0xfffffff0ull) return sfail(c, -EOVERFLOW, "more than 2^32 stack ids");
  k_sl_store_ids<<<blocks(2 * static_cast<size_t>(n_tags)), 256, 0, st>>>(n_tags, d_tcreate, d_rank, d_idbase, id0, d_g,
                                                                          c->p_id);
  Ids ids;
  ids.pt_id = c->p_id;
  for (int k = 0; k < 2; ++k) {
    ids.e_id[k] = c->e_id[k];
    ids.a_id[k] = c->a_id[k];
  }
  k_sl_slice_flags<<<blocks(N), 256, 0, st>>>(ev, N, d_eff, d_rin, d_prev, d_cf, d_sflag, c->p_cum, c->p_in, c->p_out, d_g);
  t = tmp_bytes;
  cub::DeviceScan::ExclusiveSum(d_tmp, t, d_sflag, d_soff, static_cast<int>(N), st);
  k_sl_emit<<<blocks(N), 256, 0, st>>>(ev, N, d_prev, d_cf, d_sflag, d_soff, ids, d_slices);
  k_sl_unit_counts<<<blocks(n_units + 1), 256, 0, st>>>(d_ub, n_units, N, d_soff, d_sflag, d_slice_begin);
  // what the next window starts from
  k_sl_final_active<<<blocks(n_units), 256, 0, st>>>(ev, d_ub, n_units, d_rank, d_epoch, d_eff, d_rin, d_cf,
                                                     final_epoch, c->d_carried);
  if (final_epoch > 0) k_sl_clear_dormant<<<blocks(n_tags), 256, 0, st>>>(n_tags, final_epoch, d_last_epoch, c->p_flags);
  // the stack table, by id: flag the ids that belong to a state, scan, place
  const uint32_t n_ids = static_cast<uint32_t>(c->next_id);
  SK(cudaMemsetAsync(d_idflag, 0, (static_cast<size_t>(n_ids) + 1) * 4, st));
  k_sl_stacks<<<blocks(n_slots), 256, 0, st>>>(n_tags, ids, c->p_flags, c->p_cum, c->p_in, c->p_out, d_idflag, nullptr,
                                               d_stacks);
  t = tmp_bytes;
  size_t t_ids = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, t_ids, d_idflag, d_idpos, static_cast<int>(n_ids + 1), st);
  if (t_ids > tmp_bytes) return sfail(c, -EIO, "internal: scan scratch too small");
  cub::DeviceScan::ExclusiveSum(d_tmp, t, d_idflag, d_idpos, static_cast<int>(n_ids + 1), st);
  k_sl_stacks<<<blocks(n_slots), 256, 0, st>>>(n_tags, ids, c->p_flags, c->p_cum, c->p_in, c->p_out, d_idflag, d_idpos,
                                               d_stacks);
  SK(cudaGetLastError());
  SK(cudaEventRecord(c->ev1, st));
  uint32_t n_states = 0;
  SK(cudaMemcpyAsync(&n_states, d_idpos + n_ids, 4, cudaMemcpyDeviceToHost, st));
  SK(cudaMemcpyAsync(&gh, d_g, sizeof(gh), cudaMemcpyDeviceToHost, st));
  c->slice_begin.assign(n_units + 1, 0);
  SK(cudaMemcpyAsync(c->slice_begin.data(), d_slice_begin, (n_units + 1) * sizeof(uint64_t), cudaMemcpyDeviceToHost, st));
  hbtgpu_cs_event last_ev;
  SK(cudaMemcpyAsync(&last_ev, ev + last_idx, sizeof(last_ev), cudaMemcpyDeviceToHost, st));
  SK(cudaStreamSynchronize(st));
  SK(cudaEventElapsedTime(&c->last_ms, c->ev0, c->ev1));
  if (Nreal) c->last_ts = std::max<uint64_t>(c->last_ts, last_ev.tstamp);  // (the last in merge order is a real event)
  c->depth += depth_last;
  c->total.num_processed += Nreal;  // every event is processed once (none is older than its predecessor)
  c->total.num_error_start += gh.n_we_start;
  c->total.num_error_end += gh.n_we_end;
  c->total.num_dropped += gh.n_dropped + gh.n_logic;
  c->total.num_logic_errors += gh.n_logic;
  c->total.num_slices_completed += gh.n_completed;
  if (stats) *stats = c->total;
  if (last_ts) *last_ts = c->last_ts;
  if (n_slices)
    for (uint32_t u = 0; u < n_units; ++u) n_slices[u] = c->slice_begin[u + 1] - c->slice_begin[u];
  c->n_stacks = n_states;
  c->d_stacks = d_stacks;
  if (n_stacks) *n_stacks = c->n_stacks;
  c->d_slices = d_slices;
  c->d_slice_begin = d_slice_begin;
  c->have = true;
  return 0;
}

int hbtgpu_slicer_get_slices(hbtgpu_slicer_ctx* c, uint32_t unit, hbtgpu_slice* out, uint64_t cap) {
  if (!c || !out) return -EINVAL;
  if (!c->have) return sfail(c, -EINVAL, "no results: call hbtgpu_slicer_run first");
  if (unit + 1 >= c->slice_begin.size()) return sfail(c, -EINVAL, "unit %u out of range", unit);
  const uint64_t n = c->slice_begin[unit + 1] - c->slice_begin[unit];
  if (cap < n) return sfail(c, -ENOSPC, "%llu slices, room for %llu", (unsigned long long)n, (unsigned long long)cap);
  SK(cudaSetDevice(c->device));
  if (n) SK(cudaMemcpy(out, c->d_slices + c->slice_begin[unit], n * sizeof(hbtgpu_slice), cudaMemcpyDeviceToHost));
  return 0;
}

int hbtgpu_slicer_get_stacks(hbtgpu_slicer_ctx* c, hbtgpu_stack_stats* out, uint64_t cap) {
  if (!c || !out) return -EINVAL;
  if (!c->have) return sfail(c, -EINVAL, "no results: call hbtgpu_slicer_run first");
  if (cap < c->n_stacks) return sfail(c, -ENOSPC, "%llu stacks, room for %llu", (unsigned long long)c->n_stacks, (unsigned long long)cap);
  SK(cudaSetDevice(c->device));
  if (c->n_stacks) SK(cudaMemcpy(out, c->d_stacks, c->n_stacks * sizeof(hbtgpu_stack_stats), cudaMemcpyDeviceToHost));
  return 0;
}

int hbtgpu_slicer_slices_device(hbtgpu_slicer_ctx* c, const hbtgpu_slice** slices, const uint64_t** slice_begin) {
  if (!c || !slices || !slice_begin) return -EINVAL;
  if (!c->have) return sfail(c, -EINVAL, "no results: call hbtgpu_slicer_run first");
  *slices = c->d_slices;
  *slice_begin = c->d_slice_begin;
  return 0;
}

}  // extern "C"
