// hbtgpu_islice.cu -- count intervals x tagstack slices on the GPU.
//
// Reference: IntervalSlicer::makeIntervalIntersectingSlices (hbt/src/tagstack/IntervalSlicer.cpp:
// 52-188) called row by row from TraceMonitor::releaseCountData_ (hbt/src/mon/TraceMonitor.cpp:
// 52-96) with a consumer mark that advances by what each row consumed.  It is a serial walk:
// skip slices that end before the interval, take slices that start before its end, trim the
// first and the last, and the next row starts where this one stopped.
//
// Parallel form.  Let t_j, e_j = t_j + d_j be start and end of slice j.  When a unit's slices
// are time ordered and disjoint (d_j > 0, t_j >= e_{j-1}: what tagstack::Slicer writes) and its
// rows move forward in time (next tstamp >= this end: what CpuCountSampleGenerator writes),
// e and t are strictly increasing, so for the interval [T, Tend):
//   * the walk's skip loop stops at p = max(mark, lb) with lb = #{j : e_j < T}  (binary search);
//   * the take loop stops at r = max(p, #{j : t_j < Tend})   (trim)
//                      or at r = max(p, #{j : e_j <= Tend})  (no trim);
//   * "hole in the cover" is G[hi] - G[p] != 0, G = inclusive prefix sum of the gap lengths
//     t_j - e_{j-1}; the same prefix gives the taken slices' total duration for the
//     reference's consistency assertion (:175-183);
//   * the mark after a row that produced slices is r, or r - 1 when the last slice runs past
//     Tend; a row that produced none leaves it alone.
// The mark can exceed lb only by one, and only when a slice ends exactly at T and the previous
// row consumed it (then the row starts at lb + 1; otherwise that slice is taken with length 0,
// which the reference does too).  So a row depends on its predecessor only if a slice ends
// exactly on its start, and a chain of such rows ends at the first row without one: a row in
// that situation re-evaluates the (at most LOOKBACK) rows of its chain; every other row is
// independent.  That makes the work a grid over (unit, tile of NT rows):
//   k_islice_validate  (unit, tile of VT slices): order / overlap / zero length, tile-local G
//   k_islice_vscan     per unit: exclusive scan of the tiles' gap sums (G = local + base)
//   k_islice_bounds    one thread per row tile: the slice window the tile's rows can touch
//   k_islice_rows      (unit, tile of NT rows): stages the tile's slice window {t, e, G} in
//                      shared memory (one coalesced pass; windows beyond WMULT * NT slices are
//                      searched in global memory), resolves every row: first slice, count,
//                      new mark, throws; tile totals;
//   k_islice_rscan     per unit: exclusive scan of the tile totals, the unit's summary; units
//                      that broke a precondition anywhere are walked here by one thread exactly
//                      as the reference does (exact_walk), so results are the reference's on
//                      ANY input
//                      and copies + trims the tile's slices into the unit's output region (one
//                      slice per thread and step, the owning row found by binary search over
//                      the tile's offsets in shared memory).  A tile writes from p + k of its
//                      first row on: dense inside, at most one unused slot per row between
//                      tiles, so no scan across tiles is needed before the copy
//   k_islice_emit      the same copy for units that took the serial walk
//   k_islice_pack      packs a unit's slices densely when the host asks for them
// Scratch: G (8 B per slice), per-tile sums, the per-row plan (first, count, offset).
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
#include <cub/device/device_scan.cuh>

#include "hbtgpu.h"

namespace {

#ifndef HBTGPU_ISLICE_NT
#define HBTGPU_ISLICE_NT 128
#endif
constexpr int NT = HBTGPU_ISLICE_NT;  // threads per CTA = rows per row tile
constexpr int NW = NT / 32;
constexpr int VT_SHIFT = 12;
constexpr int VT = 1 << VT_SHIFT;  // slices per validation tile
// slices of a row tile's window staged in shared memory: WMULT * NT, WMULT in {4, 8, 16} picked
// per call from the slices-per-row density (a bigger window costs resident CTAs)
#ifndef HBTGPU_EMIT_ILP
#define HBTGPU_EMIT_ILP 1
#endif
constexpr int EMIT_ILP = HBTGPU_EMIT_ILP;  // output slices a thread has in flight in the copy loop
constexpr int LOOKBACK = 8;        // longest chain of rows that start exactly on a slice end
constexpr uint8_t TT_ANALYSIS = 1;  // Slice::TransitionType::Analysis, Slicer.h:36
constexpr uint64_t NONE = ~0ull;

enum : uint32_t { UF_BAD_SLICES = 1, UF_BAD_ROWS = 2, UF_CHAIN = 4, UF_MARK = 8 };

struct Unit {
  uint64_t s0, ns;  // first slice, number of slices
  uint64_t r0, K;   // first row, number of rows
  uint64_t out0;    // the unit's output region
  uint64_t out_cap;
  uint64_t mark;
  uint32_t vt0, rt0;  // first validation tile, first row tile
};

struct UState {
  uint32_t flags;
  uint32_t pad;
  unsigned long long n_stop;  // first row the stop rule cuts
  unsigned long long err_k;   // first row the reference throws on
  unsigned long long max_c;   // largest new mark
};

struct Row {
  uint64_t T, D;
};

__device__ __forceinline__ ulonglong2 ld_td(const hbtgpu_slice* s) {
  return *reinterpret_cast<const ulonglong2*>(s);  // {tstamp, duration}
}
__device__ __forceinline__ uint64_t slice_end(const hbtgpu_slice* s) {
  const ulonglong2 v = ld_td(s);
  return v.x + v.y;
}
__device__ __forceinline__ Row ld_row(const uint64_t* rows, uint64_t stride, uint64_t k) {
  const uint64_t* p = rows + k * stride;
  return Row{p[0], p[1]};
}

// last unit whose first tile is <= tile (units without tiles share their successor's first tile)
template <bool ROWS>
__device__ __forceinline__ uint32_t find_unit(const Unit* units, uint32_t n, uint32_t tile) {
  uint32_t lo = 0, hi = n;
  while (hi - lo > 1) {
    const uint32_t mid = (lo + hi) >> 1;
    if ((ROWS ? units[mid].rt0 : units[mid].vt0) <= tile) lo = mid;
    else hi = mid;
  }
  return lo;
}

// A unit's slices in global memory.  G(j) = tile-local prefix + the tile's base.
struct GAcc {
  const hbtgpu_slice* S;
  const uint64_t* Gl;
  const uint64_t* vbase;
  __device__ __forceinline__ uint64_t t(uint32_t j) const { return S[j].tstamp; }
  __device__ __forceinline__ uint64_t e(uint32_t j) const { return slice_end(S + j); }
  __device__ __forceinline__ uint64_t G(uint32_t j) const { return Gl[j] + vbase[j >> VT_SHIFT]; }
};
// The window [lo, lo + W) of them in shared memory.  Every index the parallel form touches lies
// inside for rows and slices that meet the preconditions; for input that is being rejected
// anyway the clamp keeps the reads in bounds.
struct WAcc {
  const uint64_t *wt, *we, *wG;
  uint32_t lo, Wm1;
  __device__ __forceinline__ uint32_t at(uint32_t j) const { return min(j - lo, Wm1); }
  __device__ __forceinline__ uint64_t t(uint32_t j) const { return wt[at(j)]; }
  __device__ __forceinline__ uint64_t e(uint32_t j) const { return we[at(j)]; }
  __device__ __forceinline__ uint64_t G(uint32_t j) const { return wG[at(j)]; }
};

// first j in [a, b) with e_j >= x (b if none) = #{j : e_j < x} when the answer lies in [a, b]
template <class A>
__device__ __forceinline__ uint32_t first_end_ge(const A& acc, uint32_t a, uint32_t b, uint64_t x) {
  while (a < b) {
    const uint32_t mid = (a + b) >> 1;
    if (acc.e(mid) < x) a = mid + 1;
    else b = mid;
  }
  return a;
}
template <class A>
__device__ __forceinline__ uint32_t first_end_gt(const A& acc, uint32_t a, uint32_t b, uint64_t x) {
  while (a < b) {
    const uint32_t mid = (a + b) >> 1;
    if (acc.e(mid) <= x) a = mid + 1;
    else b = mid;
  }
  return a;
}
template <class A>
__device__ __forceinline__ uint32_t first_start_ge(const A& acc, uint32_t a, uint32_t b, uint64_t x) {
  while (a < b) {
    const uint32_t mid = (a + b) >> 1;
    if (acc.t(mid) < x) a = mid + 1;
    else b = mid;
  }
  return a;
}

// ---- block primitives (NT threads, all must call) ------------------------------------------
// inclusive prefix sum; *total = sum over the block
__device__ __forceinline__ uint64_t block_incl_sum_u64(uint64_t v, uint64_t* sh, uint64_t* total) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (int d = 1; d < 32; d <<= 1) {
    const uint64_t o = __shfl_up_sync(0xffffffffu, v, d);
    if (lane >= d) v += o;
  }
  __syncthreads();
  if (lane == 31) sh[w] = v;
  __syncthreads();
  if (w == 0) {  // the warp totals, scanned by the first warp
    uint64_t t = lane < NW ? sh[lane] : 0;
    for (int d = 1; d < NW; d <<= 1) {
      const uint64_t o = __shfl_up_sync(0xffffffffu, t, d);
      if (lane >= d) t += o;
    }
    if (lane < NW) sh[lane] = t;
  }
  __syncthreads();
  *total = sh[NW - 1];
  return v + (w ? sh[w - 1] : 0);
}

// exclusive prefix sum of 32-bit values; *total = sum over the block (sh: NW words)
__device__ __forceinline__ uint32_t block_excl_sum_u32(uint32_t v, uint32_t* sh, uint32_t* total) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  uint32_t s = v;
  for (int d = 1; d < 32; d <<= 1) {
    const uint32_t o = __shfl_up_sync(0xffffffffu, s, d);
    if (lane >= d) s += o;
  }
  __syncthreads();
  if (lane == 31) sh[w] = s;
  __syncthreads();
  if (w == 0) {
    uint32_t t = lane < NW ? sh[lane] : 0;
    for (int d = 1; d < NW; d <<= 1) {
      const uint32_t o = __shfl_up_sync(0xffffffffu, t, d);
      if (lane >= d) t += o;
    }
    if (lane < NW) sh[lane] = t;
  }
  __syncthreads();
  *total = sh[NW - 1];
  return s - v + (w ? sh[w - 1] : 0);
}

// One slice of an interval's answer: the i-th of cnt, trimmed as IntervalSlicer.cpp:146-173 trims
// front() and back().
__device__ __forceinline__ void trim_and_store(const uint4 (&q)[2], uint32_t i, uint32_t cnt, uint64_t T, uint64_t Tend,
                                               bool trim, hbtgpu_slice* dst) {
  union {
    uint4 q[2];
    hbtgpu_slice s;
  } u;
  u.q[0] = q[0];
  u.q[1] = q[1];
  if (trim) {
    const uint64_t end = u.s.tstamp + u.s.duration;
    if (i == 0 && u.s.tstamp < T) {
      u.s.duration = end - T;
      u.s.tstamp = T;
      u.s.swin_type = TT_ANALYSIS;
    }
    if (i == cnt - 1 && end > Tend) {
      u.s.duration -= end - Tend;
      u.s.swout_type = TT_ANALYSIS;
    }
  }
  uint4* d = reinterpret_cast<uint4*>(dst);
  d[0] = u.q[0];
  d[1] = u.q[1];
}
__device__ __forceinline__ void emit_one(const hbtgpu_slice* S, uint64_t j, uint32_t i, uint32_t cnt,
                                         uint64_t T, uint64_t Tend, bool trim, hbtgpu_slice* dst) {
  const uint4* src = reinterpret_cast<const uint4*>(S + j);
  const uint4 q[2] = {src[0], src[1]};
  trim_and_store(q, i, cnt, T, Tend, trim, dst);
}

// The reference's walk, row by row.  Fills the plan (first slice, count, output offset per row); nothing is copied.
__device__ void exact_walk(const hbtgpu_slice* S, uint64_t ns, const uint64_t* rows, uint64_t stride,
                           uint64_t K, uint64_t stop_ts, bool trim, bool abort_missing, uint64_t mark,
                           uint32_t* ioff, uint32_t* icnt, uint32_t* ifirst, hbtgpu_islice_unit* R) {
  uint64_t total = 0, k = 0;
  int64_t status = 0;
  uint64_t err_k = NONE;
  for (; k < K; ++k) {
    const Row rw = ld_row(rows, stride, k);
    const uint64_t T = rw.T, D = rw.D, Tend = T + D;
    if (Tend >= stop_ts) break;  // TraceMonitor.cpp:63-70
    bool thrown = (D == 0);      // IntervalSlicer.cpp:64
    bool null_res = false;
    uint64_t cur = min(mark, ns), consumed = 0, p = 0, n = 0;
    if (!thrown) {
      while (cur < ns && slice_end(S + cur) < T) {  // :76-82
        if (S[cur].duration == 0) {
          thrown = true;
          break;
        }
        ++cur;
        ++consumed;
      }
    }
    if (!thrown) {
      if (cur >= ns) null_res = true;                                   // :85-89
      else if (abort_missing && S[cur].tstamp > T) null_res = true;     // :91-95
    }
    if (!thrown && !null_res) {
      p = cur;
      uint64_t last_end = S[cur].tstamp, sumdur = 0;
      while (cur < ns && S[cur].tstamp < Tend) {  // :103-122
        const ulonglong2 td = ld_td(S + cur);
        if (abort_missing && last_end != td.x) {
          null_res = true;
          break;
        }
        if (!trim && td.x + td.y > Tend) break;
        sumdur += td.y;
        ++n;
        last_end = td.x + td.y;
        ++cur;
        ++consumed;
      }
      if (!null_res) {
        if (n == 0) null_res = true;                                   // :128-133
        else if (abort_missing && last_end < Tend) null_res = true;    // :139-143
        else {
          if (trim) {  // :146-159
            const ulonglong2 f = ld_td(S + p);
            if (f.x < T) sumdur -= f.y - (f.x + f.y - T);
          }
          if (last_end > Tend) --consumed;                  // :161-164
          if (trim && last_end > Tend) sumdur -= last_end - Tend;  // :167-173
          if (sumdur > D) thrown = true;                               // :175-177
          if (trim && abort_missing && sumdur != D) thrown = true;     // :178-183
        }
      }
    }
    if (thrown) {
      status = -EINVAL;
      err_k = k;
      break;
    }
    if (total + n > 0xffffffffull) {
      status = -EOVERFLOW;
      err_k = k;
      break;
    }
    ioff[k] = static_cast<uint32_t>(total);
    if (null_res) {
      icnt[k] = 0;
      ifirst[k] = 0;
    } else {
      icnt[k] = static_cast<uint32_t>(n);
      ifirst[k] = static_cast<uint32_t>(p);
      total += n;
      mark += consumed;  // TraceMonitor.cpp:89
    }
  }
  R->n_intervals = k;
  R->n_slices = total;
  R->mark = mark;
  R->status = status;
  R->err_interval = err_k;
  R->exact_path = 1;
  R->pad = 0;
}


struct Eval {
  bool null_res, thrown;
  uint32_t p, r, c;
};

// One row of the parallel form, started at slice p (lb or lb + 1).  x = where the take loop
// stops, q = #{t_j < Tend} (see the file comment).
template <class A>
__device__ __forceinline__ Eval eval_row(const A& a, uint32_t ns, uint64_t T, uint64_t D, uint64_t Tend, uint32_t p,
                                         uint32_t x, uint32_t q, bool trim, bool abort_missing) {
  Eval e{true, false, p, p, 0};
  if (p >= ns) return e;                       // IntervalSlicer.cpp:85-89
  const uint64_t tp = a.t(p);
  if (abort_missing && tp > T) return e;       // :91-95
  const uint32_t r = max(p, x);
  if (r == p) return e;                        // :128-133
  const uint64_t Gp = a.G(p);
  if (abort_missing) {
    // the walk compares every slice it looks at with its predecessor's end (:106-110), including
    // the one the no-trim break (:113-115) stops on
    uint32_t hi = trim ? r - 1 : min(r, max(p, q) - 1);
    hi = min(hi, ns - 1);  // (only reached with rows or slices that are being rejected anyway)
    if (a.G(hi) != Gp) return e;
  }
  const uint64_t er = a.e(r - 1);
  if (abort_missing && er < Tend) return e;    // :139-143
  e.null_res = false;
  e.r = r;
  e.c = r - (er > Tend ? 1u : 0u);             // :161-164
  if (!trim) {
    const uint64_t sumdur = (er - tp) - (a.G(r - 1) - Gp);
    if (sumdur > D) e.thrown = true;           // :175-177
  }
  return e;
}

// What a row's evaluation starts from: lb = #{e_j < T}, q = #{t_j < Tend}, x = where the take
// loop stops, hasB = a slice ends exactly at T.
struct RowInfo {
  uint64_t T, D;
  uint32_t lb, q, x;
  bool hasB;
};

__device__ RowInfo row_info_global(const GAcc& g, uint32_t ns, const Row rw, bool trim) {
  RowInfo ri;
  ri.T = rw.T;
  ri.D = rw.D;
  const uint64_t Tend = rw.T + rw.D;
  ri.lb = first_end_ge(g, 0, ns, rw.T);
  ri.hasB = ri.lb < ns && g.e(ri.lb) == rw.T;
  ri.q = first_start_ge(g, 0, ns, Tend);
  ri.x = trim ? ri.q : first_end_gt(g, 0, ns, Tend);
  return ri;
}

// Does row k (which starts exactly on a slice end) find that slice already consumed?  Walks
// back to the first row of its chain, then forward again.  info_at(j) hands out a row's
// searches (from shared memory for rows of this tile, k0 and later; computed in global memory
// for earlier ones), acc serves the slices of this tile's rows.
template <class A, class INFO>
__device__ uint32_t resolve_chain(const A& acc, const GAcc& g, uint32_t ns, INFO info_at, uint64_t k, uint64_t k0,
                                  uint64_t mark, bool trim, bool abort_missing, bool* give_up) {
  uint64_t j = k;
  uint32_t bsel = 0;
  int depth = 0;
  for (;;) {
    if (j == 0) {
      bsel = mark > info_at(0).lb ? 1u : 0u;
      break;
    }
    if (depth >= LOOKBACK) {
      *give_up = true;
      return 0;
    }
    --j;
    ++depth;
    if (!info_at(j).hasB) break;  // that row starts at its lb whatever came before
  }
  for (uint64_t i = j; i < k; ++i) {
    const RowInfo ri = info_at(i);
    const uint64_t Tend = ri.T + ri.D;
    const Eval e = i >= k0 ? eval_row(acc, ns, ri.T, ri.D, Tend, ri.lb + bsel, ri.x, ri.q, trim, abort_missing)
                           : eval_row(g, ns, ri.T, ri.D, Tend, ri.lb + bsel, ri.x, ri.q, trim, abort_missing);
    bsel = (!e.null_res && !e.thrown && e.c > info_at(i + 1).lb) ? 1u : 0u;
  }
  return bsel;
}

// ---- (unit, tile of VT slices): validation and the tile-local gap prefix ---------------------
__global__ void __launch_bounds__(NT)
k_islice_validate(const hbtgpu_slice* __restrict__ slices, const Unit* __restrict__ units, uint32_t n_units,
                  UState* __restrict__ ustate, uint64_t* __restrict__ Gl_all, uint64_t* __restrict__ vt_tot) {
  __shared__ uint64_t sh_red[32];
  const int tid = threadIdx.x;
  const uint32_t u = find_unit<false>(units, n_units, blockIdx.x);
  const Unit U = units[u];
  const hbtgpu_slice* S = slices + U.s0;
  uint64_t* Gl = Gl_all + U.s0;
  const uint32_t ns = static_cast<uint32_t>(U.ns);
  const uint32_t j0 = (blockIdx.x - U.vt0) << VT_SHIFT;
  uint64_t carry = 0;
  int bad = 0;
  for (uint32_t i = 0; i < VT && j0 + i < ns; i += NT) {
    const uint32_t j = j0 + i + tid;
    uint64_t gap = 0;
    if (j < ns) {
      const ulonglong2 td = ld_td(S + j);
      if (td.y == 0 || td.x + td.y < td.x) bad = 1;
      if (j > 0) {
        const uint64_t pe = slice_end(S + j - 1);
        if (td.x < pe) bad = 1;
        gap = td.x - pe;
      }
    }
    uint64_t tot;
    const uint64_t incl = block_incl_sum_u64(gap, sh_red, &tot);
    if (j < ns) Gl[j] = carry + incl;
    carry += tot;
  }
  if (__syncthreads_or(bad) && tid == 0) atomicOr(&ustate[u].flags, UF_BAD_SLICES);
  if (tid == 0) vt_tot[blockIdx.x] = carry;
}

// ---- per unit: bases of the validation tiles (one warp) ---------------------------------------
__global__ void __launch_bounds__(32)
k_islice_vscan(const Unit* __restrict__ units, const uint64_t* __restrict__ vt_tot, uint64_t* __restrict__ vt_base) {
  const Unit U = units[blockIdx.x];
  const uint32_t n = static_cast<uint32_t>((U.ns + VT - 1) >> VT_SHIFT);
  const int lane = threadIdx.x;
  uint64_t carry = 0;
  for (uint32_t i0 = 0; i0 < n; i0 += 32) {
    const uint32_t i = i0 + lane;
    const uint64_t v = i < n ? vt_tot[U.vt0 + i] : 0;
    uint64_t s = v;
    for (int d = 1; d < 32; d <<= 1) {
      const uint64_t o = __shfl_up_sync(0xffffffffu, s, d);
      if (lane >= d) s += o;
    }
    if (i < n) vt_base[U.vt0 + i] = carry + s - v;
    carry += __shfl_sync(0xffffffffu, s, 31);
  }
}

// ---- (unit, tile of NT rows): every row's first slice, count and new mark ---------------------
// per row tile: the slice window its rows can touch, [lo, ql + 2) with lo = #{e_j < T of its
// first row} and ql = #{t_j < end of its last row}.  One thread per tile.
__global__ void __launch_bounds__(128)
k_islice_bounds(const hbtgpu_slice* __restrict__ slices, const uint64_t* __restrict__ rows_all, uint64_t stride,
                const Unit* __restrict__ units, uint32_t n_units, uint32_t n_tiles, uint32_t* __restrict__ tile_lo,
                uint32_t* __restrict__ tile_ql) {
  const uint32_t tile = blockIdx.x * blockDim.x + threadIdx.x;
  if (tile >= n_tiles) return;
  const uint32_t u = find_unit<true>(units, n_units, tile);
  const Unit U = units[u];
  const uint64_t* rows = rows_all + U.r0 * stride;
  const uint32_t ns = static_cast<uint32_t>(U.ns);
  const uint64_t k0 = static_cast<uint64_t>(tile - U.rt0) * NT;
  const uint64_t kl = min(U.K, k0 + NT) - 1;
  const GAcc g{slices + U.s0, nullptr, nullptr};
  const Row rl = ld_row(rows, stride, kl);
  tile_lo[tile] = first_end_ge(g, 0, ns, ld_row(rows, stride, k0).T);
  tile_ql[tile] = first_start_ge(g, 0, ns, rl.T + rl.D);
}

template <int WMULT, int MIN_CTAS>
__global__ void __launch_bounds__(NT, MIN_CTAS)
k_islice_rows(const hbtgpu_slice* __restrict__ slices, const uint64_t* __restrict__ rows_all, uint64_t stride,
              const Unit* __restrict__ units, uint32_t n_units, UState* __restrict__ ustate,
              const uint64_t* __restrict__ Gl_all, const uint64_t* __restrict__ vt_base,
              const uint32_t* __restrict__ tile_lo, const uint32_t* __restrict__ tile_ql,
              uint32_t* __restrict__ ioff_all, uint32_t* __restrict__ icnt_all, uint32_t* __restrict__ ifirst_all,
              uint32_t* __restrict__ rt_tot, hbtgpu_slice* __restrict__ out_all, uint64_t stop_ts, uint32_t flags) {
  constexpr int WMAX = WMULT * NT;
  extern __shared__ __align__(16) uint64_t sh_win[];  // wt[WMAX], we[WMAX], wG[WMAX]
  __shared__ uint64_t sh_red[32];
  __shared__ uint64_t sh_T[NT], sh_D[NT];
  __shared__ uint32_t sh_stop, sh_err, sh_maxc, sh_flags, sh_firstact;
  __shared__ uint32_t sh_excl[NT + 1];
  __shared__ uint32_t sh_lb[NT], sh_q[NT], sh_x[NT];
  __shared__ uint8_t sh_hb[NT];

  const int tid = threadIdx.x;
  const uint32_t u = find_unit<true>(units, n_units, blockIdx.x);
  const Unit U = units[u];
  UState* st = ustate + u;
  if ((flags & HBTGPU_ISLICE_FORCE_EXACT) || (st->flags & UF_BAD_SLICES)) return;  // the serial walk does the unit
  const hbtgpu_slice* S = slices + U.s0;
  const uint64_t* rows = rows_all + U.r0 * stride;
  const bool trim = flags & HBTGPU_ISLICE_TRIM, abort_missing = flags & HBTGPU_ISLICE_ABORT_ON_MISSING;
  const uint32_t ns = static_cast<uint32_t>(U.ns);
  const uint64_t K = U.K;
  const uint64_t k0 = static_cast<uint64_t>(blockIdx.x - U.rt0) * NT;
  const uint64_t k = k0 + tid;
  const bool live = k < K;
  const GAcc g{S, Gl_all + U.s0, vt_base + U.vt0};

  uint64_t T = 0, D = 0, Tend = 0;
  uint32_t uflags = 0;
  bool cut = false;
  if (live) {
    const Row rw = ld_row(rows, stride, k);
    T = rw.T;
    D = rw.D;
    Tend = T + D;
    cut = Tend >= stop_ts;  // TraceMonitor.cpp:63-70
    // rows must move forward in time (all of them: then the stop rule cuts a suffix)
    if (Tend < T || (k + 1 < K && rows[(k + 1) * stride] < Tend)) uflags |= UF_BAD_ROWS;
  }
  sh_T[tid] = T;
  sh_D[tid] = D;
  if (tid == 0) {
    sh_firstact = NT;
    sh_stop = 0xffffffffu;
    sh_err = 0xffffffffu;
    sh_maxc = 0;
    sh_flags = 0;
  }
  const bool zero_len = live && !cut && D == 0;  // IntervalSlicer.cpp:64
  const bool act = live && !cut && D != 0;
  Eval e{true, false, 0, 0, 0};
  // One row: its searches (q and x by binary search; lb follows from the previous row's q when
  // the rows touch, as count rows do), then the chain question, then the evaluation.
  auto process = [&](const auto& acc, uint32_t a, uint32_t b) {
    uint32_t lb = 0, q = 0, x = 0;
    bool hasB = false;
    if (act) {
      q = first_start_ge(acc, a, b, Tend);
      x = trim ? q : first_end_gt(acc, a, b, Tend);
    }
    sh_q[tid] = act ? q : 0xffffffffu;
    __syncthreads();
    if (act) {
      const uint32_t qp = tid > 0 ? sh_q[tid - 1] : 0xffffffffu;
      if (qp != 0xffffffffu && sh_T[tid - 1] + sh_D[tid - 1] == T)
        lb = (qp > a && acc.e(qp - 1) >= T) ? qp - 1 : qp;  // at most one slice straddles T
      else
        lb = first_end_ge(acc, a, b, T);
      hasB = lb < ns && acc.e(lb) == T;
    }
    __syncthreads();
    sh_q[tid] = q;
    sh_lb[tid] = lb;
    sh_x[tid] = x;
    sh_hb[tid] = hasB;
    __syncthreads();
    if (!act) return;
    uint32_t bsel = 0;
    if (k == 0 && U.mark > lb + (hasB ? 1u : 0u)) uflags |= UF_MARK;  // mark ahead of the first row's slices
    if (hasB) {
      auto info_at = [&](uint64_t kk) {
        if (kk >= k0) {
          const uint32_t i = static_cast<uint32_t>(kk - k0);
          return RowInfo{sh_T[i], sh_D[i], sh_lb[i], sh_q[i], sh_x[i], sh_hb[i] != 0};
        }
        return row_info_global(g, ns, ld_row(rows, stride, kk), trim);
      };
      bool give_up = false;
      bsel = resolve_chain(acc, g, ns, info_at, k, k0, U.mark, trim, abort_missing, &give_up);
      if (give_up) uflags |= UF_CHAIN;
    }
    e = eval_row(acc, ns, T, D, Tend, lb + bsel, x, q, trim, abort_missing);
  };
  if (__syncthreads_or(act)) {
    const uint32_t lo = tile_lo[blockIdx.x], ql = tile_ql[blockIdx.x];
    const uint32_t hiw = min(ns, ql + 2);
    const uint32_t W = hiw > lo ? hiw - lo : 0;
    if (W != 0 && W <= WMAX) {
      uint64_t *wt = sh_win, *we = sh_win + WMAX, *wG = sh_win + 2 * WMAX;
      for (uint32_t i = tid; i < W; i += NT) {
        const ulonglong2 td = ld_td(S + lo + i);
        wt[i] = td.x;
        we[i] = td.x + td.y;
      }
      __syncthreads();
      // gap prefix over the window: each thread owns c consecutive slices
      const uint32_t c = (W + NT - 1) / NT;  // <= WMAX / NT
      const uint32_t i0 = tid * c;
      uint64_t loc[WMAX / NT];
      uint64_t sum = 0;
#pragma unroll
      for (uint32_t j = 0; j < WMAX / NT; ++j) {
        const uint32_t i = i0 + j;
        if (j < c && i > 0 && i < W) sum += wt[i] - we[i - 1];
        loc[j] = sum;
      }
      uint64_t tot;
      const uint64_t base = block_incl_sum_u64(sum, sh_red, &tot) - sum;
#pragma unroll
      for (uint32_t j = 0; j < WMAX / NT; ++j)
        if (j < c && i0 + j < W) wG[i0 + j] = base + loc[j];
      __syncthreads();
      process(WAcc{wt, we, wG, lo, W - 1}, lo, min(ns, ql + 1));
    } else {
      process(g, 0u, ns);
    }
  }
  const bool counted = act && !e.null_res && !e.thrown;
  const uint32_t cnt = counted ? e.r - e.p : 0u;
  // tile sums: warp reductions, then one shared atomic per warp
  {
    const uint32_t m1 = __reduce_min_sync(0xffffffffu, cut ? static_cast<uint32_t>(tid) : 0xffffffffu);
    const uint32_t m2 = __reduce_min_sync(0xffffffffu, (zero_len || (act && e.thrown)) ? static_cast<uint32_t>(tid) : 0xffffffffu);
    const uint32_t mx = __reduce_max_sync(0xffffffffu, counted ? e.c : 0u);
    const uint32_t fl = __reduce_or_sync(0xffffffffu, uflags);
    const uint32_t fa = __reduce_min_sync(0xffffffffu, act ? static_cast<uint32_t>(tid) : 0xffffffffu);
    if ((tid & 31) == 0) {
      if (m1 != 0xffffffffu) atomicMin(&sh_stop, m1);
      if (m2 != 0xffffffffu) atomicMin(&sh_err, m2);
      if (mx) atomicMax(&sh_maxc, mx);
      if (fl) atomicOr(&sh_flags, fl);
      if (fa != 0xffffffffu) atomicMin(&sh_firstact, fa);
    }
  }
  // The tile's place in the unit's output region.  Rows take slices [p_k, r_k) and r_k <= p_{k+1} + 1,
  // so a tile that starts writing at p + k (p, k of its first row) ends before the next tile's start:
  // dense inside the tile, at most one unused slot per row between tiles, no scan across tiles.
  uint32_t tile_total;
  const uint32_t excl = block_excl_sum_u32(cnt, reinterpret_cast<uint32_t*>(sh_red), &tile_total);
  sh_excl[tid] = excl;
  if (tid == 0) sh_excl[NT] = tile_total;
  sh_lb[tid] = e.p;  // (every chain walk is done: block_excl_sum_u32 has barriers)
  __syncthreads();
  const uint32_t fa = sh_firstact;
  const uint64_t base = fa < NT ? static_cast<uint64_t>(sh_lb[fa]) + (k0 + fa) : 0;
  if (live) {
    ioff_all[U.r0 + k] = static_cast<uint32_t>(base + excl);
    icnt_all[U.r0 + k] = cnt;
    ifirst_all[U.r0 + k] = counted ? e.p : 0u;
  }
  if (base + tile_total <= U.out_cap) {
    // one slice per thread and step, EMIT_ILP steps in flight: the owning row by binary search over
    // the tile's offsets, then a 32-byte copy with the first / last slice of a row trimmed
    hbtgpu_slice* out = out_all + U.out0 + base;
    for (uint32_t o0 = tid; o0 < tile_total; o0 += EMIT_ILP * NT) {
      uint32_t row[EMIT_ILP];
#pragma unroll
      for (int j = 0; j < EMIT_ILP; ++j) {
        const uint32_t o = min(o0 + j * NT, tile_total - 1);
        uint32_t lo = 0, hi = NT;  // last row with sh_excl[row] <= o
        while (hi - lo > 1) {
          const uint32_t mid = (lo + hi) >> 1;
          if (sh_excl[mid] <= o) lo = mid;
          else hi = mid;
        }
        row[j] = lo;
      }
      uint4 q[EMIT_ILP][2];
#pragma unroll
      for (int j = 0; j < EMIT_ILP; ++j) {
        const uint32_t o = min(o0 + j * NT, tile_total - 1);
        const uint4* src = reinterpret_cast<const uint4*>(S + sh_lb[row[j]] + (o - sh_excl[row[j]]));
        q[j][0] = src[0];
        q[j][1] = src[1];
      }
#pragma unroll
      for (int j = 0; j < EMIT_ILP; ++j) {
        const uint32_t o = o0 + j * NT;
        if (o < tile_total) {
          const uint32_t r_ = row[j], i = o - sh_excl[r_], c = sh_excl[r_ + 1] - sh_excl[r_];
          trim_and_store(q[j], i, c, sh_T[r_], sh_T[r_] + sh_D[r_], trim, out + o);
        }
      }
    }
  } else if (tid == 0) {
    atomicOr(&sh_flags, UF_BAD_ROWS);  // only rows that do not move forward in time can get here
  }
  __syncthreads();
  if (tid == 0) {
    rt_tot[blockIdx.x] = tile_total;
    if (sh_stop != 0xffffffffu) atomicMin(&st->n_stop, static_cast<unsigned long long>(k0 + sh_stop));
    if (sh_err != 0xffffffffu) atomicMin(&st->err_k, static_cast<unsigned long long>(k0 + sh_err));
    if (sh_maxc) atomicMax(&st->max_c, static_cast<unsigned long long>(sh_maxc));
    if (sh_flags) atomicOr(&st->flags, sh_flags);
  }
}

// ---- per unit: bases of the row tiles, the summary; the serial walk for units that need it ----
__global__ void __launch_bounds__(NT)
k_islice_rscan(const hbtgpu_slice* __restrict__ slices, const uint64_t* __restrict__ rows_all, uint64_t stride,
               const Unit* __restrict__ units, const UState* __restrict__ ustate, const uint32_t* __restrict__ rt_tot,
               uint64_t* __restrict__ rt_base, uint32_t* __restrict__ ioff_all, uint32_t* __restrict__ icnt_all,
               uint32_t* __restrict__ ifirst_all, hbtgpu_islice_unit* __restrict__ results, uint64_t stop_ts,
               uint32_t flags) {
  __shared__ uint64_t sh_red[32];
  __shared__ hbtgpu_islice_unit sh_R;
  const int tid = threadIdx.x;
  const Unit U = units[blockIdx.x];
  const UState st = ustate[blockIdx.x];
  const uint32_t n_tiles = static_cast<uint32_t>((U.K + NT - 1) / NT);
  const bool trim = flags & HBTGPU_ISLICE_TRIM, abort_missing = flags & HBTGPU_ISLICE_ABORT_ON_MISSING;
  if ((flags & HBTGPU_ISLICE_FORCE_EXACT) || st.flags != 0) {
    if (tid == 0) {
      exact_walk(slices + U.s0, U.ns, rows_all + U.r0 * stride, stride, U.K, stop_ts, trim, abort_missing, U.mark,
                 ioff_all + U.r0, icnt_all + U.r0, ifirst_all + U.r0, &sh_R);
      sh_R.pad = st.flags;
      results[blockIdx.x] = sh_R;
    }
    __syncthreads();
    for (uint32_t i = tid; i < n_tiles; i += NT) {
      const uint64_t kk = static_cast<uint64_t>(i) * NT;
      rt_base[U.rt0 + i] = kk < sh_R.n_intervals ? ioff_all[U.r0 + kk] : sh_R.n_slices;
    }
    return;
  }
  uint64_t carry = 0;
  for (uint32_t i0 = 0; i0 < n_tiles; i0 += NT) {
    const uint32_t i = i0 + tid;
    const uint64_t v = i < n_tiles ? rt_tot[U.rt0 + i] : 0;
    uint64_t tot;
    const uint64_t incl = block_incl_sum_u64(v, sh_red, &tot);
    if (i < n_tiles) rt_base[U.rt0 + i] = carry + incl - v;
    carry += tot;
  }
  if (tid == 0) {
    const uint64_t n_int = min(U.K, static_cast<uint64_t>(st.n_stop));
    const bool thrown = st.err_k < n_int;
    hbtgpu_islice_unit R;
    R.n_intervals = thrown ? st.err_k : n_int;
    R.n_slices = carry;
    R.mark = max(U.mark, static_cast<uint64_t>(st.max_c));
    R.status = thrown ? -EINVAL : (carry > 0xffffffffull ? -EOVERFLOW : 0);
    R.err_interval = thrown ? st.err_k : NONE;
    R.exact_path = 0;
    R.pad = 0;
    results[blockIdx.x] = R;
  }
}

// ---- (unit, tile of NT rows) of units that took the serial walk: output offsets, copy + trim ---
__global__ void __launch_bounds__(NT)
k_islice_emit(const hbtgpu_slice* __restrict__ slices, const uint64_t* __restrict__ rows_all, uint64_t stride,
              const Unit* __restrict__ units, uint32_t n_units, const hbtgpu_islice_unit* __restrict__ results,
              const uint64_t* __restrict__ rt_base, uint32_t* __restrict__ ioff_all,
              const uint32_t* __restrict__ icnt_all, const uint32_t* __restrict__ ifirst_all,
              hbtgpu_slice* __restrict__ out_all, uint32_t flags) {
  __shared__ uint64_t sh_red[32];
  __shared__ uint64_t sh_T[NT], sh_Tend[NT];
  __shared__ uint32_t sh_excl[NT + 1], sh_p[NT];
  const int tid = threadIdx.x;
  const uint32_t u = find_unit<true>(units, n_units, blockIdx.x);
  const Unit U = units[u];
  const uint64_t n = results[u].n_intervals;
  const uint64_t k0 = static_cast<uint64_t>(blockIdx.x - U.rt0) * NT;
  if (!results[u].exact_path) return;  // the rows pass has written the others already
  if (k0 >= n || results[u].status == -EOVERFLOW) return;
  const hbtgpu_slice* S = slices + U.s0;
  const uint64_t* rows = rows_all + U.r0 * stride;
  const bool trim = flags & HBTGPU_ISLICE_TRIM;
  const uint64_t k = k0 + tid;
  const bool in = k < n;
  const uint32_t cnt = in ? icnt_all[U.r0 + k] : 0u;
  uint64_t tile_total;
  const uint64_t incl = block_incl_sum_u64(cnt, sh_red, &tile_total);
  const uint32_t excl = static_cast<uint32_t>(incl - cnt);
  const uint64_t base = rt_base[blockIdx.x];
  sh_excl[tid] = excl;
  if (tid == 0) sh_excl[NT] = static_cast<uint32_t>(tile_total);
  if (in) {
    ioff_all[U.r0 + k] = static_cast<uint32_t>(base + excl);
    sh_p[tid] = ifirst_all[U.r0 + k];
    if (cnt) {
      const Row rw = ld_row(rows, stride, k);
      sh_T[tid] = rw.T;
      sh_Tend[tid] = rw.T + rw.D;
    }
  }
  __syncthreads();
  hbtgpu_slice* out = out_all + U.out0 + base;
  for (uint32_t o = tid; o < tile_total; o += NT) {
    uint32_t lo = 0, hi = NT;  // last row with sh_excl[row] <= o
    while (hi - lo > 1) {
      const uint32_t mid = (lo + hi) >> 1;
      if (sh_excl[mid] <= o) lo = mid;
      else hi = mid;
    }
    const uint32_t i = o - sh_excl[lo];
    emit_one(S, sh_p[lo] + i, i, sh_excl[lo + 1] - sh_excl[lo], sh_T[lo], sh_Tend[lo], trim, out + o);
  }
}


// a unit's slices, tile after tile, packed densely (for the host getter)
__global__ void __launch_bounds__(256)
k_islice_pack(const Unit U, const uint32_t* __restrict__ ioff_all, const uint32_t* __restrict__ rt_tot,
              const uint64_t* __restrict__ rt_base, const hbtgpu_slice* __restrict__ out_all,
              hbtgpu_slice* __restrict__ dst) {
  const uint32_t tile = U.rt0 + blockIdx.x;
  const uint32_t n = rt_tot[tile];
  if (n == 0) return;
  const uint4* src = reinterpret_cast<const uint4*>(out_all + U.out0 + ioff_all[U.r0 + static_cast<uint64_t>(blockIdx.x) * NT]);
  uint4* d = reinterpret_cast<uint4*>(dst + rt_base[tile]);
  for (uint32_t i = threadIdx.x; i < 2 * n; i += blockDim.x) d[i] = src[i];
}


#include "hbtgpu_islice_bin_kernels.inc"

}  // namespace

struct hbtgpu_islice_ctx {
  int device = 0;
  std::string err;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev2 = nullptr, ev3 = nullptr;
  float last_ms = 0.f;
  // device buffers (grown on demand)
  hbtgpu_slice* d_slices = nullptr;  size_t cap_slices = 0;
  uint64_t* d_rows = nullptr;        size_t cap_rows = 0;
  uint64_t* d_G = nullptr;           size_t cap_G = 0;
  uint64_t* d_vt_tot = nullptr;      size_t cap_vt_tot = 0;
  uint64_t* d_vt_base = nullptr;     size_t cap_vt_base = 0;
  uint32_t* d_rt_tot = nullptr;      size_t cap_rt_tot = 0;
  uint64_t* d_rt_base = nullptr;     size_t cap_rt_base = 0;
  uint32_t* d_tile_lo = nullptr;     size_t cap_tile_lo = 0;
  uint32_t* d_tile_ql = nullptr;     size_t cap_tile_ql = 0;
  hbtgpu_slice* d_tmp = nullptr;     size_t cap_tmp = 0;
  uint32_t* d_ioff = nullptr;        size_t cap_ioff = 0;
  uint32_t* d_icnt = nullptr;        size_t cap_icnt = 0;
  uint32_t* d_ifirst = nullptr;      size_t cap_ifirst = 0;
  hbtgpu_slice* d_out = nullptr;     size_t cap_out = 0;
  Unit* d_units = nullptr;           size_t cap_units = 0;
  UState* d_ustate = nullptr;        size_t cap_ustate = 0;
  hbtgpu_islice_unit* d_res = nullptr; size_t cap_res = 0;
  uint64_t* d_base = nullptr;        size_t cap_base = 0;
  // binner (f1)
  unsigned long long* d_bin_tab = nullptr; size_t cap_bin_tab = 0;  // key, dur, obs, nint: 4 x (cap + 1)
  uint32_t* d_bin_bid = nullptr;     size_t cap_bin_bid = 0;
  uint8_t* d_has = nullptr;          size_t cap_has = 0;
  uint32_t* d_bt_tot = nullptr;      size_t cap_bt_tot = 0;
  uint64_t* d_bt_base = nullptr;     size_t cap_bt_base = 0;
  BinEntry* d_bin_list = nullptr;    size_t cap_bin_list = 0;
  uint32_t* d_bin_up = nullptr;      size_t cap_bin_up = 0;
  unsigned long long* d_bin_M = nullptr; size_t cap_bin_M = 0;
  unsigned long long* d_bin_ctr = nullptr; size_t cap_bin_ctr = 0;
  uint8_t* d_sel = nullptr;          size_t cap_sel = 0;
  hbtgpu_bin_freq* d_bin_freqs = nullptr; size_t cap_bin_freqs = 0;
  uint64_t* d_ext_ids = nullptr;     size_t cap_ext_ids = 0;
  uint32_t* d_ext_bids = nullptr;    size_t cap_ext_bids = 0;
  hbtgpu_bin_summary bin_summary{};
  bool have_bin = false;
  // last call
  std::vector<Unit> units;
  std::vector<UState> ustate;
  std::vector<hbtgpu_islice_unit> res;
  bool have = false;
};

namespace {

int ifail(hbtgpu_islice_ctx* c, int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  if (c) c->err = buf;
  return code;
}

#define IK(call)                                                                                    \
  do {                                                                                              \
    cudaError_t e_ = (call);                                                                        \
    if (e_ != cudaSuccess)                                                                          \
      return ifail(c, -EIO, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
  } while (0)

template <class T>
int grow(hbtgpu_islice_ctx* c, T** p, size_t* cap, size_t need) {
  if (need <= *cap && *p) return 0;
  const size_t ncap = std::max<size_t>(std::max(need, *cap + *cap / 2), 16);
  if (*p) {
    cudaFree(*p);
    *p = nullptr;
    *cap = 0;
  }
  IK(cudaMalloc(p, ncap * sizeof(T)));
  *cap = ncap;
  return 0;
}

constexpr size_t rows_smem(int wmult) { return 3 * static_cast<size_t>(wmult) * NT * sizeof(uint64_t); }

}  // namespace

extern "C" {

hbtgpu_islice_ctx* hbtgpu_islice_create(int device) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || device < 0 || device >= n) return nullptr;  // no CPU fallback
  if (cudaSetDevice(device) != cudaSuccess) return nullptr;
  auto* c = new hbtgpu_islice_ctx();
  c->device = device;
  if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess ||
      cudaEventCreate(&c->ev0) != cudaSuccess || cudaEventCreate(&c->ev1) != cudaSuccess ||
      cudaEventCreate(&c->ev2) != cudaSuccess || cudaEventCreate(&c->ev3) != cudaSuccess ||
      cudaFuncSetAttribute(k_islice_rows<4, 12>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                           static_cast<int>(rows_smem(4))) != cudaSuccess ||
      cudaFuncSetAttribute(k_islice_rows<8, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                           static_cast<int>(rows_smem(8))) != cudaSuccess ||
      cudaFuncSetAttribute(k_islice_rows<16, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                           static_cast<int>(rows_smem(16))) != cudaSuccess) {
    hbtgpu_islice_destroy(c);
    return nullptr;
  }
  return c;
}

void hbtgpu_islice_destroy(hbtgpu_islice_ctx* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  cudaFree(c->d_slices); cudaFree(c->d_rows); cudaFree(c->d_G); cudaFree(c->d_vt_tot); cudaFree(c->d_vt_base);
  cudaFree(c->d_rt_tot); cudaFree(c->d_rt_base); cudaFree(c->d_tile_lo); cudaFree(c->d_tile_ql); cudaFree(c->d_tmp); cudaFree(c->d_ioff); cudaFree(c->d_icnt); cudaFree(c->d_ifirst);
  cudaFree(c->d_out); cudaFree(c->d_units); cudaFree(c->d_ustate); cudaFree(c->d_res); cudaFree(c->d_base);
  cudaFree(c->d_bin_tab); cudaFree(c->d_bin_bid); cudaFree(c->d_has); cudaFree(c->d_bt_tot); cudaFree(c->d_bt_base);
  cudaFree(c->d_bin_list); cudaFree(c->d_bin_up); cudaFree(c->d_bin_freqs); cudaFree(c->d_ext_ids); cudaFree(c->d_ext_bids); cudaFree(c->d_bin_M); cudaFree(c->d_bin_ctr); cudaFree(c->d_sel);
  for (cudaEvent_t e : {c->ev0, c->ev1, c->ev2, c->ev3})
    if (e) cudaEventDestroy(e);
  if (c->stream) cudaStreamDestroy(c->stream);
  delete c;
}

const char* hbtgpu_islice_last_error(hbtgpu_islice_ctx* c) { return c ? c->err.c_str() : "no context"; }

float hbtgpu_islice_last_ms(const hbtgpu_islice_ctx* c) { return c ? c->last_ms : 0.f; }

int hbtgpu_islice_release(hbtgpu_islice_ctx* c, const hbtgpu_slice* slices, const uint64_t* slice_begin,
                          const uint64_t* rows, uint64_t row_stride, const uint64_t* row_begin,
                          uint32_t n_units, uint64_t stop_ts, uint32_t flags, uint64_t* marks,
                          hbtgpu_islice_unit* units_out) {
  if (!c) return -EINVAL;
  c->have = false;
  c->have_bin = false;
  if (!slice_begin || !row_begin || !marks || n_units == 0) return ifail(c, -EINVAL, "null argument or no units");
  if (row_stride < 2) return ifail(c, -EINVAL, "row_stride must be >= 2 (tstamp, duration, ...)");
  IK(cudaSetDevice(c->device));
  const uint64_t s_base = slice_begin[0], r_base = row_begin[0];
  const uint64_t total_s = slice_begin[n_units] - s_base, total_r = row_begin[n_units] - r_base;
  c->units.assign(n_units, Unit{});
  uint64_t n_vt = 0, n_rt = 0, out_run = 0;
  for (uint32_t u = 0; u < n_units; ++u) {
    if (slice_begin[u + 1] < slice_begin[u] || row_begin[u + 1] < row_begin[u])
      return ifail(c, -EINVAL, "begin arrays must not decrease (unit %u)", u);
    Unit& U = c->units[u];
    U.s0 = slice_begin[u] - s_base;
    U.ns = slice_begin[u + 1] - slice_begin[u];
    U.r0 = row_begin[u] - r_base;
    U.K = row_begin[u + 1] - row_begin[u];
    if (U.ns >= 0xfffffff0ull || U.K >= 0xfffffff0ull) return ifail(c, -EINVAL, "unit %u: 2^32 or more slices or rows", u);
    U.out0 = out_run;
    U.out_cap = U.ns + U.K + NT;  // the rows pass leaves at most one unused slot per row
    out_run += U.out_cap;
    U.mark = marks[u];
    U.vt0 = static_cast<uint32_t>(n_vt);
    U.rt0 = static_cast<uint32_t>(n_rt);
    n_vt += (U.ns + VT - 1) >> VT_SHIFT;
    n_rt += (U.K + NT - 1) / NT;
    if (n_vt >= 0x7fffffffull || n_rt >= 0x7fffffffull) return ifail(c, -EINVAL, "too many tiles");
  }
  if ((total_s && !slices) || (total_r && !rows)) return ifail(c, -EINVAL, "null slices or rows");

  // inputs
  const hbtgpu_slice* d_slices = nullptr;
  const uint64_t* d_rows = nullptr;
  uint64_t stride = row_stride;
  if (flags & HBTGPU_ISLICE_DEVICE_INPUT) {
    d_slices = slices + s_base;
    d_rows = rows + r_base * row_stride;
    if (reinterpret_cast<uintptr_t>(d_slices) % 16)
      return ifail(c, -EINVAL, "device slices must be 16-byte aligned");
  } else {
    if (int rc = grow(c, &c->d_slices, &c->cap_slices, total_s + 1)) return rc;
    if (int rc = grow(c, &c->d_rows, &c->cap_rows, 2 * total_r + 2)) return rc;
    if (total_s)
      IK(cudaMemcpyAsync(c->d_slices, slices + s_base, total_s * sizeof(hbtgpu_slice), cudaMemcpyHostToDevice,
                         c->stream));
    if (total_r) {  // only {tstamp, duration} of every row travel
      if (row_stride == 2) {
        IK(cudaMemcpyAsync(c->d_rows, rows + r_base * 2, total_r * 16, cudaMemcpyHostToDevice, c->stream));
      } else {
        IK(cudaMemcpy2DAsync(c->d_rows, 16, rows + r_base * row_stride, row_stride * 8, 16, total_r,
                             cudaMemcpyHostToDevice, c->stream));
      }
    }
    d_slices = c->d_slices;
    d_rows = c->d_rows;
    stride = 2;
  }
  if (int rc = grow(c, &c->d_G, &c->cap_G, total_s + 1)) return rc;
  if (int rc = grow(c, &c->d_vt_tot, &c->cap_vt_tot, n_vt + 2)) return rc;
  if (int rc = grow(c, &c->d_vt_base, &c->cap_vt_base, n_vt + 2)) return rc;
  if (int rc = grow(c, &c->d_rt_tot, &c->cap_rt_tot, n_rt + 1)) return rc;
  if (int rc = grow(c, &c->d_rt_base, &c->cap_rt_base, n_rt + 1)) return rc;
  if (int rc = grow(c, &c->d_tile_lo, &c->cap_tile_lo, n_rt + 1)) return rc;
  if (int rc = grow(c, &c->d_tile_ql, &c->cap_tile_ql, n_rt + 1)) return rc;
  if (int rc = grow(c, &c->d_ioff, &c->cap_ioff, total_r + 1)) return rc;
  if (int rc = grow(c, &c->d_icnt, &c->cap_icnt, total_r + 1)) return rc;
  if (int rc = grow(c, &c->d_ifirst, &c->cap_ifirst, total_r + 1)) return rc;
  if (int rc = grow(c, &c->d_units, &c->cap_units, n_units)) return rc;
  if (int rc = grow(c, &c->d_ustate, &c->cap_ustate, n_units)) return rc;
  if (int rc = grow(c, &c->d_res, &c->cap_res, n_units)) return rc;
  if (int rc = grow(c, &c->d_out, &c->cap_out, out_run + 1)) return rc;
  c->ustate.assign(n_units, UState{0u, 0u, NONE, NONE, 0ull});
  IK(cudaMemcpyAsync(c->d_units, c->units.data(), n_units * sizeof(Unit), cudaMemcpyHostToDevice, c->stream));
  IK(cudaMemcpyAsync(c->d_ustate, c->ustate.data(), n_units * sizeof(UState), cudaMemcpyHostToDevice, c->stream));

  IK(cudaEventRecord(c->ev0, c->stream));
  if (n_vt) {
    k_islice_validate<<<static_cast<unsigned>(n_vt), NT, 0, c->stream>>>(d_slices, c->d_units, n_units, c->d_ustate,
                                                                         c->d_G, c->d_vt_tot);
    k_islice_vscan<<<n_units, 32, 0, c->stream>>>(c->d_units, c->d_vt_tot, c->d_vt_base);
  }
  if (n_rt) {
    k_islice_bounds<<<static_cast<unsigned>((n_rt + 127) / 128), 128, 0, c->stream>>>(
        d_slices, d_rows, stride, c->d_units, n_units, static_cast<uint32_t>(n_rt), c->d_tile_lo, c->d_tile_ql);
    // window size by density: a tile of NT rows touches about NT * slices-per-row slices
    const double density = static_cast<double>(total_s) / static_cast<double>(total_r ? total_r : 1);
#define ROWS_ARGS d_slices, d_rows, stride, c->d_units, n_units, c->d_ustate, c->d_G, c->d_vt_base, c->d_tile_lo, \
                  c->d_tile_ql, c->d_ioff, c->d_icnt, c->d_ifirst, c->d_rt_tot, c->d_out, stop_ts, flags
    if (density < 3.0)
      k_islice_rows<4, 12><<<static_cast<unsigned>(n_rt), NT, rows_smem(4), c->stream>>>(ROWS_ARGS);
    else if (density < 6.5)
      k_islice_rows<8, 8><<<static_cast<unsigned>(n_rt), NT, rows_smem(8), c->stream>>>(ROWS_ARGS);
    else
      k_islice_rows<16, 4><<<static_cast<unsigned>(n_rt), NT, rows_smem(16), c->stream>>>(ROWS_ARGS);
#undef ROWS_ARGS
  }
  k_islice_rscan<<<n_units, NT, 0, c->stream>>>(d_slices, d_rows, stride, c->d_units, c->d_ustate, c->d_rt_tot,
                                                c->d_rt_base, c->d_ioff, c->d_icnt, c->d_ifirst, c->d_res, stop_ts,
                                                flags);
  IK(cudaGetLastError());
  IK(cudaEventRecord(c->ev1, c->stream));
  c->res.assign(n_units, hbtgpu_islice_unit{});
  IK(cudaMemcpyAsync(c->res.data(), c->d_res, n_units * sizeof(hbtgpu_islice_unit), cudaMemcpyDeviceToHost,
                     c->stream));
  IK(cudaStreamSynchronize(c->stream));

  // units that took the serial walk: their slices are copied now, behind the regular regions
  // when they do not fit their own
  bool any_exact = false, moved = false;
  for (uint32_t u = 0; u < n_units; ++u) {
    if (!c->res[u].exact_path || c->res[u].status == -EOVERFLOW || c->res[u].n_slices == 0) continue;
    any_exact = true;
    if (c->res[u].n_slices > c->units[u].out_cap) {
      c->units[u].out0 = out_run;
      c->units[u].out_cap = c->res[u].n_slices;
      out_run += c->res[u].n_slices;
      moved = true;
    }
  }
  if (moved) {
    if (out_run + 1 > c->cap_out) {  // grow, keeping what the rows pass wrote
      hbtgpu_slice* q = nullptr;
      const size_t ncap = out_run + 1;
      IK(cudaMalloc(&q, ncap * sizeof(hbtgpu_slice)));
      cudaError_t e = cudaMemcpyAsync(q, c->d_out, c->cap_out * sizeof(hbtgpu_slice), cudaMemcpyDeviceToDevice, c->stream);
      if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
      if (e != cudaSuccess) {
        cudaFree(q);
        return ifail(c, -EIO, "device copy failed: %s", cudaGetErrorString(e));
      }
      cudaFree(c->d_out);
      c->d_out = q;
      c->cap_out = ncap;
    }
    IK(cudaMemcpyAsync(c->d_units, c->units.data(), n_units * sizeof(Unit), cudaMemcpyHostToDevice, c->stream));
  }
  IK(cudaEventRecord(c->ev2, c->stream));
  if (n_rt && any_exact)
    k_islice_emit<<<static_cast<unsigned>(n_rt), NT, 0, c->stream>>>(d_slices, d_rows, stride, c->d_units, n_units,
                                                                     c->d_res, c->d_rt_base, c->d_ioff, c->d_icnt,
                                                                     c->d_ifirst, c->d_out, flags);
  IK(cudaGetLastError());
  IK(cudaEventRecord(c->ev3, c->stream));
  IK(cudaStreamSynchronize(c->stream));
  float a = 0.f, b = 0.f;
  IK(cudaEventElapsedTime(&a, c->ev0, c->ev1));
  IK(cudaEventElapsedTime(&b, c->ev2, c->ev3));
  c->last_ms = a + b;

  for (uint32_t u = 0; u < n_units; ++u) {
    c->res[u].pad = 0;
    if (c->res[u].status == 0) marks[u] = c->res[u].mark;
    if (units_out) units_out[u] = c->res[u];
  }
  c->have = true;
  return 0;
}

int hbtgpu_islice_get_units(hbtgpu_islice_ctx* c, hbtgpu_islice_unit* units, uint32_t cap) {
  if (!c || !units) return -EINVAL;
  if (!c->have) return ifail(c, -EINVAL, "no results: call hbtgpu_islice_release first");
  if (cap < c->res.size()) return ifail(c, -ENOSPC, "%zu units, room for %u", c->res.size(), cap);
  memcpy(units, c->res.data(), c->res.size() * sizeof(hbtgpu_islice_unit));
  return 0;
}

int hbtgpu_islice_get(hbtgpu_islice_ctx* c, uint32_t unit, uint32_t* counts, uint64_t counts_cap,
                      hbtgpu_slice* slices, uint64_t slices_cap) {
  if (!c) return -EINVAL;
  if (!c->have) return ifail(c, -EINVAL, "no results: call hbtgpu_islice_release first");
  if (unit >= c->res.size()) return ifail(c, -EINVAL, "unit %u out of range", unit);
  IK(cudaSetDevice(c->device));
  const Unit& U = c->units[unit];
  const hbtgpu_islice_unit& R = c->res[unit];
  if (counts) {
    if (counts_cap < R.n_intervals) return ifail(c, -ENOSPC, "%llu intervals, room for %llu",
                                                 (unsigned long long)R.n_intervals, (unsigned long long)counts_cap);
    if (R.n_intervals)
      IK(cudaMemcpyAsync(counts, c->d_icnt + U.r0, R.n_intervals * sizeof(uint32_t), cudaMemcpyDeviceToHost,
                         c->stream));
  }
  if (slices) {
    if (slices_cap < R.n_slices) return ifail(c, -ENOSPC, "%llu slices, room for %llu",
                                              (unsigned long long)R.n_slices, (unsigned long long)slices_cap);
    if (R.n_slices) {
      const hbtgpu_slice* src = c->d_out + U.out0;  // the serial walk writes densely
      if (!R.exact_path) {                          // the rows pass leaves unused slots between tiles
        if (int rc = grow(c, &c->d_tmp, &c->cap_tmp, R.n_slices)) return rc;
        const unsigned n_tiles = static_cast<unsigned>((U.K + NT - 1) / NT);
        k_islice_pack<<<n_tiles, 256, 0, c->stream>>>(U, c->d_ioff, c->d_rt_tot, c->d_rt_base, c->d_out, c->d_tmp);
        IK(cudaGetLastError());
        src = c->d_tmp;
      }
      IK(cudaMemcpyAsync(slices, src, R.n_slices * sizeof(hbtgpu_slice), cudaMemcpyDeviceToHost, c->stream));
    }
  }
  IK(cudaStreamSynchronize(c->stream));
  return 0;
}

int hbtgpu_islice_device_view(hbtgpu_islice_ctx* c, hbtgpu_islice_view* view) {
  if (!c || !view) return -EINVAL;
  if (!c->have) return ifail(c, -EINVAL, "no results: call hbtgpu_islice_release first");
  IK(cudaSetDevice(c->device));
  const size_t n = c->units.size();
  if (int rc = grow(c, &c->d_base, &c->cap_base, n)) return rc;
  std::vector<uint64_t> base(n);
  for (size_t u = 0; u < n; ++u) base[u] = c->units[u].out0;
  IK(cudaMemcpyAsync(c->d_base, base.data(), n * sizeof(uint64_t), cudaMemcpyHostToDevice, c->stream));
  IK(cudaStreamSynchronize(c->stream));
  view->slices = c->d_out;
  view->unit_base = c->d_base;
  view->interval_offset = c->d_ioff;
  view->interval_count = c->d_icnt;
  view->n_units = static_cast<uint32_t>(n);
  return 0;
}

#include "hbtgpu_islice_bin_host.inc"

}  // extern "C"
