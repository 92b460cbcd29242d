// hbtgpu_islice.cu -- count intervals x tagstack slices on the GPU.
//
// Reference: IntervalSlicer::makeIntervalIntersectingSlices (hbt/src/tagstack/IntervalSlicer.cpp:
// 52-188) called row by row from TraceMonitor::releaseCountData_ (hbt/src/mon/TraceMonitor.cpp:
// 52-96) with a consumer mark that advances by what each row consumed.  It is a serial walk:
// skip slices that end before the interval, take slices that start before its end, trim the
// first and the last, and the next row starts where this one stopped.
//
// Parallel form (one CTA per comp unit, a tile of NT rows per step).  Let t_j, e_j = t_j + d_j
// be start and end of slice j.  When the unit's slices are time ordered and disjoint
// (d_j > 0, t_j >= e_{j-1}: what tagstack::Slicer writes) and its rows move forward in time
// (d > 0, next tstamp >= this end: what CpuCountSampleGenerator writes), e and t are strictly
// increasing, so for the interval [T, Tend):
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
// which the reference does too).  So every row is evaluated for both starts, "does row k+1
// start one further" is a function {0,1} -> {0,1} of the same question for row k, and an
// inclusive scan under function composition settles all rows of a tile at once.  A prefix sum
// of the slice counts gives every row its place in the output, and the tile's slices are
// copied by all threads (one slice each per step, the owning row found by binary search in
// shared memory).
//
// Units that break the preconditions, or whose mark is ahead of the first row, are walked by
// one thread exactly as the reference does (exact_walk) and emitted by k_islice_emit_exact, so
// the results are the reference's on any input.  Scratch: G (8 B per slice).
#include <cuda_runtime.h>
#include <errno.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <string>
#include <vector>

#include "hbtgpu.h"

namespace {

constexpr int NT = 512;  // threads per CTA = rows per tile
constexpr int NW = NT / 32;
constexpr uint8_t TT_ANALYSIS = 1;  // Slice::TransitionType::Analysis, Slicer.h:36
constexpr uint64_t NONE = ~0ull;

struct Unit {
  uint64_t s0, ns;  // first slice, number of slices
  uint64_t r0, K;   // first row, number of rows
  uint64_t out0, out_cap;
  uint64_t mark;
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

// #{j in [0, n) : e_j < x}
__device__ __forceinline__ uint32_t count_end_lt(const hbtgpu_slice* S, uint32_t n, uint64_t x) {
  uint32_t lo = 0, hi = n;
  while (lo < hi) {
    const uint32_t mid = (lo + hi) >> 1;
    if (slice_end(S + mid) < x) lo = mid + 1;
    else hi = mid;
  }
  return lo;
}
// #{j : e_j <= x}
__device__ __forceinline__ uint32_t count_end_le(const hbtgpu_slice* S, uint32_t n, uint64_t x) {
  uint32_t lo = 0, hi = n;
  while (lo < hi) {
    const uint32_t mid = (lo + hi) >> 1;
    if (slice_end(S + mid) <= x) lo = mid + 1;
    else hi = mid;
  }
  return lo;
}
// #{j : t_j < x}
__device__ __forceinline__ uint32_t count_start_lt(const hbtgpu_slice* S, uint32_t n, uint64_t x) {
  uint32_t lo = 0, hi = n;
  while (lo < hi) {
    const uint32_t mid = (lo + hi) >> 1;
    if (S[mid].tstamp < x) lo = mid + 1;
    else hi = mid;
  }
  return lo;
}

// ---- block primitives (NT threads, all must call) ------------------------------------------
__device__ __forceinline__ uint64_t block_min_u64(uint64_t v, uint64_t* sh) {
  for (int d = 16; d; d >>= 1) v = min(v, __shfl_xor_sync(0xffffffffu, v, d));
  __syncthreads();
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
  __syncthreads();
  uint64_t r = sh[0];
  for (int w = 1; w < NW; ++w) r = min(r, sh[w]);
  return r;
}
__device__ __forceinline__ uint64_t block_max_u64(uint64_t v, uint64_t* sh) {
  for (int d = 16; d; d >>= 1) v = max(v, __shfl_xor_sync(0xffffffffu, v, d));
  __syncthreads();
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
  __syncthreads();
  uint64_t r = sh[0];
  for (int w = 1; w < NW; ++w) r = max(r, sh[w]);
  return r;
}
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
  uint64_t base = 0, tot = 0;
  for (int i = 0; i < NW; ++i) {
    if (i < w) base += sh[i];
    tot += sh[i];
  }
  *total = tot;
  return v + base;
}
// {0,1} -> {0,1} functions packed as f(0) | f(1) << 1; then(a, b)(v) = b(a(v))
__device__ __forceinline__ uint32_t fn_then(uint32_t a, uint32_t b) {
  return ((b >> (a & 1u)) & 1u) | (((b >> ((a >> 1) & 1u)) & 1u) << 1);
}
constexpr uint32_t FN_ID = 2u;
// EXCLUSIVE scan under composition: the composition of the functions of all earlier threads
__device__ __forceinline__ uint32_t block_excl_fn(uint32_t f, uint64_t* sh) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (int d = 1; d < 32; d <<= 1) {
    const uint32_t o = __shfl_up_sync(0xffffffffu, f, d);
    if (lane >= d) f = fn_then(o, f);
  }
  __syncthreads();
  if (lane == 31) sh[w] = f;
  __syncthreads();
  uint32_t base = FN_ID;
  for (int i = 0; i < w; ++i) base = fn_then(base, static_cast<uint32_t>(sh[i]));
  uint32_t ex = __shfl_up_sync(0xffffffffu, f, 1);
  if (lane == 0) ex = FN_ID;
  return fn_then(base, ex);
}

// One slice of an interval's answer: slice j, the i-th of cnt, trimmed as IntervalSlicer.cpp:
// 146-173 trims front() and back().
__device__ __forceinline__ void emit_one(const hbtgpu_slice* S, uint64_t j, uint32_t i, uint32_t cnt,
                                         uint64_t T, uint64_t Tend, bool trim, hbtgpu_slice* dst) {
  const uint4* src = reinterpret_cast<const uint4*>(S + j);
  union {
    uint4 q[2];
    hbtgpu_slice s;
  } u;
  u.q[0] = src[0];
  u.q[1] = src[1];
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

__global__ void __launch_bounds__(NT)
k_islice_unit(const hbtgpu_slice* __restrict__ slices, const uint64_t* __restrict__ rows_all, uint64_t stride,
              const Unit* __restrict__ units, hbtgpu_islice_unit* __restrict__ results, uint64_t* __restrict__ Gall,
              uint32_t* __restrict__ ioff_all, uint32_t* __restrict__ icnt_all, uint32_t* __restrict__ ifirst_all,
              hbtgpu_slice* __restrict__ out_all, uint64_t stop_ts, uint32_t flags) {
  __shared__ uint64_t sh_red[32];
  __shared__ uint64_t sh_T[NT], sh_Tend[NT];
  __shared__ uint32_t sh_excl[NT + 1], sh_p[NT], sh_lb[NT + 1];
  __shared__ uint32_t sh_flag;

  const int tid = threadIdx.x;
  const Unit U = units[blockIdx.x];
  const hbtgpu_slice* S = slices + U.s0;
  uint64_t* G = Gall + U.s0;
  const uint64_t* rows = rows_all + U.r0 * stride;
  uint32_t* ioff = ioff_all + U.r0;
  uint32_t* icnt = icnt_all + U.r0;
  uint32_t* ifirst = ifirst_all + U.r0;
  hbtgpu_slice* out = out_all + U.out0;
  hbtgpu_islice_unit* R = results + blockIdx.x;
  const bool trim = flags & HBTGPU_ISLICE_TRIM, abort_missing = flags & HBTGPU_ISLICE_ABORT_ON_MISSING;
  const uint32_t ns = static_cast<uint32_t>(U.ns);
  const uint64_t K = U.K;

  // ---- pass 1, rows: where the stop rule cuts, the first zero length, the first row that
  // wraps or whose successor starts before it ends
  uint64_t n_int = K, z = K, mw = K, mp = K;
  for (uint64_t k = tid; k < K; k += NT) {
    const Row rw = ld_row(rows, stride, k);
    const uint64_t Tend = rw.T + rw.D;
    if (Tend >= stop_ts) n_int = min(n_int, k);
    if (rw.D == 0) z = min(z, k);
    if (Tend < rw.T) mw = min(mw, k);
    if (k + 1 < K && rows[(k + 1) * stride] < Tend) mp = min(mp, k);
  }
  n_int = block_min_u64(n_int, sh_red);
  z = block_min_u64(z, sh_red);
  mw = block_min_u64(mw, sh_red);
  mp = block_min_u64(mp, sh_red);
  const uint64_t n_eff = min(n_int, z);  // rows the fast path may evaluate; row z (if < n_int) throws
  bool slow = (flags & HBTGPU_ISLICE_FORCE_EXACT) || mw < n_eff || mp + 1 < n_eff;

  // ---- pass 2, slices: order / overlap / zero length, and the gap prefix G
  {
    uint64_t carry = 0;
    int bad = 0;
    for (uint32_t base = 0; base < ns; base += NT) {
      const uint32_t j = base + tid;
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
      if (j < ns) G[j] = carry + incl;
      carry += tot;
    }
    if (__syncthreads_or(bad)) slow = true;
  }

  // ---- the mark must not be ahead of the first row's slices
  if (!slow && n_eff > 0) {
    if (tid == 0) {
      const Row rw = ld_row(rows, stride, 0);
      sh_flag = U.mark > count_end_le(S, ns, rw.T);
    }
    __syncthreads();
    if (sh_flag) slow = true;
  }
  if (slow) {
    if (tid == 0)
      exact_walk(S, ns, rows, stride, K, stop_ts, trim, abort_missing, U.mark, ioff, icnt, ifirst, R);
    return;
  }

  // ---- tiles of NT rows
  uint64_t mark = U.mark;      // uniform
  uint64_t run_off = 0;        // slices written so far, uniform
  uint64_t err_k = NONE;       // first row that throws, uniform
  if (z < n_int) err_k = z;
  for (uint64_t k0 = 0; k0 < n_eff && err_k > k0; k0 += NT) {
    const uint64_t k = k0 + tid;
    const bool live = k < n_eff;
    uint64_t T = 0, D = 0, Tend = 0;
    uint32_t lb = ns, x = ns, q = ns;
    bool hasB = false;
    if (live) {
      const Row rw = ld_row(rows, stride, k);
      T = rw.T;
      D = rw.D;
      Tend = T + D;
      lb = count_end_lt(S, ns, T);
      hasB = lb < ns && slice_end(S + lb) == T;
      q = count_start_lt(S, ns, Tend);
      x = trim ? q : count_end_le(S, ns, Tend);
    }
    __syncthreads();  // previous tile's readers of the shared arrays are done
    sh_lb[tid] = lb;
    if (tid == 0) sh_lb[NT] = ns;

    auto eval = [&](uint32_t p) {
      Eval e{true, false, p, p, 0};
      if (p >= ns) return e;
      const ulonglong2 tdp = ld_td(S + p);
      if (abort_missing && tdp.x > T) return e;
      const uint32_t r = max(p, x);
      if (r == p) return e;
      const uint64_t Gp = G[p];
      if (abort_missing) {
        // the walk compares every slice it looks at with its predecessor's end, including the
        // one the no-trim break stops on
        const uint32_t hi = trim ? r - 1 : min(r, max(p, q) - 1);
        if (G[hi] != Gp) return e;
      }
      const uint64_t er = slice_end(S + r - 1);
      if (abort_missing && er < Tend) return e;
      e.null_res = false;
      e.r = r;
      e.c = r - (er > Tend ? 1u : 0u);
      if (!trim) {
        const uint64_t sumdur = (er - tdp.x) - (G[r - 1] - Gp);
        if (sumdur > D) e.thrown = true;
      }
      return e;
    };
    Eval eA{true, false, 0, 0, 0}, eB{true, false, 0, 0, 0};
    if (live) {
      eA = eval(lb);
      eB = hasB ? eval(lb + 1) : eA;
    }
    __syncthreads();
    // does the next row start one past ITS lb?  Only if this row produced slices and its new
    // mark is beyond the next row's lb.
    const uint32_t lb_next = sh_lb[tid + 1];
    const bool last_in_tile = tid == NT - 1;
    uint32_t f = 0;
    if (live && !last_in_tile) {
      const uint32_t g0 = (!eA.null_res && !eA.thrown && eA.c > lb_next) ? 1u : 0u;
      const uint32_t g1 = (!eB.null_res && !eB.thrown && eB.c > lb_next) ? 1u : 0u;
      f = g0 | (g1 << 1);
    }
    const uint32_t b_first = (k0 < n_eff && mark > sh_lb[0]) ? 1u : 0u;
    const uint32_t ex = block_excl_fn(f, sh_red);
    const uint32_t b = (ex >> b_first) & 1u;
    const Eval e = b ? eB : eA;

    // first throwing row of the tile (rows after it do not count)
    uint64_t my_err = (live && e.thrown) ? k : NONE;
    my_err = block_min_u64(my_err, sh_red);
    if (my_err < err_k) err_k = my_err;
    const bool counted = live && k < err_k && !e.null_res;
    const uint32_t cnt = counted ? e.r - e.p : 0u;

    uint64_t tile_total;
    const uint64_t incl = block_incl_sum_u64(cnt, sh_red, &tile_total);
    const uint32_t excl = static_cast<uint32_t>(incl - cnt);
    sh_excl[tid] = excl;
    if (tid == 0) sh_excl[NT] = static_cast<uint32_t>(tile_total);
    sh_p[tid] = e.p;
    sh_T[tid] = T;
    sh_Tend[tid] = Tend;
    if (live && k < err_k) {
      ioff[k] = static_cast<uint32_t>(run_off + excl);
      icnt[k] = cnt;
      ifirst[k] = counted ? e.p : 0u;
    }
    const uint64_t new_mark = block_max_u64(counted ? static_cast<uint64_t>(e.c) : 0ull, sh_red);
    mark = max(mark, new_mark);
    __syncthreads();

    // copy the tile's slices: output slot -> owning row by binary search over sh_excl
    if (run_off + tile_total <= U.out_cap) {
      for (uint32_t o = tid; o < tile_total; o += NT) {
        uint32_t lo = 0, hi = NT;  // last idx with sh_excl[idx] <= o
        while (hi - lo > 1) {
          const uint32_t mid = (lo + hi) >> 1;
          if (sh_excl[mid] <= o) lo = mid;
          else hi = mid;
        }
        const uint32_t i = o - sh_excl[lo];
        const uint32_t c = sh_excl[lo + 1] - sh_excl[lo];
        emit_one(S, sh_p[lo] + i, i, c, sh_T[lo], sh_Tend[lo], trim, out + run_off + o);
      }
    } else {
      err_k = min(err_k, k0);  // cannot happen for valid fast-path input (bound ns + K); reported below
      if (tid == 0) R->pad = 1;
    }
    run_off += tile_total;
  }

  if (tid == 0) {
    R->n_intervals = err_k != NONE ? err_k : n_eff;
    R->n_slices = run_off;
    R->mark = mark;
    R->status = err_k != NONE ? -EINVAL : 0;
    R->err_interval = err_k;
    R->exact_path = 0;
  }
}

// Emission for units that took the serial walk: one warp per row, lanes over its slices.
__global__ void __launch_bounds__(256)
k_islice_emit_exact(const hbtgpu_slice* __restrict__ slices, const uint64_t* __restrict__ rows_all, uint64_t stride,
                    const Unit* __restrict__ units, const uint32_t* __restrict__ unit_ids,
                    const hbtgpu_islice_unit* __restrict__ results, const uint32_t* __restrict__ ioff_all,
                    const uint32_t* __restrict__ icnt_all, const uint32_t* __restrict__ ifirst_all,
                    hbtgpu_slice* __restrict__ out_all, uint32_t flags) {
  const uint32_t u = unit_ids[blockIdx.x];
  const Unit U = units[u];
  const hbtgpu_slice* S = slices + U.s0;
  const uint64_t* rows = rows_all + U.r0 * stride;
  const bool trim = flags & HBTGPU_ISLICE_TRIM;
  const uint64_t n = results[u].n_intervals;
  const int lane = threadIdx.x & 31;
  for (uint64_t k = threadIdx.x >> 5; k < n; k += blockDim.x >> 5) {
    const uint32_t cnt = icnt_all[U.r0 + k];
    if (cnt == 0) continue;
    const Row rw = ld_row(rows, stride, k);
    const uint64_t first = ifirst_all[U.r0 + k];
    hbtgpu_slice* dst = out_all + U.out0 + ioff_all[U.r0 + k];
    for (uint32_t i = lane; i < cnt; i += 32)
      emit_one(S, first + i, i, cnt, rw.T, rw.T + rw.D, trim, dst + i);
  }
}

}  // namespace

struct hbtgpu_islice_ctx {
  int device = 0;
  std::string err;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  float last_ms = 0.f;
  // device buffers (grown on demand)
  hbtgpu_slice* d_slices = nullptr;  size_t cap_slices = 0;
  uint64_t* d_rows = nullptr;        size_t cap_rows = 0;
  uint64_t* d_G = nullptr;           size_t cap_G = 0;
  uint32_t* d_ioff = nullptr;        size_t cap_ioff = 0;
  uint32_t* d_icnt = nullptr;        size_t cap_icnt = 0;
  uint32_t* d_ifirst = nullptr;      size_t cap_ifirst = 0;
  hbtgpu_slice* d_out = nullptr;     size_t cap_out = 0;
  Unit* d_units = nullptr;           size_t cap_units = 0;
  hbtgpu_islice_unit* d_res = nullptr; size_t cap_res = 0;
  uint32_t* d_ids = nullptr;         size_t cap_ids = 0;
  uint64_t* d_base = nullptr;        size_t cap_base = 0;
  // last call
  std::vector<Unit> units;
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

// grow *p to need elements; keep = elements to preserve
template <class T>
int grow(hbtgpu_islice_ctx* c, T** p, size_t* cap, size_t need, size_t keep = 0) {
  if (need <= *cap && *p) return 0;
  const size_t ncap = std::max<size_t>(std::max(need, *cap + *cap / 2), 16);
  T* q = nullptr;
  IK(cudaMalloc(&q, ncap * sizeof(T)));
  if (keep && *p) {
    cudaError_t e = cudaMemcpyAsync(q, *p, keep * sizeof(T), cudaMemcpyDeviceToDevice, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    if (e != cudaSuccess) {
      cudaFree(q);
      return ifail(c, -EIO, "device copy failed: %s", cudaGetErrorString(e));
    }
  }
  if (*p) cudaFree(*p);
  *p = q;
  *cap = ncap;
  return 0;
}

}  // namespace

extern "C" {

hbtgpu_islice_ctx* hbtgpu_islice_create(int device) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || device < 0 || device >= n) return nullptr;  // no CPU fallback
  if (cudaSetDevice(device) != cudaSuccess) return nullptr;
  auto* c = new hbtgpu_islice_ctx();
  c->device = device;
  if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess ||
      cudaEventCreate(&c->ev0) != cudaSuccess || cudaEventCreate(&c->ev1) != cudaSuccess) {
    delete c;
    return nullptr;
  }
  return c;
}

void hbtgpu_islice_destroy(hbtgpu_islice_ctx* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  cudaFree(c->d_slices); cudaFree(c->d_rows); cudaFree(c->d_G); cudaFree(c->d_ioff); cudaFree(c->d_icnt);
  cudaFree(c->d_ifirst); cudaFree(c->d_out); cudaFree(c->d_units); cudaFree(c->d_res); cudaFree(c->d_ids);
  cudaFree(c->d_base);
  if (c->ev0) cudaEventDestroy(c->ev0);
  if (c->ev1) cudaEventDestroy(c->ev1);
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
  if (!slice_begin || !row_begin || !marks || n_units == 0) return ifail(c, -EINVAL, "null argument or no units");
  if (row_stride < 2) return ifail(c, -EINVAL, "row_stride must be >= 2 (tstamp, duration, ...)");
  IK(cudaSetDevice(c->device));
  const uint64_t s_base = slice_begin[0], r_base = row_begin[0];
  const uint64_t total_s = slice_begin[n_units] - s_base, total_r = row_begin[n_units] - r_base;
  c->units.assign(n_units, Unit{});
  uint64_t out_run = 0;
  for (uint32_t u = 0; u < n_units; ++u) {
    if (slice_begin[u + 1] < slice_begin[u] || row_begin[u + 1] < row_begin[u])
      return ifail(c, -EINVAL, "begin arrays must not decrease (unit %u)", u);
    Unit& U = c->units[u];
    U.s0 = slice_begin[u] - s_base;
    U.ns = slice_begin[u + 1] - slice_begin[u];
    U.r0 = row_begin[u] - r_base;
    U.K = row_begin[u + 1] - row_begin[u];
    if (U.ns + U.K >= 0xffffffffull) return ifail(c, -EINVAL, "unit %u: more than 2^32 slices + rows", u);
    U.out0 = out_run;
    U.out_cap = U.ns + U.K;  // the parallel path writes at most one extra slice per row
    U.mark = marks[u];
    out_run += U.out_cap;
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
  if (int rc = grow(c, &c->d_ioff, &c->cap_ioff, total_r + 1)) return rc;
  if (int rc = grow(c, &c->d_icnt, &c->cap_icnt, total_r + 1)) return rc;
  if (int rc = grow(c, &c->d_ifirst, &c->cap_ifirst, total_r + 1)) return rc;
  if (int rc = grow(c, &c->d_out, &c->cap_out, out_run + 1)) return rc;
  if (int rc = grow(c, &c->d_units, &c->cap_units, n_units)) return rc;
  if (int rc = grow(c, &c->d_res, &c->cap_res, n_units)) return rc;
  IK(cudaMemcpyAsync(c->d_units, c->units.data(), n_units * sizeof(Unit), cudaMemcpyHostToDevice, c->stream));
  IK(cudaMemsetAsync(c->d_res, 0, n_units * sizeof(hbtgpu_islice_unit), c->stream));

  IK(cudaEventRecord(c->ev0, c->stream));
  k_islice_unit<<<n_units, NT, 0, c->stream>>>(d_slices, d_rows, stride, c->d_units, c->d_res, c->d_G, c->d_ioff,
                                               c->d_icnt, c->d_ifirst, c->d_out, stop_ts, flags);
  IK(cudaGetLastError());
  c->res.assign(n_units, hbtgpu_islice_unit{});
  IK(cudaMemcpyAsync(c->res.data(), c->d_res, n_units * sizeof(hbtgpu_islice_unit), cudaMemcpyDeviceToHost,
                     c->stream));
  IK(cudaStreamSynchronize(c->stream));

  // units that took the serial walk: place their slices (behind the regular region when they
  // do not fit their own) and copy them
  std::vector<uint32_t> ids;
  bool moved = false;
  for (uint32_t u = 0; u < n_units; ++u) {
    if (!c->res[u].exact_path) {
      if (c->res[u].pad) return ifail(c, -EIO, "internal: unit %u overflowed its output region", u);
      continue;
    }
    if (c->res[u].status == -EOVERFLOW) continue;
    if (c->res[u].n_slices == 0) continue;
    ids.push_back(u);
    if (c->res[u].n_slices > c->units[u].out_cap) {
      c->units[u].out0 = out_run;
      c->units[u].out_cap = c->res[u].n_slices;
      out_run += c->res[u].n_slices;
      moved = true;
    }
  }
  if (!ids.empty()) {
    if (moved) {
      if (int rc = grow(c, &c->d_out, &c->cap_out, out_run + 1, c->cap_out)) return rc;
      IK(cudaMemcpyAsync(c->d_units, c->units.data(), n_units * sizeof(Unit), cudaMemcpyHostToDevice, c->stream));
    }
    if (int rc = grow(c, &c->d_ids, &c->cap_ids, ids.size())) return rc;
    IK(cudaMemcpyAsync(c->d_ids, ids.data(), ids.size() * sizeof(uint32_t), cudaMemcpyHostToDevice, c->stream));
    k_islice_emit_exact<<<static_cast<unsigned>(ids.size()), 256, 0, c->stream>>>(
        d_slices, d_rows, stride, c->d_units, c->d_ids, c->d_res, c->d_ioff, c->d_icnt, c->d_ifirst, c->d_out, flags);
    IK(cudaGetLastError());
  }
  IK(cudaEventRecord(c->ev1, c->stream));
  IK(cudaStreamSynchronize(c->stream));
  IK(cudaEventElapsedTime(&c->last_ms, c->ev0, c->ev1));

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
    if (R.n_slices)
      IK(cudaMemcpyAsync(slices, c->d_out + U.out0, R.n_slices * sizeof(hbtgpu_slice), cudaMemcpyDeviceToHost,
                         c->stream));
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

}  // extern "C"
