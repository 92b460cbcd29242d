// hbtgpu_kernels.cuh -- sm_100a device code for the hbt perf_event ring decode +
// aggregate path.  See DESIGN.md for the algorithm; reference semantics:
//   record walk      hbt/src/perf_event/PerfEventsGroup.h:1394-1589 (consume)
//   sample handling  hbt/src/perf_event/PerCpuCountSampleGenerator.h:110-202
//   scaling/reducers hbt/src/perf_event/PerfEventsGroup.h:467-481,
//                    hbt/src/perf_event/BuiltinMetrics.cpp:88-118
//
// Kernel inventory
//   k_decode<SPEC,MODE>  one CTA per chunk of one image; tiles of 16 KiB are staged
//                        into shared memory with TMA bulk copies (double buffered);
//                        every 8-byte word is speculatively parsed as a record
//                        header; the true chain is recovered inside the tile
//                        (contiguity shortcut, else pointer jumping); samples are
//                        linked to their predecessor with a block scan and the
//                        per-(cpu,bucket) sums are formed as a telescoped segmented
//                        reduction (one RED.64 set per bucket boundary, not per row).
//                        SPEC=true: the chunk's entry point and predecessor sample
//                        are speculated from a 2 KiB halo before the chunk.
//   k_verify             proves every speculation (entry, predecessor, gap) equal to
//                        what the previous chunk actually produced; any mismatch or
//                        decode error raises NEED_EXACT.
//   k_resolve            (only if NEED_EXACT) warp-per-stream sequential walk that
//                        yields exact chunk entries and the first error per stream.
//   k_clear / k_decode<false,...>  (only if NEED_EXACT) recompute from exact entries.
//   k_derive             multiplex scaling + metric reducers per (cpu,bucket).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "hbtgpu.h"

namespace hbtgpu {

#ifndef HBT_NT
#define HBT_NT 192
#endif
constexpr int NT = HBT_NT;  // threads per CTA
constexpr int NWARP = NT / 32;
// A tile is sized per context so that the records of the configured group that start in it
// fill one pass of the uniform path exactly (PASS_RECORDS lanes, see uniform_segment): with a
// fixed 16 KiB tile a 104-byte record left 15 % of the lanes idle and a 72-byte record needed a
// second, mostly empty pass.  tile_bytes = min(TILE_MAX_BYTES, PASS_RECORDS * record bytes)
// rounded down to 16 (TMA).  TILE_MAX_BYTES keeps 5 CTAs per SM: 5 * (fixed 5.5 KiB +
// 2 * (19392 + 576) + 1 KiB reserved) <= 228 KiB.
constexpr int TILE_MAX_BYTES = 19392;
constexpr int PASS_RECORDS = (NT / 32) * 31 - 1;  // 185 for NT = 192
constexpr int OV_BYTES = 576;  // >= 40 + 8*64: any SAMPLE/READ/LOST whose header is in the tile is fully staged
__host__ __device__ inline uint32_t tile_bytes_for(uint32_t sample_words) {
  const uint64_t b = static_cast<uint64_t>(PASS_RECORDS) * sample_words * 8ull;
  return static_cast<uint32_t>((b < TILE_MAX_BYTES ? b : TILE_MAX_BYTES) & ~15ull);
}
constexpr int HALO_BYTES = NT * 8;  // one halo word per thread (scout_start, quick scout)
constexpr int HALO_WORDS = HALO_BYTES / 8;
// The general path (classify + pointer jumping + scan) works on sub-tiles of GT_WORDS words so
// that its per-candidate arrays stay small (shared memory decides how many CTAs fit on an SM).
#ifndef HBT_GT_WORDS
#define HBT_GT_WORDS 384
#endif
constexpr int GT_WORDS = HBT_GT_WORDS;
constexpr int MAXC = GT_WORDS;    // worst case: every word is a candidate header
constexpr int WPT = GT_WORDS / NT;  // words classified per thread (4)
static_assert(WPT * NWARP <= 32 && WPT >= 1, "classify() scans one (round, warp) count per lane");

constexpr uint16_t IDX_NONE = 0xFFFF;
constexpr uint16_t NX_EXIT = 0xFFFF;  // chain leaves the tile
constexpr uint16_t NX_DEAD = 0xFFFE;  // chain lands on a word that is no valid header
constexpr uint32_t BKT_OOR = 0xFFFFFFFFu;

constexpr uint64_t PRED_CARRY_IN = ~0ull - 1;  // predecessor is the stream's carry-in state
constexpr uint64_t PRED_UNKNOWN = ~0ull - 2;   // speculation could not find it

// record types (linux/perf_event.h), as switched on at PerfEventsGroup.h:1439
enum : uint32_t { REC_LOST = 2, REC_THROTTLE = 5, REC_UNTHROTTLE = 6, REC_READ = 8, REC_SAMPLE = 9 };

// per-candidate flags
enum : uint8_t { FL_REACH = 1, FL_SAMPLE = 2, FL_LOST = 4, FL_ROW = 8 };

// chunk flags
enum : uint32_t {
  CH_FIRST_OF_STREAM = 1,
  CH_FIRST_OF_IMAGE = 2,
  CH_LAST_OF_STREAM = 4,
  CH_LAST_OF_IMAGE = 8,
  CH_HALO_EXACT_START = 16,  // halo begins at an image start: its chain start is known
};

// device flag words
enum { FLAG_NEED_EXACT = 0, FLAG_INTERNAL = 1, FLAG_WHY = 2, FLAG_WHERE = 3, FLAG_COUNT = 8 };

// why the speculative pass was disproved (first reporter wins; diagnostics only)
enum : uint32_t {
  WHY_HALO_NO_START = 1, WHY_HALO_DEAD = 2, WHY_HALO_ERR = 3, WHY_HALO_NO_SAMPLE = 4,
  WHY_IMAGE_SEAM = 5, WHY_TILE_DEAD = 6, WHY_TILE_ERR = 7, WHY_PAST_END = 8,
  WHY_V_ENTRY = 10, WHY_V_PRED = 11, WHY_V_GAP = 12, WHY_V_IMAGE_END = 13, WHY_V_FIRST = 14
};

__device__ __forceinline__ void raise_exact(uint32_t* flags, uint32_t why, uint32_t where) {
  if (atomicCAS(&flags[FLAG_WHY], 0u, why) == 0u) flags[FLAG_WHERE] = where;
  atomicExch(&flags[FLAG_NEED_EXACT], 1u);
}

enum { MODE_AGG = 0, MODE_EMIT = 1 };

struct ChunkDesc {
  uint64_t img_base;    // arena offset of the image
  uint64_t img_len;     // bytes
  uint64_t begin, end;  // chunk range, image-relative bytes (begin multiple of the tile size)
  uint64_t stream_pos;  // stream byte offset of the image start
  uint64_t halo_off;    // arena offset (16-aligned) of the halo
  uint32_t halo_len;    // bytes (multiple of 8); 0: none
  uint32_t stream;
  uint32_t flags;
  uint32_t image;
};

struct ChunkIO {
  // what the chunk assumed on entry (SPEC: written by k_decode; exact: written by k_resolve)
  uint64_t in_entry;  // arena offset of the first record at/after chunk begin
  uint64_t in_pred;   // arena offset of the predecessor SAMPLE | PRED_CARRY_IN | PRED_UNKNOWN
  uint64_t limit;     // exact mode: decode nothing at/after this arena offset
  uint32_t in_gap;    // a LOST lies between predecessor and chunk
  uint32_t pad0;
  // what the chunk produced
  uint64_t out_exit;  // arena offset of the first record after the chunk
  uint64_t out_last;  // arena offset of the last SAMPLE up to chunk end (pass-through of in_pred if none)
  uint32_t out_gap;   // a LOST lies after out_last
  uint32_t n_records;
  uint32_t n_rows;
  uint32_t pad1;
  uint64_t row_base;  // MODE_EMIT: global row index of the chunk's first row
  uint64_t rec_base;  // MODE_EMIT: global record index of the chunk's first record
};

struct ImageDesc {
  uint64_t base, len, stream_pos;
  uint32_t stream, first_chunk, n_chunks, pad;
};

struct StreamDesc {
  uint32_t first_image, n_images;
};

// per-stream counters as u64 slots
enum {
  C_LOST = 0, C_THROTTLE, C_UNTHROTTLE, C_SAMPLE, C_READ, C_RECORDS, C_UNKNOWN, C_ROWS, C_OOR,
  C_ERR, C_ERR_OFF, C_LAST_TS, C_LAST_SAMPLE_TS, C_SLOTS = 16
};

// Shard mode: the records in front of a shard's first SAMPLE, as consume() accounts for them.
// Needed when that SAMPLE turns out to go back in time against the previous shard: the walk
// stalls there (PerCpuCountSampleGenerator.h:156-163) and only these records were consumed.
struct HeadExtra {
  uint64_t records, lost, throttle, unthrottle, read;
};

struct Params {
  const uint8_t* arena;
  const ChunkDesc* chunks;
  ChunkIO* io;
  const ImageDesc* images;
  const StreamDesc* streams;
  const hbtgpu_carry* carry_in;
  hbtgpu_carry* carry_out;
  hbtgpu_head_stub* head;
  struct HeadExtra* head_extra;  // shard mode: what the walk met before the first SAMPLE
  uint64_t* counters;  // [stream][C_SLOTS]
  uint64_t* acc;       // [stream][bucket][F]
  uint32_t* flags;
  uint64_t* rows;      // MODE_EMIT
  uint64_t* row_times; // MODE_EMIT: {tstamp, duration} of every row again, densely (what the interval slicer reads)
  uint64_t* rec_off;   // MODE_EMIT
  uint32_t* rec_type;
  uint32_t* rec_size;
  uint64_t t0, bucket_ns, tsc_mask, time_zero;
  uint32_t n_events, n_buckets, n_chunks, n_streams;
  uint32_t time_shift, time_mult;
  uint32_t force_general, external_carry;
  // SAMPLE layout: {hdr, [ip, pid|tid,] time, nr, time_enabled, time_running, value[nr]}.
  // pre = words between the header and `time` (0 for sample_type TIME|READ as hbt opens it,
  // 2 with IP|TID in front); fix = bytes of a SAMPLE without its values (40 + 8*pre).
  uint32_t pre, fix;
  uint32_t tile_bytes;    // tile_bytes_for(5 + pre + n_events)
  // IP|TID format only: open-addressing tables key -> sample count (key ~0 = empty slot; a
  // sample whose key is ~0 itself is counted in attr_special[0] / [1])
  uint64_t* tid_keys;
  unsigned long long* tid_cnt;
  uint64_t* ip_keys;
  unsigned long long* ip_cnt;
  unsigned long long* attr_special;  // [0] tid key ~0, [1] ip ~0, [2] tid overflow, [3] ip overflow
  uint32_t tid_mask, ip_mask;
  // bucket index = (ts - t0) / bucket_ns without a 64-bit division (set by the host)
  uint64_t range_ns;    // n_buckets * bucket_ns, saturated
  uint64_t bucket_magic;  // floor(2^64 / bucket_ns) (bucket_ns not a power of two)
  uint32_t bucket_shift;  // log2(bucket_ns) when it is a power of two, else 0xFFFFFFFF
  uint32_t bucket_f64;    // 1: bucket_ns < 2^32 and range_ns < 2^53, so d / bucket_ns may go through f64
  double bucket_inv;      // 1.0 / bucket_ns
  uint32_t s_magic;       // ceil(2^32 / (5 + n_events)): x / S == (x * s_magic) >> 32 for x < 2^16
};

// ------------------------------------------------------------------------------------
// small device helpers

__device__ __forceinline__ uint64_t umin64(uint64_t a, uint64_t b) { return a < b ? a : b; }

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  do {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
  } while (!ok);
}
// TMA 1-D bulk copy global -> shared, completion on an mbarrier (SASS: UBLKCP).
// The images are read exactly once, so their lines are marked evict-first in L2: what stays
// resident instead are the accumulator records the REDs update (zeroed just before by the
// memset), which saves their fetch from and a second write-back to DRAM.
__device__ __forceinline__ uint64_t l2_evict_first_policy() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ void tma_load_1d(void* dst, const void* src, uint32_t bytes,
                                            uint64_t* bar) {
#ifdef HBT_NO_L2_HINT
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
      ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
#else
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
      ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)), "l"(l2_evict_first_policy())
      : "memory");
#endif
}

// PerfEventsGroup::kernelTimeFromTsc, PerfEventsGroup.h:1645-1658
__device__ __forceinline__ uint64_t tsc_to_ktime(uint64_t tsc, uint32_t shift, uint32_t mult,
                                                 uint64_t zero) {
  const uint64_t rem_flags = (1ull << shift) - 1;  // tsc_rem_flags_, :1193
  const uint64_t quot = tsc >> shift;
  const uint64_t rem = tsc & rem_flags;
  return zero + quot * mult + ((rem * mult) >> shift);
}
// PerfEventsGroup::tscFromKernelTime, PerfEventsGroup.h:1661-1678
__device__ __forceinline__ uint64_t ktime_to_tsc(uint64_t kt, uint32_t shift, uint32_t mult,
                                                 uint64_t zero) {
  const uint64_t delta = kt - zero;
  const uint64_t quot = delta / mult;
  const uint64_t rem = delta % mult;
  return (quot << shift) + (rem << shift) / mult;
}
// GroupReadValues::getCount, PerfEventsGroup.h:467-481.  Explicit rounding intrinsics so
// that nvcc can neither contract nor reorder: u64->f64 RN, mul RN, div RN, f64->u64 RZ.
__device__ __forceinline__ uint64_t get_count(uint64_t count, uint64_t te, uint64_t tr) {
  if (te == 0) return 0;
  if (tr == 0) return 0;
  const double c = __ull2double_rn(count), e = __ull2double_rn(te), r = __ull2double_rn(tr);
  const double q = __ddiv_rn(__dmul_rn(c, e), r);
  // q >= 2^64 is undefined behaviour in the reference's static_cast<uint64_t>; its x86-64
  // build (cvttsd2si on q - 2^63, xor the sign bit) yields 0 there, and so do we.
  if (!(q < 18446744073709551616.0)) return 0;
  return __double2ull_rz(q);
}

// ------------------------------------------------------------------------------------
// shared memory of k_decode

struct Carry {
  uint64_t ts, te, tr;                // previous SAMPLE: time, time_enabled, time_running
  uint64_t last[HBTGPU_MAX_EVENTS];   // ... and its (transformed) values
  uint64_t last_off;                  // arena offset of that SAMPLE (or in_pred)
  uint32_t valid, gap;                // a predecessor exists / a LOST came after it
};

struct __align__(128) Smem {
  uint16_t pos[MAXC];             // candidate word positions, ascending
  union {
    struct {                      // chain phase
      uint16_t jump[2][MAXC];
      uint16_t map[GT_WORDS];     // word -> candidate index
    } ch;
    struct {                      // decode phase
      uint32_t bkt[MAXC];
      uint16_t pred[MAXC];
    } de;
  } u;
  uint16_t next[MAXC];
  uint8_t fl[MAXC];
  // carry = generator state up to the current tile.  Two copies: the uniform fast path reads
  // one and writes the other, which saves a block barrier per tile.
  Carry c[2];
  // scratch
  uint64_t bar[2];
  uint32_t wcnt[WPT * NWARP + 1];
  uint32_t scan_a[NWARP], scan_b[NWARP], scan_c[NWARP];
  uint32_t e_idx, term_tgt, term_kind, red_key;
  uint32_t idx_n, idx_end, idx_gap, idx_bad, idx_other;  // index_tile: samples listed, chain position after the tile,
                                              // a LOST after the last sample, word of a record consume() stalls on
  uint32_t tile_err;
  uint32_t ix[NWARP][8];          // index_tile: what each warp's walk found (IX_*)
  uint32_t cnt[C_SLOTS];          // chunk-local counters (u32), lost is u64 below
  unsigned long long lost_sum;
};

__device__ __forceinline__ bool is_header(uint64_t w) {
  const uint32_t hi = static_cast<uint32_t>(w >> 32);
  // size = hi >> 16: valid iff >= 8 and a multiple of 8
  return (hi & 0xFFF80000u) != 0 && (hi & 0x00070000u) == 0;
}
__device__ __forceinline__ uint32_t rec_words(uint64_t w) { return static_cast<uint32_t>(w >> 51); }

// exclusive block scan of (a: per-half max via vmaxu2, b: sum).  Returns exclusive prefix of
// this thread and block totals through tot_a / tot_b.  Contains two __syncthreads.
__device__ __forceinline__ void block_scan(Smem& s, uint32_t a, uint32_t b, uint32_t& ex_a,
                                           uint32_t& ex_b, uint32_t& tot_a, uint32_t& tot_b) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t ia = a, ib = b;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const uint32_t ya = __shfl_up_sync(0xffffffffu, ia, d);
    const uint32_t yb = __shfl_up_sync(0xffffffffu, ib, d);
    if (lane >= d) {
      ia = __vmaxu2(ia, ya);
      ib += yb;
    }
  }
  if (lane == 31) {
    s.scan_a[warp] = ia;
    s.scan_b[warp] = ib;
  }
  __syncthreads();
  uint32_t wa = 0, wb = 0, ta = 0, tb = 0;
#pragma unroll
  for (int w = 0; w < NWARP; ++w) {
    const uint32_t xa = s.scan_a[w], xb = s.scan_b[w];
    if (w < warp) {
      wa = __vmaxu2(wa, xa);
      wb += xb;
    }
    ta = __vmaxu2(ta, xa);
    tb += xb;
  }
  uint32_t pa = __shfl_up_sync(0xffffffffu, ia, 1);
  uint32_t pb = __shfl_up_sync(0xffffffffu, ib, 1);
  if (lane == 0) {
    pa = 0;
    pb = 0;
  }
  ex_a = __vmaxu2(wa, pa);
  ex_b = wb + pb;
  tot_a = ta;
  tot_b = tb;
  __syncthreads();
}

// Classify words [0, wt) of `buf` as candidate headers and compact their positions in
// ascending order into s.pos.  Returns K.  Contains __syncthreads.
__device__ __forceinline__ uint32_t classify(Smem& s, const uint64_t* buf, uint32_t wt) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t ballots[WPT];
#pragma unroll
  for (int i = 0; i < WPT; ++i) {
    const uint32_t w = i * NT + threadIdx.x;
    const bool c = (w < wt) && is_header(buf[w]);
    ballots[i] = __ballot_sync(0xffffffffu, c);
    if (lane == 0) s.wcnt[i * NWARP + warp] = __popc(ballots[i]);
  }
  __syncthreads();
  if (warp == 0) {
    // exclusive scan of the (up to 32) counts in (round, warp) order: one per lane
    const uint32_t v = lane < WPT * NWARP ? s.wcnt[lane] : 0;
    uint32_t incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t y = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += y;
    }
    if (lane < WPT * NWARP) s.wcnt[lane] = incl - v;
    if (lane == 31) s.wcnt[WPT * NWARP] = incl;
  }
  __syncthreads();
  const uint32_t lt = (1u << lane) - 1;
#pragma unroll
  for (int i = 0; i < WPT; ++i) {
    if ((ballots[i] >> lane) & 1u) {
      const uint32_t idx = s.wcnt[i * NWARP + warp] + __popc(ballots[i] & lt);
      s.pos[idx] = static_cast<uint16_t>(i * NT + threadIdx.x);
    }
  }
  const uint32_t K = s.wcnt[WPT * NWARP];
  __syncthreads();
  return K;
}

// Build word->candidate map and the successor of every candidate (into s.u.ch.jump[0]).
__device__ __forceinline__ void build_links(Smem& s, const uint64_t* buf, uint32_t wt, uint32_t K) {
  for (uint32_t w = threadIdx.x; w < wt; w += NT) s.u.ch.map[w] = IDX_NONE;
  __syncthreads();
  for (uint32_t j = threadIdx.x; j < K; j += NT) s.u.ch.map[s.pos[j]] = static_cast<uint16_t>(j);
  __syncthreads();
  for (uint32_t j = threadIdx.x; j < K; j += NT) {
    const uint32_t p = s.pos[j];
    const uint32_t tgt = p + rec_words(buf[p]);
    uint16_t nx;
    if (tgt >= wt) {
      nx = NX_EXIT;
    } else {
      const uint16_t m = s.u.ch.map[tgt];
      nx = (m == IDX_NONE) ? NX_DEAD : m;
    }
    s.u.ch.jump[0][j] = nx;
  }
  __syncthreads();
}

// Recover the chain that starts at word `entry` inside a tile of `wt` valid words with K
// candidates: sets FL_REACH in s.fl[j] for the candidates on it, and reports its end:
//   s.term_kind = 0: leaves the tile, next header at word s.term_tgt (>= wt)
//   s.term_kind = 1: lands on an invalid header at word s.term_tgt (or the entry itself is none)
// Contiguity shortcut first; pointer jumping otherwise.
__device__ __forceinline__ void recover_chain(Smem& s, const uint64_t* buf, uint32_t wt, uint32_t K,
                                              uint32_t entry, bool force_general) {
  if (threadIdx.x == 0) {
    s.e_idx = IDX_NONE;
    s.term_kind = 1;
    s.term_tgt = entry;
  }
  __syncthreads();
  for (uint32_t j = threadIdx.x; j < K; j += NT) {
    s.fl[j] = 0;
    if (s.pos[j] == entry) s.e_idx = j;
  }
  __syncthreads();
  const uint32_t e = s.e_idx;
  if (e == IDX_NONE) return;  // entry is not a valid header: term_kind=1 at entry

  // Shortcut: from e on, every candidate's record ends exactly where the next candidate
  // begins (and the last one leaves the tile).  Then the chain is pos[e..K).
  int good = 1;
  if (!force_general) {
    for (uint32_t j = e + threadIdx.x; j < K; j += NT) {
      const uint32_t p = s.pos[j];
      const uint32_t tgt = p + rec_words(buf[p]);
      const uint32_t nxt_pos = (j + 1 < K) ? s.pos[j + 1] : 0xFFFFFFFFu;
      if (j + 1 < K ? (tgt != nxt_pos) : (tgt < wt)) good = 0;
    }
  } else {
    good = 0;
  }
  good = __syncthreads_and(good);
  if (good) {
    for (uint32_t j = e + threadIdx.x; j < K; j += NT) s.fl[j] = FL_REACH;
    if (threadIdx.x == 0) {
      const uint32_t p = s.pos[K - 1];
      s.term_kind = 0;
      s.term_tgt = p + rec_words(buf[p]);
    }
    __syncthreads();
    return;
  }

  // General: pointer jumping.  After round r every candidate within 2^(r+1) hops of e is marked.
  build_links(s, buf, wt, K);
  if (threadIdx.x == 0) s.fl[e] = FL_REACH;
  __syncthreads();
  int cur = 0;
  for (;;) {
    int marked = 0;
    for (uint32_t j = threadIdx.x; j < K; j += NT) {
      const uint16_t jn = s.u.ch.jump[cur][j];
      if (s.fl[j] && jn < NX_DEAD && !s.fl[jn]) {
        s.fl[jn] = FL_REACH;
        marked = 1;
      }
    }
    marked = __syncthreads_or(marked);
    if (!marked) break;
    for (uint32_t j = threadIdx.x; j < K; j += NT) {
      const uint16_t jn = s.u.ch.jump[cur][j];
      s.u.ch.jump[cur ^ 1][j] = (jn < NX_DEAD) ? s.u.ch.jump[cur][jn] : jn;
    }
    cur ^= 1;
    __syncthreads();
  }
  // the unique reachable candidate whose successor is a sentinel ends the chain
  for (uint32_t j = threadIdx.x; j < K; j += NT) {
    if (s.fl[j]) {
      const uint32_t p = s.pos[j];
      const uint32_t tgt = p + rec_words(buf[p]);
      if (tgt >= wt) {
        s.term_kind = 0;
        s.term_tgt = tgt;
      } else if (s.u.ch.map[tgt] == IDX_NONE) {
        s.term_kind = 1;
        s.term_tgt = tgt;
      }
    }
  }
  __syncthreads();
}

// For the halo: choose the candidate whose chain (a) leaves the halo, (b) holds the most
// records that look exactly like what this capture contains (a SAMPLE of the configured group,
// a 24-byte LOST, a 32-byte THROTTLE/UNTHROTTLE) and (c) is longest.  Criterion (b) matters for
// counters whose values happen to carry a plausible size in their top 16 bits: such a counter
// forms a false chain in lock step with the true one, equally long, but made of records of
// nonsense types.  A wrong pick is only a lost bet: k_verify refutes it.
// K <= HALO_WORDS <= NT so one candidate per thread.  Returns the word position or ~0.
template <int PRE>
__device__ __forceinline__ uint32_t scout_start(Smem& s, const Params& P, const uint64_t* buf,
                                                uint32_t wt, uint32_t K) {
  build_links(s, buf, wt, K);
  const uint32_t j = threadIdx.x;
  uint32_t t = 0, h = 0, own = 0;  // h, own: (plausible records << 16) | hops
  if (j < K) {
    const uint32_t p = s.pos[j];
    const uint64_t hd = buf[p];
    const uint32_t type = static_cast<uint32_t>(hd), size = static_cast<uint32_t>(hd >> 48);
    const uint32_t n = P.n_events;
    const bool plausible = (type == REC_SAMPLE && size == (40 + 8 * PRE) + 8 * n && buf[p + PRE + 2] == n) ||
                           (type == REC_LOST && size == 24) ||
                           ((type == REC_THROTTLE || type == REC_UNTHROTTLE) && size == 32);
    own = (plausible ? 0x10000u : 0u) | 1u;
    const uint16_t nx = s.u.ch.jump[0][j];
    t = (nx >= NX_DEAD) ? j : nx;
    h = (nx >= NX_DEAD) ? 0 : own;  // weight of [j, t[j]); the terminal's own weight is added last
  }
  // t[j] -> terminal candidate of j's chain (jump[1] holds t, bkt holds h)
  for (int r = 0; r < 9; ++r) {
    if (j < K) {
      s.u.ch.jump[1][j] = static_cast<uint16_t>(t);
      s.next[j] = static_cast<uint16_t>(h & 0xFFFFu);
      s.fl[j] = static_cast<uint8_t>(h >> 16);  // plausible count < 256
    }
    __syncthreads();
    uint32_t t2 = t, h2 = 0;
    if (j < K) {
      t2 = s.u.ch.jump[1][t];
      h2 = s.next[t] | (static_cast<uint32_t>(s.fl[t]) << 16);
    }
    __syncthreads();
    t = t2;
    h += h2;
  }
  // add the terminal's own weight: publish `own`, read the terminal's
  if (j < K) {
    s.next[j] = static_cast<uint16_t>(own & 0xFFFFu);
    s.fl[j] = static_cast<uint8_t>(own >> 16);
  }
  if (threadIdx.x == 0) s.red_key = 0;
  __syncthreads();
  if (j < K && s.u.ch.jump[0][t] == NX_EXIT) {
    h += s.next[t] | (static_cast<uint32_t>(s.fl[t]) << 16);
    const uint32_t plaus = h >> 16, hops = h & 0xFFFFu;
    // most plausible records, then longest chain, then earliest start
    const uint32_t key = (min(plaus, 1023u) << 20) | (min(hops, 1023u) << 10) | (1023u - j);
    atomicMax(&s.red_key, key + 1);
  }
  __syncthreads();
  const uint32_t key = s.red_key;
  __syncthreads();
  if (key == 0) return 0xFFFFFFFFu;
  return s.pos[1023u - ((key - 1) & 1023u)];
}

__device__ __forceinline__ uint32_t bucket_of(const Params& P, uint64_t ts) {
  const uint64_t d = ts - P.t0;                // wraps to a huge value when ts < t0 ...
  if (d >= P.range_ns) return BKT_OOR;         // ... which this catches too (range_ns <= 2^64 - 1)
  if (P.bucket_shift != 0xFFFFFFFFu) return static_cast<uint32_t>(d >> P.bucket_shift);
  if (P.bucket_f64) {
    // d < 2^53 converts exactly; the product is within 2^-20 of d / bucket_ns (two roundings of
    // 2^-53 relative, quotient < 2^32), so its floor is the quotient or one off: fix it up
    // exactly with integers.  Half the instructions of the 64-bit multiply-high below.
    uint32_t q32 = __double2uint_rz(__dmul_rn(__ull2double_rz(d), P.bucket_inv));
    const uint64_t t = static_cast<uint64_t>(q32) * static_cast<uint32_t>(P.bucket_ns);
    if (t > d) --q32;
    else if (d - t >= P.bucket_ns) ++q32;
    return q32;
  }
  // d / bucket_ns by multiply-high with floor(2^64 / bucket_ns): one below the quotient at most
  uint64_t q = __umul64hi(d, P.bucket_magic);
  if (d - q * P.bucket_ns >= P.bucket_ns) ++q;
  return static_cast<uint32_t>(q);
}

__device__ __forceinline__ uint64_t xform(const Params& P, uint32_t e, uint64_t v) {
  if (P.tsc_mask == 0) return v;  // uniform: no event is transformed (the usual case)
  return ((P.tsc_mask >> e) & 1ull) ? tsc_to_ktime(v, P.time_shift, P.time_mult, P.time_zero) : v;
}

__device__ __forceinline__ void red_add(uint64_t* p, uint64_t v) {
#ifdef HBT_DEBUG_NO_RED
  if (v == 0x123456789ull)
#endif
  atomicAdd(reinterpret_cast<unsigned long long*>(p), static_cast<unsigned long long>(v));
}

constexpr uint64_t ATTR_EMPTY = ~0ull;

// 32-bit arithmetic only: the decode kernel is short of issue slots and a 64-bit multiply is
// four instructions.  attr_hash0 alone places a key in the per-CTA cache.
__device__ __forceinline__ uint32_t attr_hash0(uint64_t k) {
  return static_cast<uint32_t>(k) * 0x9E3779B1u ^ static_cast<uint32_t>(k >> 32) * 0x85EBCA6Bu;
}
__device__ __forceinline__ uint32_t attr_hash(uint64_t k) {
  uint32_t x = attr_hash0(k);
  x ^= x >> 15;
  x *= 0x2C1B3C6Du;
  x ^= x >> 12;
  x *= 0x297A2D39u;
  x ^= x >> 15;
  return x;
}

// count += c for `key` (!= ATTR_EMPTY) starting at its home slot, whose key word the caller has
// already loaded (so that several lookups can be in flight at once)
__device__ __forceinline__ void attr_add_from(uint64_t* keys, unsigned long long* cnt, uint32_t mask, uint64_t key,
                                              unsigned long long c, unsigned long long* overflow, uint32_t slot,
                                              uint64_t k) {
  for (uint32_t probe = 0; probe <= mask; ++probe) {
    if (probe) k = reinterpret_cast<volatile uint64_t*>(keys)[slot];
    if (k == ATTR_EMPTY) {
      k = atomicCAS(reinterpret_cast<unsigned long long*>(keys + slot), ATTR_EMPTY, key);
      if (k == ATTR_EMPTY) k = key;
    }
    if (k == key) {
      atomicAdd(cnt + slot, c);
      return;
    }
    slot = (slot + 1) & mask;
  }
  atomicAdd(overflow, c);  // table full: reported, never silently dropped
}
// count += c for `key` in an open-addressing table (linear probing, insert on first use)
__device__ __forceinline__ void attr_add(uint64_t* keys, unsigned long long* cnt, uint32_t mask, uint64_t key,
                                         unsigned long long c, unsigned long long* special,
                                         unsigned long long* overflow) {
  if (key == ATTR_EMPTY) {
    atomicAdd(special, c);
    return;
  }
  const uint32_t slot = attr_hash(key) & mask;
  attr_add_from(keys, cnt, mask, key, c, overflow, slot, reinterpret_cast<volatile uint64_t*>(keys)[slot]);
}

// Per-CTA front of the two attribution tables (AGG kernels of the IP|TID layout only; it sits
// behind Smem in dynamic shared memory, which leaves 4 CTAs per SM).  A chunk's samples are
// counted here first and the distinct keys are flushed once at the end of the chunk, so a hot
// thread or a hot instruction costs one global atomic per chunk instead of one per sample.
// Keys that find no slot in two probes go straight to the global table.  The decode kernel is
// close to issue-bound, so this path is kept short: a multiplicative hash, two unrolled
// probes, no warp-level matching (shared-memory atomics resolve duplicates within a warp).
constexpr uint32_t ATTR_C_TID_LOG2 = 8, ATTR_C_IP_LOG2 = 10;
struct AttrCache {
  uint64_t ip_key[1u << ATTR_C_IP_LOG2];
  uint64_t tid_key[1u << ATTR_C_TID_LOG2];  // pid | tid << 32
  uint32_t ip_cnt[1u << ATTR_C_IP_LOG2];
  uint32_t tid_cnt[1u << ATTR_C_TID_LOG2];
};
constexpr size_t ATTR_CACHE_OFF = (sizeof(Smem) + 127) & ~static_cast<size_t>(127);
constexpr size_t ATTR_SMEM_BYTES = ATTR_CACHE_OFF + sizeof(AttrCache);
__device__ __forceinline__ AttrCache& attr_cache() {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  return *reinterpret_cast<AttrCache*>(smem_raw + ATTR_CACHE_OFF);
}
template <bool IP>
__device__ __forceinline__ void attr_global(const Params& P, uint64_t key, unsigned long long c) {
  if (IP) attr_add(P.ip_keys, P.ip_cnt, P.ip_mask, key, c, P.attr_special + 1, P.attr_special + 3);
  else attr_add(P.tid_keys, P.tid_cnt, P.tid_mask, key, c, P.attr_special + 0, P.attr_special + 2);
}
// true if the per-CTA cache took the sample; false: the caller owes the global table one count
template <bool IP>
__device__ __forceinline__ bool attr_cache_try(uint64_t key) {
  constexpr uint32_t LOG2 = IP ? ATTR_C_IP_LOG2 : ATTR_C_TID_LOG2;
  AttrCache& A = attr_cache();
  uint64_t* const keys = IP ? A.ip_key : A.tid_key;
  uint32_t* const cnts = IP ? A.ip_cnt : A.tid_cnt;
  if (key == ATTR_EMPTY) return false;
  uint32_t slot = attr_hash0(key) >> (32 - LOG2);
#pragma unroll
  for (uint32_t probe = 0; probe < 2; ++probe) {
    uint64_t k = reinterpret_cast<volatile uint64_t*>(keys)[slot];
    if (k == ATTR_EMPTY) {
      k = atomicCAS(reinterpret_cast<unsigned long long*>(keys + slot), ATTR_EMPTY, key);
      if (k == ATTR_EMPTY) k = key;
    }
    if (k == key) {
      atomicAdd(cnts + slot, 1u);
      return true;
    }
    slot = (slot + 1) & ((1u << LOG2) - 1);
  }
  return false;
}
template <bool IP>
__device__ __forceinline__ void attr_count(const Params& P, uint64_t key) {
  if (!attr_cache_try<IP>(key)) attr_global<IP>(P, key, 1);
}
__device__ __forceinline__ void attr_cache_init() {
  AttrCache& A = attr_cache();
  for (uint32_t i = threadIdx.x; i < (1u << ATTR_C_IP_LOG2); i += NT) {
    A.ip_key[i] = ATTR_EMPTY;
    A.ip_cnt[i] = 0;
  }
  for (uint32_t i = threadIdx.x; i < (1u << ATTR_C_TID_LOG2); i += NT) {
    A.tid_key[i] = ATTR_EMPTY;
    A.tid_cnt[i] = 0;
  }
}
// after a __syncthreads: every distinct key of the chunk goes to the global tables once
__device__ __forceinline__ void attr_cache_flush(const Params& P) {
  AttrCache& A = attr_cache();
  for (uint32_t i = threadIdx.x; i < (1u << ATTR_C_IP_LOG2); i += NT)
    if (A.ip_key[i] != ATTR_EMPTY) attr_global<true>(P, A.ip_key[i], A.ip_cnt[i]);
  for (uint32_t i = threadIdx.x; i < (1u << ATTR_C_TID_LOG2); i += NT)
    if (A.tid_key[i] != ATTR_EMPTY) attr_global<false>(P, A.tid_key[i], A.tid_cnt[i]);
}

struct TileCtx {
  uint64_t tile_abs;    // arena offset of word 0 of the buffer
  uint64_t stream_off;  // stream byte offset of word 0 of the buffer
  uint32_t stream;
  uint32_t rows_before, recs_before;  // MODE_EMIT: already emitted in this chunk
  uint64_t row_base, rec_base;
  uint64_t* acc_s;      // MODE_AGG: accumulator records of this stream (computed once per chunk)
};

// Decode the reachable records of one staged tile and fold them into the carry, the
// counters and (if AGG) the bucket sums.  SCOUT: only the carry is updated.
template <bool SCOUT, int MODE, int PRE>
__device__ __forceinline__ void decode_tile(Smem& s, const Params& P, const uint64_t* buf,
                                            uint32_t K, TileCtx& tc, uint32_t cur) {
  Carry& C = s.c[cur];
  const uint32_t n = P.n_events;
  const uint32_t ipt = (K + NT - 1) / NT;  // candidates per thread (blocked)
  const uint32_t j0 = threadIdx.x * ipt;
  const uint32_t j1 = min(j0 + ipt, K);

  // pass 0: type checks, counters, scan inputs
  uint32_t a = 0, b = 0;
  uint32_t n_thr = 0, n_unthr = 0, n_read = 0, n_rec = 0;
  unsigned long long lost = 0;
  for (uint32_t j = j0; j < j1; ++j) {
    uint8_t f = s.fl[j];
    if (f & FL_REACH) {
      const uint32_t p = s.pos[j];
      const uint64_t h = buf[p];
      const uint32_t type = static_cast<uint32_t>(h);
      const uint32_t size = static_cast<uint32_t>(h >> 48);
      ++n_rec;
      if (type == REC_SAMPLE) {
        // recordSize check (:1502-1507) + nr == kNumEvents (fenced DCHECK, gen :133)
        const uint64_t nr = (size + 16 >= (40u + 8u * PRE)) ? buf[p + PRE + 2] : ~0ull;
        if (size < (40u + 8u * PRE) || nr != ((size - (40u + 8u * PRE)) >> 3) || nr != n) {
          atomicMin(&s.tile_err, p);
        } else {
          f |= FL_SAMPLE;
          a = __vmaxu2(a, (j + 1) << 16);
          b += 1;
          s.next[j] = IDX_NONE;
          if (!SCOUT && MODE == MODE_AGG && PRE == 2) {
            attr_count<true>(P, buf[p + 1]);
            attr_count<false>(P, buf[p + 2]);
          }
        }
      } else if (type == REC_LOST) {
        if (size < 24) {
          atomicMin(&s.tile_err, p);
        } else {
          f |= FL_LOST;
          a = __vmaxu2(a, j + 1);
          lost += buf[p + 2];  // num_lost (:1448)
        }
      } else if (type == REC_READ) {
        const uint64_t nr = (size >= 24) ? buf[p + 2] : ~0ull;
        if (size < 40 || nr != ((size - 40u) >> 3)) atomicMin(&s.tile_err, p);  // :1518-1523
        else ++n_read;
      } else if (type == REC_THROTTLE) {
        ++n_thr;
      } else if (type == REC_UNTHROTTLE) {
        ++n_unthr;
      }
      s.fl[j] = f;
    }
  }
  uint32_t ex_a, ex_b, tot_a, tot_b;
  block_scan(s, a, b, ex_a, ex_b, tot_a, tot_b);  // also orders the fl/next writes above

  // pass 1: link every sample to its predecessor
  const bool c_valid = C.valid != 0, c_gap = C.gap != 0;
  const uint64_t c_ts = C.ts;
  uint32_t run_a = ex_a, run_b = ex_b;
  uint32_t n_rows = 0, n_unknown = 0, n_sample = 0;
  for (uint32_t j = j0; j < j1; ++j) {
    const uint8_t f = s.fl[j];
    if (f & FL_SAMPLE) {
      const uint32_t prev_s = run_a >> 16, prev_l = run_a & 0xFFFFu;
      const bool in_tile = prev_s != 0;
      const bool gap = in_tile ? (prev_l > prev_s) : (c_gap || prev_l != 0);
      const bool link = (in_tile || c_valid) && !gap;
      const uint32_t p = s.pos[j];
      const uint64_t ts = buf[p + PRE + 1];
      const uint64_t pts = in_tile ? buf[s.pos[prev_s - 1] + PRE + 1] : c_ts;
      if (link && pts > ts) atomicMin(&s.tile_err, p);  // time regression (gen :156-163)
      if (!SCOUT) {
        s.u.de.bkt[j] = bucket_of(P, ts);
        s.u.de.pred[j] = in_tile ? static_cast<uint16_t>(prev_s - 1) : IDX_NONE;
        if (in_tile) s.next[prev_s - 1] = static_cast<uint16_t>(j);
        if (link) {
          s.fl[j] = f | FL_ROW;
          ++n_rows;
        } else {
          ++n_unknown;
        }
        ++n_sample;
      }
      run_a = __vmaxu2(run_a, (j + 1) << 16);
      run_b += 1;
    } else if (f & FL_LOST) {
      run_a = __vmaxu2(run_a, j + 1);
    }
  }
  if (!SCOUT) {
    // counters: warp-aggregate then shared atomics
    const uint32_t vals[7] = {n_thr, n_unthr, n_sample, n_read, n_rec, n_unknown, n_rows};
    const int slots[7] = {C_THROTTLE, C_UNTHROTTLE, C_SAMPLE, C_READ, C_RECORDS, C_UNKNOWN, C_ROWS};
#pragma unroll
    for (int k = 0; k < 7; ++k) {
      const uint32_t v = __reduce_add_sync(0xffffffffu, vals[k]);
      if ((threadIdx.x & 31) == 0 && v) atomicAdd(&s.cnt[slots[k]], v);
    }
    if (__any_sync(0xffffffffu, lost != 0)) {
      unsigned long long l = lost;
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) l += __shfl_xor_sync(0xffffffffu, l, d);
      if ((threadIdx.x & 31) == 0) atomicAdd(&s.lost_sum, l);
    }
  }
  __syncthreads();

  // MODE_EMIT needs the rank of each row / record inside the tile
  uint32_t ex_row = 0, ex_rec = 0;
  if (!SCOUT && MODE == MODE_EMIT) {
    uint32_t r = 0, q = 0;
    for (uint32_t j = j0; j < j1; ++j) {
      r += (s.fl[j] & FL_ROW) ? 1u : 0u;
      q += (s.fl[j] & FL_REACH) ? 1u : 0u;
    }
    uint32_t d0, d1, t0_, t1_;
    block_scan(s, 0, r | (q << 16), d0, d1, t0_, t1_);
    (void)d0;
    (void)t0_;
    ex_row = d1 & 0xFFFFu;
    ex_rec = d1 >> 16;
    if (threadIdx.x == 0) {
      // totals for the next tile (t1_ packs rows | records<<16; both <= 2048)
      s.scan_c[0] = t1_ & 0xFFFFu;
      s.scan_c[1] = t1_ >> 16;
    }
  }

  // pass 2: telescoped bucket sums / row emission
  if (!SCOUT) {
    uint32_t ord = ex_b;  // ordinal of the sample inside the tile; predecessor's is ord-1
    for (uint32_t j = j0; j < j1; ++j) {
      const uint8_t f = s.fl[j];
      if (MODE == MODE_EMIT && (f & FL_REACH)) {
        const uint32_t p = s.pos[j];
        const uint64_t h = buf[p];
        const uint64_t gi = tc.rec_base + tc.recs_before + ex_rec;
        P.rec_off[gi] = tc.stream_off + 8ull * p;
        P.rec_type[gi] = static_cast<uint32_t>(h);
        P.rec_size[gi] = static_cast<uint32_t>(h >> 48);
        ++ex_rec;
      }
      if (f & FL_SAMPLE) {
        if (f & FL_ROW) {
          const uint32_t p = s.pos[j];
          const uint32_t bk = s.u.de.bkt[j];
          const uint16_t pj = s.u.de.pred[j];
          const uint16_t nj = s.next[j];
          if (MODE == MODE_AGG) {
            const bool pred_cancel =
                pj != IDX_NONE && (s.fl[pj] & FL_ROW) && s.u.de.bkt[pj] == bk;
            const bool next_cancel =
                nj != IDX_NONE && (s.fl[nj] & FL_ROW) && s.u.de.bkt[nj] == bk;
            if (bk == BKT_OOR) {
              atomicAdd(&s.cnt[C_OOR], 1u);
            } else if (!pred_cancel || !next_cancel) {
              uint64_t* base = P.acc + (static_cast<uint64_t>(tc.stream) * P.n_buckets + bk) * (n + 3);
              const uint32_t pp = (pj != IDX_NONE) ? s.pos[pj] : 0;
              // field order: delta[0..n), dTE, dTR, n_rows
              for (uint32_t e = 0; e < n + 3; ++e) {
                uint64_t plus, minus;
                if (e < n) {
                  plus = xform(P, e, buf[p + PRE + 5 + e]);
                  minus = (pj != IDX_NONE) ? xform(P, e, buf[pp + PRE + 5 + e]) : C.last[e];
                } else if (e == n) {
                  plus = buf[p + PRE + 3];
                  minus = (pj != IDX_NONE) ? buf[pp + PRE + 3] : C.te;
                } else if (e == n + 1) {
                  plus = buf[p + PRE + 4];
                  minus = (pj != IDX_NONE) ? buf[pp + PRE + 4] : C.tr;
                } else {
                  plus = ord;
                  minus = static_cast<uint64_t>(ord) - 1;
                }
                const uint64_t v = (next_cancel ? 0 : plus) - (pred_cancel ? 0 : minus);
                red_add(base + e, v);
              }
            }
          } else {  // MODE_EMIT: the row as CpuCountSampleGenerator writes it (:165-182)
            const uint32_t pp = (pj != IDX_NONE) ? s.pos[pj] : 0;
            const uint64_t ts = buf[p + PRE + 1];
            const uint64_t pts = (pj != IDX_NONE) ? buf[pp + PRE + 1] : C.ts;
            uint64_t* row = P.rows + (tc.row_base + tc.rows_before + ex_row) * (n + 2);
            row[0] = pts;
            row[1] = ts - pts;
            reinterpret_cast<ulonglong2*>(P.row_times)[tc.row_base + tc.rows_before + ex_row] = make_ulonglong2(pts, ts - pts);
            for (uint32_t e = 0; e < n; ++e) {
              const uint64_t cur = xform(P, e, buf[p + PRE + 5 + e]);
              const uint64_t prv = (pj != IDX_NONE) ? xform(P, e, buf[pp + PRE + 5 + e]) : C.last[e];
              row[2 + e] = cur - prv;
            }
            ++ex_row;
          }
        }
        ++ord;
      }
    }
  }
  __syncthreads();  // all reads of the carry are done

  // fold the tile into the carry
  const uint32_t last_s = tot_a >> 16, last_l = tot_a & 0xFFFFu;
  if (last_s != 0) {
    const uint32_t p = s.pos[last_s - 1];
    for (uint32_t e = threadIdx.x; e < n; e += NT) C.last[e] = xform(P, e, buf[p + PRE + 5 + e]);
    if (threadIdx.x == 0) {
      C.ts = buf[p + PRE + 1];
      C.te = buf[p + PRE + 3];
      C.tr = buf[p + PRE + 4];
      C.valid = 1;
      C.gap = last_l > last_s;
      C.last_off = tc.tile_abs + 8ull * p;
    }
  } else if (last_l != 0) {
    if (threadIdx.x == 0) C.gap = 1;
  }
  if (MODE == MODE_EMIT && !SCOUT) {
    __syncthreads();
    tc.rows_before += s.scan_c[0];
    tc.recs_before += s.scan_c[1];
  }
  __syncthreads();
}

// General path, MODE_AGG: check and count the reachable records of one staged (sub-)tile as
// decode_tile does, but instead of aggregating them list the SAMPLEs exactly as index_tile does
// (s.pos[k] = word position, s.fl[k] = a LOST lies between sample k-1 and k; s.idx_n, s.idx_gap),
// so that ONE aggregation routine (uniform_segment<..., IDX = true>) serves every path.
// Leaves the carry alone except for "a LOST and no SAMPLE behind it".  Contains barriers.
template <int PRE>
__device__ __forceinline__ void list_tile(Smem& s, const Params& P, const uint64_t* buf, uint32_t K,
                                          uint32_t cur) {
  const uint32_t n = P.n_events;
  const uint32_t ipt = (K + NT - 1) / NT;  // candidates per thread (blocked)
  const uint32_t j0 = threadIdx.x * ipt;
  const uint32_t j1 = min(j0 + ipt, K);
  if (threadIdx.x == 0) s.red_key = 0;  // number of samples with a LOST in front
  // pass 0: type checks, counters of everything that is not a SAMPLE, scan inputs
  uint32_t a = 0, b = 0;
  uint32_t n_thr = 0, n_unthr = 0, n_read = 0, n_other = 0;
  unsigned long long lost = 0;
  for (uint32_t j = j0; j < j1; ++j) {
    uint8_t f = s.fl[j];
    if (f & FL_REACH) {
      const uint32_t p = s.pos[j];
      const uint64_t h = buf[p];
      const uint32_t type = static_cast<uint32_t>(h);
      const uint32_t size = static_cast<uint32_t>(h >> 48);
      if (type == REC_SAMPLE) {
        const uint64_t nr = (size + 16 >= (40u + 8u * PRE)) ? buf[p + PRE + 2] : ~0ull;
        if (size < (40u + 8u * PRE) || nr != ((size - (40u + 8u * PRE)) >> 3) || nr != n) {
          atomicMin(&s.tile_err, p);
        } else {
          f |= FL_SAMPLE;
          a = __vmaxu2(a, (j + 1) << 16);
          b += 1;
        }
      } else {
        ++n_other;
        if (type == REC_LOST) {
          if (size < 24) {
            atomicMin(&s.tile_err, p);
          } else {
            f |= FL_LOST;
            a = __vmaxu2(a, j + 1);
            lost += buf[p + 2];  // num_lost (:1448)
          }
        } else if (type == REC_READ) {
          const uint64_t nr = (size >= 24) ? buf[p + 2] : ~0ull;
          if (size < 40 || nr != ((size - 40u) >> 3)) atomicMin(&s.tile_err, p);  // :1518-1523
          else ++n_read;
        } else if (type == REC_THROTTLE) {
          ++n_thr;
        } else if (type == REC_UNTHROTTLE) {
          ++n_unthr;
        }
      }
      s.fl[j] = f;
    }
  }
  uint32_t ex_a, ex_b, tot_a, tot_b;
  block_scan(s, a, b, ex_a, ex_b, tot_a, tot_b);
  // pass 1: every sample's rank among the samples and whether a LOST precedes it inside the tile
  uint32_t run_a = ex_a, ord = ex_b, flagged = 0;
  for (uint32_t j = j0; j < j1; ++j) {
    const uint8_t f = s.fl[j];
    if (f & FL_SAMPLE) {
      const uint32_t prev_s = run_a >> 16, prev_l = run_a & 0xFFFFu;
      const bool lost_before = prev_s != 0 ? (prev_l > prev_s) : (prev_l != 0);
      s.u.de.bkt[ord] = static_cast<uint32_t>(s.pos[j]) | (lost_before ? 0x10000u : 0u);
      flagged += lost_before ? 1u : 0u;
      run_a = __vmaxu2(run_a, (j + 1) << 16);
      ++ord;
    } else if (f & FL_LOST) {
      run_a = __vmaxu2(run_a, j + 1);
    }
  }
  {
    const uint32_t vals[5] = {n_thr, n_unthr, n_read, n_other, flagged};
#pragma unroll
    for (int k = 0; k < 5; ++k) {
      const uint32_t v = __reduce_add_sync(0xffffffffu, vals[k]);
      if ((threadIdx.x & 31) == 0 && v) {
        if (k == 0) atomicAdd(&s.cnt[C_THROTTLE], v);
        else if (k == 1) atomicAdd(&s.cnt[C_UNTHROTTLE], v);
        else if (k == 2) atomicAdd(&s.cnt[C_READ], v);
        else if (k == 3) atomicAdd(&s.cnt[C_RECORDS], v);
        else atomicAdd(&s.red_key, v);
      }
    }
    if (__any_sync(0xffffffffu, lost != 0)) {
      unsigned long long l = lost;
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) l += __shfl_xor_sync(0xffffffffu, l, d);
      if ((threadIdx.x & 31) == 0) atomicAdd(&s.lost_sum, l);
    }
  }
  __syncthreads();  // the list is complete; pos / fl of the candidates are no longer needed
  const uint32_t ns = tot_b, last_s = tot_a >> 16, last_l = tot_a & 0xFFFFu;
  for (uint32_t k = threadIdx.x; k < ns; k += NT) {
    const uint32_t v = s.u.de.bkt[k];
    s.pos[k] = static_cast<uint16_t>(v);
    s.fl[k] = static_cast<uint8_t>(v >> 16);
  }
  if (threadIdx.x == 0) {
    s.idx_n = ns;
    s.idx_gap = (last_l > last_s ? 1u : 0u) | (s.red_key << 1);  // bit 0: a LOST behind the last sample
    s.idx_bad = 0xFFFFFFFFu;
    if (ns == 0 && last_l != 0) s.c[cur].gap = 1;  // no sample behind the LOST: the carry remembers it
  }
  __syncthreads();
}

// Index of a mixed tile (MODE_AGG).  A tile that holds other records than configured SAMPLEs
// (THROTTLE / UNTHROTTLE pairs under a high sample rate, LOST) used to be cut at each of them
// and every piece cost a full pass.  Instead warp 0 walks the chain once -- runs of SAMPLEs 32
// at a time (lane l tests the record at p + l*S), any other record by its size, with the
// checks and the accounting of consume() (:1440-1450 LOST, :1476-1487 THROTTLE, :1515-1530
// READ, :1565-1570 anything else) -- and lists the SAMPLEs: s.pos[j] = word position,
// s.fl[j] = 1 if a LOST lies between sample j-1 and j (its row does not exist, gen :110-124).
// One pass of uniform_segment<..., IDX = true> over the list then does what the plain pass does
// for a contiguous run.  Must be called by all threads; contains barriers.
enum { IX_ENTRY = 0, IX_END, IX_NREC, IX_FLAGS, IX_OTHER, IX_BAD, IX_THR_UNTHR, IX_READ };
enum : uint32_t { IXF_GAP_END = 1, IXF_GAP_FIRST = 2, IXF_OVERFLOW = 4 };
constexpr uint32_t IX_LIST = 64;  // listed samples per warp (a tile holds at most PASS_RECORDS = 185)

// One warp walks the chain from word `p` until it reaches `stop` (or a record consume() stalls on):
// runs of configured SAMPLEs 32 at a time (lane l tests the record at p + l*S), any other record by
// its size with consume()'s checks.  The SAMPLEs go to out[k] = position | (a LOST lies between
// sample k-1 and k) << 16; res[] gets the IX_* summary; lost_out the sum of num_lost.
// DIRECT: the list goes straight to s.pos / s.fl (a single walker), res[IX_READ + ...] unchanged,
// and the number of flagged samples is returned in res[IX_ENTRY].
template <int PRE, int NEV, bool DIRECT = false>
__device__ __forceinline__ void index_walk(const Params& P, const uint64_t* buf, uint32_t wt, uint32_t p,
                                           uint32_t stop, uint32_t* out, uint32_t cap, uint32_t* res,
                                           unsigned long long* lost_out, Smem* sm = nullptr) {
  const uint32_t lane = threadIdx.x & 31;
  const uint32_t n = NEV ? static_cast<uint32_t>(NEV) : P.n_events;
  const uint32_t S = 5 + PRE + n;
  const uint64_t want = static_cast<uint64_t>(REC_SAMPLE) | (static_cast<uint64_t>(S * 8) << 48);
  uint32_t nrec = 0, bad = 0xFFFFFFFFu, n_other = 0, n_thr = 0, n_unthr = 0, n_read = 0, flags = 0, n_flagged = 0;
  unsigned long long lost = 0;
  bool gap = false;  // a LOST since the last listed sample
  (void)wt;
  while (p < stop) {
    const uint32_t q = p + lane * S;
    const bool ok = q < stop && (buf[q] & 0xFFFF0000FFFFFFFFull) == want && buf[q + PRE + 2] == n;
    const uint32_t okm = __ballot_sync(0xffffffffu, ok);
    const uint32_t run = okm == 0xFFFFFFFFu ? 32u : static_cast<uint32_t>(__ffs(~okm) - 1);
    if (run > 0) {
      if (nrec + run > cap) {
        flags |= IXF_OVERFLOW;
        break;
      }
      if (DIRECT) {
        if (lane < run) {
          sm->pos[nrec + lane] = static_cast<uint16_t>(q);
          sm->fl[nrec + lane] = (lane == 0 && gap) ? 1 : 0;
        }
        n_flagged += gap ? 1 : 0;
      } else if (lane < run) {
        out[nrec + lane] = q | ((lane == 0 && gap) ? 0x10000u : 0u);
      }
      if (nrec == 0 && gap) flags |= IXF_GAP_FIRST;
      nrec += run;
      gap = false;
      p += run * S;
      continue;
    }
    const uint64_t h = buf[p];
    const uint32_t type = static_cast<uint32_t>(h), size = static_cast<uint32_t>(h >> 48);
    // what consume() stalls on (or the fences of DESIGN.md): the general path reports it
    if (size < 8 || (size & 7) || type == REC_SAMPLE || (type == REC_LOST && size < 24) ||
        (type == REC_READ && (size < 40 || buf[p + 2] != ((size - 40u) >> 3)))) {
      bad = p;
      break;
    }
    n_other += 1;
    if (type == REC_LOST) {
      lost += buf[p + 2];
      gap = true;
    } else if (type == REC_THROTTLE) {
      n_thr += 1;
    } else if (type == REC_UNTHROTTLE) {
      n_unthr += 1;
    } else if (type == REC_READ) {
      n_read += 1;
    }
    p += size >> 3;
  }
  if (lane == 0) {
    if (DIRECT) res[IX_ENTRY] = n_flagged;
    res[IX_END] = p;
    res[IX_NREC] = nrec;
    res[IX_FLAGS] = flags | (gap ? IXF_GAP_END : 0u);
    res[IX_OTHER] = n_other;
    res[IX_BAD] = bad;
    res[IX_THR_UNTHR] = n_thr | (n_unthr << 16);
    res[IX_READ] = n_read;
    *lost_out = lost;
  }
}

// Index of a mixed tile (MODE_AGG).  A tile that holds other records than configured SAMPLEs
// (THROTTLE / UNTHROTTLE pairs under a high sample rate, LOST) is not cut at each of them:
// the chain from word `ew` to the end of the tile is walked once and the SAMPLEs are listed --
// s.pos[j] = word position, s.fl[j] = 1 if a LOST lies between sample j-1 and j (its row does
// not exist, gen :110-124) -- with the checks and the accounting of consume() (:1440-1450 LOST,
// :1476-1487 THROTTLE, :1515-1530 READ, :1565-1570 anything else); one pass of
// uniform_segment<..., IDX = true> over the list then does what the plain pass does for a run.
// All NWARP warps walk at once: warp 0 from `ew`, warp w from the first word of its sixth of the
// tile that looks like a configured SAMPLE (a guess), each up to the next warp's entry.  The
// guesses are proven the way k_verify proves chunks: warp w's entry must be exactly where the
// walk before it ended.  From the first entry that is not, one warp walks the rest.
// parallel = false: only warp 0 walks (cheaper when the tile holds one or two other records, as in
// BASELINE config 3: fewer barriers, no guesses to prove).
template <int PRE, int NEV>
__device__ __forceinline__ void index_tile(Smem& s, const Params& P, const uint64_t* buf, uint32_t wt,
                                           uint32_t ew, uint32_t cur, bool parallel) {
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t n = NEV ? static_cast<uint32_t>(NEV) : P.n_events;
  const uint32_t S = 5 + PRE + n;
  const uint64_t want = static_cast<uint64_t>(REC_SAMPLE) | (static_cast<uint64_t>(S * 8) << 48);
  unsigned long long* const lostw = reinterpret_cast<unsigned long long*>(s.u.de.pred);  // [NWARP], 8-byte aligned
  uint32_t* const list = s.u.de.bkt;                                                      // [NWARP][IX_LIST]
  if (!parallel) {
    // one walker, one barrier
    if (warp == 0) {
      index_walk<PRE, NEV, true>(P, buf, wt, ew, wt, nullptr, MAXC, s.ix[0], &lostw[0], &s);
      __syncwarp();
      const uint32_t nr = s.ix[0][IX_NREC];
      const uint32_t f = s.ix[0][IX_ENTRY];
      if (lane == 0) {
        const bool gap = (s.ix[0][IX_FLAGS] & IXF_GAP_END) != 0;
        s.cnt[C_RECORDS] += s.ix[0][IX_OTHER];
        s.cnt[C_THROTTLE] += s.ix[0][IX_THR_UNTHR] & 0xFFFFu;
        s.cnt[C_UNTHROTTLE] += s.ix[0][IX_THR_UNTHR] >> 16;
        s.cnt[C_READ] += s.ix[0][IX_READ];
        s.lost_sum += lostw[0];
        s.idx_n = nr;
        s.idx_end = s.ix[0][IX_END];
        s.idx_bad = s.ix[0][IX_BAD];
        s.idx_other = s.ix[0][IX_OTHER];
        s.idx_gap = (gap ? 1u : 0u) | (f << 1);
        if (nr == 0 && gap) s.c[cur].gap = 1;  // no sample behind the last LOST of this tile: the carry remembers it
      }
    }
    __syncthreads();
    return;
  }
  // ---- entries
  const uint32_t span = wt > ew ? wt - ew : 0;
  const uint32_t sub = (span + NWARP - 1) / NWARP;
  uint32_t entry = 0xFFFFFFFFu;
  if (warp == 0) {
    entry = ew;
  } else if (sub >= S) {
    const uint32_t lo = ew + warp * sub, hi = min(lo + sub, wt);
    for (uint32_t q0 = lo; q0 < hi && entry == 0xFFFFFFFFu; q0 += 32) {
      const uint32_t q = q0 + lane;
      const bool m = q < hi && q + PRE + 2 < wt + OV_BYTES / 8 && (buf[q] & 0xFFFF0000FFFFFFFFull) == want &&
                     buf[q + PRE + 2] == n;
      const uint32_t bal = __ballot_sync(0xffffffffu, m);
      if (bal) entry = q0 + (__ffs(bal) - 1);
    }
  }
  if (lane == 0) s.ix[warp][IX_ENTRY] = entry;
  __syncthreads();
  // ---- walks
  if (entry != 0xFFFFFFFFu) {
    uint32_t stop = wt;
    for (uint32_t w = warp + 1; w < NWARP; ++w) {
      const uint32_t e = s.ix[w][IX_ENTRY];
      if (e != 0xFFFFFFFFu) {
        stop = e;
        break;
      }
    }
    index_walk<PRE, NEV>(P, buf, wt, entry, stop, list + warp * IX_LIST, IX_LIST, s.ix[warp], &lostw[warp]);
  }
  __syncthreads();
  // ---- proof of the guesses, and what is good of the lists (every thread computes the same)
  uint32_t good = 0;          // warps 0..good-1 are consistent
  uint32_t pos_end = ew;      // where the proven chain stands
  uint32_t total = 0, others = 0, bad = 0xFFFFFFFFu;
  uint32_t base_of_mine = 0;  // where this warp's list goes
  bool gap = false, mine_good = false, first_gap_mine = false;
  for (uint32_t w = 0; w < NWARP; ++w) {
    const uint32_t e = s.ix[w][IX_ENTRY];
    if (e == 0xFFFFFFFFu) {  // no guess: the walk before it goes on through this warp's part
      good = w + 1;
      continue;
    }
    if (e != pos_end || bad != 0xFFFFFFFFu) break;
    const uint32_t fl = s.ix[w][IX_FLAGS];
    if (fl & IXF_OVERFLOW) break;
    const uint32_t nr = s.ix[w][IX_NREC];
    if (w == warp) {
      mine_good = true;
      base_of_mine = total;
      first_gap_mine = gap;
    }
    gap = nr > 0 ? (fl & IXF_GAP_END) != 0 : (gap || (fl & IXF_GAP_END) != 0);
    total += nr;
    others += s.ix[w][IX_OTHER];
    pos_end = s.ix[w][IX_END];
    bad = s.ix[w][IX_BAD];
    good = w + 1;
  }
  // ---- the lists of the proven walks, in order; counters
  if (mine_good) {
    const uint32_t nr = s.ix[warp][IX_NREC];
    for (uint32_t k = lane; k < nr; k += 32) {
      const uint32_t v = list[warp * IX_LIST + k];
      s.pos[base_of_mine + k] = static_cast<uint16_t>(v);
      s.fl[base_of_mine + k] = static_cast<uint8_t>((v >> 16) | ((k == 0 && first_gap_mine) ? 1u : 0u));
    }
    if (lane == 0) {
      atomicAdd(&s.cnt[C_RECORDS], s.ix[warp][IX_OTHER]);
      const uint32_t tu = s.ix[warp][IX_THR_UNTHR];
      if (tu & 0xFFFFu) atomicAdd(&s.cnt[C_THROTTLE], tu & 0xFFFFu);
      if (tu >> 16) atomicAdd(&s.cnt[C_UNTHROTTLE], tu >> 16);
      if (s.ix[warp][IX_READ]) atomicAdd(&s.cnt[C_READ], s.ix[warp][IX_READ]);
      if (lostw[warp]) atomicAdd(&s.lost_sum, lostw[warp]);
    }
  }
  __syncthreads();
  // ---- a refuted guess: one warp walks the rest of the tile from where the proven chain stands
  const bool rest = good < NWARP && bad == 0xFFFFFFFFu && pos_end < wt;
  if (rest && warp == 0) {
    index_walk<PRE, NEV>(P, buf, wt, pos_end, wt, list, MAXC, s.ix[0], &lostw[0]);
    __syncwarp();
    const uint32_t nr = s.ix[0][IX_NREC];
    for (uint32_t k = lane; k < nr; k += 32) {
      const uint32_t v = list[k];
      s.pos[total + k] = static_cast<uint16_t>(v);
      s.fl[total + k] = static_cast<uint8_t>((v >> 16) | ((k == 0 && gap) ? 1u : 0u));
    }
    if (lane == 0) {
      s.cnt[C_RECORDS] += s.ix[0][IX_OTHER];
      s.cnt[C_THROTTLE] += s.ix[0][IX_THR_UNTHR] & 0xFFFFu;
      s.cnt[C_UNTHROTTLE] += s.ix[0][IX_THR_UNTHR] >> 16;
      s.cnt[C_READ] += s.ix[0][IX_READ];
      s.lost_sum += lostw[0];
    }
  }
  __syncthreads();
  if (rest) {
    const uint32_t nr = s.ix[0][IX_NREC], fl = s.ix[0][IX_FLAGS];
    gap = nr > 0 ? (fl & IXF_GAP_END) != 0 : (gap || (fl & IXF_GAP_END) != 0);
    total += nr;
    others += s.ix[0][IX_OTHER];
    pos_end = s.ix[0][IX_END];
    bad = s.ix[0][IX_BAD];
  }
  // samples without a row = the flagged entries of the final list
  uint32_t f = 0;
  for (uint32_t k = threadIdx.x; k < total; k += NT) f += s.fl[k] ? 1u : 0u;
  f = __reduce_add_sync(0xffffffffu, f);
  if (threadIdx.x == 0) {
    s.idx_n = total;
    s.idx_end = pos_end;
    s.idx_bad = bad;
    s.idx_other = others;
    s.idx_gap = gap ? 1u : 0u;  // bit 0: a LOST behind the last sample; the count of flagged samples is added below
    if (total == 0 && gap) s.c[cur].gap = 1;  // no sample behind the last LOST of this tile: the carry remembers it
  }
  __syncthreads();
  if (lane == 0 && f) atomicAdd(&s.idx_gap, f << 1);
  __syncthreads();
}

// Fast path for a tile in which every record from `entry` on is a SAMPLE of the configured
// group (type 9, size 40+8n, nr == n).  Then record j sits at entry + j*S and each equality
// test IS the chain link (record j's size carries the walk to record j+1), so the chain is
// proven without parsing the other words; the predecessor of sample j is j-1 and there is no
// LOST, so no scan is needed either.
//
// Telescoping.  Row j adds V_j - V_{j-1} to bucket(j), V = (value[], te, tr, ordinal).  So the
// vector V_j is used twice: +V_j into bucket(j) (if row j exists) and -V_j into bucket(j+1)
// (if row j+1 exists); when both exist and the buckets are equal the two cancel.  Only records
// at a bucket transition (and the tile's two edges) touch global memory.
//
// Warp-local: a warp takes 31 consecutive records on lanes 0..30 (record -1 is the carry),
// lane 31 holds the right neighbour, so transitions are found with shuffles and the only block
// barrier is the vote that the whole tile is uniform.  The carry is read from s.c[cur] and the
// new one written to s.c[cur^1]; the caller's end-of-tile barrier publishes it.
// Returns how many records from `entry` on were proven and committed: all of them for a
// uniform tile, else the run up to the first record that is something else (possibly 0); the
// caller deals with that record (odd_record) and calls again behind it.  Nothing is committed
// for records at or after the first mismatch.
// MODE_EMIT writes, instead of bucket sums, the rows exactly as CpuCountSampleGenerator writes
// them to its output ring ([tstamp, duration, delta...], gen :165-182) and the record table;
// a warp assembles its (consecutive) rows in shared memory and copies them out coalesced.
constexpr int EMIT_MAX_ROW_WORDS = 16;                     // rows of up to 14 events go through the stage
constexpr int EMIT_STAGE_WORDS = 31 * EMIT_MAX_ROW_WORDS;  // per warp
constexpr size_t EMIT_STAGE_BYTES = static_cast<size_t>(NWARP) * EMIT_STAGE_WORDS * 8;

// Dynamic shared memory of k_decode: [Smem][EMIT row staging | AttrCache | nothing][two tile
// buffers of tile_bytes + OV_BYTES each].  The tile buffers come last because their size is a
// launch parameter.
template <int MODE, int PRE>
__host__ __device__ constexpr size_t decode_data_off() {
  return (sizeof(Smem) + (MODE == MODE_EMIT ? EMIT_STAGE_BYTES : (PRE == 2 ? sizeof(AttrCache) : 0)) + 127) &
         ~static_cast<size_t>(127);
}
template <int MODE, int PRE>
__host__ __device__ constexpr size_t decode_smem_bytes(uint32_t tile_bytes) {
  return decode_data_off<MODE, PRE>() + 2 * (static_cast<size_t>(tile_bytes) + OV_BYTES);
}

// IDX: the records are the s.idx_n samples index_tile listed (MODE_AGG only): positions from
// s.pos, s.fl[j] = a LOST lies before sample j; they are valid by construction, so no vote.
template <int MODE, int PRE, int NEV, bool IDX = false>
__device__ __forceinline__ uint32_t uniform_segment(Smem& s, const Params& P, const uint64_t* buf,
                                                    uint32_t wt, uint32_t entry, TileCtx& tc,
                                                    uint32_t cur, uint32_t& k_all) {
  const Carry& C = s.c[cur];
  // NEV != 0: the kernel is specialised for this group size (the strides, the field map of the
  // accumulator record and the division by the record size become constants)
  const uint32_t n = NEV ? static_cast<uint32_t>(NEV) : P.n_events;
  constexpr uint32_t pre = PRE;
  const uint32_t S = 5 + pre + n;  // words per SAMPLE
  const uint64_t want = static_cast<uint64_t>(REC_SAMPLE) | (static_cast<uint64_t>(S * 8) << 48);
  // records whose header lies inside the tile: ceil((wt - entry) / S) by multiply-shift
  const int Kall = NEV ? static_cast<int>((wt - entry + S - 1) / S)
                       : static_cast<int>((static_cast<uint64_t>(wt - entry + S - 1) * P.s_magic) >> 32);
  k_all = static_cast<uint32_t>(IDX ? s.idx_n : Kall);
  int K = min(IDX ? static_cast<int>(s.idx_n) : Kall, PASS_RECORDS);  // never more by construction of the tile size
  // word position of record j, and whether its row is missing because of a LOST before it
  auto posof = [&](int j) -> uint32_t { return IDX ? static_cast<uint32_t>(s.pos[j]) : entry + j * S; };
  auto gapof = [&](int j) -> bool { return IDX && j >= 0 && s.fl[j] != 0; };
  const bool link0 = C.valid != 0 && C.gap == 0;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  constexpr int OWN = 31;  // records owned per warp-iteration (lanes 0..30)
  // the tile is sized so that K + 1 <= NWARP * OWN items: one pass (tile_bytes_for)
  constexpr int MAXG = 1;
  const uint32_t F = n + 3;
  const bool any_tsc = P.tsc_mask != 0;

  // pass 1: check every record, compute its bucket
  uint32_t bk_[MAXG];
  int first_bad = 0x7FFFFFFF;   // first record index that is not such a SAMPLE (this lane's view)
  int regress = 0x7FFFFFFF;     // first record index whose row would go back in time
#pragma unroll
  for (int g = 0; g < MAXG; ++g) {
    const int j0 = (g * NWARP + warp) * OWN - 1;  // first item of this warp-iteration
    if (j0 >= K) break;                           // warp-uniform: nothing here
    const int jr = j0 + lane;                     // record index; -1 is the carry
    uint64_t ts = 0;
    bk_[g] = BKT_OOR;
    if (jr >= 0 && jr < K) {
      const uint32_t p = posof(jr);
      // misc (bits 32..47) is free; type and size must match, and so must nr (three loads
      // issued together, one combined test: no short-circuit branches)
      const uint64_t hd = buf[p], nr = buf[p + pre + 2];
      ts = buf[p + pre + 1];
      const bool mismatch = ((hd & 0xFFFF0000FFFFFFFFull) != want) | (nr != n);
      if (!IDX && lane < 31 && mismatch) first_bad = min(first_bad, jr);
      bk_[g] = bucket_of(P, ts);
    } else if (jr == -1) {
      ts = C.ts;
    }
    const uint64_t nts = __shfl_down_sync(0xffffffffu, ts, 1);
    // row jr+1 exists and goes back in time (gen :156-163)
    if (lane < 31 && jr + 1 < K && (jr >= 0 || link0) && !gapof(jr + 1) && ts > nts) regress = min(regress, jr + 1);
  }
  if (!IDX && !__syncthreads_and(first_bad == 0x7FFFFFFF)) {
    // not uniform to the end: agree on the first record that breaks the run
    if (threadIdx.x == 0) s.red_key = static_cast<uint32_t>(Kall);
    __syncthreads();
    if (first_bad != 0x7FFFFFFF) atomicMin(&s.red_key, static_cast<uint32_t>(first_bad));
    __syncthreads();
    K = static_cast<int>(s.red_key);
    __syncthreads();  // red_key may be reused right away
    if (K == 0) return 0;
  }
  if (IDX && K == 0) return 0;
  if (regress < K) atomicMin(&s.tile_err, posof(regress));

  if (MODE == MODE_EMIT) {
    // rows of this segment are consecutive: record jr's row has ordinal jr - (link0 ? 0 : 1)
    uint64_t* stage = reinterpret_cast<uint64_t*>(reinterpret_cast<unsigned char*>(&s) + sizeof(Smem)) +
                      warp * EMIT_STAGE_WORDS;
    const uint32_t RW = n + 2;
    const int skip = link0 ? 0 : 1;
#pragma unroll
    for (int g = 0; g < MAXG; ++g) {
      const int j0 = (g * NWARP + warp) * OWN - 1;
      if (j0 >= K) break;
      const int jr = j0 + lane;
      if (lane < 31 && jr >= 0 && jr < K) {  // record table
        const uint64_t gi = tc.rec_base + tc.recs_before + jr;
        P.rec_off[gi] = tc.stream_off + 8ull * (entry + jr * S);
        P.rec_type[gi] = REC_SAMPLE;
        P.rec_size[gi] = 8 * S;
      }
      const int jlo = max(max(j0, 0), skip), jhi = min(j0 + 31, K);  // records with a row here
      if (lane < 31 && jr >= jlo && jr < jhi) {  // the compact copy of the row's times: 16 bytes per lane, consecutive
        const uint32_t p = entry + jr * S;
        const uint64_t ts = buf[p + pre + 1];
        const uint64_t pts = jr ? buf[p - S + pre + 1] : C.ts;
        reinterpret_cast<ulonglong2*>(P.row_times)[tc.row_base + tc.rows_before + static_cast<uint64_t>(jr - skip)] =
            make_ulonglong2(pts, ts - pts);
      }
      if (RW > EMIT_MAX_ROW_WORDS) {
        // wide rows (more than 14 events) do not fit the stage and do not need it: the warp's
        // rows are one contiguous run of words in the output; lane l computes words l, l+32, ...
        // of that run straight from the tile
        if (jhi > jlo) {
          uint64_t* dst = P.rows + (tc.row_base + tc.rows_before + static_cast<uint64_t>(jlo - skip)) * RW;
          const uint32_t words = static_cast<uint32_t>(jhi - jlo) * RW;
          for (uint32_t o = lane; o < words; o += 32) {
            const uint32_t r = o / RW, w = o - r * RW;
            const int j = jlo + static_cast<int>(r);
            const uint32_t p = entry + j * S;
            uint64_t v;
            if (w < 2) {
              const uint64_t ts = buf[p + pre + 1];
              const uint64_t pts = j ? buf[p - S + pre + 1] : C.ts;
              v = w == 0 ? pts : ts - pts;
            } else {
              const uint32_t e = w - 2;
              const uint64_t cv = xform(P, e, buf[p + pre + 5 + e]);
              const uint64_t pv = j ? xform(P, e, buf[p - S + pre + 5 + e]) : C.last[e];
              v = cv - pv;
            }
            dst[o] = v;
          }
        }
        continue;
      }
      __syncwarp();  // the previous copy-out has read the stage
      if (lane < 31 && jr >= jlo && jr < jhi) {
        const uint32_t p = entry + jr * S;
        uint64_t* row = stage + (jr - jlo) * RW;
        const uint64_t ts = buf[p + pre + 1];
        const uint64_t pts = jr ? buf[p - S + pre + 1] : C.ts;
        row[0] = pts;
        row[1] = ts - pts;
        for (uint32_t e = 0; e < n; ++e) {
          const uint64_t cv = xform(P, e, buf[p + pre + 5 + e]);
          const uint64_t pv = jr ? xform(P, e, buf[p - S + pre + 5 + e]) : C.last[e];
          row[2 + e] = cv - pv;
        }
      }
      __syncwarp();
      if (jhi > jlo) {
        uint64_t* dst = P.rows + (tc.row_base + tc.rows_before + static_cast<uint64_t>(jlo - skip)) * RW;
        const uint32_t words = static_cast<uint32_t>(jhi - jlo) * RW;
        for (uint32_t w = lane; w < words; w += 32) dst[w] = stage[w];
      }
    }
    tc.rows_before += static_cast<uint32_t>(K - skip);
    tc.recs_before += static_cast<uint32_t>(K);
  }

  if (MODE == MODE_AGG && pre == 2) {
    // who was sampled and where
#pragma unroll
    for (int g = 0; g < MAXG; ++g) {
      const int jr = (g * NWARP + warp) * OWN - 1 + lane;
      if (lane < 31 && jr >= 0 && jr < K) {
        const uint32_t p = posof(jr);
        attr_count<true>(P, buf[p + 1]);
        attr_count<false>(P, buf[p + 2]);
      }
    }
  }

  // pass 2
  uint32_t n_oor = 0;
  // lane -> field of the accumulator record (delta[0..n), dTE, dTR, n_rows) and where that
  // field sits in a SAMPLE (value[e] at word 5+e, time_enabled at 3, time_running at 4)
  const bool two = F <= 16;
  const uint32_t e0 = two ? (lane & 15) : lane;
  const uint32_t woff0 = pre + (e0 < n ? 5 + e0 : (e0 == n ? 3 : 4));
  uint64_t* const acc_s = tc.acc_s;                       // this stream's accumulator records
  uint64_t* const acc_lane = acc_s + e0;                  // ... and this lane's field in bucket 0
  const bool lane_tsc = any_tsc && e0 < n && ((P.tsc_mask >> e0) & 1ull);  // field e0 is transformed
#pragma unroll
  for (int g = 0; g < MAXG; ++g) {
    if (MODE != MODE_AGG) break;
    const int j0 = (g * NWARP + warp) * OWN - 1;
    if (j0 >= K) break;
    const int jr = j0 + lane;
    const uint32_t bk = bk_[g];
    const uint32_t bk_next = __shfl_down_sync(0xffffffffu, bk, 1);
    const bool own = lane < 31 && jr < K;                    // jr >= -1 always
    const bool rowA = (jr >= 1 || (jr == 0 && link0)) && !gapof(jr);            // row jr exists
    const bool rowB = jr + 1 < K && (jr >= 0 || link0) && !gapof(jr + 1);       // row jr+1 exists
    const bool cancel = rowA && rowB && bk == bk_next;
    const bool doA = own && rowA && !cancel && bk != BKT_OOR;
    const bool doB = own && rowB && !cancel && bk_next != BKT_OOR;
    n_oor += __popc(__ballot_sync(0xffffffffu, own && rowA && bk == BKT_OOR));
    uint32_t mask = __ballot_sync(0xffffffffu, doA || doB);
    const uint32_t flags = (doA ? 1u : 0u) | (doB ? 2u : 0u);
    // one group of lanes per such record, one lane per field: the REDs of a record go to
    // consecutive addresses
    while (mask) {
      int src = __ffs(mask) - 1;
      mask &= mask - 1;
      if (two) {
        const int b1 = mask ? __ffs(mask) - 1 : -1;
        mask &= mask - 1;  // no-op when mask is already 0
        if (lane >= 16) src = b1;
      }
      const int srcl = src < 0 ? 0 : src;
      const uint32_t pk = __shfl_sync(0xffffffffu, flags, srcl);
      const uint32_t bkA = __shfl_sync(0xffffffffu, bk, srcl);
      const uint32_t bkB = __shfl_sync(0xffffffffu, bk_next, srcl);
      if (src >= 0) {
        const int j = j0 + src;
        const uint32_t p = j < 0 ? 0u : posof(j);
        uint32_t e = e0, woff = woff0;
        for (;;) {
          if (e < F) {
            uint64_t v;
            if (e == n + 2) {
              v = static_cast<uint64_t>(static_cast<int64_t>(j));  // ordinal; -1 for the carry
            } else if (j < 0) {
              v = e < n ? C.last[e] : (e == n ? C.te : C.tr);  // carry values are already transformed
            } else {
              v = buf[p + woff];
              if (two ? lane_tsc : (any_tsc && e < n)) v = xform(P, e, v);
            }
            uint64_t* const fa = two ? acc_lane : acc_s + e;
            if (pk & 1u) red_add(fa + static_cast<uint64_t>(bkA) * F, v);
            if (pk & 2u) red_add(fa + static_cast<uint64_t>(bkB) * F, 0ull - v);
          }
          if (two || e + 32 >= F) break;
          e += 32;  // n > 13: a lane serves several fields
          woff = pre + (e < n ? 5 + e : (e == n ? 3 : 4));
        }
      }
    }
  }
  if (lane == 0 && n_oor) atomicAdd(&s.cnt[C_OOR], n_oor);
  // the new carry (other copy): last record of the tile
  if (warp == NWARP - 1) {
    Carry& N = s.c[cur ^ 1];
    const uint32_t p = posof(K - 1);
    for (uint32_t e = lane; e < n; e += 32) N.last[e] = xform(P, e, buf[p + pre + 5 + e]);
    if (lane == 0) {
      N.ts = buf[p + pre + 1];
      N.te = buf[p + pre + 3];
      N.tr = buf[p + pre + 4];
      N.valid = 1;
      N.gap = IDX ? (s.idx_gap & 1u) : 0;  // a LOST behind the last sample of an indexed tile
      N.last_off = tc.tile_abs + 8ull * p;
      // samples without a row: behind a LOST (indexed tiles), or the first one after an invalid state
      const uint32_t rowless = (IDX ? (s.idx_gap >> 1) : 0) + ((!link0 && !gapof(0)) ? 1u : 0u);
      s.cnt[C_SAMPLE] += K;
      s.cnt[C_RECORDS] += K;
      s.cnt[C_ROWS] += K - rowless;
      s.cnt[C_UNKNOWN] += rowless;
    }
  }
  return static_cast<uint32_t>(K);
}

// The record at word `ew` is not a SAMPLE of the configured group.  If it is a well-formed
// record of another type, account for it exactly as consume() does (LOST: sum num_lost and
// invalidate the interval, :1440-1450 / gen :110-124; THROTTLE/UNTHROTTLE: count, :1476-1487;
// READ: size check, :1515-1530; anything else: skip by size, :1565-1570) and return its size in
// words; else return 0 and let the general path (which owns the error reporting) look at it.
template <int MODE>
__device__ __forceinline__ uint32_t odd_record(Smem& s, const Params& P, const uint64_t* buf, uint32_t ew,
                                               uint32_t cur, TileCtx& tc) {
  const uint64_t h = buf[ew];
  const uint32_t type = static_cast<uint32_t>(h), size = static_cast<uint32_t>(h >> 48);
  if (size < 8 || (size & 7) || type == REC_SAMPLE) return 0;
  if (type == REC_LOST && size < 24) return 0;
  if (type == REC_READ && (size < 40 || buf[ew + 2] != ((size - 40u) >> 3))) return 0;
  if (threadIdx.x == 0) {
    s.cnt[C_RECORDS] += 1;
    if (type == REC_LOST) {
      s.lost_sum += buf[ew + 2];
      s.c[cur].gap = 1;
    } else if (type == REC_THROTTLE) {
      s.cnt[C_THROTTLE] += 1;
    } else if (type == REC_UNTHROTTLE) {
      s.cnt[C_UNTHROTTLE] += 1;
    } else if (type == REC_READ) {
      s.cnt[C_READ] += 1;
    }
    if (MODE == MODE_EMIT) {
      const uint64_t gi = tc.rec_base + tc.recs_before;
      P.rec_off[gi] = tc.stream_off + 8ull * ew;
      P.rec_type[gi] = type;
      P.rec_size[gi] = size;
    }
  }
  if (MODE == MODE_EMIT) tc.recs_before += 1;
  return size >> 3;
}

// Initialise the carry from a predecessor SAMPLE in global memory / the stream carry-in.
template <int PRE>
__device__ __forceinline__ void load_carry(Smem& s, const Params& P, uint32_t stream, uint64_t pred,
                                           uint32_t gap) {
  Carry& C = s.c[0];
  const uint32_t n = P.n_events;
  if (pred == PRED_CARRY_IN) {
    const hbtgpu_carry& c = P.carry_in[stream];
    for (uint32_t e = threadIdx.x; e < n; e += NT) C.last[e] = c.last[e];
    if (threadIdx.x == 0) {
      C.ts = c.tstamp;
      C.te = c.te;
      C.tr = c.tr;
      C.valid = c.valid != 0;
      C.gap = gap;
      C.last_off = PRED_CARRY_IN;
    }
  } else if (pred == PRED_UNKNOWN) {
    if (threadIdx.x == 0) {
      C.valid = 0;
      C.gap = 1;
      C.last_off = PRED_UNKNOWN;
    }
  } else {
    const uint64_t* r = reinterpret_cast<const uint64_t*>(P.arena + pred);
    for (uint32_t e = threadIdx.x; e < n; e += NT) C.last[e] = xform(P, e, r[PRE + 5 + e]);
    if (threadIdx.x == 0) {
      C.ts = r[PRE + 1];
      C.te = r[PRE + 3];
      C.tr = r[PRE + 4];
      C.valid = 1;
      C.gap = gap;
      C.last_off = pred;
    }
  }
}

#ifndef HBT_ODD_FIRST
#define HBT_ODD_FIRST 2  // other records of a tile followed one by one before the rest of it is indexed
#endif
#ifndef HBT_MIXED_MIN
#define HBT_MIXED_MIN 2
#endif
#ifndef HBT_MINB
#define HBT_MINB 5
#endif
template <bool SPEC, int MODE, int PRE, int NEV = 0>
__global__ void __launch_bounds__(NT, (PRE == 2 && MODE == MODE_AGG) ? 4 : HBT_MINB) k_decode(const Params P) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  Smem& s = *reinterpret_cast<Smem*>(smem_raw);
  if (!SPEC && MODE == MODE_AGG && P.flags[FLAG_NEED_EXACT] == 0) return;  // exact pass not needed
  const uint32_t tb = P.tile_bytes;
  const uint32_t stage_bytes = tb + OV_BYTES;
  uint64_t* const data0 = reinterpret_cast<uint64_t*>(smem_raw + decode_data_off<MODE, PRE>());
  uint64_t* const data1 = data0 + stage_bytes / 8;

  const uint32_t c = blockIdx.x;
  const ChunkDesc cd = P.chunks[c];
  ChunkIO& io = P.io[c];

  if (threadIdx.x == 0) {
    mbar_init(&s.bar[0], 1);
    mbar_init(&s.bar[1], 1);
    fence_barrier_init();
    s.tile_err = 0xFFFFFFFFu;
    s.lost_sum = 0;
  }
  if (threadIdx.x < C_SLOTS) s.cnt[threadIdx.x] = 0;
  if (MODE == MODE_AGG && PRE == 2) attr_cache_init();
  __syncthreads();

  uint32_t phase_bits = 0;  // parity of the two TMA barriers (a bit each: no local array)
  uint32_t cur = 0;  // which copy of the carry is current
  uint32_t bad = 0;  // fast mode: why this chunk's result is untrustworthy (WHY_*), 0 if fine

  // ---- where does the chain enter this chunk, and who is the predecessor sample? ----
  uint64_t entry_abs;
  const uint64_t chunk_abs = cd.img_base + cd.begin;
  uint64_t limit_abs = cd.img_base + cd.img_len;
  const uint64_t img_end_abs = cd.img_base + cd.img_len;
  const uint32_t avail_rel = static_cast<uint32_t>(umin64((img_end_abs - chunk_abs + 15ull) & ~15ull, 0xFFFFFFF0ull));
  const uint8_t* const chunk_ptr = P.arena + chunk_abs;
  auto issue = [&](uint32_t tile_rel, uint32_t b) {
    const uint32_t bytes = min(avail_rel - tile_rel, stage_bytes);
    mbar_expect_tx(&s.bar[b], bytes);
    tma_load_1d(b ? data1 : data0, chunk_ptr + tile_rel, bytes, &s.bar[b]);
  };
  // tile t lives in buffer (t + boff) & 1; with a halo in buffer 0 the first tile is fetched
  // into buffer 1 at once, so the two DRAM latencies at the start of a chunk overlap
  uint32_t boff = 0, n_issued = 0, n_waited = 0;
  if (SPEC) {
    if (cd.halo_len == 0) {
      entry_abs = chunk_abs;
      load_carry<PRE>(s, P, cd.stream, PRED_CARRY_IN, 0);
      __syncthreads();
    } else {
      // stage the halo and scout it
      const uint32_t hw = cd.halo_len / 8;
      if (threadIdx.x == 0) {
        // like a tile, the halo is staged with an overlap so that a record whose header is in
        // the halo but whose body runs into the chunk is fully readable (a halo taken from the
        // previous image ends on a record boundary and needs none)
        uint32_t bytes = (cd.halo_len + 15u) & ~15u;
        if (!(cd.flags & CH_FIRST_OF_IMAGE)) {
          const uint64_t avail = (cd.img_base + cd.img_len - cd.halo_off + 15ull) & ~15ull;
          bytes = static_cast<uint32_t>(umin64(cd.halo_len + OV_BYTES, avail));
        }
        mbar_expect_tx(&s.bar[0], bytes);
        tma_load_1d(data0, P.arena + cd.halo_off, bytes, &s.bar[0]);
        issue(0, 1);
      }
      boff = 1;
      n_issued = 1;
      mbar_wait(&s.bar[0], phase_bits & 1u);
      phase_bits ^= 1u;
      const uint64_t* hb = data0;
      load_carry<PRE>(s, P, cd.stream, PRED_UNKNOWN, 0);
      entry_abs = chunk_abs;
      // Quick scout: the last word of the halo that looks like a SAMPLE header of the configured
      // group, confirmed by the record before it, and whose record reaches the chunk.  It is a
      // guess like any other -- k_verify proves or refutes it against the previous chunk.
      bool quick = false;
      if (!P.force_general && !(cd.flags & CH_FIRST_OF_IMAGE)) {
        const uint32_t n = NEV ? static_cast<uint32_t>(NEV) : P.n_events, pre = PRE, S = 5 + pre + n;
        const uint64_t want = static_cast<uint64_t>(REC_SAMPLE) | (static_cast<uint64_t>(S * 8) << 48);
        if (threadIdx.x == 0) s.red_key = 0;
        __syncthreads();
        const uint32_t w = threadIdx.x;  // hw <= HALO_WORDS == NT
        const bool m = w < hw && (hb[w] & 0xFFFF0000FFFFFFFFull) == want && hb[w + pre + 2] == n;
        const uint32_t bal = __ballot_sync(0xffffffffu, m);
        if ((threadIdx.x & 31) == 0 && bal) atomicMax(&s.red_key, (threadIdx.x & ~31u) + (31 - __clz(bal)) + 1);
        __syncthreads();
        const uint32_t wl = s.red_key;  // 1 + last matching word, 0 if none
        if (wl != 0) {
          const uint32_t p = wl - 1;
          const bool confirmed = p >= S && (hb[p - S] & 0xFFFF0000FFFFFFFFull) == want &&
                                 hb[p - S + pre + 2] == n && hb[p - S + pre + 1] <= hb[p + pre + 1];
          if (confirmed && p + S >= hw) {
            quick = true;
            entry_abs = cd.halo_off + 8ull * (p + S);
            Carry& C0 = s.c[0];
            for (uint32_t e = threadIdx.x; e < n; e += NT) C0.last[e] = xform(P, e, hb[p + pre + 5 + e]);
            if (threadIdx.x == 0) {
              C0.ts = hb[p + pre + 1];
              C0.te = hb[p + pre + 3];
              C0.tr = hb[p + pre + 4];
              C0.valid = 1;
              C0.gap = 0;
              C0.last_off = cd.halo_off + 8ull * p;
            }
          }
        }
        __syncthreads();
      }
      uint32_t K = 0;
      uint32_t start = 0xFFFFFFFFu;
      if (!quick) {
        K = classify(s, hb, hw);
        if (K > 0) {
          if (cd.flags & CH_HALO_EXACT_START) start = 0;
          else start = scout_start<PRE>(s, P, hb, hw, K);
        }
      }
      __syncthreads();
      if (quick) {
        // nothing more to do
      } else if (start == 0xFFFFFFFFu) {
        bad = WHY_HALO_NO_START;
      } else {
        recover_chain(s, hb, hw, K, start, P.force_general != 0);
        if (s.term_kind != 0) {
          bad = WHY_HALO_DEAD;
        } else {
          entry_abs = cd.halo_off + 8ull * s.term_tgt;
          TileCtx tc{cd.halo_off, 0, cd.stream, 0, 0, 0, 0, nullptr};
          decode_tile<true, MODE_AGG, PRE>(s, P, hb, K, tc, 0);
          if (s.tile_err != 0xFFFFFFFFu) bad = WHY_HALO_ERR;
          else if (!s.c[0].valid) bad = WHY_HALO_NO_SAMPLE;  // predecessor unknown
        }
      }
      __syncthreads();
      if (threadIdx.x == 0) s.tile_err = 0xFFFFFFFFu;
      // a halo inside the previous image must end exactly on this image's first byte
      if ((cd.flags & CH_FIRST_OF_IMAGE) && !bad && entry_abs != cd.halo_off + cd.halo_len) bad = WHY_IMAGE_SEAM;
      if (cd.flags & CH_FIRST_OF_IMAGE) entry_abs = chunk_abs;
      __syncthreads();
    }
    if (threadIdx.x == 0) {
      io.in_entry = entry_abs;
      io.in_pred = s.c[0].last_off;
      io.in_gap = s.c[0].gap;
      io.limit = ~0ull;
    }
  } else {
    entry_abs = io.in_entry;
    limit_abs = min(limit_abs, io.limit);
    load_carry<PRE>(s, P, cd.stream, io.in_pred, io.in_gap);
    __syncthreads();
  }

  // ---- tiles ----  (positions are chunk-relative bytes in 32 bits: a chunk is <= 1 MiB and a
  // record can leave it by at most 64 KiB)
  const uint64_t chunk_end_abs = min(cd.img_base + cd.end, limit_abs);
  TileCtx tc{0, 0, cd.stream, 0, 0, MODE == MODE_EMIT ? io.row_base : 0, MODE == MODE_EMIT ? io.rec_base : 0,
             P.acc + static_cast<uint64_t>(cd.stream) * P.n_buckets * (P.n_events + 3)};
  const bool any_work = entry_abs < chunk_end_abs;  // also false for exact-mode "skip" (~0)
  const uint32_t chunk_len = any_work ? static_cast<uint32_t>(chunk_end_abs - chunk_abs) : 0;
  const uint32_t limit_rel = static_cast<uint32_t>(umin64(limit_abs - chunk_abs, 0xFFFFFFF0ull));
  uint32_t next_rel = any_work ? static_cast<uint32_t>(entry_abs - chunk_abs) : 0;

  if (any_work && n_issued == 0) {
    if (threadIdx.x == 0) issue(0, boff);
    n_issued = 1;
  }
  bool mixed_prev = false;  // the previous tile held other records than configured SAMPLEs
  bool dense_prev = false;  // ... half a dozen or more
  for (uint32_t tile_rel = 0, b = boff; tile_rel < chunk_len; tile_rel += tb, b ^= 1) {
    if (tile_rel + tb < chunk_len) {
      if (threadIdx.x == 0) issue(tile_rel + tb, b ^ 1);
      ++n_issued;
    }
    mbar_wait(&s.bar[b], (phase_bits >> b) & 1u);
    phase_bits ^= 1u << b;
    ++n_waited;
    const uint64_t* buf = b ? data1 : data0;
    const uint32_t wt = min(tb, chunk_len - tile_rel) / 8;
    uint32_t tile_kind = 0;
#ifdef HBT_DBG_LOAD_ONLY
    if (buf[threadIdx.x] == 0x123456789ull) s.cnt[0] = 1;
    next_rel = tile_rel + 8 * wt;
#endif
    if (next_rel < tile_rel + 8 * wt) {
      const uint32_t entry = (next_rel - tile_rel) / 8;
      const uint64_t tile_abs = chunk_abs + tile_rel;
      tc.tile_abs = tile_abs;
      if (MODE == MODE_EMIT) tc.stream_off = cd.stream_pos + (tile_abs - cd.img_base);
      uint32_t term_kind = 0, term_tgt = 0;
      bool done = false;
      uint32_t ew = entry;  // chain position, tile-relative words
      if (!P.force_general) {
        // runs of configured SAMPLEs in parallel, the few other records in between one by one;
        // a tile with many of those goes to the general path
        const uint32_t S = 5 + PRE + (NEV ? static_cast<uint32_t>(NEV) : P.n_events);
        for (int odd = 0;;) {
          // (after a mixed tile the next one is indexed right away: the plain attempt would
          // only prove the records up to the first other one)
          if (!(MODE == MODE_AGG && mixed_prev)) {
            uint32_t k_all = 0;
            const uint32_t k = uniform_segment<MODE, PRE, NEV>(s, P, buf, wt, ew, tc, cur, k_all);
            if (k) cur ^= 1;
            ew += k * S;
            if (k == k_all) {  // ran to the end of the tile
              done = true;
              break;
            }
            __syncthreads();  // the segment's carry is published
          }
          if (MODE == MODE_AGG && (odd >= HBT_ODD_FIRST || mixed_prev)) {
            // the tile holds other records too: index the rest of it once (index_tile), then
            // one pass over the listed samples, however many other records lie in between
            index_tile<PRE, NEV>(s, P, buf, wt, ew, cur, dense_prev);
            uint32_t n_idx = 0;
            const uint32_t ki = uniform_segment<MODE, PRE, NEV, true>(s, P, buf, wt, ew, tc, cur, n_idx);
            if (ki) cur ^= 1;
            ew = s.idx_end;
            // (the next tile is indexed right away only if this one held several other records:
            // after a lone LOST the next tile is most likely plain -- cfg3 76 -> 82 %, cfg5 78 -> 83 %)
            mixed_prev = s.idx_other >= HBT_MIXED_MIN;
            dense_prev = s.idx_other >= 6;  // many other records: the next tile is indexed by all warps
            if (s.idx_bad != 0xFFFFFFFFu) {
              // a record consume() stalls on: everything before it is accounted for, the general
              // path looks at it and owns the error reporting
              ew = s.idx_bad;
              __syncthreads();
              break;
            }
            done = true;
            break;
          }
          const uint32_t sz = (++odd <= 8) ? odd_record<MODE>(s, P, buf, ew, cur, tc) : 0;
          if (sz == 0) break;
          ew += sz;
          __syncthreads();  // thread 0's update of the carry (gap) is visible
          if (ew >= wt) {
            done = true;
            break;
          }
        }
        term_tgt = ew;
      }
      if (!done) {
        for (uint32_t w0 = 0; w0 < wt && term_kind == 0; w0 += GT_WORDS) {
          const uint32_t wh = min(static_cast<uint32_t>(GT_WORDS), wt - w0);
          if (ew >= w0 + wh) continue;  // a long record jumps over this sub-tile
          const uint64_t* sb = buf + w0;
          const uint32_t K = classify(s, sb, wh);
          recover_chain(s, sb, wh, K, ew - w0, P.force_general != 0);
          const uint32_t tk = s.term_kind, tt = s.term_tgt;
          tc.tile_abs = tile_abs + 8ull * w0;
          if (MODE == MODE_EMIT) tc.stream_off = cd.stream_pos + (tile_abs - cd.img_base) + 8ull * w0;
          if (MODE == MODE_AGG) {
            // one aggregation routine for every path: list the samples, then the indexed pass
            list_tile<PRE>(s, P, sb, K, cur);
            uint32_t n_idx = 0;
            const uint32_t ki = uniform_segment<MODE, PRE, NEV, true>(s, P, sb, wh, 0, tc, cur, n_idx);
            if (ki) cur ^= 1;
            __syncthreads();  // the new carry and s.tile_err are published
          } else {
            decode_tile<false, MODE, PRE>(s, P, sb, K, tc, cur);
          }
          term_kind = (tk != 0 || s.tile_err != 0xFFFFFFFFu) ? 1u + tk : 0u;
          ew = w0 + tt;
        }
        term_tgt = ew;
      }
      // decided after the barrier below: s.tile_err may still be written by other warps
      tile_kind = term_kind;
      next_rel = tile_rel + 8 * term_tgt;
    }
    __syncthreads();  // buffer b is free for the tile after next; publishes the new carry
    if (tile_kind != 0 || s.tile_err != 0xFFFFFFFFu) {
      bad = tile_kind > 1 ? WHY_TILE_DEAD : WHY_TILE_ERR;
      break;
    }
    if (next_rel > limit_rel) {  // record runs past the image end (or the exact-mode limit)
      bad = WHY_PAST_END;
      break;
    }
  }
  // never exit with a bulk copy still in flight into this CTA's shared memory
  if (n_waited < n_issued) mbar_wait(&s.bar[(n_waited + boff) & 1], (phase_bits >> ((n_waited + boff) & 1)) & 1u);
  const uint64_t next_abs = any_work ? chunk_abs + next_rel : entry_abs;

  // ---- publish ----
  __syncthreads();
  if (MODE == MODE_AGG && PRE == 2) attr_cache_flush(P);
  const Carry& C = s.c[cur];
  // (the chunk's stream and flags are read again here rather than kept in registers through the
  // tile loop, which has none to spare)
  const uint32_t c_end = blockIdx.x;
  const uint32_t cd_stream = *reinterpret_cast<const volatile uint32_t*>(&P.chunks[c_end].stream);
  const uint32_t cd_flags = *reinterpret_cast<const volatile uint32_t*>(&P.chunks[c_end].flags);
  if (threadIdx.x == 0) {
    ChunkIO& ioe = P.io[c_end];
    ioe.out_exit = next_abs;
    ioe.out_last = C.last_off;
    ioe.out_gap = C.gap;
    ioe.n_records = s.cnt[C_RECORDS];
    ioe.n_rows = s.cnt[C_ROWS];
    if (bad) {
      if (SPEC) raise_exact(P.flags, bad, c_end);
      else atomicExch(&P.flags[FLAG_INTERNAL], bad);
    }
  }
  if (MODE == MODE_AGG) {
    uint64_t* ctr = P.counters + static_cast<uint64_t>(cd_stream) * C_SLOTS;
    if (threadIdx.x < C_SLOTS && threadIdx.x != C_LOST && s.cnt[threadIdx.x])
      red_add(&ctr[threadIdx.x], s.cnt[threadIdx.x]);
    if (threadIdx.x == 0 && s.lost_sum) red_add(&ctr[C_LOST], s.lost_sum);
    // stream carry-out: in exact mode k_resolve owns it
    if (SPEC && (cd_flags & CH_LAST_OF_STREAM)) {
      hbtgpu_carry& co = P.carry_out[cd_stream];
      for (uint32_t e = threadIdx.x; e < P.n_events; e += NT) co.last[e] = C.last[e];
      if (threadIdx.x == 0) {
        const bool v = C.valid && !C.gap;
        co.valid = v;
        co.tstamp = v ? C.ts : ~0ull;
        co.te = C.te;
        co.tr = C.tr;
        ctr[C_LAST_TS] = co.tstamp;
        ctr[C_LAST_SAMPLE_TS] = C.valid ? C.ts : 0;  // time of the last SAMPLE, LOST behind it or not
      }
    }
  }
}

// ------------------------------------------------------------------------------------
// k_verify: one thread per chunk.  Induction over stream order: chunk 0 of a stream is exact
// by construction; chunk c is exact if what it assumed equals what chunk c-1 produced.
__global__ void k_verify(const Params P) {
  const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= P.n_chunks) return;
  const ChunkDesc cd = P.chunks[c];
  const ChunkIO& io = P.io[c];
  uint32_t why = 0;
  if (cd.flags & CH_FIRST_OF_STREAM) {
    if (!(io.in_entry == cd.img_base + cd.begin && io.in_pred == PRED_CARRY_IN && io.in_gap == 0))
      why = WHY_V_FIRST;
  } else {
    const ChunkDesc pd = P.chunks[c - 1];
    const ChunkIO& pio = P.io[c - 1];
    if (cd.flags & CH_FIRST_OF_IMAGE) {
      if (!(pio.out_exit == pd.img_base + pd.img_len && io.in_entry == cd.img_base)) why = WHY_V_ENTRY;
    } else if (pio.out_exit != io.in_entry) {
      why = WHY_V_ENTRY;
    }
    if (!why && pio.out_last != io.in_pred) why = WHY_V_PRED;
    if (!why && pio.out_gap != io.in_gap) why = WHY_V_GAP;
  }
  if (!why && (cd.flags & CH_LAST_OF_IMAGE) && io.out_exit != cd.img_base + cd.img_len)
    why = WHY_V_IMAGE_END;
  if (why) raise_exact(P.flags, why, c);
}

// ------------------------------------------------------------------------------------
// k_resolve: exact sequential walk, one warp per stream (lane 0 walks; the warp stages
// 4 KiB windows into shared memory so the dependent loads hit on-chip memory).
constexpr int RES_WIN_BYTES = 4096;
constexpr int RES_OV_BYTES = 64;  // header + the fields the walk reads (time, nr / num_lost)
constexpr int RES_THREADS = 32;   // one walker (lane 0) per block

// k_resolve: exact sequential walk, one walker per stream.  Windows of 4 KiB (+64 B overlap) are
// brought into shared memory with TMA bulk copies, double buffered, so the dependent loads of
// the walk hit on-chip memory while the next window is in flight.
__global__ void __launch_bounds__(RES_THREADS) k_resolve(const Params P) {
  if (P.flags[FLAG_NEED_EXACT] == 0) return;
  __shared__ __align__(128) uint64_t win[2][(RES_WIN_BYTES + RES_OV_BYTES) / 8];
  __shared__ __align__(8) uint64_t bar[2];
  const uint32_t sidx = blockIdx.x;
  if (sidx >= P.n_streams) return;
  // The whole warp runs the walk with identical state (reads of shared memory are broadcasts);
  // lane 0 alone issues copies and writes results.  Runs of configured SAMPLEs are skipped 32
  // records at a time: lane l tests the record at pos + l*S.
  const uint32_t lane = threadIdx.x;
  const bool writer = lane == 0;
  if (writer) {
    mbar_init(&bar[0], 1);
    mbar_init(&bar[1], 1);
    fence_barrier_init();
  }
  __syncwarp();
  uint32_t phase[2] = {0, 0};
  const uint32_t pre = P.pre;
  const uint32_t S8 = 8 * (5 + pre + P.n_events);  // bytes of a configured SAMPLE
  const uint64_t want = static_cast<uint64_t>(REC_SAMPLE) | (static_cast<uint64_t>(S8) << 48);
  const StreamDesc sd = P.streams[sidx];
  const uint32_t n = P.n_events;
  uint64_t* ctr = P.counters + static_cast<uint64_t>(sidx) * C_SLOTS;

  const hbtgpu_carry& cin = P.carry_in[sidx];
  bool valid = cin.valid != 0;        // count_.tstamp != kInvalidTimeStamp
  uint64_t st_ts = cin.tstamp;        // count_.tstamp
  uint64_t last_off = PRED_CARRY_IN;  // last SAMPLE seen
  bool gap = false;                   // LOST after last_off
  int64_t err = 0;
  uint64_t err_off = 0;
  bool stopped = false;

  for (uint32_t ii = 0; ii < sd.n_images; ++ii) {
    const ImageDesc im = P.images[sd.first_image + ii];
    uint32_t ch = im.first_chunk;
    const uint32_t ch_end = im.first_chunk + im.n_chunks;
    uint64_t next_begin = ch < ch_end ? P.chunks[ch].begin : ~0ull;
    uint64_t pos = 0;  // image-relative
    const uint8_t* const img = P.arena + im.base;
    auto issue = [&](uint64_t wb, uint32_t b) {
      const uint64_t avail = (im.len - wb + 15ull) & ~15ull;
      const uint32_t bytes = static_cast<uint32_t>(umin64(avail, RES_WIN_BYTES + RES_OV_BYTES));
      mbar_expect_tx(&bar[b], bytes);
      tma_load_1d(win[b], img + wb, bytes, &bar[b]);
    };
    uint32_t issued = 0, waited = 0;
    if (!stopped && im.len > 0) {
      if (writer) issue(0, 0);
      issued = 1;
    }
    for (uint64_t wb = 0; !stopped && wb < im.len; wb += RES_WIN_BYTES) {
      const uint32_t b = static_cast<uint32_t>(wb / RES_WIN_BYTES) & 1;
      __syncwarp();  // every lane is done with buffer b^1
      if (wb + RES_WIN_BYTES < im.len) {
        if (writer) issue(wb + RES_WIN_BYTES, b ^ 1);
        ++issued;
      }
      mbar_wait(&bar[b], phase[b]);
      phase[b] ^= 1;
      ++waited;
      const uint64_t* wd = win[b];
      const uint64_t wend = umin64(wb + RES_WIN_BYTES, im.len);
      while (pos < wend) {
        // ---- a run of configured SAMPLEs: 32 at a time ----
        {
          const uint64_t q = pos + static_cast<uint64_t>(lane) * S8;
          bool ok = q < wend && q + S8 <= im.len;
          uint64_t tsl = 0;
          if (ok) {
            const uint32_t qi = static_cast<uint32_t>((q - wb) / 8);
            ok = (wd[qi] & 0xFFFF0000FFFFFFFFull) == want && wd[qi + pre + 2] == n;
            tsl = wd[qi + pre + 1];
          }
          const uint64_t pts = __shfl_up_sync(0xffffffffu, tsl, 1);
          if (lane == 0) ok = ok && !(valid && !gap && st_ts > tsl);
          else ok = ok && pts <= tsl;
          const uint32_t okm = __ballot_sync(0xffffffffu, ok);
          const uint32_t run = okm == 0xFFFFFFFFu ? 32u : static_cast<uint32_t>(__ffs(~okm) - 1);  // leading lanes that hold
          if (run > 0) {
            const uint64_t last_rec = pos + static_cast<uint64_t>(run - 1) * S8;
            while (next_begin <= last_rec) {  // chunks that begin inside the run
              const uint64_t j = next_begin <= pos ? 0 : (next_begin - pos + S8 - 1) / S8;
              if (writer) {
                ChunkIO& io = P.io[ch];
                io.in_entry = im.base + pos + j * S8;
                io.in_pred = j ? im.base + pos + (j - 1) * S8 : last_off;
                io.in_gap = j ? 0u : static_cast<uint32_t>(gap);
                io.limit = ~0ull;
              }
              ++ch;
              next_begin = ch < ch_end ? P.chunks[ch].begin : ~0ull;
            }
            st_ts = __shfl_sync(0xffffffffu, tsl, run - 1);
            valid = true;
            gap = false;
            last_off = im.base + last_rec;
            pos = last_rec + S8;
            continue;
          }
        }
        // ---- any other record: one step ----
        const uint32_t wi = static_cast<uint32_t>((pos - wb) / 8);
        const uint64_t h = wd[wi];
        const uint32_t type = static_cast<uint32_t>(h);
        const uint32_t size = static_cast<uint32_t>(h >> 48);
        // hand the chunks that begin at or before this record their entry
        while (pos >= next_begin) {
          if (writer) {
            ChunkIO& io = P.io[ch];
            io.in_entry = im.base + pos;
            io.in_pred = last_off;
            io.in_gap = gap;
            io.limit = ~0ull;
          }
          ++ch;
          next_begin = ch < ch_end ? P.chunks[ch].begin : ~0ull;
        }
        bool bad = size < 8 || (size & 7) || pos + size > im.len;  // fences, see DESIGN.md
        uint64_t ts = 0;
        if (!bad && type == REC_SAMPLE) {
          const uint64_t nr = (size + 16 >= P.fix) ? wd[wi + pre + 2] : ~0ull;
          ts = wd[wi + pre + 1];
          bad = size < P.fix || nr != ((size - P.fix) >> 3) || nr != n;
        }
        if (!bad && type == REC_READ) {
          const uint64_t nr = (size >= 24) ? wd[wi + 2] : ~0ull;
          bad = size < 40 || nr != ((size - 40u) >> 3);
        }
        if (!bad && type == REC_LOST) bad = size < 24;
        if (!bad && type == REC_SAMPLE && valid && !gap && st_ts > ts) {
          bad = true;     // time regression: state is invalidated (gen :156-163)
          valid = false;
        }
        if (bad) {
          err = -1;  // -EPERM
          err_off = im.stream_pos + pos;
          stopped = true;
          break;
        }
        if (type == REC_SAMPLE) {
          st_ts = ts;
          valid = true;
          gap = false;
          last_off = im.base + pos;
        } else if (type == REC_LOST) {
          gap = true;
        }
        pos += size;
      }
    }
    // never leave a bulk copy in flight
    if (waited < issued) {
      const uint32_t b = waited & 1;
      mbar_wait(&bar[b], phase[b]);
      phase[b] ^= 1;
    }
    __syncwarp();
    // chunks the walk never reached (after an error, or jumped over by a long record)
    const uint64_t stop_abs = im.base + pos;
    for (uint32_t k = im.first_chunk + lane; k < ch_end; k += 32) {
      ChunkIO& io = P.io[k];
      if (k >= ch) {
        io.in_entry = stopped ? ~0ull : stop_abs;
        io.in_pred = last_off;
        io.in_gap = gap;
      }
      io.limit = stopped ? stop_abs : ~0ull;
    }
  }
  // images after an error were visited above with stopped=true: in_entry = ~0, nothing decoded
  if (!writer) return;
  hbtgpu_carry& co = P.carry_out[sidx];
  const bool v = valid && !gap;
  if (last_off == PRED_CARRY_IN) {
    for (uint32_t e = 0; e < n; ++e) co.last[e] = cin.last[e];
    co.te = cin.te;
    co.tr = cin.tr;
  } else {
    const uint64_t* r = reinterpret_cast<const uint64_t*>(P.arena + last_off);
    for (uint32_t e = 0; e < n; ++e) co.last[e] = xform(P, e, r[pre + 5 + e]);
    co.te = r[pre + 3];
    co.tr = r[pre + 4];
  }
  co.valid = v;
  co.tstamp = v ? st_ts : ~0ull;
  ctr[C_ERR] = static_cast<uint64_t>(err);
  ctr[C_ERR_OFF] = err_off;
  ctr[C_LAST_TS] = co.tstamp;
  ctr[C_LAST_SAMPLE_TS] = last_off == PRED_CARRY_IN ? 0 : st_ts;
}

// zero result buffers when the exact path has to recompute them
__global__ void k_clear(const Params P, uint64_t acc_words) {
  if (P.flags[FLAG_NEED_EXACT] == 0) return;
  const uint64_t stride = static_cast<uint64_t>(gridDim.x) * blockDim.x;
  const uint64_t i0 = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  for (uint64_t i = i0; i < acc_words; i += stride) P.acc[i] = 0;
  const uint64_t nctr = static_cast<uint64_t>(P.n_streams) * C_SLOTS;
  for (uint64_t i = i0; i < nctr; i += stride)
    if ((i % C_SLOTS) < C_ERR) P.counters[i] = 0;
  if (P.pre == 2) {
    for (uint64_t i = i0; i <= P.tid_mask; i += stride) {
      P.tid_keys[i] = ATTR_EMPTY;
      P.tid_cnt[i] = 0;
    }
    for (uint64_t i = i0; i <= P.ip_mask; i += stride) {
      P.ip_keys[i] = ATTR_EMPTY;
      P.ip_cnt[i] = 0;
    }
    if (i0 < 4) P.attr_special[i0] = 0;
  }
}

// The bucket sums start at zero.  Written from the end to the start: what is written last stays
// in L2, and the decode starts with the first cpu's records -- its first REDs find their lines
// there instead of reading them back from DRAM.
__global__ void __launch_bounds__(256) k_zero_desc(ulonglong2* p, uint64_t n16) {
  const uint64_t stride = static_cast<uint64_t>(gridDim.x) * blockDim.x;
  for (uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n16; i += stride)
    p[n16 - 1 - i] = make_ulonglong2(0, 0);
}

// what a run starts from, in one launch: zeroed counters and flags, carry_out := carry_in for
// streams without images (shard mode: every stream starts from an invalid state)
__global__ void __launch_bounds__(256) k_prologue(const Params P, hbtgpu_carry* carry_in, uint32_t force_exact) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < P.n_streams * C_SLOTS) P.counters[i] = 0;
  if (i < FLAG_COUNT) P.flags[i] = (i == FLAG_NEED_EXACT && force_exact) ? 1u : 0u;
  constexpr uint32_t W = sizeof(hbtgpu_carry) / 8;
  if (i < P.n_streams * W) {
    uint64_t* in = reinterpret_cast<uint64_t*>(carry_in);
    uint64_t* out = reinterpret_cast<uint64_t*>(P.carry_out);
    if (P.external_carry) in[i] = 0;
    out[i] = in[i];
  }
}

// ------------------------------------------------------------------------------------
// top-K of the sampled-IP table by (count descending, ip ascending): compact the table; the host
// side sorts the pairs (hbtgpu_get_top_ips).
__global__ void k_attr_compact(const uint64_t* keys, const unsigned long long* cnt, uint32_t mask,
                               uint64_t* out_key, unsigned long long* out_cnt, unsigned int* n_out) {
  const uint64_t stride = static_cast<uint64_t>(gridDim.x) * blockDim.x;
  for (uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i <= mask; i += stride) {
    const uint64_t k = keys[i];
    if (k != ATTR_EMPTY) {
      const unsigned int o = atomicAdd(n_out, 1u);
      out_key[o] = k;
      out_cnt[o] = cnt[i];
    }
  }
}

// HBTGPU_F_ACCUMULATE_ATTR: add one run's table to the totals (after the exact path, if it ran,
// has settled the run's counts)
__global__ void k_attr_merge(const uint64_t* keys, const unsigned long long* cnt, uint32_t mask, uint64_t* tot_keys,
                             unsigned long long* tot_cnt, uint32_t tot_mask, unsigned long long* overflow) {
  const uint64_t stride = static_cast<uint64_t>(gridDim.x) * blockDim.x;
  for (uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i <= mask; i += stride) {
    const uint64_t k = __ldcv(keys + i);  // (the source may be a peer's memory: never through a stale line)
    if (k == ATTR_EMPTY) continue;
    const uint32_t slot = attr_hash(k) & tot_mask;
    attr_add_from(tot_keys, tot_cnt, tot_mask, k, __ldcv(cnt + i), overflow, slot, reinterpret_cast<volatile uint64_t*>(tot_keys)[slot]);
  }
}
__global__ void k_attr_merge_special(const unsigned long long* run, unsigned long long* tot) {
  if (threadIdx.x < 4) tot[threadIdx.x] += __ldcv(run + threadIdx.x);
}

// ------------------------------------------------------------------------------------
// Shard mode (HBTGPU_F_EXTERNAL_CARRY): head stubs and boundary rows.

// One thread per stream: walk from the stream start to its first SAMPLE (normally the very
// first record).  The stream start is a record boundary, so this walk is exact.  The records it
// steps over are counted as consume() counts them (HeadExtra).
__global__ void k_head(const Params P) {
  const uint32_t sidx = blockIdx.x * blockDim.x + threadIdx.x;
  if (sidx >= P.n_streams) return;
  hbtgpu_head_stub& hs = P.head[sidx];
  HeadExtra hx{0, 0, 0, 0, 0};
  hs.has_sample = 0;
  hs.lost_before = 0;
  hs.offset = 0;
  const StreamDesc sd = P.streams[sidx];
  const uint32_t n = P.n_events;
  bool done = false;
  for (uint32_t ii = 0; ii < sd.n_images && !done; ++ii) {
    const ImageDesc im = P.images[sd.first_image + ii];
    uint64_t pos = 0;
    while (pos < im.len) {
      const uint64_t* r = reinterpret_cast<const uint64_t*>(P.arena + im.base + pos);
      const uint64_t h = r[0];
      const uint32_t type = static_cast<uint32_t>(h), size = static_cast<uint32_t>(h >> 48);
      if (size < 8 || (size & 7) || pos + size > im.len) {  // the decode reports the error
        done = true;
        break;
      }
      if (type == REC_SAMPLE) {
        const uint32_t pre = P.pre;
        if (!(size < P.fix || r[pre + 2] != ((size - P.fix) >> 3) || r[pre + 2] != n)) {
          hs.tstamp = r[pre + 1];
          hs.te = r[pre + 3];
          hs.tr = r[pre + 4];
          for (uint32_t e = 0; e < n; ++e) hs.value[e] = xform(P, e, r[pre + 5 + e]);
          hs.offset = im.stream_pos + pos;
          hs.has_sample = 1;
        }
        done = true;
        break;
      }
      if (type == REC_LOST) {
        if (size < 24) {
          done = true;
          break;
        }
        hs.lost_before = 1;
        hx.lost += r[2];
      } else if (type == REC_READ) {
        if (size < 40 || r[2] != ((size - 40u) >> 3)) {
          done = true;
          break;
        }
        hx.read += 1;
      } else if (type == REC_THROTTLE) {
        hx.throttle += 1;
      } else if (type == REC_UNTHROTTLE) {
        hx.unthrottle += 1;
      }
      hx.records += 1;
      pos += size;
    }
  }
  if (P.head_extra) P.head_extra[sidx] = hx;
}

// ---- packed stubs: what one rank tells the others about its shard, per cpu ----------------
// u64 words: [0] flags  [1] stream offset of the first SAMPLE  [2] bytes of the shard's stream
//            [3..5] first SAMPLE: time, time_enabled, time_running
//            [6..8] state after the shard: count_.tstamp (~0 if invalid), te, tr of the last SAMPLE
//            [9] time of the last SAMPLE  [10..10+n) values of the first SAMPLE  [10+n..10+2n) last_[]
enum : uint64_t { SF_HAS_SAMPLE = 1, SF_LOST_BEFORE = 2, SF_TAIL_VALID = 4, SF_STALLED = 8 };
enum { SW_FLAGS = 0, SW_HEAD_OFF, SW_STREAM_LEN, SW_HEAD_TS, SW_HEAD_TE, SW_HEAD_TR, SW_TAIL_TS, SW_TAIL_TE,
       SW_TAIL_TR, SW_LAST_SAMPLE_TS, SW_VALUES };
__host__ __device__ inline uint32_t stub_words(uint32_t n_events) { return SW_VALUES + 2 * n_events; }

// word w of the packed stub of stream `cpu` (any thread may ask for any word)
__device__ __forceinline__ uint64_t stub_word(const Params& P, uint32_t cpu, uint32_t w) {
  const hbtgpu_head_stub& hs = P.head[cpu];
  const hbtgpu_carry& co = P.carry_out[cpu];
  const uint64_t* ctr = P.counters + static_cast<uint64_t>(cpu) * C_SLOTS;
  const uint32_t n = P.n_events;
  switch (w) {
    case SW_FLAGS:
      return (hs.has_sample ? SF_HAS_SAMPLE : 0) | (hs.lost_before ? SF_LOST_BEFORE : 0) |
             (co.valid ? SF_TAIL_VALID : 0) | (ctr[C_ERR] ? SF_STALLED : 0);
    case SW_HEAD_OFF: return hs.offset;
    case SW_STREAM_LEN: {
      const StreamDesc sd = P.streams[cpu];
      uint64_t len = 0;
      for (uint32_t i = 0; i < sd.n_images; ++i) len += P.images[sd.first_image + i].len;
      return len;
    }
    case SW_HEAD_TS: return hs.tstamp;
    case SW_HEAD_TE: return hs.te;
    case SW_HEAD_TR: return hs.tr;
    case SW_TAIL_TS: return co.tstamp;
    case SW_TAIL_TE: return co.te;
    case SW_TAIL_TR: return co.tr;
    case SW_LAST_SAMPLE_TS: return ctr[C_LAST_SAMPLE_TS];
    default: return w < SW_VALUES + n ? hs.value[w - SW_VALUES] : co.last[w - SW_VALUES - n];
  }
}

// this rank's block of packed stubs -> dst (device memory of this GPU or, over NVLink, of a peer)
__global__ void __launch_bounds__(256) k_pack_stubs(const Params P, uint64_t* dst) {
  const uint32_t sw = stub_words(P.n_events);
  const uint32_t total = P.n_streams * sw;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const uint32_t cpu = i / sw;
    dst[i] = stub_word(P, cpu, i - cpu * sw);
  }
}

// Replay of the generator's state over the shards in rank order: one thread per cpu.
// gathered = `world` blocks of n_streams packed stubs.  Mirrors handleRecordLost /
// handleRecordSample (PerCpuCountSampleGenerator.h:110-124,135-163,198-199): a shard with a
// SAMPLE determines the state by itself; one without passes the previous state through,
// invalidated if it holds a LOST.  The walk of a cpu STALLS for good at the first record it
// fails on (consume() keeps returning the error, :1505,1521, gen :162): at a shard's own decode
// error, or at a shard's first SAMPLE when it goes back in time against the state before it.
// Shards behind a stall contribute nothing.  Outputs for `rank`: the state in front of its shard
// and dead[cpu] (1: an earlier shard stalled); for every rank r: the bucket range its rows can
// have touched (ranges[r * n_streams + cpu] = lo | hi << 32, lo > hi: none), if ranges != null.
// tid / nthr: the calling thread's index among the nthr threads (1 or a warp) that work on this
// cpu together; all of them follow the same control flow, the copies of last_[] are split.
__device__ __forceinline__ void fold_cpu(const Params& P, uint32_t cpu, const uint64_t* gathered, uint32_t world,
                                         uint32_t rank, const hbtgpu_carry* init, hbtgpu_carry* prev_out,
                                         uint32_t* dead_out, uint64_t* ranges, uint32_t tid, uint32_t nthr) {
  const uint32_t ns = P.n_streams;
  const uint32_t n = P.n_events, sw = stub_words(n);
  // the state is (valid, tstamp) plus a pointer to where te / tr / last[] come from
  bool valid = init[cpu].valid != 0;
  uint64_t tstamp = init[cpu].tstamp;
  const uint64_t* src = nullptr;  // packed stub whose tail holds te, tr, last[]; null: init
  bool dead = false;
  auto emit = [&]() {  // the state as it is now -> prev_out[cpu]
    hbtgpu_carry& d = prev_out[cpu];
    if (tid == 0) {
      d.valid = valid ? 1 : 0;
      d.tstamp = valid ? tstamp : ~0ull;
      d.te = src ? src[SW_TAIL_TE] : init[cpu].te;
      d.tr = src ? src[SW_TAIL_TR] : init[cpu].tr;
      dead_out[cpu] = dead ? 1u : 0u;
    }
    for (uint32_t e = tid; e < n; e += nthr) d.last[e] = src ? src[SW_VALUES + n + e] : init[cpu].last[e];
  };
  for (uint32_t r = 0; r < world; ++r) {
    if (r == rank) {
      emit();
      if (!ranges) return;
    }
    const uint64_t* st = gathered + (static_cast<size_t>(r) * ns + cpu) * sw;
    const uint64_t fl = st[SW_FLAGS];
    uint32_t lo = 1, hi = 0;
    if (!dead) {
      const bool has = (fl & SF_HAS_SAMPLE) != 0, lostb = (fl & SF_LOST_BEFORE) != 0;
      const uint64_t head_ts = st[SW_HEAD_TS];
      if (has && !lostb && valid && tstamp > head_ts) {
        dead = true;  // stalls on this shard's first SAMPLE: nothing of the shard is aggregated
      } else {
        if (has) {
          // rows of the shard lie between its first and its last SAMPLE
          const uint64_t last_ts = st[SW_LAST_SAMPLE_TS];
          const uint64_t d0 = head_ts - P.t0, d1 = last_ts - P.t0;
          const bool below0 = head_ts < P.t0, below1 = last_ts < P.t0;
          if (!below1 && !(!below0 && d0 >= P.range_ns)) {
            lo = below0 ? 0u : bucket_of(P, head_ts);
            hi = d1 >= P.range_ns ? P.n_buckets - 1 : bucket_of(P, last_ts);
          }
          src = st;
          valid = (fl & SF_TAIL_VALID) != 0;
          tstamp = st[SW_TAIL_TS];
        } else if (lostb) {
          valid = false;
        }
        if (fl & SF_STALLED) dead = true;  // the shard's own walk stalled: later shards are never reached
      }
    }
    if (ranges && tid == 0)
      ranges[static_cast<size_t>(r) * ns + cpu] = static_cast<uint64_t>(lo) | (static_cast<uint64_t>(hi) << 32);
  }
  if (rank >= world) emit();  // the state after the last shard (no rank owns it)
}

__global__ void k_fold(const Params P, const uint64_t* gathered, uint32_t world, uint32_t rank,
                       const hbtgpu_carry* init, hbtgpu_carry* prev_out, uint32_t* dead_out,
                       uint64_t* ranges) {
  const uint32_t cpu = blockIdx.x * blockDim.x + threadIdx.x;
  if (cpu < P.n_streams) fold_cpu(P, cpu, gathered, world, rank, init, prev_out, dead_out, ranges, 0, 1);
}

// One block per stream: add the row that straddles the shard boundary -- or, where the walk of
// this cpu stalled before or at this shard's first SAMPLE, take the shard's contribution out.
// dead: per cpu, 1 if an earlier shard stalled (null: none did).
// tid / nthr: the calling thread's index among the nthr threads that work on this stream together
__device__ __forceinline__ void boundary_cpu(const Params& P, uint32_t sidx, const hbtgpu_carry* prev_tails,
                                             const uint32_t* dead, uint64_t* touched_cell, uint32_t tid,
                                             uint32_t nthr) {
  if (tid == 0) touched_cell[sidx] = ~0ull;  // which (cpu, bucket) cell got the boundary row
  const hbtgpu_head_stub& hs = P.head[sidx];
  const hbtgpu_carry& pt = prev_tails[sidx];
  uint64_t* ctr = P.counters + static_cast<uint64_t>(sidx) * C_SLOTS;
  const uint32_t n = P.n_events;
  const bool is_dead = dead && dead[sidx];
  const bool row = hs.has_sample && !hs.lost_before && pt.valid;
  // handleRecordSample emits a row iff the state is valid when the SAMPLE arrives (:135,165-182)
  const bool reached = !(ctr[C_ERR] != 0 && ctr[C_ERR_OFF] <= hs.offset);  // the shard's own walk got to that SAMPLE
  const bool regress = !is_dead && row && reached && pt.tstamp > hs.tstamp;  // (:156-163)
  // every thread has read the counters before thread 0 rewrites them
  if (nthr == 32) __syncwarp();
  else if (nthr > 1) __syncthreads();
  if (is_dead || regress) {
    // consume() never gets past the stall: of this shard only the records in front of its first
    // SAMPLE (regress) or nothing at all (dead) were walked
    uint64_t* acc = P.acc + static_cast<uint64_t>(sidx) * P.n_buckets * (n + 3);
    const uint64_t words = static_cast<uint64_t>(P.n_buckets) * (n + 3);
    for (uint64_t i = tid; i < words; i += nthr) acc[i] = 0;
    if (tid == 0) {
      HeadExtra hx{0, 0, 0, 0, 0};
      if (regress && P.head_extra) hx = P.head_extra[sidx];
      ctr[C_LOST] = hx.lost;
      ctr[C_THROTTLE] = hx.throttle;
      ctr[C_UNTHROTTLE] = hx.unthrottle;
      ctr[C_READ] = hx.read;
      ctr[C_RECORDS] = hx.records;
      ctr[C_SAMPLE] = 0;
      ctr[C_UNKNOWN] = 0;
      ctr[C_ROWS] = 0;
      ctr[C_OOR] = 0;
      ctr[C_ERR] = static_cast<uint64_t>(-1);
      ctr[C_ERR_OFF] = regress ? hs.offset : 0;
      // the state stays what it was in front of the shard; a regression invalidates count_.tstamp
      hbtgpu_carry& co = P.carry_out[sidx];
      co.valid = regress ? 0 : pt.valid;
      co.tstamp = (regress || !pt.valid) ? ~0ull : pt.tstamp;
      co.te = pt.te;
      co.tr = pt.tr;
      ctr[C_LAST_TS] = co.tstamp;
      touched_cell[sidx] = ~0ull - 1;  // "the whole row changed": derive_after_boundaries re-derives everything
    }
    for (uint32_t e = tid; e < n; e += nthr) P.carry_out[sidx].last[e] = pt.last[e];
    return;
  }
  if (!hs.has_sample) {
    // no SAMPLE in the shard: the state in front of it passes through (a LOST invalidates it)
    hbtgpu_carry& co = P.carry_out[sidx];
    for (uint32_t e = tid; e < n; e += nthr) co.last[e] = pt.last[e];
    if (tid == 0) {
      const bool v = pt.valid && !hs.lost_before;
      co.valid = v;
      co.tstamp = v ? pt.tstamp : ~0ull;
      co.te = pt.te;
      co.tr = pt.tr;
      ctr[C_LAST_TS] = co.tstamp;
    }
    return;
  }
  if (!row || !reached) return;
  const uint32_t bk = bucket_of(P, hs.tstamp);
  if (tid == 0) {
    ctr[C_ROWS] += 1;
    ctr[C_UNKNOWN] -= 1;  // the shard counted its first SAMPLE as a re-base
    if (bk == BKT_OOR) ctr[C_OOR] += 1;
  }
  if (bk == BKT_OOR) return;
  if (tid == 0) touched_cell[sidx] = static_cast<uint64_t>(sidx) * P.n_buckets + bk;
  uint64_t* base = P.acc + (static_cast<uint64_t>(sidx) * P.n_buckets + bk) * (n + 3);
  for (uint32_t e = tid; e < n + 3; e += nthr) {
    uint64_t v;
    if (e < n) v = hs.value[e] - pt.last[e];
    else if (e == n) v = hs.te - pt.te;
    else if (e == n + 1) v = hs.tr - pt.tr;
    else v = 1;
    red_add(base + e, v);
  }
}

__global__ void __launch_bounds__(128) k_boundary(const Params P, const hbtgpu_carry* prev_tails,
                                                  const uint32_t* dead, uint64_t* touched_cell) {
  boundary_cpu(P, blockIdx.x, prev_tails, dead, touched_cell, threadIdx.x, blockDim.x);
}

// ------------------------------------------------------------------------------------
// k_derive: per (cpu,bucket): scaled = getCount(acc), metrics = reducers(scaled)
struct DeriveParams {
  const uint64_t* acc;
  uint64_t* scaled;
  double* metric;
  uint32_t* valid;
  uint64_t n_cells;  // cpus * buckets
  uint32_t n_events, n_metrics;
  uint64_t n_magic;  // ceil(2^32 / n_events): x / n == (x * n_magic) >> 32 for x < 2^16
  hbtgpu_metric metrics[HBTGPU_MAX_METRICS];
};

__device__ __forceinline__ double add_reduce_int(const uint64_t* v, uint32_t k) {
  // std::accumulate(begin, end, 0): int accumulator (BuiltinMetrics.cpp:94)
  int32_t acc = 0;
  for (uint32_t i = 0; i < k; ++i)
    acc = static_cast<int32_t>(static_cast<uint32_t>(static_cast<uint64_t>(static_cast<int64_t>(acc)) + v[i]));
  return static_cast<double>(acc);
}

// metrics of one cell from its scaled counts `sc` (any address space) and accumulated time_running
__device__ __forceinline__ uint32_t derive_metrics(const DeriveParams& D, const uint64_t* sc, uint64_t tr,
                                                   double* out_row) {
  uint32_t vmask = 0;
  for (uint32_t m = 0; m < D.n_metrics; ++m) {
    const hbtgpu_metric& md = D.metrics[m];
    double out = 0.0;
    bool ok = false;
    if (md.kind == HBTGPU_METRIC_RATE) {
      // getRateReducer: exactly two values, denominator != 0 (BuiltinMetrics.cpp:104-116)
      if (md.n_idx == 2) {
        const uint64_t num = sc[md.idx[0]], den = sc[md.idx[1]];
        if (den != 0) {
          out = __ddiv_rn(__ull2double_rn(num), __ull2double_rn(den));
          ok = true;
        }
      }
    } else if (md.kind == HBTGPU_METRIC_ADD) {
      if (md.n_idx > 0) {
        uint64_t v[HBTGPU_MAX_METRIC_EVENTS];
        for (uint32_t k = 0; k < md.n_idx; ++k) v[k] = sc[md.idx[k]];
        out = add_reduce_int(v, md.n_idx);
        ok = true;
      }
    } else if (md.kind == HBTGPU_METRIC_PER_SECOND_MEGA) {
      // float(double(count) * 1e3 / double(time_running)) (PerfMonitor.cpp:63-66)
      const double c = __ull2double_rn(sc[md.idx[0]]);
      const float f = __double2float_rn(__ddiv_rn(__dmul_rn(c, 1e3), __ull2double_rn(tr)));
      out = static_cast<double>(f);
      ok = true;
    }
    out_row[m] = out;
    if (ok) vmask |= 1u << m;
  }
  return vmask;
}

// generic form: one thread per cell, straight from global memory (any n_events)
__global__ void __launch_bounds__(256) k_derive(const DeriveParams D) {
  const uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= D.n_cells) return;
  const uint32_t n = D.n_events;
  const uint64_t* a = D.acc + i * (n + 3);
  const uint64_t te = a[n], tr = a[n + 1];
  uint64_t* sc = D.scaled + i * n;
  for (uint32_t e = 0; e < n; ++e) sc[e] = get_count(a[e], te, tr);
  D.valid[i] = derive_metrics(D, sc, tr, D.metric + i * D.n_metrics);
}

// the same for the cells a shard-boundary row was added to after the full derive had already
// run: one block per cpu, cells[cpu] = the cell, ~0 for none, ~0 - 1 for "the whole row of this
// cpu changed" (its walk stalled at or before this shard's first SAMPLE)
__device__ __forceinline__ void derive_cell(const DeriveParams& D, uint64_t i) {
  const uint32_t n = D.n_events;
  const uint64_t* a = D.acc + i * (n + 3);
  const uint64_t te = a[n], tr = a[n + 1];
  uint64_t* sc = D.scaled + i * n;
  for (uint32_t e = 0; e < n; ++e) sc[e] = get_count(a[e], te, tr);
  D.valid[i] = derive_metrics(D, sc, tr, D.metric + i * D.n_metrics);
}
__global__ void __launch_bounds__(128) k_derive_cells(const DeriveParams D, const uint64_t* cells, uint32_t n_buckets) {
  const uint64_t i = cells[blockIdx.x];
  if (i == ~0ull) return;
  if (i == ~0ull - 1) {
    const uint64_t row = static_cast<uint64_t>(blockIdx.x) * n_buckets;
    for (uint32_t b = threadIdx.x; b < n_buckets; b += blockDim.x) derive_cell(D, row + b);
  } else if (threadIdx.x == 0) {
    derive_cell(D, i);
  }
}

// tiled form: a block stages the accumulators of blockDim.x cells with coalesced loads, each
// thread scales its cell in shared memory, the scaled counts leave coalesced again.  The host
// picks CB (a power of two <= 256) so that the tile fits ~40 KiB of shared memory.
constexpr int DERIVE_CELLS = 256;
template <uint32_t CB>  // cells per block = threads per block (compile-time: the copy loops unroll)
__global__ void __launch_bounds__(CB) k_derive_tiled(const DeriveParams D, const uint32_t* only_if = nullptr) {
  if (only_if != nullptr && *only_if == 0) return;  // (the pass after the exact path: it did not run)
  extern __shared__ __align__(16) uint64_t dsm[];  // [CB][n + 3]
  const uint32_t n = D.n_events, F = n + 3;
  // last cells first: what the decode wrote last is still in L2
  const uint64_t cell0 = static_cast<uint64_t>(gridDim.x - 1 - blockIdx.x) * CB;
  const uint32_t cells = static_cast<uint32_t>(umin64(CB, D.n_cells - cell0));
  const uint64_t* src = D.acc + cell0 * F;
  for (uint32_t i = threadIdx.x; i < cells * F; i += CB) dsm[i] = src[i];
  __syncthreads();
  if (threadIdx.x < cells) {
    uint64_t* a = dsm + threadIdx.x * F;
    const uint64_t te = a[n], tr = a[n + 1];
    for (uint32_t e = 0; e < n; ++e) a[e] = get_count(a[e], te, tr);
    const uint64_t i = cell0 + threadIdx.x;
    D.valid[i] = derive_metrics(D, a, tr, D.metric + i * D.n_metrics);
  }
  __syncthreads();
  uint64_t* dst = D.scaled + cell0 * n;
  for (uint32_t o = threadIdx.x; o < cells * n; o += CB) {
    const uint32_t c = static_cast<uint32_t>((static_cast<uint64_t>(o) * D.n_magic) >> 32);  // o / n, o < 2^16
    dst[o] = dsm[c * F + (o - c * n)];
  }
}

// ------------------------------------------------------------------------------------
// elementwise arithmetic kernels (device pointers)
__global__ void k_get_count(const uint64_t* c, const uint64_t* te, const uint64_t* tr, uint64_t* out,
                            size_t n) {
  const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < n) out[i] = get_count(c[i], te[i], tr[i]);
}
__global__ void k_rate(const uint64_t* num, const uint64_t* den, double* out, uint8_t* valid, size_t n) {
  const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const bool ok = den[i] != 0;
  valid[i] = ok;
  out[i] = ok ? __ddiv_rn(__ull2double_rn(num[i]), __ull2double_rn(den[i])) : 0.0;
}
__global__ void k_add_reduce(const uint64_t* vals, size_t k, double* out, size_t n) {
  const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int32_t acc = 0;
  for (size_t j = 0; j < k; ++j)
    acc = static_cast<int32_t>(static_cast<uint32_t>(static_cast<uint64_t>(static_cast<int64_t>(acc)) + vals[i * k + j]));
  out[i] = static_cast<double>(acc);
}
__global__ void k_tsc(const uint64_t* in, uint64_t* out, size_t n, uint32_t shift, uint32_t mult,
                      uint64_t zero, int inverse) {
  const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  out[i] = inverse ? ktime_to_tsc(in[i], shift, mult, zero) : tsc_to_ktime(in[i], shift, mult, zero);
}
__global__ void k_per_second_mega(const uint64_t* c, const uint64_t* tr, float* out, size_t n) {
  const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  out[i] = __double2float_rn(__ddiv_rn(__dmul_rn(__ull2double_rn(c[i]), 1e3), __ull2double_rn(tr[i])));
}
__global__ void k_running_ratio(const uint64_t* te, const uint64_t* tr, float* ratio, uint64_t* ri,
                                uint8_t* mux, size_t n) {
  const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  // getRunningRatio returns te/tr despite its comment (:513-522)
  ratio[i] = te[i] == 0 ? 1.0f : __fdiv_rn(__ull2float_rn(te[i]), __ull2float_rn(tr[i]));
  ri[i] = te[i] == 0 ? 100ull : (100ull * tr[i]) / te[i];
  mux[i] = te[i] != 0 && tr[i] != te[i];
}

}  // namespace hbtgpu
