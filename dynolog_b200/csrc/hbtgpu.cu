// hbtgpu.cu -- host side of libhbtgpu.so (C ABI in include/hbtgpu.h).
//
// Owns the device arena, the chunk table and the launch sequence
//   memset -> k_decode<spec> -> k_verify -> [k_resolve -> k_clear -> k_decode<exact>] -> k_derive
// where the bracketed kernels return immediately unless the speculation was disproved.
// No CPU fallback exists: every result is produced by the kernels in hbtgpu_kernels.cuh.
#include "hbtgpu_kernels.cuh"

#include <cub/device/device_radix_sort.cuh>

#include <errno.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <cmath>
#include <math.h>
#include <mutex>
#include <string>
#include <vector>

using namespace hbtgpu;

namespace {

thread_local std::string g_create_error;

struct HostImage {
  uint64_t base, len;  // arena offset, bytes
  uint32_t stream;
};

struct PeerState;                  // hbtgpu_peer.inc
void peer_release(hbtgpu_ctx* ctx);
int peer_poll_error(hbtgpu_ctx* ctx);

}  // namespace

struct hbtgpu_ctx {
  hbtgpu_config cfg;
  int device = 0;
  cudaStream_t stream = nullptr;
  cudaStream_t d2h_stream = nullptr;  // result copies that overlap the next submit
  cudaEvent_t ev_results = nullptr, ev_d2h = nullptr;
  // shard mode: the head stubs need only the images, so k_head runs beside the decode
  cudaStream_t aux_stream = nullptr;
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  bool d2h_pending = false;
  cudaEvent_t ev_run0 = nullptr, ev_run1 = nullptr;
  // one event pair per hbtgpu_run around the decode+aggregate kernel, harvested at sync
  static constexpr int kEvRing = 128;
  cudaEvent_t ev_dec0[kEvRing] = {nullptr}, ev_dec1[kEvRing] = {nullptr};
  int ev_pending = 0;
  std::string err;

  // device memory
  uint8_t* arena = nullptr;
  uint64_t arena_cap = 0, arena_used = 0;
  hbtgpu_carry *d_carry_in = nullptr, *d_carry_out = nullptr, *d_carry_init = nullptr;
  hbtgpu_head_stub* d_head = nullptr;
  uint64_t* d_counters = nullptr;
  uint64_t* d_acc = nullptr;
  uint64_t* d_acc_total = nullptr;  // hbtgpu_merge: sums over all ranks
  uint64_t* acc_cur = nullptr;      // what getters / derive read: d_acc, or d_acc_total after a merge
  uint64_t* d_scaled = nullptr;
  double* d_metric = nullptr;
  uint32_t* d_valid = nullptr;
  uint32_t* d_flags = nullptr;
  // IP|TID format: attribution tables
  uint64_t *d_tid_keys = nullptr, *d_ip_keys = nullptr;
  unsigned long long *d_tid_cnt = nullptr, *d_ip_cnt = nullptr, *d_attr_special = nullptr;
  uint32_t tid_slots = 0, ip_slots = 0;
  // HBTGPU_F_ACCUMULATE_ATTR: the totals over the runs since the last hbtgpu_attr_clear
  uint64_t *d_tid_keys_tot = nullptr, *d_ip_keys_tot = nullptr;
  unsigned long long *d_tid_cnt_tot = nullptr, *d_ip_cnt_tot = nullptr, *d_attr_special_tot = nullptr;
  // hbtgpu_attr_merge: the totals over all ranks
  uint64_t *d_tid_keys_all = nullptr, *d_ip_keys_all = nullptr;
  unsigned long long *d_tid_cnt_all = nullptr, *d_ip_cnt_all = nullptr, *d_attr_special_all = nullptr;
  // every set of tables is one allocation {tid keys, tid counts, ip keys, ip counts, special[4]}
  // (one IPC handle per set): attr_carve
  uint64_t *d_attr_block = nullptr, *d_attr_block_tot = nullptr, *d_attr_block_all = nullptr;
  bool attr_all_valid = false;  // the getters read the merged tables (until the next run)
  bool accumulate = false;
  ChunkDesc* d_chunks = nullptr;
  ChunkIO* d_io = nullptr;
  ImageDesc* d_images = nullptr;
  StreamDesc* d_streams = nullptr;
  size_t cap_chunks = 0, cap_io = 0, cap_images = 0;
  uint64_t *d_rows = nullptr, *d_rec_off = nullptr;
  uint64_t* d_row_times = nullptr;  // [row][2]: tstamp, duration again, densely
  size_t cap_row_times = 0;
  uint32_t *d_rec_type = nullptr, *d_rec_size = nullptr;
  size_t cap_rows = 0, cap_rec_off = 0, cap_rec_type = 0, cap_rec_size = 0;
  void* d_math = nullptr;
  size_t cap_math = 0;
  uint64_t* d_series = nullptr;  // hbtgpu_series_reduce results
  uint64_t* d_boundary_cell = nullptr;  // [max_cpus] cell a shard-boundary row was added to, or ~0
  HeadExtra* d_head_extra = nullptr;    // [max_cpus] shard mode: records in front of the first SAMPLE
  uint32_t* d_dead = nullptr;           // [max_cpus] shard mode: an earlier shard of this cpu stalled
  uint64_t* d_stub_block = nullptr;     // this rank's packed stubs (host getters)
  PeerState* peer = nullptr;     // peer-memory exchange (hbtgpu_peer.inc)
  bool derived = false;                 // k_derive has run over the current accumulators
  size_t cap_series = 0;

  // host bookkeeping
  std::vector<hbtgpu_carry> carry_in;
  std::vector<HostImage> images;  // submission order
  bool table_dirty = true;
  bool ran = false, emitted = false;
  std::vector<ChunkDesc> chunks;
  std::vector<ImageDesc> img_descs;
  std::vector<StreamDesc> stream_descs;
  std::vector<uint64_t> stream_row_base, stream_rec_base, stream_rows, stream_recs;
  uint64_t chunk_bytes = 0;
  uint64_t acc_words = 0, n_cells = 0;
  hbtgpu_stats stats{};
  std::vector<uint64_t> h_counters;
  std::vector<hbtgpu_carry> h_carry_out;
  bool counters_valid = false;
  uint32_t h_flags[FLAG_COUNT] = {0};
};

namespace {

int fail(hbtgpu_ctx* c, int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  if (c) c->err = buf;
  else g_create_error = buf;
  return code;
}

#define CK(call)                                                                              \
  do {                                                                                        \
    cudaError_t e_ = (call);                                                                  \
    if (e_ != cudaSuccess)                                                                    \
      return fail(ctx, -EIO, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, \
                  __LINE__);                                                                  \
  } while (0)

template <class T>
int ensure(hbtgpu_ctx* ctx, T** p, size_t* cap, size_t need) {
  if (need <= *cap && *p) return 0;
  // allocate first, then swap: a failed cudaMalloc leaves the old buffer and its capacity intact
  const size_t ncap = std::max(need, *cap * 2);
  T* fresh = nullptr;
  CK(cudaMalloc(reinterpret_cast<void**>(&fresh), ncap * sizeof(T)));
  if (*p) {
    const cudaError_t e = cudaFree(*p);
    if (e != cudaSuccess) {
      cudaFree(fresh);
      return fail(ctx, -EIO, "cudaFree failed: %s", cudaGetErrorString(e));
    }
  }
  *p = fresh;
  *cap = ncap;
  return 0;
}

constexpr uint64_t kImageAlign = 128;

// Split every image into chunks; order = stream, then submission order, then position.
// tile size of a context (see tile_bytes_for) and a chunk size that is a whole number of tiles
uint32_t cfg_tile_bytes(const hbtgpu_config& cfg) {
  return tile_bytes_for(5u + (cfg.sample_format == HBTGPU_SAMPLE_IP_TID_TIME_READ ? 2u : 0u) + cfg.n_events);
}
uint32_t ctx_tile_bytes(const hbtgpu_ctx* ctx) { return cfg_tile_bytes(ctx->cfg); }
uint64_t chunk_bytes_for(uint64_t target, uint64_t cap, uint32_t tile) {
  const uint64_t kmax = std::max<uint64_t>(1, (cap + tile / 2) / tile);
  const uint64_t k = std::min(kmax, std::max<uint64_t>(1, (target + tile - 1) / tile));
  return k * tile;
}

int build_table(hbtgpu_ctx* ctx) {
  const uint32_t ns = ctx->cfg.max_cpus;
  ctx->chunks.clear();
  ctx->img_descs.clear();
  ctx->stream_descs.assign(ns, StreamDesc{0, 0});
  uint64_t total = 0;
  for (auto& im : ctx->images) total += im.len;
  // enough chunks to fill 148 SMs several times over, but not smaller than one tile
  uint64_t target = total / (148ull * 12ull);
  const uint64_t cap = 512ull << 10;
  uint64_t cb = chunk_bytes_for(target, cap, ctx_tile_bytes(ctx));
  {
    // 148 SMs x 5 resident CTAs take 740 chunks at a time: among the chunk sizes around the
    // target pick the one whose last wave is fullest (4032 chunks were 5.45 waves)
    const uint64_t tile = ctx_tile_bytes(ctx);
    const uint64_t k0 = cb / tile;
    double best = -1.0;
    uint64_t best_k = k0;
    for (uint64_t k = std::max<uint64_t>(1, k0 - k0 / 4); k <= k0 + k0 / 4 && k * tile <= cap + tile / 2; ++k) {
      uint64_t n = 0;
      for (auto& im : ctx->images) n += (im.len + k * tile - 1) / (k * tile);
      if (n < 740) continue;
      const double waves = static_cast<double>(n) / 740.0;
      const double fill = waves / std::ceil(waves);
      if (fill > best + 1e-9) {
        best = fill;
        best_k = k;
      }
    }
    cb = best_k * tile;
  }
  ctx->chunk_bytes = cb;

  std::vector<std::vector<uint32_t>> per_stream(ns);
  for (uint32_t i = 0; i < ctx->images.size(); ++i) per_stream[ctx->images[i].stream].push_back(i);
  for (uint32_t s = 0; s < ns; ++s) {
    ctx->stream_descs[s].first_image = static_cast<uint32_t>(ctx->img_descs.size());
    ctx->stream_descs[s].n_images = static_cast<uint32_t>(per_stream[s].size());
    uint64_t spos = 0;
    const HostImage* prev = nullptr;
    const size_t first_chunk_of_stream = ctx->chunks.size();
    for (uint32_t ii : per_stream[s]) {
      const HostImage& im = ctx->images[ii];
      ImageDesc id{im.base, im.len, spos, s, static_cast<uint32_t>(ctx->chunks.size()), 0, 0};
      for (uint64_t b = 0; b < im.len; b += cb) {
        ChunkDesc cd{};
        cd.img_base = im.base;
        cd.img_len = im.len;
        cd.begin = b;
        cd.end = std::min(im.len, b + cb);
        cd.stream_pos = spos;
        cd.stream = s;
        cd.image = static_cast<uint32_t>(ctx->img_descs.size());
        cd.flags = 0;
        if (b == 0) cd.flags |= CH_FIRST_OF_IMAGE;
        if (cd.end == im.len) cd.flags |= CH_LAST_OF_IMAGE;
        if (b > 0) {
          const uint64_t hb = b > HALO_BYTES ? b - HALO_BYTES : 0;
          cd.halo_off = im.base + hb;
          cd.halo_len = static_cast<uint32_t>(b - hb);
          if (hb == 0) cd.flags |= CH_HALO_EXACT_START;
        } else if (prev) {
          uint64_t hb = prev->len > HALO_BYTES ? prev->len - HALO_BYTES : 0;
          if (hb == 0) cd.flags |= CH_HALO_EXACT_START;
          if (hb & 15ull) hb += 8;  // TMA source must be 16-byte aligned (and the halo <= HALO_BYTES)
          cd.halo_off = prev->base + hb;
          cd.halo_len = static_cast<uint32_t>(prev->len - hb);
        } else {
          cd.halo_len = 0;
        }
        ctx->chunks.push_back(cd);
        ++id.n_chunks;
      }
      ctx->img_descs.push_back(id);
      spos += im.len;
      prev = &im;
    }
    if (ctx->chunks.size() > first_chunk_of_stream) {
      ctx->chunks[first_chunk_of_stream].flags |= CH_FIRST_OF_STREAM;
      ctx->chunks.back().flags |= CH_LAST_OF_STREAM;
    }
  }
  const size_t nc = ctx->chunks.size(), ni = ctx->img_descs.size();
  if (int r = ensure(ctx, &ctx->d_chunks, &ctx->cap_chunks, std::max<size_t>(nc, 1))) return r;
  if (int r = ensure(ctx, &ctx->d_io, &ctx->cap_io, std::max<size_t>(nc, 1))) return r;
  if (int r = ensure(ctx, &ctx->d_images, &ctx->cap_images, std::max<size_t>(ni, 1))) return r;
  if (nc) CK(cudaMemcpyAsync(ctx->d_chunks, ctx->chunks.data(), nc * sizeof(ChunkDesc),
                             cudaMemcpyHostToDevice, ctx->stream));
  if (ni) CK(cudaMemcpyAsync(ctx->d_images, ctx->img_descs.data(), ni * sizeof(ImageDesc),
                             cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaMemcpyAsync(ctx->d_streams, ctx->stream_descs.data(), ns * sizeof(StreamDesc),
                     cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaMemsetAsync(ctx->d_io, 0, std::max<size_t>(nc, 1) * sizeof(ChunkIO), ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));  // host vectors may be rebuilt after this
  ctx->table_dirty = false;
  ctx->stats.n_chunks = static_cast<uint32_t>(nc);
  ctx->stats.image_bytes = total;
  return 0;
}

// IP|TID format: words of one set of attribution tables, and the tables inside such a block
size_t attr_block_words(const hbtgpu_ctx* ctx) { return 2ull * ctx->tid_slots + 2ull * ctx->ip_slots + 4; }
void attr_carve(const hbtgpu_ctx* ctx, uint64_t* block, uint64_t** tid_keys, unsigned long long** tid_cnt,
                uint64_t** ip_keys, unsigned long long** ip_cnt, unsigned long long** special) {
  *tid_keys = block;
  *tid_cnt = reinterpret_cast<unsigned long long*>(block + ctx->tid_slots);
  *ip_keys = block + 2ull * ctx->tid_slots;
  *ip_cnt = reinterpret_cast<unsigned long long*>(block + 2ull * ctx->tid_slots + ctx->ip_slots);
  *special = reinterpret_cast<unsigned long long*>(block + 2ull * ctx->tid_slots + 2ull * ctx->ip_slots);
}

Params make_params(hbtgpu_ctx* ctx) {
  Params P{};
  P.arena = ctx->arena;
  P.chunks = ctx->d_chunks;
  P.io = ctx->d_io;
  P.images = ctx->d_images;
  P.streams = ctx->d_streams;
  P.carry_in = ctx->d_carry_in;
  P.carry_out = ctx->d_carry_out;
  P.head = ctx->d_head;
  P.head_extra = ctx->d_head_extra;
  P.counters = ctx->d_counters;
  P.acc = ctx->d_acc;
  P.flags = ctx->d_flags;
  P.rows = ctx->d_rows;
  P.row_times = ctx->d_row_times;
  P.rec_off = ctx->d_rec_off;
  P.rec_type = ctx->d_rec_type;
  P.rec_size = ctx->d_rec_size;
  P.t0 = ctx->cfg.t0;
  P.bucket_ns = ctx->cfg.bucket_ns;
  P.tsc_mask = ctx->cfg.tsc_mask;
  P.time_zero = ctx->cfg.time_zero;
  P.n_events = ctx->cfg.n_events;
  P.n_buckets = ctx->cfg.n_buckets;
  P.n_chunks = static_cast<uint32_t>(ctx->chunks.size());
  P.n_streams = ctx->cfg.max_cpus;
  P.time_shift = ctx->cfg.time_shift;
  P.time_mult = ctx->cfg.time_mult;
  P.force_general = (ctx->cfg.flags & HBTGPU_F_FORCE_GENERAL) ? 1 : 0;
  P.external_carry = (ctx->cfg.flags & HBTGPU_F_EXTERNAL_CARRY) ? 1 : 0;
  {
    const unsigned __int128 range = static_cast<unsigned __int128>(ctx->cfg.n_buckets) * ctx->cfg.bucket_ns;
    P.range_ns = range > ~0ull ? ~0ull : static_cast<uint64_t>(range);
    P.bucket_magic = static_cast<uint64_t>((static_cast<unsigned __int128>(1) << 64) / ctx->cfg.bucket_ns);
    const uint64_t b = ctx->cfg.bucket_ns;
    P.bucket_shift = (b & (b - 1)) == 0 ? static_cast<uint32_t>(__builtin_ctzll(b)) : 0xFFFFFFFFu;
    P.bucket_f64 = (b < (1ull << 32) && P.range_ns < (1ull << 53)) ? 1u : 0u;
    P.bucket_inv = 1.0 / static_cast<double>(b);
    P.pre = ctx->cfg.sample_format == HBTGPU_SAMPLE_IP_TID_TIME_READ ? 2 : 0;
    P.fix = 40 + 8 * P.pre;
    P.tile_bytes = ctx_tile_bytes(ctx);
    P.tid_keys = ctx->d_tid_keys;
    P.tid_cnt = ctx->d_tid_cnt;
    P.ip_keys = ctx->d_ip_keys;
    P.ip_cnt = ctx->d_ip_cnt;
    P.attr_special = ctx->d_attr_special;
    P.tid_mask = ctx->tid_slots ? ctx->tid_slots - 1 : 0;
    P.ip_mask = ctx->ip_slots ? ctx->ip_slots - 1 : 0;
    const uint64_t S = 5ull + P.pre + ctx->cfg.n_events;
    P.s_magic = static_cast<uint32_t>(((1ull << 32) + S - 1) / S);
  }
  return P;
}

// the attribution tables the getters read: all ranks' (after hbtgpu_attr_merge), this context's
// totals over its runs (HBTGPU_F_ACCUMULATE_ATTR), or the last run's
struct AttrView {
  const uint64_t *tid_keys, *ip_keys;
  const unsigned long long *tid_cnt, *ip_cnt, *special;
};
AttrView attr_view(const hbtgpu_ctx* ctx) {
  if (ctx->attr_all_valid)
    return {ctx->d_tid_keys_all, ctx->d_ip_keys_all, ctx->d_tid_cnt_all, ctx->d_ip_cnt_all, ctx->d_attr_special_all};
  if (ctx->accumulate)
    return {ctx->d_tid_keys_tot, ctx->d_ip_keys_tot, ctx->d_tid_cnt_tot, ctx->d_ip_cnt_tot, ctx->d_attr_special_tot};
  return {ctx->d_tid_keys, ctx->d_ip_keys, ctx->d_tid_cnt, ctx->d_ip_cnt, ctx->d_attr_special};
}

DeriveParams derive_params(hbtgpu_ctx* ctx) {
  DeriveParams D{};
  D.acc = ctx->acc_cur;
  D.scaled = ctx->d_scaled;
  D.metric = ctx->d_metric;
  D.valid = ctx->d_valid;
  D.n_cells = ctx->n_cells;
  D.n_events = ctx->cfg.n_events;
  D.n_metrics = ctx->cfg.n_metrics;
  memcpy(D.metrics, ctx->cfg.metrics, sizeof(D.metrics));
  D.n_magic = ((1ull << 32) + D.n_events - 1) / D.n_events;
  return D;
}

// st: the stream (default: the context's); only_if: launch a pass that returns at once unless
// *only_if != 0 (the exact path has replaced the sums)
int launch_derive(hbtgpu_ctx* ctx, cudaStream_t st = nullptr, const uint32_t* only_if = nullptr) {
  if (!st) st = ctx->stream;
  if (ctx->d2h_pending && st == ctx->stream) {  // do not overwrite tensors that are still being copied out
    CK(cudaStreamWaitEvent(ctx->stream, ctx->ev_d2h, 0));
    ctx->d2h_pending = false;
  }
  const DeriveParams D = derive_params(ctx);
  // cells per block: the largest power of two <= 256 whose tile fits 40 KiB of shared memory
  const uint32_t F = D.n_events + 3;
  uint32_t cb = DERIVE_CELLS;
  while (cb > 32 && static_cast<size_t>(cb) * F * 8 > (40u << 10)) cb >>= 1;
  const uint64_t blocks = (ctx->n_cells + cb - 1) / cb;
  const size_t smem = static_cast<size_t>(cb) * F * 8;
  const unsigned grid = static_cast<unsigned>(blocks);
  if (cb == 256) k_derive_tiled<256><<<grid, 256, smem, st>>>(D, only_if);
  else if (cb == 128) k_derive_tiled<128><<<grid, 128, smem, st>>>(D, only_if);
  else if (cb == 64) k_derive_tiled<64><<<grid, 64, smem, st>>>(D, only_if);
  else k_derive_tiled<32><<<grid, 32, smem, st>>>(D, only_if);
  CK(cudaGetLastError());
  ctx->derived = true;
  return 0;
}

// After shard-boundary rows were added: a full derive, or -- if the caller already derived
// after hbtgpu_run (to overlap it with the stub exchange) -- only the cells those rows touched.
int derive_after_boundaries(hbtgpu_ctx* ctx) {
  if (!ctx->derived) return launch_derive(ctx);
  const uint32_t ns = ctx->cfg.max_cpus;
  k_derive_cells<<<ns, 128, 0, ctx->stream>>>(derive_params(ctx), ctx->d_boundary_cell, ctx->cfg.n_buckets);
  CK(cudaGetLastError());
  return 0;
}

int fetch_counters(hbtgpu_ctx* ctx) {
  if (ctx->counters_valid) return 0;
  const size_t n = static_cast<size_t>(ctx->cfg.max_cpus) * C_SLOTS;
  ctx->h_counters.resize(n);
  CK(cudaMemcpyAsync(ctx->h_counters.data(), ctx->d_counters, n * 8, cudaMemcpyDeviceToHost,
                     ctx->stream));
  CK(cudaMemcpyAsync(ctx->h_flags, ctx->d_flags, sizeof(ctx->h_flags), cudaMemcpyDeviceToHost,
                     ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  ctx->counters_valid = true;
  ctx->stats.exact_runs = ctx->h_flags[FLAG_NEED_EXACT] ? 1 : 0;
  ctx->stats.exact_reason = ctx->h_flags[FLAG_WHY];
  ctx->stats.exact_chunk = ctx->h_flags[FLAG_WHERE];
  // getLastTimeStamp() is the carried state: count_.tstamp or kInvalidTimeStamp
  ctx->h_carry_out.resize(ctx->cfg.max_cpus);
  CK(cudaMemcpyAsync(ctx->h_carry_out.data(), ctx->d_carry_out,
                     ctx->cfg.max_cpus * sizeof(hbtgpu_carry), cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  if (ctx->h_flags[FLAG_INTERNAL])
    return fail(ctx, -EIO, "internal inconsistency: exact pass met an unresolved record (code %u)",
                ctx->h_flags[FLAG_INTERNAL]);
  return 0;
}

// Second pass that writes rows and the record table, using the (verified) chunk entries.
int ensure_emitted(hbtgpu_ctx* ctx) {
  if (!ctx->ran) return fail(ctx, -EINVAL, "hbtgpu_run has not been called");
  if (ctx->emitted) return 0;
  if (int r = fetch_counters(ctx)) return r;
  const size_t nc = ctx->chunks.size();
  const uint32_t ns = ctx->cfg.max_cpus;
  ctx->stream_row_base.assign(ns, 0);
  ctx->stream_rec_base.assign(ns, 0);
  ctx->stream_rows.assign(ns, 0);
  ctx->stream_recs.assign(ns, 0);
  uint64_t rows = 0, recs = 0;
  if (nc) {
    std::vector<ChunkIO> io(nc);
    CK(cudaMemcpyAsync(io.data(), ctx->d_io, nc * sizeof(ChunkIO), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    uint32_t cur = 0xFFFFFFFFu;
    for (size_t c = 0; c < nc; ++c) {
      const uint32_t s = ctx->chunks[c].stream;
      if (s != cur) {
        cur = s;
        ctx->stream_row_base[s] = rows;
        ctx->stream_rec_base[s] = recs;
      }
      io[c].row_base = rows;
      io[c].rec_base = recs;
      rows += io[c].n_rows;
      recs += io[c].n_records;
      ctx->stream_rows[s] += io[c].n_rows;
      ctx->stream_recs[s] += io[c].n_records;
    }
    CK(cudaMemcpyAsync(ctx->d_io, io.data(), nc * sizeof(ChunkIO), cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
  }
  const uint32_t n = ctx->cfg.n_events;
  if (int r = ensure(ctx, &ctx->d_rows, &ctx->cap_rows, std::max<uint64_t>(rows, 1) * (n + 2))) return r;
  if (int r = ensure(ctx, &ctx->d_row_times, &ctx->cap_row_times, std::max<uint64_t>(rows, 1) * 2)) return r;
  {
    if (int r = ensure(ctx, &ctx->d_rec_off, &ctx->cap_rec_off, std::max<uint64_t>(recs, 1))) return r;
    if (int r = ensure(ctx, &ctx->d_rec_type, &ctx->cap_rec_type, std::max<uint64_t>(recs, 1))) return r;
    if (int r = ensure(ctx, &ctx->d_rec_size, &ctx->cap_rec_size, std::max<uint64_t>(recs, 1))) return r;
  }
  if (nc) {
    Params P = make_params(ctx);
    CK(cudaEventRecord(ctx->ev_run0, ctx->stream));
    if (P.pre) k_decode<false, MODE_EMIT, 2><<<static_cast<unsigned>(nc), NT, decode_smem_bytes<MODE_EMIT, 2>(P.tile_bytes), ctx->stream>>>(P);
    else k_decode<false, MODE_EMIT, 0><<<static_cast<unsigned>(nc), NT, decode_smem_bytes<MODE_EMIT, 0>(P.tile_bytes), ctx->stream>>>(P);
    CK(cudaGetLastError());
    CK(cudaEventRecord(ctx->ev_run1, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaEventElapsedTime(&ctx->stats.last_emit_ms, ctx->ev_run0, ctx->ev_run1));
  }
  ctx->emitted = true;
  return 0;
}

template <class K>
int set_smem(hbtgpu_ctx* ctx, K kernel, size_t bytes) {
  CK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(bytes)));
  return 0;
}

}  // namespace

namespace {
struct MathBuf {
  hbtgpu_ctx* ctx;
  uint8_t* base;
  size_t off = 0;
  template <class T>
  T* take(size_t n) {
    T* p = reinterpret_cast<T*>(base + off);
    off += (n * sizeof(T) + 255) / 256 * 256;
    return p;
  }
};
int math_reserve(hbtgpu_ctx* ctx, size_t bytes, MathBuf* mb) {
  CK(cudaSetDevice(ctx->device));
  uint8_t* p = static_cast<uint8_t*>(ctx->d_math);
  if (int r = ensure(ctx, &p, &ctx->cap_math, bytes)) {
    ctx->d_math = p;
    return r;
  }
  ctx->d_math = p;
  mb->ctx = ctx;
  mb->base = p;
  return 0;
}
inline size_t pad256(size_t b) { return (b + 255) / 256 * 256; }
inline unsigned nblk(size_t n) { return static_cast<unsigned>((n + 255) / 256); }
}  // namespace

extern "C" {

const char* hbtgpu_version(void) { return "hbtgpu 0.1 (sm_100a)"; }

const char* hbtgpu_last_error(const hbtgpu_ctx* ctx) {
  return ctx ? ctx->err.c_str() : g_create_error.c_str();
}

int hbtgpu_create(const hbtgpu_config* cfg, hbtgpu_ctx** out) {
  hbtgpu_ctx* ctx = nullptr;  // for CK/fail before the context exists
  if (!cfg || !out) return fail(ctx, -EINVAL, "null argument");
  *out = nullptr;
  // HBT_ARG_CHECK-style argument validation (hbt/src/common/Defs.h:275-293 throws; the ABI returns)
  if (cfg->n_events == 0 || cfg->n_events > HBTGPU_MAX_EVENTS)
    return fail(ctx, -EINVAL, "n_events must be in [1,%d]", HBTGPU_MAX_EVENTS);
  if (cfg->max_cpus == 0 || cfg->max_cpus > 4096) return fail(ctx, -EINVAL, "max_cpus out of range");
  if (cfg->n_buckets == 0 || cfg->n_buckets >= 0xFFFFFFFEu) return fail(ctx, -EINVAL, "n_buckets out of range");
  if (cfg->bucket_ns == 0) return fail(ctx, -EINVAL, "bucket_ns must be >= 1");
  if (cfg->n_metrics > HBTGPU_MAX_METRICS) return fail(ctx, -EINVAL, "too many metrics");
  if (cfg->time_shift > 63) return fail(ctx, -EINVAL, "time_shift must be < 64");
  if (cfg->sample_format > HBTGPU_SAMPLE_IP_TID_TIME_READ) return fail(ctx, -EINVAL, "unknown sample_format");
  if (cfg->tsc_mask != 0 && cfg->time_mult == 0) return fail(ctx, -EINVAL, "time_mult must be != 0");
  for (uint32_t m = 0; m < cfg->n_metrics; ++m) {
    const hbtgpu_metric& md = cfg->metrics[m];
    if (md.n_idx > HBTGPU_MAX_METRIC_EVENTS) return fail(ctx, -EINVAL, "metric %u: too many events", m);
    if (md.kind == HBTGPU_METRIC_PER_SECOND_MEGA && md.n_idx != 1)
      return fail(ctx, -EINVAL, "metric %u: PER_SECOND_MEGA takes one event", m);
    for (uint32_t k = 0; k < md.n_idx; ++k)
      if (md.idx[k] >= cfg->n_events) return fail(ctx, -EINVAL, "metric %u: event index out of range", m);
  }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
    return fail(ctx, -ENODEV, "no CUDA device: libhbtgpu has no CPU fallback");
  if (cfg->device < 0 || cfg->device >= ndev) return fail(ctx, -ENODEV, "device %d not present", cfg->device);

  ctx = new hbtgpu_ctx();
  ctx->cfg = *cfg;
  ctx->device = cfg->device;
  auto bail = [&](int code) {
    g_create_error = ctx->err;
    hbtgpu_destroy(ctx);
    return code;
  };
#define CKC(call)                                                                         \
  do {                                                                                    \
    cudaError_t e_ = (call);                                                              \
    if (e_ != cudaSuccess) {                                                              \
      fail(ctx, -EIO, "%s failed: %s", #call, cudaGetErrorString(e_));                    \
      return bail(-EIO);                                                                  \
    }                                                                                     \
  } while (0)
  CKC(cudaSetDevice(ctx->device));
  CKC(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
  CKC(cudaStreamCreateWithFlags(&ctx->d2h_stream, cudaStreamNonBlocking));
  CKC(cudaEventCreateWithFlags(&ctx->ev_results, cudaEventDisableTiming));
  CKC(cudaEventCreateWithFlags(&ctx->ev_d2h, cudaEventDisableTiming));
  CKC(cudaStreamCreateWithFlags(&ctx->aux_stream, cudaStreamNonBlocking));
  CKC(cudaEventCreateWithFlags(&ctx->ev_fork, cudaEventDisableTiming));
  CKC(cudaEventCreateWithFlags(&ctx->ev_join, cudaEventDisableTiming));
  CKC(cudaEventCreate(&ctx->ev_run0));
  CKC(cudaEventCreate(&ctx->ev_run1));
  for (int i = 0; i < hbtgpu_ctx::kEvRing; ++i) {
    CKC(cudaEventCreate(&ctx->ev_dec0[i]));
    CKC(cudaEventCreate(&ctx->ev_dec1[i]));
  }
  const uint32_t ns = cfg->max_cpus, n = cfg->n_events;
  ctx->n_cells = static_cast<uint64_t>(ns) * cfg->n_buckets;
  ctx->acc_words = ctx->n_cells * (n + 3);
  ctx->arena_cap = (cfg->arena_bytes + kImageAlign - 1) / kImageAlign * kImageAlign;
  // slack: image padding plus the staged overlap / 16-byte rounding never leaves the allocation
  CKC(cudaMalloc(reinterpret_cast<void**>(&ctx->arena), ctx->arena_cap + kImageAlign * ns + TILE_MAX_BYTES + OV_BYTES));
  CKC(cudaMemsetAsync(ctx->arena, 0, ctx->arena_cap + kImageAlign * ns + TILE_MAX_BYTES + OV_BYTES, ctx->stream));
  CKC(cudaMalloc(reinterpret_cast<void**>(&ctx->d_carry_in), ns * sizeof(hbtgpu_carry)));
  CKC(cudaMalloc(reinterpret_cast<void**>(&ctx->d_carry_out), ns * sizeof(hbtgpu_carry)));
  CKC(cudaMalloc(reinterpret_cast<void**>(&ctx->d_carry_init), ns * sizeof(hbtgpu_carry)));
  CKC(cudaMemsetAsync(ctx->d_carry_init, 0, ns * sizeof(hbtgpu_carry), ctx->stream));
  CKC(cudaMalloc(reinterpret_cast<void**>(&ctx->d_head), ns * sizeof(hbtgpu_head_stub)));
  CKC(cudaMalloc(reinterpret_cast<void**>(&ctx->d_counters), ns * C_SLOTS * 8));
  CKC(cudaMalloc(reinterpret_cast<void**>(&ctx->d_acc), ctx->acc_words * 8));
  ctx->acc_cur = ctx->d_acc;
  CKC(cudaMalloc(reinterpret_cast<void**>(&ctx->d_scaled), ctx->n_cells * n * 8));
  CKC(cudaMalloc(reinterpret_cast<void**>(&ctx->d_metric),
                 std::max<uint64_t>(ctx->n_cells * cfg->n_metrics, 1) * 8));
  CKC(cudaMalloc(reinterpret_cast<void**>(&ctx->d_valid), ctx->n_cells * 4));
  CKC(cudaMalloc(reinterpret_cast<void**>(&ctx->d_flags), FLAG_COUNT * 4));
  CKC(cudaMalloc(reinterpret_cast<void**>(&ctx->d_streams), ns * sizeof(StreamDesc)));
  CKC(cudaMalloc(reinterpret_cast<void**>(&ctx->d_boundary_cell), ns * 8ull));
  CKC(cudaMalloc(reinterpret_cast<void**>(&ctx->d_head_extra), ns * sizeof(HeadExtra)));
  CKC(cudaMalloc(reinterpret_cast<void**>(&ctx->d_dead), ns * 4ull));
  CKC(cudaMemsetAsync(ctx->d_head_extra, 0, ns * sizeof(HeadExtra), ctx->stream));
  CKC(cudaMemsetAsync(ctx->d_dead, 0, ns * 4ull, ctx->stream));
  if (cfg->sample_format == HBTGPU_SAMPLE_IP_TID_TIME_READ) {
    const uint32_t tl = cfg->tid_table_log2 ? cfg->tid_table_log2 : 16, il = cfg->ip_table_log2 ? cfg->ip_table_log2 : 22;
    if (tl < 4 || tl > 28 || il < 4 || il > 28) {
      fail(ctx, -EINVAL, "table_log2 out of range [4,28]");
      return bail(-EINVAL);
    }
    ctx->tid_slots = 1u << tl;
    ctx->ip_slots = 1u << il;
    CKC(cudaMalloc(reinterpret_cast<void**>(&ctx->d_attr_block), attr_block_words(ctx) * 8));
    attr_carve(ctx, ctx->d_attr_block, &ctx->d_tid_keys, &ctx->d_tid_cnt, &ctx->d_ip_keys, &ctx->d_ip_cnt, &ctx->d_attr_special);
    if (cfg->flags & HBTGPU_F_ACCUMULATE_ATTR) {
      ctx->accumulate = true;
      CKC(cudaMalloc(reinterpret_cast<void**>(&ctx->d_attr_block_tot), attr_block_words(ctx) * 8));
      attr_carve(ctx, ctx->d_attr_block_tot, &ctx->d_tid_keys_tot, &ctx->d_tid_cnt_tot, &ctx->d_ip_keys_tot,
                 &ctx->d_ip_cnt_tot, &ctx->d_attr_special_tot);
      CKC(cudaMemsetAsync(ctx->d_tid_keys_tot, 0xFF, ctx->tid_slots * 8ull, ctx->stream));
      CKC(cudaMemsetAsync(ctx->d_tid_cnt_tot, 0, ctx->tid_slots * 8ull, ctx->stream));
      CKC(cudaMemsetAsync(ctx->d_ip_keys_tot, 0xFF, ctx->ip_slots * 8ull, ctx->stream));
      CKC(cudaMemsetAsync(ctx->d_ip_cnt_tot, 0, ctx->ip_slots * 8ull, ctx->stream));
      CKC(cudaMemsetAsync(ctx->d_attr_special_tot, 0, 4 * 8, ctx->stream));
    }
  }
  CKC(cudaMemsetAsync(ctx->d_carry_in, 0, ns * sizeof(hbtgpu_carry), ctx->stream));
  CKC(cudaMemsetAsync(ctx->d_carry_out, 0, ns * sizeof(hbtgpu_carry), ctx->stream));
  CKC(cudaMemsetAsync(ctx->d_head, 0, ns * sizeof(hbtgpu_head_stub), ctx->stream));
  CKC(cudaMemsetAsync(ctx->d_counters, 0, ns * C_SLOTS * 8, ctx->stream));
  CKC(cudaMemsetAsync(ctx->d_acc, 0, ctx->acc_words * 8, ctx->stream));
  CKC(cudaMemsetAsync(ctx->d_flags, 0, FLAG_COUNT * 4, ctx->stream));
  CKC(cudaStreamSynchronize(ctx->stream));
  ctx->carry_in.assign(ns, hbtgpu_carry{});
  if (set_smem(ctx, k_decode<true, MODE_AGG, 0>, decode_smem_bytes<MODE_AGG, 0>(TILE_MAX_BYTES)) ||
      set_smem(ctx, k_decode<true, MODE_AGG, 0, 8>, decode_smem_bytes<MODE_AGG, 0>(TILE_MAX_BYTES)) ||
      set_smem(ctx, k_decode<true, MODE_AGG, 0, 4>, decode_smem_bytes<MODE_AGG, 0>(TILE_MAX_BYTES)) ||
      set_smem(ctx, k_decode<false, MODE_AGG, 0>, decode_smem_bytes<MODE_AGG, 0>(TILE_MAX_BYTES)) ||
      set_smem(ctx, k_decode<false, MODE_EMIT, 0>, decode_smem_bytes<MODE_EMIT, 0>(TILE_MAX_BYTES)) ||
      set_smem(ctx, k_decode<true, MODE_AGG, 2>, decode_smem_bytes<MODE_AGG, 2>(TILE_MAX_BYTES)) ||
      set_smem(ctx, k_decode<true, MODE_AGG, 2, 4>, decode_smem_bytes<MODE_AGG, 2>(TILE_MAX_BYTES)) ||
      set_smem(ctx, k_decode<false, MODE_AGG, 2>, decode_smem_bytes<MODE_AGG, 2>(TILE_MAX_BYTES)) ||
      set_smem(ctx, k_decode<false, MODE_EMIT, 2>, decode_smem_bytes<MODE_EMIT, 2>(TILE_MAX_BYTES)))
    return bail(-EIO);
#undef CKC
  *out = ctx;
  return 0;
}

void hbtgpu_destroy(hbtgpu_ctx* ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  if (ctx->stream) cudaStreamSynchronize(ctx->stream);
  peer_release(ctx);
  void* ptrs[] = {ctx->arena, ctx->d_carry_in, ctx->d_carry_out, ctx->d_carry_init, ctx->d_head, ctx->d_counters,
                  ctx->d_acc, ctx->d_scaled, ctx->d_metric, ctx->d_valid, ctx->d_flags,
                  ctx->d_chunks, ctx->d_io, ctx->d_images, ctx->d_streams, ctx->d_rows, ctx->d_row_times,
                  ctx->d_rec_off, ctx->d_rec_type, ctx->d_rec_size, ctx->d_math, ctx->d_attr_block,
                  ctx->d_attr_block_tot, ctx->d_attr_block_all, ctx->d_series, ctx->d_boundary_cell,
                  ctx->d_head_extra, ctx->d_dead, ctx->d_stub_block, ctx->d_acc_total};
  for (void* p : ptrs)
    if (p) cudaFree(p);
  for (cudaEvent_t e : {ctx->ev_run0, ctx->ev_run1})
    if (e) cudaEventDestroy(e);
  for (int i = 0; i < hbtgpu_ctx::kEvRing; ++i) {
    if (ctx->ev_dec0[i]) cudaEventDestroy(ctx->ev_dec0[i]);
    if (ctx->ev_dec1[i]) cudaEventDestroy(ctx->ev_dec1[i]);
  }
  if (ctx->d2h_stream) {
    cudaStreamSynchronize(ctx->d2h_stream);
    cudaStreamDestroy(ctx->d2h_stream);
  }
  if (ctx->aux_stream) {
    cudaStreamSynchronize(ctx->aux_stream);
    cudaStreamDestroy(ctx->aux_stream);
  }
  for (cudaEvent_t e : {ctx->ev_results, ctx->ev_d2h, ctx->ev_fork, ctx->ev_join})
    if (e) cudaEventDestroy(e);
  if (ctx->stream) cudaStreamDestroy(ctx->stream);
  delete ctx;
}

int hbtgpu_set_carry(hbtgpu_ctx* ctx, uint32_t cpu, const hbtgpu_carry* carry) {
  if (!ctx || !carry) return -EINVAL;
  if (cpu >= ctx->cfg.max_cpus) return fail(ctx, -EINVAL, "cpu %u out of range", cpu);
  CK(cudaSetDevice(ctx->device));
  ctx->carry_in[cpu] = *carry;
  CK(cudaMemcpyAsync(&ctx->d_carry_in[cpu], &ctx->carry_in[cpu], sizeof(hbtgpu_carry),
                     cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaMemcpyAsync(&ctx->d_carry_init[cpu], &ctx->carry_in[cpu], sizeof(hbtgpu_carry),
                     cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  return 0;
}

static int reserve_image(hbtgpu_ctx* ctx, uint32_t cpu, size_t n, uint64_t* base) {
  if (cpu >= ctx->cfg.max_cpus) return fail(ctx, -EINVAL, "cpu %u out of range", cpu);
  if (n % 8) return fail(ctx, -EINVAL, "image size %zu is not a multiple of 8 (records are u64-aligned)", n);
  const uint64_t b = (ctx->arena_used + kImageAlign - 1) / kImageAlign * kImageAlign;
  if (b + n > ctx->arena_cap + kImageAlign * ctx->cfg.max_cpus)
    return fail(ctx, -ENOSPC, "arena full: %llu + %zu > %llu bytes",
                (unsigned long long)b, n, (unsigned long long)ctx->arena_cap);
  *base = b;
  return 0;
}

int hbtgpu_submit_image(hbtgpu_ctx* ctx, uint32_t cpu, const void* first, size_t n1,
                        const void* second, size_t n2) {
  if (!ctx) return -EINVAL;
  if ((n1 && !first) || (n2 && !second)) return fail(ctx, -EINVAL, "null slice");
  const size_t n = n1 + n2;
  if (n == 0) return 0;
  CK(cudaSetDevice(ctx->device));
  uint64_t base;
  if (int r = reserve_image(ctx, cpu, n, &base)) return r;
  if (n1) CK(cudaMemcpyAsync(ctx->arena + base, first, n1, cudaMemcpyHostToDevice, ctx->stream));
  if (n2) CK(cudaMemcpyAsync(ctx->arena + base + n1, second, n2, cudaMemcpyHostToDevice, ctx->stream));
  // "copied before the call returns" holds for pageable memory (staged by the driver); for
  // pinned memory the copy is asynchronous, so wait: the ring may be overwritten right after.
  cudaPointerAttributes at{};
  bool pinned = cudaPointerGetAttributes(&at, first ? first : second) == cudaSuccess &&
                at.type == cudaMemoryTypeHost;
  (void)cudaGetLastError();
  if (pinned) CK(cudaStreamSynchronize(ctx->stream));
  ctx->images.push_back(HostImage{base, n, cpu});
  ctx->arena_used = base + n;
  ctx->table_dirty = true;
  return 0;
}

int hbtgpu_submit_images(hbtgpu_ctx* ctx, size_t n, const uint32_t* cpus, const void* const* images,
                         const size_t* sizes) {
  if (!ctx) return -EINVAL;
  if (n && (!cpus || !images || !sizes)) return fail(ctx, -EINVAL, "null argument");
  CK(cudaSetDevice(ctx->device));
  // validate everything first so that a failure leaves the context unchanged
  uint64_t used = ctx->arena_used;
  for (size_t i = 0; i < n; ++i) {
    if (cpus[i] >= ctx->cfg.max_cpus) return fail(ctx, -EINVAL, "cpu %u out of range", cpus[i]);
    if (sizes[i] % 8) return fail(ctx, -EINVAL, "image size %zu is not a multiple of 8", sizes[i]);
    if (sizes[i] && !images[i]) return fail(ctx, -EINVAL, "null image");
    if (!sizes[i]) continue;
    const uint64_t b = (used + kImageAlign - 1) / kImageAlign * kImageAlign;
    if (b + sizes[i] > ctx->arena_cap + kImageAlign * ctx->cfg.max_cpus)
      return fail(ctx, -ENOSPC, "arena full: %llu + %zu > %llu bytes", (unsigned long long)b, sizes[i],
                  (unsigned long long)ctx->arena_cap);
    used = b + sizes[i];
  }
  for (size_t i = 0; i < n; ++i) {
    if (!sizes[i]) continue;
    uint64_t base;
    if (int r = reserve_image(ctx, cpus[i], sizes[i], &base)) return r;
    CK(cudaMemcpyAsync(ctx->arena + base, images[i], sizes[i], cudaMemcpyHostToDevice, ctx->stream));
    ctx->images.push_back(HostImage{base, sizes[i], cpus[i]});
    ctx->arena_used = base + sizes[i];
  }
  ctx->table_dirty = true;
  CK(cudaStreamSynchronize(ctx->stream));  // the caller may overwrite the rings after this
  return 0;
}

int hbtgpu_submit_image_device(hbtgpu_ctx* ctx, uint32_t cpu, const void* device_ptr, size_t n) {
  if (!ctx) return -EINVAL;
  if (n == 0) return 0;
  if (!device_ptr) return fail(ctx, -EINVAL, "null image");
  CK(cudaSetDevice(ctx->device));
  uint64_t base;
  if (int r = reserve_image(ctx, cpu, n, &base)) return r;
  CK(cudaMemcpyAsync(ctx->arena + base, device_ptr, n, cudaMemcpyDeviceToDevice, ctx->stream));
  ctx->images.push_back(HostImage{base, n, cpu});
  ctx->arena_used = base + n;
  ctx->table_dirty = true;
  return 0;
}

int hbtgpu_reset(hbtgpu_ctx* ctx, int keep_carry) {
  if (!ctx) return -EINVAL;
  CK(cudaSetDevice(ctx->device));
  const uint32_t ns = ctx->cfg.max_cpus;
  if (keep_carry && ctx->ran) {
    CK(cudaMemcpyAsync(ctx->carry_in.data(), ctx->d_carry_out, ns * sizeof(hbtgpu_carry),
                       cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaMemcpyAsync(ctx->d_carry_in, ctx->carry_in.data(), ns * sizeof(hbtgpu_carry),
                       cudaMemcpyHostToDevice, ctx->stream));
  }
  ctx->images.clear();
  ctx->arena_used = 0;
  ctx->table_dirty = true;
  ctx->ran = false;
  ctx->emitted = false;
  ctx->counters_valid = false;
  // result tensors are cleared by the next hbtgpu_run; nothing here waits for result copies
  // that are still in flight, so the next submit can overlap them
  return 0;
}

int hbtgpu_run(hbtgpu_ctx* ctx) {
  if (!ctx) return -EINVAL;
  CK(cudaSetDevice(ctx->device));
  if (ctx->table_dirty)
    if (int r = build_table(ctx)) return r;
  const uint32_t ns = ctx->cfg.max_cpus;
  const size_t nc = ctx->chunks.size();
  const bool force_exact = (ctx->cfg.flags & HBTGPU_F_FORCE_EXACT) != 0;
  uint32_t launches = 0;
  bool derive_forked = false;
  Params P = make_params(ctx);

  if (ctx->d2h_pending) CK(cudaStreamWaitEvent(ctx->stream, ctx->ev_d2h, 0));  // acc may still be copied out
  CK(cudaEventRecord(ctx->ev_run0, ctx->stream));
  if (ctx->acc_words & 1) CK(cudaMemsetAsync(ctx->d_acc + ctx->acc_words - 1, 0, 8, ctx->stream));
  k_zero_desc<<<148 * 8, 256, 0, ctx->stream>>>(reinterpret_cast<ulonglong2*>(ctx->d_acc), ctx->acc_words / 2);
  const bool external = (ctx->cfg.flags & HBTGPU_F_EXTERNAL_CARRY) != 0;
  if (external) {
    // what lies in front of every cpu's first SAMPLE (the head stubs): read from the images alone,
    // so it runs on a second stream beside the decode and is joined at the end of the run
    CK(cudaEventRecord(ctx->ev_fork, ctx->stream));
    CK(cudaStreamWaitEvent(ctx->aux_stream, ctx->ev_fork, 0));
    k_head<<<(ns + 127) / 128, 128, 0, ctx->aux_stream>>>(P);
    CK(cudaGetLastError());
    CK(cudaEventRecord(ctx->ev_join, ctx->aux_stream));
    ++launches;
  }
  // counters, flags, carry-out of streams without images; shard mode: every stream starts from
  // an invalid state (the boundary row is added later)
  {
    const uint32_t words = ns * static_cast<uint32_t>(sizeof(hbtgpu_carry) / 8);
    k_prologue<<<(words + 255) / 256, 256, 0, ctx->stream>>>(P, ctx->d_carry_in, force_exact ? 1u : 0u);
    ++launches;
  }
  if (ctx->tid_slots) {  // attribution tables start empty
    CK(cudaMemsetAsync(ctx->d_tid_keys, 0xFF, ctx->tid_slots * 8ull, ctx->stream));
    CK(cudaMemsetAsync(ctx->d_tid_cnt, 0, ctx->tid_slots * 8ull, ctx->stream));
    CK(cudaMemsetAsync(ctx->d_ip_keys, 0xFF, ctx->ip_slots * 8ull, ctx->stream));
    CK(cudaMemsetAsync(ctx->d_ip_cnt, 0, ctx->ip_slots * 8ull, ctx->stream));
    CK(cudaMemsetAsync(ctx->d_attr_special, 0, 4 * 8, ctx->stream));
  }
  ctx->derived = false;
  ctx->attr_all_valid = false;
  ctx->acc_cur = ctx->d_acc;
  const bool timed = ctx->ev_pending < hbtgpu_ctx::kEvRing;
  if (nc) {
    if (timed) CK(cudaEventRecord(ctx->ev_dec0[ctx->ev_pending], ctx->stream));
    if (!force_exact) {
      if (P.pre && P.n_events == 4) k_decode<true, MODE_AGG, 2, 4><<<static_cast<unsigned>(nc), NT, decode_smem_bytes<MODE_AGG, 2>(P.tile_bytes), ctx->stream>>>(P);
      else if (P.pre) k_decode<true, MODE_AGG, 2><<<static_cast<unsigned>(nc), NT, decode_smem_bytes<MODE_AGG, 2>(P.tile_bytes), ctx->stream>>>(P);
      else if (P.n_events == 8) k_decode<true, MODE_AGG, 0, 8><<<static_cast<unsigned>(nc), NT, decode_smem_bytes<MODE_AGG, 0>(P.tile_bytes), ctx->stream>>>(P);
      else if (P.n_events == 4) k_decode<true, MODE_AGG, 0, 4><<<static_cast<unsigned>(nc), NT, decode_smem_bytes<MODE_AGG, 0>(P.tile_bytes), ctx->stream>>>(P);
      else k_decode<true, MODE_AGG, 0><<<static_cast<unsigned>(nc), NT, decode_smem_bytes<MODE_AGG, 0>(P.tile_bytes), ctx->stream>>>(P);
      ++launches;
    }
    if (timed) CK(cudaEventRecord(ctx->ev_dec1[ctx->ev_pending++], ctx->stream));
    if (!force_exact && !external) {
      // The derive step starts on the second stream as soon as the sums are there, beside the
      // verification and the (nearly always idle) launches of the exact path; if that path does
      // run, a second derive pass follows it (below), else that pass returns at once.
      CK(cudaEventRecord(ctx->ev_fork, ctx->stream));
      CK(cudaStreamWaitEvent(ctx->aux_stream, ctx->ev_fork, 0));
      if (int r = launch_derive(ctx, ctx->aux_stream)) return r;
      CK(cudaEventRecord(ctx->ev_join, ctx->aux_stream));
      derive_forked = true;
      ++launches;
    }
    if (!force_exact) {
      k_verify<<<static_cast<unsigned>((nc + 255) / 256), 256, 0, ctx->stream>>>(P);
      ++launches;
    }
    k_resolve<<<ns, RES_THREADS, 0, ctx->stream>>>(P);
    k_clear<<<148 * 4, 256, 0, ctx->stream>>>(P, force_exact ? 0 : ctx->acc_words);
    if (P.pre) k_decode<false, MODE_AGG, 2><<<static_cast<unsigned>(nc), NT, decode_smem_bytes<MODE_AGG, 2>(P.tile_bytes), ctx->stream>>>(P);
    else k_decode<false, MODE_AGG, 0><<<static_cast<unsigned>(nc), NT, decode_smem_bytes<MODE_AGG, 0>(P.tile_bytes), ctx->stream>>>(P);
    launches += 3;
    CK(cudaGetLastError());
  }
  if (ctx->accumulate && nc) {
    k_attr_merge<<<148 * 4, 256, 0, ctx->stream>>>(ctx->d_tid_keys, ctx->d_tid_cnt, ctx->tid_slots - 1, ctx->d_tid_keys_tot,
                                                  ctx->d_tid_cnt_tot, ctx->tid_slots - 1, ctx->d_attr_special_tot + 2);
    k_attr_merge<<<148 * 8, 256, 0, ctx->stream>>>(ctx->d_ip_keys, ctx->d_ip_cnt, ctx->ip_slots - 1, ctx->d_ip_keys_tot,
                                                  ctx->d_ip_cnt_tot, ctx->ip_slots - 1, ctx->d_attr_special_tot + 3);
    k_attr_merge_special<<<1, 32, 0, ctx->stream>>>(ctx->d_attr_special, ctx->d_attr_special_tot);
    CK(cudaGetLastError());
    launches += 3;
  }
  if (external) {
    CK(cudaStreamWaitEvent(ctx->stream, ctx->ev_join, 0));
  } else if (derive_forked) {
    CK(cudaStreamWaitEvent(ctx->stream, ctx->ev_join, 0));
    if (int r = launch_derive(ctx, ctx->stream, ctx->d_flags + FLAG_NEED_EXACT)) return r;
    ++launches;
  } else {
    if (int r = launch_derive(ctx)) return r;
    ++launches;
  }
  CK(cudaEventRecord(ctx->ev_run1, ctx->stream));
  ctx->stats.kernel_launches = launches;
  ctx->ran = true;
  ctx->emitted = false;
  ctx->counters_valid = false;
  return 0;
}

int hbtgpu_sync(hbtgpu_ctx* ctx) {
  if (!ctx) return -EINVAL;
  CK(cudaSetDevice(ctx->device));
  CK(cudaStreamSynchronize(ctx->stream));
  CK(cudaStreamSynchronize(ctx->d2h_stream));
  if (int r = peer_poll_error(ctx)) return r;
  if (ctx->ran) CK(cudaEventElapsedTime(&ctx->stats.last_run_ms, ctx->ev_run0, ctx->ev_run1));
  for (int i = 0; i < ctx->ev_pending; ++i) {
    float ms = 0.f;
    CK(cudaEventElapsedTime(&ms, ctx->ev_dec0[i], ctx->ev_dec1[i]));
    ctx->stats.last_decode_ms = ms;
    ctx->stats.decode_ms_sum += ms;
    ctx->stats.decode_runs += 1;
  }
  ctx->ev_pending = 0;
  return 0;
}

int hbtgpu_derive(hbtgpu_ctx* ctx) {
  if (!ctx) return -EINVAL;
  CK(cudaSetDevice(ctx->device));
  return launch_derive(ctx);
}

int hbtgpu_get_counters(hbtgpu_ctx* ctx, uint32_t cpu, hbtgpu_counters* out) {
  if (!ctx || !out) return -EINVAL;
  if (cpu >= ctx->cfg.max_cpus) return fail(ctx, -EINVAL, "cpu %u out of range", cpu);
  if (!ctx->ran) return fail(ctx, -EINVAL, "hbtgpu_run has not been called");
  CK(cudaSetDevice(ctx->device));
  if (int r = fetch_counters(ctx)) return r;
  const uint64_t* c = &ctx->h_counters[static_cast<size_t>(cpu) * C_SLOTS];
  memset(out, 0, sizeof(*out));
  out->lost = c[C_LOST];
  out->throttle = c[C_THROTTLE];
  out->unthrottle = c[C_UNTHROTTLE];
  out->sample = c[C_SAMPLE];
  out->read = c[C_READ];
  // comm/exit/fork/aux/itrace_start/switch_cpu_wide stay 0: Sampling mode has void record
  // types for them (PerfEventsGroup.h:241-308), consume() only skips them by size
  out->num_records = c[C_RECORDS];
  out->num_unknown_count_interval = c[C_UNKNOWN];
  out->num_rows = c[C_ROWS];
  out->rows_out_of_range = c[C_OOR];
  out->err = c[C_ERR] ? -EPERM : 0;
  // data_tail after the walk: the stalled record on error, else the whole stream
  uint64_t stream_len = 0;
  for (const HostImage& im : ctx->images)
    if (im.stream == cpu) stream_len += im.len;
  out->err_offset = c[C_ERR] ? c[C_ERR_OFF] : stream_len;
  out->last_tstamp = ctx->h_carry_out[cpu].valid ? ctx->h_carry_out[cpu].tstamp : ~0ull;
  return 0;
}

int hbtgpu_get_carry(hbtgpu_ctx* ctx, uint32_t cpu, hbtgpu_carry* out) {
  if (!ctx || !out) return -EINVAL;
  if (cpu >= ctx->cfg.max_cpus) return fail(ctx, -EINVAL, "cpu %u out of range", cpu);
  if (!ctx->ran) return fail(ctx, -EINVAL, "hbtgpu_run has not been called");
  CK(cudaSetDevice(ctx->device));
  CK(cudaMemcpyAsync(out, &ctx->d_carry_out[cpu], sizeof(hbtgpu_carry), cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  return 0;
}

int hbtgpu_get_rows(hbtgpu_ctx* ctx, uint32_t cpu, uint64_t* rows, size_t cap_rows, size_t* n_rows) {
  if (!ctx || !n_rows) return -EINVAL;
  if (cpu >= ctx->cfg.max_cpus) return fail(ctx, -EINVAL, "cpu %u out of range", cpu);
  CK(cudaSetDevice(ctx->device));
  if (int r = ensure_emitted(ctx)) return r;
  const uint64_t have = ctx->stream_rows[cpu];
  *n_rows = have;
  const uint64_t k = std::min<uint64_t>(have, cap_rows);
  if (k && rows) {
    const uint32_t w = ctx->cfg.n_events + 2;
    CK(cudaMemcpyAsync(rows, ctx->d_rows + ctx->stream_row_base[cpu] * w, k * w * 8,
                       cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
  }
  return 0;
}

int hbtgpu_rows_device(hbtgpu_ctx* ctx, const uint64_t** rows, uint64_t* row_begin, uint32_t* row_stride) {
  if (!ctx || !rows || !row_begin) return -EINVAL;
  CK(cudaSetDevice(ctx->device));
  if (int r = ensure_emitted(ctx)) return r;
  // the row table is dense, stream after stream in cpu order (the chunks are laid out that way)
  uint64_t run = 0;
  for (uint32_t cpu = 0; cpu < ctx->cfg.max_cpus; ++cpu) {
    row_begin[cpu] = run;
    if (ctx->stream_rows[cpu] && ctx->stream_row_base[cpu] != run)
      return fail(ctx, -ENOTSUP, "rows of cpu %u are not where the dense layout puts them", cpu);
    run += ctx->stream_rows[cpu];
  }
  row_begin[ctx->cfg.max_cpus] = run;
  *rows = ctx->d_rows;
  if (row_stride) *row_stride = ctx->cfg.n_events + 2;
  return 0;
}

int hbtgpu_row_times_device(hbtgpu_ctx* ctx, const uint64_t** times, uint64_t* row_begin) {
  if (!ctx || !times) return -EINVAL;
  const uint64_t* rows = nullptr;
  std::vector<uint64_t> rb(ctx->cfg.max_cpus + 1);
  if (int r = hbtgpu_rows_device(ctx, &rows, row_begin ? row_begin : rb.data(), nullptr)) return r;
  *times = ctx->d_row_times;
  return 0;
}

int hbtgpu_get_records(hbtgpu_ctx* ctx, uint32_t cpu, uint64_t* off, uint32_t* type, uint32_t* size,
                       size_t cap, size_t* n_records) {
  if (!ctx || !n_records) return -EINVAL;
  if (cpu >= ctx->cfg.max_cpus) return fail(ctx, -EINVAL, "cpu %u out of range", cpu);
  CK(cudaSetDevice(ctx->device));
  if (int r = ensure_emitted(ctx)) return r;
  const uint64_t have = ctx->stream_recs[cpu];
  *n_records = have;
  const uint64_t k = std::min<uint64_t>(have, cap);
  const uint64_t b = ctx->stream_rec_base[cpu];
  if (k) {
    if (off) CK(cudaMemcpyAsync(off, ctx->d_rec_off + b, k * 8, cudaMemcpyDeviceToHost, ctx->stream));
    if (type) CK(cudaMemcpyAsync(type, ctx->d_rec_type + b, k * 4, cudaMemcpyDeviceToHost, ctx->stream));
    if (size) CK(cudaMemcpyAsync(size, ctx->d_rec_size + b, k * 4, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
  }
  return 0;
}

int hbtgpu_get_tid_histogram(hbtgpu_ctx* ctx, uint32_t* pid, uint32_t* tid, uint64_t* count, size_t cap,
                             size_t* n) {
  if (!ctx || !n) return -EINVAL;
  if (!ctx->tid_slots) return fail(ctx, -EINVAL, "context was not created with HBTGPU_SAMPLE_IP_TID_TIME_READ");
  if (!ctx->ran) return fail(ctx, -EINVAL, "hbtgpu_run has not been called");
  CK(cudaSetDevice(ctx->device));
  if (int r = fetch_counters(ctx)) return r;  // surfaces exact-path / internal errors first
  std::vector<uint64_t> keys(ctx->tid_slots), cnt(ctx->tid_slots);
  unsigned long long special[4];
  const AttrView av = attr_view(ctx);
  CK(cudaMemcpyAsync(keys.data(), av.tid_keys, ctx->tid_slots * 8ull, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaMemcpyAsync(cnt.data(), av.tid_cnt, ctx->tid_slots * 8ull, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaMemcpyAsync(special, av.special, sizeof(special), cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  if (special[2]) return fail(ctx, -ENOSPC, "(pid,tid) table overflowed: raise hbtgpu_config.tid_table_log2");
  size_t k = 0;
  auto put = [&](uint64_t key, uint64_t c) {
    if (k < cap) {
      if (pid) pid[k] = static_cast<uint32_t>(key);
      if (tid) tid[k] = static_cast<uint32_t>(key >> 32);
      if (count) count[k] = c;
    }
    ++k;
  };
  for (uint32_t i = 0; i < ctx->tid_slots; ++i)
    if (keys[i] != ~0ull) put(keys[i], cnt[i]);
  if (special[0]) put(~0ull, special[0]);
  *n = k;
  return 0;
}

int hbtgpu_get_top_ips(hbtgpu_ctx* ctx, uint32_t k, uint64_t* ip, uint64_t* count, size_t* n) {
  if (!ctx || !n) return -EINVAL;
  if (!ctx->ip_slots) return fail(ctx, -EINVAL, "context was not created with HBTGPU_SAMPLE_IP_TID_TIME_READ");
  if (!ctx->ran) return fail(ctx, -EINVAL, "hbtgpu_run has not been called");
  if (k > 65536) return fail(ctx, -EINVAL, "k must be <= 65536");
  CK(cudaSetDevice(ctx->device));
  if (int r = fetch_counters(ctx)) return r;
  *n = 0;
  if (k == 0) return 0;
  // scratch: compacted (key, count) of the occupied slots, twice (the sorts below go from one copy
  // to the other), the sort's own, a counter
  const size_t slots = ctx->ip_slots;
  size_t tmp_sort = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, tmp_sort, static_cast<uint64_t*>(nullptr), static_cast<uint64_t*>(nullptr),
                                  static_cast<unsigned long long*>(nullptr), static_cast<unsigned long long*>(nullptr),
                                  static_cast<int>(slots + 1), 0, 64, ctx->stream);
  const size_t bytes = (slots + 1) * 32 + tmp_sort + 256;
  MathBuf mb;
  if (int r = math_reserve(ctx, bytes + 8 * 256, &mb)) return r;
  uint64_t* ckey = mb.take<uint64_t>(slots + 1);
  unsigned long long* ccnt = mb.take<unsigned long long>(slots + 1);
  uint64_t* skey = mb.take<uint64_t>(slots + 1);
  unsigned long long* scnt = mb.take<unsigned long long>(slots + 1);
  uint8_t* d_tmp = mb.take<uint8_t>(tmp_sort);
  unsigned int* d_m = mb.take<unsigned int>(1);
  CK(cudaMemsetAsync(d_m, 0, 4, ctx->stream));
  const AttrView av = attr_view(ctx);
  k_attr_compact<<<148 * 8, 256, 0, ctx->stream>>>(av.ip_keys, av.ip_cnt, ctx->ip_slots - 1, ckey, ccnt, d_m);
  unsigned int m = 0;
  unsigned long long special[4];
  CK(cudaMemcpyAsync(&m, d_m, 4, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaMemcpyAsync(special, av.special, sizeof(special), cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  if (special[3]) return fail(ctx, -ENOSPC, "IP table overflowed: raise hbtgpu_config.ip_table_log2");
  if (special[1]) {  // samples whose ip equals the empty-slot marker live in a side counter
    const uint64_t key = ~0ull;
    CK(cudaMemcpyAsync(ckey + m, &key, 8, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(ccnt + m, &special[1], 8, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    ++m;
  }
  const uint32_t kk = std::min<uint32_t>(k, m);
  if (kk) {
    // Order (count descending, ip ascending): a radix sort by ip, then a STABLE one by count from
    // the top -- equal counts keep the order of their ips.  Two sorts of a few hundred thousand
    // pairs, whatever k is (round 1 ran k rounds of arg-max over the table: 1.3 ms at k = 100,
    // most of a second at the largest k the call takes).  Library sort (cub), as in the dictionary.
    size_t t = tmp_sort;
    cub::DeviceRadixSort::SortPairs(d_tmp, t, ckey, skey, ccnt, scnt, static_cast<int>(m), 0, 64, ctx->stream);
    t = tmp_sort;
    cub::DeviceRadixSort::SortPairsDescending(d_tmp, t, scnt, ccnt, skey, ckey, static_cast<int>(m), 0, 64, ctx->stream);
    CK(cudaGetLastError());
    if (ip) CK(cudaMemcpyAsync(ip, ckey, kk * 8ull, cudaMemcpyDeviceToHost, ctx->stream));
    if (count) CK(cudaMemcpyAsync(count, ccnt, kk * 8ull, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
  }
  *n = kk;
  return 0;
}

int hbtgpu_attr_clear(hbtgpu_ctx* ctx) {
  if (!ctx) return -EINVAL;
  if (!ctx->accumulate) return fail(ctx, -EINVAL, "context was not created with HBTGPU_F_ACCUMULATE_ATTR");
  CK(cudaSetDevice(ctx->device));
  CK(cudaMemsetAsync(ctx->d_tid_keys_tot, 0xFF, ctx->tid_slots * 8ull, ctx->stream));
  CK(cudaMemsetAsync(ctx->d_tid_cnt_tot, 0, ctx->tid_slots * 8ull, ctx->stream));
  CK(cudaMemsetAsync(ctx->d_ip_keys_tot, 0xFF, ctx->ip_slots * 8ull, ctx->stream));
  CK(cudaMemsetAsync(ctx->d_ip_cnt_tot, 0, ctx->ip_slots * 8ull, ctx->stream));
  CK(cudaMemsetAsync(ctx->d_attr_special_tot, 0, 4 * 8, ctx->stream));
  ctx->attr_all_valid = false;
  return 0;
}

int hbtgpu_get_buckets_device(hbtgpu_ctx* ctx, hbtgpu_bucket_view* v) {
  if (!ctx || !v) return -EINVAL;
  v->acc = ctx->acc_cur;
  v->scaled = ctx->d_scaled;
  v->metric = ctx->d_metric;
  v->valid = ctx->d_valid;
  v->max_cpus = ctx->cfg.max_cpus;
  v->n_buckets = ctx->cfg.n_buckets;
  v->n_events = ctx->cfg.n_events;
  v->n_metrics = ctx->cfg.n_metrics;
  return 0;
}

int hbtgpu_get_buckets_host(hbtgpu_ctx* ctx, uint64_t* acc, uint64_t* scaled, double* metric,
                            uint32_t* valid) {
  if (!ctx) return -EINVAL;
  CK(cudaSetDevice(ctx->device));
  const uint32_t n = ctx->cfg.n_events;
  if (acc) CK(cudaMemcpyAsync(acc, ctx->acc_cur, ctx->acc_words * 8, cudaMemcpyDeviceToHost, ctx->stream));
  if (scaled) CK(cudaMemcpyAsync(scaled, ctx->d_scaled, ctx->n_cells * n * 8, cudaMemcpyDeviceToHost, ctx->stream));
  if (metric && ctx->cfg.n_metrics)
    CK(cudaMemcpyAsync(metric, ctx->d_metric, ctx->n_cells * ctx->cfg.n_metrics * 8,
                       cudaMemcpyDeviceToHost, ctx->stream));
  if (valid) CK(cudaMemcpyAsync(valid, ctx->d_valid, ctx->n_cells * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  return 0;
}

int hbtgpu_get_buckets_host_async(hbtgpu_ctx* ctx, uint64_t* acc, uint64_t* scaled, double* metric,
                                  uint32_t* valid) {
  if (!ctx) return -EINVAL;
  CK(cudaSetDevice(ctx->device));
  const uint32_t n = ctx->cfg.n_events;
  CK(cudaEventRecord(ctx->ev_results, ctx->stream));
  CK(cudaStreamWaitEvent(ctx->d2h_stream, ctx->ev_results, 0));
  cudaStream_t s2 = ctx->d2h_stream;
  if (acc) CK(cudaMemcpyAsync(acc, ctx->acc_cur, ctx->acc_words * 8, cudaMemcpyDeviceToHost, s2));
  if (scaled) CK(cudaMemcpyAsync(scaled, ctx->d_scaled, ctx->n_cells * n * 8, cudaMemcpyDeviceToHost, s2));
  if (metric && ctx->cfg.n_metrics)
    CK(cudaMemcpyAsync(metric, ctx->d_metric, ctx->n_cells * ctx->cfg.n_metrics * 8, cudaMemcpyDeviceToHost, s2));
  if (valid) CK(cudaMemcpyAsync(valid, ctx->d_valid, ctx->n_cells * 4, cudaMemcpyDeviceToHost, s2));
  CK(cudaEventRecord(ctx->ev_d2h, s2));
  ctx->d2h_pending = true;
  return 0;
}

int hbtgpu_get_stubs(hbtgpu_ctx* ctx, hbtgpu_head_stub* head, hbtgpu_carry* tail) {
  if (!ctx || !head || !tail) return -EINVAL;
  if (!ctx->ran) return fail(ctx, -EINVAL, "hbtgpu_run has not been called");
  if (!(ctx->cfg.flags & HBTGPU_F_EXTERNAL_CARRY))
    return fail(ctx, -EINVAL, "context was not created with HBTGPU_F_EXTERNAL_CARRY");
  CK(cudaSetDevice(ctx->device));
  const uint32_t ns = ctx->cfg.max_cpus;
  CK(cudaMemcpyAsync(head, ctx->d_head, ns * sizeof(hbtgpu_head_stub), cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaMemcpyAsync(tail, ctx->d_carry_out, ns * sizeof(hbtgpu_carry), cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  return 0;
}

void hbtgpu_compose_tail(const hbtgpu_carry* prev, const hbtgpu_head_stub* head,
                         const hbtgpu_carry* own_tail, hbtgpu_carry* out) {
  if (head->has_sample) {
    *out = *own_tail;  // the last SAMPLE (and any LOST after it) decides, whatever came before
  } else {
    hbtgpu_carry t = *prev;  // last_[] survives; a LOST invalidates count_.tstamp
    if (head->lost_before) {
      t.valid = 0;
      t.tstamp = ~0ull;
    }
    *out = t;
  }
}

int hbtgpu_apply_boundaries(hbtgpu_ctx* ctx, const hbtgpu_carry* prev_tails) {
  if (!ctx || !prev_tails) return -EINVAL;
  if (!ctx->ran) return fail(ctx, -EINVAL, "hbtgpu_run has not been called");
  if (!(ctx->cfg.flags & HBTGPU_F_EXTERNAL_CARRY))
    return fail(ctx, -EINVAL, "context was not created with HBTGPU_F_EXTERNAL_CARRY");
  CK(cudaSetDevice(ctx->device));
  const uint32_t ns = ctx->cfg.max_cpus;
  // d_carry_in is free in shard mode (the decode used an invalid state): stage prev_tails there
  CK(cudaMemcpyAsync(ctx->d_carry_in, prev_tails, ns * sizeof(hbtgpu_carry), cudaMemcpyHostToDevice, ctx->stream));
  Params P = make_params(ctx);
  k_boundary<<<ns, 128, 0, ctx->stream>>>(P, ctx->d_carry_in, nullptr, ctx->d_boundary_cell);
  CK(cudaGetLastError());
  // streams without a SAMPLE pass the previous state through: fix their carry-out on the host side
  CK(cudaStreamSynchronize(ctx->stream));  // prev_tails may be reused by the caller after this
  ctx->counters_valid = false;
  ctx->emitted = false;
  return derive_after_boundaries(ctx);
}

size_t hbtgpu_stub_bytes(const hbtgpu_ctx* ctx) {
  return ctx ? static_cast<size_t>(ctx->cfg.max_cpus) * stub_words(ctx->cfg.n_events) * 8 : 0;
}

int hbtgpu_export_stubs_device(hbtgpu_ctx* ctx, void* device_dst) {
  if (!ctx || !device_dst) return -EINVAL;
  if (!ctx->ran) return fail(ctx, -EINVAL, "hbtgpu_run has not been called");
  if (!(ctx->cfg.flags & HBTGPU_F_EXTERNAL_CARRY))
    return fail(ctx, -EINVAL, "context was not created with HBTGPU_F_EXTERNAL_CARRY");
  CK(cudaSetDevice(ctx->device));
  const uint32_t total = ctx->cfg.max_cpus * stub_words(ctx->cfg.n_events);
  k_pack_stubs<<<(total + 255) / 256, 256, 0, ctx->stream>>>(make_params(ctx), static_cast<uint64_t*>(device_dst));
  CK(cudaGetLastError());
  ctx->stats.kernel_launches += 1;
  return 0;
}

int hbtgpu_apply_boundaries_gathered(hbtgpu_ctx* ctx, const void* device_gathered, uint32_t world,
                                     uint32_t rank) {
  if (!ctx || !device_gathered) return -EINVAL;
  if (rank >= world) return fail(ctx, -EINVAL, "rank %u >= world %u", rank, world);
  if (!ctx->ran) return fail(ctx, -EINVAL, "hbtgpu_run has not been called");
  if (!(ctx->cfg.flags & HBTGPU_F_EXTERNAL_CARRY))
    return fail(ctx, -EINVAL, "context was not created with HBTGPU_F_EXTERNAL_CARRY");
  CK(cudaSetDevice(ctx->device));
  const uint32_t ns = ctx->cfg.max_cpus;
  Params P = make_params(ctx);
  // d_carry_in is free in shard mode (the decode used an invalid state): it receives the fold
  k_fold<<<(ns + 127) / 128, 128, 0, ctx->stream>>>(P, static_cast<const uint64_t*>(device_gathered), world, rank,
                                                    ctx->d_carry_init, ctx->d_carry_in, ctx->d_dead, nullptr);
  k_boundary<<<ns, 128, 0, ctx->stream>>>(P, ctx->d_carry_in, ctx->d_dead, ctx->d_boundary_cell);
  CK(cudaGetLastError());
  ctx->counters_valid = false;
  ctx->emitted = false;
  ctx->stats.kernel_launches += 3;
  return derive_after_boundaries(ctx);
}

int hbtgpu_get_prev_tail(hbtgpu_ctx* ctx, uint32_t cpu, hbtgpu_carry* out) {
  if (!ctx || !out) return -EINVAL;
  if (cpu >= ctx->cfg.max_cpus) return fail(ctx, -EINVAL, "cpu %u out of range", cpu);
  if (!(ctx->cfg.flags & HBTGPU_F_EXTERNAL_CARRY))
    return fail(ctx, -EINVAL, "context was not created with HBTGPU_F_EXTERNAL_CARRY");
  CK(cudaSetDevice(ctx->device));
  CK(cudaMemcpyAsync(out, &ctx->d_carry_in[cpu], sizeof(hbtgpu_carry), cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  return 0;
}

int hbtgpu_acc_device(hbtgpu_ctx* ctx, void** device_ptr, size_t* bytes) {
  if (!ctx || !device_ptr || !bytes) return -EINVAL;
  *device_ptr = ctx->d_acc;
  *bytes = ctx->acc_words * 8;
  return 0;
}

void* hbtgpu_stream(hbtgpu_ctx* ctx) { return ctx ? ctx->stream : nullptr; }

int hbtgpu_get_stats(hbtgpu_ctx* ctx, hbtgpu_stats* out) {
  if (!ctx || !out) return -EINVAL;
  if (ctx->ran && !ctx->counters_valid) {  // exact_runs / exact_reason come from device flags
    CK(cudaSetDevice(ctx->device));
    if (int r = fetch_counters(ctx)) return r;
  }
  *out = ctx->stats;
  return 0;
}

}  // extern "C"

// ---- elementwise arithmetic ---------------------------------------------------------


extern "C" {

#define H2D(dst, src, bytes) CK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, ctx->stream))
#define D2H(dst, src, bytes) CK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, ctx->stream))

int hbtgpu_math_get_count(hbtgpu_ctx* ctx, const uint64_t* count, const uint64_t* te,
                          const uint64_t* tr, uint64_t* out, size_t n) {
  if (!ctx) return -EINVAL;
  if (n == 0) return 0;
  MathBuf mb;
  if (int r = math_reserve(ctx, 4 * pad256(n * 8), &mb)) return r;
  uint64_t *dc = mb.take<uint64_t>(n), *de = mb.take<uint64_t>(n), *dr = mb.take<uint64_t>(n),
           *dout = mb.take<uint64_t>(n);
  H2D(dc, count, n * 8);
  H2D(de, te, n * 8);
  H2D(dr, tr, n * 8);
  k_get_count<<<nblk(n), 256, 0, ctx->stream>>>(dc, de, dr, dout, n);
  D2H(out, dout, n * 8);
  CK(cudaStreamSynchronize(ctx->stream));
  return 0;
}

int hbtgpu_math_rate(hbtgpu_ctx* ctx, const uint64_t* num, const uint64_t* den, double* out,
                     uint8_t* valid, size_t n) {
  if (!ctx) return -EINVAL;
  if (n == 0) return 0;
  MathBuf mb;
  if (int r = math_reserve(ctx, 3 * pad256(n * 8) + pad256(n), &mb)) return r;
  uint64_t *dn = mb.take<uint64_t>(n), *dd = mb.take<uint64_t>(n);
  double* dout = mb.take<double>(n);
  uint8_t* dv = mb.take<uint8_t>(n);
  H2D(dn, num, n * 8);
  H2D(dd, den, n * 8);
  k_rate<<<nblk(n), 256, 0, ctx->stream>>>(dn, dd, dout, dv, n);
  D2H(out, dout, n * 8);
  D2H(valid, dv, n);
  CK(cudaStreamSynchronize(ctx->stream));
  return 0;
}

int hbtgpu_math_add_reduce(hbtgpu_ctx* ctx, const uint64_t* vals, size_t k, double* out, size_t n) {
  if (!ctx) return -EINVAL;
  if (n == 0) return 0;
  MathBuf mb;
  if (int r = math_reserve(ctx, pad256(n * k * 8) + pad256(n * 8), &mb)) return r;
  uint64_t* dv = mb.take<uint64_t>(n * k);
  double* dout = mb.take<double>(n);
  H2D(dv, vals, n * k * 8);
  k_add_reduce<<<nblk(n), 256, 0, ctx->stream>>>(dv, k, dout, n);
  D2H(out, dout, n * 8);
  CK(cudaStreamSynchronize(ctx->stream));
  return 0;
}

static int math_tsc(hbtgpu_ctx* ctx, const uint64_t* in, uint64_t* out, size_t n, int inverse) {
  if (!ctx) return -EINVAL;
  if (n == 0) return 0;
  if (inverse && ctx->cfg.time_mult == 0) return fail(ctx, -EINVAL, "time_mult must be != 0");
  MathBuf mb;
  if (int r = math_reserve(ctx, 2 * pad256(n * 8), &mb)) return r;
  uint64_t *di = mb.take<uint64_t>(n), *dout = mb.take<uint64_t>(n);
  H2D(di, in, n * 8);
  k_tsc<<<nblk(n), 256, 0, ctx->stream>>>(di, dout, n, ctx->cfg.time_shift, ctx->cfg.time_mult,
                                          ctx->cfg.time_zero, inverse);
  D2H(out, dout, n * 8);
  CK(cudaStreamSynchronize(ctx->stream));
  return 0;
}
int hbtgpu_math_tsc_to_ktime(hbtgpu_ctx* ctx, const uint64_t* tsc, uint64_t* out, size_t n) {
  return math_tsc(ctx, tsc, out, n, 0);
}
int hbtgpu_math_ktime_to_tsc(hbtgpu_ctx* ctx, const uint64_t* kt, uint64_t* out, size_t n) {
  return math_tsc(ctx, kt, out, n, 1);
}

int hbtgpu_math_per_second_mega(hbtgpu_ctx* ctx, const uint64_t* count, const uint64_t* tr,
                                float* out, size_t n) {
  if (!ctx) return -EINVAL;
  if (n == 0) return 0;
  MathBuf mb;
  if (int r = math_reserve(ctx, 2 * pad256(n * 8) + pad256(n * 4), &mb)) return r;
  uint64_t *dc = mb.take<uint64_t>(n), *dr = mb.take<uint64_t>(n);
  float* dout = mb.take<float>(n);
  H2D(dc, count, n * 8);
  H2D(dr, tr, n * 8);
  k_per_second_mega<<<nblk(n), 256, 0, ctx->stream>>>(dc, dr, dout, n);
  D2H(out, dout, n * 4);
  CK(cudaStreamSynchronize(ctx->stream));
  return 0;
}

int hbtgpu_math_running_ratio(hbtgpu_ctx* ctx, const uint64_t* te, const uint64_t* tr, float* ratio,
                              uint64_t* ratio_int100, uint8_t* has_mux, size_t n) {
  if (!ctx) return -EINVAL;
  if (n == 0) return 0;
  MathBuf mb;
  if (int r = math_reserve(ctx, 3 * pad256(n * 8) + pad256(n * 4) + pad256(n), &mb)) return r;
  uint64_t *de = mb.take<uint64_t>(n), *dr = mb.take<uint64_t>(n), *dri = mb.take<uint64_t>(n);
  float* drat = mb.take<float>(n);
  uint8_t* dm = mb.take<uint8_t>(n);
  H2D(de, te, n * 8);
  H2D(dr, tr, n * 8);
  k_running_ratio<<<nblk(n), 256, 0, ctx->stream>>>(de, dr, drat, dri, dm, n);
  D2H(ratio, drat, n * 4);
  D2H(ratio_int100, dri, n * 8);
  D2H(has_mux, dm, n);
  CK(cudaStreamSynchronize(ctx->stream));
  return 0;
}

}  // extern "C"


// =======================================================================================
// hbtgpu_decode_image_noalloc: the decode pipeline for callers that cannot allocate.
//
// Everything the kernels need besides the caller's image and accumulator lives in this one
// file-scope device array.  In host code the NAME of a __device__ variable is only a handle for
// the runtime, not an address a kernel could use; the usable address is fetched once per device
// with cudaGetSymbolAddress (below) and handed to the kernels through Params.
// =======================================================================================
namespace {

constexpr uint32_t kNaMaxChunks = 1024;
struct NaScratch {                       // layout of g_noalloc_scratch
  uint32_t flags[FLAG_COUNT];            //   must start at zero each call
  uint64_t counters[C_SLOTS];            //   must start at zero each call
  hbtgpu_carry carry_in, carry_out;
  hbtgpu_head_stub head;
  StreamDesc stream;
  ImageDesc image;
  ChunkDesc chunks[kNaMaxChunks];
  ChunkIO io[kNaMaxChunks];              //   must start at zero each call
};
constexpr size_t kNaScratchBytes = (sizeof(NaScratch) + 255) / 256 * 256;

static __device__ __align__(256) unsigned char g_noalloc_scratch[kNaScratchBytes];

// Fills the tables for one image of one stream and clears what has to start from zero.
__global__ void k_na_setup(NaScratch* sc, uint64_t image_bytes, uint64_t chunk_bytes,
                           uint32_t n_chunks, hbtgpu_carry carry_in) {
  const uint32_t tid = threadIdx.x;
  if (tid < FLAG_COUNT) sc->flags[tid] = 0;
  if (tid < C_SLOTS) sc->counters[tid] = 0;
  if (tid == 0) {
    sc->carry_in = carry_in;
    sc->carry_out = carry_in;  // an empty image leaves the state as it was
    sc->stream = StreamDesc{0, image_bytes ? 1u : 0u};
    sc->image = ImageDesc{0, image_bytes, 0, 0, 0, n_chunks, 0};
  }
  for (uint32_t c = tid; c < n_chunks; c += blockDim.x) {
    ChunkDesc cd{};
    cd.img_base = 0;
    cd.img_len = image_bytes;
    cd.begin = static_cast<uint64_t>(c) * chunk_bytes;
    cd.end = cd.begin + chunk_bytes < image_bytes ? cd.begin + chunk_bytes : image_bytes;
    cd.stream_pos = 0;
    cd.stream = 0;
    cd.image = 0;
    cd.flags = (c == 0 ? (CH_FIRST_OF_IMAGE | CH_FIRST_OF_STREAM) : 0u) |
               (c + 1 == n_chunks ? (CH_LAST_OF_IMAGE | CH_LAST_OF_STREAM) : 0u);
    if (c > 0) {
      const uint64_t hb = cd.begin > HALO_BYTES ? cd.begin - HALO_BYTES : 0;
      cd.halo_off = hb;
      cd.halo_len = static_cast<uint32_t>(cd.begin - hb);
      if (hb == 0) cd.flags |= CH_HALO_EXACT_START;
    }
    sc->chunks[c] = cd;
    ChunkIO z{};
    sc->io[c] = z;
  }
}

thread_local char g_na_error[512];
int na_fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_na_error, sizeof(g_na_error), fmt, ap);
  va_end(ap);
  return code;
}

constexpr int kNaMaxDevices = 64;
NaScratch* g_na_ptr[kNaMaxDevices] = {nullptr};  // device address of the array, per device
std::mutex g_na_mutex;

// Once per device, with that device current: resolve the array's device address.
int na_device_setup(int dev, NaScratch** out) {
  std::lock_guard<std::mutex> lock(g_na_mutex);
  if (!g_na_ptr[dev]) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess || !p)
      return na_fail(-EIO, "cudaGetSymbolAddress(scratch) failed on device %d: %s", dev,
                     cudaGetErrorString(e));
    for (auto k : {(const void*)k_decode<true, MODE_AGG, 0>, (const void*)k_decode<true, MODE_AGG, 0, 8>,
                   (const void*)k_decode<true, MODE_AGG, 0, 4>, (const void*)k_decode<false, MODE_AGG, 0>}) {
      e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               static_cast<int>(decode_smem_bytes<MODE_AGG, 0>(TILE_MAX_BYTES)));
      if (e != cudaSuccess)
        return na_fail(-EIO, "cudaFuncSetAttribute failed on device %d: %s", dev, cudaGetErrorString(e));
    }
    g_na_ptr[dev] = static_cast<NaScratch*>(p);
  }
  *out = g_na_ptr[dev];
  return 0;
}

}  // namespace

extern "C" {

const char* hbtgpu_noalloc_last_error(void) { return g_na_error; }

int hbtgpu_decode_image_noalloc(const hbtgpu_config* cfg, const hbtgpu_carry* carry_in,
                                const void* device_image, size_t image_bytes,
                                uint64_t* device_acc, hbtgpu_counters* counters_out,
                                hbtgpu_carry* carry_out, void* cuda_stream) {
#define NCK(call)                                                                        \
  do {                                                                                   \
    cudaError_t e_ = (call);                                                             \
    if (e_ != cudaSuccess) return na_fail(-EIO, "%s failed: %s", #call, cudaGetErrorString(e_)); \
  } while (0)
  g_na_error[0] = 0;
  if (!cfg || !device_acc || !counters_out) return na_fail(-EINVAL, "null argument");
  if (image_bytes && !device_image) return na_fail(-EINVAL, "null image");
  if (cfg->n_events == 0 || cfg->n_events > HBTGPU_MAX_EVENTS)
    return na_fail(-EINVAL, "n_events must be in [1,%d]", HBTGPU_MAX_EVENTS);
  if (cfg->n_buckets == 0 || cfg->n_buckets >= 0xFFFFFFFEu || cfg->bucket_ns == 0)
    return na_fail(-EINVAL, "n_buckets / bucket_ns out of range");
  if (cfg->time_shift > 63 || (cfg->tsc_mask != 0 && cfg->time_mult == 0))
    return na_fail(-EINVAL, "bad TSC conversion parameters");
  if (cfg->sample_format != HBTGPU_SAMPLE_TIME_READ)
    return na_fail(-EINVAL, "only HBTGPU_SAMPLE_TIME_READ is supported here (the IP|TID tables need allocations)");
  if (image_bytes > HBTGPU_NOALLOC_MAX_IMAGE_BYTES)
    return na_fail(-E2BIG,
                   "image of %zu bytes exceeds HBTGPU_NOALLOC_MAX_IMAGE_BYTES (%llu): this entry "
                   "point works out of a fixed static scratch array; use a context for larger images",
                   image_bytes, (unsigned long long)HBTGPU_NOALLOC_MAX_IMAGE_BYTES);
  if (image_bytes % 8) return na_fail(-EINVAL, "image size %zu is not a multiple of 8", image_bytes);
  if (reinterpret_cast<uintptr_t>(device_image) % 16)
    return na_fail(-EINVAL, "device_image must be 16-byte aligned");
  if (cfg->device < 0 || cfg->device >= kNaMaxDevices) return na_fail(-ENODEV, "device %d out of range", cfg->device);
  NCK(cudaSetDevice(cfg->device));
  NaScratch* sc = nullptr;
  if (int r = na_device_setup(cfg->device, &sc)) return r;
  cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);

  // chunks: as many as fill the GPU, never more than the static table holds
  const uint32_t tile = cfg_tile_bytes(*cfg);
  uint64_t cb = chunk_bytes_for(image_bytes / (148ull * 12ull), 512ull << 10, tile);
  while ((image_bytes + cb - 1) / cb > kNaMaxChunks) cb <<= 1;
  const uint32_t nc = static_cast<uint32_t>((image_bytes + cb - 1) / cb);

  const uint32_t n = cfg->n_events;
  const uint64_t acc_words = static_cast<uint64_t>(cfg->n_buckets) * (n + 3);
  hbtgpu_carry cin{};
  if (carry_in) cin = *carry_in;
  NCK(cudaMemsetAsync(device_acc, 0, acc_words * 8, st));
  k_na_setup<<<1, 256, 0, st>>>(sc, image_bytes, cb, nc, cin);

  Params P{};
  P.arena = static_cast<const uint8_t*>(device_image);
  P.chunks = sc->chunks;
  P.io = sc->io;
  P.images = &sc->image;
  P.streams = &sc->stream;
  P.carry_in = &sc->carry_in;
  P.carry_out = &sc->carry_out;
  P.head = &sc->head;
  P.counters = sc->counters;
  P.acc = device_acc;
  P.flags = sc->flags;
  P.t0 = cfg->t0;
  P.bucket_ns = cfg->bucket_ns;
  P.tsc_mask = cfg->tsc_mask;
  P.time_zero = cfg->time_zero;
  P.n_events = n;
  P.n_buckets = cfg->n_buckets;
  P.n_chunks = nc;
  P.n_streams = 1;
  P.time_shift = cfg->time_shift;
  P.time_mult = cfg->time_mult;
  P.force_general = (cfg->flags & HBTGPU_F_FORCE_GENERAL) ? 1 : 0;
  {
    const unsigned __int128 range = static_cast<unsigned __int128>(cfg->n_buckets) * cfg->bucket_ns;
    P.range_ns = range > ~0ull ? ~0ull : static_cast<uint64_t>(range);
    P.bucket_magic = static_cast<uint64_t>((static_cast<unsigned __int128>(1) << 64) / cfg->bucket_ns);
    const uint64_t b = cfg->bucket_ns;
    P.bucket_shift = (b & (b - 1)) == 0 ? static_cast<uint32_t>(__builtin_ctzll(b)) : 0xFFFFFFFFu;
    P.bucket_f64 = (b < (1ull << 32) && P.range_ns < (1ull << 53)) ? 1u : 0u;
    P.bucket_inv = 1.0 / static_cast<double>(b);
    P.pre = cfg->sample_format == HBTGPU_SAMPLE_IP_TID_TIME_READ ? 2 : 0;
    P.fix = 40 + 8 * P.pre;
    P.tile_bytes = tile;
    const uint64_t S = 5ull + P.pre + n;
    P.s_magic = static_cast<uint32_t>(((1ull << 32) + S - 1) / S);
  }
  if (nc) {
    const bool force_exact = (cfg->flags & HBTGPU_F_FORCE_EXACT) != 0;
    if (force_exact) NCK(cudaMemsetAsync(&sc->flags[FLAG_NEED_EXACT], 1, 1, st));
    if (!force_exact) {
      // the same kernels hbtgpu_run launches: the builds specialised for 8- and 4-event groups where they apply
      if (n == 8) k_decode<true, MODE_AGG, 0, 8><<<nc, NT, decode_smem_bytes<MODE_AGG, 0>(P.tile_bytes), st>>>(P);
      else if (n == 4) k_decode<true, MODE_AGG, 0, 4><<<nc, NT, decode_smem_bytes<MODE_AGG, 0>(P.tile_bytes), st>>>(P);
      else k_decode<true, MODE_AGG, 0><<<nc, NT, decode_smem_bytes<MODE_AGG, 0>(P.tile_bytes), st>>>(P);
      k_verify<<<(nc + 255) / 256, 256, 0, st>>>(P);
    }
    k_resolve<<<1, RES_THREADS, 0, st>>>(P);
    k_clear<<<148 * 4, 256, 0, st>>>(P, force_exact ? 0 : acc_words);
    k_decode<false, MODE_AGG, 0><<<nc, NT, decode_smem_bytes<MODE_AGG, 0>(P.tile_bytes), st>>>(P);
  }
  NCK(cudaGetLastError());

  // results: three small device-to-host copies into stack / caller memory
  uint64_t ctr[C_SLOTS];
  uint32_t flags[FLAG_COUNT];
  hbtgpu_carry cout_;
  NCK(cudaMemcpyAsync(ctr, sc->counters, sizeof(ctr), cudaMemcpyDeviceToHost, st));
  NCK(cudaMemcpyAsync(flags, sc->flags, sizeof(flags), cudaMemcpyDeviceToHost, st));
  NCK(cudaMemcpyAsync(&cout_, &sc->carry_out, sizeof(cout_), cudaMemcpyDeviceToHost, st));
  NCK(cudaStreamSynchronize(st));
  if (flags[FLAG_INTERNAL])
    return na_fail(-EIO, "internal inconsistency: exact pass met an unresolved record (code %u)",
                   flags[FLAG_INTERNAL]);
  memset(counters_out, 0, sizeof(*counters_out));
  counters_out->lost = ctr[C_LOST];
  counters_out->throttle = ctr[C_THROTTLE];
  counters_out->unthrottle = ctr[C_UNTHROTTLE];
  counters_out->sample = ctr[C_SAMPLE];
  counters_out->read = ctr[C_READ];
  counters_out->num_records = ctr[C_RECORDS];
  counters_out->num_unknown_count_interval = ctr[C_UNKNOWN];
  counters_out->num_rows = ctr[C_ROWS];
  counters_out->rows_out_of_range = ctr[C_OOR];
  counters_out->err = ctr[C_ERR] ? -EPERM : 0;
  counters_out->err_offset = ctr[C_ERR] ? ctr[C_ERR_OFF] : image_bytes;
  counters_out->last_tstamp = cout_.valid ? cout_.tstamp : ~0ull;
  if (carry_out) *carry_out = cout_;
  return 0;
#undef NCK
}

}  // extern "C"

#include "hbtgpu_peer.inc"
#include "hbtgpu_series.inc"
#include "hbtcap.inc"
#include "hbtgpu_cs_host.inc"
