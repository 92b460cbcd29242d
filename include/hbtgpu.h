/*
 * hbtgpu.h -- C ABI of libhbtgpu.so: B200 (sm_100a) decode + aggregate of hbt
 * perf_event ring-buffer images.
 *
 * This is the drop-in boundary for ONE path of dynolog/hbt: what
 * PerfEventsGroup<CpuCountSampleGenerator, mode::Sampling>::consume() and its
 * handlers compute from the bytes of a per-CPU perf_event ring, plus the
 * GroupReadValues / BuiltinMetrics arithmetic applied to the result.  The
 * reference has no FFI layer (its boundary is a C++ template surface), so every
 * entry point cites the reference code it replaces (paths relative to the
 * dynolog tree).  dynolog_b200/shim/HbtGpuShim.h re-exposes these entry points
 * with the reference's class shapes; INTEGRATION.md shows the binding.
 *
 * Conventions (mirroring hbt):
 *   - return 0 on success or a NEGATIVE errno (hbt's hot path returns negative
 *     errno in ssize_t: -EPERM bad record / time regression,
 *     hbt/src/perf_event/PerfEventsGroup.h:1505,1521 and
 *     hbt/src/perf_event/PerCpuCountSampleGenerator.h:162).  CUDA failures map to
 *     -EIO; hbtgpu_last_error() returns the text.  No exceptions cross the ABI.
 *   - one context = one CUDA device + one stream; not thread-safe; callers
 *     serialise, exactly as they do around a generator
 *     (hbt/src/mon/TraceCollector.h:390).
 *   - all pointers are HOST pointers unless the name says "device".
 *   - there is NO CPU fallback: without a CUDA device hbtgpu_create fails.
 */
#ifndef HBTGPU_H
#define HBTGPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HBTGPU_MAX_EVENTS 64 /* CpuCountSampleGenerator::kMaxNumEvents, PerCpuCountSampleGenerator.h:44 */
#define HBTGPU_MAX_METRICS 8
#define HBTGPU_MAX_METRIC_EVENTS 8

/* hbtgpu_config.flags */
#define HBTGPU_F_FORCE_EXACT 0x1u   /* always run the exact (sequential-resolver) path */
#define HBTGPU_F_FORCE_GENERAL 0x2u /* always run pointer-jumping, never the contiguous-chain shortcut */
#define HBTGPU_F_EXTERNAL_CARRY 0x4u /* shard mode: predecessor of each stream lives on another GPU (see hbtgpu_get_stubs) */
#define HBTGPU_F_ACCUMULATE_ATTR 0x8u /* IP|TID format: the (pid,tid) histogram and the IP counts add up over runs
                                        (a capture larger than the arena is streamed through in pieces: submit,
                                        run, hbtgpu_reset(keep_carry = 1), ...) until hbtgpu_attr_clear */

/* hbtgpu_config.sample_format: layout of PERF_RECORD_SAMPLE.
 * TIME_READ is what hbt's Sampling mode opens (sample_type = TIME|READ,
 * hbt/src/perf_event/PerfEventsGroup.h:961,973-977): {hdr, time, nr, te, tr, value[nr]}.
 * IP_TID_TIME_READ is a spec extension for BASELINE config 4 (hbt never requests IP|TID in
 * Sampling mode; layout from the kernel ABI, linux/perf_event.h PERF_RECORD_SAMPLE):
 * {hdr, ip, pid:u32, tid:u32, time, nr, te, tr, value[nr]}; it additionally yields a
 * per-(pid,tid) sample histogram and per-IP sample counts (hbtgpu_get_tid_histogram,
 * hbtgpu_get_top_ips). */
#define HBTGPU_SAMPLE_TIME_READ 0u
#define HBTGPU_SAMPLE_IP_TID_TIME_READ 1u

/* metric kinds (hbtgpu_metric.kind) */
#define HBTGPU_METRIC_RATE 1 /* getRateReducer, hbt/src/perf_event/BuiltinMetrics.cpp:102-118 */
#define HBTGPU_METRIC_ADD 2  /* getAddReducer (int accumulator), BuiltinMetrics.cpp:88-97 */
#define HBTGPU_METRIC_PER_SECOND_MEGA 3 /* mips / mega_cycles_per_second, dynolog/src/PerfMonitor.cpp:63-72 */

typedef struct hbtgpu_ctx hbtgpu_ctx;

/* One derived metric over the multiplex-scaled counts of a bucket: the device
 * form of MetricDesc::reducer (hbt/src/perf_event/Metrics.h:46-214) as applied
 * by GroupReadValues::getReducedCount (PerfEventsGroup.h:502-511). */
typedef struct {
  uint32_t kind;
  uint32_t n_idx;                          /* RATE: 2 (num, denom); PER_SECOND_MEGA: 1 */
  uint32_t idx[HBTGPU_MAX_METRIC_EVENTS];  /* event indices */
} hbtgpu_metric;

typedef struct {
  int32_t device;      /* CUDA ordinal */
  uint32_t n_events;   /* events in the group = nr of every SAMPLE (kNumEvents, PerCpuCountSampleGenerator.h:47) */
  uint64_t tsc_mask;   /* bit i: event i has EventValueTransforms::TscToKernelTime (hbt/src/perf_event/PmuEvent.h:248-256) */
  uint32_t time_shift; /* perf_event_mmap_page time_shift/mult/zero = TscConversionParams (PerfEventsGroup.h:829-841) */
  uint32_t time_mult;
  uint64_t time_zero;
  uint32_t max_cpus;   /* streams: cpu ids 0..max_cpus-1 (kMaxCpus = 512, hbt/src/common/System.h:204) */
  uint32_t n_buckets;  /* time buckets per cpu (>= 1) */
  uint64_t t0;         /* bucket b covers [t0 + b*bucket_ns, t0 + (b+1)*bucket_ns) */
  uint64_t bucket_ns;  /* >= 1 */
  uint64_t arena_bytes; /* device capacity for submitted images (bytes) */
  uint32_t n_metrics;
  uint32_t flags;
  hbtgpu_metric metrics[HBTGPU_MAX_METRICS];
  uint32_t sample_format;   /* HBTGPU_SAMPLE_* */
  uint32_t tid_table_log2;  /* IP|TID format: slots of the (pid,tid) table = 2^this (0: 2^16) */
  uint32_t ip_table_log2;   /* IP|TID format: slots of the sampled-IP table  = 2^this (0: 2^22) */
  uint32_t reserved;
} hbtgpu_config;

/* Generator state that crosses image boundaries: CpuCountSampleGenerator's
 * count_.tstamp / last_[] (PerCpuCountSampleGenerator.h:222-228), plus the
 * previous sample's time_enabled/time_running for the bucket dTE/dTR sums.
 * valid=1,last=0 is the state enableImpl() leaves (:91-106); valid=0 is the
 * state after a PERF_RECORD_LOST (:110-124). */
typedef struct {
  uint64_t valid;
  uint64_t tstamp;
  uint64_t te, tr;
  uint64_t last[HBTGPU_MAX_EVENTS]; /* already transformed values */
} hbtgpu_carry;

/* Per-cpu counters: the 11 getNumRecord*() of PerfEventsGroup.h:791-824 (same
 * order as the members :887-897) and the generator's bookkeeping. */
typedef struct {
  uint64_t lost;       /* SUM of num_lost, not #LOST records (:1448) */
  uint64_t comm, exit, throttle, unthrottle, fork, sample, read, aux, itrace_start,
      switch_cpu_wide;
  uint64_t num_records;                /* every record walked, any type (:1569) */
  uint64_t num_unknown_count_interval; /* PerCpuCountSampleGenerator.h:152 */
  uint64_t num_rows;                   /* rows written to the output ring (:182) */
  uint64_t rows_out_of_range;          /* rows whose bucket is outside [0, n_buckets) */
  int64_t err;                         /* 0 or the negative errno consume() would return */
  uint64_t err_offset;                 /* stream byte offset of the record consume() stalls on */
  uint64_t last_tstamp;                /* getLastTimeStamp(), PerCpuCountSampleGenerator.h:209-211 */
} hbtgpu_counters;

/* Layout of the result tensors (row-major, packed):
 *   acc    [max_cpus][n_buckets][n_events+3] u64: sum delta[e], sum dTE, sum dTR, n_rows
 *   scaled [max_cpus][n_buckets][n_events]   u64: GroupReadValues::getCount on the accumulated triple
 *   metric [max_cpus][n_buckets][n_metrics]  f64
 *   valid  [max_cpus][n_buckets]             u32 bitmask: bit m set when metric m has a value
 *          (std::optional has_value of the reducer) */
typedef struct {
  const uint64_t* acc;
  const uint64_t* scaled;
  const double* metric;
  const uint32_t* valid;
  uint32_t max_cpus, n_buckets, n_events, n_metrics;
} hbtgpu_bucket_view;

/* ---- lifetime --------------------------------------------------------------------- */

/* Allocates the device arena, result tensors and a stream.  Replaces the
 * constructor + open() of PerCpuCountSampleGenerator
 * (PerCpuCountSampleGenerator.h:234-294) for the purposes of this path. */
int hbtgpu_create(const hbtgpu_config* cfg, hbtgpu_ctx** out);
void hbtgpu_destroy(hbtgpu_ctx* ctx);
const char* hbtgpu_last_error(const hbtgpu_ctx* ctx); /* ctx may be NULL: error of the last failed create */

/* Initial generator state per cpu; default is valid=0 (as after a LOST).
 * Replaces enableImpl() (PerCpuCountSampleGenerator.h:91-106), with now()
 * passed explicitly. */
int hbtgpu_set_carry(hbtgpu_ctx* ctx, uint32_t cpu, const hbtgpu_carry* carry);

/* ---- input ------------------------------------------------------------------------- */

/* Append one ring image of `cpu`: the bytes PerfEventsGroup::onCpuDataBufferRead
 * hands to its callback (PerfEventsGroup.h:1592-1637, RbDataSlices :45-62):
 * `first` then the optional wrapped `second` slice.  The image must start and
 * end on record boundaries (data_tail .. data_head).  Images of one cpu are
 * decoded in submission order as one stream.  The bytes are copied (H2D) before
 * the call returns, so the caller may consume the ring immediately. */
int hbtgpu_submit_image(hbtgpu_ctx* ctx, uint32_t cpu, const void* first, size_t n1,
                        const void* second, size_t n2);

/* Several images in one call (e.g. every cpu's slice from TraceCollector::onCpusDataBufferRead,
 * hbt/src/mon/TraceCollector.h:406-420): same contract as n calls of hbtgpu_submit_image in
 * array order, but the copies are queued back to back and waited for once. */
int hbtgpu_submit_images(hbtgpu_ctx* ctx, size_t n, const uint32_t* cpus, const void* const* images,
                         const size_t* sizes);

/* Same, for an image that already lives in device memory (copied D2D). */
int hbtgpu_submit_image_device(hbtgpu_ctx* ctx, uint32_t cpu, const void* device_ptr, size_t n);

/* Forget all submitted images, counters and bucket sums (carry-in is kept as
 * last set).  With keep_carry != 0 each cpu's carry-in becomes the carry-out
 * of the previous run, i.e. the next images continue the same streams. */
int hbtgpu_reset(hbtgpu_ctx* ctx, int keep_carry);

/* ---- the hot path ------------------------------------------------------------------ */

/* Decode every submitted image and aggregate: the device form of calling
 * consume() until -ENODATA on every cpu (PerCpuSampleGeneratorBase::accumUntil,
 * hbt/src/perf_event/PerCpuSampleGeneratorBase.h:69-94) followed by the bucket
 * sums, multiplex scaling and metric reducers.  Asynchronous on the context's
 * stream; may be called repeatedly on the same images (results are recomputed).
 * A per-cpu decode error does not fail the call: see hbtgpu_counters.err. */
int hbtgpu_run(hbtgpu_ctx* ctx);
int hbtgpu_sync(hbtgpu_ctx* ctx);

/* ---- results ----------------------------------------------------------------------- */

/* getNumRecord*(), getLastTimeStamp(), getNumEntries() of one cpu. */
int hbtgpu_get_counters(hbtgpu_ctx* ctx, uint32_t cpu, hbtgpu_counters* out);
/* State after the last record: feed to hbtgpu_set_carry of the next window. */
int hbtgpu_get_carry(hbtgpu_ctx* ctx, uint32_t cpu, hbtgpu_carry* out);

/* Rows of one cpu as CpuCountSampleGenerator writes them to its output ring:
 * [tstamp, duration, delta_0..n-1] u64 (PerCpuCountSampleGenerator.h:34,165-182;
 * consumer callback shape hbt/src/mon/TraceMonitor.h:35-108).  *n_rows is
 * set to the number available; at most `cap_rows` are copied. */
int hbtgpu_get_rows(hbtgpu_ctx* ctx, uint32_t cpu, uint64_t* rows, size_t cap_rows, size_t* n_rows);
/* The same rows where they were written, for consumers on the GPU (hbtgpu_islice_release with
 * HBTGPU_ISLICE_DEVICE_INPUT): cpu c's rows are rows[row_begin[c] .. row_begin[c+1]), n_events + 2
 * words each.  *rows is a device pointer; row_begin is a HOST array of max_cpus + 1 entries that
 * the call fills.  Valid until the next hbtgpu_run / hbtgpu_reset. */
int hbtgpu_rows_device(hbtgpu_ctx* ctx, const uint64_t** rows, uint64_t* row_begin, uint32_t* row_stride);
/* {tstamp, duration} of the same rows once more, two words per row, densely: all that
 * IntervalSlicer::makeIntervalIntersectingSlices reads of a row (TraceMonitor.cpp:52-96), so
 * hbtgpu_islice_release(..., rows = times, row_stride = 2, ...) touches 16 bytes per row instead of
 * every 64-byte segment of the full rows.  row_begin as above (may be NULL). */
int hbtgpu_row_times_device(hbtgpu_ctx* ctx, const uint64_t** times, uint64_t* row_begin);

/* Record table of one cpu: (stream byte offset, type, size) of every record the
 * walk visits, in order -- what consume()'s loop steps through (:1415-1572). */
int hbtgpu_get_records(hbtgpu_ctx* ctx, uint32_t cpu, uint64_t* off, uint32_t* type, uint32_t* size,
                       size_t cap, size_t* n_records);

/* HBTGPU_SAMPLE_IP_TID_TIME_READ only (BASELINE config 4; spec extension, no reference code).
 * Per-(pid,tid) sample histogram over all cpus: one entry per distinct (pid,tid) seen in a
 * SAMPLE the walk counted, in unspecified order.  *n is set to the number of entries; at most
 * `cap` are written.  -ENOSPC if the table (hbtgpu_config.tid_table_log2) overflowed. */
int hbtgpu_get_tid_histogram(hbtgpu_ctx* ctx, uint32_t* pid, uint32_t* tid, uint64_t* count, size_t cap,
                             size_t* n);
/* The k most-sampled instruction pointers over all cpus, ordered by (count descending, ip
 * ascending).  *n = min(k, distinct IPs).  -ENOSPC if the IP table overflowed. */
int hbtgpu_get_top_ips(hbtgpu_ctx* ctx, uint32_t k, uint64_t* ip, uint64_t* count, size_t* n);

/* HBTGPU_F_ACCUMULATE_ATTR: start the histogram and the IP counts from zero again. */
int hbtgpu_attr_clear(hbtgpu_ctx* ctx);

/* Bucket tensors.  _device returns pointers into the context's device memory
 * (valid until destroy; contents change on run/reset); _host copies them into
 * caller buffers (any may be NULL). */
int hbtgpu_get_buckets_device(hbtgpu_ctx* ctx, hbtgpu_bucket_view* view);
int hbtgpu_get_buckets_host(hbtgpu_ctx* ctx, uint64_t* acc, uint64_t* scaled, double* metric,
                            uint32_t* valid);

/* Asynchronous form of hbtgpu_get_buckets_host: the device-to-host copies are queued on a
 * second stream behind the last run and may overlap the next hbtgpu_submit_image(s) (PCIe is
 * full duplex); the buffers (ideally pinned) are complete after hbtgpu_sync, and the next run's
 * derive step waits for the copies before it overwrites the tensors. */
int hbtgpu_get_buckets_host_async(hbtgpu_ctx* ctx, uint64_t* acc, uint64_t* scaled, double* metric,
                                  uint32_t* valid);

/* Re-run only the derive step (scale + metrics) from the current acc tensor,
 * e.g. after partial aggregates of other GPUs were summed into it. */
int hbtgpu_derive(hbtgpu_ctx* ctx);

/* ---- multi-GPU (time-window sharding) ---------------------------------------------- */

/* A capture is sharded by TIME WINDOW: rank r holds window r of every cpu (images start on
 * record boundaries, so no boundary search crosses GPUs).  With HBTGPU_F_EXTERNAL_CARRY a
 * context decodes its shard as if every stream started from an invalid state (as after a
 * LOST): the first SAMPLE of a stream only re-bases.  What is then missing is, per cpu, the ONE
 * row that straddles the shard boundary.  Protocol (no data-path collective):
 *   1. every rank: hbtgpu_run, then hbtgpu_get_stubs -> head[cpu], tail[cpu]
 *   2. all_gather of the stubs (about 1 KB per cpu and rank)
 *   3. every rank folds the tails of the ranks before it, per cpu, with hbtgpu_compose_tail
 *      (start value: the state enableImpl() leaves, i.e. what a single-GPU run passes to
 *      hbtgpu_set_carry)
 *   4. every rank: hbtgpu_apply_boundaries(prev_tails) adds the boundary rows to its own
 *      buckets, fixes the counters and runs the derive step.
 * The union of the ranks' results then equals the single-stream result bit for bit. */
typedef struct {
  uint64_t has_sample;  /* the shard's stream contains at least one SAMPLE */
  uint64_t lost_before; /* a LOST precedes the first SAMPLE (or: any LOST, if there is no SAMPLE) */
  uint64_t offset;      /* stream byte offset of the first SAMPLE */
  uint64_t tstamp, te, tr;
  uint64_t value[HBTGPU_MAX_EVENTS]; /* transformed values of the first SAMPLE */
} hbtgpu_head_stub;

/* head and tail: arrays of max_cpus entries (host). tail[cpu] is the carry-out of the shard. */
int hbtgpu_get_stubs(hbtgpu_ctx* ctx, hbtgpu_head_stub* head, hbtgpu_carry* tail);
/* Host-only, no GPU work: generator state after a shard, given the state before it (prev),
 * the shard's head stub and the shard's own carry-out (computed from an invalid state).
 * Mirrors handleRecordLost / handleRecordSample: a shard with a SAMPLE determines the state by
 * itself; one without keeps prev, invalidated if it contains a LOST
 * (PerCpuCountSampleGenerator.h:110-124,135-154,198-199). */
void hbtgpu_compose_tail(const hbtgpu_carry* prev, const hbtgpu_head_stub* head,
                         const hbtgpu_carry* own_tail, hbtgpu_carry* out);
/* prev_tails: max_cpus entries (host), the state before this shard per cpu.  Adds the row
 * first_sample - prev_tail to the bucket of the first SAMPLE where handleRecordSample would
 * have emitted it (:165-182), updates num_rows / num_unknown_count_interval / err, then runs
 * the derive step. Call once per hbtgpu_run. */
int hbtgpu_apply_boundaries(hbtgpu_ctx* ctx, const hbtgpu_carry* prev_tails);

/* The same exchange without host round trips: the stubs stay in device memory, a collective
 * (e.g. ncclAllGather through torch.distributed, enqueued on hbtgpu_stream) moves them, and the
 * fold runs in a kernel.  A rank's block is opaque to the caller (packed: 10 + 2 * n_events words
 * per cpu).  Unlike the host protocol above, the device fold also reproduces what the reference
 * does after an error: a cpu's walk stalls for good at the first record consume() fails on
 * (hbt/src/perf_event/PerfEventsGroup.h:1505,1521, PerCpuCountSampleGenerator.h:156-163), so the
 * shards behind it -- and a shard whose first SAMPLE goes back in time against the previous shard
 * -- contribute nothing: their sums for that cpu are cleared and their counters say err = -EPERM.
 *   hbtgpu_stub_bytes            size of one rank's stub block
 *   hbtgpu_export_stubs_device   pack this rank's block into device_dst (async, context stream)
 *   hbtgpu_apply_boundaries_gathered
 *                                device_gathered = `world` blocks in rank order; replays the
 *                                generator state over the ranks starting from the carry set with
 *                                hbtgpu_set_carry, adds the boundary rows, runs the derive step
 *                                (all async on the context stream).  If hbtgpu_derive was
 *                                called after hbtgpu_run (say, while the collective was in
 *                                flight on another stream), only the cells the boundary rows
 *                                touched are derived again. */
size_t hbtgpu_stub_bytes(const hbtgpu_ctx* ctx);
int hbtgpu_export_stubs_device(hbtgpu_ctx* ctx, void* device_dst);
int hbtgpu_apply_boundaries_gathered(hbtgpu_ctx* ctx, const void* device_gathered, uint32_t world,
                                     uint32_t rank);

/* ---- the exchange and the merge over peer memory (NVLink), no collective library ---------
 *
 * SURVEY.md 8(b)/(e) asks for hbtgpu_merge: "only the small per-shard partial aggregates are
 * merged".  There is no reference code for this (hbt is one process on one host,
 * hbt/src/mon/TraceCollector.h:501-550).  One process per GPU (or several contexts in one
 * process): every rank exports a handle, the handles are exchanged by any means (a few bytes:
 * torch.distributed.all_gather_object, MPI, a file), every rank attaches, and from then on a step
 * is   hbtgpu_run -> hbtgpu_exchange_boundaries -> [hbtgpu_merge]   with no host round trip:
 *   hbtgpu_exchange_boundaries  stores this rank's packed stubs into every peer's mailbox,
 *                               waits for everybody's, then does what
 *                               hbtgpu_apply_boundaries_gathered does
 *   hbtgpu_merge                every rank ends up with the sums of ALL ranks: the bucket ranges
 *                               each rank's rows can have touched are known from the stubs, so a
 *                               rank reads exactly those from its peers' accumulators (ordinary
 *                               loads over NVLink) and runs the derive step on the totals.
 *                               Afterwards the bucket getters return the totals (until the next
 *                               hbtgpu_run); hbtgpu_acc_device keeps returning this rank's own
 *                               partial sums.  Requires the same t0 / bucket_ns / n_buckets /
 *                               metrics on all ranks.  Integer sums: bit-identical to one
 *                               context over the whole capture, whatever the order.
 *   hbtgpu_attr_merge           IP|TID format (last bullet): every rank ends up
 *                               with the (pid,tid) histogram and the sampled-IP counts of ALL
 *                               ranks.  A rank offers its totals over its runs
 *                               (HBTGPU_F_ACCUMULATE_ATTR) or else its last run's tables; every
 *                               rank inserts every rank's entries into an empty table of its own
 *                               (ordinary loads over NVLink) -- the full merge, exact: no top-K'
 *                               candidates, no recount.  Afterwards hbtgpu_get_tid_histogram /
 *                               hbtgpu_get_top_ips return the merged tables (until the next
 *                               hbtgpu_run or hbtgpu_attr_clear).  Same table sizes on all ranks;
 *                               a merged table that overflows is reported by the getters
 *                               (-ENOSPC), never silently truncated.  Needs no
 *                               hbtgpu_exchange_boundaries first.
 * All ranks must issue the same sequence of these calls.  A rank that waits longer than 4 s for
 * a peer gives up: the next hbtgpu_sync returns -ETIMEDOUT (nothing hangs).
 * Contexts must have HBTGPU_F_EXTERNAL_CARRY.  hbtgpu_destroy detaches. */
#define HBTGPU_PEER_HANDLE_BYTES 256
int hbtgpu_peer_export(hbtgpu_ctx* ctx, uint32_t world, void* handle /* HBTGPU_PEER_HANDLE_BYTES */);
/* handles: world * HBTGPU_PEER_HANDLE_BYTES, in rank order (this rank's own included) */
int hbtgpu_peer_attach(hbtgpu_ctx* ctx, uint32_t world, uint32_t rank, const void* handles);
/* the same for contexts that live in this process (peers[rank] is ignored) */
int hbtgpu_peer_attach_local(hbtgpu_ctx* ctx, uint32_t world, uint32_t rank, hbtgpu_ctx* const* peers);
int hbtgpu_peer_detach(hbtgpu_ctx* ctx);
#define HBTGPU_EXCHANGE_NO_DERIVE 0x1u /* hbtgpu_merge follows and derives from the totals */
int hbtgpu_exchange_boundaries(hbtgpu_ctx* ctx, uint32_t flags);
int hbtgpu_merge(hbtgpu_ctx* ctx);
int hbtgpu_attr_merge(hbtgpu_ctx* ctx);

/* Shard mode, after the boundaries were applied (any of the three ways): the generator state in
 * front of this rank's shard, as the fold over the earlier ranks found it. */
int hbtgpu_get_prev_tail(hbtgpu_ctx* ctx, uint32_t cpu, hbtgpu_carry* out);
/* Device address/size of the acc tensor, for a collective (NCCL all-reduce over u64 as i64). */
int hbtgpu_acc_device(hbtgpu_ctx* ctx, void** device_ptr, size_t* bytes);
/* The CUDA stream the context launches on (a cudaStream_t). */
void* hbtgpu_stream(hbtgpu_ctx* ctx);

/* ---- elementwise arithmetic of the path (host in, host out; computed on the GPU) --- */

/* GroupReadValues::getCount (PerfEventsGroup.h:467-481). */
int hbtgpu_math_get_count(hbtgpu_ctx* ctx, const uint64_t* count, const uint64_t* te,
                          const uint64_t* tr, uint64_t* out, size_t n);
/* getRateReducer (BuiltinMetrics.cpp:102-118); valid[i]=0 where it returns nullopt. */
int hbtgpu_math_rate(hbtgpu_ctx* ctx, const uint64_t* num, const uint64_t* den, double* out,
                     uint8_t* valid, size_t n);
/* getAddReducer (BuiltinMetrics.cpp:88-97) over n rows of k values. */
int hbtgpu_math_add_reduce(hbtgpu_ctx* ctx, const uint64_t* vals, size_t k, double* out, size_t n);
/* kernelTimeFromTsc / tscFromKernelTime (PerfEventsGroup.h:1645-1678) with the context's params. */
int hbtgpu_math_tsc_to_ktime(hbtgpu_ctx* ctx, const uint64_t* tsc, uint64_t* out, size_t n);
int hbtgpu_math_ktime_to_tsc(hbtgpu_ctx* ctx, const uint64_t* kt, uint64_t* out, size_t n);
/* float(double(count)*1e3/double(time_running)) (dynolog/src/PerfMonitor.cpp:63-72). */
int hbtgpu_math_per_second_mega(hbtgpu_ctx* ctx, const uint64_t* count, const uint64_t* tr,
                                float* out, size_t n);
/* getRunningRatio / getRunningRatioInteger<100> / hasMultiplexing (PerfEventsGroup.h:513-540). */
int hbtgpu_math_running_ratio(hbtgpu_ctx* ctx, const uint64_t* te, const uint64_t* tr, float* ratio,
                              uint64_t* ratio_int100, uint8_t* has_mux, size_t n);

/* ---- introspection ------------------------------------------------------------------ */

typedef struct {
  uint64_t image_bytes;   /* algorithmic bytes: sum of image sizes */
  uint32_t n_chunks;      /* CTAs of the decode kernel */
  uint32_t exact_runs;    /* times the last run fell back to the exact path (0 or 1) */
  uint32_t kernel_launches; /* kernels launched by the last hbtgpu_run */
  float last_run_ms;      /* CUDA-event time of the last hbtgpu_run (after hbtgpu_sync) */
  float last_decode_ms;   /* ... of its decode+aggregate kernel alone */
  uint32_t exact_reason;  /* diagnostics: why the speculative pass was disproved (0: it held) */
  uint32_t exact_chunk;   /* ... and the first chunk that reported it */
  double decode_ms_sum;   /* CUDA-event time of the decode+aggregate kernel, summed over ... */
  uint64_t decode_runs;   /* ... this many runs since create (harvested at hbtgpu_sync) */
  float last_emit_ms;     /* CUDA-event time of the last row / record-table emission pass */
  uint32_t pad;
} hbtgpu_stats;
int hbtgpu_get_stats(hbtgpu_ctx* ctx, hbtgpu_stats* out);

const char* hbtgpu_version(void);

/* ---- for callers that must not allocate ----------------------------------------------- */

/* Largest image hbtgpu_decode_image_noalloc accepts. */
#define HBTGPU_NOALLOC_MAX_IMAGE_BYTES (64ull << 20)

/* Decode + aggregate ONE ring image of ONE cpu without allocating anything, neither on the
 * device nor on the host.  (No counterpart in the reference: hbt never touches a GPU.)
 *
 * Who needs this: dynolog is an always-on daemon that shares the GPU with the training job it
 * watches.  That job typically owns almost all device memory, so a cudaMalloc from the
 * monitor's sampling loop can fail at any moment, and even when it succeeds cudaMalloc and
 * cudaFree synchronise the whole device, i.e. they stall the very job being observed.  The
 * same holds for a caller that runs inside somebody else's memory budget (a plugin loaded into
 * the trainer process) and for one that has frozen its memory pool.  Such callers create no
 * context; they hand in device buffers they already own and call this one function.
 *
 * The working memory the decode kernels need (chunk table, per-chunk hand-over records, flags,
 * carry, counters; about 160 KiB) lives in a static __device__ array inside the library, which
 * every GPU that loads the library holds until the process exits.  The price of that fixed
 * array is the size limit above: a larger image is refused with -E2BIG and a message, nothing
 * is truncated.
 *
 *   cfg           device, n_events, tsc_*, time_*, t0, bucket_ns, n_buckets and flags are used;
 *                 max_cpus, arena_bytes and the metrics are ignored
 *   carry_in      host pointer, may be NULL (= invalid state, as after a LOST)
 *   device_image  device pointer, 16-byte aligned; the bytes of one image (start and end on
 *                 record boundaries); must be readable up to image_bytes rounded up to 16
 *   device_acc    device pointer to n_buckets * (n_events + 3) u64; overwritten with the bucket
 *                 sums (same layout as hbtgpu_bucket_view.acc for one cpu)
 *   counters_out  host pointer; carry_out host pointer or NULL
 *   cuda_stream   a cudaStream_t (NULL = default stream); the call returns after the stream
 *                 has drained, results are then complete
 *
 * First call on a device: looks up the address of the static array (cudaGetSymbolAddress) and
 * raises the kernels' shared-memory limit.  Make that first call before any stream capture.
 * All calls on one device share the one array and it keeps what the previous call left; every
 * call re-initialises the parts that must start from zero.  Two calls whose GPU work overlaps
 * (different streams, different threads) therefore race on it: serialise them. */
int hbtgpu_decode_image_noalloc(const hbtgpu_config* cfg, const hbtgpu_carry* carry_in,
                                const void* device_image, size_t image_bytes,
                                uint64_t* device_acc, hbtgpu_counters* counters_out,
                                hbtgpu_carry* carry_out, void* cuda_stream);
/* Message of the last failed hbtgpu_decode_image_noalloc on this thread. */
const char* hbtgpu_noalloc_last_error(void);

/* ---- series reductions over the buckets (the sink side) --------------
 * dynolog keeps every logged metric in a facebook::dynolog::MetricSeries<T> and reduces it
 * before handing it to a Logger (dynolog/src/metric_frame/MetricSeries.h): avg<R> :207-213,
 * percentile :215-226, diff :228-234, rate<R> :198-205,246-263.  Here the series of (cpu, event)
 * is scaled[cpu][first_bucket .. first_bucket + count)[event] of the last hbtgpu_run (T =
 * uint64_t), reduced on the GPU with the reference's arithmetic (the f64 mean adds left to
 * right).  Outputs are host arrays [max_cpus][n_events]; any may be NULL.
 * period_ns / duration_ns: std::chrono::steady_clock::duration ticks of rate().
 * -EINVAL for percentage outside [0, 1] (the reference would index past the end). */
int hbtgpu_series_reduce(hbtgpu_ctx* ctx, uint32_t first_bucket, uint32_t count, float percentage,
                         int64_t period_ns, int64_t duration_ns, double* avg_f64, uint64_t* avg_u64,
                         uint64_t* percentile, uint64_t* diff, double* rate_f64, uint64_t* rate_u64);

/* ---- capture container and Logger sink ----------------------------------
 *
 * The reference defines no capture file format: ring images exist only as the borrowed slices
 * that PerfEventsGroup::onCpuDataBufferRead hands to a callback (PerfEventsGroup.h:45-64,
 * 1592-1637).  hbtcap_* is the smallest container that lets a daemon write what it read and a GPU
 * box decode it later: what the decoder has to know about the group (events with their
 * nicknames and TscToKernelTime flags, hbt/src/perf_event/PmuEvent.h:248-256; the TSC conversion
 * parameters of perf_event_mmap_page, PerfEventsGroup.h:829-841; the SAMPLE layout), the images
 * (128-byte aligned) and an index (window, cpu) -> (offset, length).  Host code only.
 * Writer: hbtcap_create, hbtcap_append per (window, cpu) in any order, hbtcap_finish.
 * Reader: hbtcap_open (mmap; -EINVAL for anything that is not a well-formed capture),
 * hbtcap_config fills the decoder's configuration from it, hbtcap_submit hands the images of the
 * windows [window_lo, window_hi) to a context in time order (hbtgpu_submit_images). */
#define HBTCAP_NAME_BYTES 32
typedef struct {
  uint32_t n_events, sample_format, n_cpus, n_windows;
  uint32_t time_shift, time_mult;
  uint64_t time_zero, tsc_mask;
  char event_names[HBTGPU_MAX_EVENTS][HBTCAP_NAME_BYTES]; /* nicknames, NUL-terminated */
} hbtcap_desc;
typedef struct hbtcap hbtcap;
int hbtcap_create(const char* path, const hbtcap_desc* desc, hbtcap** out);
int hbtcap_append(hbtcap* cap, uint32_t window, uint32_t cpu, const void* first, size_t n1, const void* second,
                  size_t n2);
int hbtcap_finish(hbtcap* cap); /* writes index and header, closes, frees */
int hbtcap_open(const char* path, hbtcap** out);
int hbtcap_describe(const hbtcap* cap, hbtcap_desc* out);
int hbtcap_config(const hbtcap* cap, hbtgpu_config* cfg); /* n_events, sample_format, tsc_*, time_*; raises max_cpus */
int hbtcap_image(const hbtcap* cap, uint32_t window, uint32_t cpu, const void** bytes, size_t* n);
int hbtcap_submit(const hbtcap* cap, hbtgpu_ctx* ctx, uint32_t window_lo, uint32_t window_hi);
void hbtcap_close(hbtcap* cap);
const char* hbtcap_last_error(const hbtcap* cap);

/* dynolog::Logger (dynolog/src/Logger.h:26-48) as C callbacks: logUint / logFloat / finalize.
 * hbtgpu_log_series reduces every (cpu, event) series of scaled counts over the buckets
 * [first_bucket, first_bucket + count) on the GPU (hbtgpu_series_reduce, i.e. MetricSeries avg /
 * percentile / diff / rate, dynolog/src/metric_frame/MetricSeries.h:198-263) and publishes one log
 * entry per cpu: keys "<nickname>_avg.cpu<N>", "_rate" (logFloat), "_p<percent>", "_diff"
 * (logUint) -- "<metric>.<entity>", the form dynolog::splitKey takes apart -- then finalize(). */
typedef struct {
  void* user;
  void (*log_uint)(void* user, const char* key, uint64_t value);
  void (*log_float)(void* user, const char* key, float value);
  void (*finalize)(void* user);
} hbtgpu_logger;
int hbtgpu_log_series(hbtgpu_ctx* ctx, const hbtgpu_logger* logger, const char* const* event_names,
                      uint32_t first_bucket, uint32_t count, float percentage, int64_t period_ns,
                      int64_t duration_ns);

/* ---- ContextSwitch-mode decode -------------------------------------
 *
 * Replaces, for captured ring images of hbt's ContextSwitch mode (sample_type TIME|TID with
 * sample_id_all, hbt/src/perf_event/PerfEventsGroup.h:963-970; record layouts :105-191),
 *   PerfEventsGroup<CpuThreadSwitchGenerator, mode::ContextSwitch>::consume      :1394-1589
 *   CpuThreadSwitchGenerator::handleRecord{Lost,Comm,Exit,Fork,SwitchCpuWide}
 *       hbt/src/perf_event/PerCpuThreadSwitchGenerator.h:219-446
 * The GPU turns records into the 16-byte tagstack::Event stream; the tid -> virtual id
 * dictionary (PerThreadStats, PerCpuThreadSwitchGenerator.h:52-135), which is order-dependent
 * and shared by all CPUs, is the host function hbtgpu_cs_assign_vids below.
 */

/* tagstack::Event, hbt/src/tagstack/Event.h:28-56: same size, field order and type codes */
typedef struct {
  uint64_t tstamp;
  uint8_t type;          /* 2 ThreadCreation, 3 ThreadDestruction, 4 SwitchIn, 5 SwitchOutPreempt,
                            6 SwitchOutYield, 7 WriteErrorsStart, 8 WriteErrorsEnd */
  uint8_t level;         /* tid_level; 255 (kInvalidLevel) for the WriteErrors events */
  uint16_t comp_unit_id; /* cpu */
  uint32_t tag;          /* as decoded: the raw tid; after hbtgpu_cs_assign_vids: the virtual id.
                            0xFFFFFFFF (kNA) for events that carry none */
} hbtgpu_cs_event;

/* A record that changes the thread dictionary and is not a plain switch-in: LOST (clears the
 * active tids, :243-247), COMM, EXIT, FORK.  event_index is the position, in the cpu's event
 * array, of the event this record wrote (COMM writes none: the position its next event has). */
typedef struct {
  uint64_t tstamp;
  uint64_t event_index;
  uint32_t type;         /* perf record type: 2 LOST, 3 COMM, 4 EXIT, 7 FORK */
  uint32_t misc;
  int32_t pid, ppid, tid, ptid;
  uint32_t name_len;     /* COMM: bytes of the comm field (size - 32) */
  uint32_t pad;
  char name[16];         /* COMM: first 16 bytes of the comm field */
} hbtgpu_cs_op;

/* CpuThreadSwitchGenerator::last_tstamp_ / in_write_errors_ (:418-421) */
typedef struct {
  uint64_t last_tstamp;
  uint32_t in_write_errors;
  uint32_t pad;
} hbtgpu_cs_state;

typedef struct {
  int32_t device;
  uint32_t max_cpus;
  uint32_t tid_level;    /* tagstack::Level of the thread ids (ctor argument, :187-197) */
  uint32_t flags;        /* HBTGPU_F_FORCE_EXACT */
  uint64_t arena_bytes;  /* device memory for submitted images */
} hbtgpu_cs_config;

typedef struct hbtgpu_cs_ctx hbtgpu_cs_ctx;

hbtgpu_cs_ctx* hbtgpu_cs_create(const hbtgpu_cs_config* cfg);
void hbtgpu_cs_destroy(hbtgpu_cs_ctx* ctx);
const char* hbtgpu_cs_last_error(hbtgpu_cs_ctx* ctx);
/* Generator state a cpu starts from (default: last_tstamp 0, not in write errors, as after
 * construction).  After hbtgpu_cs_run the state to carry into the next window is
 * hbtgpu_cs_get_state. */
int hbtgpu_cs_set_state(hbtgpu_cs_ctx* ctx, uint32_t cpu, const hbtgpu_cs_state* st);
/* Append one ring image (whole records, oldest first) of one cpu; host memory, copied
 * asynchronously (pin it for speed).  Images of a cpu form its stream in submission order. */
int hbtgpu_cs_submit_image(hbtgpu_cs_ctx* ctx, uint32_t cpu, const void* bytes, uint64_t len);
/* The same for the two borrowed slices onCpuDataBufferRead hands out when the unread span wraps
 * around the end of the ring (RbDataSlices, PerfEventsGroup.h:45-62,1592-1637): first ++ second is
 * one image.  The cut may fall anywhere, also inside a record (consume() reassembles such a
 * record through its aux buffer, :1425-1437); here both slices land back to back in the arena. */
int hbtgpu_cs_submit_image_slices(hbtgpu_cs_ctx* ctx, uint32_t cpu, const void* first, uint64_t n1,
                                  const void* second, uint64_t n2);
/* Forget the submitted images (states are kept). */
int hbtgpu_cs_reset(hbtgpu_cs_ctx* ctx);
/* Decode everything submitted since the last reset.  Asynchronous on the context's stream
 * except for one read-back of the event totals. */
int hbtgpu_cs_run(hbtgpu_cs_ctx* ctx);
int hbtgpu_cs_sync(hbtgpu_cs_ctx* ctx);
/* getNumRecord*() (:791-824) of the cpu; num_rows = events written; err / err_offset as for
 * the Sampling path (-EPERM at the first record that trips a fence). */
int hbtgpu_cs_get_counters(hbtgpu_cs_ctx* ctx, uint32_t cpu, hbtgpu_counters* out);
int hbtgpu_cs_get_state(hbtgpu_cs_ctx* ctx, uint32_t cpu, hbtgpu_cs_state* out);
int hbtgpu_cs_num_events(hbtgpu_cs_ctx* ctx, uint32_t cpu, uint64_t* n_events, uint64_t* n_ops);
/* Events of a cpu in ring order, tag = raw tid; pids (may be NULL) gets, per event, the pid
 * the record carried (SwitchIn: sample_id.pid; ThreadCreation/Destruction: pid). */
int hbtgpu_cs_get_events(hbtgpu_cs_ctx* ctx, uint32_t cpu, hbtgpu_cs_event* out, uint32_t* pids,
                         uint64_t cap, uint64_t* n);
int hbtgpu_cs_get_ops(hbtgpu_cs_ctx* ctx, uint32_t cpu, hbtgpu_cs_op* out, uint64_t cap, uint64_t* n);
/* Device views for consumers that stay on the GPU: events / pids of all cpus, cpu c at
 * [base[c], base[c+1]). */
int hbtgpu_cs_events_device(hbtgpu_cs_ctx* ctx, const hbtgpu_cs_event** events, const uint32_t** pids,
                            const uint64_t** base);
typedef struct {
  uint64_t image_bytes;
  uint32_t n_chunks, exact_runs, exact_reason, exact_chunk;
  float last_count_ms, last_emit_ms, last_run_ms;
  uint32_t kernel_launches;
} hbtgpu_cs_stats;
int hbtgpu_cs_get_stats(hbtgpu_cs_ctx* ctx, hbtgpu_cs_stats* out);

/* Host mirror of PerThreadStats (PerCpuThreadSwitchGenerator.h:52-135): assigns virtual ids in
 * encounter order.  Feed it one cpu's window at a time, cpus in the order the reference's
 * consumer loop visits them; events[i].tag is rewritten from tid to vid in place.  /proc
 * lookups (cmdline, comm of threads first seen in a switch) are the caller's business. */
typedef struct hbtgpu_cs_threads hbtgpu_cs_threads;
typedef struct {
  uint32_t vid;
  int32_t ppid, ptid, pid, tid;
  uint32_t n_names;
  uint64_t fork_ts, end_ts;  /* ~0 (kInvalidTimeStamp) when unset */
} hbtgpu_cs_thread;
typedef struct {
  uint32_t vid;
  uint8_t is_exec;           /* always 1: `misc | PERF_RECORD_MISC_COMM_EXEC`, :273 */
  uint8_t len;
  uint8_t pad[2];
  uint64_t tstamp;
  char name[16];
} hbtgpu_cs_name;
hbtgpu_cs_threads* hbtgpu_cs_threads_create(void);
void hbtgpu_cs_threads_destroy(hbtgpu_cs_threads* t);
/* -EINVAL for input on which the reference's handlers would throw. */
int hbtgpu_cs_assign_vids(hbtgpu_cs_threads* t, hbtgpu_cs_event* events, const uint32_t* pids,
                          uint64_t n_events, const hbtgpu_cs_op* ops, uint64_t n_ops);
uint64_t hbtgpu_cs_threads_count(const hbtgpu_cs_threads* t);
uint64_t hbtgpu_cs_threads_name_count(const hbtgpu_cs_threads* t);
/* threads by vid; names grouped by thread, each thread's in series (time) order */
int hbtgpu_cs_threads_get(const hbtgpu_cs_threads* t, hbtgpu_cs_thread* threads, uint64_t thread_cap,
                          hbtgpu_cs_name* names, uint64_t name_cap);

/* The same dictionary on the GPU (for windows whose events should not make the round trip
 * through one host thread first): after hbtgpu_cs_run, rewrites the tags of ALL cpus of the
 * window from tid to virtual id in place on the device, cpus in index order, and keeps the
 * dictionary (active tids, thread table) on the device between windows.  Same results as
 * feeding hbtgpu_cs_assign_vids cpu by cpu.  -EINVAL for COMM/EXIT/FORK of tid 0 and for a FORK
 * of an already forked active tid; the reference's time-order assertions (:125-131,262-265) are
 * not checked here. */
typedef struct hbtgpu_cs_vids hbtgpu_cs_vids;
hbtgpu_cs_vids* hbtgpu_cs_vids_create(int device);
void hbtgpu_cs_vids_destroy(hbtgpu_cs_vids* v);
const char* hbtgpu_cs_vids_last_error(hbtgpu_cs_vids* v);
int hbtgpu_cs_assign_vids_device(hbtgpu_cs_ctx* ctx, hbtgpu_cs_vids* v);
uint64_t hbtgpu_cs_vids_count(const hbtgpu_cs_vids* v);
uint64_t hbtgpu_cs_vids_name_count(const hbtgpu_cs_vids* v);
float hbtgpu_cs_vids_last_ms(const hbtgpu_cs_vids* v);   /* device time of the last call */
/* thread table and names as hbtgpu_cs_threads_get returns them */
int hbtgpu_cs_vids_get(hbtgpu_cs_vids* v, hbtgpu_cs_thread* threads, uint64_t thread_cap,
                       hbtgpu_cs_name* names, uint64_t name_cap);

/* ---- Count intervals x tagstack slices (IntervalSlicer side) ---------
 *
 * Replaces, for a whole window of count rows at once,
 *   IntervalSlicer::makeIntervalIntersectingSlices   hbt/src/tagstack/IntervalSlicer.cpp:52-188
 *   the row loop of TraceMonitor::releaseCountData_  hbt/src/mon/TraceMonitor.cpp:52-96
 *   IntervalSlicer::consumeSlices                    IntervalSlicer.cpp:195-217 (one row)
 * For every comp unit (CPU) u the caller hands over the slices its IntervalSlicer holds
 * (slices[slice_begin[u] .. slice_begin[u+1]), in ring order) and the count rows of the window
 * (rows[row_begin[u] .. row_begin[u+1]), row_stride u64 words each, word 0 = tstamp, word 1 =
 * duration: the rows hbtgpu_get_rows returns).  Row by row, in order, exactly as the reference:
 * stop in front of the first row with tstamp + duration >= stop_ts; otherwise intersect
 * [tstamp, tstamp + duration) with the slices not yet consumed (marks[u] = the consumer mark,
 * an index into the unit's slices), trimming the first and last slice to the interval
 * (HBTGPU_ISLICE_TRIM -> TransitionType::Analysis) and giving up on the interval when the
 * slices do not cover it (HBTGPU_ISLICE_ABORT_ON_MISSING); the mark advances by the slices the
 * interval consumed.  Results stay on the device until fetched.
 *
 * The reference walks slice by slice; here every interval finds its slices by binary search and
 * the chain of marks is resolved with a parallel scan.  Units whose input
 * breaks what that needs (slices out of order or overlapping, rows going back in time, ...) are
 * walked exactly as the reference does by one thread, so results are the reference's on ANY
 * input.  Where the reference throws (zero-length interval, zero-length slice in the skipped
 * prefix, durations exceeding the interval: IntervalSlicer.cpp:64,77,175-183) the unit's status
 * is -EINVAL and err_interval names the row.
 */

/* tagstack::slice::Slice, hbt/src/tagstack/Slicer.h:30-49: same size and field order */
typedef struct {
  uint64_t tstamp;
  uint64_t duration;
  uint64_t stack_id;
  uint8_t swin_type;  /* Slice::TransitionType, Slicer.h:31-43: 0 NA, 1 Analysis, 2 ThreadPreempted, 3 ThreadYield, 4 PhaseChange */
  uint8_t swout_type;
  uint8_t pad[6];
} hbtgpu_slice;

#define HBTGPU_ISLICE_TRIM 0x1u
#define HBTGPU_ISLICE_ABORT_ON_MISSING 0x2u
#define HBTGPU_ISLICE_DEVICE_INPUT 0x4u  /* slices and rows are device pointers (the begin arrays and marks stay host) */
#define HBTGPU_ISLICE_FORCE_EXACT 0x8u   /* walk every unit serially (the fallback path), for tests */

typedef struct {
  uint64_t n_intervals;   /* rows processed (the stop_ts rule, TraceMonitor.cpp:63-70) */
  uint64_t n_slices;      /* slices in all its intervals */
  uint64_t mark;          /* consumer mark afterwards */
  int64_t status;         /* 0, or -EINVAL where the reference throws */
  uint64_t err_interval;  /* the row it throws on (~0 if none) */
  uint32_t exact_path;    /* 1 if the unit took the serial walk */
  uint32_t pad;
} hbtgpu_islice_unit;

typedef struct hbtgpu_islice_ctx hbtgpu_islice_ctx;
hbtgpu_islice_ctx* hbtgpu_islice_create(int device);
void hbtgpu_islice_destroy(hbtgpu_islice_ctx* ctx);
const char* hbtgpu_islice_last_error(hbtgpu_islice_ctx* ctx);
/* marks: [n_units] in/out.  units: [n_units] out (may be NULL; hbtgpu_islice_get_units later). */
int hbtgpu_islice_release(hbtgpu_islice_ctx* ctx, const hbtgpu_slice* slices, const uint64_t* slice_begin,
                          const uint64_t* rows, uint64_t row_stride, const uint64_t* row_begin,
                          uint32_t n_units, uint64_t stop_ts, uint32_t flags, uint64_t* marks,
                          hbtgpu_islice_unit* units);
int hbtgpu_islice_get_units(hbtgpu_islice_ctx* ctx, hbtgpu_islice_unit* units, uint32_t cap);
/* One unit's PerIntervalSlices (IntervalSlicer.h:17): counts[k] = slices of interval k (0 = the
 * empty entry pushed for a nullptr result, TraceMonitor.cpp:82-85), slices = all of them in
 * interval order.  Either pair may be NULL/0. */
int hbtgpu_islice_get(hbtgpu_islice_ctx* ctx, uint32_t unit, uint32_t* counts, uint64_t counts_cap,
                      hbtgpu_slice* slices, uint64_t slices_cap);
/* Device view for consumers on the GPU: unit u's intervals are interval_offset/interval_count
 * [row_begin[u] + k] (offset relative to unit_base[u]), interval k's slices
 * slices[unit_base[u] + interval_offset .. + interval_count).  Consecutive intervals need not be
 * adjacent in memory (unused slots may lie between them); hbtgpu_islice_get packs. */
typedef struct {
  const hbtgpu_slice* slices;
  const uint64_t* unit_base;        /* [n_units] */
  const uint32_t* interval_offset;  /* [total rows] */
  const uint32_t* interval_count;   /* [total rows] */
  uint32_t n_units;
} hbtgpu_islice_view;
int hbtgpu_islice_device_view(hbtgpu_islice_ctx* ctx, hbtgpu_islice_view* view);
float hbtgpu_islice_last_ms(const hbtgpu_islice_ctx* ctx); /* device time of the kernels of the last release */

/* ---- Stack-id frequencies and the interval x bin duration matrix ----
 *
 * Replaces, over the per-interval slices the last hbtgpu_islice_release left on the device,
 *   CuMonData::stackIdFreqs / MonData::stackIdFreqs   hbt/src/mon/MonData.h:236-256,382-398
 *   binmap::TagStackIdBinner (constructor)            hbt/src/mon/MonData.h:528-607
 *   binmap::BinMapper / IntervalBinMatrix<uint64_t>   :425-523
 * Per stack id: total duration, number of slices, number of distinct intervals (per unit) it
 * shows up in.  A stack id with num_intervals >= min_distinct_intervals and duration >=
 * min_duration gets a bin of its own, all others share bin 0 (ignored_bid); own bins are
 * numbered in ascending stack_id order (the reference numbers them in the iteration order of a
 * std::unordered_map, i.e. unspecified).  The matrix has one row per count interval of the
 * selected units and one column per bin; as in the reference every unit starts adding at row 0,
 * intervals without slices are passed over, zero-length slices are not added, and an interval
 * only moves the unit to its next row if its durations added up to something (:563-597).
 * unit_selected: [n_units] 0/1 (the reference's cpu_set), NULL = all.  max_stack_ids bounds the
 * device dictionary (-ENOSPC beyond), max_matrix_bytes the dense matrix (-E2BIG beyond; the
 * summary and the frequencies are valid then).  -EINVAL if a selected unit's release failed. */
typedef struct {
  uint64_t stack_id, duration, num_obs, num_intervals;  /* SliceFreq, MonData.h:31-51 */
  uint32_t bid, pad;                                    /* BinMapper::binkey_to_bid */
} hbtgpu_bin_freq;
typedef struct {
  uint64_t num_rows;       /* IntervalBinMatrix::kNumIntervals = MonData::numCountIntervals */
  uint64_t num_bins;       /* kNumBids, the ignored bin included */
  uint64_t num_stack_ids;  /* distinct stack ids = entries of hbtgpu_islice_bin_get's freqs */
  uint64_t rows_skipped;   /* intervals with slices whose durations are all 0 (num_skipped, :565,595) */
} hbtgpu_bin_summary;
int hbtgpu_islice_bin(hbtgpu_islice_ctx* ctx, const uint8_t* unit_selected, uint64_t min_distinct_intervals,
                      uint64_t min_duration, uint64_t max_stack_ids, uint64_t max_matrix_bytes,
                      hbtgpu_bin_summary* summary);
/* The same with the numbering of the bins taken from outside, for comp units spread over several
 * GPUs (no counterpart in the reference, which bins one MonData in one process): every rank
 * runs hbtgpu_islice_bin with thresholds 0 / 0 or reads its frequencies another way, the ranks
 * add the frequencies of equal stack ids up, apply the thresholds to the sums and number the bins
 * (dynolog_b200/shard.py: merge_bin_freqs), then each rank builds its matrix with the agreed
 * numbering -- stack_ids strictly ascending, bids[i] < num_bins, num_rows = the intervals of all
 * ranks -- and the matrices are summed (one integer all-reduce: every unit starts at row 0 in the
 * reference too).  -EINVAL if a local stack id is missing from the numbering. */
int hbtgpu_islice_bin_with(hbtgpu_islice_ctx* ctx, const uint8_t* unit_selected, const uint64_t* stack_ids,
                           const uint32_t* bids, uint64_t n_ids, uint64_t num_bins, uint64_t num_rows,
                           uint64_t max_stack_ids, uint64_t max_matrix_bytes, hbtgpu_bin_summary* summary);
/* freqs in ascending stack_id order; durations = the matrix, num_rows x num_bins, row-major.  Either may be NULL. */
int hbtgpu_islice_bin_get(hbtgpu_islice_ctx* ctx, hbtgpu_bin_freq* freqs, uint64_t freq_cap, uint64_t* durations,
                          uint64_t durations_cap);
const uint64_t* hbtgpu_islice_bin_durations_device(hbtgpu_islice_ctx* ctx); /* the matrix where it was built */

/* ---- tagstack events -> slices (Slicer side) ----------------------------
 *
 * Replaces tagstack::slice::Slicer<>::processNext (hbt/src/tagstack/Slicer.h:336-459, with
 * createStackState_ :560-584, activateStackState* :515-634, handleWriteErrors_ :668-703,
 * handleThreadLifetime_ :705-745, handleLastState_ :747-787, completeSlice_ :529-558) for one
 * window of events of all comp units at once: unit u's events (16-byte tagstack::Event records,
 * what hbtgpu_cs_get_events / hbtgpu_cs_events_device hand out after the thread dictionary has
 * run) at events[event_begin[u] .. event_begin[u+1]), each unit in stream order.  The merge of
 * the units (Event::happenedBefore: tstamp, then comp unit, tagstack/Event.h:217-223), the stack
 * dictionary with ids in order of creation, the dormant lists and the resets are computed on
 * the GPU.  A call starts an empty Slicer unless HBTGPU_SLICER_CONTINUE is
 * set: then the window follows the previous call's (same units) -- active states with their
 * switch-in times, the dictionary with its ids and dormant flags, the write-error depth, the
 * statistics and the last timestamp carry over, exactly as one Slicer fed both windows would
 * have them.  stats / the stack table / last_ts are cumulative since the last fresh start.
 *
 * Two forms, both on the device.  The PARALLEL form (scans and sorts over all events at once) takes
 * switch events at level 0 (SwitchIn / SwitchOut*), thread lifetime events (those at another
 * level are dropped as logic errors, as in the reference, :711-720) and WriteErrors* without a
 * WriteErrorsEnd in front of its Start; a switch-in on a unit whose active state is a thread
 * resets everything, as in the reference (:839-868).  The GENERAL form takes every event the
 * reference takes -- phases (Start / End) and switches at any of the six levels -- by keeping the
 * reference's order: the units are merged on the GPU and one thread walks the merged events with
 * the stack dictionary in device memory (about a microsecond per event: phase events are few).
 * A fresh window whose events need the general form takes it by itself; HBTGPU_SLICER_GENERAL asks
 * for it outright, and a Slicer that continues over windows stays on the form its first window
 * took (a later window with phases on a parallel Slicer is -ENOTSUP and leaves the state as it
 * was: give HBTGPU_SLICER_GENERAL from the first window on).
 *
 * Refused by both forms: a unit's events that go back in time, or start before the previous
 * window's last timestamp (-ENOTSUP: the merge would not be the reference's); comp_unit_id that
 * differs from the unit an event was filed under (-EINVAL, the reference throws, :351-352);
 * unknown event types and levels >= 6 (-EINVAL, HBT_DCHECK_LT :431).  Refused by the general
 * form with -ENOTSUP: an event at a level above the active stack's height -- there the reference
 * copies tags it never initialised (Stack::copyTruncated / copyFillUpTo, TagStack.h:103-118) and
 * its answer depends on what the memory held.  After an error the Slicer's state is undefined
 * unless stated otherwise: start afresh.  There is no CPU path.
 */
#define HBTGPU_SLICER_DEVICE_INPUT 0x1u /* events is a device pointer (event_begin stays host) */
#define HBTGPU_SLICER_CONTINUE 0x2u     /* this window follows the previous call's: keep the Slicer's state */
#define HBTGPU_SLICER_GENERAL 0x4u      /* take the general path (below) whatever the events are */
typedef struct {  /* SlicerStats, Slicer.h:276-285 */
  uint64_t num_processed, num_error_start, num_error_end, num_error_out_sync, num_dropped, num_out_of_order,
      num_slices_completed, num_logic_errors;
} hbtgpu_slicer_stats;
typedef struct {  /* one entry of getTagStackStatsDump(), Slicer.h:160-193,478-485 */
  uint64_t stack_id, parent_stack_id;  /* parent ~0 (kInvalidTagStackId) for the empty stacks */
  uint32_t tags[6];                    /* tags[i >= num_set_tags] = ~0 (kNA) */
  uint8_t num_set_tags, known_num_levels, pad[6];
  uint64_t cum_active_time, num_swins, num_swouts;
} hbtgpu_stack_stats;
typedef struct hbtgpu_slicer_ctx hbtgpu_slicer_ctx;
hbtgpu_slicer_ctx* hbtgpu_slicer_create(int device);
void hbtgpu_slicer_destroy(hbtgpu_slicer_ctx* ctx);
const char* hbtgpu_slicer_last_error(hbtgpu_slicer_ctx* ctx);
/* n_slices: [n_units] slices written for each unit; any output pointer may be NULL. */
int hbtgpu_slicer_run(hbtgpu_slicer_ctx* ctx, const hbtgpu_cs_event* events, const uint64_t* event_begin,
                      uint32_t n_units, uint32_t flags, hbtgpu_slicer_stats* stats, uint64_t* n_slices,
                      uint64_t* n_stacks, uint64_t* last_ts);
int hbtgpu_slicer_get_slices(hbtgpu_slicer_ctx* ctx, uint32_t unit, hbtgpu_slice* out, uint64_t cap);
int hbtgpu_slicer_get_stacks(hbtgpu_slicer_ctx* ctx, hbtgpu_stack_stats* out, uint64_t cap); /* by stack_id */
/* Where the slices are: unit u at slices[slice_begin[u] .. slice_begin[u+1]) (both device pointers;
 * the layout hbtgpu_islice_release takes with HBTGPU_ISLICE_DEVICE_INPUT). */
int hbtgpu_slicer_slices_device(hbtgpu_slicer_ctx* ctx, const hbtgpu_slice** slices, const uint64_t** slice_begin);
float hbtgpu_slicer_last_ms(const hbtgpu_slicer_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif /* HBTGPU_H */
