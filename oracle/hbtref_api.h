/*
 * oracle/hbtref_api.h -- plain-C interface shared by the two CPU checkers:
 *
 *   oracle/_ref/libhbtref.so   ("reference": the REAL hbt code, compiled from
 *                               /root/reference by oracle/Makefile, driven by
 *                               oracle/ref_harness.cpp)
 *   oracle/libhbtoracle.so     ("port": oracle/hbt_oracle.c, a plain-C
 *                               restatement that also runs where
 *                               /root/reference does not exist)
 *
 * THIS IS TEST INFRASTRUCTURE. Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load these libraries.
 * The product (dynolog_b200/, include/hbtgpu.h) never links or calls them.
 *
 * Both libraries export the same symbols with prefix hbtref_ / hbtoracle_.
 * Field meanings follow the reference:
 *   counters  -> PerfEventsGroup::getNumRecord*()    hbt/src/perf_event/PerfEventsGroup.h:791-824
 *   carry     -> CpuCountSampleGenerator::count_.tstamp / last_[]
 *                                                   hbt/src/perf_event/PerCpuCountSampleGenerator.h:222-228
 *   rows      -> [tstamp, duration, delta_0..n-1]    PerCpuCountSampleGenerator.h:34,69,165-182
 *   buckets   -> SURVEY.md section 8 row A7 (spec extension)
 */
#ifndef HBTREF_API_H
#define HBTREF_API_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HBTREF_MAX_EVENTS 64

typedef struct {
  uint32_t n_events;    /* kNumEvents */
  uint32_t n_buckets;   /* buckets per stream */
  uint64_t tsc_mask;    /* bit i: event i has EventValueTransforms::TscToKernelTime */
  uint32_t time_shift;  /* perf_event_mmap_page::time_shift */
  uint32_t time_mult;   /* perf_event_mmap_page::time_mult  */
  uint64_t time_zero;   /* perf_event_mmap_page::time_zero  */
  uint64_t t0;          /* bucket 0 starts here (ns) */
  uint64_t bucket_ns;   /* bucket width */
  uint32_t pre;         /* words between header and `time`: 0 = TIME|READ (hbt), 2 = IP|TID|TIME|READ */
  uint32_t pad;         /*   (pre != 0 is a spec extension: only the port implements it)          */
} hbtref_params;

typedef struct {
  uint64_t valid;   /* count_.tstamp != kInvalidTimeStamp */
  uint64_t tstamp;  /* count_.tstamp */
  uint64_t te, tr;  /* previous sample's time_enabled / time_running (A7) */
  uint64_t last[HBTREF_MAX_EVENTS]; /* last_[] (already transformed) */
} hbtref_carry;

typedef struct {
  uint64_t lost, comm, exit, throttle, unthrottle, fork, sample, read, aux,
      itrace_start, switch_cpu_wide;   /* the 11 getNumRecord*() */
  uint64_t num_records;                /* records walked (all types) */
  uint64_t num_unknown_count_interval; /* PerCpuCountSampleGenerator.h:152 */
  uint64_t num_rows;                   /* rows emitted */
  uint64_t rows_out_of_range;          /* rows whose bucket is outside [0,n_buckets) */
  int64_t err;                         /* 0, or the negative errno consume() returned */
  uint64_t err_offset;                 /* data_tail when the walk stopped */
  uint64_t last_tstamp;                /* getLastTimeStamp() */
} hbtref_counters;

#define HBTREF_DECL(prefix)                                                        \
  /* Walk one per-CPU stream exactly as repeated consume() calls would. Any     */ \
  /* output pointer may be NULL. bucket_acc is [n_buckets][n_events+3] u64:     */ \
  /* sum delta[0..n), sum dTE, sum dTR, n_rows; it is ACCUMULATED into.         */ \
  int prefix##_decode_stream(const hbtref_params* p, const hbtref_carry* in,       \
      hbtref_carry* out, const uint8_t* bytes, uint64_t len, hbtref_counters* ctr, \
      uint64_t* rec_off, uint32_t* rec_type, uint32_t* rec_size, uint64_t rec_cap, \
      uint64_t* rows, uint64_t rows_cap, uint64_t* bucket_acc);                    \
  /* GroupReadValues::getCount over n_buckets x n_events (A6/A7).               */ \
  void prefix##_scale_buckets(const hbtref_params* p, const uint64_t* bucket_acc,  \
      uint64_t* scaled);                                                           \
  /* GroupReadValues::getCount (PerfEventsGroup.h:467-481), elementwise.        */ \
  void prefix##_get_count(const uint64_t* count, const uint64_t* te,               \
      const uint64_t* tr, uint64_t* out, uint64_t n);                              \
  /* getRateReducer (BuiltinMetrics.cpp:102-118): returns validity per element. */ \
  void prefix##_rate(const uint64_t* num, const uint64_t* den, double* out,        \
      uint8_t* valid, uint64_t n);                                                 \
  /* getAddReducer (BuiltinMetrics.cpp:88-97) over rows of k values.            */ \
  void prefix##_add_reduce(const uint64_t* vals, uint64_t k, double* out,          \
      uint64_t n);                                                                 \
  /* kernelTimeFromTsc / tscFromKernelTime (PerfEventsGroup.h:1645-1678).       */ \
  void prefix##_tsc_to_ktime(const hbtref_params* p, const uint64_t* tsc,          \
      uint64_t* out, uint64_t n);                                                  \
  void prefix##_ktime_to_tsc(const hbtref_params* p, const uint64_t* kt,           \
      uint64_t* out, uint64_t n);                                                  \
  /* PerfMonitor::log mips formula (dynolog/src/PerfMonitor.cpp:63-76).         */ \
  void prefix##_per_second_mega(const uint64_t* count, const uint64_t* tr,         \
      float* out, uint64_t n);                                                     \
  /* getRunningRatio / getRunningRatioInteger<100> / hasMultiplexing            */ \
  /* (PerfEventsGroup.h:513-540).                                               */ \
  void prefix##_running_ratio(const uint64_t* te, const uint64_t* tr, float* ratio,\
      uint64_t* ratio_int100, uint8_t* has_mux, uint64_t n);                       \
  /* Timed multi-threaded CPU path over a whole capture: n_streams streams,     */ \
  /* each decoded + bucketised (raw sums and n_rows). Returns seconds.          */ \
  double prefix##_run_capture(const hbtref_params* p, uint32_t n_streams,          \
      const uint8_t* const* bytes, const uint64_t* lens, const hbtref_carry* in,   \
      uint32_t n_threads, uint32_t batch, uint64_t* bucket_acc,                    \
      hbtref_counters* ctrs);                                                      \
  const char* prefix##_describe(void);

HBTREF_DECL(hbtref)
HBTREF_DECL(hbtoracle)

/* BASELINE config 4 (spec extension, NO reference code: parity UNPINNED).  Walk one stream
 * of IP|TID|TIME|READ samples ({hdr, ip, pid:u32, tid:u32, time, nr, te, tr, value[nr]},
 * linux/perf_event.h PERF_RECORD_SAMPLE field order) and append (pid|tid<<32) and ip of every
 * SAMPLE the walk counts (same stop rules as hbtoracle_decode_stream).  Returns #samples. */
uint64_t hbtoracle_collect_ip_tid(const hbtref_params* p, const hbtref_carry* in, const uint8_t* bytes,
    uint64_t len, uint64_t* tid_keys, uint64_t* ips, uint64_t cap);

#ifdef __cplusplus
}
#endif
#endif
