/*
 * oracle/hbt_oracle.c -- plain-C CPU restatement of the hbt perf_event ring
 * decode + aggregate path.  TEST INFRASTRUCTURE ONLY (see oracle/hbtref_api.h):
 * it is the checker for the CUDA path and the "port" CPU baseline; the product
 * never links it.
 *
 * Parity status: PINNED.  tests/test_oracle_pinning.py checks every function
 * here against (a) oracle/_ref/libhbtref.so = the real reference compiled from
 * /root/reference (differential, seeded streams incl. error paths) whenever
 * that library is present, (b) the golden vectors in tests/golden/ generated
 * from it (tests/golden/make_golden.py), and (c) the reference's own
 * known-answer tests (hbt/src/mon/tests/MonitorMockTest.cpp:86-88,121-126).
 * The per-bucket attribution rule (A7) is a spec extension with no reference
 * code; see DESIGN.md.
 *
 * Each function cites the reference lines it follows.
 */
#define _GNU_SOURCE
#include <errno.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "hbtref_api.h"

/* linux/perf_event.h record type codes used by consume()'s switch */
enum {
  REC_LOST = 2, REC_COMM = 3, REC_EXIT = 4, REC_THROTTLE = 5, REC_UNTHROTTLE = 6,
  REC_FORK = 7, REC_READ = 8, REC_SAMPLE = 9, REC_AUX = 11, REC_ITRACE_START = 12,
  REC_SWITCH_CPU_WIDE = 15
};

#define INVALID_TS UINT64_MAX /* kInvalidTimeStamp, hbt/src/common/System.h:106 */

static inline uint64_t ld64(const uint8_t* p) {
  uint64_t v;
  memcpy(&v, p, 8);
  return v;
}

/* PerfEventsGroup::kernelTimeFromTsc, hbt/src/perf_event/PerfEventsGroup.h:1645-1658 */
static inline uint64_t tsc_to_ktime(const hbtref_params* p, uint64_t tsc) {
  const uint64_t rem_flags = (((uint64_t)1) << p->time_shift) - 1; /* :1193 */
  uint64_t quot = tsc >> p->time_shift;
  uint64_t rem = tsc & rem_flags;
  return p->time_zero + quot * p->time_mult + ((rem * p->time_mult) >> p->time_shift);
}

/* PerfEventsGroup::tscFromKernelTime, PerfEventsGroup.h:1661-1678 */
static inline uint64_t ktime_to_tsc(const hbtref_params* p, uint64_t kt) {
  uint64_t delta = kt - p->time_zero;
  uint64_t quot = delta / p->time_mult;
  uint64_t rem = delta % p->time_mult;
  return (quot << p->time_shift) + (rem << p->time_shift) / p->time_mult;
}

/* GroupReadValues::getCount, PerfEventsGroup.h:467-481 */
static inline uint64_t get_count(uint64_t count, uint64_t te, uint64_t tr) {
  if (te == 0) return 0;
  if (tr == 0) return 0;
  /* x86-64 cvttsd2si semantics for out-of-range: the reference is UB there;  */
  /* the harness never feeds such values to the comparison (see tests).       */
  return (uint64_t)((double)count * (double)te / (double)tr);
}

const char* hbtoracle_describe(void) {
  return "plain-C restatement of hbt consume()+CpuCountSampleGenerator (oracle/hbt_oracle.c)";
}

/*
 * The walk.  Mirrors PerfEventsGroup::consume (PerfEventsGroup.h:1394-1589)
 * called repeatedly until -ENODATA, with CpuCountSampleGenerator's handlers
 * (PerCpuCountSampleGenerator.h:110-202) inlined.
 */
static uint64_t walk(const hbtref_params* p, const hbtref_carry* in,
    hbtref_carry* out, const uint8_t* bytes, uint64_t len, hbtref_counters* ctr,
    uint64_t* rec_off, uint32_t* rec_type, uint32_t* rec_size, uint64_t rec_cap,
    uint64_t* rows, uint64_t rows_cap, uint64_t* bucket_acc,
    uint64_t* sink_tid, uint64_t* sink_ip, uint64_t sink_cap) {
  uint64_t sink_n = 0;
  const uint32_t n = p->n_events, F = n + 3;
  const uint32_t pb = 8 * p->pre;        /* extra bytes before `time` (spec extension, 0 for hbt) */
  const uint32_t fix = 40 + pb;          /* sizeof(Sample) without values */
  hbtref_counters c;
  memset(&c, 0, sizeof(c));
  /* generator state: count_.tstamp, last_[] (:222-228) + te/tr for A7 */
  uint64_t st_ts = in->valid ? in->tstamp : INVALID_TS;
  uint64_t last[HBTREF_MAX_EVENTS];
  uint64_t pte = in->te, ptr_ = in->tr;
  for (uint32_t i = 0; i < n; ++i) last[i] = in->last[i];

  uint64_t tail = 0, nrec = 0, nrows = 0;
  while (tail < len) { /* :1415 */
    const uint8_t* r = bytes + tail;
    uint64_t hdr = ld64(r);
    uint32_t type = (uint32_t)hdr;
    uint32_t size = (uint32_t)(hdr >> 48);
    /* Fences (inputs on which the reference itself misbehaves):   */
    /* size==0 spins consume() (:1415,:1570); size%8!=0 or a record running past  */
    /* data_head walks the reference off the record grid.  Reported as -EPERM at  */
    /* this record, nothing after it is decoded.                                  */
    if (size < 8 || (size & 7) || tail + size > len) {
      c.err = -EPERM;
      goto done;
    }
    switch (type) { /* :1439 */
      case REC_LOST: /* :1440-1450; handler :110-124 */
        if (size < 24) { /* fence: the reference would read num_lost past the record */
          c.err = -EPERM;
          goto done;
        }
        st_ts = INVALID_TS;
        c.lost += ld64(r + 16); /* num_lost, not #records (:1448); size>=24 fenced below */
        break;
      case REC_THROTTLE: c.throttle++; break;     /* :1476-1481 */
      case REC_UNTHROTTLE: c.unthrottle++; break; /* :1482-1487 */
      case REC_SAMPLE: { /* :1499-1514 */
        uint64_t nr = size + 16 >= fix ? ld64(r + pb + 16) : 0;
        if (size != fix + nr * 8) { /* recordSize(), :346-350, :1502-1507 */
          c.err = -EPERM;
          goto done;
        }
        /* Fence: nr != kNumEvents is only a debug DCHECK in the reference (:133), */
        /* which then reads kNumEvents values anyway; here it is a hard error.     */
        if (nr != n) {
          c.err = -EPERM;
          goto done;
        }
        uint64_t ts = ld64(r + pb + 8), te = ld64(r + pb + 24), tr = ld64(r + pb + 32);
        if (sink_cap) { /* config-4 extension: remember who was sampled, and where */
          if (sink_n < sink_cap) {
            sink_tid[sink_n] = ld64(r + 16);
            sink_ip[sink_n] = ld64(r + 8);
          }
          sink_n++;
        }
        if (st_ts == INVALID_TS) { /* handler :135-154 */
          st_ts = ts;
          for (uint32_t i = 0; i < n; ++i) {
            uint64_t v = ld64(r + fix + 8 * i);
            if ((p->tsc_mask >> i) & 1) v = tsc_to_ktime(p, v);
            last[i] = v;
          }
          pte = te;
          ptr_ = tr;
          c.num_unknown_count_interval++;
          c.sample++;
          break;
        }
        if (st_ts > ts) { /* :156-163 */
          st_ts = INVALID_TS;
          c.err = -EPERM;
          goto done;
        }
        uint64_t row[2 + HBTREF_MAX_EVENTS];
        row[0] = st_ts;
        row[1] = ts - st_ts; /* :165 */
        for (uint32_t i = 0; i < n; ++i) { /* :166-179 */
          uint64_t v = ld64(r + fix + 8 * i);
          if ((p->tsc_mask >> i) & 1) v = tsc_to_ktime(p, v);
          row[2 + i] = v - last[i];
          last[i] = v;
        }
        if (rows && nrows < rows_cap) memcpy(rows + nrows * (2 + n), row, 8 * (2 + n));
        nrows++;
        if (bucket_acc) { /* A7: whole row to the bucket of the sample's own time */
          if (ts >= p->t0 && (ts - p->t0) / p->bucket_ns < p->n_buckets) {
            uint64_t* a = bucket_acc + ((ts - p->t0) / p->bucket_ns) * F;
            for (uint32_t i = 0; i < n; ++i) a[i] += row[2 + i];
            a[n] += te - pte;      /* GroupReadValues::diff then accum, :548-573 */
            a[n + 1] += tr - ptr_;
            a[n + 2] += 1;
          } else {
            c.rows_out_of_range++;
          }
        }
        pte = te;
        ptr_ = tr;
        st_ts = ts; /* :199 */
        c.sample++;
      } break;
      case REC_READ: { /* :1515-1530; handler is a no-op :126-130 */
        uint64_t nr = size >= 24 ? ld64(r + 16) : 0;
        if (size != 40 + nr * 8) {
          c.err = -EPERM;
          goto done;
        }
        c.read++;
      } break;
      /* COMM/EXIT/FORK/AUX/ITRACE_START/SWITCH_CPU_WIDE: Sampling mode has   */
      /* void record types, so the case bodies compile away (:1451-1474,      */
      /* :1488-1498,:1531-1563): skipped by size, no counter.                 */
      default: break; /* unknown: log + skip by size (:1565-1567) */
    }
    if (rec_off && nrec < rec_cap) {
      rec_off[nrec] = tail;
      rec_type[nrec] = type;
      rec_size[nrec] = size;
    }
    nrec++;       /* :1569 every record counts */
    tail += size; /* :1570 */
  }
done:
  c.num_records = nrec;
  c.num_rows = nrows;
  c.err_offset = tail;
  c.last_tstamp = st_ts;
  if (ctr) *ctr = c;
  if (out) {
    memset(out, 0, sizeof(*out));
    out->valid = st_ts != INVALID_TS;
    out->tstamp = st_ts;
    out->te = pte;
    out->tr = ptr_;
    for (uint32_t i = 0; i < n; ++i) out->last[i] = last[i];
  }
  return sink_n;
}

int hbtoracle_decode_stream(const hbtref_params* p, const hbtref_carry* in,
    hbtref_carry* out, const uint8_t* bytes, uint64_t len, hbtref_counters* ctr,
    uint64_t* rec_off, uint32_t* rec_type, uint32_t* rec_size, uint64_t rec_cap,
    uint64_t* rows, uint64_t rows_cap, uint64_t* bucket_acc) {
  walk(p, in, out, bytes, len, ctr, rec_off, rec_type, rec_size, rec_cap, rows, rows_cap, bucket_acc,
       NULL, NULL, 0);
  return 0;
}

uint64_t hbtoracle_collect_ip_tid(const hbtref_params* p, const hbtref_carry* in, const uint8_t* bytes,
    uint64_t len, uint64_t* tid_keys, uint64_t* ips, uint64_t cap) {
  /* a time-regressed SAMPLE is not counted by consume(); the sink runs after that check? No:   */
  /* it is recorded when the record passes the size checks, so drop the last entry on -EPERM    */
  /* from the regression rule to stay consistent with `sample` (see below).                     */
  hbtref_counters c;
  uint64_t k = walk(p, in, NULL, bytes, len, &c, NULL, NULL, NULL, 0, NULL, 0, NULL, tid_keys, ips,
                    cap ? cap : 1);
  return k > c.sample ? c.sample : k;
}

/* accum()-then-getCount() per bucket: PerPerfEventsGroupBase.h:182-194 + :467-481 */
void hbtoracle_scale_buckets(const hbtref_params* p, const uint64_t* bucket_acc,
    uint64_t* scaled) {
  const uint32_t n = p->n_events, F = n + 3;
  for (uint64_t b = 0; b < p->n_buckets; ++b) {
    const uint64_t* a = bucket_acc + b * F;
    for (uint32_t i = 0; i < n; ++i) scaled[b * n + i] = get_count(a[i], a[n], a[n + 1]);
  }
}

void hbtoracle_get_count(const uint64_t* count, const uint64_t* te, const uint64_t* tr,
    uint64_t* out, uint64_t n) {
  for (uint64_t i = 0; i < n; ++i) out[i] = get_count(count[i], te[i], tr[i]);
}

/* getRateReducer, BuiltinMetrics.cpp:102-118 */
void hbtoracle_rate(const uint64_t* num, const uint64_t* den, double* out,
    uint8_t* valid, uint64_t n) {
  for (uint64_t i = 0; i < n; ++i) {
    if (den[i] == 0) {
      valid[i] = 0;
      out[i] = 0.0;
    } else {
      valid[i] = 1;
      out[i] = (double)num[i] / (double)den[i];
    }
  }
}

/* getAddReducer, BuiltinMetrics.cpp:88-97: std::accumulate(..., 0) -- the init
 * is an int, so every partial sum is converted back to int (mod 2^32, signed). */
void hbtoracle_add_reduce(const uint64_t* vals, uint64_t k, double* out, uint64_t n) {
  for (uint64_t i = 0; i < n; ++i) {
    int32_t acc = 0;
    for (uint64_t j = 0; j < k; ++j)
      acc = (int32_t)(uint32_t)((uint64_t)(int64_t)acc + vals[i * k + j]);
    out[i] = (double)acc;
  }
}

void hbtoracle_tsc_to_ktime(const hbtref_params* p, const uint64_t* tsc, uint64_t* out,
    uint64_t n) {
  for (uint64_t i = 0; i < n; ++i) out[i] = tsc_to_ktime(p, tsc[i]);
}

void hbtoracle_ktime_to_tsc(const hbtref_params* p, const uint64_t* kt, uint64_t* out,
    uint64_t n) {
  for (uint64_t i = 0; i < n; ++i) out[i] = ktime_to_tsc(p, kt[i]);
}

/* dynolog/src/PerfMonitor.cpp:63-66 (mips) and :69-72 (mega_cycles_per_second) */
void hbtoracle_per_second_mega(const uint64_t* count, const uint64_t* tr, float* out,
    uint64_t n) {
  for (uint64_t i = 0; i < n; ++i)
    out[i] = (float)((double)count[i] * 1e3 / (double)tr[i]);
}

/* getRunningRatio :513-522 (te/tr despite its comment), getRunningRatioInteger<100>
 * :527-535, hasMultiplexing :538-540 */
void hbtoracle_running_ratio(const uint64_t* te, const uint64_t* tr, float* ratio,
    uint64_t* ratio_int100, uint8_t* has_mux, uint64_t n) {
  for (uint64_t i = 0; i < n; ++i) {
    ratio[i] = te[i] == 0 ? 1.0f : (float)te[i] / (float)tr[i];
    ratio_int100[i] = te[i] == 0 ? 100 : (100 * tr[i]) / te[i];
    has_mux[i] = te[i] != 0 && tr[i] != te[i];
  }
}

/* ---- timed whole-capture path (port): same walk, streams split over threads ---- */
typedef struct {
  const hbtref_params* p;
  uint32_t n_streams;
  const uint8_t* const* bytes;
  const uint64_t* lens;
  const hbtref_carry* in;
  uint64_t* bucket_acc;
  hbtref_counters* ctrs;
  volatile uint32_t* next;
} run_args;

static void* run_worker(void* vp) {
  run_args* a = (run_args*)vp;
  const uint32_t F = a->p->n_events + 3;
  for (;;) {
    uint32_t s = __atomic_fetch_add(a->next, 1, __ATOMIC_RELAXED);
    if (s >= a->n_streams) break;
    hbtref_counters c;
    hbtoracle_decode_stream(a->p, &a->in[s], NULL, a->bytes[s], a->lens[s], &c, NULL, NULL,
        NULL, 0, NULL, 0,
        a->bucket_acc ? a->bucket_acc + (uint64_t)s * a->p->n_buckets * F : NULL);
    if (a->ctrs) a->ctrs[s] = c;
  }
  return NULL;
}

double hbtoracle_run_capture(const hbtref_params* p, uint32_t n_streams,
    const uint8_t* const* bytes, const uint64_t* lens, const hbtref_carry* in,
    uint32_t n_threads, uint32_t batch, uint64_t* bucket_acc, hbtref_counters* ctrs) {
  (void)batch;
  if (n_threads == 0) n_threads = 1;
  if (n_threads > n_streams) n_threads = n_streams;
  volatile uint32_t next = 0;
  run_args a = {p, n_streams, bytes, lens, in, bucket_acc, ctrs, &next};
  pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * n_threads);
  struct timespec t0, t1;
  clock_gettime(CLOCK_MONOTONIC, &t0);
  for (uint32_t t = 0; t < n_threads; ++t) pthread_create(&th[t], NULL, run_worker, &a);
  for (uint32_t t = 0; t < n_threads; ++t) pthread_join(th[t], NULL);
  clock_gettime(CLOCK_MONOTONIC, &t1);
  free(th);
  return (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
}

/* ---- series reductions (sink side) ---------------------------------
 * Restates facebook::dynolog::MetricSeries<uint64_t> (dynolog/src/metric_frame/MetricSeries.h):
 *   avg<R>      std::accumulate(begin, end, R{0}) / (end - begin)   :207-213 (difference is int)
 *   percentile  nth_element at lround(percentage * (size - 1)), float arithmetic   :215-226
 *   diff        *(end - 1) - *begin                                                 :228-234
 *   rate<R>     integral: value * (period / duration) or value / (duration / period);
 *               floating: R(value) * double(period in us) / (duration in us)        :246-263
 * Pinned against the real class by tests/test_series_oracle.py. */
#include <math.h>
#include <stdlib.h>

#include "hbtseries_api.h"

static int cmp_u64(const void* a, const void* b) {
  const uint64_t x = *(const uint64_t*)a, y = *(const uint64_t*)b;
  return x < y ? -1 : x > y;
}

int hbtoracle_series_reduce(const uint64_t* series, uint64_t n_series, uint64_t len, float percentage,
    int64_t period_ns, int64_t duration_ns, double* avg_f64, uint64_t* avg_u64, uint64_t* percentile,
    uint64_t* diff, double* rate_f64, uint64_t* rate_u64) {
  if (len == 0 || percentage < 0.f || percentage > 1.f || period_ns <= 0 || duration_ns <= 0) return -1;
  uint64_t* copy = percentile ? (uint64_t*)malloc(len * sizeof(uint64_t)) : NULL;
  const int count = (int)len;
  for (uint64_t s = 0; s < n_series; ++s) {
    const uint64_t* v = series + s * len;
    double accd = 0.0;
    uint64_t accu = 0;
    for (uint64_t i = 0; i < len; ++i) {
      accd = accd + (double)v[i];
      accu += v[i];
    }
    if (avg_f64) avg_f64[s] = accd / count;
    if (avg_u64) avg_u64[s] = accu / (uint64_t)count;
    if (percentile) {
      memcpy(copy, v, len * sizeof(uint64_t));
      qsort(copy, len, sizeof(uint64_t), cmp_u64);
      const long nth = lroundf(percentage * (float)(len - 1));
      percentile[s] = copy[nth];
    }
    const uint64_t d = v[len - 1] - v[0];
    if (diff) diff[s] = d;
    if (rate_f64) rate_f64[s] = (double)d * (double)(period_ns / 1000) / (double)(duration_ns / 1000);
    if (rate_u64)
      rate_u64[s] = period_ns > duration_ns ? d * (uint64_t)(period_ns / duration_ns)
                                            : d / (uint64_t)(duration_ns / period_ns);
  }
  free(copy);
  return 0;
}
