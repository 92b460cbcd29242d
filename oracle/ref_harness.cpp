// oracle/ref_harness.cpp -- drives the UNMODIFIED reference (hbt) code over
// synthetic ring images.  TEST INFRASTRUCTURE ONLY (see oracle/hbtref_api.h).
//
// Built by oracle/Makefile into oracle/_ref/libhbtref.so with the reference
// headers and sources included/compiled from where they lie under
// /root/reference (nothing is copied).  What is "real reference" here:
//   * PerfEventsGroup<CpuCountSampleGenerator, mode::Sampling>::consume
//       hbt/src/perf_event/PerfEventsGroup.h:1394-1589
//   * CpuCountSampleGenerator::handleRecord{Sample,Lost,Read}
//       hbt/src/perf_event/PerCpuCountSampleGenerator.h:110-202
//   * ringbuffer::Consumer (row drain)  hbt/src/ringbuffer/Consumer.h
//   * GroupReadValues::{diff,accum,getCount,getRunningRatio*,hasMultiplexing}
//       hbt/src/perf_event/PerfEventsGroup.h:386-586
//   * getRateReducer / getAddReducer    hbt/src/perf_event/BuiltinMetrics.cpp:88-118
//   * kernelTimeFromTsc / tscFromKernelTime  PerfEventsGroup.h:1645-1678
// What is harness glue (the reference has no such code): the fake
// perf_event_mmap_page, the per-bucket attribution rule and the
// PerfMonitor::log formula (dynolog/src/PerfMonitor.cpp:63-76, which cannot be
// linked without glog/json; it is one expression and restated below).
//
// Access to protected members uses -fno-access-control (flag only; sources untouched).

#include <vector>  // hbt/src/common/Defs.h:27 uses std::vector without including it

#include "hbt/src/common/Defs.h"
#ifdef HBTREF_NOLOG
// Variant with hbt's per-record stderr logging compiled out (Defs.h is
// #pragma once, so these redefinitions survive the later includes).
#undef HBT_LOG_WARNING
#undef HBT_LOG_ERROR
struct HbtrefNullLog {
  template <class T>
  HbtrefNullLog& operator<<(const T&) { return *this; }
  HbtrefNullLog& operator<<(std::ostream& (*)(std::ostream&)) { return *this; }
};
#define HBT_LOG_WARNING() HbtrefNullLog()
#define HBT_LOG_ERROR() HbtrefNullLog()
#endif

#include "hbt/src/perf_event/BuiltinMetrics.h"
#include "hbt/src/perf_event/PerCpuCountSampleGenerator.h"
#include "hbt/src/ringbuffer/Consumer.h"

#include <atomic>
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <thread>

#include "hbtref_api.h"

namespace hp = facebook::hbt::perf_event;
namespace hr = facebook::hbt::ringbuffer;
using Gen = hp::CpuCountSampleGenerator;
using RV = hp::GroupReadValues<hp::mode::Sampling>;

namespace {

struct NullBuf : std::streambuf {
  int overflow(int c) override { return c; }
  std::streamsize xsputn(const char*, std::streamsize n) override { return n; }
};

// "stderr to /dev/null": hbt logs through std::cerr (Defs.h:177-181); formatting
// still happens, only the write is dropped.
std::atomic<int> g_silence_depth{0};
NullBuf g_nullbuf;
std::streambuf* g_old = nullptr;
void silence_begin() {
  if (g_silence_depth.fetch_add(1) == 0) g_old = std::cerr.rdbuf(&g_nullbuf);
}
void silence_end() {
  if (g_silence_depth.fetch_sub(1) == 1) std::cerr.rdbuf(g_old);
}

// A generator pointed at a fake perf_event_mmap_page in ordinary memory.
struct FakeGen {
  std::unique_ptr<Gen> gen;
  uint8_t* base = nullptr;
  size_t data_size = 0;
  volatile perf_event_mmap_page* page = nullptr;

  FakeGen(const hbtref_params& p, size_t max_len, size_t rb_rows) {
    hp::EventConfs confs(p.n_events);
    std::vector<std::string> names;
    for (uint32_t i = 0; i < p.n_events; ++i) {
      confs[i].id = "ev" + std::to_string(i);
      confs[i].transforms.flags = ((p.tsc_mask >> i) & 1)
          ? hp::EventValueTransforms::TscToKernelTime
          : hp::EventValueTransforms::None;
      names.push_back(confs[i].id);
    }
    gen = std::make_unique<Gen>(names, 0, -1, confs, rb_rows, 0);
    data_size = 4096;
    while (data_size < max_len) data_size <<= 1;
    base = static_cast<uint8_t*>(aligned_alloc(4096, 4096 + data_size));
    memset(base, 0, 4096);
    page = reinterpret_cast<volatile perf_event_mmap_page*>(base);
    page->data_offset = 4096;
    page->data_size = data_size;
    page->time_shift = static_cast<uint16_t>(p.time_shift);
    page->time_mult = p.time_mult;
    page->time_zero = p.time_zero;
    page->cap_user_time = 1;
    page->cap_user_time_zero = 1;
    gen->mmap_base_ = base;
    gen->event_fds_[0] = 3;  // any fd >= 0 makes isOpen() true; never used by consume
    gen->tsc_rem_flags_ = ((((uint64_t)1) << p.time_shift) - 1);  // as mmap_ does, :1193
  }
  ~FakeGen() {
    gen->event_fds_[0] = -1;  // dtor must not close() a fake fd (:939-943)
    gen->mmap_base_ = nullptr;
    gen.reset();
    free(base);
  }
  void setCarry(const hbtref_carry& c, uint32_t n) {
    gen->count_.tstamp = c.valid ? c.tstamp : facebook::hbt::kInvalidTimeStamp;
    for (uint32_t i = 0; i < n; ++i) gen->last_[i] = c.last[i];
  }
  void load(const uint8_t* bytes, size_t len) {
    memcpy(base + 4096, bytes, len);
    page->data_tail = 0;
    page->data_head = len;
  }
  uint8_t* data() { return base + 4096; }
};

void export_counters(const Gen& g, hbtref_counters* c) {
  c->lost = g.getNumRecordLost();
  c->comm = g.getNumRecordComm();
  c->exit = g.getNumRecordExit();
  c->throttle = g.getNumRecordThrottle();
  c->unthrottle = g.getNumRecordUnThrottle();
  c->fork = g.getNumRecordFork();
  c->sample = g.getNumRecordSample();
  c->read = g.getNumRecordRead();
  c->aux = g.getNumRecordAux();
  c->itrace_start = g.getNumRecordItraceStart();
  c->switch_cpu_wide = g.getNumRecordSwitchCpuWide();
  c->num_unknown_count_interval = g.num_unknown_count_interval_;
  c->last_tstamp = g.getLastTimeStamp();
}

}  // namespace

extern "C" {

const char* hbtref_describe(void) {
#ifdef HBTREF_NOLOG
  return "reference hbt code compiled from /root/reference (hbt logging compiled out)";
#else
  return "reference hbt code compiled from /root/reference (as shipped, cerr to null)";
#endif
}

int hbtref_decode_stream(const hbtref_params* p, const hbtref_carry* in,
    hbtref_carry* out, const uint8_t* bytes, uint64_t len, hbtref_counters* ctr,
    uint64_t* rec_off, uint32_t* rec_type, uint32_t* rec_size, uint64_t rec_cap,
    uint64_t* rows, uint64_t rows_cap, uint64_t* bucket_acc) {
  const uint32_t n = p->n_events;
  const uint32_t F = n + 3;
  if (p->pre != 0) return -1;  // IP|TID samples are a spec extension: hbt has no such mode
  silence_begin();
  FakeGen fg(*p, len, 4);
  Gen& g = *fg.gen;
  fg.setCarry(*in, n);
  fg.load(bytes, len);
  hr::Consumer<Gen::RbExtraData> cons(g.getOutput());

  // te/tr of the previous SAMPLE, tracked with the reference's own value type.
  RV prev(n), cur(n);
  prev.t->time_enabled = in->te;
  prev.t->time_running = in->tr;
  for (uint32_t i = 0; i < n; ++i) prev.t->count[i] = in->last[i];

  std::vector<RV> acc;  // one GroupReadValues per bucket, filled with accum()
  std::vector<uint64_t> acc_rows;
  if (bucket_acc) {
    acc.reserve(p->n_buckets);
    for (uint32_t b = 0; b < p->n_buckets; ++b) acc.emplace_back(n);
    acc_rows.assign(p->n_buckets, 0);
  }

  hbtref_counters c;
  memset(&c, 0, sizeof(c));
  std::vector<uint64_t> row(2 + n);
  uint64_t nrec = 0, nrows = 0;
  for (;;) {
    const uint64_t off = fg.page->data_tail;
    ssize_t ret = g.consume(1);
    if (ret == -ENODATA) break;
    if (ret < 0) {
      c.err = ret;
      break;
    }
    const auto* hdr = reinterpret_cast<const perf_event_header*>(fg.data() + off);
    if (rec_off && nrec < rec_cap) {
      rec_off[nrec] = off;
      rec_type[nrec] = hdr->type;
      rec_size[nrec] = hdr->size;
    }
    ++nrec;
    const bool is_sample = hdr->type == PERF_RECORD_SAMPLE;
    if (is_sample) {
      // read_format block {nr, te, tr, value[]} has GroupReadValues::T's layout (:389-397)
      memcpy(cur.t, fg.data() + off + 16, 8 * (3 + n));
    }
    if (g.getNumEntries() > 0) {
      // this record emitted a row; drain it the way TraceMonitor does (TraceMonitor.h:71-105)
      ssize_t r = cons.copy(g.kWriteByteSize, row.data());
      if (r < 0) abort();
      if (rows && nrows < rows_cap) memcpy(rows + nrows * (2 + n), row.data(), 8 * (2 + n));
      ++nrows;
      if (bucket_acc) {
        RV d = cur.diff(prev);  // reference diff(): dTE, dTR (and raw deltas)
        for (uint32_t i = 0; i < n; ++i) {
          if (!((p->tsc_mask >> i) & 1) && d.t->count[i] != row[2 + i]) abort();  // self-check
          d.t->count[i] = row[2 + i];  // the row's (possibly transformed) delta is authoritative
        }
        const uint64_t ts = row[0] + row[1];  // the sample's own time
        if (ts >= p->t0 && (ts - p->t0) / p->bucket_ns < p->n_buckets) {
          const uint64_t b = (ts - p->t0) / p->bucket_ns;
          acc[b].accum(d);
          acc_rows[b] += 1;
        } else {
          c.rows_out_of_range += 1;
        }
      }
    }
    if (is_sample) memcpy(prev.t, cur.t, 8 * (3 + n));
  }
  export_counters(g, &c);
  c.num_records = nrec;
  c.num_rows = nrows;
  c.err_offset = fg.page->data_tail;
  if (ctr) *ctr = c;
  if (out) {
    memset(out, 0, sizeof(*out));
    out->valid = g.count_.tstamp != facebook::hbt::kInvalidTimeStamp;
    out->tstamp = g.count_.tstamp;
    out->te = prev.t->time_enabled;
    out->tr = prev.t->time_running;
    for (uint32_t i = 0; i < n; ++i) out->last[i] = g.last_[i];
  }
  if (bucket_acc) {
    for (uint32_t b = 0; b < p->n_buckets; ++b) {
      uint64_t* a = bucket_acc + (uint64_t)b * F;
      for (uint32_t i = 0; i < n; ++i) a[i] += acc[b].t->count[i];
      a[n] += acc[b].t->time_enabled;
      a[n + 1] += acc[b].t->time_running;
      a[n + 2] += acc_rows[b];
    }
  }
  silence_end();
  return 0;
}

void hbtref_scale_buckets(const hbtref_params* p, const uint64_t* bucket_acc,
    uint64_t* scaled) {
  const uint32_t n = p->n_events, F = n + 3;
  RV rv(n);
  for (uint64_t b = 0; b < p->n_buckets; ++b) {
    const uint64_t* a = bucket_acc + b * F;
    for (uint32_t i = 0; i < n; ++i) rv.t->count[i] = a[i];
    rv.t->time_enabled = a[n];
    rv.t->time_running = a[n + 1];
    for (uint32_t i = 0; i < n; ++i) scaled[b * n + i] = rv.getCount(i);
  }
}

void hbtref_get_count(const uint64_t* count, const uint64_t* te, const uint64_t* tr,
    uint64_t* out, uint64_t n) {
  RV rv(1);
  for (uint64_t i = 0; i < n; ++i) {
    rv.t->count[0] = count[i];
    rv.t->time_enabled = te[i];
    rv.t->time_running = tr[i];
    out[i] = rv.getCount(0);
  }
}

void hbtref_rate(const uint64_t* num, const uint64_t* den, double* out,
    uint8_t* valid, uint64_t n) {
  const auto& red = hp::getRateReducer();
  std::vector<uint64_t> v(2);
  for (uint64_t i = 0; i < n; ++i) {
    v[0] = num[i];
    v[1] = den[i];
    auto r = red(v);
    valid[i] = r.has_value();
    out[i] = r.value_or(0.0);
  }
}

void hbtref_add_reduce(const uint64_t* vals, uint64_t k, double* out, uint64_t n) {
  const auto& red = hp::getAddReducer();
  std::vector<uint64_t> v(k);
  for (uint64_t i = 0; i < n; ++i) {
    for (uint64_t j = 0; j < k; ++j) v[j] = vals[i * k + j];
    out[i] = red(v).value_or(0.0);
  }
}

void hbtref_tsc_to_ktime(const hbtref_params* p, const uint64_t* tsc, uint64_t* out,
    uint64_t n) {
  FakeGen fg(*p, 4096, 4);
  for (uint64_t i = 0; i < n; ++i) out[i] = fg.gen->kernelTimeFromTsc(tsc[i]);
}

void hbtref_ktime_to_tsc(const hbtref_params* p, const uint64_t* kt, uint64_t* out,
    uint64_t n) {
  FakeGen fg(*p, 4096, 4);
  for (uint64_t i = 0; i < n; ++i) out[i] = fg.gen->tscFromKernelTime(kt[i]);
}

void hbtref_per_second_mega(const uint64_t* count, const uint64_t* tr, float* out,
    uint64_t n) {
  // dynolog/src/PerfMonitor.cpp:63-66 (restated: the file needs glog + nlohmann/json)
  for (uint64_t i = 0; i < n; ++i)
    out[i] = static_cast<float>(
        static_cast<double>(count[i]) * 1e3 / static_cast<double>(tr[i]));
}

void hbtref_running_ratio(const uint64_t* te, const uint64_t* tr, float* ratio,
    uint64_t* ratio_int100, uint8_t* has_mux, uint64_t n) {
  RV rv(1);
  for (uint64_t i = 0; i < n; ++i) {
    rv.t->time_enabled = te[i];
    rv.t->time_running = tr[i];
    ratio[i] = rv.getRunningRatio();
    ratio_int100[i] = rv.getRunningRatioInteger<100>();
    has_mux[i] = rv.hasMultiplexing();
  }
}

// Timed CPU path: reference consume(batch) -> rows -> Consumer drain -> bucket
// sums.  One std::thread per group of streams (legal: Sampling generators
// share no state, PerCpuCountSampleGenerator.h:262-272).  The reference's rows
// carry no time_enabled/time_running, so this path sums only delta[] and
// n_rows per bucket (it does LESS work than the GPU path).
double hbtref_run_capture(const hbtref_params* p, uint32_t n_streams,
    const uint8_t* const* bytes, const uint64_t* lens, const hbtref_carry* in,
    uint32_t n_threads, uint32_t batch, uint64_t* bucket_acc, hbtref_counters* ctrs) {
  const uint32_t n = p->n_events, F = n + 3;
  if (n_threads == 0) n_threads = 1;
  if (n_threads > n_streams) n_threads = n_streams;
  silence_begin();
  const size_t rb_rows = (size_t)batch + 16;
  // Set-up (fake mmap pages with the image bytes in place) is not timed.
  std::vector<std::unique_ptr<FakeGen>> gens(n_streams);
  {
    std::vector<std::thread> th;
    std::atomic<uint32_t> next{0};
    for (uint32_t t = 0; t < n_threads; ++t)
      th.emplace_back([&] {
        for (uint32_t s; (s = next.fetch_add(1)) < n_streams;) {
          gens[s] = std::make_unique<FakeGen>(*p, lens[s], rb_rows);
          gens[s]->setCarry(in[s], n);
          gens[s]->load(bytes[s], lens[s]);
        }
      });
    for (auto& t : th) t.join();
  }
  auto t_begin = std::chrono::steady_clock::now();
  {
    std::vector<std::thread> th;
    std::atomic<uint32_t> next{0};
    for (uint32_t t = 0; t < n_threads; ++t)
      th.emplace_back([&] {
        for (uint32_t s; (s = next.fetch_add(1)) < n_streams;) {
          Gen& g = *gens[s]->gen;
          hr::Consumer<Gen::RbExtraData> cons(g.getOutput());
          uint64_t* acc = bucket_acc ? bucket_acc + (uint64_t)s * p->n_buckets * F : nullptr;
          hbtref_counters c;
          memset(&c, 0, sizeof(c));
          for (;;) {
            ssize_t ret = g.consume(batch);
            // drain rows
            if (cons.startTx() >= 0) {
              for (;;) {
                auto [sz, ptr] = cons.readInTx(g.kWriteByteSize);
                if (sz < 0) break;
                const uint64_t* row = static_cast<const uint64_t*>(ptr);
                ++c.num_rows;
                if (acc) {
                  const uint64_t ts = row[0] + row[1];
                  if (ts >= p->t0 && (ts - p->t0) / p->bucket_ns < p->n_buckets) {
                    uint64_t* a = acc + ((ts - p->t0) / p->bucket_ns) * F;
                    for (uint32_t i = 0; i < n; ++i) a[i] += row[2 + i];
                    a[n + 2] += 1;
                  } else {
                    ++c.rows_out_of_range;
                  }
                }
              }
              (void)cons.commitTx();
            }
            if (ret == -ENODATA) break;
            if (ret < 0) {
              c.err = ret;
              break;
            }
            c.num_records += (uint64_t)ret;
          }
          export_counters(g, &c);
          c.err_offset = gens[s]->page->data_tail;
          if (ctrs) ctrs[s] = c;
        }
      });
    for (auto& t : th) t.join();
  }
  auto t_end = std::chrono::steady_clock::now();
  gens.clear();
  silence_end();
  return std::chrono::duration<double>(t_end - t_begin).count();
}

}  // extern "C"

// =====================================================================================
// ContextSwitch-mode decode: the real
// PerfEventsGroup<CpuThreadSwitchGenerator, mode::ContextSwitch>::consume and the handlers of
// hbt/src/perf_event/PerCpuThreadSwitchGenerator.h:219-407 over a fake mmap page.  Harness
// glue: the fake page, draining the output ring (ringbuffer::Consumer) after every consume
// call, exporting PerThreadStats::info.  The third-party `pfs` /proc reader that
// ProcFsHelpers.h includes is replaced by oracle/stubs/pfs/procfs.hpp (cmdline := nullopt).
#include "hbt/src/perf_event/PerCpuThreadSwitchGenerator.h"

#include "hbtcs_api.h"

namespace {
using CsGen = hp::CpuThreadSwitchGenerator;
static_assert(sizeof(hbtcs_event) == sizeof(facebook::hbt::tagstack::Event));

struct FakeCsGen {
  std::unique_ptr<CsGen> gen;
  uint8_t* base = nullptr;
  volatile perf_event_mmap_page* page = nullptr;
  FakeCsGen(uint32_t cpu, uint32_t tid_level, std::shared_ptr<hp::PerThreadStats> stats, size_t len,
            size_t rb_events) {
    gen = std::make_unique<CsGen>(static_cast<facebook::hbt::CpuId>(cpu), -1,
                                  static_cast<facebook::hbt::tagstack::Level>(tid_level), stats, rb_events);
    size_t data_size = 4096;
    while (data_size < len) data_size <<= 1;
    base = static_cast<uint8_t*>(aligned_alloc(4096, 4096 + data_size));
    memset(base, 0, 4096);
    page = reinterpret_cast<volatile perf_event_mmap_page*>(base);
    page->data_offset = 4096;
    page->data_size = data_size;
    gen->mmap_base_ = base;
    gen->event_fds_[0] = 3;  // isOpen(); never used by consume
  }
  ~FakeCsGen() {
    gen->event_fds_[0] = -1;
    gen->mmap_base_ = nullptr;
    gen.reset();
    free(base);
  }
  void load(const uint8_t* bytes, size_t len) {
    memcpy(base + 4096, bytes, len);
    page->data_tail = 0;
    page->data_head = len;
  }
};
}  // namespace

extern "C" int hbtref_cs_run(uint32_t n_streams, const uint8_t* const* bytes, const uint64_t* lens,
    uint32_t tid_level, uint32_t batch, const hbtcs_state* in, hbtcs_state* out,
    hbtcs_event* events, const uint64_t* ev_off, const uint64_t* ev_cap, uint64_t* n_events,
    hbtref_counters* ctrs, hbtcs_thread* threads, uint64_t thread_cap, uint64_t* n_threads,
    hbtcs_name* names, uint64_t name_cap, uint64_t* n_names) {
  silence_begin();
  auto stats = std::make_shared<hp::PerThreadStats>();
  if (batch == 0) batch = 1;
  for (uint32_t s = 0; s < n_streams; ++s) {
    // every record writes at most two events (a pending WriteErrorsEnd and its own)
    FakeCsGen fg(s, tid_level, stats, lens[s], 2ull * batch + 16);
    CsGen& g = *fg.gen;
    if (in) {
      g.last_tstamp_ = in[s].last_tstamp;
      g.in_write_errors_ = in[s].in_write_errors != 0;
    }
    fg.load(bytes[s], lens[s]);
    hr::Consumer<CsGen::RbExtraData> cons(g.getOutput());
    hbtref_counters c;
    memset(&c, 0, sizeof(c));
    uint64_t nev = 0;
    for (;;) {
      const uint64_t tail0 = fg.page->data_tail;
      ssize_t ret = g.consume(batch);
      // consume() returns -ENODATA instead of the count when it drains the ring (:1574-1588),
      // so the records it took are counted here from how far data_tail moved (glue)
      for (uint64_t t = tail0; t < fg.page->data_tail;) {
        t += reinterpret_cast<const perf_event_header*>(fg.base + 4096 + t)->size;
        c.num_records += 1;
      }
      while (g.getNumEntries() > 0) {
        hbtcs_event ev;
        if (cons.copy(sizeof(ev), &ev) < 0) abort();
        if (events && nev < ev_cap[s]) events[ev_off[s] + nev] = ev;
        ++nev;
      }
      if (ret == -ENODATA) break;
      if (ret < 0) {
        c.err = ret;
        break;
      }
    }
    c.lost = g.getNumRecordLost();
    c.comm = g.getNumRecordComm();
    c.exit = g.getNumRecordExit();
    c.throttle = g.getNumRecordThrottle();
    c.unthrottle = g.getNumRecordUnThrottle();
    c.fork = g.getNumRecordFork();
    c.sample = g.getNumRecordSample();
    c.read = g.getNumRecordRead();
    c.aux = g.getNumRecordAux();
    c.itrace_start = g.getNumRecordItraceStart();
    c.switch_cpu_wide = g.getNumRecordSwitchCpuWide();
    c.num_rows = nev;
    c.err_offset = fg.page->data_tail;
    c.last_tstamp = g.getLastTimeStamp();
    if (ctrs) ctrs[s] = c;
    if (n_events) n_events[s] = nev;
    if (out) {
      out[s].last_tstamp = g.last_tstamp_;
      out[s].in_write_errors = g.in_write_errors_ ? 1 : 0;
      out[s].pad = 0;
    }
  }
  // threads by vid (vids are 1..next_vid-1)
  uint64_t nt = 0, nn = 0;
  for (facebook::hbt::tagstack::Tag vid = 1; vid < stats->next_vid; ++vid) {
    auto it = stats->info.find(vid);
    if (it == stats->info.end()) continue;
    const hp::ThreadInfo& ti = it->second;
    uint32_t cnt = 0;
    for (const auto& p : ti.names.getPoints()) {
      if (names && nn < name_cap) {
        hbtcs_name& o = names[nn];
        memset(&o, 0, sizeof(o));
        o.vid = vid;
        o.is_exec = p.value.is_exec;
        o.len = static_cast<uint8_t>(std::min<size_t>(p.value.new_name.size(), 16));
        o.tstamp = p.tstamp;
        memcpy(o.name, p.value.new_name.data(), o.len);
      }
      ++nn;
      ++cnt;
    }
    if (threads && nt < thread_cap) {
      hbtcs_thread& o = threads[nt];
      o.vid = vid;
      o.ppid = ti.ppid;
      o.ptid = ti.ptid;
      o.pid = ti.pid;
      o.tid = ti.tid;
      o.n_names = cnt;
      o.fork_ts = ti.fork_timestamp;
      o.end_ts = ti.end_timestamp;
    }
    ++nt;
  }
  if (n_threads) *n_threads = nt;
  if (n_names) *n_names = nn;
  silence_end();
  return 0;
}

// =====================================================================================
// Series reductions (sink side): the real
// facebook::dynolog::MetricSeries<uint64_t> (dynolog/src/metric_frame/MetricSeries.h).
// Glue: filling the series with addSample, mapping the template arguments.
#include <string>

#include "dynolog/src/metric_frame/MetricSeries.h"
#include "hbtseries_api.h"

extern "C" int hbtref_series_reduce(const uint64_t* series, uint64_t n_series, uint64_t len,
    float percentage, int64_t period_ns, int64_t duration_ns, double* avg_f64, uint64_t* avg_u64,
    uint64_t* percentile, uint64_t* diff, double* rate_f64, uint64_t* rate_u64) {
  using facebook::dynolog::MetricSeries;
  if (len == 0 || percentage < 0.f || percentage > 1.f || period_ns <= 0 || duration_ns <= 0) return -1;
  const std::chrono::steady_clock::duration period{period_ns}, duration{duration_ns};
  for (uint64_t s = 0; s < n_series; ++s) {
    MetricSeries<uint64_t> ms(len, "series", "");
    for (uint64_t i = 0; i < len; ++i) ms.addSample(series[s * len + i]);
    if (avg_f64) avg_f64[s] = ms.avg<double>();
    if (avg_u64) avg_u64[s] = ms.avg<uint64_t>();
    if (percentile) percentile[s] = ms.percentile(percentage);
    if (diff) diff[s] = ms.diff();
    if (rate_f64) rate_f64[s] = ms.rate<double>(period, duration);
    if (rate_u64) rate_u64[s] = ms.rate<uint64_t>(period, duration);
  }
  return 0;
}
