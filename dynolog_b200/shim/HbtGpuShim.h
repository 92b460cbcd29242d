// HbtGpuShim.h -- header-only C++ shim over the C ABI (include/hbtgpu.h) that keeps the
// reference's class shapes for this path, so that code written against hbt reads the same:
//
//   reference (hbt/src/perf_event/...)                      shim (namespace hbtgpu_shim)
//   DataSlice / RbDataSlices      PerfEventsGroup.h:45-62   DataSlice / RbDataSlices
//   CpuCountSampleGenerator       PerCpuCountSampleGenerator.h:27-231
//     consume(max) -> ssize_t       PerfEventsGroup.h:1394      CpuCountSampleGenerator::consume
//     getNumRecord{Lost,...}()      :791-824                    same names
//     getLastTimeStamp()            gen :209-211                same
//     getNumEntries()               gen :86-89                  same (rows available)
//   PerCpuCountSampleGenerator    gen :234-347              PerCpuCountSampleGenerator
//     getCpuGenerator(cpu)          PerCpuBase.h                same
//     accumUntil(batch, stop_ts)    PerCpuSampleGeneratorBase.h:69-94   same (batch and stop_ts honoured record by record)
//   GroupReadValues               PerfEventsGroup.h:386-586  GroupReadValues (arithmetic through hbtgpu_math_*)
//   dynolog::Logger               dynolog/src/Logger.h:26-48  Logger + logSeries (MetricSeries reductions of the buckets)
//   TraceMonitor::applyToCountSamplesAndConsume  mon/TraceMonitor.h:35-108   applyToCountSamples
//   CpuThreadSwitchGenerator      PerCpuThreadSwitchGenerator.h:168-446   CpuThreadSwitchGenerator
//     consume / getNumRecord* / getLastTimeStamp / getNumEntries         same names
//   PerCpuThreadSwitchGenerator   :449-489                  PerCpuThreadSwitchGenerator
//     onCpusThreadSwitchBufferRead :480-484                   onCpuThreadSwitchBufferRead(cpu, slices)
//     getThreadsInfo()             :476-478                   same (ThreadInfo without /proc strings)
//   tagstack::Event               tagstack/Event.h:28-56    Event (= hbtgpu_cs_event)
//   tagstack::slice::Slice        tagstack/Slicer.h:30-84   Slice (= hbtgpu_slice) + makeSlice / tstampEnd
//   tagstack::PerIntervalSlices   tagstack/IntervalSlicer.h:17-82   PerIntervalSlices
//   tagstack::IntervalSlicer      tagstack/IntervalSlicer.h:92-238  IntervalSlicer
//     write / registerConsumerMark / getMark / numActiveSlices / numProcessedSlices /
//     discardConsumed / makeIntervalIntersectingSlices / consumeSlices          same names
//   TraceMonitor::releaseCountData_ row loop  mon/TraceMonitor.cpp:52-96   releaseCountRows (all CPUs, one GPU call)
//   tagstack::slice::Slicer       tagstack/Slicer.h:302-869  Slicer (a window of level-0 events per call)
//     getStats / getTagStackStatsDump / getLastTimeStamp / getSlices         same names
//
// Error conventions follow the reference: the hot path returns negative errno in ssize_t
// (-ENODATA drained, -EPERM bad record / time regression), argument and environment errors
// throw (std::invalid_argument like HBT_ARG_CHECK, std::runtime_error like HBT_THROW_ASSERT).
// The work itself happens on the GPU behind hbtgpu_run; there is no CPU path in here.
#pragma once

#include <errno.h>
#include <sys/types.h>

#include <algorithm>
#include <cstdint>
#include <cstring>
#include <deque>
#include <functional>
#include <map>
#include <memory>
#include <optional>
#include <stdexcept>
#include <string>
#include <vector>

extern "C" {
#include "hbtgpu.h"
}

namespace hbtgpu_shim {

using CpuId = int;
using TimeStamp = uint64_t;
constexpr TimeStamp kInvalidTimeStamp = ~0ull;  // hbt/src/common/System.h:106

struct DataSlice {
  DataSlice(const void* from, size_t size) : from{from}, size{size} {}
  const void* from;
  size_t size;
};

struct RbDataSlices {
  RbDataSlices(const void* p1, size_t s1) : first{p1, s1}, second{std::nullopt} {}
  RbDataSlices(const void* p1, size_t s1, const void* p2, size_t s2)
      : first{p1, s1}, second{DataSlice{p2, s2}} {}
  DataSlice first;
  std::optional<DataSlice> second;
};

struct TscConversionParams {  // hbt/src/common/System.h:184-188
  uint16_t time_shift = 0;
  uint32_t time_mult = 1;
  uint64_t time_zero = 0;
};

class PerCpuCountSampleGenerator;

// Per-CPU view; same accessor names as the reference's generator.  The counters are lifetime
// counters, as in the reference (PerfEventsGroup.h:887-897): they grow by what consume() walks.
class CpuCountSampleGenerator {
 public:
  /// PerfEventsGroup::consume (:1394-1589): walks up to max_num_records records of what was
  /// handed over for this CPU.  Same return convention (:1572-1588): the number of records, or
  /// -ENODATA when the ring drained before max_num_records, or the errno of the record the walk
  /// stalls on (-EPERM).  The GPU has decoded the whole window by then; this releases its
  /// results record by record, so rows and counters become visible exactly when the reference
  /// makes them visible.
  ssize_t consume(unsigned max_num_records);

  uint64_t getNumRecordLost() const { return c_.lost; }
  uint64_t getNumRecordComm() const { return c_.comm; }
  uint64_t getNumRecordExit() const { return c_.exit; }
  uint64_t getNumRecordThrottle() const { return c_.throttle; }
  uint64_t getNumRecordUnThrottle() const { return c_.unthrottle; }
  uint64_t getNumRecordFork() const { return c_.fork; }
  uint64_t getNumRecordSample() const { return c_.sample; }
  uint64_t getNumRecordRead() const { return c_.read; }
  uint64_t getNumRecordAux() const { return c_.aux; }
  uint64_t getNumRecordItraceStart() const { return c_.itrace_start; }
  uint64_t getNumRecordSwitchCpuWide() const { return c_.switch_cpu_wide; }
  uint64_t getNumUnknownCountInterval() const { return c_.num_unknown_count_interval; }
  /// count_.tstamp (gen :209-211): kInvalidTimeStamp after a LOST or a time regression
  TimeStamp getLastTimeStamp() const { return tstamp_; }
  /// rows in the output ring: released by consume(), not yet taken by applyToCountSamples
  size_t getNumEntries() const { return out_.size(); }
  CpuId getCpu() const { return cpu_; }

 private:
  friend class PerCpuCountSampleGenerator;
  CpuCountSampleGenerator(PerCpuCountSampleGenerator* owner, CpuId cpu) : owner_{owner}, cpu_{cpu} {}
  PerCpuCountSampleGenerator* owner_;
  CpuId cpu_;
  hbtgpu_counters c_{};                 // lifetime
  TimeStamp tstamp_ = kInvalidTimeStamp;  // generator state behind the last consumed record
  // the current window: the unconsumed tail of earlier images followed by the new ones
  std::vector<uint8_t> bytes_;          // host copy (timestamps / values at record offsets)
  size_t in_arena_ = 0;                 // bytes_[0, in_arena_) were decoded by the last run
  hbtgpu_carry start_{};                // generator state in front of bytes_[0]
  std::vector<uint64_t> rec_off_;       // record table of the decoded window
  std::vector<uint32_t> rec_type_, rec_size_;
  std::vector<uint64_t> rows_;          // its rows, n + 2 words each
  size_t rec_cur_ = 0, row_cur_ = 0;    // consumed so far
  size_t last_sample_ = SIZE_MAX;       // record index of the last consumed SAMPLE of this window
  int64_t err_ = 0;                     // the window's walk stalls behind the last record of the table
  uint64_t err_off_ = 0;
  std::deque<std::vector<uint64_t>> out_;
};

class PerCpuCountSampleGenerator {
 public:
  struct Options {
    int device = 0;
    uint32_t num_cpus = 1;
    std::vector<bool> tsc_to_kernel_time;  // per event: EventValueTransforms::TscToKernelTime
    TscConversionParams tsc;
    TimeStamp t0 = 0;
    uint64_t bucket_ns = 1'000'000;
    uint32_t n_buckets = 1;
    size_t arena_bytes = 64u << 20;
    std::vector<hbtgpu_metric> metrics;
  };

  PerCpuCountSampleGenerator(const std::vector<std::string>& event_names, const Options& opt)
      : event_names_{event_names}, opt_{opt} {
    if (event_names.empty() || event_names.size() > HBTGPU_MAX_EVENTS)
      throw std::invalid_argument("number of events must be in [1, 64]");  // HBT_ARG_CHECK_LE, gen :76
    hbtgpu_config cfg{};
    cfg.device = opt.device;
    cfg.n_events = static_cast<uint32_t>(event_names.size());
    for (size_t i = 0; i < opt.tsc_to_kernel_time.size() && i < event_names.size(); ++i)
      if (opt.tsc_to_kernel_time[i]) cfg.tsc_mask |= 1ull << i;
    tsc_mask_ = cfg.tsc_mask;
    cfg.time_shift = opt.tsc.time_shift;
    cfg.time_mult = opt.tsc.time_mult;
    cfg.time_zero = opt.tsc.time_zero;
    cfg.max_cpus = opt.num_cpus;
    cfg.n_buckets = opt.n_buckets;
    cfg.t0 = opt.t0;
    cfg.bucket_ns = opt.bucket_ns;
    cfg.arena_bytes = opt.arena_bytes;
    cfg.n_metrics = static_cast<uint32_t>(opt.metrics.size());
    if (cfg.n_metrics > HBTGPU_MAX_METRICS) throw std::invalid_argument("too many metrics");
    for (uint32_t m = 0; m < cfg.n_metrics; ++m) cfg.metrics[m] = opt.metrics[m];
    int rc = hbtgpu_create(&cfg, &ctx_);
    if (rc == -EINVAL) throw std::invalid_argument(hbtgpu_last_error(nullptr));
    if (rc != 0) throw std::runtime_error(hbtgpu_last_error(nullptr));
    for (uint32_t c = 0; c < opt.num_cpus; ++c) gens_.push_back(CpuCountSampleGenerator(this, static_cast<CpuId>(c)));
  }
  ~PerCpuCountSampleGenerator() { hbtgpu_destroy(ctx_); }
  PerCpuCountSampleGenerator(const PerCpuCountSampleGenerator&) = delete;
  PerCpuCountSampleGenerator& operator=(const PerCpuCountSampleGenerator&) = delete;

  size_t getNumEvents() const { return event_names_.size(); }
  const std::vector<std::string>& getEventNicknames() const { return event_names_; }

  /// State enableImpl() leaves (gen :91-106): interval starts at `now`, last_ = 0.
  void enable(CpuId cpu, TimeStamp now) {
    auto& g = gens_.at(static_cast<size_t>(cpu));
    g.start_ = hbtgpu_carry{};
    g.start_.valid = 1;
    g.start_.tstamp = now;
    g.tstamp_ = now;
    dirty_ = true;
  }

  /// The callback to hand to PerfEventsGroup::onCpuDataBufferRead (:1592-1637), consume = true:
  /// the bytes are copied, the ring may move on.  They join the CPU's stream behind whatever an
  /// earlier accumUntil left unconsumed.
  void onCpuDataBufferRead(CpuId cpu, const RbDataSlices& s) {
    auto& g = gens_.at(static_cast<size_t>(cpu));
    const auto* p1 = static_cast<const uint8_t*>(s.first.from);
    g.bytes_.insert(g.bytes_.end(), p1, p1 + s.first.size);
    if (s.second) {
      const auto* p2 = static_cast<const uint8_t*>(s.second->from);
      g.bytes_.insert(g.bytes_.end(), p2, p2 + s.second->size);
    }
    dirty_ = true;
  }

  CpuCountSampleGenerator& getCpuGenerator(CpuId cpu) { return gens_.at(static_cast<size_t>(cpu)); }

  /// PerCpuSampleGeneratorBase::accumUntil (:69-94), statement by statement: per CPU, consume
  /// batches while the generator's last timestamp is <= stop_ts; -ENODATA ends the CPU quietly,
  /// any other error throws (HBT_THROW_ASSERT_IF, :76-79).  Returns the number of CPUs that
  /// finished without an error.  (As in the reference, a generator whose state is invalid --
  /// after a LOST or a time regression -- reports kInvalidTimeStamp and is not consumed.)
  unsigned accumUntil(unsigned batch_size, TimeStamp stop_ts) {
    if (batch_size == 0) throw std::invalid_argument("batch_size must be > 0");  // HBT_ARG_CHECK_GT
    unsigned num_done = 0;
    for (auto& g : gens_) {
      ssize_t ret = 0;
      while (g.getLastTimeStamp() <= stop_ts) {
        ret = g.consume(batch_size);
        if (ret < 0 && ret != -ENODATA && ret != -EAGAIN)
          throw std::runtime_error("Unexpected error processing batch: " + std::to_string(ret));
        if (ret == -ENODATA) {
          ret = 0;
          break;
        }
      }
      if (ret >= 0) ++num_done;
    }
    return num_done;
  }

  /// TraceMonitor::applyToCountSamplesAndConsume (mon/TraceMonitor.h:35-108):
  /// func(cpu, count_idx, nicknames, tstamp, duration, counts, n) for every row in the output
  /// rings, which are emptied.
  using CountFunc = std::function<void(CpuId, size_t, const std::vector<std::string>&, TimeStamp,
                                       TimeStamp, const uint64_t*, size_t)>;
  void applyToCountSamples(const CountFunc& func) {
    const size_t n = getNumEvents();
    for (auto& g : gens_) {
      size_t i = 0;
      while (!g.out_.empty()) {
        const std::vector<uint64_t>& r = g.out_.front();
        func(g.cpu_, i++, event_names_, r[0], r[1], r.data() + 2, n);
        g.out_.pop_front();
      }
    }
  }

  /// Bucketed aggregates (scaled counts, metrics, validity) of the window as the GPU decoded it
  /// (everything handed over and not yet consumed by an earlier window; stop_ts does not apply).
  void getBuckets(std::vector<uint64_t>* acc, std::vector<uint64_t>* scaled, std::vector<double>* metric,
                  std::vector<uint32_t>* valid) {
    decode();
    const size_t cells = static_cast<size_t>(opt_.num_cpus) * opt_.n_buckets, n = getNumEvents();
    if (acc) acc->resize(cells * (n + 3));
    if (scaled) scaled->resize(cells * n);
    if (metric) metric->resize(cells * opt_.metrics.size());
    if (valid) valid->resize(cells);
    check(hbtgpu_get_buckets_host(ctx_, acc ? acc->data() : nullptr, scaled ? scaled->data() : nullptr,
                                  metric && !metric->empty() ? metric->data() : nullptr,
                                  valid ? valid->data() : nullptr));
  }

  hbtgpu_ctx* raw() { return ctx_; }

 private:
  friend class CpuCountSampleGenerator;
  void check(int rc) {
    if (rc == -EINVAL) throw std::invalid_argument(hbtgpu_last_error(ctx_));
    if (rc != 0) throw std::runtime_error(hbtgpu_last_error(ctx_));
  }
  static uint64_t word(const std::vector<uint8_t>& b, size_t off) {
    uint64_t v;
    std::memcpy(&v, b.data() + off, 8);
    return v;
  }

  // A new window on the GPU: every CPU's stream restarts at its first unconsumed record, from
  // the generator state in front of it; the arena holds nothing older (no O(history) re-decode).
  void decode() {
    if (!dirty_) return;
    const size_t n = getNumEvents();
    std::vector<uint64_t> raw_vals, xf_vals;  // values of carried samples that need the TSC transform
    for (auto& g : gens_) {
      if (g.rec_cur_ > 0 || g.in_arena_ > 0) {
        // the carry behind the consumed records: time / validity from the live state, te / tr /
        // last_[] from the last consumed SAMPLE of the window (else unchanged)
        if (g.last_sample_ != SIZE_MAX) {
          const size_t off = static_cast<size_t>(g.rec_off_[g.last_sample_]);
          g.start_.te = word(g.bytes_, off + 24);
          g.start_.tr = word(g.bytes_, off + 32);
          for (size_t e = 0; e < n; ++e) {
            g.start_.last[e] = word(g.bytes_, off + 40 + 8 * e);
            if ((tsc_mask_ >> e) & 1) raw_vals.push_back(g.start_.last[e]);
          }
        }
        const size_t cut = g.rec_cur_ < g.rec_off_.size() ? static_cast<size_t>(g.rec_off_[g.rec_cur_])
                                                          : (g.err_ ? static_cast<size_t>(g.err_off_) : g.in_arena_);
        g.bytes_.erase(g.bytes_.begin(), g.bytes_.begin() + static_cast<std::ptrdiff_t>(cut));
      }
      g.start_.valid = g.tstamp_ != kInvalidTimeStamp;
      g.start_.tstamp = g.tstamp_;
    }
    if (!raw_vals.empty()) {  // carried values are stored transformed (gen :145-150)
      xf_vals.resize(raw_vals.size());
      check(hbtgpu_math_tsc_to_ktime(ctx_, raw_vals.data(), xf_vals.data(), raw_vals.size()));
      size_t k = 0;
      for (auto& g : gens_)
        if ((g.rec_cur_ > 0 || g.in_arena_ > 0) && g.last_sample_ != SIZE_MAX)
          for (size_t e = 0; e < n; ++e)
            if ((tsc_mask_ >> e) & 1) g.start_.last[e] = xf_vals[k++];
    }
    check(hbtgpu_reset(ctx_, 0));
    for (auto& g : gens_) {
      check(hbtgpu_set_carry(ctx_, static_cast<uint32_t>(g.cpu_), &g.start_));
      if (!g.bytes_.empty())
        check(hbtgpu_submit_image(ctx_, static_cast<uint32_t>(g.cpu_), g.bytes_.data(), g.bytes_.size(), nullptr, 0));
      g.in_arena_ = g.bytes_.size();
    }
    check(hbtgpu_run(ctx_));
    check(hbtgpu_sync(ctx_));
    for (auto& g : gens_) {
      hbtgpu_counters c{};
      check(hbtgpu_get_counters(ctx_, static_cast<uint32_t>(g.cpu_), &c));
      g.err_ = c.err;
      g.err_off_ = c.err_offset;
      size_t nr = 0, rows = 0;
      check(hbtgpu_get_records(ctx_, static_cast<uint32_t>(g.cpu_), nullptr, nullptr, nullptr, 0, &nr));
      g.rec_off_.resize(nr);
      g.rec_type_.resize(nr);
      g.rec_size_.resize(nr);
      if (nr)
        check(hbtgpu_get_records(ctx_, static_cast<uint32_t>(g.cpu_), g.rec_off_.data(), g.rec_type_.data(),
                                 g.rec_size_.data(), nr, &nr));
      check(hbtgpu_get_rows(ctx_, static_cast<uint32_t>(g.cpu_), nullptr, 0, &rows));
      g.rows_.resize(rows * (n + 2));
      if (rows) check(hbtgpu_get_rows(ctx_, static_cast<uint32_t>(g.cpu_), g.rows_.data(), rows, &rows));
      g.rec_cur_ = g.row_cur_ = 0;
      g.last_sample_ = SIZE_MAX;
    }
    dirty_ = false;
  }

  std::vector<std::string> event_names_;
  Options opt_;
  uint64_t tsc_mask_ = 0;
  hbtgpu_ctx* ctx_ = nullptr;
  std::vector<CpuCountSampleGenerator> gens_;
  bool dirty_ = true;
};

inline ssize_t CpuCountSampleGenerator::consume(unsigned max_num_records) {
  owner_->decode();
  const size_t n = owner_->getNumEvents();
  unsigned num_records = 0;
  while (num_records < max_num_records && rec_cur_ < rec_off_.size()) {
    const uint32_t type = rec_type_[rec_cur_];
    const size_t off = static_cast<size_t>(rec_off_[rec_cur_]);
    switch (type) {
      case 9: {  // PERF_RECORD_SAMPLE -> handleRecordSample (gen :132-202)
        ++c_.sample;
        if (tstamp_ == kInvalidTimeStamp) {
          ++c_.num_unknown_count_interval;
        } else {
          const uint64_t* r = &rows_[row_cur_ * (n + 2)];
          out_.emplace_back(r, r + n + 2);
          ++row_cur_;
          ++c_.num_rows;
        }
        tstamp_ = PerCpuCountSampleGenerator::word(bytes_, off + 8);
        last_sample_ = rec_cur_;
        break;
      }
      case 2:  // PERF_RECORD_LOST (:1440-1450, gen :110-124)
        c_.lost += PerCpuCountSampleGenerator::word(bytes_, off + 16);
        tstamp_ = kInvalidTimeStamp;
        break;
      case 5: ++c_.throttle; break;
      case 6: ++c_.unthrottle; break;
      case 8: ++c_.read; break;
      default: break;  // skipped by size (:1565-1567)
    }
    ++c_.num_records;
    ++num_records;
    ++rec_cur_;
  }
  if (rec_cur_ == rec_off_.size() && num_records < max_num_records) {
    if (err_ < 0) {
      // the walk stalls on the next record.  A well-formed SAMPLE there means time went back:
      // the reference invalidates count_.tstamp and returns -EPERM (gen :156-163); the next
      // consume() meets the same SAMPLE with an invalid state and re-bases on it.
      const size_t off = static_cast<size_t>(err_off_);
      if (off + 24 <= bytes_.size()) {
        const uint64_t h = PerCpuCountSampleGenerator::word(bytes_, off);
        const bool sample_ok = static_cast<uint32_t>(h) == 9 && (h >> 48) == 40 + 8 * n &&
                               PerCpuCountSampleGenerator::word(bytes_, off + 16) == n;
        if (sample_ok && tstamp_ != kInvalidTimeStamp) {
          tstamp_ = kInvalidTimeStamp;
          owner_->dirty_ = true;  // decode again from the stalled record, state invalid
        }
      }
      c_.err = err_;
      c_.err_offset = err_off_;
      return static_cast<ssize_t>(err_);
    }
    return -ENODATA;  // drained (:1574-1576)
  }
  return static_cast<ssize_t>(num_records);
}

/// GroupReadValues<mode::Counting> (PerfEventsGroup.h:386-586) with its arithmetic computed by the
/// library (hbtgpu_math_*, the same device functions the bucket path uses).  accum / diff are
/// wrapping integer adds (:548-573).
class GroupReadValues {
 public:
  GroupReadValues(hbtgpu_ctx* ctx, size_t num_events) : ctx_{ctx}, count_(num_events, 0) {}
  size_t getNumEvents() const { return count_.size(); }
  void setTimeEnabled(uint64_t v) { time_enabled_ = v; }
  void setTimeRunning(uint64_t v) { time_running_ = v; }
  void setCount(size_t i, uint64_t v) { count_.at(i) = v; }
  uint64_t getTimeEnabled() const { return time_enabled_; }
  uint64_t getTimeRunning() const { return time_running_; }
  uint64_t getRawCount(size_t i) const { return count_.at(i); }
  void zero() {
    time_enabled_ = time_running_ = 0;
    std::fill(count_.begin(), count_.end(), 0);
  }
  /// :467-481
  uint64_t getCount(size_t i) const {
    uint64_t out = 0;
    ck(hbtgpu_math_get_count(ctx_, &count_.at(i), &time_enabled_, &time_running_, &out, 1));
    return out;
  }
  std::vector<uint64_t> getCounts() const {
    const size_t n = count_.size();
    std::vector<uint64_t> te(n, time_enabled_), tr(n, time_running_), out(n);
    ck(hbtgpu_math_get_count(ctx_, count_.data(), te.data(), tr.data(), out.data(), n));
    return out;
  }
  /// :513-522 (returns te / tr, as the reference does despite its comment)
  float getRunningRatio() const {
    float r = 0;
    uint64_t ri = 0;
    uint8_t m = 0;
    ck(hbtgpu_math_running_ratio(ctx_, &time_enabled_, &time_running_, &r, &ri, &m, 1));
    return r;
  }
  /// getRunningRatioInteger<100> (:527-535)
  uint64_t getRunningRatioInteger100() const {
    float r = 0;
    uint64_t ri = 0;
    uint8_t m = 0;
    ck(hbtgpu_math_running_ratio(ctx_, &time_enabled_, &time_running_, &r, &ri, &m, 1));
    return ri;
  }
  bool hasMultiplexing() const { return time_enabled_ != 0 && time_running_ != time_enabled_; }  // :538-540
  void accum(const GroupReadValues& o) {  // :548-556
    time_enabled_ += o.time_enabled_;
    time_running_ += o.time_running_;
    for (size_t i = 0; i < count_.size(); ++i) count_[i] += o.count_.at(i);
  }
  void diff(const GroupReadValues& o) {  // :558-573
    time_enabled_ -= o.time_enabled_;
    time_running_ -= o.time_running_;
    for (size_t i = 0; i < count_.size(); ++i) count_[i] -= o.count_.at(i);
  }
  /// getReducedCount (:502-511) with getRateReducer (BuiltinMetrics.cpp:102-118)
  std::optional<double> getRateReducedCount() const {
    if (count_.size() == 1) return static_cast<double>(getCount(0));
    if (count_.size() != 2) return std::nullopt;
    const std::vector<uint64_t> v = getCounts();
    double out = 0;
    uint8_t valid = 0;
    ck(hbtgpu_math_rate(ctx_, &v[0], &v[1], &out, &valid, 1));
    return valid ? std::optional<double>(out) : std::nullopt;
  }
  /// ... with getAddReducer (BuiltinMetrics.cpp:88-97, int accumulator)
  std::optional<double> getAddReducedCount() const {
    if (count_.size() == 1) return static_cast<double>(getCount(0));
    const std::vector<uint64_t> v = getCounts();
    double out = 0;
    ck(hbtgpu_math_add_reduce(ctx_, v.data(), v.size(), &out, 1));
    return out;
  }

 private:
  void ck(int rc) const {
    if (rc != 0) throw std::runtime_error(hbtgpu_last_error(ctx_));
  }
  hbtgpu_ctx* ctx_;
  uint64_t time_enabled_ = 0, time_running_ = 0;
  std::vector<uint64_t> count_;
};

/// dynolog::Logger (dynolog/src/Logger.h:26-48): the interface a sink implements.
class Logger {
 public:
  virtual ~Logger() = default;
  virtual void logInt(const std::string& key, int64_t val) = 0;
  virtual void logFloat(const std::string& key, float val) = 0;
  virtual void logUint(const std::string& key, uint64_t val) = 0;
  virtual void logStr(const std::string& key, const std::string& val) = 0;
  virtual void finalize() = 0;
};

/// The bucket series of the current window, reduced on the GPU the way dynolog's MetricSeries
/// reduces a metric before it is logged (avg, percentile, diff, rate) and published through a
/// Logger: one entry per CPU, keys "<nickname>_avg.cpu<N>" etc. (hbtgpu_log_series).
inline void logSeries(PerCpuCountSampleGenerator& gen, Logger& logger, uint32_t first_bucket, uint32_t count,
                      float percentage, int64_t period_ns, int64_t duration_ns) {
  gen.getBuckets(nullptr, nullptr, nullptr, nullptr);  // decodes what was handed over, if it has not been yet
  hbtgpu_logger lg{};
  lg.user = &logger;
  lg.log_uint = [](void* u, const char* key, uint64_t v) { static_cast<Logger*>(u)->logUint(key, v); };
  lg.log_float = [](void* u, const char* key, float v) { static_cast<Logger*>(u)->logFloat(key, v); };
  lg.finalize = [](void* u) { static_cast<Logger*>(u)->finalize(); };
  std::vector<const char*> names;
  for (const auto& nm : gen.getEventNicknames()) names.push_back(nm.c_str());
  const int rc = hbtgpu_log_series(gen.raw(), &lg, names.data(), first_bucket, count, percentage, period_ns, duration_ns);
  if (rc == -EINVAL) throw std::invalid_argument(hbtgpu_last_error(gen.raw()));
  if (rc != 0) throw std::runtime_error(hbtgpu_last_error(gen.raw()));
}

// ---- ContextSwitch mode ---------------------------------------------------------------------
using Event = hbtgpu_cs_event;  // layout of tagstack::Event

struct ThreadNameChange {  // PerCpuThreadSwitchGenerator.h:24-34
  TimeStamp tstamp;
  std::string new_name;
  bool is_exec;
};
struct ThreadInfo {  // :36-47 (cmdline and /proc comm are the caller's lookups)
  uint32_t vid;
  pid_t ppid, ptid, pid, tid;
  TimeStamp fork_timestamp, end_timestamp;
  std::vector<ThreadNameChange> names;
};

class PerCpuThreadSwitchGenerator;

class CpuThreadSwitchGenerator {
 public:
  ssize_t consume(unsigned max_num_records);
  uint64_t getNumRecordLost() const { return c_.lost; }
  uint64_t getNumRecordComm() const { return c_.comm; }
  uint64_t getNumRecordExit() const { return c_.exit; }
  uint64_t getNumRecordFork() const { return c_.fork; }
  uint64_t getNumRecordThrottle() const { return c_.throttle; }
  uint64_t getNumRecordUnThrottle() const { return c_.unthrottle; }
  uint64_t getNumRecordSwitchCpuWide() const { return c_.switch_cpu_wide; }
  TimeStamp getLastTimeStamp() const { return c_.last_tstamp; }
  size_t getNumEntries() const { return events_.size(); }
  // the events the reference would have written to its output ring, tags already virtual ids
  const std::vector<Event>& events() const { return events_; }

 private:
  friend class PerCpuThreadSwitchGenerator;
  PerCpuThreadSwitchGenerator* owner_ = nullptr;
  CpuId cpu_ = 0;
  hbtgpu_counters c_{};
  std::vector<Event> events_;
};

class PerCpuThreadSwitchGenerator {
 public:
  PerCpuThreadSwitchGenerator(int num_cpus, uint8_t tid_level, uint64_t arena_bytes = 64ull << 20, int device = 0) {
    if (num_cpus <= 0) throw std::invalid_argument("num_cpus must be positive");
    hbtgpu_cs_config cfg{};
    cfg.device = device;
    cfg.max_cpus = static_cast<uint32_t>(num_cpus);
    cfg.tid_level = tid_level;
    cfg.arena_bytes = arena_bytes;
    ctx_ = hbtgpu_cs_create(&cfg);
    if (!ctx_) throw std::runtime_error(hbtgpu_cs_last_error(nullptr));
    threads_ = hbtgpu_cs_threads_create();
    gens_.resize(num_cpus);
    for (int c = 0; c < num_cpus; ++c) {
      gens_[c].owner_ = this;
      gens_[c].cpu_ = c;
    }
  }
  ~PerCpuThreadSwitchGenerator() {
    hbtgpu_cs_threads_destroy(threads_);
    hbtgpu_cs_destroy(ctx_);
  }
  PerCpuThreadSwitchGenerator(const PerCpuThreadSwitchGenerator&) = delete;
  PerCpuThreadSwitchGenerator& operator=(const PerCpuThreadSwitchGenerator&) = delete;

  CpuThreadSwitchGenerator& getCpuGenerator(CpuId cpu) { return gens_.at(cpu); }

  // what the callback of onCpusThreadSwitchBufferRead(callback, consume=true) hands over
  void onCpuThreadSwitchBufferRead(CpuId cpu, const RbDataSlices& s) {
    if (decoded_) {  // a new window: forget the previous images, keep the generators' state
      check(hbtgpu_cs_reset(ctx_));
      decoded_ = false;
    }
    // a wrapped ring is cut wherever the byte count says, usually inside a record: both slices
    // become ONE image (consume() reassembles such a record, PerfEventsGroup.h:1425-1437)
    check(hbtgpu_cs_submit_image_slices(ctx_, static_cast<uint32_t>(cpu), s.first.from, s.first.size,
                                        s.second ? s.second->from : nullptr, s.second ? s.second->size : 0));
  }

  std::vector<ThreadInfo> getThreadsInfo() const {
    const uint64_t nt = hbtgpu_cs_threads_count(threads_), nn = hbtgpu_cs_threads_name_count(threads_);
    std::vector<hbtgpu_cs_thread> th(nt);
    std::vector<hbtgpu_cs_name> nm(nn);
    hbtgpu_cs_threads_get(threads_, th.data(), nt, nm.data(), nn);
    std::vector<ThreadInfo> out;
    size_t k = 0;
    for (const auto& t : th) {
      ThreadInfo ti{t.vid, t.ppid, t.ptid, t.pid, t.tid, t.fork_ts, t.end_ts, {}};
      for (uint32_t i = 0; i < t.n_names; ++i, ++k)
        ti.names.push_back(ThreadNameChange{nm[k].tstamp, std::string(nm[k].name, nm[k].len), nm[k].is_exec != 0});
      out.push_back(std::move(ti));
    }
    return out;
  }

 private:
  friend class CpuThreadSwitchGenerator;
  void check(int rc) {
    if (rc == -EINVAL) throw std::invalid_argument(hbtgpu_cs_last_error(ctx_));
    if (rc != 0) throw std::runtime_error(hbtgpu_cs_last_error(ctx_));
  }
  // decode the window on the GPU, then run the thread dictionary over the cpus in index order
  void decode() {
    if (decoded_) return;
    check(hbtgpu_cs_run(ctx_));
    check(hbtgpu_cs_sync(ctx_));
    std::vector<uint32_t> pids;
    std::vector<hbtgpu_cs_op> ops;
    for (auto& g : gens_) {
      const uint32_t cpu = static_cast<uint32_t>(g.cpu_);
      uint64_t ne = 0, no = 0;
      check(hbtgpu_cs_num_events(ctx_, cpu, &ne, &no));
      g.events_.resize(ne);
      pids.resize(ne);
      ops.resize(no);
      check(hbtgpu_cs_get_events(ctx_, cpu, g.events_.data(), pids.data(), ne, &ne));
      check(hbtgpu_cs_get_ops(ctx_, cpu, ops.data(), no, &no));
      if (hbtgpu_cs_assign_vids(threads_, g.events_.data(), pids.data(), ne, ops.data(), no) != 0)
        throw std::runtime_error("thread records outside the reference's contract (HBT_THROW_ASSERT there)");
      check(hbtgpu_cs_get_counters(ctx_, cpu, &g.c_));
      hbtgpu_cs_state st{};
      check(hbtgpu_cs_get_state(ctx_, cpu, &st));
      check(hbtgpu_cs_set_state(ctx_, cpu, &st));  // last_tstamp_ / in_write_errors_ carry over
    }
    decoded_ = true;
  }

  hbtgpu_cs_ctx* ctx_ = nullptr;
  hbtgpu_cs_threads* threads_ = nullptr;
  std::vector<CpuThreadSwitchGenerator> gens_;
  bool decoded_ = false;
};

inline ssize_t CpuThreadSwitchGenerator::consume(unsigned max_num_records) {
  if (max_num_records == 0) return 0;
  owner_->decode();
  if (c_.err < 0) return static_cast<ssize_t>(c_.err);
  return -ENODATA;  // drained (:1574-1588)
}

// ---- tagstack slices x count intervals -------------------------------------------------------

using Slice = hbtgpu_slice;
enum class TransitionType : uint8_t { NA = 0, Analysis, ThreadPreempted, ThreadYield, PhaseChange };  // Slicer.h:31-43
inline Slice makeSlice(TimeStamp tstamp, TimeStamp duration, uint64_t stack_id, TransitionType swin,
                       TransitionType swout) {  // Slice::makeSlice, Slicer.h:51-63
  Slice s{};
  s.tstamp = tstamp;
  s.duration = duration;
  s.stack_id = stack_id;
  s.swin_type = static_cast<uint8_t>(swin);
  s.swout_type = static_cast<uint8_t>(swout);
  return s;
}
inline bool operator==(const Slice& a, const Slice& b) {  // Slicer.h:65-69
  return a.tstamp == b.tstamp && a.duration == b.duration && a.stack_id == b.stack_id &&
         a.swin_type == b.swin_type && a.swout_type == b.swout_type;
}
using Slices = std::vector<Slice>;

class PerIntervalSlices : public std::deque<Slices> {  // IntervalSlicer.h:17-82
 public:
  size_t countSlices() const {
    size_t n = 0;
    for (const auto& s : *this) n += s.size();
    return n;
  }
  size_t numIntervals() const { return size(); }
};

// One comp unit's slice store with consumer marks.  The slices stay in host memory (the Slicer
// that writes them is host code); every query ships the live ones to the GPU.  For whole
// windows use releaseCountRows below: all CPUs and all rows in one call.
class IntervalSlicer {
 public:
  IntervalSlicer(size_t min_num_slices, size_t num_slices_drop_if_full, int device = 0)
      : drop_if_full_{num_slices_drop_if_full} {
    if (min_num_slices == 0) throw std::invalid_argument("min_num_slices must be positive");
    capacity_ = 1;
    while (capacity_ * sizeof(Slice) < min_num_slices * sizeof(Slice)) capacity_ <<= 1;  // nextPow2 of the byte size, RingBuffer.h:69
    ctx_ = hbtgpu_islice_create(device);
    if (!ctx_) throw std::runtime_error("hbtgpu_islice_create failed: no CUDA device (there is no CPU path)");
  }
  ~IntervalSlicer() { hbtgpu_islice_destroy(ctx_); }
  IntervalSlicer(const IntervalSlicer&) = delete;
  IntervalSlicer& operator=(const IntervalSlicer&) = delete;

  ssize_t write(const Slice& slice) {  // IntervalSlicer.h:115-135
    if (slice.tstamp == kInvalidTimeStamp) throw std::invalid_argument("Invalid timestamp");
    if (slices_.size() >= capacity_) {
      if (drop_if_full_ == 0) return -ENOSPC;
      const size_t n = std::min(drop_if_full_, slices_.size());
      slices_.erase(slices_.begin(), slices_.begin() + n);
      num_dropped_ += n;
    }
    slices_.push_back(slice);
    return sizeof(Slice);
  }
  void registerConsumerMark(const std::string& key) {
    if (marks_.count(key)) throw std::invalid_argument("Consumer with key " + key + " already registered.");
    marks_[key] = num_processed_slices_;
  }
  void updateConsumerMark(const std::string& key, size_t new_offset) { getMark(key) = new_offset; }
  void removeConsumerMark(const std::string& key) { marks_.erase(key); }
  size_t& getMark(const std::string& key) {
    auto it = marks_.find(key);
    if (it == marks_.end()) throw std::invalid_argument("No consumer mark for key: \"" + key + "\"");
    return it->second;
  }
  size_t numActiveSlices() const noexcept { return slices_.size(); }
  size_t numProcessedSlices() const noexcept { return num_processed_slices_; }
  void discardConsumed() {  // IntervalSlicer.h:178-207
    if (marks_.empty()) return;
    size_t m = ~size_t{0};
    for (const auto& kv : marks_) m = std::min(m, kv.second);
    const size_t first = num_processed_slices_ + num_dropped_;  // absolute index of slices_[0]
    if (m > first) {
      const size_t n = std::min(m - first, slices_.size());
      slices_.erase(slices_.begin(), slices_.begin() + n);
      num_processed_slices_ += n;
    }
  }

  // IntervalSlicer.cpp:52-188.  offset counts slices since the slicer was made, as the marks do.
  std::pair<std::unique_ptr<Slices>, size_t> makeIntervalIntersectingSlices(size_t offset, TimeStamp tstamp,
                                                                            TimeStamp duration, bool trim,
                                                                            bool abort_on_missing_slices) {
    const uint64_t row[2] = {tstamp, duration};
    std::vector<uint32_t> counts;
    Slices out;
    size_t mark = offset;
    // stop_ts: the row loop's cut does not exist here; ~0 is never reached by a wrapping sum + 1
    releaseRows(row, 1, 2, ~0ull, trim, abort_on_missing_slices, &mark, &counts, &out, /*no_stop=*/true);
    if (counts.empty() || counts[0] == 0) return {nullptr, 0};
    return {std::make_unique<Slices>(std::move(out)), mark - offset};
  }

  std::unique_ptr<PerIntervalSlices> consumeSlices(const std::string& key, uint64_t tstamp, uint64_t duration,
                                                   bool trim, bool abort_on_missing_slices) {  // :195-217
    auto pis = std::make_unique<PerIntervalSlices>();
    auto& offset = getMark(key);
    auto [slices, n] = makeIntervalIntersectingSlices(offset, tstamp, duration, trim, abort_on_missing_slices);
    if (slices == nullptr) pis->push_back({});
    else pis->push_back(std::move(*slices));
    offset += n;
    discardConsumed();
    return pis;
  }

  // The row loop of TraceMonitor::releaseCountData_ (TraceMonitor.cpp:52-96) for this unit:
  // rows = n_rows x row_stride u64 (word 0 tstamp, word 1 duration).  Returns the rows used.
  size_t releaseCountRows(const std::string& key, const uint64_t* rows, size_t n_rows, size_t row_stride,
                          TimeStamp stop_ts, PerIntervalSlices* pislices) {
    std::vector<uint32_t> counts;
    Slices out;
    const size_t n = releaseRows(rows, n_rows, row_stride, stop_ts, true, true, &getMark(key), &counts, &out, false);
    size_t o = 0;
    for (size_t k = 0; k < n; ++k) {
      pislices->emplace_back(out.begin() + o, out.begin() + o + counts[k]);
      o += counts[k];
    }
    discardConsumed();
    return n;
  }

 private:
  size_t releaseRows(const uint64_t* rows, size_t n_rows, size_t stride, TimeStamp stop_ts, bool trim, bool abort_missing,
                     size_t* mark, std::vector<uint32_t>* counts, Slices* out, bool no_stop) {
    const size_t first = num_processed_slices_ + num_dropped_;
    // seekInTx clamps the offset into [tail, head] (ringbuffer/Consumer.h:281)
    uint64_t rel = *mark > first ? *mark - first : 0;
    const uint64_t rel_in = rel;
    const uint64_t sb[2] = {0, slices_.size()}, rb[2] = {0, n_rows};
    hbtgpu_islice_unit u{};
    const uint32_t flags = (trim ? HBTGPU_ISLICE_TRIM : 0u) | (abort_missing ? HBTGPU_ISLICE_ABORT_ON_MISSING : 0u);
    std::vector<uint64_t> one;
    if (no_stop && n_rows == 1 && rows[0] + rows[1] == ~0ull) {
      // an interval that ends exactly at ~0 would meet the stop rule; it cannot hold slices
      // that the reference would not also reject (tstamp ~0 is kInvalidTimeStamp)
      counts->assign(1, 0);
      return 1;
    }
    const int rc = hbtgpu_islice_release(ctx_, slices_.data(), sb, rows, stride, rb, 1, stop_ts, flags, &rel, &u);
    if (rc != 0) throw std::runtime_error(hbtgpu_islice_last_error(ctx_));
    if (u.status == -EINVAL) throw std::runtime_error("IntervalSlicer: inconsistent interval (HBT_THROW_ASSERT / HBT_ARG_CHECK in the reference)");
    if (u.status != 0) throw std::runtime_error("IntervalSlicer: too many slices for one call");
    counts->resize(u.n_intervals);
    out->resize(u.n_slices);
    if (hbtgpu_islice_get(ctx_, 0, counts->data(), counts->size(), out->data(), out->size()) != 0)
      throw std::runtime_error(hbtgpu_islice_last_error(ctx_));
    *mark += rel - rel_in;
    return u.n_intervals;
  }

  hbtgpu_islice_ctx* ctx_ = nullptr;
  std::vector<Slice> slices_;  // the live slices, oldest first
  size_t capacity_ = 0, drop_if_full_ = 0;
  size_t num_processed_slices_ = 0, num_dropped_ = 0;
  std::map<std::string, size_t> marks_;
};

// ---- tagstack events -> slices -----------------------------------------------------------------

using SlicerStats = hbtgpu_slicer_stats;    // Slicer.h:276-285
using TagStackStats = hbtgpu_stack_stats;   // Slicer.h:160-193 (stack = tags[0 .. num_set_tags), known_num_levels)

// tagstack::slice::Slicer for the events the thread-switch generators emit (level 0).  The
// reference pulls events from a merged stream until stop_ts (processNext); here a window of
// per-CPU events is handed over at once and the slices are written to the per-CPU producers
// (anything with write(const Slice&), e.g. IntervalSlicer) exactly as completeSlice_ would
// (Slicer.h:529-558).  Successive windows continue each other, as the reference's one Slicer does.
template <class TSlices>
class Slicer {
 public:
  using TGroupedSlices = std::vector<std::shared_ptr<TSlices>>;  // GroupedSlices<TSlices>, Slicer.h:273-274
  explicit Slicer(std::shared_ptr<TGroupedSlices> grouped_slices, int device = 0) : grouped_slices_{grouped_slices} {
    if (!grouped_slices || grouped_slices->empty()) throw std::invalid_argument("grouped_slices must be populated");
    ctx_ = hbtgpu_slicer_create(device);
    if (!ctx_) throw std::runtime_error("hbtgpu_slicer_create failed: no CUDA device (there is no CPU path)");
  }
  ~Slicer() { hbtgpu_slicer_destroy(ctx_); }
  Slicer(const Slicer&) = delete;
  Slicer& operator=(const Slicer&) = delete;

  // Returns the number of processed events.  events[cpu] = that CPU's events in stream order.
  size_t processWindow(const std::vector<std::vector<Event>>& events) {
    if (events.size() != grouped_slices_->size()) throw std::invalid_argument("one event vector per slice producer");
    std::vector<uint64_t> begin(events.size() + 1, 0);
    std::vector<Event> all;
    for (size_t u = 0; u < events.size(); ++u) {
      all.insert(all.end(), events[u].begin(), events[u].end());
      begin[u + 1] = all.size();
    }
    std::vector<uint64_t> n_slices(events.size());
    uint64_t n_stacks = 0;
    const uint64_t before = stats_.num_processed;  // the statistics are cumulative, processNext's result is not
    const int rc = hbtgpu_slicer_run(ctx_, all.data(), begin.data(), static_cast<uint32_t>(events.size()),
                                     started_ ? HBTGPU_SLICER_CONTINUE : 0u, &stats_, n_slices.data(), &n_stacks, &last_ts_);
    if (rc != 0) throw std::runtime_error(hbtgpu_slicer_last_error(ctx_));
    started_ = true;
    for (size_t u = 0; u < events.size(); ++u) {
      std::vector<Slice> out(n_slices[u]);
      if (!out.empty() && hbtgpu_slicer_get_slices(ctx_, static_cast<uint32_t>(u), out.data(), out.size()) != 0)
        throw std::runtime_error(hbtgpu_slicer_last_error(ctx_));
      for (const auto& s : out)
        if ((*grouped_slices_)[u]->write(s) < 0)
          throw std::runtime_error("Buffer back-pressure when generating slices is not yet supported.");  // :551-553
    }
    stacks_.resize(n_stacks);
    if (n_stacks && hbtgpu_slicer_get_stacks(ctx_, stacks_.data(), n_stacks) != 0)
      throw std::runtime_error(hbtgpu_slicer_last_error(ctx_));
    return stats_.num_processed - before;
  }

  auto getSlices() { return grouped_slices_; }
  TimeStamp getLastTimeStamp() const { return last_ts_; }
  SlicerStats getStats() const { return stats_; }
  // by stack_id, as the std::map of the reference iterates
  const std::vector<TagStackStats>& getTagStackStatsDump() const { return stacks_; }

 private:
  hbtgpu_slicer_ctx* ctx_ = nullptr;
  std::shared_ptr<TGroupedSlices> grouped_slices_;
  SlicerStats stats_{};
  std::vector<TagStackStats> stacks_;
  TimeStamp last_ts_ = 0;
  bool started_ = false;
};

}  // namespace hbtgpu_shim
