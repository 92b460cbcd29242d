// Reference-shaped test of the C++ shim (dynolog_b200/shim/HbtGpuShim.h): written the way
// hbt/src/perf_event/tests/PerCpuGeneratorsTest.cpp:53-88 drives a generator -- feed the ring,
// consume(), check ret >= 0 || ret == -ENODATA and the getNumRecord*() counters -- but on a
// synthetic ring image with known answers (the survey's probe: 10 SAMPLE + LOST(3) + THROTTLE
// => samples 10, lost 3, throttle 1, 9 rows).
#include <cassert>
#include <cstdio>
#include <cstring>
#include <vector>

#include "HbtGpuShim.h"

using namespace hbtgpu_shim;

static void put(std::vector<uint8_t>& b, uint64_t v) {
  for (int i = 0; i < 8; ++i) b.push_back((v >> (8 * i)) & 255);
}
static uint64_t hdr(uint32_t type, uint32_t size) { return type | (uint64_t(size) << 48); }

#define EXPECT(cond)                                                       \
  do {                                                                     \
    if (!(cond)) {                                                         \
      std::fprintf(stderr, "EXPECT failed: %s (line %d)\n", #cond, __LINE__); \
      return 1;                                                            \
    }                                                                      \
  } while (0)

#define EXPECT_V(cond)                                                      \
  do {                                                                      \
    if (!(cond)) std::fprintf(stderr, "EXPECT failed: %s (line %d)\n", #cond, __LINE__); \
  } while (0)

int main() {
  std::vector<uint8_t> ring;
  for (int i = 0; i < 10; ++i) {
    put(ring, hdr(9, 72));
    put(ring, 1000 + 100 * i);  // time
    put(ring, 4);               // nr
    put(ring, 100 * i);         // time_enabled
    put(ring, 50 * i);          // time_running: multiplexed at 50 %
    for (int e = 0; e < 4; ++e) put(ring, 10 * i * (e + 1));
    if (i == 4) {
      put(ring, hdr(2, 24));  // LOST
      put(ring, 7);
      put(ring, 3);
    }
  }
  put(ring, hdr(5, 32));  // THROTTLE
  put(ring, 5);
  put(ring, 7);
  put(ring, 9);

  PerCpuCountSampleGenerator::Options opt;
  opt.num_cpus = 2;
  opt.t0 = 1000;
  opt.bucket_ns = 300;
  opt.n_buckets = 4;
  opt.metrics.push_back(hbtgpu_metric{HBTGPU_METRIC_RATE, 2, {1, 0}});
  PerCpuCountSampleGenerator gen({"cycles", "instructions", "llc_misses", "branch_misses"}, opt);
  gen.enable(0, 900);

  // the ring wrapped: hand over the image as two slices, as onCpuDataBufferRead would (:1626-1632)
  const size_t cut = 72 * 3;
  gen.onCpuDataBufferRead(0, RbDataSlices(ring.data(), cut, ring.data() + cut, ring.size() - cut));

  auto& g0 = gen.getCpuGenerator(0);
  ssize_t ret = g0.consume(1000);
  EXPECT(ret >= 0 || ret == -ENODATA);  // PerCpuGeneratorsTest.cpp:77
  EXPECT(g0.getNumRecordSample() == 10);
  EXPECT(g0.getNumRecordLost() == 3);  // SUM of num_lost
  EXPECT(g0.getNumRecordThrottle() == 1);
  EXPECT(g0.getNumRecordUnThrottle() == 0);
  EXPECT(g0.getNumEntries() == 9);
  EXPECT(g0.getLastTimeStamp() == 1900);
  EXPECT(gen.getCpuGenerator(1).getNumRecordSample() == 0);
  EXPECT(gen.accumUntil(100, 0) == 2);

  size_t n_rows = 0;
  uint64_t first[6] = {0};
  gen.applyToCountSamples([&](CpuId cpu, size_t idx, const std::vector<std::string>& names, TimeStamp ts,
                              TimeStamp dur, const uint64_t* counts, size_t n) {
    if (cpu == 0 && idx == 1) {
      first[0] = ts;
      first[1] = dur;
      for (size_t i = 0; i < n && i < 4; ++i) first[2 + i] = counts[i];
    }
    (void)names;
    ++n_rows;
  });
  EXPECT(n_rows == 9);
  EXPECT(first[0] == 1000 && first[1] == 100 && first[2] == 10 && first[5] == 40);

  std::vector<uint64_t> acc, scaled;
  std::vector<double> metric;
  std::vector<uint32_t> valid;
  gen.getBuckets(&acc, &scaled, &metric, &valid);
  // bucket 0 of cpu 0 = [1000,1300): rows at 1000 (from 900), 1100, 1200 -> cycles 0+10+10,
  // dTE = 200, dTR = 100 -> scaled = 20 * 200 / 100 = 40 (GroupReadValues::getCount)
  EXPECT(acc[0] == 20 && acc[4] == 200 && acc[5] == 100 && acc[6] == 3);
  EXPECT(scaled[0] == 40 && scaled[1] == 80);
  EXPECT(valid[0] == 1 && metric[0] == 2.0);  // instructions / cycles

  // a malformed record stalls the walk with -EPERM, like consume() (:1502-1507)
  PerCpuCountSampleGenerator bad({"a", "b"}, PerCpuCountSampleGenerator::Options{});
  std::vector<uint8_t> r2;
  put(r2, hdr(9, 56));
  put(r2, 10);
  put(r2, 2);
  put(r2, 1);
  put(r2, 1);
  put(r2, 5);
  put(r2, 6);
  put(r2, hdr(9, 64));  // size says 3 values, nr says 2
  for (int i = 0; i < 7; ++i) put(r2, 2);
  bad.onCpuDataBufferRead(0, RbDataSlices(r2.data(), r2.size()));
  EXPECT(bad.getCpuGenerator(0).consume(10) == -EPERM);
  EXPECT(bad.getCpuGenerator(0).getNumRecordSample() == 1);
  bool threw = false;
  try {
    bad.accumUntil(10, 1000);  // last timestamp 10 <= stop_ts: consume() runs into the bad record
  } catch (const std::runtime_error&) {
    threw = true;  // accumUntil throws on unexpected errors (:76-79)
  }
  EXPECT(threw);
  threw = false;
  try {
    PerCpuCountSampleGenerator none({}, PerCpuCountSampleGenerator::Options{});
  } catch (const std::invalid_argument&) {
    threw = true;
  }
  EXPECT(threw);

  // ---- accumUntil(batch, stop_ts), PerCpuSampleGeneratorBase.h:69-94: a CPU is consumed batch by
  // batch while its last timestamp is <= stop_ts; what lies behind stays in the ring for the next call
  {
    auto put_sample = [&](std::vector<uint8_t>& b, uint64_t ts, uint64_t v) {
      put(b, hdr(9, 56));
      put(b, ts);
      put(b, 2);
      put(b, ts);
      put(b, ts);
      put(b, v);
      put(b, 2 * v);
    };
    PerCpuCountSampleGenerator::Options o2;
    o2.num_cpus = 1;
    o2.t0 = 0;
    o2.bucket_ns = 1000;
    o2.n_buckets = 8;
    PerCpuCountSampleGenerator g2({"a", "b"}, o2);
    g2.enable(0, 50);
    std::vector<uint8_t> w1;
    for (int i = 1; i <= 10; ++i) put_sample(w1, 100 * i, 10 * i);  // samples at 100 .. 1000
    g2.onCpuDataBufferRead(0, RbDataSlices(w1.data(), w1.size()));
    auto& c0 = g2.getCpuGenerator(0);
    // batch 3, stop at 450: batches [100,200,300] (last 300 <= 450) and [400,500,600] (last 600 > 450)
    EXPECT(g2.accumUntil(3, 450) == 1);
    EXPECT(c0.getNumRecordSample() == 6 && c0.getNumEntries() == 6 && c0.getLastTimeStamp() == 600);
    // nothing more happens while stop_ts is behind the last timestamp
    EXPECT(g2.accumUntil(3, 450) == 1 && c0.getNumRecordSample() == 6);
    size_t seen = 0;
    uint64_t last_ts = 0, last_delta = 0;
    g2.applyToCountSamples([&](CpuId, size_t idx, const std::vector<std::string>&, TimeStamp ts, TimeStamp dur,
                               const uint64_t* counts, size_t) {
      EXPECT_V(idx == seen);
      ++seen;
      last_ts = ts + dur;
      last_delta = counts[0];
    });
    EXPECT(seen == 6 && last_ts == 600 && last_delta == 10 && c0.getNumEntries() == 0);
    // the next window arrives while four records are still unconsumed: a daemon's loop
    std::vector<uint8_t> w2;
    for (int i = 11; i <= 14; ++i) put_sample(w2, 100 * i, 10 * i);
    const size_t cut = 56 + 24;  // the ring wrapped inside the second record
    g2.onCpuDataBufferRead(0, RbDataSlices(w2.data(), cut, w2.data() + cut, w2.size() - cut));
    EXPECT(g2.accumUntil(100, 100000) == 1);  // drains: -ENODATA ends the CPU quietly
    EXPECT(c0.getNumRecordSample() == 14 && c0.getNumEntries() == 8 && c0.getLastTimeStamp() == 1400);
    seen = 0;
    uint64_t first_ts = 0, first_delta = 0;
    g2.applyToCountSamples([&](CpuId, size_t idx, const std::vector<std::string>&, TimeStamp ts, TimeStamp,
                               const uint64_t* counts, size_t) {
      if (idx == 0) {
        first_ts = ts;
        first_delta = counts[1];
      }
      ++seen;
    });
    EXPECT(seen == 8 && first_ts == 600 && first_delta == 20);  // row 700: continues from the sample at 600
    // a third window: only the new bytes are decoded (lifetime counters keep growing)
    std::vector<uint8_t> w3;
    put_sample(w3, 1500, 150);
    put(w3, hdr(2, 24));  // LOST: count_.tstamp becomes invalid ...
    put(w3, 7);
    put(w3, 5);
    put_sample(w3, 1600, 160);
    g2.onCpuDataBufferRead(0, RbDataSlices(w3.data(), w3.size()));
    EXPECT(g2.accumUntil(2, 100000) == 1);
    // ... and an invalid timestamp (the maximum) is never <= stop_ts: the reference stops consuming
    // this CPU right there, and so does the shim
    EXPECT(c0.getNumRecordSample() == 15 && c0.getNumRecordLost() == 5 && c0.getLastTimeStamp() == kInvalidTimeStamp);
    EXPECT(c0.consume(10) == -ENODATA && c0.getNumRecordSample() == 16 && c0.getNumUnknownCountInterval() == 1);
    EXPECT(c0.getLastTimeStamp() == 1600 && c0.getNumEntries() == 1);

    // time going back: -EPERM once, count_.tstamp invalidated, the next consume() re-bases (gen :156-163)
    PerCpuCountSampleGenerator g3({"a", "b"}, o2);
    g3.enable(0, 50);
    std::vector<uint8_t> w4;
    put_sample(w4, 100, 1);
    put_sample(w4, 200, 2);
    put_sample(w4, 150, 3);
    put_sample(w4, 300, 4);
    g3.onCpuDataBufferRead(0, RbDataSlices(w4.data(), w4.size()));
    auto& d0 = g3.getCpuGenerator(0);
    EXPECT(d0.consume(10) == -EPERM && d0.getNumRecordSample() == 2 && d0.getLastTimeStamp() == kInvalidTimeStamp);
    EXPECT(d0.consume(10) == -ENODATA && d0.getNumRecordSample() == 4 && d0.getNumEntries() == 3);
    EXPECT(d0.getNumUnknownCountInterval() == 1 && d0.getLastTimeStamp() == 300);

    // the sink: MetricSeries reductions of the buckets through a dynolog::Logger-shaped object
    struct Collect : Logger {
      std::map<std::string, uint64_t> u;
      std::map<std::string, float> f;
      int entries = 0;
      void logInt(const std::string&, int64_t) override {}
      void logFloat(const std::string& k, float v) override { f[k] = v; }
      void logUint(const std::string& k, uint64_t v) override { u[k] = v; }
      void logStr(const std::string&, const std::string&) override {}
      void finalize() override { ++entries; }
    } sink;
    logSeries(g3, sink, 0, o2.n_buckets, 0.5f, 1000, 1000);
    EXPECT(sink.entries == 1 && sink.u.count("a_diff.cpu0") == 1 && sink.f.count("b_avg.cpu0") == 1);

    // GroupReadValues (PerfEventsGroup.h:386-586) over the library's arithmetic; known answers of
    // hbt/src/mon/tests/MonitorMockTest.cpp:86-88,121-126
    GroupReadValues rv(g3.raw(), 2);
    rv.setTimeEnabled(4);
    rv.setTimeRunning(2);
    rv.setCount(0, 200);
    rv.setCount(1, 100);
    EXPECT(rv.getCount(0) == 400 && rv.hasMultiplexing() && rv.getRunningRatioInteger100() == 50);
    EXPECT(rv.getRateReducedCount().value() == 2.0);
    GroupReadValues rw(g3.raw(), 2);
    rw.setTimeEnabled(1);
    rw.setTimeRunning(1);
    rw.setCount(0, 100);
    EXPECT(rw.getCount(0) == 100 && !rw.hasMultiplexing());
    rv.accum(rw);
    EXPECT(rv.getRawCount(0) == 300 && rv.getTimeEnabled() == 5);
    rv.diff(rw);
    EXPECT(rv.getRawCount(0) == 200 && rv.getTimeRunning() == 2);  // PerCpuGeneratorsTest.cpp:195-232
  }
  // ---- ContextSwitch mode: the stream of tests/test_cs_oracle.py::test_known_answer_small ----
  {
    const uint32_t T = 5000001;
    std::vector<uint8_t> cs;
    auto sid = [&](uint32_t pid, uint32_t tid, uint64_t ts) {
      put(cs, pid | (uint64_t(tid) << 32));
      put(cs, ts);
    };
    auto sw = [&](uint64_t ts, uint32_t tid, uint32_t misc) {
      put(cs, 15 | (uint64_t(misc) << 32) | (32ull << 48));
      put(cs, 0);
      sid(1, tid, ts);
    };
    sw(1000, T, (1u << 13) | (1u << 14));  // out, preempted
    sw(1100, T + 1, 0);                    // in
    put(cs, hdr(2, 40));                   // LOST
    put(cs, 7);
    put(cs, 7);
    sid(0, 0, 1500);
    sw(1600, T + 1, 1u << 13);             // out, yield
    put(cs, hdr(7, 48));                   // FORK
    put(cs, (T + 8) | (1ull << 32));
    put(cs, (T + 9) | (1ull << 32));
    put(cs, 1700);
    sid(T + 8, T + 9, 1700);
    put(cs, hdr(4, 48));                   // EXIT
    put(cs, (T + 8) | (1ull << 32));
    put(cs, (T + 9) | (1ull << 32));
    put(cs, 1900);
    sid(T + 8, T + 9, 1900);
    sw(2000, 0, 0);                        // idle in
    PerCpuThreadSwitchGenerator tsg(2, 3);
    // the ring wrapped in the middle of the FORK record (a wrap almost never falls on a
    // record boundary; consume() reassembles the record, PerfEventsGroup.h:1425-1437)
    const size_t wrap = 32 + 32 + 40 + 32 + 20;
    tsg.onCpuThreadSwitchBufferRead(0, RbDataSlices(cs.data(), wrap, cs.data() + wrap, cs.size() - wrap));
    auto& t0 = tsg.getCpuGenerator(0);
    ssize_t r = t0.consume(1000);
    EXPECT(r >= 0 || r == -ENODATA);
    EXPECT(t0.getNumRecordSwitchCpuWide() == 4 && t0.getNumRecordLost() == 7);
    EXPECT(t0.getNumRecordFork() == 1 && t0.getNumRecordExit() == 1);
    EXPECT(t0.getNumEntries() == 8 && t0.getLastTimeStamp() == 2000);
    const auto& ev = t0.events();
    EXPECT(ev[0].type == 5 && ev[0].tag == 0xFFFFFFFFu && ev[0].level == 3);   // SwitchOutPreempt
    EXPECT(ev[1].type == 4 && ev[1].tag == 1);                                  // SwitchIn, vid 1
    EXPECT(ev[2].type == 7 && ev[2].tstamp == 1101 && ev[2].level == 255);      // WriteErrorsStart
    EXPECT(ev[3].type == 8 && ev[3].tstamp == 1600);                            // WriteErrorsEnd
    EXPECT(ev[4].type == 6);                                                    // SwitchOutYield
    EXPECT(ev[5].type == 2 && ev[5].tag == 2);                                  // ThreadCreation, vid 2
    EXPECT(ev[6].type == 3 && ev[6].tag == 2);                                  // ThreadDestruction
    EXPECT(ev[7].type == 4 && ev[7].tag == 0);                                  // tid 0 is never remapped
    auto info = tsg.getThreadsInfo();
    EXPECT(info.size() == 2 && info[1].tid == pid_t(T + 9) && info[1].fork_timestamp == 1700 &&
           info[1].end_timestamp == 1900 && info[0].ppid == -1);
    EXPECT(tsg.getCpuGenerator(1).getNumEntries() == 0);
  }

  // ---- tagstack::IntervalSlicer, after hbt/src/tagstack/tests/IntervalSlicerTest.cpp ----------
  {
    const auto A = TransitionType::Analysis;
    {  // EmptySlices :13-22
      IntervalSlicer islicer{20, 0};
      auto [s, n] = islicer.makeIntervalIntersectingSlices(0, 0, 100, false, false);
      EXPECT(s == nullptr && n == 0 && islicer.numProcessedSlices() == 0);
      islicer.discardConsumed();
    }
    {  // SingleConsumerMultiple :124-204
      IntervalSlicer islicer{20, 0};
      auto s1 = makeSlice(0, 99, 1, A, A);
      EXPECT(islicer.write(s1) >= 0);
      {
        auto [s, n] = islicer.makeIntervalIntersectingSlices(0, 0, 100, true, true);
        EXPECT(s == nullptr && n == 0);
      }
      auto s2 = makeSlice(99, 2, 1, A, A);
      EXPECT(islicer.write(s2) >= 0);
      {
        auto [s, n] = islicer.makeIntervalIntersectingSlices(0, 0, 100, true, true);
        EXPECT(s != nullptr && n == 1 && s->size() == 2 && s->at(0) == s1 && s->at(1) == makeSlice(99, 1, 1, A, A));
      }
      {
        auto [s, n] = islicer.makeIntervalIntersectingSlices(1, 0, 101, true, true);
        EXPECT(s == nullptr && n == 0);
      }
      auto [s, n] = islicer.makeIntervalIntersectingSlices(0, 100, 1, true, true);
      EXPECT(s != nullptr && n == 2 && s->size() == 1 && s->at(0) == makeSlice(100, 1, 1, A, A));
    }
    {  // MultipleSlicerInSequence :206-300 (first part)
      IntervalSlicer islicer{20, 0};
      islicer.registerConsumerMark("c1");
      islicer.registerConsumerMark("c2");
      auto s1 = makeSlice(0, 99, 1, A, A);
      EXPECT(islicer.write(s1) >= 0 && islicer.numActiveSlices() == 1);
      {
        auto s = islicer.consumeSlices("c1", 0, 100, true, true);
        EXPECT(islicer.getMark("c1") == 0 && s->size() == 1 && s->at(0).size() == 0 && islicer.numProcessedSlices() == 0);
      }
      auto s2 = makeSlice(99, 2, 33, TransitionType::ThreadPreempted, TransitionType::PhaseChange);
      EXPECT(islicer.write(s2) >= 0 && islicer.numActiveSlices() == 2);
      {
        auto s = islicer.consumeSlices("c2", 0, 99, true, true);
        EXPECT(islicer.getMark("c2") == 1 && islicer.numProcessedSlices() == 0 && s->at(0).at(0) == s1);
      }
      {
        auto s = islicer.consumeSlices("c1", 0, 99, true, true);
        EXPECT(s->at(0).at(0) == s1 && islicer.getMark("c1") == 1 && islicer.numActiveSlices() == 1);
      }
      {
        auto s = islicer.consumeSlices("c1", 100, 1, true, true);
        EXPECT(s->size() == 1 && s->at(0).size() == 1 &&
               s->at(0).at(0) == makeSlice(100, 1, 33, A, TransitionType::PhaseChange));
      }
    }
    {  // a window of count rows in one call (TraceMonitor::releaseCountData_'s loop)
      IntervalSlicer islicer{64, 0};
      islicer.registerConsumerMark("counts");
      for (uint64_t i = 0; i < 10; ++i)
        EXPECT(islicer.write(makeSlice(1000 + 30 * i, 30, i % 3, TransitionType::ThreadYield, TransitionType::ThreadPreempted)) >= 0);
      const uint64_t rows[4][3] = {{1000, 45, 7}, {1045, 45, 7}, {1090, 60, 7}, {1150, 500, 7}};
      PerIntervalSlices pis;
      const size_t used = islicer.releaseCountRows("counts", &rows[0][0], 4, 3, 1600, &pis);
      EXPECT(used == 3 && pis.numIntervals() == 3 && pis.countSlices() == 2 + 2 + 2);
      EXPECT(pis[0][0] == makeSlice(1000, 30, 0, TransitionType::ThreadYield, TransitionType::ThreadPreempted));
      EXPECT(pis[0][1] == makeSlice(1030, 15, 1, TransitionType::ThreadYield, A));
      EXPECT(pis[1][0] == makeSlice(1045, 15, 1, A, TransitionType::ThreadPreempted));
      EXPECT(pis[2][1] == makeSlice(1120, 30, 1, TransitionType::ThreadYield, TransitionType::ThreadPreempted));
      EXPECT(islicer.getMark("counts") == 5 && islicer.numProcessedSlices() == 5 && islicer.numActiveSlices() == 5);
    }
  }

  // ---- tagstack::slice::Slicer, after TagStackSlice.1LevelsOnePhaseWithThreadSwOut -----------
  // (hbt/src/tagstack/tests/SlicerTest.cpp:65-160), slices written into an IntervalSlicer as
  // TraceMonitor wires them, then one count interval taken out of it
  {
    auto mk = [](uint64_t ts, uint8_t type, uint32_t tag) {
      Event e{};
      e.tstamp = ts;
      e.type = type;
      e.level = 0;
      e.comp_unit_id = 0;
      e.tag = tag;
      return e;
    };
    const uint8_t kIn = 4, kOutPreempt = 5;
    std::vector<std::vector<Event>> events(1);
    events[0] = {mk(20, kIn, 2), mk(30, kOutPreempt, ~0u), mk(30, kIn, 4), mk(50, kOutPreempt, ~0u),
                 mk(60, kIn, 1), mk(80, kOutPreempt, ~0u)};
    auto grouped = std::make_shared<Slicer<IntervalSlicer>::TGroupedSlices>();
    grouped->push_back(std::make_shared<IntervalSlicer>(64, 0));
    Slicer<IntervalSlicer> slicer{grouped};
    EXPECT(slicer.processWindow(events) == 6);
    const auto& st = slicer.getTagStackStatsDump();
    EXPECT(st.size() == 5);                                                       // :97
    EXPECT(st[0].num_set_tags == 0 && st[0].known_num_levels == 0);               // Stack(0, false), id 0
    EXPECT(st[1].stack_id == 1 && st[1].tags[0] == 2 && st[1].known_num_levels == 0);
    EXPECT(st[2].stack_id == 2 && st[2].num_set_tags == 0 && st[2].known_num_levels == 1);
    EXPECT(st[3].stack_id == 3 && st[3].tags[0] == 4 && st[4].stack_id == 4 && st[4].tags[0] == 1);
    EXPECT(slicer.getStats().num_slices_completed == 5 && slicer.getLastTimeStamp() == 80);
    auto& islicer = *(*slicer.getSlices())[0];
    EXPECT(islicer.numActiveSlices() == 4);                                       // :108
    auto [s, n] = islicer.makeIntervalIntersectingSlices(0, 25, 30, true, true);  // [25, 55)
    EXPECT(s != nullptr && s->size() == 3 && n == 2);
    EXPECT(s->at(0) == makeSlice(25, 5, 1, TransitionType::Analysis, TransitionType::ThreadPreempted));
    EXPECT(s->at(1) == makeSlice(30, 20, 3, TransitionType::ThreadYield, TransitionType::ThreadPreempted));
    EXPECT(s->at(2) == makeSlice(50, 5, 2, TransitionType::ThreadPreempted, TransitionType::Analysis));
    // the next window continues: thread 4 comes back after the idle stretch that began at 80
    events[0] = {mk(95, kIn, 4), mk(120, kOutPreempt, ~0u)};
    EXPECT(slicer.processWindow(events) == 2);
    EXPECT(islicer.numActiveSlices() == 6 && slicer.getStats().num_slices_completed == 7);
    EXPECT(slicer.getTagStackStatsDump().size() == 5 && slicer.getTagStackStatsDump()[3].num_swins == 2);
  }
  std::puts("SHIM OK");
  return 0;
}
