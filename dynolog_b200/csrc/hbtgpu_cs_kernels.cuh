// hbtgpu_cs_kernels.cuh -- ContextSwitch-mode decode on the GPU.
//
// Replaces, for captured ring images, the device-independent part of
//   PerfEventsGroup<CpuThreadSwitchGenerator, mode::ContextSwitch>::consume
//       hbt/src/perf_event/PerfEventsGroup.h:1394-1589
//   CpuThreadSwitchGenerator::handleRecord{Lost,Comm,Exit,Fork,SwitchCpuWide}, pendingRecordLost_
//       hbt/src/perf_event/PerCpuThreadSwitchGenerator.h:219-446
// i.e. records -> 16-byte tagstack::Event stream (hbt/src/tagstack/Event.h:28-56) + counters.
// The tid -> virtual-id dictionary (PerThreadStats :52-135) is an order-dependent state machine
// shared by all CPUs; it stays on the host (hbtgpu_cs_assign_vids), so events leave the GPU with
// the raw tid in `tag` and the records that drive the dictionary besides SWITCH-in (LOST, COMM,
// EXIT, FORK) are exported as "thread ops".
//
// What an event depends on: its own record, and of the previous *handled* record (LOST, COMM,
// EXIT, FORK, SWITCH_CPU_WIDE) only the type (a LOST makes the next record write
// WriteErrorsEnd first) and the time (a LOST writes WriteErrorsStart(last_tstamp_ + 1)).  So
// the same scheme as the Sampling path applies: CTA per chunk, chunk entry and predecessor
// speculated from a halo, proven by k_verify, exact walk (k_cs_resolve) as the fallback.
//
// In a context-switch capture nearly every record is a 32-byte SWITCH_CPU_WIDE.  A tile is
// processed as runs of those -- every thread tests the header at entry + 32*j; equality IS the
// chain link, so the first mismatch ends the proven run -- and the records in between (LOST,
// COMM, EXIT, FORK, THROTTLE, anything else) are followed one by one by thread 0.
//
// Two passes over the images: COUNT (events / ops per chunk, counters, speculation) and, after
// a scan of the per-chunk counts, EMIT (events written to their final, per-CPU contiguous
// position).  Both stream 16 KiB tiles through shared memory with TMA bulk copies.
#pragma once

#include "hbtgpu_kernels.cuh"

namespace hbtgpu {

enum : uint32_t { REC_COMM = 3, REC_EXIT = 4, REC_FORK = 7, REC_SWITCH_CPU_WIDE = 15 };
constexpr uint32_t MISC_SWITCH_OUT = 1u << 13, MISC_SWITCH_OUT_PREEMPT = 1u << 14;
constexpr uint32_t CS_KNA = 0xFFFFFFFFu;       // tagstack::kNA
constexpr uint32_t CS_INVALID_LEVEL = 255;     // tagstack::kInvalidLevel

constexpr int CS_TILE_BYTES = 16384;
constexpr int CS_OV_BYTES = 64;      // a 32-byte SWITCH whose header is in the tile is fully staged
constexpr int CS_STAGE_WORDS = (CS_TILE_BYTES + CS_OV_BYTES) / 8;
constexpr int CS_HALO_BYTES = 1024;  // <= NT * 8: one halo word per thread
constexpr int CS_RPT = (CS_TILE_BYTES / 32 + NT - 1) / NT;  // SWITCH records per thread and tile
// A stretch of a tile that is dense with other records is cut into elements -- runs of 32-byte
// SWITCH records and single other records -- by one warp, and the elements are then handled side
// by side (cs_mixed_stretch).
constexpr int CS_MAXE = 96;                 // elements per stretch
constexpr uint32_t CS_MIXED_AFTER = 2;      // other records in a tile before the rest of it goes that way
constexpr uint16_t CS_E_OTHER = 0xFFFF;     // CsSmem::e_len of an element that is one other record
static_assert(CS_HALO_BYTES <= NT * 8, "scout takes one halo word per thread");

// per-stream counters (u64 slots of Params::counters)
enum {
  CSC_LOST = 0, CSC_COMM, CSC_EXIT, CSC_THROTTLE, CSC_UNTHROTTLE, CSC_FORK, CSC_SWITCH, CSC_RECORDS,
  CSC_EVENTS, CSC_OPS, CSC_ERR = 12, CSC_ERR_OFF = 13, CSC_LAST_TS = 14
};
static_assert(CSC_OPS < CSC_ERR && CSC_LAST_TS < C_SLOTS, "counter slots overlap");

struct CsOut {
  const hbtgpu_cs_state* state_in;   // [stream]
  hbtgpu_cs_state* state_out;        // [stream]
  hbtgpu_cs_event* events;           // all streams; stream s at ev_base[s]
  uint32_t* pids;                    // parallel to events: sample_id.pid / record pid
  hbtgpu_cs_op* ops;                 // all streams; stream s at op_base[s]
  uint64_t* ev_base;                 // [stream + 1]
  uint64_t* op_base;                 // [stream + 1]
  uint32_t level;                    // tid_level
};

struct CsSmem {
  uint64_t data[2][CS_STAGE_WORDS];
  uint64_t bar[2];
  uint64_t pred_ts, pred_off;   // previous handled record: time, arena offset (or PRED_CARRY_IN)
  uint64_t entry;               // speculated chunk entry (arena offset)
  unsigned long long lost_sum;
  uint32_t pred_lost;           // ... and whether it was a LOST (in_write_errors_)
  uint32_t n_ev, n_ops;         // events / ops of this chunk so far
  uint32_t seg_k, ew, bad, wl;
  uint32_t wmin[NWARP];         // per warp: first record of the run that is no 32-byte SWITCH
  // cs_mixed_stretch: the elements (word position in the tile, records of a run or CS_E_OTHER),
  // what each leaves behind (E_*), its events / ops (counts, then exclusive prefix sums)
  uint16_t e_pos[CS_MAXE], e_len[CS_MAXE];
  uint8_t e_fl[CS_MAXE];
  uint32_t e_ev[CS_MAXE + 1], e_op[CS_MAXE + 1];
  uint64_t e_ts[CS_MAXE];
  uint32_t e_n, e_end;
  uint32_t cnt[16];
};

__device__ __forceinline__ bool cs_is_switch32(uint64_t h) {
  return (h & 0xFFFF0000FFFFFFFFull) == (static_cast<uint64_t>(REC_SWITCH_CPU_WIDE) | (32ull << 48));
}
__device__ __forceinline__ bool cs_handled(uint32_t type) {
  return type == REC_LOST || type == REC_COMM || type == REC_EXIT || type == REC_FORK ||
         type == REC_SWITCH_CPU_WIDE;
}
// smallest size at which the handler's fixed-offset reads stay inside the record (fence)
__device__ __forceinline__ uint32_t cs_min_size(uint32_t type) {
  return type == REC_LOST ? 40u : (type == REC_COMM || type == REC_SWITCH_CPU_WIDE) ? 32u
       : (type == REC_EXIT || type == REC_FORK) ? 48u : 8u;
}
// time the handlers take from a record (r.sample_id.tstamp, or r.time for EXIT/FORK)
__device__ __forceinline__ uint64_t cs_time(const uint64_t* r, uint32_t type, uint32_t size) {
  return type == REC_LOST ? r[4] : type == REC_COMM ? r[size / 8 - 1] : r[3];  // EXIT/FORK/SWITCH: word 3
}
__device__ __forceinline__ void cs_store_event(const CsOut& O, uint64_t idx, uint64_t ts, uint32_t type,
                                               uint32_t level, uint32_t cu, uint32_t tag, uint32_t pid) {
  // one 16-byte store: {u64 tstamp; u8 type; u8 level; u16 comp_unit_id; u32 tag}
  const uint64_t lo = static_cast<uint64_t>(type) | (static_cast<uint64_t>(level) << 8) |
                      (static_cast<uint64_t>(cu) << 16) | (static_cast<uint64_t>(tag) << 32);
  *reinterpret_cast<ulonglong2*>(&O.events[idx]) = make_ulonglong2(ts, lo);
  O.pids[idx] = pid;
}

// One record that is not a 32-byte SWITCH, handled by thread 0 exactly as consume() + the
// handlers do.  `r` points at the record (shared memory, or the arena for one longer than the tile).
// Returns its size in bytes, or 0 if it trips a fence (size 0 / not a multiple of 8 / past the
// image end / too short for its handler).
template <bool EMIT>
__device__ __forceinline__ uint32_t cs_odd_record(CsSmem& s, const CsOut& O, const uint64_t* r, uint64_t abs_off,
                                                  uint64_t room, uint32_t cu, uint64_t ev0, uint64_t op0) {
  const uint64_t h = r[0];
  const uint32_t type = static_cast<uint32_t>(h), misc = static_cast<uint32_t>(h >> 32) & 0xFFFFu;
  const uint32_t size = static_cast<uint32_t>(h >> 48);
  if (size < 8 || (size & 7) || size > room || size < cs_min_size(type)) return 0;
  s.cnt[CSC_RECORDS] += 1;
  if (type == REC_THROTTLE) s.cnt[CSC_THROTTLE] += 1;
  if (type == REC_UNTHROTTLE) s.cnt[CSC_UNTHROTTLE] += 1;
  if (!cs_handled(type)) return size;  // SAMPLE/READ/AUX/ITRACE_START are void in this mode; unknown: skipped
  const uint64_t ts = cs_time(r, type, size);
  if (s.pred_lost) {  // pendingRecordLost_ :424-435
    if (EMIT) cs_store_event(O, ev0 + s.n_ev, ts, 8, CS_INVALID_LEVEL, cu, CS_KNA, 0);
    s.n_ev += 1;
  }
  const uint64_t last = s.pred_lost ? ts : s.pred_ts;  // last_tstamp_ when the handler body runs
  uint32_t etype = 0, tag = CS_KNA, pid = 0, level = O.level;
  uint64_t ets = ts;
  hbtgpu_cs_op op{};
  bool has_op = true;
  op.type = type;
  op.misc = misc;
  op.tstamp = ts;
  if (type == REC_LOST) {  // :219-253
    etype = 7;
    ets = last + 1;
    level = CS_INVALID_LEVEL;
    s.lost_sum += r[2];
    op.pid = static_cast<int32_t>(r[3]);
    op.tid = static_cast<int32_t>(r[3] >> 32);
  } else if (type == REC_COMM) {  // :255-292
    s.cnt[CSC_COMM] += 1;
    op.pid = static_cast<int32_t>(r[1]);
    op.tid = static_cast<int32_t>(r[1] >> 32);
    op.name_len = size - 32;
    if (EMIT) {
      const uint32_t nw = min(op.name_len / 8, 2u);
      for (uint32_t i = 0; i < nw; ++i) reinterpret_cast<uint64_t*>(op.name)[i] = r[2 + i];
    }
  } else if (type == REC_EXIT || type == REC_FORK) {  // :294-372
    etype = type == REC_EXIT ? 3 : 2;
    s.cnt[type == REC_EXIT ? CSC_EXIT : CSC_FORK] += 1;
    op.pid = static_cast<int32_t>(r[1]);
    op.ppid = static_cast<int32_t>(r[1] >> 32);
    op.tid = static_cast<int32_t>(r[2]);
    op.ptid = static_cast<int32_t>(r[2] >> 32);
    tag = static_cast<uint32_t>(op.tid);
    pid = static_cast<uint32_t>(op.pid);
  } else {  // SWITCH_CPU_WIDE of an unusual size :374-414
    has_op = false;
    s.cnt[CSC_SWITCH] += 1;
    if (misc & MISC_SWITCH_OUT) {
      etype = (misc & MISC_SWITCH_OUT_PREEMPT) ? 5 : 6;
    } else {
      etype = 4;
      pid = static_cast<uint32_t>(r[2]);
      tag = static_cast<uint32_t>(r[2] >> 32);
    }
  }
  if (has_op) {
    op.event_index = s.n_ev;  // chunk-relative here; EMIT adds the chunk's base
    if (EMIT) {
      op.event_index += ev0 - O.ev_base[cu];
      O.ops[op0 + s.n_ops] = op;
    }
    s.n_ops += 1;
  }
  if (etype) {
    if (EMIT) cs_store_event(O, ev0 + s.n_ev, ets, etype, level, cu, tag, pid);
    s.n_ev += 1;
  }
  s.pred_ts = ts;
  s.pred_off = abs_off;
  s.pred_lost = type == REC_LOST;
  return size;
}

enum : uint8_t { E_HANDLED = 1, E_LOST = 2, E_EVENT = 4, E_OP = 8 };

// The stretch of a tile from word `ew` on, as elements handled side by side.  What the walk of
// consume() + the handlers does record after record (cs_odd_record, the SWITCH runs of cs_chunk)
// depends, for one record, only on the handled record before it (was it a LOST, its time) and on
// how many events / ops came before: warp 0 finds the elements (the one serial part: a record's
// size says where the next one starts -- a header read and a vote per element), every element
// looks back for its predecessor, a scan gives the output positions, and then all other records
// and all runs are written at once.  Stops in front of anything it does not take (a record that
// trips a fence or is not fully staged, the CS_MAXE-th element): the caller goes on from the
// word returned, record by record if need be.  Block-uniform; returns the new chain position.
template <bool EMIT>
__device__ __forceinline__ uint32_t cs_mixed_stretch(CsSmem& s, const CsOut& O, const uint64_t* buf, uint32_t wt, uint32_t ew,
                                                     uint32_t staged, uint64_t tile_abs, uint64_t limit_abs, uint32_t cu,
                                                     uint64_t ev0, uint64_t op0) {
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // ---- 1. the elements ----
  if (warp == 0) {
    uint32_t w = ew, n = 0;
    bool other_next = true;  // (the caller found that the record at ew is no 32-byte SWITCH)
    while (w < wt && n < CS_MAXE) {
      if (!other_next) {
        // a run of SWITCH records from w on: 32 at a step
        uint32_t run = 0;
        bool tile_end = false;
        for (;;) {
          const uint32_t Kall = (wt - (w + 4 * run) + 3) / 4;
          const uint32_t p = w + 4 * run + 4 * lane;
          const bool ok = lane < Kall && cs_is_switch32(buf[p]) && tile_abs + 8ull * p + 32 <= limit_abs;
          const uint32_t nok = ~__ballot_sync(0xffffffffu, ok);
          const uint32_t k = nok ? static_cast<uint32_t>(__ffs(nok) - 1) : 32u;
          run += k;
          tile_end = k == Kall;
          if (k < 32 || tile_end || run >= 0xFF00u) break;
        }
        if (run) {
          if (lane == 0) {
            s.e_pos[n] = static_cast<uint16_t>(w);
            s.e_len[n] = static_cast<uint16_t>(run);
          }
          ++n;
          w += 4 * run;
          if (tile_end || run >= 0xFF00u) continue;  // (the loop condition ends the walk, or the run goes on)
          if (n == CS_MAXE) break;
        }
        // the record at w is no 32-byte SWITCH (or runs past the end of the image)
      }
      other_next = false;
      // another record at w: does the walk take it?  (the fences of cs_odd_record)
      const uint64_t h = buf[w];
      const uint32_t type = static_cast<uint32_t>(h), size = static_cast<uint32_t>(h >> 48);
      const uint64_t room = limit_abs - (tile_abs + 8ull * w);
      if (size < 8 || (size & 7) || size > room || size < cs_min_size(type) || w + size / 8 > staged) break;
      if (lane == 0) {
        s.e_pos[n] = static_cast<uint16_t>(w);
        s.e_len[n] = CS_E_OTHER;
      }
      ++n;
      w += size / 8;
    }
    if (lane == 0) {
      s.e_n = n;
      s.e_end = w;
    }
  }
  __syncthreads();
  const uint32_t n = s.e_n;
  if (n == 0) return ew;
  // ---- 2. what every element leaves behind ----
  const uint32_t k = threadIdx.x;
  uint32_t pos = 0, len = 0, type = 0, size = 0, misc = 0, fl = 0;
  uint64_t ts = 0;
  if (k < n) {
    pos = s.e_pos[k];
    len = s.e_len[k];
    if (len != CS_E_OTHER) {
      fl = E_HANDLED;
      ts = buf[pos + 4 * (len - 1) + 3];  // the run's last record
    } else {
      const uint64_t h = buf[pos];
      type = static_cast<uint32_t>(h);
      misc = static_cast<uint32_t>(h >> 32) & 0xFFFFu;
      size = static_cast<uint32_t>(h >> 48);
      if (cs_handled(type)) {
        ts = cs_time(buf + pos, type, size);
        fl = E_HANDLED | (type == REC_LOST ? E_LOST : 0) | (type != REC_COMM ? E_EVENT : 0) |
             (type != REC_SWITCH_CPU_WIDE ? E_OP : 0);
      }
    }
    s.e_fl[k] = static_cast<uint8_t>(fl);
    s.e_ts[k] = ts;
  }
  __syncthreads();
  // ---- 3. the handled record before it, events and ops it writes ----
  uint32_t pl = 0;      // in_write_errors_ when the element is reached
  uint64_t pts = 0;     // last_tstamp_
  if (k < n) {
    int j = static_cast<int>(k) - 1;
    while (j >= 0 && !(s.e_fl[j] & E_HANDLED)) --j;
    if (j >= 0) {
      pl = (s.e_fl[j] & E_LOST) ? 1u : 0u;
      pts = s.e_ts[j];
    } else {
      pl = s.pred_lost;
      pts = s.pred_ts;
    }
    uint32_t cev = 0, cop = 0;
    if (fl & E_HANDLED) {
      cev = pl + (len != CS_E_OTHER ? len : ((fl & E_EVENT) ? 1u : 0u));
      cop = (fl & E_OP) ? 1u : 0u;
    }
    s.e_ev[k] = cev;
    s.e_op[k] = cop;
  }
  __syncthreads();
  if (warp == 0) {  // exclusive prefix sums over the (at most 96) elements, three per lane
    uint32_t a[3], b[3], sa = 0, sb = 0;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const uint32_t e = 3 * lane + i;
      a[i] = e < n ? s.e_ev[e] : 0;
      b[i] = e < n ? s.e_op[e] : 0;
      sa += a[i];
      sb += b[i];
    }
    uint32_t xa = sa, xb = sb;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t ya = __shfl_up_sync(0xffffffffu, xa, o), yb = __shfl_up_sync(0xffffffffu, xb, o);
      if (lane >= static_cast<uint32_t>(o)) {
        xa += ya;
        xb += yb;
      }
    }
    uint32_t ra = xa - sa, rb = xb - sb;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const uint32_t e = 3 * lane + i;
      if (e < n) {
        s.e_ev[e] = ra;
        s.e_op[e] = rb;
      }
      ra += a[i];
      rb += b[i];
    }
    if (lane == 31) {
      s.e_ev[CS_MAXE] = xa;  // totals
      s.e_op[CS_MAXE] = xb;
    }
  }
  __syncthreads();
  const uint32_t nev0 = s.n_ev, nop0 = s.n_ops;
  // ---- 4. the other records: counters, events, ops (one thread each) ----
  if (k < n && len == CS_E_OTHER) {
    const uint64_t* r = buf + pos;
    atomicAdd(&s.cnt[CSC_RECORDS], 1u);
    if (type == REC_THROTTLE) atomicAdd(&s.cnt[CSC_THROTTLE], 1u);
    if (type == REC_UNTHROTTLE) atomicAdd(&s.cnt[CSC_UNTHROTTLE], 1u);
    if (fl & E_HANDLED) {
      const uint32_t e_at = nev0 + s.e_ev[k];
      if (pl && EMIT) cs_store_event(O, ev0 + e_at, ts, 8, CS_INVALID_LEVEL, cu, CS_KNA, 0);  // pendingRecordLost_ :424-435
      const uint64_t last = pl ? ts : pts;  // last_tstamp_ when the handler body runs
      uint32_t etype = 0, tag = CS_KNA, pid = 0, level = O.level;
      uint64_t ets = ts;
      hbtgpu_cs_op op{};
      op.type = type;
      op.misc = misc;
      op.tstamp = ts;
      if (type == REC_LOST) {  // :219-253
        etype = 7;
        ets = last + 1;
        level = CS_INVALID_LEVEL;
        atomicAdd(&s.lost_sum, static_cast<unsigned long long>(r[2]));
        op.pid = static_cast<int32_t>(r[3]);
        op.tid = static_cast<int32_t>(r[3] >> 32);
      } else if (type == REC_COMM) {  // :255-292
        atomicAdd(&s.cnt[CSC_COMM], 1u);
        op.pid = static_cast<int32_t>(r[1]);
        op.tid = static_cast<int32_t>(r[1] >> 32);
        op.name_len = size - 32;
        if (EMIT) {
          const uint32_t nw = min(op.name_len / 8, 2u);
          for (uint32_t i = 0; i < nw; ++i) reinterpret_cast<uint64_t*>(op.name)[i] = r[2 + i];
        }
      } else if (type == REC_EXIT || type == REC_FORK) {  // :294-372
        etype = type == REC_EXIT ? 3 : 2;
        atomicAdd(&s.cnt[type == REC_EXIT ? CSC_EXIT : CSC_FORK], 1u);
        op.pid = static_cast<int32_t>(r[1]);
        op.ppid = static_cast<int32_t>(r[1] >> 32);
        op.tid = static_cast<int32_t>(r[2]);
        op.ptid = static_cast<int32_t>(r[2] >> 32);
        tag = static_cast<uint32_t>(op.tid);
        pid = static_cast<uint32_t>(op.pid);
      } else {  // SWITCH_CPU_WIDE of an unusual size :374-414
        atomicAdd(&s.cnt[CSC_SWITCH], 1u);
        if (misc & MISC_SWITCH_OUT) {
          etype = (misc & MISC_SWITCH_OUT_PREEMPT) ? 5 : 6;
        } else {
          etype = 4;
          pid = static_cast<uint32_t>(r[2]);
          tag = static_cast<uint32_t>(r[2] >> 32);
        }
      }
      if (EMIT) {
        if (fl & E_OP) {
          op.event_index = e_at + pl + ev0 - O.ev_base[cu];
          O.ops[op0 + nop0 + s.e_op[k]] = op;
        }
        if (etype) cs_store_event(O, ev0 + e_at + pl, ets, etype, level, cu, tag, pid);
      }
    }
  }
  // ---- 5. the runs: a warp per run, 32 records a step ----
  for (uint32_t e = warp; e < n; e += NWARP) {
    const uint32_t rl = s.e_len[e];
    if (rl == CS_E_OTHER) continue;
    const uint32_t rp = s.e_pos[e];
    if (lane == 0) {
      atomicAdd(&s.cnt[CSC_SWITCH], rl);
      atomicAdd(&s.cnt[CSC_RECORDS], rl);
    }
    if (EMIT) {
      // (its predecessor: as in step 3)
      int j = static_cast<int>(e) - 1;
      while (j >= 0 && !(s.e_fl[j] & E_HANDLED)) --j;
      const uint32_t rpl = j >= 0 ? ((s.e_fl[j] & E_LOST) ? 1u : 0u) : s.pred_lost;
      const uint32_t e_at = nev0 + s.e_ev[e];
      for (uint32_t q = lane; q < rl; q += 32) {
        const uint32_t p = rp + 4 * q;
        const uint64_t h = buf[p];
        const uint32_t m2 = static_cast<uint32_t>(h >> 32) & 0xFFFFu;
        const uint64_t id = buf[p + 2], t2 = buf[p + 3];
        const bool out = (m2 & MISC_SWITCH_OUT) != 0;
        const uint32_t et = out ? ((m2 & MISC_SWITCH_OUT_PREEMPT) ? 5u : 6u) : 4u;
        if (q == 0 && rpl) cs_store_event(O, ev0 + e_at, t2, 8, CS_INVALID_LEVEL, cu, CS_KNA, 0);
        cs_store_event(O, ev0 + e_at + rpl + q, t2, et, O.level, cu, out ? CS_KNA : static_cast<uint32_t>(id >> 32),
                       out ? 0u : static_cast<uint32_t>(id));
      }
    }
  }
  __syncthreads();  // everybody has read the state the stretch started from
  // ---- 6. the state behind the stretch ----
  if (threadIdx.x == 0) {
    int j = static_cast<int>(n) - 1;
    while (j >= 0 && !(s.e_fl[j] & E_HANDLED)) --j;
    if (j >= 0) {
      const uint32_t jl = s.e_len[j], jp = s.e_pos[j];
      s.pred_ts = s.e_ts[j];
      s.pred_lost = (s.e_fl[j] & E_LOST) ? 1u : 0u;
      s.pred_off = tile_abs + 8ull * (jl == CS_E_OTHER ? jp : jp + 4 * (jl - 1));
    }
    s.n_ev = nev0 + s.e_ev[CS_MAXE];
    s.n_ops = nop0 + s.e_op[CS_MAXE];
  }
  const uint32_t end = s.e_end;
  __syncthreads();
  return end;
}

// state of the generator after the handled record at arena offset `off`
__device__ __forceinline__ void cs_pred_state(const Params& P, const CsOut& O, uint32_t stream, uint64_t off,
                                              uint64_t& ts, uint32_t& lost) {
  if (off == PRED_CARRY_IN) {
    ts = O.state_in[stream].last_tstamp;
    lost = O.state_in[stream].in_write_errors != 0;
  } else {
    const uint64_t* r = reinterpret_cast<const uint64_t*>(P.arena + off);
    const uint32_t type = static_cast<uint32_t>(r[0]), size = static_cast<uint32_t>(r[0] >> 48);
    ts = cs_time(r, type, size);
    lost = type == REC_LOST;
  }
}

constexpr uint32_t WHY_CS_NO_TRIPLE = 20, WHY_CS_HALO_WALK = 21, WHY_CS_SEAM = 22, WHY_CS_RECORD = 23,
                   WHY_CS_PAST_END = 24;

// One chunk, one pass (COUNT or EMIT).  Called once per CTA by k_cs, twice (COUNT, look-back,
// EMIT) by k_cs_fused.  ev_base / op_base: where this chunk's events / ops go (EMIT).
template <bool SPEC, bool EMIT>
__device__ __forceinline__ void cs_chunk(const Params& P, const CsOut& O, CsSmem& s, const uint32_t c,
                                         const uint64_t ev_base, const uint64_t op_base) {
  const ChunkDesc cd = P.chunks[c];
  ChunkIO& io = P.io[c];
  const uint32_t cu = cd.stream;
  if (threadIdx.x == 0) {
    mbar_init(&s.bar[0], 1);
    mbar_init(&s.bar[1], 1);
    fence_barrier_init();
    s.lost_sum = 0;
    s.n_ev = 0;
    s.n_ops = 0;
    s.bad = 0;
    s.wl = 0;
  }
  if (threadIdx.x < 16) s.cnt[threadIdx.x] = 0;
  __syncthreads();

  uint32_t phase_bits = 0;  // parity of the two TMA barriers
  const uint64_t chunk_abs = cd.img_base + cd.begin;
  const uint64_t img_end_abs = cd.img_base + cd.img_len;
  uint64_t limit_abs = img_end_abs;
  const uint32_t avail_rel = static_cast<uint32_t>(umin64((img_end_abs - chunk_abs + 15ull) & ~15ull, 0xFFFFFFF0ull));
  const uint8_t* const chunk_ptr = P.arena + chunk_abs;
  auto issue = [&](uint32_t tile_rel, uint32_t b) {
    const uint32_t bytes = min(avail_rel - tile_rel, static_cast<uint32_t>(CS_TILE_BYTES + CS_OV_BYTES));
    mbar_expect_tx(&s.bar[b], bytes);
    tma_load_1d(s.data[b], chunk_ptr + tile_rel, bytes, &s.bar[b]);
  };

  // ---- entry and predecessor ----
  uint64_t entry_abs = chunk_abs;
  uint32_t boff = 0, n_issued = 0, n_waited = 0;
  if (SPEC) {
    uint64_t pred = PRED_CARRY_IN;
    if (cd.halo_len != 0) {
      // Scout: the last three consecutive 32-byte SWITCH headers in the halo mark a point of
      // the chain (a guess: k_verify proves it); thread 0 follows the chain from there to the
      // chunk, remembering the last handled record.
      const uint32_t hw = cd.halo_len / 8;
      if (threadIdx.x == 0) {
        const uint32_t bytes = (cd.halo_len + 15u) & ~15u;
        mbar_expect_tx(&s.bar[0], bytes);
        tma_load_1d(s.data[0], P.arena + cd.halo_off, bytes, &s.bar[0]);
        issue(0, 1);
      }
      boff = 1;
      n_issued = 1;
      mbar_wait(&s.bar[0], phase_bits & 1u);
      phase_bits ^= 1u;
      const uint64_t* hb = s.data[0];
      const uint32_t w = threadIdx.x;
      const bool m = w >= 8 && w < hw && cs_is_switch32(hb[w]) && cs_is_switch32(hb[w - 4]) &&
                     cs_is_switch32(hb[w - 8]);
      const uint32_t bal = __ballot_sync(0xffffffffu, m);
      if ((threadIdx.x & 31) == 0 && bal) atomicMax(&s.wl, (threadIdx.x & ~31u) + (31 - __clz(bal)) + 1);
      __syncthreads();
      if (threadIdx.x == 0) {
        uint32_t bad = 0;
        if (cd.flags & CH_HALO_EXACT_START) s.wl = 1;  // the halo starts at the image start
        if (s.wl == 0) {
          bad = WHY_CS_NO_TRIPLE;
        } else {
          uint32_t p = s.wl - 1;
          uint64_t pr = (cd.flags & CH_HALO_EXACT_START) ? PRED_UNKNOWN : 0;
          while (p < hw) {
            const uint64_t h = hb[p];
            const uint32_t type = static_cast<uint32_t>(h), size = static_cast<uint32_t>(h >> 48);
            if (size < 8 || (size & 7)) {
              bad = WHY_CS_HALO_WALK;
              break;
            }
            if (cs_handled(type)) pr = cd.halo_off + 8ull * p;
            p += size / 8;
          }
          // a halo that starts at the image start may hold no handled record at all: then the
          // predecessor is whatever preceded the image, which only the exact walk knows
          if (!bad && pr == PRED_UNKNOWN) bad = WHY_CS_NO_TRIPLE;
          pred = pr;
          entry_abs = cd.halo_off + 8ull * p;
          if ((cd.flags & CH_FIRST_OF_IMAGE) && !bad) {
            if (entry_abs != cd.halo_off + cd.halo_len) bad = WHY_CS_SEAM;
            entry_abs = chunk_abs;
          }
        }
        s.bad = bad;
        s.pred_off = pred;
        s.entry = entry_abs;
      }
      __syncthreads();
      pred = s.pred_off;
      entry_abs = s.entry;
    }
    if (threadIdx.x == 0) {
      io.in_entry = entry_abs;
      io.in_pred = pred;
      io.in_gap = 0;
      io.limit = ~0ull;
      s.pred_off = pred;
    }
  } else {
    entry_abs = io.in_entry;
    limit_abs = min(limit_abs, io.limit);
    if (threadIdx.x == 0) s.pred_off = io.in_pred;
  }
  __syncthreads();
  uint32_t bad = s.bad;
  if (threadIdx.x == 0) {
    uint64_t ts = 0;
    uint32_t lost = 0;
    if (!bad && s.pred_off != PRED_UNKNOWN) cs_pred_state(P, O, cu, s.pred_off, ts, lost);
    s.pred_ts = ts;
    s.pred_lost = lost;
  }
  __syncthreads();

  // ---- tiles ----
  const uint64_t ev0 = EMIT ? ev_base : 0, op0 = EMIT ? op_base : 0;
  const uint64_t chunk_end_abs = min(cd.img_base + cd.end, limit_abs);
  const bool any_work = !bad && entry_abs < chunk_end_abs;
  const uint32_t chunk_len = any_work ? static_cast<uint32_t>(chunk_end_abs - chunk_abs) : 0;
  const uint32_t limit_rel = static_cast<uint32_t>(umin64(limit_abs - chunk_abs, 0xFFFFFFF0ull));
  uint32_t next_rel = any_work ? static_cast<uint32_t>(entry_abs - chunk_abs) : 0;
  if (any_work && n_issued == 0) {
    if (threadIdx.x == 0) issue(0, boff);
    n_issued = 1;
  }
  for (uint32_t tile_rel = 0, b = boff; tile_rel < chunk_len; tile_rel += CS_TILE_BYTES, b ^= 1) {
    if (tile_rel + CS_TILE_BYTES < chunk_len) {
      if (threadIdx.x == 0) issue(tile_rel + CS_TILE_BYTES, b ^ 1);
      ++n_issued;
    }
    mbar_wait(&s.bar[b], (phase_bits >> b) & 1u);
    phase_bits ^= 1u << b;
    ++n_waited;
    const uint64_t* buf = s.data[b];
    const uint32_t wt = min(static_cast<uint32_t>(CS_TILE_BYTES), chunk_len - tile_rel) / 8;
    if (next_rel < tile_rel + 8 * wt) {
      uint32_t ew = (next_rel - tile_rel) / 8;  // chain position, tile-relative words
      bool mixed = false;  // the tile has shown another record: look one record per thread ahead only
      uint32_t n_other = 0;  // other records met in this tile, one by one
      for (;;) {
        // ---- a run of 32-byte SWITCH records from ew on ----
        const uint32_t Kall = (wt - ew + 3) / 4;
        // (a tile with other records in it is walked in runs that end at the next of them: testing
        // the whole rest of the tile for every run would be quadratic)
        const uint32_t Kscan = mixed ? min(Kall, static_cast<uint32_t>(NT)) : Kall;
        uint32_t first_bad = 0xFFFFFFFFu;
        uint64_t hd[CS_RPT];
#pragma unroll
        for (int i = 0; i < CS_RPT; ++i) {
          const uint32_t j = threadIdx.x + i * NT;
          hd[i] = 0;
          if (j < Kscan) {
            hd[i] = buf[ew + 4 * j];
            // the last record of the image may be cut short: it is not part of the run
            const bool whole = chunk_abs + tile_rel + 8ull * (ew + 4 * j) + 32 <= limit_abs;
            if (!(cs_is_switch32(hd[i]) && whole)) first_bad = min(first_bad, j);
          }
        }
        // where the run ends: one barrier (a warp's minimum by __reduce_min_sync, the block's from
        // the six words) -- a tile with many other records pays this once per record
        {
          const uint32_t wm = __reduce_min_sync(0xffffffffu, first_bad);
          if ((threadIdx.x & 31) == 0) s.wmin[threadIdx.x >> 5] = wm;
        }
        __syncthreads();
        uint32_t K = Kscan;
#pragma unroll
        for (int w = 0; w < NWARP; ++w) K = min(K, s.wmin[w]);
        uint32_t e_base_of_run = 0;
        if (K > 0) {
          const uint32_t pl = s.pred_lost;      // read before thread 0 updates it below
          const uint32_t e_base = s.n_ev + pl;  // chunk-relative index of record 0's own event
          e_base_of_run = e_base;
          if (EMIT) {
#pragma unroll
            for (int i = 0; i < CS_RPT; ++i) {
              const uint32_t j = threadIdx.x + i * NT;
              if (j < K) {
                const uint32_t p = ew + 4 * j;
                const uint32_t misc = static_cast<uint32_t>(hd[i] >> 32) & 0xFFFFu;
                const uint64_t id = buf[p + 2], ts = buf[p + 3];
                const bool out = (misc & MISC_SWITCH_OUT) != 0;
                const uint32_t et = out ? ((misc & MISC_SWITCH_OUT_PREEMPT) ? 5u : 6u) : 4u;
                if (j == 0 && pl) cs_store_event(O, ev0 + s.n_ev, ts, 8, CS_INVALID_LEVEL, cu, CS_KNA, 0);
                cs_store_event(O, ev0 + e_base + j, ts, et, O.level, cu,
                               out ? CS_KNA : static_cast<uint32_t>(id >> 32), out ? 0u : static_cast<uint32_t>(id));
              }
            }
          }
        }
        __syncthreads();  // everyone has read pred_lost / n_ev / wmin
        if (K > 0) {
          if (threadIdx.x == 0) {
            const uint32_t p = ew + 4 * (K - 1);
            s.pred_ts = buf[p + 3];
            s.pred_off = chunk_abs + tile_rel + 8ull * p;
            s.pred_lost = 0;
            s.n_ev = e_base_of_run + K;
            s.cnt[CSC_SWITCH] += K;
            s.cnt[CSC_RECORDS] += K;
          }
          ew += 4 * K;
        }
        if (K == Kall) break;  // ran to the end of the tile (block-uniform)
        if (K == Kscan) continue;  // ... or to the end of what was looked at: the run goes on
        mixed = true;
        if (n_other >= CS_MIXED_AFTER) {
          // a tile dense with other records: the rest of it as elements handled side by side
          const uint32_t staged = min(avail_rel - tile_rel, static_cast<uint32_t>(CS_TILE_BYTES + CS_OV_BYTES)) / 8;
          const uint32_t nw = cs_mixed_stretch<EMIT>(s, O, buf, wt, ew, staged, chunk_abs + tile_rel, limit_abs, cu, ev0, op0);
          if (nw != ew) {
            ew = nw;
            if (ew >= wt) break;
            continue;
          }
          // (it took nothing: the record at ew trips a fence or is not staged -- thread 0 looks at it)
        }
        ++n_other;
        // ---- the record at ew is something else: thread 0 follows it (in the same breath as the
        // run's bookkeeping above: no barrier in between) ----
        if (threadIdx.x == 0) {
          const uint64_t abs_off = chunk_abs + tile_rel + 8ull * ew;
          const uint64_t room = limit_abs - abs_off;
          uint32_t sz = 0;
          if (room >= 8) {
            // read it from the staged tile when it lies inside (nearly always), else from the arena
            const uint32_t staged = min(avail_rel - tile_rel, static_cast<uint32_t>(CS_TILE_BYTES + CS_OV_BYTES)) / 8;
            const uint32_t rw = static_cast<uint32_t>(buf[ew] >> 51);  // size / 8
            const uint64_t* r = ew + rw <= staged ? buf + ew : reinterpret_cast<const uint64_t*>(P.arena + abs_off);
            sz = cs_odd_record<EMIT>(s, O, r, abs_off, room, cu, ev0, op0);
          }
          s.seg_k = sz;
        }
        __syncthreads();
        const uint32_t sz = s.seg_k;  // (written again only behind the next run's two barriers)
        if (sz == 0) {
          bad = WHY_CS_RECORD;
          break;
        }
        ew += sz / 8;
        if (ew >= wt) break;
      }
      next_rel = tile_rel + 8 * ew;
    }
    __syncthreads();  // buffer b is free for the tile after next; pred / counts are published
    if (bad) break;
    if (next_rel > limit_rel) {
      bad = WHY_CS_PAST_END;
      break;
    }
  }
  if (n_waited < n_issued) mbar_wait(&s.bar[(n_waited + boff) & 1], (phase_bits >> ((n_waited + boff) & 1)) & 1u);
  const uint64_t next_abs = any_work ? chunk_abs + next_rel : entry_abs;

  // ---- publish ----
  __syncthreads();
  if (threadIdx.x == 0) {
    if (!EMIT) {
      io.out_exit = next_abs;
      io.out_last = s.pred_off;
      io.out_gap = 0;
      io.n_rows = s.n_ev;
      io.n_records = s.n_ops;
    }
    if (bad) {
      if (SPEC) raise_exact(P.flags, bad, c);
      else if (bad != WHY_CS_RECORD) atomicExch(&P.flags[FLAG_INTERNAL], bad);  // the exact walk stops before a bad record
    }
  }
  if (!EMIT) {
    uint64_t* ctr = P.counters + static_cast<uint64_t>(cd.stream) * C_SLOTS;
    if (threadIdx.x < 16 && threadIdx.x != CSC_LOST && s.cnt[threadIdx.x]) red_add(&ctr[threadIdx.x], s.cnt[threadIdx.x]);
    if (threadIdx.x == 0) {
      if (s.lost_sum) red_add(&ctr[CSC_LOST], s.lost_sum);
      if (s.n_ev) red_add(&ctr[CSC_EVENTS], s.n_ev);
      if (s.n_ops) red_add(&ctr[CSC_OPS], s.n_ops);
    }
  }
}

template <bool SPEC, bool EMIT>
__global__ void __launch_bounds__(NT, 5) k_cs(const Params P, const CsOut O, const uint32_t gated) {
  extern __shared__ __align__(128) unsigned char cs_smem_raw[];
  CsSmem& s = *reinterpret_cast<CsSmem*>(cs_smem_raw);
  // the exact passes run only when the speculation was disproved
  if (!SPEC && (!EMIT || gated) && P.flags[FLAG_NEED_EXACT] == 0) return;
  const uint32_t c = blockIdx.x;
  cs_chunk<SPEC, EMIT>(P, O, s, c, EMIT ? P.io[c].row_base : 0, EMIT ? P.io[c].rec_base : 0);
}

// ------------------------------------------------------------------------------------
// k_cs_resolve (only if NEED_EXACT): exact sequential walk, one warp per stream; runs of
// 32-byte SWITCH records are skipped 32 at a time (lane l tests the header at pos + 32*l).
// Windows of 8 KiB arrive in shared memory by double-buffered TMA copies, so the dependent
// loads of the walk hit on-chip memory while the next window is in flight.
// Hands every chunk its exact entry and predecessor, finds the first fence violation.
constexpr int CS_RES_WIN_BYTES = 8192;
constexpr int CS_RES_OV_BYTES = 16;  // only headers are read: one word past the window at most
__global__ void __launch_bounds__(32) k_cs_resolve(const Params P) {
  if (P.flags[FLAG_NEED_EXACT] == 0) return;
  __shared__ __align__(128) uint64_t win[2][(CS_RES_WIN_BYTES + CS_RES_OV_BYTES) / 8];
  __shared__ __align__(8) uint64_t bar[2];
  const uint32_t sidx = blockIdx.x, lane = threadIdx.x;
  if (sidx >= P.n_streams) return;
  const bool writer = lane == 0;
  if (writer) {
    mbar_init(&bar[0], 1);
    mbar_init(&bar[1], 1);
    fence_barrier_init();
  }
  __syncwarp();
  uint32_t phase_bits = 0;
  const StreamDesc sd = P.streams[sidx];
  uint64_t* ctr = P.counters + static_cast<uint64_t>(sidx) * C_SLOTS;
  uint64_t last_off = PRED_CARRY_IN;
  int64_t err = 0;
  uint64_t err_off = 0;
  bool stopped = false;
  for (uint32_t ii = 0; ii < sd.n_images; ++ii) {
    const ImageDesc im = P.images[sd.first_image + ii];
    uint32_t ch = im.first_chunk;
    const uint32_t ch_end = im.first_chunk + im.n_chunks;
    uint64_t next_begin = ch < ch_end ? P.chunks[ch].begin : ~0ull;
    uint64_t pos = 0;
    const uint8_t* const img = P.arena + im.base;
    auto issue = [&](uint64_t wb, uint32_t b) {
      const uint64_t avail = (im.len - wb + 15ull) & ~15ull;
      const uint32_t bytes = static_cast<uint32_t>(umin64(avail, CS_RES_WIN_BYTES + CS_RES_OV_BYTES));
      mbar_expect_tx(&bar[b], bytes);
      tma_load_1d(win[b], img + wb, bytes, &bar[b]);
    };
    uint32_t issued = 0, waited = 0;
    if (!stopped && im.len > 0) {
      if (writer) issue(0, 0);
      issued = 1;
    }
    for (uint64_t wb = 0; !stopped && wb < im.len; wb += CS_RES_WIN_BYTES) {
      const uint32_t b = static_cast<uint32_t>(wb / CS_RES_WIN_BYTES) & 1;
      __syncwarp();  // every lane is done with buffer b^1
      if (wb + CS_RES_WIN_BYTES < im.len) {
        if (writer) issue(wb + CS_RES_WIN_BYTES, b ^ 1);
        ++issued;
      }
      mbar_wait(&bar[b], (phase_bits >> b) & 1u);
      phase_bits ^= 1u << b;
      ++waited;
      const uint64_t* wd = win[b];
      const uint64_t wend = umin64(wb + CS_RES_WIN_BYTES, im.len);
      while (pos < wend) {
        const uint64_t q = pos + 32ull * lane;
        const bool ok = q < wend && q + 32 <= im.len && cs_is_switch32(wd[(q - wb) / 8]);
        const uint32_t okm = __ballot_sync(0xffffffffu, ok);
        const uint32_t run = okm == 0xFFFFFFFFu ? 32u : static_cast<uint32_t>(__ffs(~okm) - 1);
        if (run > 0) {
          const uint64_t last_rec = pos + 32ull * (run - 1);
          while (next_begin <= last_rec) {
            const uint64_t j = next_begin <= pos ? 0 : (next_begin - pos + 31) / 32;
            if (writer) {
              ChunkIO& io = P.io[ch];
              io.in_entry = im.base + pos + 32 * j;
              io.in_pred = j ? im.base + pos + 32 * (j - 1) : last_off;
              io.in_gap = 0;
              io.limit = ~0ull;
            }
            ++ch;
            next_begin = ch < ch_end ? P.chunks[ch].begin : ~0ull;
          }
          last_off = im.base + last_rec;
          pos = last_rec + 32;
          continue;
        }
        while (pos >= next_begin) {
          if (writer) {
            ChunkIO& io = P.io[ch];
            io.in_entry = im.base + pos;
            io.in_pred = last_off;
            io.in_gap = 0;
            io.limit = ~0ull;
          }
          ++ch;
          next_begin = ch < ch_end ? P.chunks[ch].begin : ~0ull;
        }
        uint32_t type = 0, size = 0;
        if (pos + 8 <= im.len) {
          const uint64_t h = wd[(pos - wb) / 8];
          type = static_cast<uint32_t>(h);
          size = static_cast<uint32_t>(h >> 48);
        }
        if (size < 8 || (size & 7) || pos + size > im.len || size < cs_min_size(type)) {
          err = -1;  // -EPERM
          err_off = im.stream_pos + pos;
          stopped = true;
          break;
        }
        if (cs_handled(type)) last_off = im.base + pos;
        pos += size;
      }
    }
    if (waited < issued) {  // never leave a bulk copy in flight
      const uint32_t b = waited & 1;
      mbar_wait(&bar[b], (phase_bits >> b) & 1u);
      phase_bits ^= 1u << b;
    }
    __syncwarp();
    const uint64_t stop_abs = im.base + pos;
    for (uint32_t k = im.first_chunk + lane; k < ch_end; k += 32) {
      ChunkIO& io = P.io[k];
      if (k >= ch) {
        io.in_entry = stopped ? ~0ull : stop_abs;
        io.in_pred = last_off;
        io.in_gap = 0;
      }
      io.limit = stopped ? stop_abs : ~0ull;
    }
  }
  if (writer) {
    ctr[CSC_ERR] = static_cast<uint64_t>(err);
    ctr[CSC_ERR_OFF] = err_off;
  }
}

// zero the counters when the exact path recomputes them
__global__ void k_cs_clear(const Params P) {
  if (P.flags[FLAG_NEED_EXACT] == 0) return;
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < P.n_streams * C_SLOTS && (i % C_SLOTS) < CSC_ERR) P.counters[i] = 0;
}

// Per stream: exclusive scan of the chunks' event / op counts (a warp per stream, 32 chunks per
// step), the generator state after the stream's last handled record, and the stream totals.
__global__ void __launch_bounds__(32) k_cs_scan(const Params P, const CsOut O, uint64_t* totals) {
  const uint32_t sidx = blockIdx.x, lane = threadIdx.x;
  const StreamDesc sd = P.streams[sidx];
  uint64_t ev = 0, op = 0, last = PRED_CARRY_IN;
  for (uint32_t ii = 0; ii < sd.n_images; ++ii) {
    const ImageDesc im = P.images[sd.first_image + ii];
    for (uint32_t k0 = 0; k0 < im.n_chunks; k0 += 32) {
      const uint32_t k = im.first_chunk + k0 + lane;
      const bool in = k0 + lane < im.n_chunks;
      uint32_t a = in ? P.io[k].n_rows : 0, b = in ? P.io[k].n_records : 0;
      uint32_t ia = a, ib = b;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const uint32_t ya = __shfl_up_sync(0xffffffffu, ia, d), yb = __shfl_up_sync(0xffffffffu, ib, d);
        if (lane >= d) {
          ia += ya;
          ib += yb;
        }
      }
      if (in) {
        P.io[k].row_base = ev + ia - a;  // stream-relative; k_cs_bases adds the stream's base
        P.io[k].rec_base = op + ib - b;
      }
      ev += __shfl_sync(0xffffffffu, ia, 31);
      op += __shfl_sync(0xffffffffu, ib, 31);
    }
    if (im.n_chunks) last = P.io[im.first_chunk + im.n_chunks - 1].out_last;
  }
  if (lane == 0) {
    totals[2 * sidx] = ev;
    totals[2 * sidx + 1] = op;
    uint64_t ts = 0;
    uint32_t lost = 0;
    cs_pred_state(P, O, sidx, last, ts, lost);
    O.state_out[sidx].last_tstamp = ts;
    O.state_out[sidx].in_write_errors = lost;
    O.state_out[sidx].pad = 0;
    P.counters[static_cast<uint64_t>(sidx) * C_SLOTS + CSC_LAST_TS] = ts;
  }
}

// exclusive scan of the stream totals (one warp, 32 streams per step) and rebase of every chunk
__global__ void __launch_bounds__(32) k_cs_bases(const Params P, const CsOut O, const uint64_t* totals) {
  const uint32_t lane = threadIdx.x;
  uint64_t ev = 0, op = 0;
  for (uint32_t s0 = 0; s0 < P.n_streams; s0 += 32) {
    const uint32_t s = s0 + lane;
    const uint64_t a = s < P.n_streams ? totals[2 * s] : 0, b = s < P.n_streams ? totals[2 * s + 1] : 0;
    uint64_t ia = a, ib = b;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint64_t ya = __shfl_up_sync(0xffffffffu, ia, d), yb = __shfl_up_sync(0xffffffffu, ib, d);
      if (lane >= d) {
        ia += ya;
        ib += yb;
      }
    }
    if (s < P.n_streams) {
      O.ev_base[s] = ev + ia - a;
      O.op_base[s] = op + ib - b;
    }
    ev += __shfl_sync(0xffffffffu, ia, 31);
    op += __shfl_sync(0xffffffffu, ib, 31);
  }
  if (lane == 0) {
    O.ev_base[P.n_streams] = ev;
    O.op_base[P.n_streams] = op;
  }
}
__global__ void k_cs_rebase(const Params P, const CsOut O) {
  const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= P.n_chunks) return;
  const uint32_t s = P.chunks[c].stream;
  P.io[c].row_base += O.ev_base[s];
  P.io[c].rec_base += O.op_base[s];
}

}  // namespace hbtgpu
