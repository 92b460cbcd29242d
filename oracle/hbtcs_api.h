/*
 * oracle/hbtcs_api.h -- plain-C interface of the two CPU checkers for the ContextSwitch-mode
 * decode.  TEST INFRASTRUCTURE ONLY, like hbtref_api.h.
 *
 *   hbtref_cs_run     oracle/_ref/libhbtref.so: the REAL hbt code
 *                     (PerfEventsGroup<CpuThreadSwitchGenerator, mode::ContextSwitch>::consume,
 *                     hbt/src/perf_event/PerfEventsGroup.h:1394-1589, and the handlers of
 *                     hbt/src/perf_event/PerCpuThreadSwitchGenerator.h:168-446 with their shared
 *                     PerThreadStats :52-135), driven by oracle/ref_harness_cs.cpp
 *   hbtoracle_cs_run  oracle/libhbtoracle.so: plain-C restatement (oracle/hbt_cs_oracle.c)
 *
 * Record layouts (mode::ContextSwitch, PerfEventsGroup.h:105-191; sample_type TIME|TID with
 * sample_id_all, :963-970): every record ends in sample_id {u32 pid, tid; u64 time}.
 *   LOST 40 B {hdr, id, num_lost, sample_id}; EXIT/FORK 48 B {hdr, pid, ppid, tid, ptid, time,
 *   sample_id}; SWITCH_CPU_WIDE 32 B {hdr, next_prev_pid, next_prev_tid, sample_id};
 *   COMM {hdr, pid, tid, comm[size-32], sample_id}.
 */
#ifndef HBTCS_API_H
#define HBTCS_API_H

#include <stdint.h>

#include "hbtref_api.h"

#ifdef __cplusplus
extern "C" {
#endif

/* tagstack::Event, hbt/src/tagstack/Event.h:28-56 (16 bytes, same field order) */
typedef struct {
  uint64_t tstamp;
  uint8_t type;          /* Event::Type: 2 ThreadCreation, 3 ThreadDestruction, 4 SwitchIn,
                            5 SwitchOutPreempt, 6 SwitchOutYield, 7 WriteErrorsStart, 8 WriteErrorsEnd */
  uint8_t level;         /* tid_level, or kInvalidLevel (255) for the WriteErrors events */
  uint16_t comp_unit_id; /* the CPU */
  uint32_t tag;          /* virtual thread id, or kNA (0xFFFFFFFF) */
} hbtcs_event;

/* ThreadInfo, PerCpuThreadSwitchGenerator.h:36-47 (cmdline omitted: /proc of the test host) */
typedef struct {
  uint32_t vid;
  int32_t ppid, ptid, pid, tid;
  uint32_t n_names;
  uint64_t fork_ts, end_ts; /* kInvalidTimeStamp = ~0 when unset */
} hbtcs_thread;

/* one point of ThreadInfo::names that came from a COMM record (:244-279) */
typedef struct {
  uint32_t vid;
  uint8_t is_exec;
  uint8_t len;      /* bytes of the record's comm field (size - 32), clipped to 16 */
  uint8_t pad[2];
  uint64_t tstamp;
  char name[16];
} hbtcs_name;

/* CpuThreadSwitchGenerator::last_tstamp_ / in_write_errors_ (:418-421) */
typedef struct {
  uint64_t last_tstamp;
  uint32_t in_write_errors;
  uint32_t pad;
} hbtcs_state;

#define HBTCS_DECL(prefix)                                                                   \
  /* Decode n_streams per-CPU streams (comp_unit_id = stream index): CPUs in index order,   */ \
  /* each drained with consume(batch) until -ENODATA before the next one starts, all        */ \
  /* sharing one PerThreadStats.  Events of stream s go to events + ev_off[s] (at most      */ \
  /* ev_cap[s]); threads by vid, names grouped by thread in series order.  in may be NULL   */ \
  /* (state after construction: last_tstamp 0, not in write errors).                        */ \
  int prefix##_cs_run(uint32_t n_streams, const uint8_t* const* bytes, const uint64_t* lens, \
      uint32_t tid_level, uint32_t batch, const hbtcs_state* in, hbtcs_state* out,           \
      hbtcs_event* events, const uint64_t* ev_off, const uint64_t* ev_cap,                   \
      uint64_t* n_events, hbtref_counters* ctrs, hbtcs_thread* threads,                      \
      uint64_t thread_cap, uint64_t* n_threads, hbtcs_name* names, uint64_t name_cap,        \
      uint64_t* n_names);

HBTCS_DECL(hbtref)
HBTCS_DECL(hbtoracle)

#ifdef __cplusplus
}
#endif
#endif
