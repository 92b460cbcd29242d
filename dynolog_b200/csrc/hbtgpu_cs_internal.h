// hbtgpu_cs_internal.h -- what hbtgpu_cs_vids.cu needs to see of a ContextSwitch context
// (library-internal, not part of the C ABI).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "hbtgpu.h"

struct hbtgpu_cs_view {
  int device;
  cudaStream_t stream;
  uint32_t ns;
  hbtgpu_cs_event* events;  // all cpus, cpu c at [ev_base[c], ev_base[c + 1])
  uint32_t* pids;
  hbtgpu_cs_op* ops;
  const uint64_t* d_ev_base;
  const uint64_t* d_op_base;
  uint64_t n_events, n_ops;
  bool* vids_done;          // the window's tags have been rewritten to virtual ids
};

// -EINVAL if the context has not run
__attribute__((visibility("hidden"))) int hbtgpu_cs_internal_view(hbtgpu_cs_ctx* ctx, hbtgpu_cs_view* out);
