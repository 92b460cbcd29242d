// hbtgpu_cs_vids.cu -- the thread dictionary of the ContextSwitch-mode decode on the GPU.
//
// Reference: PerThreadStats (hbt/src/perf_event/PerCpuThreadSwitchGenerator.h:52-135) as the
// handlers drive it (:255-407).  It is a serial state machine shared by all CPUs:
//   touch(tid)  [COMM, EXIT, FORK, switch-in; tid != 0]: vid = active[tid] if present, else
//               vid = next_vid++ and active[tid] = vid            (selectNextVid/createThreadInfo)
//   EXIT        additionally erases active[tid] afterwards                      (registerExit)
//   LOST        clears active                                                       (:243-247)
// in the order the consumer loop visits the records: cpu by cpu, each in ring order.
//
// Parallel form.  Number the LOSTs: epoch(x) = how many LOSTs precede x.  Within one
// (epoch, tid) group in visiting order, a touch creates a vid iff it is the group's first or
// follows an EXIT of the group; every other touch takes the vid of the latest creating touch
// before it.  vids are numbered in visiting order of the creating touches.  So:
//   1. list the touches in visiting order (switch-in / thread events of all cpus, the COMM
//      records merged in at their place), key = (epoch, tid);
//   2. stable radix sort by key (cub::DeviceRadixSort: library code, the one non-hand-written
//      step) -> groups, still in visiting order inside;
//   3. flag the creating touches, scatter the flags back to visiting order, exclusive sum ->
//      their vid numbers;
//   4. in group order an inclusive "last non-zero" scan hands every touch its vid;
//   5. write the vid into the event's tag, fill the thread table, collect the tids still
//      active at the end of the window (the seeds of the next window).
// A dictionary carried over from earlier windows enters as seed touches in front of epoch 0.
#include <cuda_runtime.h>
#include <errno.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <string>
#include <vector>

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <cub/device/device_select.cuh>
#include <cub/iterator/counting_input_iterator.cuh>

#include "hbtgpu.h"
#include "hbtgpu_cs_internal.h"

namespace {

constexpr uint32_t KIND_EVENT = 0, KIND_COMM = 1, KIND_SEED = 2;
constexpr uint32_t EV_CREATE = 2, EV_DESTROY = 3, EV_IN = 4, EV_WE_START = 7;
constexpr uint32_t REC_COMM = 3, REC_EXIT = 4, REC_FORK = 7;
constexpr uint64_t INVALID_TS = ~0ull;

enum { VF_TID0 = 1, VF_DOUBLE_FORK = 2, VF_INTERNAL = 4, VF_WIDE_TID = 8 };

struct IsLost {
  const hbtgpu_cs_event* ev;
  __device__ bool operator()(uint32_t i) const { return ev[i].type == EV_WE_START; }
};
struct IsTouch {
  const hbtgpu_cs_event* ev;
  __device__ bool operator()(uint32_t i) const {
    const uint32_t t = ev[i].type;
    return t == EV_CREATE || t == EV_DESTROY || (t == EV_IN && ev[i].tag != 0);
  }
};
struct LastNonZero {
  __device__ uint32_t operator()(uint32_t a, uint32_t b) const { return b ? b : a; }
};

__device__ __forceinline__ uint32_t lower_bound_u32(const uint32_t* a, uint32_t n, uint32_t x) {
  uint32_t lo = 0, hi = n;
  while (lo < hi) {
    const uint32_t mid = (lo + hi) >> 1;
    if (a[mid] < x) lo = mid + 1;
    else hi = mid;
  }
  return lo;  // number of entries < x
}

// global event index of every op (ops are in visiting order: cpu-major, ring order inside)
__global__ void k_vid_ops(const hbtgpu_cs_op* ops, uint32_t n_ops, const uint64_t* ev_base, const uint64_t* op_base,
                          uint32_t ns, uint32_t* op_gidx, uint32_t* comm_flag) {
  const uint32_t o = blockIdx.x * blockDim.x + threadIdx.x;
  if (o >= n_ops) return;
  uint32_t lo = 0, hi = ns;  // cpu with op_base[cpu] <= o < op_base[cpu + 1]
  while (lo + 1 < hi) {
    const uint32_t mid = (lo + hi) >> 1;
    if (op_base[mid] <= o) lo = mid;
    else hi = mid;
  }
  op_gidx[o] = static_cast<uint32_t>(ev_base[lo] + ops[o].event_index);
  comm_flag[o] = ops[o].type == REC_COMM;
}

struct VidArgs {
  const hbtgpu_cs_event* ev;
  hbtgpu_cs_event* ev_out;
  const uint32_t* pids;
  const hbtgpu_cs_op* ops;
  const uint32_t* op_gidx;   // [n_ops]
  const uint32_t* comm_idx;  // [n_comm] op indices of the COMM records, ascending
  const uint32_t* touch_idx; // [n_touch] global event indices of the touch events, ascending
  const uint32_t* lost_idx;  // [n_lost] global event indices of the WriteErrorsStart events
  const uint32_t* seed_tid;  // [n_seed]
  const uint32_t* seed_vid;
  uint32_t n_ops, n_comm, n_touch, n_lost, n_seed, n_items;
  uint32_t next_vid0;
  // items, visiting order
  uint64_t* key;      // epoch << tid_bits | tid: only the bits in use are sorted
  uint32_t tid_bits;  // 32, or fewer when every thread id fits (VF_WIDE_TID is raised when one does not)
  uint32_t* src;      // kind << 30 | ET_* << 28 | index
  uint32_t* pos_iota;
  uint32_t* touch_pos;  // [n_touch] visiting position of every touch event
  uint32_t* vid_pos;    // by visiting position: the vid the item got (k_vid_apply -> k_vid_tag)
  // sorted
  const uint64_t* key_s;
  const uint32_t* pos_s;
  uint32_t* newflag;  // by visiting position
  const uint32_t* rank;
  uint32_t* def;      // sorted order
  const uint32_t* vid_s;
  // thread table (by vid - 1)
  hbtgpu_cs_thread* info;
  hbtgpu_cs_name* names;
  unsigned int* n_names;
  uint32_t* out_tid;  // active at the end of the window
  uint32_t* out_vid;
  unsigned int* n_out;
  unsigned int* flags;
};

// src of an item: kind (2 bits) | what the touched event is (2 bits: ET_*) | index (28 bits)
enum : uint32_t { ET_OTHER = 0, ET_CREATE = 1, ET_DESTROY = 2 };
constexpr uint32_t SRC_IDX_MASK = 0x0FFFFFFFu;
__device__ __forceinline__ uint32_t src_kind(uint32_t src) { return src >> 30; }
__device__ __forceinline__ uint32_t src_et(uint32_t src) { return (src >> 28) & 3u; }
__device__ __forceinline__ uint32_t src_idx(uint32_t src) { return src & SRC_IDX_MASK; }

__device__ __forceinline__ uint32_t key_tid(const VidArgs& A, uint64_t key) {
  return static_cast<uint32_t>(A.tid_bits < 32 ? key & ((1ull << A.tid_bits) - 1) : key & 0xFFFFFFFFull);
}
__device__ __forceinline__ uint32_t comm_gidx(const VidArgs& A, uint32_t c) { return A.op_gidx[A.comm_idx[c]]; }

// items in visiting order: seeds, then touch events and COMM records merged by position
__global__ void k_vid_items(const VidArgs A) {
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < A.n_seed) {
    A.key[t] = A.seed_tid[t];  // epoch 0
    if (A.tid_bits < 32 && (A.seed_tid[t] >> A.tid_bits)) atomicOr(A.flags, VF_WIDE_TID);
    A.src[t] = (KIND_SEED << 30) | t;
    A.pos_iota[t] = t;
    return;
  }
  const uint32_t u = t - A.n_seed;
  if (u < A.n_touch) {
    const uint32_t gi = A.touch_idx[u];
    // COMM records placed before this event: their event_index is <= gi
    uint32_t lo = 0, hi = A.n_comm;
    while (lo < hi) {
      const uint32_t mid = (lo + hi) >> 1;
      if (comm_gidx(A, mid) <= gi) lo = mid + 1;
      else hi = mid;
    }
    const uint32_t p = A.n_seed + u + lo;
    const uint64_t epoch = lower_bound_u32(A.lost_idx, A.n_lost, gi);
    const hbtgpu_cs_event e = A.ev[gi];
    const uint32_t tid = e.tag;
    if (A.tid_bits < 32 && (tid >> A.tid_bits)) atomicOr(A.flags, VF_WIDE_TID);
    A.key[p] = (epoch << A.tid_bits) | tid;
    const uint32_t et = e.type == EV_CREATE ? ET_CREATE : e.type == EV_DESTROY ? ET_DESTROY : ET_OTHER;
    A.src[p] = (KIND_EVENT << 30) | (et << 28) | u;
    A.pos_iota[p] = p;
    A.touch_pos[u] = p;
  } else if (u < A.n_touch + A.n_comm) {
    const uint32_t c = u - A.n_touch;
    const uint32_t cg = comm_gidx(A, c);
    const uint32_t before = lower_bound_u32(A.touch_idx, A.n_touch, cg);  // touch events with index < cg
    const uint32_t p = A.n_seed + c + before;
    const uint64_t epoch = lower_bound_u32(A.lost_idx, A.n_lost, cg);
    const uint32_t tid = static_cast<uint32_t>(A.ops[A.comm_idx[c]].tid);
    if (tid == 0) atomicOr(A.flags, VF_TID0);
    if (A.tid_bits < 32 && (tid >> A.tid_bits)) atomicOr(A.flags, VF_WIDE_TID);
    A.key[p] = (epoch << A.tid_bits) | tid;
    A.src[p] = (KIND_COMM << 30) | c;
    A.pos_iota[p] = p;
  }
}

__device__ __forceinline__ bool item_is_exit(const VidArgs& A, uint32_t src) {
  (void)A;
  return src_kind(src) == KIND_EVENT && src_et(src) == ET_DESTROY;
}

// sorted order: which touches create a vid
__global__ void k_vid_new(const VidArgs A) {
  const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= A.n_items) return;
  const uint32_t p = A.pos_s[k];
  const uint32_t src = A.src[p];
  bool nw = false;
  if ((src >> 30) != KIND_SEED) {
    nw = k == 0 || A.key_s[k] != A.key_s[k - 1] || item_is_exit(A, A.src[A.pos_s[k - 1]]);
    if (key_tid(A, A.key_s[k]) == 0) atomicOr(A.flags, VF_TID0);  // COMM/EXIT/FORK of tid 0
  }
  A.newflag[p] = nw;
}

__global__ void k_vid_def(const VidArgs A) {
  const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= A.n_items) return;
  const uint32_t p = A.pos_s[k];
  const uint32_t src = A.src[p];
  uint32_t d = 0;
  if (src_kind(src) == KIND_SEED) d = A.seed_vid[src_idx(src)];
  else if (A.newflag[p]) d = A.next_vid0 + A.rank[p];
  A.def[k] = d;
}

// the FORK / EXIT record of an event (rare): ops are sorted by global event index
__device__ __forceinline__ const hbtgpu_cs_op* op_of_event(const VidArgs& A, uint32_t gi, uint32_t type) {
  uint32_t o = lower_bound_u32(A.op_gidx, A.n_ops, gi);
  for (; o < A.n_ops && A.op_gidx[o] == gi; ++o)
    if (A.ops[o].type == type) return &A.ops[o];
  return nullptr;
}

// creating touches write the whole ThreadInfo (createThreadInfo :60-91) ...
__global__ void k_vid_create(const VidArgs A) {
  const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= A.n_items) return;
  const uint32_t p = A.pos_s[k];
  const uint32_t src = A.src[p], kind = src_kind(src), idx = src_idx(src);
  if (kind == KIND_SEED || !A.newflag[p]) return;
  const uint32_t vid = A.vid_s[k];
  if (vid == 0) {
    atomicOr(A.flags, VF_INTERNAL);
    return;
  }
  const int32_t tid = static_cast<int32_t>(key_tid(A, A.key_s[k]));
  hbtgpu_cs_thread ti{vid, -1, -1, -1, tid, 0, INVALID_TS, INVALID_TS};
  if (kind == KIND_COMM) {
    ti.pid = A.ops[A.comm_idx[idx]].pid;
  } else {
    const uint32_t gi = A.touch_idx[idx];
    const uint32_t type = A.ev[gi].type;
    if (type == EV_IN) {
      ti.pid = static_cast<int32_t>(A.pids[gi]);
    } else {
      const hbtgpu_cs_op* op = op_of_event(A, gi, type == EV_CREATE ? REC_FORK : REC_EXIT);
      if (!op) {
        atomicOr(A.flags, VF_INTERNAL);
        return;
      }
      ti.ppid = op->ppid;
      ti.ptid = op->ptid;
      ti.pid = op->pid;
    }
  }
  A.info[vid - 1] = ti;
}

// ... then every touch applies its own effect: the event's tag, fork / end time, names, and
// whether its tid is still active after the window
__global__ void k_vid_apply(const VidArgs A) {
  const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= A.n_items) return;
  const uint32_t p = A.pos_s[k];
  const uint32_t src = A.src[p], kind = src_kind(src), idx = src_idx(src);
  const uint32_t vid = A.vid_s[k];
  const uint64_t key = A.key_s[k];
  const uint32_t tid = key_tid(A, key);
  if (vid == 0) {
    atomicOr(A.flags, VF_INTERNAL);
    return;
  }
  bool is_exit = false;
  if (kind == KIND_EVENT) {
    // the event's tag is written by k_vid_tag, in event order: from here it would be one access
    // to a random event per item, and that -- not the sort -- was 2/3 of the dictionary's time
    A.vid_pos[p] = vid;
    const uint32_t et = src_et(src);
    if (et == ET_CREATE) {
      // a FORK of a thread that already has a fork time throws in the reference (:322-325):
      // look back over this vid's run for another FORK or for the seed of a forked thread
      for (uint32_t j = k; j > 0;) {
        --j;
        if (A.key_s[j] != key || A.vid_s[j] != vid) break;
        const uint32_t sj = A.src[A.pos_s[j]];
        if (src_kind(sj) == KIND_EVENT && src_et(sj) == ET_CREATE) atomicOr(A.flags, VF_DOUBLE_FORK);
        if (src_kind(sj) == KIND_SEED && A.info[vid - 1].fork_ts != INVALID_TS) atomicOr(A.flags, VF_DOUBLE_FORK);
      }
      A.info[vid - 1].fork_ts = A.ev[A.touch_idx[idx]].tstamp;
    } else if (et == ET_DESTROY) {
      A.info[vid - 1].end_ts = A.ev[A.touch_idx[idx]].tstamp;
      is_exit = true;
    }
  } else if (kind == KIND_COMM) {
    const hbtgpu_cs_op& op = A.ops[A.comm_idx[idx]];
    const unsigned int s = atomicAdd(A.n_names, 1u);
    hbtgpu_cs_name nm{};
    nm.vid = vid;
    nm.is_exec = 1;
    nm.len = static_cast<uint8_t>(op.name_len > 16 ? 16 : op.name_len);
    nm.tstamp = op.tstamp;
    for (int i = 0; i < 16; ++i) nm.name[i] = i < nm.len ? op.name[i] : 0;
    A.names[s] = nm;
  }
  // still active after the window: last touch of its group, in the last epoch, not an EXIT
  const bool last_of_group = k + 1 == A.n_items || A.key_s[k + 1] != key;
  if (last_of_group && (key >> A.tid_bits) == A.n_lost && !is_exit) {
    const unsigned int s = atomicAdd(A.n_out, 1u);
    A.out_tid[s] = tid;
    A.out_vid[s] = vid;
  }
}

// ... and the touch events take their virtual ids, in event order
__global__ void k_vid_tag(const VidArgs A) {
  const uint32_t u = blockIdx.x * blockDim.x + threadIdx.x;
  if (u >= A.n_touch) return;
  const uint32_t vid = A.vid_pos[A.touch_pos[u]];
  if (vid) A.ev_out[A.touch_idx[u]].tag = vid;  // (0: an item without a vid, reported by k_vid_apply)
}

// switch-ins of tid 0 keep tag 0 (selectNextVid :97-101): nothing to do.  Everything else that
// carries a tid has been rewritten by k_vid_apply.

}  // namespace

struct hbtgpu_cs_vids {
  int device = 0;
  std::string err;
  uint32_t next_vid = 1;
  // dictionary carried between windows
  uint32_t *seed_tid = nullptr, *seed_vid = nullptr;
  size_t cap_seed_tid = 0, cap_seed_vid = 0;
  uint32_t n_seed = 0;
  // thread table on the device
  hbtgpu_cs_thread* info = nullptr;
  size_t cap_info = 0;
  hbtgpu_cs_name* names = nullptr;
  size_t cap_names = 0;
  uint32_t n_names = 0;
  // scratch
  void* scratch = nullptr;
  size_t cap_scratch = 0;
  unsigned int* d_small = nullptr;  // [0] n_names, [1] n_out, [2] flags, [3..5] select counts
  float last_ms = 0.f;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
};

namespace {

int vfail(hbtgpu_cs_vids* v, int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  if (v) v->err = buf;
  return code;
}

#define VK(call)                                                                                   \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess)                                                                         \
      return vfail(v, -EIO, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
  } while (0)

template <class T>
int grow(hbtgpu_cs_vids* v, T** p, size_t* cap, size_t need, size_t keep, cudaStream_t st) {
  if (need <= *cap) return 0;
  const size_t ncap = std::max(need, *cap * 2);
  T* q = nullptr;
  VK(cudaMalloc(reinterpret_cast<void**>(&q), ncap * sizeof(T)));
  if (*p && keep) VK(cudaMemcpyAsync(q, *p, keep * sizeof(T), cudaMemcpyDeviceToDevice, st));
  if (*p) {
    VK(cudaStreamSynchronize(st));
    VK(cudaFree(*p));
  }
  *p = q;
  *cap = ncap;
  return 0;
}

struct Bump {  // carve aligned arrays out of one scratch allocation
  uint8_t* base;
  size_t off = 0;
  template <class T>
  T* take(size_t n) {
    off = (off + 255) & ~static_cast<size_t>(255);
    T* p = base ? reinterpret_cast<T*>(base + off) : nullptr;
    off += n * sizeof(T);
    return p;
  }
};

}  // namespace

extern "C" {

hbtgpu_cs_vids* hbtgpu_cs_vids_create(int device) {
  if (cudaSetDevice(device) != cudaSuccess) return nullptr;
  hbtgpu_cs_vids* v = new hbtgpu_cs_vids();
  v->device = device;
  if (cudaMalloc(reinterpret_cast<void**>(&v->d_small), 8 * sizeof(unsigned int)) != cudaSuccess ||
      cudaEventCreate(&v->ev0) != cudaSuccess || cudaEventCreate(&v->ev1) != cudaSuccess) {
    delete v;
    return nullptr;
  }
  return v;
}

void hbtgpu_cs_vids_destroy(hbtgpu_cs_vids* v) {
  if (!v) return;
  cudaSetDevice(v->device);
  for (void* p : {(void*)v->seed_tid, (void*)v->seed_vid, (void*)v->info, (void*)v->names, v->scratch,
                  (void*)v->d_small})
    if (p) cudaFree(p);
  if (v->ev0) cudaEventDestroy(v->ev0);
  if (v->ev1) cudaEventDestroy(v->ev1);
  delete v;
}

const char* hbtgpu_cs_vids_last_error(hbtgpu_cs_vids* v) { return v ? v->err.c_str() : "null dictionary"; }

int hbtgpu_cs_assign_vids_device(hbtgpu_cs_ctx* ctx, hbtgpu_cs_vids* v) {
  if (!ctx || !v) return -EINVAL;
  hbtgpu_cs_view W{};
  if (int r = hbtgpu_cs_internal_view(ctx, &W)) return vfail(v, r, "hbtgpu_cs_run has not been called");
  if (W.device != v->device) return vfail(v, -EINVAL, "context and dictionary live on different devices");
  if (*W.vids_done) return vfail(v, -EINVAL, "this window's events already carry virtual ids");
  if (W.n_events >= (1ull << 30) || W.n_ops >= (1ull << 30))
    return vfail(v, -E2BIG, "window too large for 30-bit item indices (%llu events)", (unsigned long long)W.n_events);
  VK(cudaSetDevice(v->device));
  cudaStream_t st = W.stream;
  const uint32_t E = static_cast<uint32_t>(W.n_events), O = static_cast<uint32_t>(W.n_ops);
  VK(cudaEventRecord(v->ev0, st));
  VK(cudaMemsetAsync(v->d_small, 0, 8 * sizeof(unsigned int), st));

  // ---- pass A: sizes of the selections (cub two-phase calls need them) ----
  cub::CountingInputIterator<uint32_t> iota(0);
  IsLost is_lost{W.events};
  IsTouch is_touch{W.events};
  size_t tmp_sel = 0, tmp_scan = 0, tmp_sort = 0, t2 = 0;
  cub::DeviceSelect::If(nullptr, tmp_sel, iota, static_cast<uint32_t*>(nullptr), static_cast<unsigned int*>(nullptr),
                        static_cast<int>(std::max(E, O)), is_touch, st);
  cub::DeviceSelect::Flagged(nullptr, t2, iota, static_cast<uint32_t*>(nullptr), static_cast<uint32_t*>(nullptr),
                             static_cast<unsigned int*>(nullptr), static_cast<int>(std::max(E, O)), st);
  tmp_sel = std::max(tmp_sel, t2);

  // first scratch layout: selections
  size_t need0;
  {
    Bump b{nullptr};
    b.take<uint32_t>(O);      // op_gidx
    b.take<uint32_t>(O);      // comm_flag
    b.take<uint32_t>(O);      // comm_idx
    b.take<uint32_t>(E);      // touch_idx
    b.take<uint32_t>(E);      // lost_idx (upper bound)
    b.take<uint8_t>(tmp_sel);
    need0 = b.off + 256;
  }
  // (the item arrays are sized after the counts are known; reserve generously now to avoid a
  // second allocation in the common case: items <= E + O + seeds)
  const size_t N_max = static_cast<size_t>(E) + O + v->n_seed;
  cub::DeviceRadixSort::SortPairs(nullptr, tmp_sort, static_cast<uint64_t*>(nullptr), static_cast<uint64_t*>(nullptr),
                                  static_cast<uint32_t*>(nullptr), static_cast<uint32_t*>(nullptr),
                                  static_cast<int>(N_max), 0, 64, st);
  cub::DeviceScan::ExclusiveSum(nullptr, tmp_scan, static_cast<uint32_t*>(nullptr), static_cast<uint32_t*>(nullptr),
                                static_cast<int>(N_max), st);
  cub::DeviceScan::InclusiveScan(nullptr, t2, static_cast<uint32_t*>(nullptr), static_cast<uint32_t*>(nullptr),
                                 LastNonZero(), static_cast<int>(N_max), st);
  tmp_scan = std::max(tmp_scan, t2);
  size_t need1;
  {
    Bump b{nullptr};
    b.off = need0;
    b.take<uint64_t>(N_max);  // key
    b.take<uint64_t>(N_max);  // key_s
    b.take<uint32_t>(N_max);  // src
    b.take<uint32_t>(N_max);  // pos_iota
    b.take<uint32_t>(N_max);  // pos_s
    b.take<uint32_t>(N_max);  // newflag
    b.take<uint32_t>(N_max);  // rank
    b.take<uint32_t>(N_max);  // def
    b.take<uint32_t>(N_max);  // vid_s
    b.take<uint32_t>(N_max);  // out_tid
    b.take<uint32_t>(N_max);  // out_vid
    b.take<uint32_t>(N_max);  // touch_pos
    b.take<uint8_t>(std::max(tmp_sort, tmp_scan));
    need1 = b.off + 256;
  }
  if (need1 > v->cap_scratch || !v->scratch) {
    // allocate first, then swap: a failed cudaMalloc leaves the old buffer and its capacity intact
    void* fresh = nullptr;
    VK(cudaMalloc(&fresh, need1));
    if (v->scratch) cudaFree(v->scratch);
    v->scratch = fresh;
    v->cap_scratch = need1;
  }
  Bump b{static_cast<uint8_t*>(v->scratch)};
  uint32_t* op_gidx = b.take<uint32_t>(O);
  uint32_t* comm_flag = b.take<uint32_t>(O);
  uint32_t* comm_idx = b.take<uint32_t>(O);
  uint32_t* touch_idx = b.take<uint32_t>(E);
  uint32_t* lost_idx = b.take<uint32_t>(E);
  uint8_t* tmp0 = b.take<uint8_t>(tmp_sel);
  b.off = need0;
  uint64_t* key = b.take<uint64_t>(N_max);
  uint64_t* key_s = b.take<uint64_t>(N_max);
  uint32_t* src = b.take<uint32_t>(N_max);
  uint32_t* pos_iota = b.take<uint32_t>(N_max);
  uint32_t* pos_s = b.take<uint32_t>(N_max);
  uint32_t* newflag = b.take<uint32_t>(N_max);
  uint32_t* rank = b.take<uint32_t>(N_max);
  uint32_t* def = b.take<uint32_t>(N_max);
  uint32_t* vid_s = b.take<uint32_t>(N_max);
  uint32_t* out_tid = b.take<uint32_t>(N_max);
  uint32_t* out_vid = b.take<uint32_t>(N_max);
  uint32_t* touch_pos = b.take<uint32_t>(N_max);
  uint8_t* tmp1 = b.take<uint8_t>(std::max(tmp_sort, tmp_scan));

  // ---- selections ----
  if (O) {
    k_vid_ops<<<(O + 255) / 256, 256, 0, st>>>(W.ops, O, W.d_ev_base, W.d_op_base, W.ns, op_gidx, comm_flag);
    size_t t = tmp_sel;
    cub::DeviceSelect::Flagged(tmp0, t, iota, comm_flag, comm_idx, v->d_small + 3, static_cast<int>(O), st);
  }
  if (E) {
    size_t t = tmp_sel;
    cub::DeviceSelect::If(tmp0, t, iota, touch_idx, v->d_small + 4, static_cast<int>(E), is_touch, st);
    t = tmp_sel;
    cub::DeviceSelect::If(tmp0, t, iota, lost_idx, v->d_small + 5, static_cast<int>(E), is_lost, st);
  }
  VK(cudaGetLastError());
  unsigned int h_small[8] = {0};
  VK(cudaMemcpyAsync(h_small, v->d_small, sizeof(h_small), cudaMemcpyDeviceToHost, st));
  VK(cudaStreamSynchronize(st));
  const uint32_t n_comm = h_small[3], n_touch = h_small[4], n_lost = h_small[5];
  const uint32_t N = v->n_seed + n_touch + n_comm;
  if (static_cast<uint64_t>(v->n_seed) + n_touch + n_comm > SRC_IDX_MASK)
    return vfail(v, -EOVERFLOW, "more than 2^28 touches of thread ids in one window");

  // the thread table can grow by at most one entry per touch; names by one per COMM
  if (int r = grow(v, &v->info, &v->cap_info, static_cast<size_t>(v->next_vid - 1) + n_touch + n_comm + 1,
                   v->next_vid - 1, st))
    return r;
  if (int r = grow(v, &v->names, &v->cap_names, static_cast<size_t>(v->n_names) + n_comm + 1, v->n_names, st)) return r;
  VK(cudaMemcpyAsync(v->d_small, &v->n_names, sizeof(unsigned int), cudaMemcpyHostToDevice, st));

  VidArgs A{};
  A.ev = W.events;
  A.ev_out = W.events;
  A.pids = W.pids;
  A.ops = W.ops;
  A.op_gidx = op_gidx;
  A.comm_idx = comm_idx;
  A.touch_idx = touch_idx;
  A.lost_idx = lost_idx;
  A.seed_tid = v->seed_tid;
  A.seed_vid = v->seed_vid;
  A.n_ops = O;
  A.n_comm = n_comm;
  A.n_touch = n_touch;
  A.n_lost = n_lost;
  A.n_seed = v->n_seed;
  A.n_items = N;
  A.next_vid0 = v->next_vid;
  A.key = key;
  A.src = src;
  A.pos_iota = pos_iota;
  A.touch_pos = touch_pos;
  A.vid_pos = pos_iota;  // (the sort has consumed pos_iota by the time k_vid_apply writes here)
  A.key_s = key_s;
  A.pos_s = pos_s;
  A.newflag = newflag;
  A.rank = rank;
  A.def = def;
  A.vid_s = vid_s;
  A.info = v->info;
  A.names = v->names;
  A.n_names = v->d_small + 0;
  A.out_tid = out_tid;
  A.out_vid = out_vid;
  A.n_out = v->d_small + 1;
  A.flags = v->d_small + 2;

  uint32_t n_new = 0, n_out = 0;
  if (N) {
    const unsigned grid = (N + 255) / 256;
    // The sort key is (epoch, tid): radix passes only over the bits in use.  Thread ids are below
    // 2^22 on Linux (pid_max); 24 bits are tried first, and a window that holds a wider one (the
    // items kernel notices) is keyed again with all 32.
    int epoch_bits = 1;
    while ((static_cast<uint64_t>(n_lost) >> epoch_bits) != 0) ++epoch_bits;
    A.tid_bits = 24;
    k_vid_items<<<grid, 256, 0, st>>>(A);
    unsigned int wide = 0;
    VK(cudaMemcpyAsync(&wide, v->d_small + 2, sizeof(wide), cudaMemcpyDeviceToHost, st));
    VK(cudaStreamSynchronize(st));
    if (wide & VF_WIDE_TID) {
      A.tid_bits = 32;
      k_vid_items<<<grid, 256, 0, st>>>(A);
    }
    size_t t = tmp_sort;
    cub::DeviceRadixSort::SortPairs(tmp1, t, key, key_s, pos_iota, pos_s, static_cast<int>(N), 0,
                                    static_cast<int>(A.tid_bits) + epoch_bits, st);
    k_vid_new<<<grid, 256, 0, st>>>(A);
    t = tmp_scan;
    cub::DeviceScan::ExclusiveSum(tmp1, t, newflag, rank, static_cast<int>(N), st);
    k_vid_def<<<grid, 256, 0, st>>>(A);
    t = tmp_scan;
    cub::DeviceScan::InclusiveScan(tmp1, t, def, vid_s, LastNonZero(), static_cast<int>(N), st);
    k_vid_create<<<grid, 256, 0, st>>>(A);
    k_vid_apply<<<grid, 256, 0, st>>>(A);
    if (n_touch) k_vid_tag<<<(n_touch + 255) / 256, 256, 0, st>>>(A);
    VK(cudaGetLastError());
    // number of vids created = rank of the last position + its flag
    uint32_t last_rank = 0, last_flag = 0;
    VK(cudaMemcpyAsync(&last_rank, rank + (N - 1), 4, cudaMemcpyDeviceToHost, st));
    VK(cudaMemcpyAsync(&last_flag, newflag + (N - 1), 4, cudaMemcpyDeviceToHost, st));
    VK(cudaMemcpyAsync(h_small, v->d_small, sizeof(h_small), cudaMemcpyDeviceToHost, st));
    VK(cudaStreamSynchronize(st));
    n_new = last_rank + last_flag;
    n_out = h_small[1];
    if (h_small[2] & VF_INTERNAL) return vfail(v, -EIO, "internal: a touch was left without a virtual id");
    if (h_small[2] & (VF_TID0 | VF_DOUBLE_FORK))
      return vfail(v, -EINVAL, "thread records outside the reference's contract (%s)",
                   (h_small[2] & VF_TID0) ? "COMM/EXIT/FORK of tid 0" : "FORK of an already forked active tid");
    v->n_names = h_small[0];
  }
  v->next_vid += n_new;
  // the tids still active become the seeds of the next window
  if (int r = grow(v, &v->seed_tid, &v->cap_seed_tid, std::max<size_t>(n_out, 1), 0, st)) return r;
  if (int r = grow(v, &v->seed_vid, &v->cap_seed_vid, std::max<size_t>(n_out, 1), 0, st)) return r;
  if (n_out) {
    VK(cudaMemcpyAsync(v->seed_tid, out_tid, n_out * 4ull, cudaMemcpyDeviceToDevice, st));
    VK(cudaMemcpyAsync(v->seed_vid, out_vid, n_out * 4ull, cudaMemcpyDeviceToDevice, st));
  }
  v->n_seed = n_out;
  *W.vids_done = true;
  VK(cudaEventRecord(v->ev1, st));
  VK(cudaStreamSynchronize(st));
  VK(cudaEventElapsedTime(&v->last_ms, v->ev0, v->ev1));
  return 0;
}

uint64_t hbtgpu_cs_vids_count(const hbtgpu_cs_vids* v) { return v ? v->next_vid - 1 : 0; }
uint64_t hbtgpu_cs_vids_name_count(const hbtgpu_cs_vids* v) { return v ? v->n_names : 0; }
float hbtgpu_cs_vids_last_ms(const hbtgpu_cs_vids* v) { return v ? v->last_ms : 0.f; }

int hbtgpu_cs_vids_get(hbtgpu_cs_vids* v, hbtgpu_cs_thread* threads, uint64_t thread_cap, hbtgpu_cs_name* names,
                       uint64_t name_cap) {
  if (!v) return -EINVAL;
  VK(cudaSetDevice(v->device));
  const uint64_t nt = std::min<uint64_t>(v->next_vid - 1, thread_cap), nn = std::min<uint64_t>(v->n_names, name_cap);
  if (threads && nt) VK(cudaMemcpy(threads, v->info, nt * sizeof(hbtgpu_cs_thread), cudaMemcpyDeviceToHost));
  if (names && nn) {
    VK(cudaMemcpy(names, v->names, nn * sizeof(hbtgpu_cs_name), cudaMemcpyDeviceToHost));
    // grouped by thread, each thread's in series (time) order, like hbtgpu_cs_threads_get
    std::stable_sort(names, names + nn, [](const hbtgpu_cs_name& a, const hbtgpu_cs_name& b) {
      return a.vid != b.vid ? a.vid < b.vid : a.tstamp < b.tstamp;
    });
  }
  if (threads && names && nn == v->n_names) {
    for (uint64_t i = 0; i < nt; ++i) threads[i].n_names = 0;
    for (uint64_t i = 0; i < nn; ++i)
      if (names[i].vid >= 1 && names[i].vid <= nt) threads[names[i].vid - 1].n_names += 1;
  }
  return 0;
}

}  // extern "C"
