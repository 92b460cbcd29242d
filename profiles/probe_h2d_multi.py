"""Host-to-device feed with several ranks on one box (e2e ms per rank 38.7 -> 41 ->
78 -> 103 at N = 1/2/4/8).  Run under torch.distributed.run with N ranks; for k in 1, 2, 4, ... N
only ranks < k copy (the others idle), so one launch gives the whole curve:
  per-rank and aggregate H2D GB/s for  (a) one 2 GB cudaMemcpyAsync  (b) 192 copies of 10.4 MB
  from pinned memory allocated by  torch (cudaHostAlloc default)  /  cudaHostAllocWriteCombined.
Ranks are left where the launcher put them; affinity, NUMA node of the GPU and the allowed
memory nodes are printed (a container may expose one node only)."""
import ctypes as C
import json
import os
import sys
import time

import torch
import torch.distributed as dist

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
rt = C.CDLL("libcudart.so.12")
NB = 192 * 10_400_000  # 2.0 GB
dev = torch.empty(NB, dtype=torch.uint8, device="cuda")
stream = torch.cuda.Stream()


def host_alloc(flags):
    p = C.c_void_p()
    rc = rt.cudaHostAlloc(C.byref(p), C.c_size_t(NB), C.c_uint(flags))
    assert rc == 0, rc
    C.memset(p, 1, NB)  # first touch by this rank
    return p


def copy(hptr, pieces):
    step = NB // pieces
    for i in range(pieces):
        rc = rt.cudaMemcpyAsync(C.c_void_p(dev.data_ptr() + i * step), C.c_void_p(hptr.value + i * step),
                                C.c_size_t(step), C.c_int(1), C.c_void_p(stream.cuda_stream))
        assert rc == 0, rc


def info():
    out = {"rank": rank, "affinity": sorted(os.sched_getaffinity(0))[:4] + ["...", len(os.sched_getaffinity(0))]}
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(lr)
        bus = pynvml.nvmlDeviceGetPciInfo(h).busId
        bus = bus.decode() if isinstance(bus, bytes) else bus
        out["pci"] = bus
        p = f"/sys/bus/pci/devices/{bus.lower()[4:] if len(bus) > 12 else bus.lower()}/numa_node"
        out["gpu_numa_node"] = open(p).read().strip() if os.path.exists(p) else "n/a"
        out["pcie_gen_width"] = [pynvml.nvmlDeviceGetCurrPcieLinkGeneration(h), pynvml.nvmlDeviceGetCurrPcieLinkWidth(h)]
    except Exception as e:
        out["nvml"] = str(e)[:60]
    for line in open("/proc/self/status"):
        if line.startswith("Mems_allowed_list"):
            out["mems_allowed"] = line.split()[1]
    return out


infos = [None] * world
dist.all_gather_object(infos, info())
bufs = {"default": host_alloc(0), "write_combined": host_alloc(4)}
res = []
ks = [k for k in (1, 2, 4, 8, 16) if k <= world]
for kind, hp in bufs.items():
    for pieces in (1, 192):
        for k in ks:
            best = None
            for rep in range(3):
                dist.barrier()
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                if rank < k:
                    copy(hp, pieces)
                    stream.synchronize()
                dt = time.perf_counter() - t0
                t = torch.tensor([dt if rank < k else 0.0], device="cuda", dtype=torch.float64)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                best = float(t.item()) if best is None else min(best, float(t.item()))
            res.append({"memory": kind, "copies": pieces, "active_ranks": k, "per_rank_gbs": round(NB / 1e9 / best, 1),
                        "aggregate_gbs": round(k * NB / 1e9 / best, 1)})
if rank == 0:
    print(json.dumps({"world": world, "ranks": infos, "results": res}))
dist.barrier()
dist.destroy_process_group()
