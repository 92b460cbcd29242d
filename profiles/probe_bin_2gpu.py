"""The binner over comp units spread across GPUs, with the real exchange:
  torchrun --nproc-per-node 2 --master-addr 127.0.0.1 profiles/probe_bin_2gpu.py
Each rank releases its half of the units, the frequency lists are all-gathered and merged, the
bins agreed, every rank builds its matrix with hbtgpu_islice_bin_with and the matrices are summed
by an NCCL all-reduce on the device buffers they were built in.  Rank 0 checks the result against
the C restatement over all units."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import torch
import torch.distributed as dist
import slicegen
from dynolog_b200 import hbtgpu_slices as gs, shard
from oracle import pyoracle as po


class DeviceArray:
    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<i8", "data": (ptr, False), "version": 2}


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dist.init_process_group("nccl")
    r = np.random.default_rng(3)
    n_units = 64
    us = [slicegen.gen_slices(r, 20000, hole_prob=0.02, n_stacks=300) for _ in range(n_units)]
    ur = [slicegen.gen_rows(r, s, 300, stride=3, align_prob=0.05, mean_dur=2000) for s in us]
    mine = range(rank * n_units // world, (rank + 1) * n_units // world)
    isl = gs.IntervalSlicer(torch.cuda.current_device())
    sb = np.concatenate([[0], np.cumsum([len(us[u]) for u in mine])]).astype(np.uint64)
    rb = np.concatenate([[0], np.cumsum([len(ur[u]) for u in mine])]).astype(np.uint64)
    units = isl.release(np.concatenate([us[u] for u in mine]), sb, np.concatenate([ur[u] for u in mine]), rb, 2**63, 3,
                        np.zeros(len(mine), np.uint64))
    rows = torch.tensor([int(units["n_intervals"].sum())], dtype=torch.int64, device="cuda")
    dist.all_reduce(rows)
    torch.cuda.synchronize()
    t0 = time.time()
    local = isl.bin(0, 0, max_matrix_bytes=1 << 33)["freqs"]
    merged, num_bins = shard.merge_bin_freqs(shard.all_gather_bin_freqs(local), 3, 10_000)
    res = isl.bin(0, 0, numbering=(merged["stack_id"], merged["bid"], num_bins, int(rows[0])), fetch=False,
                  max_matrix_bytes=1 << 33)
    cells = res["num_rows"] * res["num_bins"]
    m = torch.as_tensor(DeviceArray(isl.bin_durations_device(), cells), device="cuda")
    shard.all_reduce_matrix(m)
    torch.cuda.synchronize()
    dt = time.time() - t0
    if rank == 0:
        uc, uo = [], []
        for u in range(n_units):
            e = po.islice_release("port", us[u], ur[u], 2**63, 3, 0)
            uc.append(e["counts"]); uo.append(e["slices"])
        exp = po.bin_durations("port", uc, uo, 3, 10_000)
        got = m.cpu().numpy().view(np.uint64).reshape(res["num_rows"], res["num_bins"])
        ok = np.array_equal(got, exp["durations"]) and np.array_equal(merged["bid"], exp["freqs"]["bid"])
        print(f"binner over {world} GPUs: {n_units} units, matrix {res['num_rows']} x {res['num_bins']}, "
              f"frequencies + agreement + matrices + NCCL all-reduce {dt*1e3:.1f} ms; equals the one-process restatement: {ok}")
        assert ok
    dist.barrier()
    isl.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
