"""Time-window sharding across GPUs: the small host-side exchange.

Rank r decodes window r of every cpu with HBTGPU_F_EXTERNAL_CARRY.  The only thing that crosses
ranks is, per cpu, a head stub and a tail (about 1 KB): one all_gather over torch.distributed
(NCCL on GPUs, gloo in the CPU tests), then each rank folds the tails of the ranks before it
with hbtgpu_compose_tail and hands the result to hbtgpu_apply_boundaries.  torch.distributed is
plumbing here; the arithmetic lives in libhbtgpu.so.
"""
import ctypes as C

import numpy as np

from . import hbtgpu


def pack_stubs(heads, tails):
    """ctypes arrays -> one uint8 numpy array [heads | tails]."""
    a = np.frombuffer(bytes(heads), dtype=np.uint8)
    b = np.frombuffer(bytes(tails), dtype=np.uint8)
    return np.concatenate([a, b])


def unpack_stubs(buf, n_cpus):
    hs = C.sizeof(hbtgpu.HeadStub) * n_cpus
    heads = (hbtgpu.HeadStub * n_cpus).from_buffer_copy(bytes(buf[:hs]))
    tails = (hbtgpu.Carry * n_cpus).from_buffer_copy(bytes(buf[hs:hs + C.sizeof(hbtgpu.Carry) * n_cpus]))
    return heads, tails


def fold_prev_tails(all_stubs, initial, rank):
    """all_stubs[r] = (heads, tails) of rank r; initial = ctypes array of the stream-start state
    per cpu.  Returns the state before rank `rank`'s shard, per cpu."""
    n_cpus = len(initial)
    prev = (hbtgpu.Carry * n_cpus)()
    for cpu in range(n_cpus):
        st = initial[cpu]
        for r in range(rank):
            heads, tails = all_stubs[r]
            st = hbtgpu.compose_tail(st, heads[cpu], tails[cpu])
        prev[cpu] = st
    return prev


def all_gather_stubs(heads, tails, device=None):
    """Exchange stubs over the default process group. Returns list over ranks of (heads, tails)."""
    import torch
    import torch.distributed as dist
    n_cpus = len(heads)
    mine = torch.from_numpy(pack_stubs(heads, tails).copy())
    if device is not None:
        mine = mine.to(device)
    world = dist.get_world_size()
    outs = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(outs, mine)
    return [unpack_stubs(o.cpu().numpy(), n_cpus) for o in outs]


# --- the binner over comp units spread across ranks -----------------------
def merge_bin_freqs(freq_lists, min_distinct_intervals, min_duration):
    """freq_lists: per rank, a structured array with stack_id / duration / num_obs / num_intervals
    (what hbtgpu_islice_bin_get returns; the units of different ranks are disjoint, so the sums are
    MonData::stackIdFreqs over all of them).  Returns (merged, num_bins): merged sorted by stack_id
    with the agreed `bid` (0 = ignored, own bins in ascending stack_id order)."""
    parts = [f for f in freq_lists if len(f)]
    if not parts:
        return np.zeros(0, freq_lists[0].dtype if freq_lists else np.uint64), 1
    allf = np.concatenate(parts)
    order = np.argsort(allf["stack_id"], kind="stable")
    allf = allf[order]
    ids, start = np.unique(allf["stack_id"], return_index=True)
    merged = np.zeros(len(ids), allf.dtype)
    merged["stack_id"] = ids
    for k in ("duration", "num_obs", "num_intervals"):
        merged[k] = np.add.reduceat(allf[k], start)          # wrapping u64 sums, as the reference adds
    own = (merged["num_intervals"] >= np.uint64(min_distinct_intervals)) & (merged["duration"] >= np.uint64(min_duration))
    merged["bid"] = np.where(own, np.cumsum(own), 0).astype(np.uint32)
    return merged, int(own.sum()) + 1


def all_gather_bin_freqs(freqs):
    """Every rank's frequency list (small: one record per distinct stack id)."""
    import torch.distributed as dist
    outs = [None] * dist.get_world_size()
    dist.all_gather_object(outs, freqs)
    return outs


def all_reduce_matrix(matrix):
    """Sum of the ranks' interval x bin matrices (u64, wrapping): `matrix` is a torch int64 tensor
    (on the GPU with NCCL, on the host with gloo), reduced in place."""
    import torch.distributed as dist
    dist.all_reduce(matrix, op=dist.ReduceOp.SUM)
    return matrix
