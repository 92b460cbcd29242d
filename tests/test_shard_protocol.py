"""CPU tests (world_size 2, gloo) of the time-window sharding protocol: the stubs are produced
from the oracle's decode of each shard, exchanged with all_gather, folded with the library's
hbtgpu_compose_tail (host-only), and the union of the shard results plus the boundary rows must
equal the single-stream result bit for bit."""
import os
import struct
import sys

import numpy as np
import pytest

import ringgen as rg
from oracle import pyoracle as po

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
N = 3
T0, BNS, NB = 1_000_000_000, 400_000, 64


def make_streams():
    """3 cpus; cut points chosen to hit the interesting boundary cases."""
    out = []
    for cpu, kw in enumerate([dict(p_lost=0.01), dict(p_lost=0.05, p_throttle=0.05), dict()]):
        data, _ = rg.random_stream(900 + cpu, 220, N, **kw)
        out.append(data)
    # cpu 2: put a LOST exactly at the start of shard 1 later (handled by cut choice below)
    return out


def cuts_for(port, data, fractions):
    offs = port.decode_stream(po.make_params(N), data, po.make_carry())["records"][0]
    idx = [int(len(offs) * f) for f in fractions]
    return [0] + [int(offs[i]) for i in idx] + [len(data)]


def head_stub_from_bytes(hbtgpu, shard):
    h = hbtgpu.HeadStub()
    pos = 0
    while pos < len(shard):
        hdr = struct.unpack_from("<Q", shard, pos)[0]
        typ, size = hdr & 0xFFFFFFFF, hdr >> 48
        if typ == rg.SAMPLE:
            ts, nr, te, tr = struct.unpack_from("<4Q", shard, pos + 8)
            vals = struct.unpack_from(f"<{N}Q", shard, pos + 40)
            h.has_sample, h.offset, h.tstamp, h.te, h.tr = 1, pos, ts, te, tr
            for i, v in enumerate(vals):
                h.value[i] = v
            return h
        if typ == rg.LOST:
            h.lost_before = 1
        pos += size
    return h


def to_gpu_carry(hbtgpu, c):
    g = hbtgpu.Carry()
    g.valid, g.tstamp, g.te, g.tr = c.valid, c.tstamp, c.te, c.tr
    for i in range(64):
        g.last[i] = c.last[i]
    return g


def boundary_row(prev, head, params):
    """handleRecordSample for the first SAMPLE of a shard, given the state before the shard."""
    if not head.has_sample or head.lost_before or not prev.valid or prev.tstamp > head.tstamp:
        return None
    m = (1 << 64) - 1
    b = (head.tstamp - params.t0) // params.bucket_ns if head.tstamp >= params.t0 else -1
    vec = [(head.value[i] - prev.last[i]) & m for i in range(N)] + \
        [(head.te - prev.te) & m, (head.tr - prev.tr) & m, 1]
    return b, vec


def _worker(rank, world, port_path, tmpdir):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from dynolog_b200 import hbtgpu, shard
    dist.init_process_group("gloo", init_method=f"file://{tmpdir}/rdzv", rank=rank, world_size=world)
    port = po.load("port")
    params = po.make_params(N, n_buckets=NB, t0=T0, bucket_ns=BNS)
    streams = make_streams()
    initial = (hbtgpu.Carry * len(streams))()
    for cpu in range(len(streams)):
        initial[cpu] = hbtgpu.make_carry(valid=True, tstamp=T0 - 7, te=3, tr=2, last=[5] * N)
    heads = (hbtgpu.HeadStub * len(streams))()
    tails = (hbtgpu.Carry * len(streams))()
    acc = np.zeros((len(streams), NB, N + 3), np.uint64)
    ctrs = []
    for cpu, data in enumerate(streams):
        c = cuts_for(port, data, [0.5])
        sh = data[c[rank]:c[rank + 1]]
        r = port.decode_stream(params, sh, po.make_carry())  # shard mode: invalid state
        acc[cpu] = r["buckets"]
        ctrs.append(r["counters"])
        heads[cpu] = head_stub_from_bytes(hbtgpu, sh.tobytes())
        tails[cpu] = to_gpu_carry(hbtgpu, r["carry_out"])
    gathered = shard.all_gather_stubs(heads, tails)
    prev = shard.fold_prev_tails(gathered, initial, rank)
    rows = unknown = 0
    for cpu in range(len(streams)):
        br = boundary_row(prev[cpu], heads[cpu], params)
        rows += ctrs[cpu]["num_rows"]
        unknown += ctrs[cpu]["num_unknown_count_interval"]
        if br is not None:
            b, vec = br
            rows += 1
            unknown -= 1
            if 0 <= b < NB:
                acc[cpu, b] += np.array(vec, np.uint64)
    t = torch.from_numpy(acc.view(np.int64).copy())
    dist.all_reduce(t)  # wrapping i64 sum == u64 sum
    cnt = torch.tensor([rows, unknown], dtype=torch.int64)
    dist.all_reduce(cnt)
    if rank == 0:
        np.save(os.path.join(tmpdir, "acc.npy"), t.numpy().view(np.uint64))
        np.save(os.path.join(tmpdir, "cnt.npy"), cnt.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_window_sharding_gloo(port_oracle, tmp_path):
    import torch.multiprocessing as mp
    pytest.importorskip("torch")
    from dynolog_b200 import hbtgpu
    if not os.path.exists(hbtgpu.LIB_PATH):
        pytest.skip("libhbtgpu.so not built")
    mp.spawn(_worker, args=(2, port_oracle.path, str(tmp_path)), nprocs=2, join=True)
    acc = np.load(tmp_path / "acc.npy")
    cnt = np.load(tmp_path / "cnt.npy")
    params = po.make_params(N, n_buckets=NB, t0=T0, bucket_ns=BNS)
    rows = unknown = 0
    for cpu, data in enumerate(make_streams()):
        full = port_oracle.decode_stream(params, data, po.make_carry(valid=True, tstamp=T0 - 7, te=3,
                                                                    tr=2, last=[5] * N))
        assert np.array_equal(acc[cpu], full["buckets"]), f"cpu {cpu}: sharded sums differ"
        rows += full["counters"]["num_rows"]
        unknown += full["counters"]["num_unknown_count_interval"]
    assert [rows, unknown] == [int(cnt[0]), int(cnt[1])]


def test_compose_tail_cases():
    """hbtgpu_compose_tail mirrors handleRecordLost/handleRecordSample on shard summaries."""
    from dynolog_b200 import hbtgpu
    if not os.path.exists(hbtgpu.LIB_PATH):
        pytest.skip("libhbtgpu.so not built")
    prev = hbtgpu.make_carry(valid=True, tstamp=100, te=9, tr=8, last=[1, 2, 3])
    own = hbtgpu.make_carry(valid=True, tstamp=500, te=90, tr=80, last=[10, 20, 30])
    h = hbtgpu.HeadStub()
    h.has_sample = 1
    out = hbtgpu.compose_tail(prev, h, own)   # a shard with a SAMPLE decides by itself
    assert (out.valid, out.tstamp, out.last[2]) == (1, 500, 30)
    h = hbtgpu.HeadStub()                      # empty / no SAMPLE, no LOST: pass through
    out = hbtgpu.compose_tail(prev, h, own)
    assert (out.valid, out.tstamp, out.te, out.last[1]) == (1, 100, 9, 2)
    h.lost_before = 1                          # no SAMPLE but a LOST: invalidated, last_ kept
    out = hbtgpu.compose_tail(prev, h, own)
    assert out.valid == 0 and out.last[0] == 1


# ---- the binner over units spread across ranks (SURVEY 8(f) f1 x 8(e)) ------------------------------
def _bin_case(seed=5, n_units=6):
    import slicegen
    from oracle import pyoracle as po
    r = np.random.default_rng(seed)
    uc, us = [], []
    for u in range(n_units):
        s = slicegen.gen_slices(r, int(r.integers(50, 400)), hole_prob=0.05, n_stacks=int(r.choice([5, 60])))
        rows = slicegen.gen_rows(r, s, int(r.integers(10, 120)), align_prob=0.2)
        res = po.islice_release("port", s, rows, 2**63, 3)
        uc.append(res["counts"])
        us.append(res["slices"])
    return uc, us


def _local_matrix_in_global_bins(po, uc, us, merged, num_bins, num_rows):
    """what hbtgpu_islice_bin_with computes on a rank, from the restatement: own numbering with
    thresholds 0 / 0, columns moved to the agreed bins, rows padded."""
    loc = po.bin_durations("port", uc, us, 0, 0)
    m = np.zeros((num_rows, num_bins), np.uint64)
    bid_of = dict(zip(merged["stack_id"].tolist(), merged["bid"].tolist()))
    for f in loc["freqs"]:
        col = loc["durations"][:, int(f["bid"])]
        m[: len(col), bid_of[int(f["stack_id"])]] += col
    return loc["freqs"], m


def test_merge_bin_freqs_matches_one_process():
    from dynolog_b200 import shard
    from oracle import pyoracle as po
    uc, us = _bin_case()
    for min_int, min_dur in [(0, 0), (2, 50), (5, 10**6)]:
        whole = po.bin_durations("port", uc, us, min_int, min_dur)
        parts = [po.bin_durations("port", uc[:2], us[:2], 0, 0)["freqs"], po.bin_durations("port", uc[2:], us[2:], 0, 0)["freqs"]]
        merged, num_bins = shard.merge_bin_freqs(parts, min_int, min_dur)
        for k in ("stack_id", "duration", "num_obs", "num_intervals", "bid"):
            assert np.array_equal(merged[k], whole["freqs"][k]), k
        assert num_bins == whole["num_bins"]
        num_rows = whole["num_rows"]
        m = sum(_local_matrix_in_global_bins(po, c, s, merged, num_bins, num_rows)[1] for c, s in ((uc[:2], us[:2]), (uc[2:], us[2:])))
        assert np.array_equal(m, whole["durations"])


def _bin_worker(rank, world, tmpdir):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from dynolog_b200 import shard
    from oracle import pyoracle as po
    dist.init_process_group("gloo", init_method=f"file://{tmpdir}/rdzv_bin", rank=rank, world_size=world)
    try:
        uc, us = _bin_case()
        mine = slice(0, 3) if rank == 0 else slice(3, None)
        whole = po.bin_durations("port", uc, us, 2, 50)
        local_freqs = po.bin_durations("port", uc[mine], us[mine], 0, 0)["freqs"]
        merged, num_bins = shard.merge_bin_freqs(shard.all_gather_bin_freqs(local_freqs), 2, 50)
        rows = torch.tensor([sum(len(c) for c in uc[mine])], dtype=torch.int64)
        dist.all_reduce(rows)
        _, m = _local_matrix_in_global_bins(po, uc[mine], us[mine], merged, num_bins, int(rows[0]))
        t = torch.from_numpy(m.view(np.int64).copy())
        shard.all_reduce_matrix(t)
        assert np.array_equal(t.numpy().view(np.uint64), whole["durations"])
        assert np.array_equal(merged["bid"], whole["freqs"]["bid"])
    finally:
        dist.destroy_process_group()


def test_two_rank_binner_gloo(port_oracle, tmp_path):
    import torch.multiprocessing as mp
    mp.spawn(_bin_worker, args=(2, str(tmp_path)), nprocs=2, join=True)
