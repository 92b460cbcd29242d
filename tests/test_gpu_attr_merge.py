"""-m gpu: BASELINE config 4 over several GPUs (last bullet) -- hbtgpu_attr_merge.
The capture is sharded by time window over ranks (contexts of one process on one GPU here; the
exchange is the same code as between processes, see tests/test_gpu_shard.py); after the merge
every rank holds the (pid,tid) histogram and the top-100 sampled IPs of the WHOLE capture.
Checker: the C restatement's walk + numpy counting (spec extension: parity unpinned by
construction, as for config 4 on one GPU)."""
import numpy as np
import pytest

from dynolog_b200 import synth
from oracle import pyoracle as po

pytestmark = pytest.mark.gpu
hbtgpu = pytest.importorskip("dynolog_b200.hbtgpu")


def _expected(port, p, windows, n):
    tids, ips = [], []
    for streams in windows:
        for s in streams:
            a, b = port.collect_ip_tid(p, s, po.make_carry())
            tids.append(a)
            ips.append(b)
    tids, ips = np.concatenate(tids), np.concatenate(ips)
    tk, tc = np.unique(tids, return_counts=True)
    hist = {(int(k & 0xFFFFFFFF), int(k >> 32)): int(c) for k, c in zip(tk, tc)}
    ik, ic = np.unique(ips, return_counts=True)
    order = np.lexsort((ik, -ic.astype(np.int64)))[:n]
    return hist, [(int(ik[i]), int(ic[i])) for i in order], len(tids)


@pytest.mark.parametrize("world,per_rank,accumulate", [(3, 1, False), (2, 2, True), (4, 2, True)])
def test_attr_merge_equals_whole_capture(port_oracle, world, per_rank, accumulate):
    n, ncpu, per_cpu = 4, 16, 1 << 20
    cap = synth.Capture(synth.make_spec(4), ncpu)
    t0, nb = synth.T_START, 4000
    p = po.make_params(n, n_buckets=nb, t0=t0, bucket_ns=1_000_000, pre=2)
    windows = [[s.copy() for s in cap.window(max_bytes_per_cpu=per_cpu)] for _ in range(world * per_rank)]
    exp_hist, exp_top, n_samples = _expected(port_oracle, p, windows, 100)
    assert n_samples > 100_000 * world
    ctxs = []
    try:
        for r in range(world):
            ctxs.append(hbtgpu.Context(n_events=n, max_cpus=ncpu, n_buckets=nb, t0=t0, bucket_ns=1_000_000,
                                       arena_bytes=ncpu * per_cpu + (1 << 20),
                                       sample_format=hbtgpu.SAMPLE_IP_TID_TIME_READ,
                                       flags=hbtgpu.F_EXTERNAL_CARRY | (hbtgpu.F_ACCUMULATE_ATTR if accumulate else 0)))
        for r, g in enumerate(ctxs):
            g.peer_attach_local(world, r, ctxs)
        for rep in range(2):
            for r, g in enumerate(ctxs):
                for k in range(per_rank):
                    g.reset(keep_carry=True)
                    g.submit_images(list(range(ncpu)), windows[r * per_rank + k])
                    g.run()
                    assert accumulate or k == 0
            for g in ctxs:   # (nothing blocks the host: the waits of one rank are met by the others)
                g.attr_merge()
            for g in ctxs:
                g.sync()
            mult = (rep + 1) if accumulate else 1     # totals keep adding up until attr_clear
            for g in ctxs:
                assert g.stats()["exact_runs"] == 0
                assert g.tid_histogram() == {k: v * mult for k, v in exp_hist.items()}
                assert g.top_ips(100) == [(ip, c * mult) for ip, c in exp_top]
        # a run on one rank: its getters are back to its own tables
        g = ctxs[0]
        g.reset(keep_carry=True)
        g.submit_images(list(range(ncpu)), windows[0])
        g.run()
        g.sync()
        own = g.tid_histogram()
        assert sum(own.values()) < sum(exp_hist.values()) * (3 if accumulate else 1)
    finally:
        for g in ctxs:
            g.close()


def test_attr_merge_needs_the_ip_tid_format():
    ctxs = [hbtgpu.Context(n_events=4, max_cpus=2, n_buckets=10, t0=0, bucket_ns=1000, arena_bytes=1 << 16,
                           flags=hbtgpu.F_EXTERNAL_CARRY) for _ in range(2)]
    try:
        for r, g in enumerate(ctxs):
            g.peer_attach_local(2, r, ctxs)
        for g in ctxs:
            g.run()
        with pytest.raises(hbtgpu.HbtGpuError, match="IP_TID"):
            ctxs[0].attr_merge()
    finally:
        for g in ctxs:
            g.close()
