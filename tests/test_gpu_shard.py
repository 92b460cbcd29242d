"""-m gpu: time-window sharding (HBTGPU_F_EXTERNAL_CARRY).  Several "ranks" are emulated as
separate contexts on one GPU: each decodes one window of every cpu, stubs are folded with
hbtgpu_compose_tail, hbtgpu_apply_boundaries adds the straddling rows; the union must equal the
oracle's single-stream result bit for bit."""
import numpy as np
import pytest

import ringgen as rg
from oracle import pyoracle as po

pytestmark = pytest.mark.gpu
hbtgpu = pytest.importorskip("dynolog_b200.hbtgpu")
from dynolog_b200 import shard  # noqa: E402
from gpu_util import to_gpu_carry  # noqa: E402

N = 4
T0, BNS = 1_000_000_000, 1_000_000


def _streams():
    s = []
    for cpu, kw in enumerate([dict(), dict(p_lost=0.01), dict(p_lost=0.03, p_throttle=0.02, mux=True),
                              dict(p_other=0.1)]):
        data, info = rg.random_stream(4000 + cpu, 9000, N, **kw)
        s.append((data, info["t_end"]))
    # cpu 4: a window that holds only non-SAMPLE records, and a LOST right at a window start
    parts = [rg.sample(T0 + 1000 * i, i, i, [i, 2 * i, 3 * i, 4 * i]) for i in range(1, 40)]
    mid = [rg.throttle(1), rg.unthrottle(2), rg.other(33, [1, 2, 3])]
    parts2 = [rg.lost(3)] + [rg.sample(T0 + 50_000 + 1000 * i, 50 + i, 50 + i, [50 + i] * 4) for i in range(40)]
    s.append(((parts, mid, parts2), T0 + 100_000))
    return s


def _cut(port, data, fr):
    offs = port.decode_stream(po.make_params(N), data, po.make_carry())["records"][0]
    return [0] + [int(offs[int(len(offs) * f)]) for f in fr] + [len(data)]


@pytest.mark.parametrize("device_fold", [False, True])
@pytest.mark.parametrize("flags", [0, hbtgpu.F_FORCE_EXACT])
def test_three_window_shards_match_single_stream(port_oracle, flags, device_fold):
    world = 3
    raw = _streams()
    streams, shards = [], []
    for cpu, (d, _) in enumerate(raw):
        if isinstance(d, tuple):
            pieces = [rg.to_u8(b"".join(p)) for p in d]
            streams.append(np.concatenate(pieces))
            shards.append(pieces)
        else:
            c = _cut(port_oracle, d, [0.31, 0.77])
            streams.append(d)
            shards.append([d[c[i]:c[i + 1]] for i in range(world)])
    t_end = max(t for _, t in raw)
    nb = int((t_end - T0) // BNS) + 2
    params = po.make_params(N, n_buckets=nb, t0=T0, bucket_ns=BNS)
    ncpu = len(streams)
    init = [po.make_carry(valid=(cpu != 1), tstamp=T0 - 3, te=1, tr=1, last=[2] * N) for cpu in range(ncpu)]
    initial = (hbtgpu.Carry * ncpu)()
    for cpu in range(ncpu):
        initial[cpu] = to_gpu_carry(init[cpu])
    ctxs = []
    try:
        for r in range(world):
            total = sum(len(shards[c][r]) for c in range(ncpu))
            g = hbtgpu.Context(n_events=N, max_cpus=ncpu, n_buckets=nb, t0=T0, bucket_ns=BNS,
                               arena_bytes=total + (1 << 16),
                               flags=flags | hbtgpu.F_EXTERNAL_CARRY)
            for c in range(ncpu):
                g.submit_image(c, shards[c][r])
            g.run()
            g.sync()
            ctxs.append(g)
        stubs = [g.get_stubs() for g in ctxs]
        # device-side variant of the same exchange: blocks "gathered" into one device buffer
        import torch
        sb = ctxs[0].stub_bytes()
        gathered = torch.empty(world * sb, dtype=torch.uint8, device="cuda:0")
        for r, g in enumerate(ctxs):
            g.export_stubs_device(gathered.data_ptr() + r * sb)
            g.sync()
        acc = np.zeros((ncpu, nb, N + 3), np.uint64)
        rows = np.zeros(ncpu, np.int64)
        unknown = np.zeros(ncpu, np.int64)
        lost = np.zeros(ncpu, np.int64)
        for r, g in enumerate(ctxs):
            if device_fold:
                for c in range(ncpu):
                    g.set_carry(c, initial[c])
                g.apply_boundaries_gathered(gathered.data_ptr(), world, r)
            else:
                prev = shard.fold_prev_tails(stubs, initial, r)
                g.apply_boundaries(prev)
            g.sync()
            acc += g.buckets(want_scaled=False, want_metric=False, want_valid=False)["acc"]
            for c in range(ncpu):
                k = g.counters(c)
                rows[c] += k["num_rows"]
                unknown[c] += k["num_unknown_count_interval"]
                lost[c] += k["lost"]
        final = shard.fold_prev_tails(stubs, initial, world)  # state after the last shard
    finally:
        for g in ctxs:
            g.close()
    for c in range(ncpu):
        exp = port_oracle.decode_stream(params, streams[c], init[c])
        assert np.array_equal(acc[c], exp["buckets"]), f"cpu {c}: sharded bucket sums differ"
        assert rows[c] == exp["counters"]["num_rows"], f"cpu {c}"
        assert unknown[c] == exp["counters"]["num_unknown_count_interval"], f"cpu {c}"
        assert lost[c] == exp["counters"]["lost"]
        co = exp["carry_out"]
        assert int(final[c].valid) == int(co.valid)
        if co.valid:
            assert int(final[c].tstamp) == int(co.tstamp)
        assert [int(final[c].last[i]) for i in range(N)] == [int(co.last[i]) for i in range(N)]


def test_derive_before_boundaries_equals_derive_after():
    """A caller may derive right after hbtgpu_run (to overlap it with the stub exchange);
    hbtgpu_apply_boundaries* then re-derives only the cells the boundary rows touched.  Every
    output must equal the plain order (boundaries, then one full derive)."""
    n, ncpu, nb = 4, 5, 300
    streams = [rg.random_stream(200 + i, 3000, n, mux=True, p_lost=0.01 if i == 3 else 0.0)[0] for i in range(ncpu)]
    prev = (hbtgpu.Carry * ncpu)()
    for c in range(ncpu):
        prev[c] = to_gpu_carry(po.make_carry(valid=(c != 2), tstamp=999_000_000, te=5, tr=3, last=[1] * n))
    outs = []
    for early in (False, True):
        with hbtgpu.Context(n_events=n, max_cpus=ncpu, n_buckets=nb, t0=1_000_000_000, bucket_ns=1_200_000,
                            arena_bytes=sum(len(s) for s in streams) + (1 << 16), flags=hbtgpu.F_EXTERNAL_CARRY,
                            metrics=[(hbtgpu.METRIC_RATE, [1, 0])]) as g:
            for c, s in enumerate(streams):
                g.submit_image(c, s)
            g.run()
            if early:
                g.derive()
            g.apply_boundaries(prev)
            g.sync()
            outs.append(g.buckets())
    for k in ("acc", "scaled", "metric", "valid"):
        assert np.array_equal(outs[0][k].view(np.uint8), outs[1][k].view(np.uint8)), k
    assert outs[0]["scaled"].any()


# ---- exchange + merge over peer memory (hbtgpu_peer_*, hbtgpu_exchange_boundaries, hbtgpu_merge) ----
def _run_peer_world(shards, init, nb, n=N, flags=0, metrics=(), merge=True, bucket_ns=BNS):
    """shards[cpu][rank] = bytes.  One context per rank on this GPU, attached to each other as
    peers; returns per rank (buckets dict, [counters per cpu], [carry per cpu])."""
    ncpu, world = len(shards), len(shards[0])
    ctxs, out = [], []
    try:
        for r in range(world):
            total = sum(len(shards[c][r]) for c in range(ncpu))
            g = hbtgpu.Context(n_events=n, max_cpus=ncpu, n_buckets=nb, t0=T0, bucket_ns=bucket_ns,
                               arena_bytes=total + (1 << 16), metrics=metrics,
                               flags=flags | hbtgpu.F_EXTERNAL_CARRY)
            for c in range(ncpu):
                g.set_carry(c, to_gpu_carry(init[c]))
                g.submit_image(c, shards[c][r])
            ctxs.append(g)
        for r, g in enumerate(ctxs):
            g.peer_attach_local(world, r, ctxs)
        for rep in range(2):  # twice: the mailbox copies alternate, results must not change
            for g in ctxs:
                g.run()
            for g in ctxs:  # nothing below blocks the host: the waits of one rank are met by the others
                g.exchange_boundaries(no_derive=merge)
            if merge:
                for g in ctxs:
                    g.merge()
            for g in ctxs:
                g.sync()
        for g in ctxs:
            out.append((g.buckets(), [g.counters(c) for c in range(ncpu)], [g.carry(c) for c in range(ncpu)]))
    finally:
        for g in ctxs:
            g.close()
    return out


@pytest.mark.parametrize("flags", [0, hbtgpu.F_FORCE_EXACT])
def test_peer_exchange_and_merge_match_single_stream(port_oracle, flags):
    world = 3
    raw = _streams()
    streams, shards = [], []
    for cpu, (d, _) in enumerate(raw):
        if isinstance(d, tuple):
            pieces = [rg.to_u8(b"".join(p)) for p in d]
            streams.append(np.concatenate(pieces))
            shards.append(pieces)
        else:
            c = _cut(port_oracle, d, [0.31, 0.77])
            streams.append(d)
            shards.append([d[c[i]:c[i + 1]] for i in range(world)])
    t_end = max(t for _, t in raw)
    nb = int((t_end - T0) // BNS) + 2
    params = po.make_params(N, n_buckets=nb, t0=T0, bucket_ns=BNS)
    ncpu = len(streams)
    init = [po.make_carry(valid=(cpu != 1), tstamp=T0 - 3, te=1, tr=1, last=[2] * N) for cpu in range(ncpu)]
    res = _run_peer_world(shards, init, nb, flags=flags, metrics=[(hbtgpu.METRIC_RATE, [1, 0])])
    for c in range(ncpu):
        exp = port_oracle.decode_stream(params, streams[c], init[c])
        scaled = port_oracle.scale_buckets(params, exp["buckets"])
        for r in range(world):  # every rank holds the totals
            assert np.array_equal(res[r][0]["acc"][c], exp["buckets"]), f"cpu {c} rank {r}: merged sums differ"
            assert np.array_equal(res[r][0]["scaled"][c], scaled), f"cpu {c} rank {r}: scaled counts differ"
        for k in ("num_rows", "num_unknown_count_interval", "lost", "num_records", "sample", "throttle"):
            assert sum(res[r][1][c][k] for r in range(world)) == exp["counters"][k], (c, k)
        co = exp["carry_out"]  # the last rank's carry-out is the stream's
        last = res[world - 1][2][c]
        assert int(last.valid) == int(co.valid)
        if co.valid:
            assert int(last.tstamp) == int(co.tstamp)
        assert [int(last.last[i]) for i in range(N)] == [int(co.last[i]) for i in range(N)]


def test_stalled_walk_masks_the_shards_behind_it(port_oracle):
    """The reference stalls a cpu for good at the first record consume() fails on
    (PerfEventsGroup.h:1505, PerCpuCountSampleGenerator.h:156-163).  Sharded, the ranks behind the
    stall must contribute nothing, and a shard whose first SAMPLE goes back in time against the
    previous shard must contribute only the records in front of that SAMPLE."""
    world = 3

    def run(k0, k1):
        return [rg.sample(T0 + 1000 * i, i, i, [i, 2 * i, 3 * i, 4 * i]) for i in range(k0, k1)]

    shards = []
    # cpu 0: a SAMPLE with a wrong size in the middle of shard 1
    shards.append([b"".join(run(1, 40)),
                   b"".join(run(40, 60) + [rg.sample(T0 + 60_000, 60, 60, [1, 2, 3, 4], size=64)] + run(61, 80)),
                   b"".join(run(80, 120))])
    # cpu 1: time goes back exactly at the boundary between shard 1 and shard 2
    shards.append([b"".join(run(1, 40)), b"".join(run(40, 80)), b"".join(run(70, 110))])
    # cpu 2: the same with THROTTLE / UNTHROTTLE / another record in front of the offending SAMPLE
    shards.append([b"".join(run(1, 40)), b"".join(run(40, 80)),
                   b"".join([rg.throttle(1), rg.unthrottle(2), rg.other(33, [5, 6])] + run(75, 110))])
    # cpu 3: time goes back at the boundary, but a LOST in front re-bases: no stall
    shards.append([b"".join(run(1, 40)), b"".join(run(40, 80)), b"".join([rg.lost(9)] + run(70, 110))])
    # cpu 4: healthy
    shards.append([b"".join(run(1, 40)), b"".join(run(40, 80)), b"".join(run(80, 120))])
    # cpu 5: time goes back at the FIRST boundary: shards 1 and 2 are both behind the stall
    shards.append([b"".join(run(10, 40)), b"".join(run(5, 80)), b"".join(run(80, 120))])
    shards = [[rg.to_u8(x) for x in s] for s in shards]
    ncpu = len(shards)
    nb = 4
    params = po.make_params(N, n_buckets=nb, t0=T0, bucket_ns=50_000)
    init = [po.make_carry(valid=True, tstamp=T0, te=0, tr=0, last=[0] * N) for _ in range(ncpu)]
    res = _run_peer_world(shards, init, nb, bucket_ns=50_000)
    for c in range(ncpu):
        exp = port_oracle.decode_stream(params, np.concatenate(shards[c]), init[c])
        for r in range(world):
            assert np.array_equal(res[r][0]["acc"][c], exp["buckets"]), f"cpu {c} rank {r}: merged sums differ"
        for k in ("num_rows", "num_unknown_count_interval", "lost", "num_records", "sample", "throttle", "unthrottle"):
            assert sum(res[r][1][c][k] for r in range(world)) == exp["counters"][k], (c, k)
        assert (exp["counters"]["err"] != 0) == any(res[r][1][c]["err"] != 0 for r in range(world)), c
    # the ranks behind a stall say so
    assert res[2][1][0]["err"] != 0 and res[2][1][0]["num_records"] == 0
    assert res[2][1][1]["err"] != 0 and res[2][1][2]["num_records"] == 3
    assert res[1][1][5]["err"] != 0 and res[2][1][5]["err"] != 0
    assert res[2][1][3]["err"] == 0 and res[2][1][4]["err"] == 0
