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
