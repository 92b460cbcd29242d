"""-m gpu: corner cases of the bucket aggregation -- streams without images or without SAMPLEs,
sparse samples, times in front of and behind the bucket range, buckets from one sample wide to
the whole stream, time going back behind a LOST, wide groups, every record mix on every path.
Every case compares ALL cells (sums, scaled counts, metrics, validity) and the counters with the
oracle, twice in a row (a second run must recompute, not accumulate).  (Written while an
aggregation by plain stores -- every bucket written once, derived in the decode kernel -- was
tried: bit-exact, but 0.60 ms against 0.37 ms for the telescoped REDs on BASELINE config 2,
because its per-warp bookkeeping costs more instructions than the zeroed accumulator costs
bandwidth)"""
import numpy as np
import pytest

import ringgen as rg
from oracle import pyoracle as po

pytestmark = pytest.mark.gpu
hbtgpu = pytest.importorskip("dynolog_b200.hbtgpu")
from gpu_util import to_gpu_carry  # noqa: E402

T0 = 1_000_000_000


def _check(port, streams, carries, n, nb, bns, t0=T0, flags=0, expect_exact=None, metrics=((1, [1, 0]),)):
    params = po.make_params(n, n_buckets=nb, t0=t0, bucket_ns=bns)
    total = sum(len(s) for s in streams)
    with hbtgpu.Context(n_events=n, max_cpus=len(streams), n_buckets=nb, t0=t0, bucket_ns=bns,
                        arena_bytes=total + 4096 * (len(streams) + 8), metrics=list(metrics), flags=flags) as g:
        for cpu, s in enumerate(streams):
            g.set_carry(cpu, to_gpu_carry(carries[cpu]))
            if len(s):
                g.submit_image(cpu, s)
        for rep in range(2):
            g.run()
            g.sync()
            b = g.buckets()
            for cpu, s in enumerate(streams):
                exp = port.decode_stream(params, s, carries[cpu], want_records=False, want_rows=False)
                assert g.counters(cpu) == exp["counters"], (rep, cpu)
                assert np.array_equal(b["acc"][cpu], exp["buckets"]), f"run {rep} cpu {cpu}: sums differ"
                sc = port.scale_buckets(params, exp["buckets"])
                assert np.array_equal(b["scaled"][cpu], sc), f"run {rep} cpu {cpu}: scaled differ"
                num, den = sc[:, 1], sc[:, 0]
                val, ok = port.rate(num, den)
                assert np.array_equal(b["valid"][cpu] & 1, ok.astype(np.uint32)), f"run {rep} cpu {cpu}: validity"
                assert np.array_equal(b["metric"][cpu][ok, 0].view(np.uint64), val[ok].view(np.uint64))
        st = g.stats()
        if expect_exact is not None:
            assert st["exact_runs"] == expect_exact, st


def _samples(ts_list, n=4, v0=0):
    out, v = [], v0
    for i, ts in enumerate(ts_list):
        v += 10 + i
        out.append(rg.sample(ts, ts - T0, (ts - T0) // 2, [v * (e + 1) for e in range(n)]))
    return rg.to_u8(b"".join(out))


def test_empty_streams_sparse_samples_and_range_edges(port_oracle):
    n, nb, bns = 4, 400, 1_000_000
    valid = po.make_carry(valid=True, tstamp=T0 + 7 * bns + 5, te=3, tr=2, last=[1, 2, 3, 4])
    invalid = po.make_carry()
    streams = [
        np.zeros(0, np.uint8),                                                   # no image at all
        rg.to_u8(rg.throttle(1) + rg.unthrottle(2) + rg.lost(3)),                # records, but no SAMPLE
        _samples([T0 + 50 * bns * k + 17 for k in range(1, 7)]),                 # 50 empty buckets between samples
        _samples([T0 - 5000, T0 - 10, T0 + 3, T0 + bns + 1]),                    # starts in front of the range
        _samples([T0 + (nb - 2) * bns, T0 + (nb - 1) * bns + 5, T0 + nb * bns + 1, T0 + (nb + 9) * bns]),  # runs out of it
        _samples([T0 + 9 * bns + k * 1000 for k in range(600)]),                 # valid carry-in in bucket 7
        _samples([T0 - 100 - k for k in range(3, 0, -1)]),                        # everything in front of the range
        _samples([T0 + nb * bns + 10 + k for k in range(5)]),                    # everything behind it
    ]
    carries = [invalid, invalid, invalid, invalid, invalid, valid, invalid, invalid]
    _check(port_oracle, streams, carries, n, nb, bns, expect_exact=0)


@pytest.mark.parametrize("bns", [1_000, 400_000_000, 2_000_000_000_000])
def test_bucket_sizes_from_one_sample_to_the_whole_stream(port_oracle, bns):
    """1 us buckets: every sample its own bucket with ~100 empty ones in between; 0.4 s: a bucket
    spans many chunks; 2000 s: the whole stream is one bucket."""
    n = 8
    streams, carries = [], []
    for cpu in range(3):
        d, info = rg.random_stream(900 + cpu, 6000, n, p_lost=0.004 * cpu, p_throttle=0.01 * cpu, mux=(cpu == 2))
        streams.append(d)
        carries.append(po.make_carry(valid=(cpu != 1), tstamp=T0, te=0, tr=0, last=[0] * n))
    nb = int(650_000_000 // bns) + 3
    _check(port_oracle, streams, carries, n, nb, bns, expect_exact=0)


def test_time_going_back_behind_a_lost(port_oracle):
    """After a LOST the generator re-bases without comparing times (gen :135-154), so a later
    bucket may come first: its cells get rows twice, far apart in the stream."""
    n, nb, bns = 4, 64, 1_000_000
    a = [T0 + k * 100_000 for k in range(1, 300)]        # buckets 0..29
    b = [T0 + 5 * bns + k * 100_000 for k in range(200)]  # back to bucket 5 behind a LOST
    v = 0
    out = []
    for ts in a:
        v += 7
        out.append(rg.sample(ts, ts - T0, ts - T0, [v, 2 * v, 3 * v, 4 * v]))
    out.append(rg.lost(4))
    for ts in b:
        v += 9
        out.append(rg.sample(ts, ts - T0 + 10**9, ts - T0 + 10**9, [v, 2 * v, 3 * v, 4 * v]))
    s = rg.to_u8(b"".join(out))
    _check(port_oracle, [s], [po.make_carry(valid=True, tstamp=T0, last=[0] * n)], n, nb, bns, expect_exact=0)


def test_wide_groups(port_oracle):
    """50 events: a lane serves two fields of the accumulator record."""
    n, nb, bns = 50, 80, 1_000_000
    d, _ = rg.random_stream(31, 700, n, p_lost=0.01)
    _check(port_oracle, [d], [po.make_carry(valid=True, tstamp=T0, last=[0] * n)], n, nb, bns, expect_exact=0)


@pytest.mark.parametrize("flags", [0, hbtgpu.F_FORCE_GENERAL, hbtgpu.F_FORCE_EXACT])
def test_mixed_records_all_paths(port_oracle, flags):
    n, nb, bns = 4, 700, 1_000_000
    streams, carries = [], []
    for cpu, kw in enumerate([dict(p_lost=0.02, p_throttle=0.05), dict(p_other=0.2, p_read=0.05),
                              dict(p_lost=0.3), dict(forge=0.02, p_lost=0.01), dict(mux=True, p_throttle=0.3)]):
        d, _ = rg.random_stream(5000 + cpu, 5000, n, **kw)
        streams.append(d)
        carries.append(po.make_carry(valid=(cpu % 2 == 0), tstamp=T0 - 9, te=4, tr=3, last=[5, 6, 7, 8]))
    _check(port_oracle, streams, carries, n, nb, bns)
