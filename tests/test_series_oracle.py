"""Series reductions (sink side): the C restatement pinned against the REAL
facebook::dynolog::MetricSeries<uint64_t> (oracle/_ref), plus the reference's own known answers
(dynolog/src/metric_frame/tests: a 1..N ramp).  CPU only."""
import numpy as np
import pytest

from oracle import pyoracle

needs_ref = pytest.mark.skipif(not pyoracle.available("reference"),
                               reason="oracle/_ref not built (needs /root/reference)")
KEYS = ("avg_f64", "avg_u64", "percentile", "diff", "rate_f64", "rate_u64")


def series_cases():
    r = np.random.default_rng(7)
    yield r.integers(0, 1 << 40, (5, 1000), dtype=np.uint64)
    yield np.cumsum(r.integers(0, 1 << 30, (7, 4097), dtype=np.uint64), axis=1)       # counters: monotone
    yield r.integers(0, 1 << 64, (3, 257), dtype=np.uint64)                           # sums wrap, f64 rounds
    yield np.full((2, 33), 12345, np.uint64)                                           # ties
    yield r.integers(0, 5, (4, 1), dtype=np.uint64)                                    # one sample


def test_known_answers():
    ramp = np.arange(1, 11, dtype=np.uint64)[None, :]          # 1..10
    o = pyoracle.series_reduce("port", ramp, 0.5, 1_000_000_000, 9_000_000_000)
    assert o["avg_f64"][0] == 5.5 and o["avg_u64"][0] == 5
    assert o["percentile"][0] == 6          # lround(0.5 * 9) = 5 (half away from zero) -> 6th smallest
    assert o["diff"][0] == 9
    assert o["rate_f64"][0] == 1.0 and o["rate_u64"][0] == 1   # 9 per 9 s, per 1 s
    o = pyoracle.series_reduce("port", ramp, 0.0, 10_000_000_000, 1_000_000_000)
    assert o["percentile"][0] == 1 and o["rate_u64"][0] == 90
    assert pyoracle.series_reduce("port", ramp, 1.0, 1000, 1000)["percentile"][0] == 10


@needs_ref
@pytest.mark.parametrize("pct,period,duration", [
    (0.5, 1_000_000_000, 10_000_000_000), (0.95, 60_000_000_000, 1_000_000_000),
    (0.0, 1_000_000, 1_000_000), (1.0, 3_000_000, 7_000_000), (0.333, 1_500_000_000, 999_999_000)])
def test_port_matches_reference(pct, period, duration):
    for series in series_cases():
        ref = pyoracle.series_reduce("reference", series, pct, period, duration)
        port = pyoracle.series_reduce("port", series, pct, period, duration)
        for k in KEYS:
            assert np.array_equal(ref[k].view(np.uint64), port[k].view(np.uint64)), (k, ref[k][:3], port[k][:3])
