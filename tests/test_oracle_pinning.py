"""Pin the plain-C oracle (oracle/hbt_oracle.c) to the reference.

(a) against golden vectors produced by the real hbt code (tests/golden/, made by
    tests/golden/make_golden.py from oracle/_ref/libhbtref.so);
(b) differentially against oracle/_ref/libhbtref.so itself, when present;
(c) against the reference's own known-answer tests:
    hbt/src/mon/tests/MonitorMockTest.cpp:86-88,121-126 ({200,te=4,tr=2} -> 400,
    {100,1,1} -> 100) and the identities of
    hbt/src/perf_event/tests/PerCpuGeneratorsTest.cpp:195-232 and
    hbt/src/perf_event/tests/ClocksTest.cpp:60-91.
"""
import errno
import os

import numpy as np
import pytest

import ringgen as rg
from cases import (GOLDEN, assert_decode_equal, golden_stream_names, load_golden_stream)
from oracle import pyoracle as po


@pytest.mark.parametrize("name", golden_stream_names())
def test_port_matches_golden_stream(port_oracle, name):
    data, params, carry, exp = load_golden_stream(name)
    got = port_oracle.decode_stream(params, data, carry)
    assert_decode_equal(got, exp, params.n_events, name)
    assert np.array_equal(port_oracle.scale_buckets(params, got["buckets"]), exp["scaled"])


def test_port_matches_golden_math(port_oracle):
    z = np.load(os.path.join(GOLDEN, "math.npz"))
    o = port_oracle
    assert np.array_equal(o.get_count(z["cnt"], z["te"], z["tr"]), z["scaled"])
    rate, valid = o.rate(z["num"], z["den"])
    assert np.array_equal(valid, z["rate_valid"])
    assert np.array_equal(rate.view(np.uint64), z["rate"].view(np.uint64))  # bit-exact f64
    assert np.array_equal(o.add_reduce(z["addv"]), z["add"])
    tp = [int(x) for x in z["tsc_params"]]
    p = po.make_params(1, time_shift=tp[0], time_mult=tp[1], time_zero=tp[2])
    assert np.array_equal(o.tsc_to_ktime(p, z["tsc"]), z["kt"])
    assert np.array_equal(o.ktime_to_tsc(p, z["kt"]), z["back"])
    assert np.array_equal(o.per_second_mega(z["scaled"], z["trn"]).view(np.uint32),
                          z["psm"].view(np.uint32))
    ratio, ri, mux = o.running_ratio(z["te"], z["tr"])
    assert np.array_equal(ratio.view(np.uint32), z["ratio"].view(np.uint32))
    assert np.array_equal(ri, z["ratio_i"])
    assert np.array_equal(mux, z["mux"])


def test_reference_known_answers(port_oracle):
    # MonitorMockTest.cpp:86-88,121-126 and the uncore twin :214-216,252-256
    assert list(port_oracle.get_count([200, 100], [4, 1], [2, 1])) == [400, 100]
    # SURVEY probe: {te=1000,tr=333,count=12345} -> 37072; te==0 or tr==0 -> 0 (:471-476)
    assert list(port_oracle.get_count([12345, 5, 5], [1000, 0, 9], [333, 9, 0])) == [37072, 0, 0]
    # PerCpuGeneratorsTest.cpp:229-231: rate == double(c0)/c1
    r, v = port_oracle.rate([300, 7], [200, 0])
    assert r[0] == 300.0 / 200.0 and v[0] and not v[1]
    # BuiltinMetrics.cpp:94 int accumulator: {3e9,3e9} -> 1.70503e+09
    assert port_oracle.add_reduce(np.array([[3_000_000_000, 3_000_000_000]], np.uint64))[0] == \
        1705032704.0
    # ClocksTest.cpp:60-91: tsc -> ktime -> tsc round trip within +-2
    p = po.make_params(1, time_shift=10, time_mult=341, time_zero=999)
    tsc = np.arange(1, 200000, 37, dtype=np.uint64) * np.uint64(1000003)
    back = port_oracle.ktime_to_tsc(p, port_oracle.tsc_to_ktime(p, tsc))
    assert np.all(np.abs(back.astype(np.int64) - tsc.astype(np.int64)) <= 4)


def _diff_accum_identity(o):
    # PerCpuGeneratorsTest.cpp:195-226: rv2.diff(rv1) then accum(rv1) gives rv2 back;
    # in stream form: sum of row deltas + carry == last sample's values.
    data, _ = rg.random_stream(11, 300, 4)
    p = po.make_params(4, n_buckets=1, t0=0, bucket_ns=1 << 62)
    r = o.decode_stream(p, data, po.make_carry(valid=True, tstamp=5, last=[0] * 4))
    tot = r["rows"][:, 2:].sum(axis=0)
    assert [int(x) for x in tot] == [int(r["carry_out"].last[i]) for i in range(4)]
    assert [int(x) for x in r["buckets"][0, :4]] == [int(x) for x in tot]
    assert int(r["buckets"][0, 4]) == int(r["carry_out"].te)
    assert int(r["buckets"][0, 6]) == r["counters"]["num_rows"]


def test_diff_accum_identity_port(port_oracle):
    _diff_accum_identity(port_oracle)


def test_diff_accum_identity_reference(ref_oracle):
    _diff_accum_identity(ref_oracle)


DIFF_CASES = [
    dict(n=4, nrec=2000, kw=dict()),
    dict(n=8, nrec=3000, kw=dict(p_lost=0.01, p_throttle=0.005, mux=True)),
    dict(n=1, nrec=1500, kw=dict(p_lost=0.05, p_other=0.05, p_read=0.05)),
    dict(n=4, nrec=2500, kw=dict(forge=0.1, p_lost=0.01)),
    dict(n=16, nrec=800, kw=dict(p_throttle=0.02, big_values=True)),
    dict(n=64, nrec=300, kw=dict(p_lost=0.02, mux=True)),
    dict(n=3, nrec=50, kw=dict(p_lost=0.9)),
    dict(n=2, nrec=0, kw=dict()),
]


@pytest.mark.parametrize("case", range(len(DIFF_CASES)))
@pytest.mark.parametrize("carry_valid", [False, True])
def test_port_vs_reference_differential(port_oracle, ref_oracle, case, carry_valid):
    c = DIFF_CASES[case]
    data, info = rg.random_stream(1000 + case, c["nrec"], c["n"], **c["kw"])
    t0 = 1_000_000_000 + 12_345
    nb = max(1, int((info["t_end"] - t0) // 250_000) - 3)
    tsc_mask = 0b101 if c["n"] >= 3 else 0
    p = po.make_params(c["n"], n_buckets=nb, t0=t0, bucket_ns=250_000, tsc_mask=tsc_mask,
                       time_shift=7, time_mult=97, time_zero=5)
    carry = po.make_carry(valid=carry_valid, tstamp=999_999_000, te=3, tr=2,
                          last=[0] * c["n"])
    a = port_oracle.decode_stream(p, data, carry)
    b = ref_oracle.decode_stream(p, data, carry)
    assert_decode_equal(a, b, c["n"], f"case {case}")
    assert np.array_equal(port_oracle.scale_buckets(p, a["buckets"]),
                          ref_oracle.scale_buckets(p, b["buckets"]))


def _err_streams():
    good = [rg.sample(1000 + 10 * i, 10 * i, 10 * i, [i, 2 * i]) for i in range(1, 8)]
    bad_size = good[:4] + [rg.sample(2000, 1, 1, [1, 2], size=64)] + good[4:]
    bad_nr_size = good[:3] + [rg.sample(2000, 1, 1, [1, 2], nr=3)] + good[3:]
    regress = good[:5] + [rg.sample(5, 1, 1, [100, 200])] + good[5:]
    regress_after_lost = good[:5] + [rg.lost(2), rg.sample(5, 1, 1, [100, 200])] + good[5:]
    bad_read = good[:2] + [rg.read_rec(1, 2, 3, 4, [5, 6], size=48)] + good[2:]
    return {"bad_size": bad_size, "bad_nr_size": bad_nr_size, "regress": regress,
            "regress_after_lost": regress_after_lost, "bad_read": bad_read}


@pytest.mark.parametrize("name", sorted(_err_streams()))
def test_error_paths_port_vs_reference(port_oracle, ref_oracle, name):
    """consume() stops before the bad record with -EPERM (PerfEventsGroup.h:1502-1507,
    1518-1523; PerCpuCountSampleGenerator.h:156-163); a regression right after a LOST
    is NOT an error (state was invalid)."""
    data = rg.to_u8(b"".join(_err_streams()[name]))
    p = po.make_params(2, n_buckets=8, t0=1000, bucket_ns=10)
    carry = po.make_carry(valid=True, tstamp=1000, last=[0, 0])
    a = port_oracle.decode_stream(p, data, carry)
    b = ref_oracle.decode_stream(p, data, carry)
    assert_decode_equal(a, b, 2, name)
    if name == "regress_after_lost":
        assert a["counters"]["err"] == 0
    else:
        assert a["counters"]["err"] == -errno.EPERM
        assert a["counters"]["err_offset"] < len(data)


def test_whole_capture_port_vs_reference(port_oracle, ref_oracle):
    streams = [rg.random_stream(50 + i, 400 + 10 * i, 4, p_lost=0.01)[0] for i in range(6)]
    p = po.make_params(4, n_buckets=64, t0=1_000_000_000, bucket_ns=1_000_000)
    carries = [po.make_carry(valid=True, tstamp=1_000_000_000, last=[0] * 4) for _ in streams]
    _, acc_p, ctr_p = port_oracle.run_capture(p, streams, carries, n_threads=3)
    _, acc_r, ctr_r = ref_oracle.run_capture(p, streams, carries, n_threads=3, batch=128)
    # the reference's rows carry no te/tr: compare raw sums and n_rows only
    assert np.array_equal(acc_p[:, :, :4], acc_r[:, :, :4])
    assert np.array_equal(acc_p[:, :, 6], acc_r[:, :, 6])
    for a, b in zip(ctr_p, ctr_r):
        for k in ("lost", "throttle", "unthrottle", "sample", "read", "num_rows",
                  "num_unknown_count_interval", "last_tstamp"):
            assert a[k] == b[k]
    for i, s in enumerate(streams):
        one = port_oracle.decode_stream(p, s, carries[i])
        assert np.array_equal(one["buckets"], acc_p[i])
