"""-m gpu: the CUDA path (through the C ABI) against the oracle and the golden vectors.

Integer results (record boundaries, counters, rows, bucket sums, scaled counts)
must be bit-exact; doubles (ratios) are compared bit-exactly too, which is
tighter than the 1 ulp the north star allows.
"""
import errno
import os

import numpy as np
import pytest

import ringgen as rg
from cases import GOLDEN, golden_stream_names, load_golden_stream
from oracle import pyoracle as po

pytestmark = pytest.mark.gpu

hbtgpu = pytest.importorskip("dynolog_b200.hbtgpu")
from gpu_util import assert_matches_oracle, gpu_decode  # noqa: E402

MODES = {"fast": 0, "general": hbtgpu.F_FORCE_GENERAL, "exact": hbtgpu.F_FORCE_EXACT,
         "exact_general": hbtgpu.F_FORCE_EXACT | hbtgpu.F_FORCE_GENERAL}


@pytest.mark.parametrize("mode", sorted(MODES))
@pytest.mark.parametrize("name", golden_stream_names())
def test_golden_streams(name, mode):
    data, params, carry, exp = load_golden_stream(name)
    got, extra = gpu_decode(params, [data], [carry], flags=MODES[mode])
    assert_matches_oracle(got[0], exp, params.n_events, f"{name}/{mode}")


def test_golden_math():
    z = np.load(os.path.join(GOLDEN, "math.npz"))
    tp = [int(x) for x in z["tsc_params"]]
    with hbtgpu.Context(n_events=1, time_shift=tp[0], time_mult=tp[1], time_zero=tp[2]) as g:
        assert np.array_equal(g.get_count(z["cnt"], z["te"], z["tr"]), z["scaled"])
        rate, valid = g.rate(z["num"], z["den"])
        assert np.array_equal(valid, z["rate_valid"])
        assert np.array_equal(rate.view(np.uint64), z["rate"].view(np.uint64))
        assert np.array_equal(g.add_reduce(z["addv"]), z["add"])
        assert np.array_equal(g.tsc_to_ktime(z["tsc"]), z["kt"])
        assert np.array_equal(g.ktime_to_tsc(z["kt"]), z["back"])
        assert np.array_equal(g.per_second_mega(z["scaled"], z["trn"]).view(np.uint32),
                              z["psm"].view(np.uint32))
        ratio, ri, mux = g.running_ratio(z["te"], z["tr"])
        assert np.array_equal(ratio.view(np.uint32), z["ratio"].view(np.uint32))
        assert np.array_equal(ri, z["ratio_i"])
        assert np.array_equal(mux, z["mux"])
        # reference known answers (MonitorMockTest.cpp:86-88,121-126)
        assert list(g.get_count([200, 100], [4, 1], [2, 1])) == [400, 100]


BIG_CASES = [
    # (n_events, n_records, kwargs) -- sizes that span many tiles and chunks
    (4, 60_000, dict()),
    (8, 40_000, dict(p_lost=0.002, p_throttle=0.001, mux=True)),
    (8, 30_000, dict(p_lost=0.01, p_other=0.02, p_read=0.01, mux=True)),
    (4, 30_000, dict(forge=0.01, p_lost=0.001)),
    (1, 50_000, dict(p_lost=0.3)),
    (64, 3_000, dict(p_lost=0.01)),
    (20, 8_000, dict(p_lost=0.002, p_throttle=0.002)),   # rows wider than the emission stage
    (2, 20_000, dict(p_other=0.5)),
]


@pytest.mark.parametrize("mode", ["fast", "general", "exact"])
@pytest.mark.parametrize("case", range(len(BIG_CASES)))
def test_multi_tile_streams(port_oracle, case, mode):
    n, nrec, kw = BIG_CASES[case]
    streams, carries, exps = [], [], []
    t0 = 1_000_000_000
    nb = 0
    for cpu in range(3):
        data, info = rg.random_stream(7000 + 10 * case + cpu, nrec + 1000 * cpu, n, **kw)
        streams.append(data)
        carries.append(po.make_carry(valid=(cpu != 1), tstamp=t0 - 1, last=[0] * n))
        nb = max(nb, int((info["t_end"] - t0) // 1_000_000) + 1)
    params = po.make_params(n, n_buckets=max(1, nb - 2), t0=t0, bucket_ns=1_000_000,
                            tsc_mask=(0b10 if n >= 2 else 0), time_shift=5, time_mult=77,
                            time_zero=3)
    got, extra = gpu_decode(params, streams, carries, flags=MODES[mode])
    for cpu in range(3):
        exp = port_oracle.decode_stream(params, streams[cpu], carries[cpu])
        exp["scaled"] = port_oracle.scale_buckets(params, exp["buckets"])
        assert_matches_oracle(got[cpu], exp, n, f"case{case}/{mode}/cpu{cpu}")
    if mode == "fast" and "forge" not in kw:
        assert extra["stats"]["exact_runs"] == 0, \
            f"speculation should hold on well-formed streams: {extra['stats']}"


def test_multi_image_streams(port_oracle):
    """Several ring images per cpu (windows): decoded as one stream with carried state."""
    n = 4
    data, info = rg.random_stream(4242, 20_000, n, p_lost=0.002)
    # cut at record boundaries
    exp_full = port_oracle.decode_stream(po.make_params(n), data, po.make_carry())
    offs = exp_full["records"][0]
    cut_recs = [0, 5, 6, 3000, 3001, 11111, len(offs)]
    cuts = [int(offs[i]) if i < len(offs) else len(data) for i in cut_recs]
    sizes = [b - a for a, b in zip(cuts[:-1], cuts[1:])]
    t0 = 1_000_000_000
    params = po.make_params(n, n_buckets=int((info["t_end"] - t0) // 500_000) + 1, t0=t0,
                            bucket_ns=500_000)
    carry = po.make_carry(valid=True, tstamp=t0, last=[0] * n)
    exp = port_oracle.decode_stream(params, data, carry)
    exp["scaled"] = port_oracle.scale_buckets(params, exp["buckets"])
    for mode in ("fast", "exact"):
        got, _ = gpu_decode(params, [data], [carry], flags=MODES[mode], split=[sizes])
        assert_matches_oracle(got[0], exp, n, f"multi-image/{mode}")


def _err_streams():
    good = [rg.sample(1000 + 10 * i, 10 * i, 10 * i, [i, 2 * i]) for i in range(1, 400)]
    k = 250
    return {
        "bad_size": good[:k] + [rg.sample(9000, 1, 1, [1, 2], size=64)] + good[k:],
        "bad_nr": good[:k] + [rg.sample(9000, 1, 1, [1, 2, 3])] + good[k:],
        "regress": good[:k] + [rg.sample(5, 1, 1, [100, 200])] + good[k:],
        "regress_after_lost": good[:k] + [rg.lost(2), rg.sample(5, 1, 1, [100, 200])] + good[k:],
        "bad_read": good[:k] + [rg.read_rec(1, 2, 3, 4, [5, 6], size=48)] + good[k:],
        "zero_size": good[:k] + [rg.header(77, 0)] + good[k:],
        "misaligned": good[:k] + [rg.header(77, 12) + b"\0" * 8] + good[k:],
        "truncated": good[:k] + [rg.sample(9000, 1, 1, [1, 2])[:40]],
        "first_record_bad": [rg.header(9, 16) + b"\0" * 8] + good,
    }


@pytest.mark.parametrize("name", sorted(_err_streams()))
def test_error_paths(port_oracle, name):
    """The walk stops at the first record consume() would stall on (-EPERM); nothing after
    it is decoded; other cpus are unaffected."""
    data = rg.to_u8(b"".join(_err_streams()[name]))
    other, _ = rg.random_stream(99, 5000, 2)
    p = po.make_params(2, n_buckets=64, t0=1000, bucket_ns=100)
    carries = [po.make_carry(valid=True, tstamp=1000, last=[0, 0]), po.make_carry()]
    got, extra = gpu_decode(p, [data, other], carries)
    for cpu, s in enumerate([data, other]):
        exp = port_oracle.decode_stream(p, s, carries[cpu])
        exp["scaled"] = port_oracle.scale_buckets(p, exp["buckets"])
        assert_matches_oracle(got[cpu], exp, 2, f"{name}/cpu{cpu}")
    if name != "regress_after_lost":
        assert got[0]["counters"]["err"] == -errno.EPERM
        assert extra["stats"]["exact_runs"] == 1


def test_empty_and_tiny(port_oracle):
    p = po.make_params(3, n_buckets=4, t0=0, bucket_ns=1000)
    streams = [np.zeros(0, np.uint8), rg.to_u8(rg.lost(5)), rg.to_u8(rg.sample(10, 1, 1, [1, 2, 3])),
               rg.to_u8(rg.sample(10, 1, 1, [1, 2, 3]) + rg.sample(20, 2, 2, [2, 4, 6]))]
    carries = [po.make_carry(valid=True, tstamp=1, te=5, tr=5, last=[1, 1, 1]) for _ in streams]
    got, _ = gpu_decode(p, streams, carries)
    for cpu, s in enumerate(streams):
        exp = port_oracle.decode_stream(p, s, carries[cpu])
        exp["scaled"] = port_oracle.scale_buckets(p, exp["buckets"])
        assert_matches_oracle(got[cpu], exp, 3, f"tiny/cpu{cpu}")


def test_huge_unknown_records(port_oracle):
    """Records up to the u16 size limit (65528 B) span several 16 KiB tiles and whole chunks."""
    parts = []
    g = rg.SplitMix64(5)
    ts = 100
    for i in range(40):
        for _ in range(g.uniform(1, 30)):
            ts += 10
            parts.append(rg.sample(ts, ts, ts, [ts, 2 * ts]))
        words = g.uniform(1, 8190)
        parts.append(rg.other(40 + i, [g.next() for _ in range(words)]))
    data = rg.to_u8(b"".join(parts))
    p = po.make_params(2, n_buckets=16, t0=0, bucket_ns=1000)
    carry = po.make_carry()
    exp = port_oracle.decode_stream(p, data, carry)
    exp["scaled"] = port_oracle.scale_buckets(p, exp["buckets"])
    for mode in ("fast", "exact"):
        got, _ = gpu_decode(p, [data], [carry], flags=MODES[mode])
        assert_matches_oracle(got[0], exp, 2, f"huge/{mode}")


def test_metrics_and_validity(port_oracle):
    """IPC-style rate, add reducer and mips over bucket sums (A7-A9, A13)."""
    n = 4
    data, info = rg.random_stream(31337, 8000, n, mux=True, p_lost=0.001)
    t0 = 1_000_000_000
    nb = int((info["t_end"] - t0) // 2_000_000) + 3
    p = po.make_params(n, n_buckets=nb, t0=t0, bucket_ns=2_000_000)
    metrics = [(hbtgpu.METRIC_RATE, [1, 0]), (hbtgpu.METRIC_RATE, [2, 1]),
               (hbtgpu.METRIC_ADD, [0, 1, 2, 3]), (hbtgpu.METRIC_PER_SECOND_MEGA, [1])]
    carry = po.make_carry(valid=True, tstamp=t0, last=[0] * n)
    got, extra = gpu_decode(p, [data], [carry], metrics=metrics, want_tables=False)
    exp = port_oracle.decode_stream(p, data, carry)
    scaled = port_oracle.scale_buckets(p, exp["buckets"])
    assert np.array_equal(got[0]["scaled"], scaled)
    m, v = extra["metric"][0], extra["valid"][0]
    r0, v0 = port_oracle.rate(scaled[:, 1], scaled[:, 0])
    r1, v1 = port_oracle.rate(scaled[:, 2], scaled[:, 1])
    add = port_oracle.add_reduce(scaled)
    assert np.array_equal((v & 1).astype(bool), v0) and np.array_equal(((v >> 1) & 1).astype(bool), v1)
    assert np.array_equal(m[:, 0].view(np.uint64)[v0], r0.view(np.uint64)[v0])
    assert np.array_equal(m[:, 1].view(np.uint64)[v1], r1.view(np.uint64)[v1])
    assert np.array_equal(m[:, 2], add)
    tr = exp["buckets"][:, n + 1]
    ok = tr != 0
    psm = port_oracle.per_second_mega(scaled[:, 1][ok], tr[ok])
    assert np.array_equal(m[:, 3][ok].astype(np.float32).view(np.uint32), psm.view(np.uint32))
    assert 0 < v0.sum() <= nb


def test_rerun_and_reset(port_oracle):
    """run() twice gives the same result; reset(keep_carry) continues the stream."""
    n = 2
    data, info = rg.random_stream(77, 6000, n, p_lost=0.003)
    full = port_oracle.decode_stream(po.make_params(n), data, po.make_carry())
    cut = int(full["records"][0][3500])
    p = po.make_params(n, n_buckets=32, t0=1_000_000_000, bucket_ns=25_000_000)
    carry = po.make_carry(valid=True, tstamp=1_000_000_000, last=[0, 0])
    from gpu_util import to_gpu_carry
    with hbtgpu.Context(n_events=n, n_buckets=32, t0=p.t0, bucket_ns=p.bucket_ns,
                        arena_bytes=len(data) + 65536) as g:
        g.set_carry(0, to_gpu_carry(carry))
        g.submit_image(0, data[:cut])
        g.run()
        g.run()
        g.sync()
        a1 = g.buckets()["acc"][0].copy()
        e1 = port_oracle.decode_stream(p, data[:cut], carry)
        assert np.array_equal(a1, e1["buckets"])
        assert g.counters(0)["num_records"] == e1["counters"]["num_records"]
        g.reset(keep_carry=True)
        g.submit_image(0, data[cut:])
        g.run()
        g.sync()
        a2 = g.buckets()["acc"][0]
        e2 = port_oracle.decode_stream(p, data[cut:], e1["carry_out"])
        assert np.array_equal(a2, e2["buckets"])
        whole = port_oracle.decode_stream(p, data, carry)
        assert np.array_equal(a1 + a2, whole["buckets"])


@pytest.mark.parametrize("mode", ["fast", "general", "exact"])
def test_ip_tid_sample_layout(port_oracle, mode):
    """BASELINE config 4 layout (IP|TID|TIME|READ; spec extension, checked against the C
    restatement only): the ordinary outputs must come out exactly as for the hbt layout."""
    n = 4
    streams, carries = [], []
    for cpu, kw in enumerate([dict(), dict(p_lost=0.01, p_throttle=0.01, mux=True), dict(p_other=0.05, p_read=0.02)]):
        data, info = rg.random_stream(31_000 + cpu, 25_000, n, ip_tid=True, **kw)
        streams.append(data)
        carries.append(po.make_carry(valid=(cpu != 2), tstamp=999_999_999, last=[0] * n))
    p = po.make_params(n, n_buckets=2600, t0=1_000_000_000, bucket_ns=1_000_000, pre=2, tsc_mask=0b100,
                       time_shift=3, time_mult=11, time_zero=7)
    got, extra = gpu_decode(p, streams, carries, flags=MODES[mode])
    for cpu in range(3):
        exp = port_oracle.decode_stream(p, streams[cpu], carries[cpu])
        exp["scaled"] = port_oracle.scale_buckets(p, exp["buckets"])
        assert_matches_oracle(got[cpu], exp, n, f"iptid/{mode}/cpu{cpu}")
    if mode == "fast":
        assert extra["stats"]["exact_runs"] == 0


def _attribution_oracle(port_oracle, p, streams, carries):
    """(pid,tid) histogram and IP counts from the C restatement's walk (spec extension: the
    counting itself is restated here with numpy)."""
    tids, ips = [], []
    for s, c in zip(streams, carries):
        a, b = port_oracle.collect_ip_tid(p, s, c)
        tids.append(a)
        ips.append(b)
    tids, ips = np.concatenate(tids), np.concatenate(ips)
    tk, tc = np.unique(tids, return_counts=True)
    hist = {(int(k & 0xFFFFFFFF), int(k >> 32)): int(c) for k, c in zip(tk, tc)}
    ik, ic = np.unique(ips, return_counts=True)
    order = np.lexsort((ik, -ic.astype(np.int64)))  # count desc, ip asc
    return hist, [(int(ik[i]), int(ic[i])) for i in order]


@pytest.mark.parametrize("mode", ["fast", "general", "exact"])
def test_tid_histogram_and_top_ips(port_oracle, mode):
    """BASELINE config 4: per-(pid,tid) histogram + top-K sampled IPs (K=100, ties by ip)."""
    n = 4
    streams, carries = [], []
    for cpu, kw in enumerate([dict(), dict(p_lost=0.01, p_throttle=0.01), dict(p_other=0.05)]):
        data, _ = rg.random_stream(41_000 + cpu, 30_000, n, ip_tid=True, **kw)
        streams.append(data)
        carries.append(po.make_carry())
    # a stream with an error in the middle: nothing after it may be counted
    good = [rg.sample(1000 + 10 * i, i, i, [i] * n, ip_tid=(0x1000 + (i % 5), 1, 2 + (i % 3))) for i in range(1, 600)]
    bad = good[:400] + [rg.sample(9, 1, 1, [0] * n, ip_tid=(0x9999, 7, 7))] + good[400:]
    streams.append(rg.to_u8(b"".join(bad)))
    carries.append(po.make_carry())
    p = po.make_params(n, n_buckets=64, t0=1_000_000_000, bucket_ns=50_000_000, pre=2)
    total = sum(len(s) for s in streams)
    with hbtgpu.Context(n_events=n, max_cpus=len(streams), n_buckets=64, t0=p.t0, bucket_ns=p.bucket_ns,
                        arena_bytes=total + (1 << 20), flags=MODES[mode],
                        sample_format=hbtgpu.SAMPLE_IP_TID_TIME_READ, tid_table_log2=10,
                        ip_table_log2=12) as g:
        g.submit_images(list(range(len(streams))), streams)
        g.run()
        g.sync()
        hist = g.tid_histogram()
        top = g.top_ips(100)
        top_all = g.top_ips(5000)
        assert g.counters(3)["err"] == -errno.EPERM and g.counters(3)["sample"] == 400
    exp_hist, exp_ips = _attribution_oracle(port_oracle, p, streams, carries)
    assert hist == exp_hist
    assert top == exp_ips[:100]
    assert top_all == exp_ips            # fewer distinct IPs than asked for: all of them, ordered
    assert sum(hist.values()) == sum(c for _, c in exp_ips)


def test_attribution_table_overflow_is_reported():
    n = 2
    data, _ = rg.random_stream(5, 4000, n, ip_tid=True)
    with hbtgpu.Context(n_events=n, n_buckets=4, t0=0, bucket_ns=1 << 40, arena_bytes=len(data) + 65536,
                        sample_format=hbtgpu.SAMPLE_IP_TID_TIME_READ, tid_table_log2=4, ip_table_log2=4) as g:
        g.submit_image(0, data)
        g.run()
        g.sync()
        with pytest.raises(hbtgpu.HbtGpuError) as e:
            g.top_ips(10)
        assert e.value.code == -errno.ENOSPC


def test_512_cpus_ragged_streams(port_oracle):
    """kMaxCpus streams (hbt/src/common/System.h:204) of ragged lengths, some empty, some with
    several images, all in one context."""
    n = 3
    ncpu = 512
    streams, splits = [], []
    for cpu in range(ncpu):
        nrec = [0, 1, 7, 40, 333, 1500][cpu % 6]
        data, _ = rg.random_stream(20_000 + cpu, nrec, n, p_lost=0.02 if cpu % 3 == 0 else 0.0,
                                   p_throttle=0.01 if cpu % 5 == 0 else 0.0)
        streams.append(data)
        if cpu % 7 == 0 and nrec >= 40:
            offs = port_oracle.decode_stream(po.make_params(n), data, po.make_carry())["records"][0]
            c1 = int(offs[len(offs) // 3])
            splits.append([c1, len(data) - c1])
        else:
            splits.append([len(data)])
    p = po.make_params(n, n_buckets=40, t0=1_000_000_000, bucket_ns=5_000_000)
    carries = [po.make_carry(valid=(cpu % 2 == 0), tstamp=999_999_999, last=[1, 2, 3]) for cpu in range(ncpu)]
    got, extra = gpu_decode(p, streams, carries, split=splits, want_tables=False)
    assert extra["stats"]["exact_runs"] == 0
    for cpu in range(ncpu):
        exp = port_oracle.decode_stream(p, streams[cpu], carries[cpu])
        exp["scaled"] = port_oracle.scale_buckets(p, exp["buckets"])
        assert_matches_oracle(got[cpu], exp, n, f"cpu{cpu}", tables=False)


def test_submit_images_batch(port_oracle):
    """hbtgpu_submit_images == the same images submitted one by one."""
    n = 2
    streams = [rg.random_stream(600 + i, 3000 + 500 * i, n, p_lost=0.01)[0] for i in range(4)]
    p = po.make_params(n, n_buckets=64, t0=1_000_000_000, bucket_ns=10_000_000)
    total = sum(len(s) for s in streams)
    with hbtgpu.Context(n_events=n, max_cpus=3, n_buckets=64, t0=p.t0, bucket_ns=p.bucket_ns,
                        arena_bytes=total + 65536) as g:
        # cpu 1 gets two images (one stream), cpu 0 and 2 one each
        g.submit_images([1, 0, 1, 2], [streams[0], streams[1], streams[2], streams[3]])
        g.run()
        g.sync()
        acc = g.buckets()["acc"]
        exp1 = port_oracle.decode_stream(p, np.concatenate([streams[0], streams[2]]), po.make_carry())
        assert np.array_equal(acc[1], exp1["buckets"])
        assert np.array_equal(acc[0], port_oracle.decode_stream(p, streams[1], po.make_carry())["buckets"])
        assert g.counters(2) == port_oracle.decode_stream(p, streams[3], po.make_carry())["counters"]
        with pytest.raises(hbtgpu.HbtGpuError):
            g.submit_images([0, 9], [streams[0], streams[1]])  # bad cpu: nothing is submitted


def test_async_result_copy_matches_sync(port_oracle):
    """hbtgpu_get_buckets_host_async: buffers complete after sync; a following run waits for it."""
    import ctypes as C
    n = 4
    a, _ = rg.random_stream(810, 20000, n, mux=True)
    b, _ = rg.random_stream(811, 15000, n, p_lost=0.01)
    nb = 2100
    with hbtgpu.Context(n_events=n, max_cpus=1, n_buckets=nb, t0=1_000_000_000, bucket_ns=1_000_000,
                        arena_bytes=len(a) + len(b) + 65536, metrics=[(hbtgpu.METRIC_RATE, [1, 0])]) as g:
        outs = []
        for data in (a, b):
            g.reset()
            g.submit_image(0, data)
            g.run()
            acc = np.zeros((1, nb, n + 3), np.uint64)
            scaled = np.zeros((1, nb, n), np.uint64)
            metric = np.zeros((1, nb, 1), np.float64)
            valid = np.zeros((1, nb), np.uint32)
            g._ck(g.L.hbtgpu_get_buckets_host_async(g.h, C.c_void_p(acc.ctypes.data),
                                                    C.c_void_p(scaled.ctypes.data),
                                                    C.c_void_p(metric.ctypes.data),
                                                    C.c_void_p(valid.ctypes.data)))
            outs.append((acc, scaled, metric, valid))
        g.sync()
        p = po.make_params(n, n_buckets=nb, t0=1_000_000_000, bucket_ns=1_000_000)
        for data, (acc, scaled, metric, valid) in zip((a, b), outs):
            exp = port_oracle.decode_stream(p, data, po.make_carry())
            assert np.array_equal(acc[0], exp["buckets"])
            assert np.array_equal(scaled[0], port_oracle.scale_buckets(p, exp["buckets"]))
        s = g.buckets()
        assert np.array_equal(s["acc"], outs[1][0]) and np.array_equal(s["metric"], outs[1][2])


def test_argument_errors():
    with pytest.raises(hbtgpu.HbtGpuError) as e:
        hbtgpu.Context(n_events=0)
    assert e.value.code == -errno.EINVAL
    with pytest.raises(hbtgpu.HbtGpuError):
        hbtgpu.Context(n_events=65)
    with hbtgpu.Context(n_events=2, arena_bytes=1024) as g:
        with pytest.raises(hbtgpu.HbtGpuError) as e:
            g.submit_image(0, np.zeros(1 << 20, np.uint8))
        assert e.value.code == -errno.ENOSPC
        with pytest.raises(hbtgpu.HbtGpuError):
            g.submit_image(5, np.zeros(8, np.uint8))
        with pytest.raises(hbtgpu.HbtGpuError):
            g.submit_image(0, np.zeros(12, np.uint8))
        with pytest.raises(hbtgpu.HbtGpuError):
            g.counters(0)  # run() not called


@pytest.mark.parametrize("bucket_ns,t0,span", [
    (1_000_000, 1_000_000_000, 1),            # f64-assisted division (bucket_ns < 2^32, range < 2^53)
    (999_983, 7, 1),                          # prime width
    (3, 0, 1),                                # tiny buckets
    ((1 << 32) - 5, 12345, 1 << 20),          # widest 32-bit width
    (1 << 20, 1 << 40, 1),                    # power of two: shift
    ((1 << 32) + 7, 99, 1 << 21),             # width >= 2^32: 64-bit multiply-high
    (1_000_003, (1 << 52) + 11, 1),           # t0 so large that range_ns >= 2^53 is not the issue, ts is
])
def test_bucket_boundaries(port_oracle, bucket_ns, t0, span):
    """Samples placed on, one before and one after bucket boundaries land in the bucket the
    integer division says (A7: bucket = (ts - t0) / bucket_ns), for every division path."""
    n, nb = 2, 5000
    r = rg.SplitMix64(bucket_ns ^ t0)
    ts_list = []
    for _ in range(1500):
        b = r.uniform(0, nb + 20)             # some beyond the range
        ts_list += [t0 + b * bucket_ns + off for off in (-1, 0, 1) if t0 + b * bucket_ns + off > 0]
        ts_list.append(t0 + r.uniform(0, nb * bucket_ns))
    ts_list = sorted(set(ts_list))
    out = bytearray()
    v = [0, 0]
    for i, ts in enumerate(ts_list):
        v = [v[0] + 3, v[1] + (i % 7)]
        out += rg.sample(ts, 10 * i, 10 * i, v)
    data = rg.to_u8(out)
    p = po.make_params(n, n_buckets=nb, t0=t0, bucket_ns=bucket_ns)
    carries = [po.make_carry(valid=True, tstamp=ts_list[0] - 1 if ts_list[0] else 0, last=[0, 0])]
    got, _ = gpu_decode(p, [data], carries)
    exp = port_oracle.decode_stream(p, data, carries[0])
    exp["scaled"] = port_oracle.scale_buckets(p, exp["buckets"])
    assert_matches_oracle(got[0], exp, n, f"bucket_ns={bucket_ns}")
    assert exp["counters"]["rows_out_of_range"] > 0


@pytest.mark.parametrize("pct,period,duration,first,count", [
    (0.5, 1_000_000_000, 10_000_000_000, 0, None), (0.95, 60_000_000_000, 1_000_000_000, 0, None),
    (0.0, 1_000_000, 1_000_000, 7, 100), (1.0, 3_000_000, 7_000_000, 300, 1), (0.333, 1_500_000_000, 999_999_000, 1, 399)])
def test_series_reductions(port_oracle, pct, period, duration, first, count):
    """MetricSeries<uint64_t> avg / percentile / diff / rate of every (cpu, event) series of scaled
    counts, bit-exact (f64 included) against the C restatement of the reference class."""
    n, nb = 8, 400
    streams = [rg.random_stream(60 + i, 6000, n, mux=(i % 2 == 0), big_values=(i == 2))[0] for i in range(3)]
    with hbtgpu.Context(n_events=n, max_cpus=3, n_buckets=nb, t0=1_000_000_000, bucket_ns=1_500_000,
                        arena_bytes=sum(len(s) for s in streams) + (1 << 16)) as g:
        for cpu, s in enumerate(streams):
            g.submit_image(cpu, s)
        g.run()
        g.sync()
        scaled = g.buckets(want_acc=False, want_metric=False, want_valid=False)["scaled"]
        got = g.series_reduce(pct, period, duration, first_bucket=first, count=count)
        with pytest.raises(hbtgpu.HbtGpuError):
            g.series_reduce(1.5, period, duration)
    cnt = nb - first if count is None else count
    series = np.ascontiguousarray(scaled[:, first:first + cnt, :].transpose(0, 2, 1)).reshape(3 * n, cnt)
    assert series.any()
    exp = po.series_reduce("port", series, pct, period, duration)
    for k, v in exp.items():
        assert np.array_equal(got[k].reshape(-1).view(np.uint64), v.view(np.uint64)), (k, got[k].reshape(-1)[:4], v[:4])
