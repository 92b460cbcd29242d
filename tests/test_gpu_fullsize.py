"""-m gpu: BASELINE.json configurations at their full sizes.  The C oracle is fast enough
(all host cores) to check them completely: every bucket sum, counter and carry bit-exact, plus
size-independent properties (telescoping identity: the sums over all buckets equal last - first
counter values; n_rows adds up; decode of window shards equals decode of the whole)."""
import numpy as np
import pytest

from oracle import pyoracle as po

pytestmark = pytest.mark.gpu
hbtgpu = pytest.importorskip("dynolog_b200.hbtgpu")
from dynolog_b200 import synth  # noqa: E402


def _run_gpu(streams, n, nb, t0, bns, flags=0, metrics=()):
    total = sum(len(s) for s in streams)
    g = hbtgpu.Context(n_events=n, max_cpus=len(streams), n_buckets=nb, t0=t0, bucket_ns=bns,
                       arena_bytes=total + 256 * len(streams) + (1 << 20), flags=flags, metrics=metrics)
    carry = hbtgpu.make_carry(valid=True, tstamp=t0, last=[0] * n)
    for cpu, s in enumerate(streams):
        g.set_carry(cpu, carry)
        g.submit_image(cpu, s)
    g.run()
    g.sync()
    return g


def _check_against_oracle(g, port, streams, n, nb, t0, bns, threads=16):
    p = po.make_params(n, n_buckets=nb, t0=t0, bucket_ns=bns)
    carries = [po.make_carry(valid=True, tstamp=t0, last=[0] * n) for _ in streams]
    _, acc, ctrs = port.run_capture(p, streams, carries, n_threads=threads)
    b = g.buckets(want_metric=False, want_valid=False)
    assert np.array_equal(b["acc"], acc), "bucket sums differ from the oracle"
    for cpu in range(len(streams)):
        assert g.counters(cpu) == ctrs[cpu], f"cpu {cpu}: counters differ"
    # scaled counts: check a slice of cpus completely
    for cpu in range(0, len(streams), max(1, len(streams) // 8)):
        assert np.array_equal(b["scaled"][cpu], port.scale_buckets(p, acc[cpu]))
    return acc, ctrs


def test_cfg1_4mib_ring_with_forged_headers(port_oracle):
    for forge in (0, 10_000):
        spec, streams = synth.capture(1, 1, max_bytes_per_cpu=4 << 20, forge_ppm=forge)
        g = _run_gpu(streams, 4, 6000, synth.T_START, 1_000_000)
        try:
            _check_against_oracle(g, port_oracle, streams, 4, 6000, synth.T_START, 1_000_000, threads=1)
            p = po.make_params(4, n_buckets=6000, t0=synth.T_START, bucket_ns=1_000_000)
            exp = port_oracle.decode_stream(p, streams[0], po.make_carry(valid=True, tstamp=synth.T_START,
                                                                        last=[0] * 4))
            for a, e in zip(g.records(0), exp["records"]):
                assert np.array_equal(a, e)  # every record boundary
            assert np.array_equal(g.rows(0), exp["rows"])
        finally:
            g.close()


def test_cfg2_192cpu_full_capture(port_oracle):
    n, nb, t0, bns = 8, 10_100, synth.T_START, 1_000_000
    spec, streams = synth.capture(2, 192, n_samples=100_000)
    g = _run_gpu(streams, n, nb, t0, bns, metrics=[(hbtgpu.METRIC_RATE, [0, 1])])
    try:
        assert g.stats()["exact_runs"] == 0
        acc, ctrs = _check_against_oracle(g, port_oracle, streams, n, nb, t0, bns)
        # telescoping identity per cpu: sum over buckets == last sample's counters (carry started at 0)
        for cpu in (0, 95, 191):
            co = g.carry(cpu)
            tot = acc[cpu].sum(axis=0, dtype=np.uint64)
            oor = ctrs[cpu]["rows_out_of_range"]
            assert ctrs[cpu]["num_rows"] == 100_000 and int(tot[n + 2]) + oor == 100_000
            if oor == 0:
                assert [int(x) for x in tot[:n]] == [int(co.last[i]) for i in range(n)]
                assert int(tot[n]) == int(co.te)
    finally:
        g.close()


def test_cfg3_multiplexed_lost_throttle_4gib(port_oracle):
    n, bns, t0 = 8, 1_000_000, synth.T_START
    per_cpu = 64 << 20
    spec, streams = synth.capture(3, 64, max_bytes_per_cpu=per_cpu)
    nb = 70_000
    g = _run_gpu(streams, n, nb, t0, bns)
    try:
        assert g.stats()["exact_runs"] == 0, g.stats()
        acc, ctrs = _check_against_oracle(g, port_oracle, streams, n, nb, t0, bns)
        assert sum(c["lost"] for c in ctrs) > 10_000 and sum(c["throttle"] for c in ctrs) > 1_000
        assert all(c["err"] == 0 for c in ctrs)
    finally:
        g.close()


def test_forged_counters_keep_speculation(port_oracle):
    """Counters whose values carry a plausible record size in their top bits form false chains in
    lock step with the true one.  Results must be exact, and the speculative pass should still
    hold (it prefers chains made of records this capture can contain)."""
    n, bns, t0 = 8, 1_000_000, synth.T_START
    spec, streams = synth.capture(3, 8, max_bytes_per_cpu=8 << 20, forge_ppm=10_000)
    g = _run_gpu(streams, n, 9000, t0, bns)
    try:
        _check_against_oracle(g, port_oracle, streams, n, 9000, t0, bns, threads=8)
        assert g.stats()["exact_runs"] == 0, g.stats()
    finally:
        g.close()
    g = _run_gpu(streams, n, 9000, t0, bns, flags=hbtgpu.F_FORCE_GENERAL)
    try:
        _check_against_oracle(g, port_oracle, streams, n, 9000, t0, bns, threads=8)
        assert g.stats()["exact_runs"] == 0, g.stats()
    finally:
        g.close()


def test_cfg5_windows_sharded_equals_whole(port_oracle):
    """Windows contiguous in time per cpu; 2 emulated ranks x 3 windows of 16 MiB x 12 cpus."""
    from dynolog_b200 import shard
    n, bns, t0, ncpu, W = 8, 1_000_000, synth.T_START, 12, 6
    cap = synth.Capture(synth.make_spec(5), ncpu)
    windows = [[s.copy() for s in cap.window(max_bytes_per_cpu=16 << 20)] for _ in range(W)]
    whole = [np.concatenate([windows[w][c] for w in range(W)]) for c in range(ncpu)]
    nb = 110_000
    p = po.make_params(n, n_buckets=nb, t0=t0, bucket_ns=bns)
    carries = [po.make_carry(valid=True, tstamp=t0, last=[0] * n) for _ in range(ncpu)]
    _, acc_ref, ctr_ref = port_oracle.run_capture(p, whole, carries, n_threads=12)
    initial = (hbtgpu.Carry * ncpu)()
    for c in range(ncpu):
        initial[c] = hbtgpu.make_carry(valid=True, tstamp=t0, last=[0] * n)
    ctxs = []
    try:
        for r in range(2):
            ws = windows[3 * r:3 * r + 3]
            total = sum(len(s) for w in ws for s in w)
            g = hbtgpu.Context(n_events=n, max_cpus=ncpu, n_buckets=nb, t0=t0, bucket_ns=bns,
                               arena_bytes=total + (1 << 20), flags=hbtgpu.F_EXTERNAL_CARRY)
            for w in ws:  # several images per cpu, submitted window by window
                for c in range(ncpu):
                    g.submit_image(c, w[c])
            g.run()
            g.sync()
            ctxs.append(g)
        stubs = [g.get_stubs() for g in ctxs]
        acc = np.zeros_like(acc_ref)
        rows = 0
        for r, g in enumerate(ctxs):
            g.apply_boundaries(shard.fold_prev_tails(stubs, initial, r))
            g.sync()
            acc += g.buckets(want_scaled=False, want_metric=False, want_valid=False)["acc"]
            rows += sum(g.counters(c)["num_rows"] for c in range(ncpu))
    finally:
        for g in ctxs:
            g.close()
    assert np.array_equal(acc, acc_ref)
    assert rows == sum(c["num_rows"] for c in ctr_ref)


def test_cfg4_ip_tid_attribution(port_oracle):
    """BASELINE config 4 shape (scaled to 32 cpus x 16 MiB = 5.6 M samples): 4096 (pid,tid)
    Zipf(1.1), 2^20 IPs Zipf(1.2), top-100.  Spec extension: oracle = C restatement + numpy."""
    n, ncpu = 4, 32
    spec, streams = synth.capture(4, ncpu, max_bytes_per_cpu=16 << 20)
    t0, nb = synth.T_START, 20_000
    total = sum(len(s) for s in streams)
    p = po.make_params(n, n_buckets=nb, t0=t0, bucket_ns=1_000_000, pre=2)
    carries = [po.make_carry(valid=True, tstamp=t0, last=[0] * n) for _ in streams]
    with hbtgpu.Context(n_events=n, max_cpus=ncpu, n_buckets=nb, t0=t0, bucket_ns=1_000_000,
                        arena_bytes=total + (1 << 20), sample_format=hbtgpu.SAMPLE_IP_TID_TIME_READ) as g:
        c = hbtgpu.make_carry(valid=True, tstamp=t0, last=[0] * n)
        for cpu in range(ncpu):
            g.set_carry(cpu, c)
        g.submit_images(list(range(ncpu)), streams)
        g.run()
        g.sync()
        assert g.stats()["exact_runs"] == 0
        hist = g.tid_histogram()
        top = g.top_ips(100)
        acc = g.buckets(want_scaled=False, want_metric=False, want_valid=False)["acc"]
        nsamp = sum(g.counters(i)["sample"] for i in range(ncpu))
    tids, ips = [], []
    for s, cin in zip(streams, carries):
        a, b = port_oracle.collect_ip_tid(p, s, cin)
        tids.append(a)
        ips.append(b)
    tids, ips = np.concatenate(tids), np.concatenate(ips)
    assert nsamp == len(tids)
    tk, tc = np.unique(tids, return_counts=True)
    assert hist == {(int(k & 0xFFFFFFFF), int(k >> 32)): int(c) for k, c in zip(tk, tc)}
    ik, ic = np.unique(ips, return_counts=True)
    order = np.lexsort((ik, -ic.astype(np.int64)))[:100]
    assert top == [(int(ik[i]), int(ic[i])) for i in order]
    _, acc_ref, _ = port_oracle.run_capture(p, streams, carries, n_threads=16)
    assert np.array_equal(acc, acc_ref)


def test_cfg2_slice_against_the_compiled_reference(ref_oracle):
    """Closes the chain GPU == restatement == reference on the box itself: 24 of cfg2's 192 CPUs
    (2.4 M records) against oracle/_ref, the unmodified hbt code (real consume() + row drain;
    its harness sums delta[] and n_rows per bucket -- the rows carry no time_enabled/running)."""
    n, nb, t0, bns, ncpu = 8, 10_100, synth.T_START, 1_000_000, 24
    spec, streams = synth.capture(2, ncpu, n_samples=100_000)
    g = _run_gpu(streams, n, nb, t0, bns)
    try:
        p = po.make_params(n, n_buckets=nb, t0=t0, bucket_ns=bns)
        carries = [po.make_carry(valid=True, tstamp=t0, last=[0] * n) for _ in streams]
        _, acc, ctrs = ref_oracle.run_capture(p, streams, carries, n_threads=8, batch=4096)
        b = g.buckets(want_scaled=False, want_metric=False, want_valid=False)["acc"]
        cols = list(range(n)) + [n + 2]
        assert np.array_equal(b[:, :, cols], acc[:, :, cols]), "bucket sums differ from the compiled reference"
        for cpu in range(ncpu):
            got = g.counters(cpu)
            for k in ("sample", "lost", "throttle", "unthrottle", "num_rows"):
                assert got[k] == ctrs[cpu][k], (cpu, k)
        # one cpu completely (te / tr sums, carry-out) through the reference's decode_stream
        exp = ref_oracle.decode_stream(p, streams[5], carries[5], want_records=False, want_rows=False)
        assert np.array_equal(b[5], exp["buckets"])
    finally:
        g.close()


def test_cfg4_attribution_accumulates_over_windows_100m_samples(port_oracle):
    """BASELINE config 4 streamed through one context in pieces (HBTGPU_F_ACCUMULATE_ATTR): 8 windows
    x 192 cpus x 6.9 MiB = 10.6 GB = 120 M samples; histogram and top-100 over ALL of them against
    the C restatement's walk (spec extension: parity unpinned by construction)."""
    n, ncpu, W, per_cpu = 4, 192, 8, 7_200_000
    cap = synth.Capture(synth.make_spec(4), ncpu)
    t0, nb = synth.T_START, 1000
    p = po.make_params(n, n_buckets=nb, t0=t0, bucket_ns=1_000_000, pre=2)
    buf = np.empty(ncpu * ((per_cpu + 127) // 128 * 128), np.uint8)
    tid_counts = np.zeros(4096, np.int64)
    ip_counts = np.zeros(1 << 20, np.int64)
    carries = [po.make_carry(valid=True, tstamp=t0, last=[0] * n) for _ in range(ncpu)]
    total = 0
    with hbtgpu.Context(n_events=n, max_cpus=ncpu, n_buckets=nb, t0=t0, bucket_ns=1_000_000,
                        arena_bytes=buf.size + (1 << 20), sample_format=hbtgpu.SAMPLE_IP_TID_TIME_READ,
                        flags=hbtgpu.F_ACCUMULATE_ATTR) as g:
        c = hbtgpu.make_carry(valid=True, tstamp=t0, last=[0] * n)
        for cpu in range(ncpu):
            g.set_carry(cpu, c)
        for w in range(W):
            streams = cap.window(max_bytes_per_cpu=per_cpu, out=buf)
            if w:
                g.reset(keep_carry=True)
            g.submit_images(list(range(ncpu)), streams)
            g.run()
            g.sync()
            assert g.stats()["exact_runs"] == 0
            total += sum(g.counters(i)["sample"] for i in range(ncpu))
            for i, s in enumerate(streams):
                tids, ips = port_oracle.collect_ip_tid(p, s, carries[i])
                tid_counts += np.bincount((tids >> np.uint64(32)).astype(np.int64) - 5000, minlength=4096)
                ip_counts += np.bincount(((ips - np.uint64(0xffffffff81000000)) >> np.uint64(4)).astype(np.int64),
                                         minlength=1 << 20)
        hist = g.tid_histogram()
        top = g.top_ips(100)
        g.attr_clear()
        g.sync()
        assert g.tid_histogram() == {}
    assert total > 100_000_000 and total == int(tid_counts.sum()) == int(ip_counts.sum())
    exp_hist = {(1000 + r // 8, 5000 + r): int(k) for r, k in enumerate(tid_counts) if k}
    assert hist == exp_hist
    order = np.lexsort((np.arange(1 << 20), -ip_counts))[:100]
    assert top == [(0xffffffff81000000 + 16 * int(r), int(ip_counts[r])) for r in order]
