"""ContextSwitch-mode decode on the GPU against the C restatement, which
tests/test_cs_oracle.py pins to the real reference.  Through the C ABI (hbtgpu_cs_*)."""
import numpy as np
import pytest

import ringgen as rg
from dynolog_b200 import hbtgpu, hbtgpu_cs
from oracle import pyoracle

pytestmark = pytest.mark.gpu

CMP_COUNTERS = ("lost", "comm", "exit", "throttle", "unthrottle", "fork", "switch_cpu_wide", "num_records",
                "num_rows", "err", "err_offset", "last_tstamp")


def check(streams, tid_level=2, states=None, flags=0, expect_exact=None):
    want = pyoracle.cs_run("port", streams, tid_level=tid_level, states=states)
    got = hbtgpu_cs.decode(streams, tid_level=tid_level, states=states, flags=flags)
    for cpu in range(len(streams)):
        x, y = got["events"][cpu], want["events"][cpu]
        assert len(x) == len(y), f"cpu {cpu}: {len(x)} vs {len(y)} events"
        for f in ("tstamp", "type", "level", "comp_unit_id", "tag"):
            bad = np.nonzero(x[f] != y[f])[0]
            assert bad.size == 0, f"cpu {cpu} event {bad[:3]} {f}: {x[f][bad[:3]]} vs {y[f][bad[:3]]}"
        for k in CMP_COUNTERS:
            assert got["counters"][cpu][k] == want["counters"][cpu][k], (cpu, k, got["counters"][cpu], want["counters"][cpu])
    assert got["states"] == want["states"]
    assert got["threads"] == want["threads"]
    assert got["names"] == want["names"]
    if expect_exact is not None:
        assert (got["stats"]["exact_runs"] > 0) == expect_exact, got["stats"]
    return got


def test_known_answer():
    T = 5_000_001
    s = rg.to_u8(
        rg.cs_switch(1000, 1, T, out=True, preempt=True) + rg.cs_switch(1100, 1, T + 1)
        + rg.cs_lost(7, 1500) + rg.cs_switch(1600, 1, T + 1, out=True)
        + rg.cs_fork(T + 8, 1, T + 9, 1, 1700) + rg.cs_comm(T + 8, T + 9, b"abc", 1800)
        + rg.cs_throttle(1850) + rg.cs_exit(T + 8, 1, T + 9, 1, 1900) + rg.cs_switch(2000, 1, 0))
    got = check([s], tid_level=3)
    assert [(int(e["tstamp"]), int(e["type"]), int(e["tag"])) for e in got["events"][0]][:4] == [
        (1000, 5, 0xFFFFFFFF), (1100, 4, 1), (1101, 7, 0xFFFFFFFF), (1600, 8, 0xFFFFFFFF)]


@pytest.mark.parametrize("seed,kw", [
    (1, dict()),
    (2, dict(p_lost=0.01)),
    (3, dict(p_life=0.05, p_comm=0.03)),
    (4, dict(p_lost=0.005, p_life=0.05, p_comm=0.03, p_throttle=0.01, p_other=0.02)),
    (5, dict(p_lost=0.2, p_life=0.1)),
])
def test_random_streams(seed, kw):
    # 6000 records per cpu: several 16 KiB chunks per stream, so chunk entries are speculated
    streams = rg.random_cs_streams(seed, 4, 6000, n_tids=48, **kw)
    check(streams)


def test_speculation_holds_on_plain_streams():
    streams = rg.random_cs_streams(11, 3, 8000, p_life=0.01, p_comm=0.01)
    check(streams, expect_exact=False)


def test_forced_exact_path():
    streams = rg.random_cs_streams(12, 3, 5000, p_lost=0.01, p_life=0.05, p_comm=0.02, p_other=0.02)
    check(streams, flags=hbtgpu.F_FORCE_EXACT, expect_exact=True)


def test_state_carry_and_empty_stream():
    streams = rg.random_cs_streams(9, 3, 3000, p_lost=0.02, p_life=0.05)
    streams[1] = streams[1][:0]
    check(streams, states=[(12345, 1), (0, 0), (777, 0)])


def test_windows_continue_across_runs():
    """Two windows decoded one after the other with the state carried equal one long stream
    (the thread dictionary is fed both windows in order)."""
    streams = rg.random_cs_streams(21, 2, 4000, p_lost=0.01, p_life=0.03, p_comm=0.02)
    want = pyoracle.cs_run("port", streams)
    ctx = hbtgpu_cs.CsContext(2, sum(len(s) for s in streams) + (1 << 16), tid_level=1)
    cuts = [(len(s) // 2) // 8 * 8 for s in streams]
    # cut on a record boundary: walk the stream
    for i, s in enumerate(streams):
        pos = 0
        while pos < cuts[i]:
            pos += int(s[pos + 6]) | int(s[pos + 7]) << 8
        cuts[i] = pos
    # the reference visits cpu 0's whole stream before cpu 1's, so feed window by window per cpu
    th = hbtgpu_cs.Threads()
    evs = [[], []]
    parts = [[s[:c], s[c:]] for s, c in zip(streams, cuts)]
    states = [(0, 0), (0, 0)]
    per_cpu = [[None, None], [None, None]]
    for w in range(2):
        ctx.reset()
        for cpu in range(2):
            ctx.set_state(cpu, *states[cpu])
            ctx.submit(cpu, parts[cpu][w])
        ctx.run()
        ctx.sync()
        for cpu in range(2):
            ev, pid = ctx.events(cpu)
            per_cpu[cpu][w] = (ev.copy(), pid.copy(), ctx.ops(cpu).copy())
            states[cpu] = ctx.state(cpu)
    for cpu in range(2):
        for w in range(2):
            ev, pid, ops = per_cpu[cpu][w]
            th.assign(ev, pid, ops)
            evs[cpu].append(ev)
    for cpu in range(2):
        x = np.concatenate(evs[cpu])
        y = want["events"][cpu]
        assert len(x) == len(y)
        for f in ("tstamp", "type", "level", "comp_unit_id", "tag"):
            assert np.array_equal(x[f], y[f]), (cpu, f)
    assert [tuple(s) for s in states] == want["states"]
    threads, names = th.table()
    assert threads == want["threads"] and names == want["names"]
    ctx.close()


def test_fences_report_first_bad_record():
    good = rg.random_cs_streams(31, 1, 3000)[0]
    for bad in (rg.header(rg.SWITCH_CPU_WIDE, 0), rg.header(rg.SWITCH_CPU_WIDE, 28) + b"\0" * 24,
                rg.header(rg.LOST, 32) + b"\0" * 24, rg.header(rg.FORK, 64) + b"\0" * 8):
        s = np.concatenate([good, rg.to_u8(bad), rg.to_u8(rg.cs_switch(9_000_000_000, 1, 0))[:0]])
        got = check([s], expect_exact=True)
        assert got["counters"][0]["err"] == -1 and got["counters"][0]["err_offset"] == len(good)


def test_many_small_images_per_cpu():
    streams = rg.random_cs_streams(41, 2, 3000, p_lost=0.01, p_life=0.02)
    want = pyoracle.cs_run("port", streams)
    ctx = hbtgpu_cs.CsContext(2, sum(len(s) for s in streams) + (1 << 18), tid_level=1)
    for cpu, s in enumerate(streams):
        pos, last = 0, 0
        k = 0
        while pos < len(s):
            pos += int(s[pos + 6]) | int(s[pos + 7]) << 8
            k += 1
            if k % 700 == 0 or pos == len(s):
                ctx.submit(cpu, s[last:pos])
                last = pos
    ctx.run()
    ctx.sync()
    th = hbtgpu_cs.Threads()
    for cpu in range(2):
        ev, pid = ctx.events(cpu)
        ev = ev.copy()
        th.assign(ev, pid, ctx.ops(cpu))
        for f in ("tstamp", "type", "tag"):
            assert np.array_equal(ev[f], want["events"][cpu][f]), (cpu, f)
        assert ctx.state(cpu) == want["states"][cpu]
    assert th.table() == (want["threads"], want["names"])
    ctx.close()


# ---- the thread dictionary on the GPU (hbtgpu_cs_assign_vids_device) -----------------------------
@pytest.mark.parametrize("seed,kw", [
    (1, dict()),
    (2, dict(p_lost=0.01)),
    (3, dict(p_life=0.05, p_comm=0.03)),
    (4, dict(p_lost=0.005, p_life=0.05, p_comm=0.03, p_throttle=0.01, p_other=0.02)),
    (5, dict(p_lost=0.2, p_life=0.1)),
    (6, dict(p_life=0.3, p_comm=0.3)),
])
def test_device_dictionary_matches_reference_order(seed, kw):
    streams = rg.random_cs_streams(seed, 4, 6000, n_tids=48, **kw)
    want = pyoracle.cs_run("port", streams, tid_level=2)
    got = hbtgpu_cs.decode(streams, tid_level=2, device_dictionary=True)
    for cpu in range(len(streams)):
        for f in ("tstamp", "type", "level", "comp_unit_id", "tag"):
            assert np.array_equal(got["events"][cpu][f], want["events"][cpu][f]), (cpu, f)
    assert got["threads"] == want["threads"]
    assert got["names"] == want["names"]


@pytest.mark.parametrize("tid_base", [(1 << 24) - 20, 2_000_000_000])  # (pid_t is signed: below 2^31)
def test_device_dictionary_with_thread_ids_wider_than_24_bits(tid_base):
    """The dictionary's sort key is keyed for 24-bit thread ids first (pid_max is 2^22); a window
    with a wider one is noticed and keyed again with all 32 bits."""
    streams = rg.random_cs_streams(11, 3, 4000, n_tids=48, p_lost=0.004, p_life=0.05, p_comm=0.03, tid_base=tid_base)
    want = pyoracle.cs_run("port", streams, tid_level=2)
    got = hbtgpu_cs.decode(streams, tid_level=2, device_dictionary=True)
    for cpu in range(len(streams)):
        for f in ("tstamp", "type", "level", "comp_unit_id", "tag"):
            assert np.array_equal(got["events"][cpu][f], want["events"][cpu][f]), (cpu, f)
    assert got["threads"] == want["threads"]
    assert got["names"] == want["names"]


def test_device_dictionary_carries_over_windows():
    """Windows decoded one after the other: the device dictionary keeps the active tids and the
    thread table, so the result equals the host mirror fed the same windows in the same order."""
    streams = rg.random_cs_streams(77, 3, 5000, n_tids=40, p_lost=0.002, p_life=0.04, p_comm=0.03)
    cuts = []
    for s in streams:
        pos, target = 0, (len(s) // 3)
        marks = []
        while pos < len(s):
            if len(marks) < 2 and pos >= target * (len(marks) + 1):
                marks.append(pos)
            pos += int(s[pos + 6]) | int(s[pos + 7]) << 8
        cuts.append([0] + marks + [len(s)])
    ctx = hbtgpu_cs.CsContext(3, sum(len(s) for s in streams) + (1 << 16), tid_level=1)
    ctx2 = hbtgpu_cs.CsContext(3, sum(len(s) for s in streams) + (1 << 16), tid_level=1)
    dev, host = hbtgpu_cs.DeviceThreads(), hbtgpu_cs.Threads()
    for w in range(3):
        for c in (ctx, ctx2):
            c.reset()
            for cpu, s in enumerate(streams):
                if w:
                    c.set_state(cpu, *states[cpu])
                c.submit(cpu, s[cuts[cpu][w]:cuts[cpu][w + 1]])
            c.run()
            c.sync()
        states = [ctx.state(cpu) for cpu in range(3)]
        dev.assign(ctx)
        for cpu in range(3):
            ev_h, pid = ctx2.events(cpu)
            ev_h = ev_h.copy()
            host.assign(ev_h, pid, ctx2.ops(cpu))
            ev_d, _ = ctx.events(cpu)
            for f in ("tstamp", "type", "tag"):
                assert np.array_equal(ev_d[f], ev_h[f]), (w, cpu, f)
    assert dev.table() == host.table()
    with pytest.raises(hbtgpu.HbtGpuError):
        dev.assign(ctx)     # the window's tags are virtual ids already
    ctx.close()
    ctx2.close()


def test_device_dictionary_rejects_out_of_contract_records():
    T = 5_000_001
    s = rg.to_u8(rg.cs_switch(1000, 1, T) + rg.cs_fork(T, 1, T + 5, 1, 1100) + rg.cs_fork(T, 1, T + 5, 1, 1200))
    ctx = hbtgpu_cs.CsContext(1, 1 << 16)
    ctx.submit(0, s)
    ctx.run()
    ctx.sync()
    with pytest.raises(hbtgpu.HbtGpuError):
        hbtgpu_cs.DeviceThreads().assign(ctx)
    ctx.close()


def test_garbage_and_hostile_payloads():
    """Random bytes, a valid stream that turns into noise, and COMM names / pid fields made of
    bytes that look like record headers: the walk stops where the restatement stops (or never
    leaves the chain), nothing hangs, nothing is decoded behind an error."""
    r = np.random.default_rng(5)
    good = rg.random_cs_streams(51, 1, 2500, p_life=0.02, p_comm=0.02)[0]
    noise = r.integers(0, 256, 40000, dtype=np.uint8)
    noise = noise[: len(noise) // 8 * 8]
    sw32 = rg.header(rg.SWITCH_CPU_WIDE, 32)                       # 8 bytes that are a SWITCH header
    hostile = bytearray()
    ts = 1_000_000_000
    for i in range(3000):
        ts += 1000
        if i % 3 == 0:   # comm field = two fake headers; pid/tid words = fake header halves
            hostile += rg.header(rg.COMM, 48) + sw32 + sw32 + sw32 + rg.sample_id(15, 0x00200000, ts)
        else:            # next_prev_pid/tid and sample_id words that read as type 15 / size 32
            hostile += rg.header(rg.SWITCH_CPU_WIDE, 32, rg.MISC_SWITCH_OUT if i % 2 else 0) \
                + bytes(sw32) + rg.sample_id(15, 0x00200000, ts)
    streams = [noise, np.concatenate([good, noise[:8000]]), rg.to_u8(hostile), good]
    want = pyoracle.cs_run("port", streams)
    for flags in (0, hbtgpu.F_FORCE_EXACT):
        ctx = hbtgpu_cs.CsContext(len(streams), sum(len(s) for s in streams) + (1 << 16), flags=flags)
        for cpu, s in enumerate(streams):
            ctx.submit(cpu, s)
        ctx.run()
        ctx.sync()
        for cpu in range(len(streams)):
            c = ctx.counters(cpu)
            for k in CMP_COUNTERS:
                assert c[k] == want["counters"][cpu][k], (flags, cpu, k, c, want["counters"][cpu])
            ev, _ = ctx.events(cpu)
            assert len(ev) == len(want["events"][cpu])
            for f in ("tstamp", "type"):
                assert np.array_equal(ev[f], want["events"][cpu][f]), (flags, cpu, f)
            assert ctx.state(cpu) == want["states"][cpu]
        ctx.close()
