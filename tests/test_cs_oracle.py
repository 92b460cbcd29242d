"""ContextSwitch-mode decode: the C restatement pinned against the REAL
reference code (oracle/_ref, PerCpuThreadSwitchGenerator.h handlers + PerfEventsGroup::consume),
plus hand-computed known answers.  CPU only."""
import numpy as np
import pytest

import ringgen as rg
from oracle import pyoracle

needs_ref = pytest.mark.skipif(not pyoracle.available("reference"),
                               reason="oracle/_ref not built (needs /root/reference)")


def assert_same(a, b):
    assert len(a["events"]) == len(b["events"])
    for cpu, (x, y) in enumerate(zip(a["events"], b["events"])):
        assert len(x) == len(y), f"cpu {cpu}: {len(x)} vs {len(y)} events"
        for f in ("tstamp", "type", "level", "comp_unit_id", "tag"):
            bad = np.nonzero(x[f] != y[f])[0]
            assert bad.size == 0, f"cpu {cpu} event {bad[:3]} field {f}: {x[f][bad[:3]]} vs {y[f][bad[:3]]}"
    for cpu, (x, y) in enumerate(zip(a["counters"], b["counters"])):
        assert x == y, f"cpu {cpu}: {x} vs {y}"
    assert a["states"] == b["states"]
    assert a["threads"] == b["threads"]
    assert a["names"] == b["names"]


def test_known_answer_small():
    """Hand-walked stream: switch pair, LOST bracket (:219-253), fork/comm/exit."""
    T = 5_000_001
    s = rg.to_u8(
        rg.cs_switch(1000, 1, T, out=True, preempt=True) + rg.cs_switch(1100, 1, T + 1)
        + rg.cs_lost(7, 1500) + rg.cs_switch(1600, 1, T + 1, out=True)
        + rg.cs_fork(T + 8, 1, T + 9, 1, 1700) + rg.cs_comm(T + 8, T + 9, b"abc", 1800)
        + rg.cs_throttle(1850) + rg.cs_exit(T + 8, 1, T + 9, 1, 1900) + rg.cs_switch(2000, 1, 0))
    r = pyoracle.cs_run("port", [s], tid_level=3)
    ev = r["events"][0]
    assert [(int(e["tstamp"]), int(e["type"]), int(e["level"]), int(e["tag"])) for e in ev] == [
        (1000, rg.EV_OUT_PREEMPT, 3, 0xFFFFFFFF),
        (1100, rg.EV_IN, 3, 1),                      # first thread seen gets vid 1
        (1101, rg.EV_WE_START, 255, 0xFFFFFFFF),     # last_tstamp_ + 1
        (1600, rg.EV_WE_END, 255, 0xFFFFFFFF),       # written by the record after the LOST
        (1600, rg.EV_OUT_YIELD, 3, 0xFFFFFFFF),
        (1700, rg.EV_CREATE, 3, 2),                  # active_tids was cleared: next vid
        (1900, rg.EV_DESTROY, 3, 2),
        (2000, rg.EV_IN, 3, 0),                      # tid 0 is never remapped
    ]
    c = r["counters"][0]
    assert (c["lost"], c["comm"], c["exit"], c["fork"], c["switch_cpu_wide"], c["throttle"]) == (7, 1, 1, 1, 4, 1)
    assert c["num_records"] == 9 and c["err"] == 0
    assert r["states"] == [(2000, 0)]
    assert r["threads"] == [(1, -1, -1, 1, T + 1, 0, 2**64 - 1, 2**64 - 1),
                            (2, 1, 1, T + 8, T + 9, 1, 1700, 1900)]
    assert r["names"] == [(2, 1800, b"abc\0\0\0\0\0", 1)]    # is_exec is always true (:273)


def test_lost_after_lost_and_state_carry():
    s = rg.to_u8(rg.cs_lost(1, 500) + rg.cs_lost(2, 900) + rg.cs_switch(1000, 1, 0))
    r = pyoracle.cs_run("port", [s], states=[(100, 0)])
    ev = r["events"][0]
    assert [(int(e["tstamp"]), int(e["type"])) for e in ev] == [
        (101, rg.EV_WE_START), (900, rg.EV_WE_END), (901, rg.EV_WE_START), (1000, rg.EV_WE_END),
        (1000, rg.EV_IN)]
    r2 = pyoracle.cs_run("port", [rg.to_u8(rg.cs_lost(1, 500))], states=[(100, 1)])
    assert [(int(e["tstamp"]), int(e["type"])) for e in r2["events"][0]] == [(500, rg.EV_WE_END), (501, rg.EV_WE_START)]
    assert r2["states"] == [(500, 1)]


def test_fences():
    good = rg.cs_switch(1000, 1, 0)
    for bad in (rg.header(rg.SWITCH_CPU_WIDE, 0), rg.header(rg.SWITCH_CPU_WIDE, 28) + b"\0" * 24,
                rg.header(rg.LOST, 32) + b"\0" * 24, rg.header(rg.FORK, 64) + b"\0" * 8):
        r = pyoracle.cs_run("port", [rg.to_u8(good + bad)])
        c = r["counters"][0]
        assert c["err"] == -1 and c["err_offset"] == 32 and c["num_records"] == 1


@needs_ref
@pytest.mark.parametrize("seed,kw", [
    (1, dict()),
    (2, dict(p_lost=0.01)),
    (3, dict(p_life=0.05, p_comm=0.03)),
    (4, dict(p_lost=0.005, p_life=0.05, p_comm=0.03, p_throttle=0.01, p_other=0.02)),
    (5, dict(p_lost=0.2, p_life=0.1)),
])
def test_port_matches_reference(seed, kw):
    streams = rg.random_cs_streams(seed, 4, 1500, n_tids=48, **kw)
    for batch in (1, 64, 100000):
        ref = pyoracle.cs_run("reference", streams, tid_level=2, batch=batch)
        port = pyoracle.cs_run("port", streams, tid_level=2, batch=batch)
        assert sum(len(e) for e in ref["events"]) > 2000
        assert_same(ref, port)


@needs_ref
def test_port_matches_reference_with_state_and_empty_stream():
    streams = rg.random_cs_streams(9, 3, 400, p_lost=0.02, p_life=0.05)
    streams[1] = streams[1][:0]
    st = [(12345, 1), (0, 0), (777, 0)]
    assert_same(pyoracle.cs_run("reference", streams, states=st), pyoracle.cs_run("port", streams, states=st))
