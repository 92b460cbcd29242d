"""GPU parity of the Slicer (events -> slices): hbtgpu_slicer_run through the
C ABI against the C restatement (pinned to the real tagstack::slice::Slicer<> by
tests/test_slicer_oracle.py) and, where oracle/_ref travelled along, against the real class itself:
every slice field, the stack table (ids, parents, tags, known levels, active time, switch counts),
the statistics, the last timestamp.  Input outside the parallel form's scope must be refused
loudly."""
import numpy as np
import pytest

from dynolog_b200 import hbtgpu_cs, hbtgpu_slices as gs, synth
from dynolog_b200.hbtgpu import HbtGpuError
from oracle import pyoracle as po
from test_slicer_oracle import CREATE, IN, NA, OUT_P, OUT_Y, START, WE_END, WE_START, events, random_events

pytestmark = pytest.mark.gpu
KINDS = ["port"] + (["reference"] if po.slices_available("reference") else [])


@pytest.fixture(scope="module")
def slicer():
    s = gs.Slicer(0)
    yield s
    s.close()


def split_units(ev, n_units):
    """merge-ordered events -> per-unit streams + the merge order the Combinator tree would produce."""
    units = [ev[ev["comp_unit_id"] == u] for u in range(n_units)]
    merged = np.concatenate(units) if n_units else ev[:0]
    order = np.lexsort((merged["comp_unit_id"], merged["tstamp"]))
    return units, merged[order]


def compare(got, exp):
    assert exp["rc"] == 0
    assert got["stats"] == exp["stats"]
    assert got["last_ts"] == exp["last_ts"]
    for u, (x, y) in enumerate(zip(got["slices"], exp["slices"])):
        assert po.slices_as_tuples(x) == po.slices_as_tuples(y), u
    a, b = got["stacks"], exp["stacks"]
    assert len(a) == len(b)
    for k in ("stack_id", "parent_stack_id", "num_set_tags", "known_num_levels", "cum_active_time", "num_swins",
              "num_swouts"):
        assert np.array_equal(a[k], b[k]), k
    assert np.array_equal(a["tags"][:, 0], b["tags"][:, 0])


def supported(ev, n_units):
    """does the window stay inside the parallel form? (no WriteErrorsEnd without a Start)"""
    depth = 0
    for e in ev:
        t = int(e["type"])
        if t == WE_END:
            if depth == 0:
                return False
            depth -= 1
        elif t == WE_START:
            depth += 1
    return True


def test_reference_vector(slicer):
    ev = events([(20, IN, 0, 2), (30, OUT_P, 0, NA), (30, IN, 0, 4), (50, OUT_P, 0, NA), (60, IN, 0, 1), (80, OUT_P, 0, NA)])
    got = slicer.run([ev])
    Y, P = po.TT_YIELD, po.TT_PREEMPTED
    assert po.slices_as_tuples(got["slices"][0]) == [(20, 10, 1, Y, P), (30, 20, 3, Y, P), (50, 10, 2, P, Y), (60, 20, 4, Y, P)]
    for kind in KINDS:
        compare(got, po.slicer_run(kind, ev, 1))


@pytest.mark.parametrize("kw", [dict(), dict(p_double_in=0.15), dict(p_double_in=0.5, p_we=0.02), dict(p_double_in=0.05, p_migrate=0.5, p_tid0=0.5),
                                dict(p_double_in=0.0), dict(p_double_in=0.0, p_we=0.0, p_migrate=0.0),
                                dict(p_double_in=0.0, p_migrate=0.5, p_create=0.2), dict(p_double_in=0.0, p_we=0.1),
                                dict(p_double_in=0.0, p_tid0=0.9), dict(p_double_in=0.0, p_same_ts=0.7),
                                dict(p_double_in=0.0, p_create=0.4, p_tid0=0.5)])
def test_random_windows(slicer, kw):
    r = np.random.default_rng(len(repr(kw)) + 1)
    done = 0
    for trial in range(60):
        n_units = int(r.integers(1, 7))
        ev = random_events(r, n_units, int(r.integers(1, 900)), int(r.choice([2, 5, 40, 300])), **kw)
        units, merged = split_units(ev, n_units)
        if not supported(merged, n_units):
            with pytest.raises(HbtGpuError):
                slicer.run(units)
            continue
        got = slicer.run(units)
        for kind in KINDS:
            compare(got, po.slicer_run(kind, merged, n_units))
        done += 1
    assert done > 20


def test_lifetime_events_at_other_levels_are_dropped(slicer):
    r = np.random.default_rng(77)
    for trial in range(20):
        n_units = int(r.integers(1, 5))
        ev = random_events(r, n_units, int(r.integers(20, 600)), 12, p_create=0.3, p_double_in=0.05)
        life = (ev["type"] == CREATE) | (ev["type"] == 3)
        ev["level"][life & (r.random(len(ev)) < 0.4)] = 2
        units, merged = split_units(ev, n_units)
        if not supported(merged, n_units):
            continue
        got = slicer.run(units)
        for kind in KINDS:
            compare(got, po.slicer_run(kind, merged, n_units))
        assert got["stats"]["num_logic_errors"] >= 0


def test_refusals(slicer):
    # (a switch-in on a unit whose active state is a thread is not refused: it resets, as in the reference)
    got = slicer.run([events([(10, IN, 0, 3), (20, IN, 0, 4), (30, OUT_P, 0, NA), (40, IN, 0, 3), (50, OUT_Y, 0, NA)])])
    for kind in KINDS:
        compare(got, po.slicer_run(kind, events([(10, IN, 0, 3), (20, IN, 0, 4), (30, OUT_P, 0, NA), (40, IN, 0, 3),
                                                 (50, OUT_Y, 0, NA)]), 1))
    # phases in a LATER window of a Slicer that began on the parallel form: refused, the state stays
    slicer.run([events([(10, IN, 0, 3)])])
    with pytest.raises(HbtGpuError, match="HBTGPU_SLICER_GENERAL"):
        slicer.run([events([(20, START, 0, 7)])], cont=True)
    got = slicer.run([events([(20, OUT_P, 0, NA)])], cont=True)
    compare(got, po.slicer_run("port", events([(10, IN, 0, 3), (20, OUT_P, 0, NA)]), 1))
    with pytest.raises(HbtGpuError, match="back in time"):
        slicer.run([events([(30, IN, 0, 3), (20, OUT_P, 0, NA)])])
    with pytest.raises(HbtGpuError, match="comp_unit_id"):
        slicer.run([events([(10, IN, 1, 3)])])
    with pytest.raises(HbtGpuError, match="comp_unit_id"):
        slicer.run([events([(10, IN, 1, 3)])], general=True)
    with pytest.raises(HbtGpuError, match="back in time"):
        slicer.run([events([(30, IN, 0, 3), (20, OUT_P, 0, NA)])], general=True)
    ev = events([(10, IN, 0, 3)])
    ev["type"][0] = 9
    with pytest.raises(HbtGpuError, match="unknown event types"):
        slicer.run([ev])
    ev = events([(10, IN, 0, 3)])
    ev["level"][0] = 6                        # HBT_DCHECK_LT(level, kNumLevels)
    with pytest.raises(HbtGpuError, match="level >= 6"):
        slicer.run([ev])
    # the marker of a carried state is not a level an event may have: as any lifetime event off level 0, dropped
    ev = events([(10, CREATE, 0, 3), (20, IN, 0, 3)])
    ev["level"][0] = 0xEE
    compare(slicer.run([ev]), po.slicer_run("port_general", ev, 1))


# ---- the general path: phases, switches at any level (hbtgpu_slicer_general.inc) -----------------------------
from test_slicer_oracle import END, events_l, random_phase_events, same_general, stack_rows  # noqa: E402

PH = 4


def run_general(slicer, ev, n_units, **kw):
    units = [ev[ev["comp_unit_id"] == u] for u in range(n_units)]
    got = slicer.run(units, **kw)
    got["rc"] = 0
    return got


def merge_order(ev):
    return ev[np.lexsort((ev["comp_unit_id"], ev["tstamp"]))] if len(ev) else ev


def test_general_reference_vectors(slicer):
    """TagStackSlice.* of hbt/src/tagstack/tests/SlicerTest.cpp: expected stack ids and slices, and the oracle."""
    Y, P = po.TT_YIELD, po.TT_PREEMPTED
    t = 100
    cases = [
        (1, [(20, IN, 1, 0, 2), (30, OUT_P, 1, 0, NA), (31, IN, 1, 0, 2), (32, OUT_P, 1, 0, NA)],
         [(0, [], 0, 0), (1, [NA], 1, 0), (2, [NA, 2], 2, 0), (3, [], 0, 1), (4, [NA], 1, 1), (5, [NA, 2], 2, 1)],
         {0: [(20, 10, 2, Y, P), (30, 1, 4, P, Y), (31, 1, 5, Y, P)]}),
        (1, [(100, IN, 0, 0, 0), (130, OUT_P, 0, 0, NA), (131, IN, 0, 0, 1), (140, START, 1, 0, 10), (160, OUT_P, 0, 0, NA),
             (170, IN, 0, 0, 1), (190, OUT_Y, 0, 0, NA)],
         [(0, [], 0, 0), (1, [0], 1, 0), (2, [], 0, 1), (3, [1], 1, 0), (4, [1], 1, 1), (5, [1, 10], 2, 1)],
         {0: [(100, 30, 1, Y, P), (130, 1, 2, P, Y), (131, 9, 3, Y, PH), (140, 20, 5, PH, P), (160, 10, 2, P, Y),
              (170, 20, 5, Y, Y)]}),
        (1, [(0, OUT_P, 0, 0, NA), (0, IN, 0, 0, t), (0, OUT_P, 0, 0, NA), (1, IN, 0, 0, t), (30, OUT_P, 0, 0, NA),
             (30, IN, 0, 0, t), (35, OUT_P, 0, 0, NA), (36, IN, 0, 0, t), (40, OUT_P, 0, 0, NA), (42, IN, 0, 0, t),
             (42, OUT_P, 0, 0, NA), (50, IN, 0, 0, t)],
         [(0, [], 0, 1), (1, [], 0, 0), (2, [t], 1, 0)],
         {0: [(0, 1, 0, P, Y), (1, 29, 2, Y, P), (30, 5, 2, Y, P), (35, 1, 0, P, Y), (36, 4, 2, Y, P), (40, 2, 0, P, Y),
              (42, 8, 0, P, Y)]}),
        (2, [(0, OUT_P, 0, 0, NA), (0, IN, 0, 0, t), (0, OUT_P, 0, 0, NA), (1, IN, 0, 1, t), (30, OUT_P, 0, 1, NA),
             (31, IN, 0, 0, t), (35, OUT_P, 0, 0, NA)],
         [(0, [], 0, 1), (1, [], 0, 0), (2, [t], 1, 0)],
         {1: [(1, 29, 2, Y, P)], 0: [(0, 31, 0, P, Y), (31, 4, 2, Y, P)]}),
    ]
    for n_units, rows, stacks, slices in cases:
        ev = events_l(rows)
        got = run_general(slicer, ev, n_units, general=True)
        assert stack_rows(got) == stacks
        for u, exp in slices.items():
            assert po.slices_as_tuples(got["slices"][u]) == exp
        same_general(got, po.slicer_run("port_general", merge_order(ev), n_units))


def cut_at_fence(ev, n_units):
    """the longest prefix the restatement accepts (past it the reference reads tags it never set)"""
    ev = merge_order(ev)
    port = po.slicer_run("port_general", ev, n_units)
    fenced = port["rc"] == -95
    if fenced:
        lo, hi = 0, len(ev)  # rc is monotone in the prefix length
        while lo < hi:
            mid = (lo + hi + 1) // 2
            if po.slicer_run("port_general", ev[:mid], n_units)["rc"] == -95:
                hi = mid - 1
            else:
                lo = mid
        ev = ev[:lo]
        port = po.slicer_run("port_general", ev, n_units)
    return ev, port, fenced


@pytest.mark.parametrize("kw", [dict(), dict(p_bad=0.0, p_we=0.0), dict(p_phase=0.7), dict(p_bad=0.1), dict(p_create=0.2),
                                dict(p_we=0.08)])
def test_general_random_phase_streams(slicer, kw):
    r = np.random.default_rng(70 + len(repr(kw)))
    n_fenced = 0
    for trial in range(25):
        n_units = int(r.integers(1, 5))
        full = random_phase_events(r, n_units, int(r.integers(1, 600)), int(r.choice([3, 6, 30])), **kw)
        ev, port, fenced = cut_at_fence(full, n_units)
        assert port["rc"] == 0
        same_general(run_general(slicer, ev, n_units), port)            # (phases: the fresh window changes over by itself)
        if po.slices_available("reference"):
            same_general(po.slicer_run("reference", ev, n_units), port)
        if fenced:                                                        # ... and the whole stream is refused as the oracle refuses it
            n_fenced += 1
            with pytest.raises(HbtGpuError, match="never set") as ei:
                run_general(slicer, merge_order(full), n_units, general=True)
            assert ei.value.code == -95
    assert kw.get("p_bad", 0.02) == 0.0 or n_fenced >= 0


def test_general_level0_streams_match_the_parallel_form(slicer):
    r = np.random.default_rng(12)
    for kw in (dict(), dict(p_double_in=0.2), dict(p_migrate=0.5, p_create=0.2), dict(p_we=0.1)):
        for trial in range(8):
            n_units = int(r.integers(1, 6))
            ev = merge_order(random_events(r, n_units, int(r.integers(1, 400)), int(r.choice([2, 5, 40])), **kw))
            exp = po.slicer_run("port_general", ev, n_units)
            same_general(run_general(slicer, ev, n_units, general=True), exp)
            if supported(ev, n_units):
                compare(run_general(slicer, ev, n_units), exp)
            else:                          # a WriteErrorsEnd without its Start: the fresh window changes over by itself
                same_general(run_general(slicer, ev, n_units), exp)


def test_general_windows_continue_each_other(slicer):
    r = np.random.default_rng(13)
    done = 0
    while done < 12:
        n_units = int(r.integers(1, 5))
        ev, exp, _ = cut_at_fence(random_phase_events(r, n_units, int(r.integers(50, 800)), int(r.choice([3, 6, 30])), p_we=0.03),
                                  n_units)
        if len(ev) < 10:
            continue
        cuts = sorted(set(int(x) for x in r.choice(ev["tstamp"], size=int(r.integers(1, 5)))))
        slices = [[] for _ in range(n_units)]
        lo = 0
        for k, cut in enumerate(cuts + [int(ev["tstamp"].max()) + 1]):
            w = ev[(ev["tstamp"] >= lo) & (ev["tstamp"] < cut)]
            got = run_general(slicer, w, n_units, cont=k > 0, general=True)
            for u in range(n_units):
                slices[u].append(got["slices"][u])
            lo = cut
        got["slices"] = [np.concatenate(x) for x in slices]
        same_general(got, exp)
        done += 1


def test_general_grows_its_dictionary(slicer):
    """more stacks than the first reservation holds: the walk stops in front of the event, the tables are
    rebuilt (hash slots and the heads of the dormant lists), the walk goes on"""
    n = 40_000
    rows = []
    for i in range(n):                 # every thread runs one phase of its own: 3 new stacks each, all parked
        t = 1000 + 10 * i
        rows += [(t, IN, 0, 0, 10 + i), (t + 2, START, 1, 0, 1 + i % 7), (t + 5, END, 1, 0, NA), (t + 7, OUT_Y, 0, 0, NA)]
    for i in range(0, n, 97):          # ... and comes back: found in its dormant list
        t = 1000 + 10 * n + 10 * i
        rows += [(t, IN, 0, 0, 10 + i), (t + 3, OUT_P, 0, 0, NA)]
    ev = events_l(rows)
    exp = po.slicer_run("port_general", ev, 1)
    assert exp["rc"] == 0 and len(exp["stacks"]) > 2 * n
    same_general(run_general(slicer, ev, 1), exp)
    half = len(ev) // 2                # the same over two windows: the second one grows tables that hold lists
    a = run_general(slicer, ev[:half], 1, general=True)
    b = run_general(slicer, ev[half:], 1, general=True, cont=True)
    b["slices"] = [np.concatenate([a["slices"][0], b["slices"][0]])]
    same_general(b, exp)


def test_decoded_capture_to_slices_on_the_device(slicer):
    """thread-switch images -> events (GPU) -> slices (GPU) -> per-interval slices (GPU): the
    slices never leave the device between the last two stages."""
    n_cpus = 8
    streams = synth.cs_capture(n_cpus, 256 * 1024, n_tids=200, life_ppm=20000, comm_ppm=0, lost_ppm=1000,
                               throttle_ppm=0, seed=5)
    dec = hbtgpu_cs.decode(streams, tid_level=0)
    units = dec["events"]
    merged = np.concatenate(units)
    merged = merged[np.lexsort((merged["comp_unit_id"], merged["tstamp"]))]
    got = slicer.run(units)
    assert got["stats"]["num_slices_completed"] > 3000 * n_cpus and got["stats"]["num_error_start"] > 0
    for kind in KINDS:
        compare(got, po.slicer_run(kind, merged, n_cpus))
    # hand the device-resident slices to the interval slicer
    import torch
    d_slices, d_begin = slicer.slices_device()
    sb = np.concatenate([[0], np.cumsum(got["n_slices"])]).astype(np.uint64)
    t0 = int(min(s["tstamp"][0] for s in got["slices"] if len(s)))
    rows = []
    for u in range(n_cpus):
        k = np.arange(200, dtype=np.uint64)
        rows.append(np.stack([np.uint64(t0) + k * np.uint64(100_000), np.full(200, 100_000, np.uint64)], axis=1))
    rb = np.concatenate([[0], np.cumsum([len(x) for x in rows])]).astype(np.uint64)
    d_rows = torch.from_numpy(np.concatenate(rows).view(np.int64)).cuda()
    torch.cuda.synchronize()
    isl = gs.IntervalSlicer(0)
    try:
        marks = np.zeros(n_cpus, np.uint64)
        flags = po.TRIM | po.ABORT_ON_MISSING
        u_res = isl.release_device(d_slices, sb, d_rows.data_ptr(), 2, rb, 2**63, flags, marks)
        assert (u_res["status"] == 0).all()
        for u in range(n_cpus):
            exp = po.islice_release("port", got["slices"][u], rows[u], 2**63, flags, 0)
            counts, out = isl.get(u)
            assert np.array_equal(counts, exp["counts"]) and po.slices_as_tuples(out) == po.slices_as_tuples(exp["slices"])
    finally:
        isl.close()


@pytest.mark.parametrize("kw", [dict(p_double_in=0.0), dict(p_double_in=0.0, p_we=0.05), dict(p_double_in=0.0, p_migrate=0.4, p_create=0.2),
                                dict(p_double_in=0.0, p_tid0=0.7, p_same_ts=0.3), dict(p_double_in=0.1), dict(p_double_in=0.3, p_we=0.03)])
def test_windows_continue_each_other(slicer, kw):
    """One stream cut into windows (HBTGPU_SLICER_CONTINUE) gives the slices, the stack table and the
    statistics of one Slicer fed the whole stream: threads running across a cut, dormant states,
    write errors open at a cut, ids continuing."""
    r = np.random.default_rng(len(repr(kw)) + 7)
    done = 0
    for trial in range(40):
        n_units = int(r.integers(1, 6))
        ev = random_events(r, n_units, int(r.integers(50, 1500)), int(r.choice([3, 10, 80])), **kw)
        units, merged = split_units(ev, n_units)
        if not supported(merged, n_units):
            continue
        exp = po.slicer_run("port", merged, n_units)
        # cut at 1-4 timestamps: every unit's events before the cut go to the earlier window
        cuts = sorted(set(int(x) for x in r.choice(merged["tstamp"], size=int(r.integers(1, 5)))))
        bounds = [0] + cuts + [2**63]
        got_slices = [[] for _ in range(n_units)]
        got = None
        for w in range(len(bounds) - 1):
            win = [u_[(u_["tstamp"] >= bounds[w]) & (u_["tstamp"] < bounds[w + 1])] for u_ in units]
            got = slicer.run(win, cont=w > 0)
            for u in range(n_units):
                got_slices[u].append(got["slices"][u])
        got["slices"] = [np.concatenate(x) for x in got_slices]
        compare(got, exp)
        done += 1
    assert done > 10


def test_large_capture(slicer):
    """64 CPUs x 2 MiB of thread-switch records (4 M events, ~170 k thread ids, LOST windows): every
    slice and the whole stack table against the restatement; in two windows as well."""
    n_cpus = 64
    streams = synth.cs_capture(n_cpus, 2 << 20, n_tids=4096, life_ppm=2000, comm_ppm=0, lost_ppm=50,
                               throttle_ppm=0, seed=11)
    units = hbtgpu_cs.decode(streams, tid_level=0, device_dictionary=True)["events"]
    merged = np.concatenate(units)
    merged = merged[np.lexsort((merged["comp_unit_id"], merged["tstamp"]))]
    exp = po.slicer_run("port", merged, n_cpus)
    assert exp["rc"] == 0 and len(merged) > 3_000_000
    got = slicer.run(units)
    for u in range(n_cpus):
        for k in ("tstamp", "duration", "stack_id", "swin_type", "swout_type"):
            assert np.array_equal(got["slices"][u][k], exp["slices"][u][k]), (u, k)
    assert got["stats"] == exp["stats"] and got["last_ts"] == exp["last_ts"]
    for k in ("stack_id", "parent_stack_id", "num_set_tags", "known_num_levels", "cum_active_time", "num_swins", "num_swouts"):
        assert np.array_equal(got["stacks"][k], exp["stacks"][k]), k
    # the same capture cut in the middle
    cut = int(np.median(merged["tstamp"]))
    first = slicer.run([u_[u_["tstamp"] < cut] for u_ in units])
    second = slicer.run([u_[u_["tstamp"] >= cut] for u_ in units], cont=True)
    for u in range(n_cpus):
        both = np.concatenate([first["slices"][u], second["slices"][u]])
        for k in ("tstamp", "duration", "stack_id", "swin_type", "swout_type"):
            assert np.array_equal(both[k], exp["slices"][u][k]), (u, k)
    assert second["stats"] == exp["stats"]
    for k in ("stack_id", "cum_active_time", "num_swins", "num_swouts"):
        assert np.array_equal(second["stacks"][k], exp["stacks"][k]), k
