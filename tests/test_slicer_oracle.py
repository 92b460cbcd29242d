"""tagstack Slicer (events -> slices): the C restatement for level-0 thread
events pinned against the REAL tagstack::slice::Slicer<> (oracle/_ref/libhbtref_slices.so), on the
shapes the thread-switch decode produces and on what breaks them (a thread running on two CPUs,
switch-in after switch-in, write errors, events going back in time, tid 0, repeated creations).
The first vector is TagStackSlice.1LevelsOnePhaseWithThreadSwOut of
hbt/src/tagstack/tests/SlicerTest.cpp:65-130.  CPU only."""
import numpy as np
import pytest

from oracle import pyoracle as po

needs_ref = pytest.mark.skipif(not po.slices_available("reference"),
                               reason="oracle/_ref/libhbtref_slices.so not built (needs /root/reference)")
START, END, CREATE, DESTROY, IN, OUT_P, OUT_Y, WE_START, WE_END = range(9)
NA = 0xFFFFFFFF


def events(rows):
    a = np.zeros(len(rows), po.EVENT_DTYPE)
    for i, (ts, typ, cu, tag) in enumerate(rows):
        a[i] = (ts, typ, 255 if typ in (WE_START, WE_END) else 0, cu, tag)
    return a


def same(a, b):
    assert a["rc"] == b["rc"]
    assert a["stats"] == b["stats"]
    assert a["last_ts"] == b["last_ts"]
    for u, (x, y) in enumerate(zip(a["slices"], b["slices"])):
        assert po.slices_as_tuples(x) == po.slices_as_tuples(y), u
    sa, sb = a["stacks"], b["stacks"]
    assert len(sa) == len(sb)
    for k in ("stack_id", "parent_stack_id", "num_set_tags", "known_num_levels", "cum_active_time",
              "num_swins", "num_swouts"):
        assert np.array_equal(sa[k], sb[k]), k
    assert np.array_equal(sa["tags"][:, 0], sb["tags"][:, 0])


def test_reference_vector_one_level():
    ev = events([(20, IN, 0, 2), (30, OUT_P, 0, NA), (30, IN, 0, 4), (50, OUT_P, 0, NA), (60, IN, 0, 1),
                 (80, OUT_P, 0, NA)])
    r = po.slicer_run("port", ev, 1)
    Y, P = po.TT_YIELD, po.TT_PREEMPTED
    # stack ids: {} unknown = 0 (parent of {2}?), {2}? = 1, {} known = 2, {4} = 3, {1} = 4 (SlicerTest.cpp:93-105)
    assert po.slices_as_tuples(r["slices"][0]) == [(20, 10, 1, Y, P), (30, 20, 3, Y, P), (50, 10, 2, P, Y),
                                                   (60, 20, 4, Y, P)]
    assert r["stats"]["num_processed"] == 6 and r["stats"]["num_slices_completed"] == 5


def random_events(r, n_units, n, n_tags, p_create=0.03, p_we=0.01, p_double_in=0.01, p_migrate=0.05,
                  p_back=0.0, p_tid0=0.2, p_same_ts=0.05):
    rows = []
    t = 1000
    running = [None] * n_units
    we = 0
    for _ in range(n):
        if r.random() > p_same_ts:
            t += int(r.integers(1, 50))
        ts = t - int(r.integers(1, 200)) if r.random() < p_back else t
        cu = int(r.integers(0, n_units))
        x = r.random()
        if x < p_create:
            rows.append((ts, CREATE if r.random() < 0.8 else DESTROY, cu, int(r.integers(0 if r.random() < 0.05 else 1, n_tags))))
        elif x < p_create + p_we:
            if we and r.random() < 0.7:
                rows.append((ts, WE_END, cu, NA)); we -= 1
            else:
                rows.append((ts, WE_START, cu, NA)); we += 1
        elif running[cu] is None or r.random() < p_double_in:
            if r.random() < p_tid0:
                tag = 0
            elif r.random() < p_migrate:
                others = [v for v in running if v]
                tag = int(r.choice(others)) if others else int(r.integers(1, n_tags))
            else:
                tag = int(r.integers(1, n_tags))
            rows.append((ts, IN, cu, tag)); running[cu] = tag
        else:
            rows.append((ts, OUT_P if r.random() < 0.5 else OUT_Y, cu, NA)); running[cu] = None
    while we:
        t += 5
        rows.append((t, WE_END, 0, NA)); we -= 1
    return events(rows)


@needs_ref
@pytest.mark.parametrize("kw", [dict(), dict(p_we=0.0, p_double_in=0.0, p_migrate=0.0), dict(p_back=0.02),
                                dict(p_double_in=0.2), dict(p_migrate=0.5, p_create=0.2), dict(p_we=0.1),
                                dict(p_tid0=0.9), dict(p_same_ts=0.7)])
def test_port_matches_reference(kw):
    r = np.random.default_rng(len(repr(kw)))
    for trial in range(25):
        n_units = int(r.integers(1, 6))
        ev = random_events(r, n_units, int(r.integers(1, 600)), int(r.choice([2, 5, 40])), **kw)
        same(po.slicer_run("reference", ev, n_units), po.slicer_run("port", ev, n_units))


def test_unsupported_events_are_refused():
    ev = events([(10, IN, 0, 3), (20, START, 0, 7)])
    ev["level"][1] = 1
    assert po.slicer_run("port", ev, 1)["rc"] == -95        # -ENOTSUP: phases need the general dictionary
    ev = events([(10, IN, 0, 3)])
    ev["level"][0] = 2
    assert po.slicer_run("port", ev, 1)["rc"] == -95


# ---- the general restatement (any level, phase events): oracle/hbt_slicer_general.c ----------------
def events_l(rows):
    """rows of (tstamp, type, level, comp_unit, tag)"""
    a = np.zeros(len(rows), po.EVENT_DTYPE)
    for i, (ts, typ, lvl, cu, tag) in enumerate(rows):
        a[i] = (ts, typ, lvl, cu, tag)
    return a


def stack_rows(r):
    return [(int(s["stack_id"]), [int(t) for t in s["tags"][:s["num_set_tags"]]], int(s["num_set_tags"]),
             int(s["known_num_levels"])) for s in r["stacks"]]


Y, P, PH = po.TT_YIELD, po.TT_PREEMPTED, 4


def test_general_reference_vectors():
    """TagStackSlice.* of hbt/src/tagstack/tests/SlicerTest.cpp with their expected stack ids and slices."""
    # :63 1LevelsOnePhaseWithThreadSwOut
    r = po.slicer_run("port_general", events_l([(20, IN, 0, 0, 2), (30, OUT_P, 0, 0, NA), (30, IN, 0, 0, 4),
                                                (50, OUT_P, 0, 0, NA), (60, IN, 0, 0, 1), (80, OUT_P, 0, 0, NA)]), 1)
    assert po.slices_as_tuples(r["slices"][0]) == [(20, 10, 1, Y, P), (30, 20, 3, Y, P), (50, 10, 2, P, Y), (60, 20, 4, Y, P)]
    # :150 2LevelsOnePhaseWithThreadSwOutAndStartOneLevelAbove: switches at level 1
    r = po.slicer_run("port_general", events_l([(20, IN, 1, 0, 2), (30, OUT_P, 1, 0, NA), (31, IN, 1, 0, 2),
                                                (32, OUT_P, 1, 0, NA)]), 1)
    assert stack_rows(r) == [(0, [], 0, 0), (1, [NA], 1, 0), (2, [NA, 2], 2, 0), (3, [], 0, 1), (4, [NA], 1, 1),
                             (5, [NA, 2], 2, 1)]
    assert po.slices_as_tuples(r["slices"][0]) == [(20, 10, 2, Y, P), (30, 1, 4, P, Y), (31, 1, 5, Y, P)]
    # :213 2LevelsOnePhaseWithThreadSwOutLong: a phase starts on top of a thread and comes back with it
    r = po.slicer_run("port_general", events_l([(100, IN, 0, 0, 0), (130, OUT_P, 0, 0, NA), (131, IN, 0, 0, 1),
                                                (140, START, 1, 0, 10), (160, OUT_P, 0, 0, NA), (170, IN, 0, 0, 1),
                                                (190, OUT_Y, 0, 0, NA)]), 1)
    assert stack_rows(r) == [(0, [], 0, 0), (1, [0], 1, 0), (2, [], 0, 1), (3, [1], 1, 0), (4, [1], 1, 1), (5, [1, 10], 2, 1)]
    assert po.slices_as_tuples(r["slices"][0]) == [(100, 30, 1, Y, P), (130, 1, 2, P, Y), (131, 9, 3, Y, PH),
                                                   (140, 20, 5, PH, P), (160, 10, 2, P, Y), (170, 20, 5, Y, Y)]
    # :323 3LevelsMultiplePhasesNoNesting: starts with a switch-out, zero-length slices are dropped
    t = 100
    r = po.slicer_run("port_general", events_l([(0, OUT_P, 0, 0, NA), (0, IN, 0, 0, t), (0, OUT_P, 0, 0, NA), (1, IN, 0, 0, t),
                                                (30, OUT_P, 0, 0, NA), (30, IN, 0, 0, t), (35, OUT_P, 0, 0, NA),
                                                (36, IN, 0, 0, t), (40, OUT_P, 0, 0, NA), (42, IN, 0, 0, t),
                                                (42, OUT_P, 0, 0, NA), (50, IN, 0, 0, t)]), 1)
    assert stack_rows(r) == [(0, [], 0, 1), (1, [], 0, 0), (2, [t], 1, 0)]
    assert po.slices_as_tuples(r["slices"][0]) == [(0, 1, 0, P, Y), (1, 29, 2, Y, P), (30, 5, 2, Y, P), (35, 1, 0, P, Y),
                                                   (36, 4, 2, Y, P), (40, 2, 0, P, Y), (42, 8, 0, P, Y)]
    # :432 3LevelsMultiplePhasesTwoStreams (merge order: tstamp, then comp unit)
    r = po.slicer_run("port_general", events_l([(0, OUT_P, 0, 0, NA), (0, IN, 0, 0, t), (0, OUT_P, 0, 0, NA), (1, IN, 0, 1, t),
                                                (30, OUT_P, 0, 1, NA), (31, IN, 0, 0, t), (35, OUT_P, 0, 0, NA)]), 2)
    assert stack_rows(r) == [(0, [], 0, 1), (1, [], 0, 0), (2, [t], 1, 0)]
    assert po.slices_as_tuples(r["slices"][1]) == [(1, 29, 2, Y, P)]
    assert po.slices_as_tuples(r["slices"][0]) == [(0, 31, 0, P, Y), (31, 4, 2, Y, P)]


def random_phase_events(r, n_units, n, n_tags, n_phases=4, p_phase=0.35, p_we=0.01, p_create=0.03, p_bad=0.02,
                        p_back=0.0):
    """Threads at level 0, phases started / ended on top of the running thread at levels 1..3 (as
    PhaseLoggers emit them); p_bad: events that break the preconditions (a Start on an occupied
    level, an End above the stack, a switch-in on a busy unit)."""
    rows, t = [], 1000
    depth = [0] * n_units      # levels in use on the unit (0: nothing runs)
    we = 0
    for _ in range(n):
        t += int(r.integers(0, 40))
        ts = t - int(r.integers(1, 100)) if r.random() < p_back else t
        cu = int(r.integers(0, n_units))
        x = r.random()
        if x < p_create:
            rows.append((ts, CREATE, 0, cu, int(r.integers(1, n_tags))))
        elif x < p_create + p_we:
            if we and r.random() < 0.7:
                rows.append((ts, WE_END, 255, cu, NA)); we -= 1
            else:
                rows.append((ts, WE_START, 255, cu, NA)); we += 1
                depth = [0] * n_units
        elif x < p_create + p_we + p_bad:
            kind = int(r.integers(0, 3))
            if kind == 0:
                rows.append((ts, START, max(1, depth[cu] - 1) if depth[cu] > 1 else 1, cu, int(r.integers(1, n_phases + 1))))
            elif kind == 1:
                rows.append((ts, END, min(3, depth[cu] + 1), cu, NA))
            else:
                rows.append((ts, IN, 0, cu, int(r.integers(0, n_tags))))
            depth = [0] * n_units if True else depth  # whatever the reference does with it, start afresh below
        elif depth[cu] == 0:
            rows.append((ts, IN, 0, cu, int(r.integers(0, n_tags)))); depth[cu] = 1
        elif r.random() < p_phase and depth[cu] < 4:
            rows.append((ts, START, depth[cu], cu, int(r.integers(1, n_phases + 1)))); depth[cu] += 1
        elif depth[cu] > 1 and r.random() < 0.5:
            depth[cu] -= 1
            rows.append((ts, END, depth[cu], cu, NA))
        else:
            rows.append((ts, OUT_P if r.random() < 0.5 else OUT_Y, 0, cu, NA)); depth[cu] = 0
    while we:
        t += 3
        rows.append((t, WE_END, 255, 0, NA)); we -= 1
    return events_l(rows)


def same_general(a, b):
    assert a["rc"] == b["rc"]
    assert a["stats"] == b["stats"]
    assert a["last_ts"] == b["last_ts"]
    for u, (x, y) in enumerate(zip(a["slices"], b["slices"])):
        assert po.slices_as_tuples(x) == po.slices_as_tuples(y), u
    assert stack_rows(a) == stack_rows(b)
    for k in ("parent_stack_id", "cum_active_time", "num_swins", "num_swouts"):
        assert np.array_equal(a["stacks"][k], b["stacks"][k]), k


@needs_ref
@pytest.mark.parametrize("kw", [dict(), dict(p_bad=0.0, p_we=0.0), dict(p_phase=0.7), dict(p_bad=0.1), dict(p_back=0.02),
                                dict(p_create=0.2), dict(p_we=0.08)])
def test_general_port_matches_reference_on_phase_streams(kw):
    r = np.random.default_rng(7 + len(repr(kw)))
    compared = 0
    for trial in range(30):
        n_units = int(r.integers(1, 5))
        ev = random_phase_events(r, n_units, int(r.integers(1, 500)), int(r.choice([3, 6, 30])), **kw)
        port = po.slicer_run("port_general", ev, n_units)
        if port["rc"] == -95:  # fenced: the reference would read tags it never set; cut the stream in front of that
            k = len(ev)
            while k and po.slicer_run("port_general", ev[:k], n_units)["rc"] == -95:
                k -= 1
            ev = ev[:k]
            port = po.slicer_run("port_general", ev, n_units)
        compared += 1
        same_general(po.slicer_run("reference", ev, n_units), port)
    assert compared == 30


@needs_ref
def test_general_port_matches_reference_on_level0_streams():
    r = np.random.default_rng(11)
    for kw in (dict(), dict(p_double_in=0.2), dict(p_migrate=0.5, p_create=0.2), dict(p_we=0.1), dict(p_back=0.02)):
        for trial in range(10):
            n_units = int(r.integers(1, 6))
            ev = random_events(r, n_units, int(r.integers(1, 400)), int(r.choice([2, 5, 40])), **kw)
            same_general(po.slicer_run("reference", ev, n_units), po.slicer_run("port_general", ev, n_units))
