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
