"""Interval/slice intersection (IntervalSlicer side): the C restatement pinned
against the REAL tagstack::IntervalSlicer (oracle/_ref/libhbtref_slices.so, compiled from
hbt/src/tagstack/IntervalSlicer.cpp) and against the reference's own hand-built vectors
(hbt/src/tagstack/tests/IntervalSlicerTest.cpp, transcribed below as (call, expectation)).
CPU only."""
import zlib

import numpy as np
import pytest

import slicegen
from oracle import pyoracle as po

needs_ref = pytest.mark.skipif(not po.slices_available("reference"),
                               reason="oracle/_ref/libhbtref_slices.so not built (needs /root/reference)")
KINDS = ["port"] + (["reference"] if po.slices_available("reference") else [])
A, PRE, YLD, PH = po.TT_ANALYSIS, po.TT_PREEMPTED, po.TT_YIELD, po.TT_PHASE
TA = po.TRIM | po.ABORT_ON_MISSING


def mk(kind, slices, offset, t, d, trim, abort):
    r, n = po.islice_make(kind, po.make_slices(slices), offset, t, d,
                          (po.TRIM if trim else 0) | (po.ABORT_ON_MISSING if abort else 0))
    return (None if r is None else po.slices_as_tuples(r)), n


# ---- IntervalSlicerTest.cpp ----------------------------------------------------------------------
@pytest.mark.parametrize("kind", KINDS)
def test_reference_vectors_single_consumer(kind):
    # EmptySlices :13-22
    assert mk(kind, [], 0, 0, 100, False, False) == (None, 0)
    s1 = (10, 20, 1, A, A)
    # SingleConsumerOneSliceNoTrim :24-43, ...Trim :45-61
    assert mk(kind, [s1], 0, 0, 100, False, False) == ([s1], 1)
    assert mk(kind, [s1], 0, 0, 100, True, True) == (None, 0)
    # MissingStart :63-81, MissingEnd :83-101, MissingEndNoFail :103-122
    assert mk(kind, [(10, 90, 1, A, A)], 0, 0, 100, True, True) == (None, 0)
    assert mk(kind, [(0, 99, 1, A, A)], 0, 0, 100, True, True) == (None, 0)
    assert mk(kind, [(0, 99, 1, A, A)], 0, 0, 100, True, False) == ([(0, 99, 1, A, A)], 1)
    # SingleConsumerMultiple :124-204
    s1, s2 = (0, 99, 1, A, A), (99, 2, 1, A, A)
    assert mk(kind, [s1], 0, 0, 100, True, True) == (None, 0)
    assert mk(kind, [s1, s2], 0, 0, 100, True, True) == ([s1, (99, 1, 1, A, A)], 1)
    assert mk(kind, [s1, s2], 1, 0, 101, True, True) == (None, 0)
    assert mk(kind, [s1, s2], 0, 99, 101, True, True) == (None, 0)
    assert mk(kind, [s1, s2], 0, 100, 102, True, True) == (None, 0)
    assert mk(kind, [s1, s2], 0, 100, 1, True, True) == ([(100, 1, 1, A, A)], 2)


@pytest.mark.parametrize("kind", KINDS)
def test_reference_vectors_two_consumers(kind):
    """MultipleSlicerInSequence :206-478: consumeSlices(key, ...) = make(mark[key], ...) and
    mark[key] += consumed (IntervalSlicer.cpp:195-217)."""
    mark = {"c1": 0, "c2": 0}
    slices = []

    def consume(key, t, d, trim, abort):
        r, n = mk(kind, slices, mark[key], t, d, trim, abort)
        mark[key] += n
        return [] if r is None else r

    s1 = (0, 99, 1, A, A)
    slices.append(s1)
    assert consume("c1", 0, 100, True, True) == [] and mark["c1"] == 0
    s2 = (99, 2, 33, PRE, PH)
    slices.append(s2)
    assert consume("c2", 0, 99, True, True) == [s1] and mark["c2"] == 1
    assert consume("c1", 0, 99, True, True) == [s1] and mark["c1"] == 1
    for key in ("c1", "c2"):
        assert consume(key, 0, 101, True, True) == []
    for key in ("c1", "c2"):
        assert consume(key, 99, 101, True, True) == []
    for key in ("c1", "c2"):
        assert consume(key, 100, 102, True, True) == []
    assert consume("c1", 100, 1, True, True) == [(100, 1, 33, A, PH)]
    s3 = (101, 1, 1, A, A)
    slices.append(s3)
    assert consume("c2", 100, 1, True, True) == [(100, 1, 33, A, PH)]
    s4 = (200, 10, 1, PH, PH)
    slices.append(s4)
    assert consume("c2", 100, 200, True, True) == []
    assert mark == {"c1": 2, "c2": 2}
    assert consume("c2", 100, 200, True, False) == [s3, s4] and mark["c2"] == 4
    assert consume("c1", 100, 105, True, False) == [s3, (200, 5, 1, PH, A)] and mark["c1"] == 3
    assert consume("c1", 100, 105, True, False) == [(200, 5, 1, PH, A)] and mark["c1"] == 3
    assert consume("c1", 100, 210, True, False) == [s4] and mark == {"c1": 4, "c2": 4}


# ---- quirks worth their own vectors ---------------------------------------------------------------
@pytest.mark.parametrize("kind", KINDS)
def test_quirks(kind):
    s = [(0, 100, 5, YLD, PRE), (100, 50, 6, PRE, YLD)]
    # a slice that ends exactly at the interval's start and was NOT consumed before is taken
    # with duration 0 after the trim (IntervalSlicer.cpp:76,148-158): `end < tstamp` is strict
    assert mk(kind, s, 0, 100, 50, True, True) == ([(100, 0, 5, A, PRE), (100, 50, 6, PRE, YLD)], 2)
    assert mk(kind, s, 1, 100, 50, True, True) == ([(100, 50, 6, PRE, YLD)], 1)
    # no trim: a slice reaching past the end stops the scan and is not consumed (:113-115)
    assert mk(kind, s, 0, 0, 120, False, False) == ([s[0]], 1)
    # no trim, first slice starts before the interval: durations add up past the interval -> throws (:175)
    with pytest.raises(ValueError):
        mk(kind, [(0, 150, 1, A, A)], 0, 100, 100, False, False)
    # zero-length interval: HBT_ARG_CHECK_GT(duration, 0) (:64)
    with pytest.raises(ValueError):
        mk(kind, s, 0, 10, 0, True, True)
    # zero-duration slice in the skipped prefix throws (:77-78), elsewhere it does not
    with pytest.raises(ValueError):
        mk(kind, [(5, 0, 1, A, A), (10, 100, 1, A, A)], 0, 20, 10, True, False)
    assert mk(kind, [(10, 100, 1, A, A), (110, 0, 2, A, A)], 0, 20, 200, True, False)[1] == 2
    # offset past the end is clamped (ringbuffer/Consumer.h:281)
    assert mk(kind, s, 7, 0, 100, True, True) == (None, 0)


# ---- restatement == real reference ----------------------------------------------------------------
def release_equal(a, b):
    assert a["rc"] == b["rc"], (a["rc"], b["rc"])
    assert a["n_intervals"] == b["n_intervals"] and a["err_interval"] == b["err_interval"]
    if a["rc"] == 0:
        assert a["mark"] == b["mark"]
        assert np.array_equal(a["counts"], b["counts"])
        assert po.slices_as_tuples(a["slices"]) == po.slices_as_tuples(b["slices"])


@needs_ref
@pytest.mark.parametrize("flags", [0, po.TRIM, po.ABORT_ON_MISSING, TA])
def test_port_matches_reference_regular(flags):
    r = np.random.default_rng(100 + flags)
    for trial in range(60):
        ns = int(r.integers(0, 300))
        s = slicegen.gen_slices(r, ns, hole_prob=float(r.choice([0.0, 0.05, 0.3])))
        rows = slicegen.gen_rows(r, s, int(r.integers(0, 120)), stride=int(r.integers(2, 6)),
                                 align_prob=float(r.choice([0.0, 0.5, 1.0])))
        stop = int(r.choice([2**63, int(rows[len(rows) // 2, 0]) if len(rows) else 5]))
        mark = int(r.choice([0, 0, 1, 5]))
        release_equal(po.islice_release("reference", s, rows, stop, flags, mark),
                      po.islice_release("port", s, rows, stop, flags, mark))


@needs_ref
@pytest.mark.parametrize("skind", ["zero_dur", "overlap", "unsorted", "shuffled", "ok"])
@pytest.mark.parametrize("rkind", ["overlap", "backwards", "zero_dur", "repeat", "ok"])
def test_port_matches_reference_hostile(skind, rkind):
    r = np.random.default_rng(zlib.crc32((skind + rkind).encode()))
    for flags in (0, po.TRIM, po.ABORT_ON_MISSING, TA):
        for trial in range(12):
            s = slicegen.hostile_slices(r, int(r.integers(1, 120)), skind)
            rows = slicegen.hostile_rows(r, np.sort(s, order="tstamp"), int(r.integers(1, 60)), rkind)
            release_equal(po.islice_release("reference", s, rows, 2**63, flags),
                          po.islice_release("port", s, rows, 2**63, flags))
