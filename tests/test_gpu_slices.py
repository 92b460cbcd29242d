"""GPU parity of the count-interval x slice intersection (IntervalSlicer side):
hbtgpu_islice_release through the C ABI against the C restatement (itself pinned to the real
tagstack::IntervalSlicer by tests/test_slices_oracle.py), bit-exact: per-interval counts, every
slice field, marks, the stop_ts cut, and the rows the reference throws on."""
import zlib

import numpy as np
import pytest

import slicegen
from dynolog_b200 import hbtgpu_slices as gs
from oracle import pyoracle as po

pytestmark = pytest.mark.gpu
A, PRE, YLD, PH = po.TT_ANALYSIS, po.TT_PREEMPTED, po.TT_YIELD, po.TT_PHASE
ALL_FLAGS = [0, po.TRIM, po.ABORT_ON_MISSING, po.TRIM | po.ABORT_ON_MISSING]


@pytest.fixture(scope="module")
def slicer():
    s = gs.IntervalSlicer(0)
    yield s
    s.close()


def fields(a):
    return np.stack([a["tstamp"], a["duration"], a["stack_id"], a["swin_type"].astype(np.uint64),
                     a["swout_type"].astype(np.uint64)], axis=1) if len(a) else np.zeros((0, 5), np.uint64)


def run_and_compare(slicer, unit_slices, unit_rows, stop_ts, flags, marks=None, expect_path=None):
    """unit_slices / unit_rows: per-unit arrays.  Checks every unit against the restatement."""
    n = len(unit_slices)
    marks = np.zeros(n, np.uint64) if marks is None else np.array(marks, np.uint64)
    marks0 = marks.copy()
    stride = unit_rows[0].shape[1]
    sb = np.concatenate([[0], np.cumsum([len(s) for s in unit_slices])]).astype(np.uint64)
    rb = np.concatenate([[0], np.cumsum([len(r) for r in unit_rows])]).astype(np.uint64)
    slices = np.concatenate(unit_slices) if n else np.zeros(0, gs.SLICE_DTYPE)
    rows = np.concatenate(unit_rows).reshape(-1, stride)
    units = slicer.release(slices, sb, rows, rb, stop_ts, flags, marks)
    assert np.array_equal(units, slicer.units())
    for u in range(n):
        exp = po.islice_release("port", unit_slices[u], unit_rows[u], stop_ts, flags & 3, int(marks0[u]))
        got = units[u]
        where = f"unit {u} flags {flags} path {got['exact_path']}"
        if exp["rc"] < 0:
            assert got["status"] == exp["rc"], where
            assert got["err_interval"] == exp["err_interval"], where
            assert marks[u] == marks0[u]
            continue
        assert got["status"] == 0, where
        assert got["n_intervals"] == exp["n_intervals"], where
        assert got["mark"] == exp["mark"] and marks[u] == exp["mark"], where
        counts, out = slicer.get(u)
        assert np.array_equal(counts, exp["counts"]), where
        assert got["n_slices"] == len(exp["slices"]), where
        assert np.array_equal(fields(out), fields(exp["slices"])), where
        if expect_path is not None:
            assert got["exact_path"] == expect_path, where
    return units


def test_reference_vectors(slicer):
    """hbt/src/tagstack/tests/IntervalSlicerTest.cpp as one-row windows."""
    s12 = po.make_slices([(0, 99, 1, A, A), (99, 2, 1, A, A)])
    cases = [
        ([], 0, 0, 100, 0), ([(10, 20, 1, A, A)], 0, 0, 100, 0), ([(10, 20, 1, A, A)], 0, 0, 100, 3),
        ([(10, 90, 1, A, A)], 0, 0, 100, 3), ([(0, 99, 1, A, A)], 0, 0, 100, 3), ([(0, 99, 1, A, A)], 0, 0, 100, 1),
    ]
    for sl, mark, t, d, flags in cases:
        run_and_compare(slicer, [po.make_slices(sl)], [np.array([[t, d]], np.uint64)], 2**63, flags, [mark])
    for mark, t, d in [(0, 0, 100), (1, 0, 101), (0, 99, 101), (0, 100, 102), (0, 100, 1)]:
        u = run_and_compare(slicer, [s12], [np.array([[t, d]], np.uint64)], 2**63, 3, [mark])
    assert u[0]["mark"] == 2 and u[0]["n_slices"] == 1          # SingleConsumerMultiple's last query
    s = po.make_slices([(99, 2, 33, PRE, PH), (101, 1, 1, A, A), (200, 10, 1, PH, PH)])
    for mark, t, d, flags in [(0, 100, 200, 3), (0, 100, 200, 1), (1, 100, 105, 1), (2, 100, 105, 1), (2, 100, 210, 1)]:
        run_and_compare(slicer, [s], [np.array([[t, d]], np.uint64)], 2**63, flags, [mark])


def test_boundary_slice_consumed_by_previous_row(slicer):
    """A slice ending exactly on a row boundary: taken with length 0 by the next row unless the
    previous row consumed it -- the one place where rows depend on each other."""
    s = po.make_slices([(0, 100, 5, YLD, PRE), (100, 50, 6, PRE, YLD), (150, 50, 7, YLD, YLD), (200, 100, 8, PH, PH)])
    rows = np.array([[0, 100], [100, 50], [150, 50], [200, 50], [250, 50]], np.uint64)
    for flags in ALL_FLAGS:
        run_and_compare(slicer, [s], [rows], 2**63, flags, expect_path=0)
        run_and_compare(slicer, [s], [rows[1:]], 2**63, flags, expect_path=0)        # first row finds it unconsumed
        run_and_compare(slicer, [s], [rows[1:]], 2**63, flags, [1], expect_path=0)   # mark already past it
        run_and_compare(slicer, [s], [rows], 2**63, flags | gs.FORCE_EXACT, expect_path=1)
    # chain of aligned rows where an aborted row leaves the boundary slice unconsumed
    s = po.make_slices([(0, 10, 1, YLD, YLD), (10, 10, 2, YLD, YLD), (25, 5, 3, YLD, YLD), (30, 10, 4, YLD, YLD)])
    rows = np.array([[0, 10], [10, 10], [20, 10], [30, 10]], np.uint64)
    for flags in ALL_FLAGS:
        run_and_compare(slicer, [s], [rows], 2**63, flags, expect_path=0)


@pytest.mark.parametrize("flags", ALL_FLAGS)
def test_regular_units_parallel_path(slicer, flags):
    r = np.random.default_rng(300 + flags)
    for trial in range(6):
        n_units = int(r.integers(1, 9))
        us, ur, marks = [], [], []
        stride = int(r.integers(2, 11))
        for u in range(n_units):
            ns = int(r.choice([0, 1, 7, 300, 2000, 5000]))
            s = slicegen.gen_slices(r, ns, hole_prob=float(r.choice([0.0, 0.02, 0.3])), max_dur=int(r.choice([3, 50, 1000])))
            rows = slicegen.gen_rows(r, s, int(r.choice([0, 1, 40, 511, 512, 513, 1500])), stride=stride,
                                     align_prob=float(r.choice([0.0, 0.3, 1.0])), gap_prob=float(r.choice([0.0, 0.05])))
            us.append(s)
            ur.append(rows)
            marks.append(0)
        allr = np.concatenate(ur)
        stop = 2**63 if trial % 2 == 0 or len(allr) == 0 else int(np.median(allr[:, 0]))
        run_and_compare(slicer, us, ur, stop, flags, marks)
        # the same through the serial walk
        run_and_compare(slicer, us, ur, stop, flags | gs.FORCE_EXACT, marks, expect_path=1)


def test_regular_units_take_the_parallel_path(slicer):
    r = np.random.default_rng(5)
    s = slicegen.gen_slices(r, 4000, hole_prob=0.05)
    rows = slicegen.gen_rows(r, s, 1300, stride=10, align_prob=0.05)   # chains of aligned rows stay short
    for flags in (po.TRIM | po.ABORT_ON_MISSING, po.TRIM, po.ABORT_ON_MISSING):
        u = run_and_compare(slicer, [s, s[:100], s], [rows, rows, rows[:1]], 2**63, flags, expect_path=0)
        assert u["n_slices"].sum() > 0 or flags == po.ABORT_ON_MISSING


@pytest.mark.parametrize("skind", ["zero_dur", "overlap", "unsorted", "shuffled", "ok"])
@pytest.mark.parametrize("rkind", ["overlap", "backwards", "zero_dur", "repeat", "ok"])
def test_hostile_input(slicer, skind, rkind):
    """Anything goes: out-of-order slices, rows going back in time, zero lengths, marks ahead."""
    r = np.random.default_rng(zlib.crc32((skind + rkind).encode()))
    for flags in ALL_FLAGS:
        us, ur, marks = [], [], []
        for u in range(6):
            s = slicegen.hostile_slices(r, int(r.integers(1, 400)), skind)
            ur.append(slicegen.hostile_rows(r, np.sort(s, order="tstamp"), int(r.integers(1, 700)), rkind))
            us.append(s)
            marks.append(int(r.choice([0, 0, 3, 1000])))
        run_and_compare(slicer, us, ur, 2**63, flags, marks)


def test_wrapping_times(slicer):
    """u64 wrap in slice ends and interval ends (TimeStamp is uint64_t; the reference just adds)."""
    top = 2**64
    s = po.make_slices([(top - 300, 100, 1, YLD, YLD), (top - 200, 150, 2, YLD, YLD), (top - 50, 100, 3, YLD, YLD)])
    rows = np.array([[top - 300, 100], [top - 200, 100], [top - 100, 200]], np.uint64)
    for flags in ALL_FLAGS:
        run_and_compare(slicer, [s], [rows], 2**64 - 1, flags)


def test_many_units_one_call(slicer):
    """192 comp units (BASELINE config 2's CPU count), mixed regular and hostile."""
    r = np.random.default_rng(11)
    us, ur = [], []
    for u in range(192):
        if u % 37 == 5:
            s = slicegen.hostile_slices(r, 200, "overlap")
        else:
            s = slicegen.gen_slices(r, int(r.integers(50, 1500)), hole_prob=0.03)
        us.append(s)
        ur.append(slicegen.gen_rows(r, np.sort(s, order="tstamp"), int(r.integers(10, 700)), stride=10, align_prob=0.2))
    units = run_and_compare(slicer, us, ur, 2**63, po.TRIM | po.ABORT_ON_MISSING)
    assert units["exact_path"].sum() < 20


def test_device_input_and_view(slicer):
    import torch
    r = np.random.default_rng(3)
    us = [slicegen.gen_slices(r, 3000, hole_prob=0.02) for _ in range(4)]
    ur = [slicegen.gen_rows(r, s, 900, stride=6, align_prob=0.3) for s in us]
    sb = np.concatenate([[0], np.cumsum([len(s) for s in us])]).astype(np.uint64)
    rb = np.concatenate([[0], np.cumsum([len(x) for x in ur])]).astype(np.uint64)
    d_s = torch.from_numpy(np.concatenate(us).view(np.uint8)).cuda()
    d_r = torch.from_numpy(np.concatenate(ur).view(np.int64)).cuda()
    torch.cuda.synchronize()
    marks = np.zeros(4, np.uint64)
    flags = po.TRIM | po.ABORT_ON_MISSING
    units = slicer.release_device(d_s.data_ptr(), sb, d_r.data_ptr(), 6, rb, 2**63, flags, marks)
    view = slicer.device_view()
    assert view.n_units == 4 and view.slices and view.interval_count
    for u in range(4):
        exp = po.islice_release("port", us[u], ur[u], 2**63, flags, 0)
        counts, out = slicer.get(u)
        assert np.array_equal(counts, exp["counts"]) and np.array_equal(fields(out), fields(exp["slices"]))
        assert units[u]["mark"] == exp["mark"]
    assert slicer.last_ms > 0


# ---- row f1: frequencies + the interval x bin matrix over the released slices ----------------------
def bin_compare(slicer, us, ur, flags, min_int, min_dur, selected=None, zero_some=False):
    n = len(us)
    stride = ur[0].shape[1]
    sb = np.concatenate([[0], np.cumsum([len(s) for s in us])]).astype(np.uint64)
    rb = np.concatenate([[0], np.cumsum([len(r) for r in ur])]).astype(np.uint64)
    marks = np.zeros(n, np.uint64)
    units = slicer.release(np.concatenate(us), sb, np.concatenate(ur).reshape(-1, stride), rb, 2**63, flags, marks)
    assert (units["status"] == 0).all()
    sel = range(n) if selected is None else [u for u in range(n) if selected[u]]
    uc, uo = [], []
    for u in sel:
        counts, out = slicer.get(u)
        uc.append(counts)
        uo.append(out)
    got = slicer.bin(min_int, min_dur, selected=selected)
    exp = po.bin_durations("port", uc, uo, min_int, min_dur)      # same numbering: ascending stack id
    for k in ("stack_id", "duration", "num_obs", "num_intervals", "bid"):
        assert np.array_equal(got["freqs"][k], exp["freqs"][k]), k
    assert got["num_rows"] == exp["num_rows"] and got["num_bins"] == exp["num_bins"]
    assert np.array_equal(got["durations"], exp["durations"])
    return got


@pytest.mark.parametrize("min_int,min_dur", [(0, 0), (1, 1), (3, 200), (1, 10**12)])
def test_bin_matrix(slicer, min_int, min_dur):
    r = np.random.default_rng(40 + min_int)
    for trial in range(4):
        n_units = int(r.integers(1, 7))
        us = [slicegen.gen_slices(r, int(r.choice([0, 5, 600, 3000])), hole_prob=0.1, n_stacks=int(r.choice([2, 40, 2000])))
              for _ in range(n_units)]
        for s in us:
            s["stack_id"] += np.uint64(r.choice([0, 1 << 45]))
            s["stack_id"][r.random(len(s)) < 0.01] = np.uint64(2**64 - 1)        # kInvalidTagStackId as a key
        ur = [slicegen.gen_rows(r, s, int(r.choice([0, 3, 700, 1100])), stride=4, align_prob=0.3) for s in us]
        got = bin_compare(slicer, us, ur, int(r.choice([1, 3])), min_int, min_dur)
        sel = (r.random(n_units) < 0.6).astype(np.uint8)
        bin_compare(slicer, us, ur, 3, min_int, min_dur, selected=sel)
    assert got["num_bins"] >= 1


def test_bin_zero_length_slices_and_limits(slicer):
    from dynolog_b200.hbtgpu import HbtGpuError
    Yd = po.TT_YIELD
    # an aligned chain leaves zero-length first slices in the intervals (trim without abort)
    s = po.make_slices([(0, 100, 5, Yd, Yd), (100, 50, 6, Yd, Yd), (150, 50, 5, Yd, Yd), (300, 10, 9, Yd, Yd)])
    rows = np.array([[100, 50], [150, 50], [200, 50], [290, 30]], np.uint64)
    got = bin_compare(slicer, [s], [rows], po.TRIM, 1, 1)
    assert got["rows_skipped"] == 0 and got["num_rows"] == 4
    with pytest.raises(HbtGpuError):
        slicer.bin(0, 0, max_stack_ids=1)              # three distinct stack ids do not fit
    with pytest.raises(HbtGpuError):
        slicer.bin(0, 0, max_matrix_bytes=8)           # 4 rows x 4 bins do not fit


def test_bin_over_two_ranks(slicer):
    """Units split between two contexts (two GPUs' worth): frequencies merged, bins agreed
    (shard.merge_bin_freqs), each side builds its matrix with hbtgpu_islice_bin_with, the sum is the
    matrix of one context holding all units."""
    from dynolog_b200 import shard
    r = np.random.default_rng(91)
    us = [slicegen.gen_slices(r, int(r.integers(100, 2000)), hole_prob=0.05, n_stacks=int(r.choice([4, 90]))) for _ in range(7)]
    ur = [slicegen.gen_rows(r, s, int(r.integers(20, 500)), stride=3, align_prob=0.2) for s in us]

    def release(ctx, lo, hi):
        sb = np.concatenate([[0], np.cumsum([len(s) for s in us[lo:hi]])]).astype(np.uint64)
        rb = np.concatenate([[0], np.cumsum([len(x) for x in ur[lo:hi]])]).astype(np.uint64)
        units = ctx.release(np.concatenate(us[lo:hi]), sb, np.concatenate(ur[lo:hi]), rb, 2**63, 3, np.zeros(hi - lo, np.uint64))
        assert (units["status"] == 0).all()
        return int(units["n_intervals"].sum())

    other = gs.IntervalSlicer(0)
    try:
        for min_int, min_dur in [(0, 0), (2, 100), (4, 10**7)]:
            release(slicer, 0, 7)
            whole = slicer.bin(min_int, min_dur)
            rows_a, rows_b = release(slicer, 0, 3), release(other, 3, 7)
            fa, fb = slicer.bin(0, 0)["freqs"], other.bin(0, 0)["freqs"]
            merged, num_bins = shard.merge_bin_freqs([fa, fb], min_int, min_dur)
            for k in ("stack_id", "duration", "num_obs", "num_intervals", "bid"):
                assert np.array_equal(merged[k], whole["freqs"][k]), k
            numbering = (merged["stack_id"], merged["bid"], num_bins, rows_a + rows_b)
            ma = slicer.bin(0, 0, numbering=numbering)
            mb = other.bin(0, 0, numbering=numbering)
            assert ma["num_bins"] == whole["num_bins"] and ma["num_rows"] == whole["num_rows"]
            assert np.array_equal(ma["durations"] + mb["durations"], whole["durations"])
        from dynolog_b200.hbtgpu import HbtGpuError
        with pytest.raises(HbtGpuError):
            slicer.bin(0, 0, numbering=(merged["stack_id"][1:], merged["bid"][1:], num_bins, rows_a + rows_b))
    finally:
        other.close()
