"""The stages in one line: thread-switch ring images -> tagstack events (GPU, f2),
count-sample ring images -> count rows (GPU, the hot path), events -> slices (GPU, f3 Slicer side;
checked here against the reference's own Slicer from oracle/_ref), count rows x slices ->
per-interval slices (GPU, f3) -> frequencies + interval x bin matrix (GPU, f1).
Checks that the formats connect (16-byte events, n+2-word rows, 32-byte slices pass from stage to
stage untouched) and that the GPU stages agree with the CPU checkers on realistic data rather
than on hand-made vectors."""
import numpy as np
import pytest

import gpu_util
from dynolog_b200 import hbtgpu_cs, hbtgpu_slices as gs, synth
from oracle import pyoracle as po

pytestmark = pytest.mark.gpu


@pytest.mark.skipif(not po.slices_available("reference"), reason="oracle/_ref/libhbtref_slices.so not built")
def test_images_to_bins():
    n_cpus = 6
    # thread switches: ~1500 switch pairs per cpu, 20 us apart, a few LOST records (write errors)
    cs_streams = synth.cs_capture(n_cpus, 96 * 1024, n_tids=64, life_ppm=20000, comm_ppm=0, lost_ppm=2000,
                                  throttle_ppm=0, seed=3)
    dec = hbtgpu_cs.decode(cs_streams, tid_level=0)
    events = dec["events"]
    assert all(len(e) > 1000 for e in events)
    for cu, e in enumerate(events):
        assert e.dtype.itemsize == 16 and (e["comp_unit_id"] == cu).all()
    # the Combinator tree's order: by (tstamp, comp unit), each cpu's own order kept
    merged = np.concatenate(events)
    order = np.lexsort((merged["comp_unit_id"], merged["tstamp"]))
    sl = po.slicer_run("reference", merged[order], n_cpus)
    assert sl["rc"] == 0 and sl["stats"]["num_slices_completed"] > 1000 * n_cpus
    assert sl["stats"]["num_error_start"] > 0               # the LOST records made it through as resets
    gsl = gs.Slicer(0)
    try:
        got_sl = gsl.run(events)
    finally:
        gsl.close()
    assert got_sl["stats"] == sl["stats"]
    for u in range(n_cpus):
        assert po.slices_as_tuples(got_sl["slices"][u]) == po.slices_as_tuples(sl["slices"][u]), u
    slices = got_sl["slices"]

    # count samples over the same stretch of time: 100 us apart, 4-event group
    spec, streams = synth.capture(1, n_cpus, n_samples=280)
    params = po.make_params(spec.n_events)
    got, _ = gpu_util.gpu_decode(params, streams)
    rows = [g["rows"] for g in got]
    assert all(r.shape[1] == spec.n_events + 2 and len(r) > 250 for r in rows)
    t_lo = min(int(s["tstamp"][0]) for s in slices)
    t_hi = max(int(s["tstamp"][-1]) for s in slices)
    assert int(rows[0][5, 0]) > t_lo and int(rows[0][5, 0]) < t_hi, "samples and switches must overlap in time"

    # rows x slices on the GPU, TraceMonitor's flags
    flags = po.TRIM | po.ABORT_ON_MISSING
    stop_ts = int(rows[0][-10, 0])
    isl = gs.IntervalSlicer(0)
    try:
        sb = np.concatenate([[0], np.cumsum([len(s) for s in slices])]).astype(np.uint64)
        rb = np.concatenate([[0], np.cumsum([len(r) for r in rows])]).astype(np.uint64)
        marks = np.zeros(n_cpus, np.uint64)
        units = isl.release(np.concatenate(slices), sb, np.concatenate(rows), rb, stop_ts, flags, marks)
        assert (units["status"] == 0).all() and (units["exact_path"] == 0).all()
        uc, uo = [], []
        for u in range(n_cpus):
            exp = po.islice_release("port", slices[u], rows[u], stop_ts, flags, 0)
            counts, out = isl.get(u)
            assert units[u]["n_intervals"] == exp["n_intervals"] and marks[u] == exp["mark"]
            assert np.array_equal(counts, exp["counts"])
            for k in ("tstamp", "duration", "stack_id", "swin_type", "swout_type"):
                assert np.array_equal(out[k], exp["slices"][k]), (u, k)
            uc.append(counts)
            uo.append(out)
        covered = sum(int((c > 0).sum()) for c in uc)
        assert covered > 100 * n_cpus, "most count intervals should be covered by slices"
        # frequencies and the matrix
        binned = isl.bin(2, 1000)
        exp = po.bin_durations("port", uc, uo, 2, 1000)
        for k in ("stack_id", "duration", "num_obs", "num_intervals", "bid"):
            assert np.array_equal(binned["freqs"][k], exp["freqs"][k]), k
        assert np.array_equal(binned["durations"], exp["durations"])
        assert binned["num_bins"] > 5
        # every covered interval's durations add up to its length (trim + abort: exact cover)
        row_sums = binned["durations"].sum(axis=1)
        assert row_sums[: int((uc[0] > 0).sum())].all()
    finally:
        isl.close()


def test_rows_and_slices_meet_on_the_device():
    """Count rows (hbtgpu_rows_device) and slices (hbtgpu_slicer_slices_device) go into
    hbtgpu_islice_release as device pointers: nothing but summaries crosses PCIe."""
    from dynolog_b200 import hbtgpu
    n_cpus = 5
    dec = hbtgpu_cs.decode(synth.cs_capture(n_cpus, 128 * 1024, n_tids=50, life_ppm=10000, comm_ppm=0, lost_ppm=500,
                                            throttle_ppm=0, seed=9), tid_level=0)
    spec, streams = synth.capture(1, n_cpus, n_samples=300)
    total = sum(len(s) for s in streams)
    ctx = hbtgpu.Context(n_events=spec.n_events, max_cpus=n_cpus, arena_bytes=total + 4096 * (n_cpus + 8) + (1 << 16))
    gsl, isl = gs.Slicer(0), gs.IntervalSlicer(0)
    try:
        for cpu, s in enumerate(streams):
            ctx.submit_image(cpu, s)
        ctx.run()
        ctx.sync()
        d_rows, row_begin, stride = ctx.rows_device()
        assert stride == spec.n_events + 2 and row_begin[-1] > 250 * n_cpus
        got_sl = gsl.run(dec["events"])
        d_slices, _ = gsl.slices_device()
        sb = np.concatenate([[0], np.cumsum(got_sl["n_slices"])]).astype(np.uint64)
        marks = np.zeros(n_cpus, np.uint64)
        flags = po.TRIM | po.ABORT_ON_MISSING
        units = isl.release_device(d_slices, sb, d_rows, stride, row_begin, 2**63, flags, marks)
        assert (units["status"] == 0).all()
        for u in range(n_cpus):
            rows = ctx.rows(u)
            assert len(rows) == row_begin[u + 1] - row_begin[u]
            exp = po.islice_release("port", got_sl["slices"][u], rows, 2**63, flags, 0)
            counts, out = isl.get(u)
            assert np.array_equal(counts, exp["counts"]) and po.slices_as_tuples(out) == po.slices_as_tuples(exp["slices"])
            assert marks[u] == exp["mark"]
        assert units["n_slices"].sum() > 500
        # the compact copy of the rows' times (hbtgpu_row_times_device) gives the same release
        from cuda.bindings import runtime as cudart
        d_times = ctx.row_times_device()
        n_rows = int(row_begin[-1])
        times = np.zeros((n_rows, 2), np.uint64)
        err, = cudart.cudaMemcpy(times.ctypes.data, d_times, 16 * n_rows, cudart.cudaMemcpyKind.cudaMemcpyDeviceToHost)
        assert int(err) == 0
        for u in range(n_cpus):
            assert np.array_equal(times[int(row_begin[u]):int(row_begin[u + 1])], ctx.rows(u)[:, :2])
        marks2 = np.zeros(n_cpus, np.uint64)
        units2 = isl.release_device(d_slices, sb, d_times, 2, row_begin, 2**63, flags, marks2)
        assert np.array_equal(units2["status"], units["status"]) and np.array_equal(units2["n_slices"], units["n_slices"])
        assert np.array_equal(marks2, marks)
        for u in range(n_cpus):
            exp = po.islice_release("port", got_sl["slices"][u], ctx.rows(u), 2**63, flags, 0)
            counts, out = isl.get(u)
            assert np.array_equal(counts, exp["counts"]) and po.slices_as_tuples(out) == po.slices_as_tuples(exp["slices"])
    finally:
        gsl.close()
        isl.close()
        ctx.close()
