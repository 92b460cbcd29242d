"""Synthetic tagstack slices and count intervals for the f3 tests (host only, numpy).

A comp unit's slices are what tagstack::slice::Slicer writes for one CPU: time ordered, never
overlapping, durations > 0 (hbt/src/tagstack/Slicer.h:529-558).  Count intervals are the rows
[tstamp, duration, ...] of CpuCountSampleGenerator (contiguous except after LOST gaps).
The `hostile` styles break those properties on purpose."""
import numpy as np

from oracle.pyoracle import SLICE_DTYPE


def gen_slices(rng, n, t0=1000, max_dur=50, hole_prob=0.1, n_stacks=7):
    s = np.zeros(n, SLICE_DTYPE)
    dur = rng.integers(1, max_dur + 1, n).astype(np.uint64)
    gap = np.where(rng.random(n) < hole_prob, rng.integers(1, max_dur + 1, n), 0).astype(np.uint64)
    start = np.uint64(t0) + np.cumsum(gap) + np.concatenate([[0], np.cumsum(dur)[:-1]]).astype(np.uint64)
    s["tstamp"] = start
    s["duration"] = dur
    s["stack_id"] = rng.integers(0, n_stacks, n)
    s["swin_type"] = rng.integers(2, 5, n)
    s["swout_type"] = rng.integers(2, 5, n)
    return s


def gen_rows(rng, slices, n, stride=2, align_prob=0.3, gap_prob=0.05, lead=20, mean_dur=None):
    """n contiguous intervals over the slices' time range; with probability align_prob an
    interval boundary is snapped onto a slice end (the case where the previous interval has
    already consumed the slice that ends exactly at this interval's start)."""
    rows = np.zeros((n, stride), np.uint64)
    if n == 0:
        return rows
    if len(slices):
        lo = int(slices["tstamp"][0]) - lead
        hi = int(slices["tstamp"][-1] + slices["duration"][-1]) + lead
        ends = (slices["tstamp"] + slices["duration"]).astype(np.uint64)
    else:
        lo, hi, ends = 100, 1000, np.zeros(0, np.uint64)
    lo = max(lo, 1)
    mean = mean_dur if mean_dur is not None else max(2, (hi - lo) // max(1, n))
    t = lo
    for k in range(n):
        d = int(rng.integers(1, 2 * mean + 1))
        if len(ends) and rng.random() < align_prob:
            j = int(np.searchsorted(ends, t + d))
            if j < len(ends) and int(ends[j]) > t:
                d = int(ends[j]) - t
        rows[k, 0] = t
        rows[k, 1] = d
        if stride > 2:
            rows[k, 2:] = rng.integers(0, 1 << 40, stride - 2)
        t += d
        if rng.random() < gap_prob:
            t += int(rng.integers(1, mean + 1))
    return rows


def hostile_slices(rng, n, kind):
    s = gen_slices(rng, n, hole_prob=0.2)
    if n == 0:
        return s
    if kind == "zero_dur":
        s["duration"][rng.integers(0, n, max(1, n // 10))] = 0
    elif kind == "overlap":
        idx = rng.integers(0, n, max(1, n // 10))
        s["duration"][idx] += rng.integers(1, 200, len(idx)).astype(np.uint64)
    elif kind == "unsorted":
        i, j = rng.integers(0, n, 2)
        s[[i, j]] = s[[j, i]]
    elif kind == "shuffled":
        rng.shuffle(s)
    return s


def hostile_rows(rng, slices, n, kind):
    rows = gen_rows(rng, slices, n)
    if n == 0:
        return rows
    if kind == "overlap":
        idx = rng.integers(0, n, max(1, n // 8))
        rows[idx, 1] += rng.integers(1, 100, len(idx)).astype(np.uint64)
    elif kind == "backwards":
        i, j = rng.integers(0, n, 2)
        rows[[i, j]] = rows[[j, i]]
    elif kind == "zero_dur":
        rows[rng.integers(n // 2, n), 1] = 0
    elif kind == "repeat":
        i = int(rng.integers(0, n))
        rows = np.concatenate([rows[: i + 1], rows[i:]])
    return rows
