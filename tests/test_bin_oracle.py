"""Stack-id frequencies and the interval x bin duration matrix: the C
restatement pinned against the REAL MonData::stackIdFreqs + binmap::TagStackIdBinner
(hbt/src/mon/MonData.h, compiled into oracle/_ref/libhbtref_slices.so).  The reference numbers
the bins in the iteration order of a std::unordered_map; everything is compared through the
stack ids.  CPU only."""
import numpy as np
import pytest

import slicegen
from oracle import pyoracle as po

needs_ref = pytest.mark.skipif(not po.slices_available("reference"),
                               reason="oracle/_ref/libhbtref_slices.so not built (needs /root/reference)")
Y = po.TT_YIELD


def same_binning(a, b):
    """a, b: results of po.bin_durations with possibly different bin numbering."""
    fa, fb = a["freqs"], b["freqs"]
    for k in ("stack_id", "duration", "num_obs", "num_intervals"):
        assert np.array_equal(fa[k], fb[k]), k
    assert np.array_equal(fa["bid"] == 0, fb["bid"] == 0)              # the same stack ids are ignored
    assert a["num_rows"] == b["num_rows"] and a["num_bins"] == b["num_bins"]
    perm = np.zeros(a["num_bins"], np.int64)                            # a's bin -> b's bin
    perm[fa["bid"]] = fb["bid"]
    assert sorted(perm[1:]) == list(range(1, a["num_bins"]))            # own bins map one to one
    assert np.array_equal(a["durations"], b["durations"][:, perm])


def random_units(r, n_units, zero_dur=0.1):
    uc, us = [], []
    for u in range(n_units):
        res = None
        s = slicegen.gen_slices(r, int(r.integers(0, 400)), hole_prob=0.1, n_stacks=int(r.choice([3, 20, 300])))
        s["stack_id"] += np.uint64(r.choice([0, 0, 1 << 40]))
        rows = slicegen.gen_rows(r, s, int(r.integers(0, 150)), align_prob=0.3)
        res = po.islice_release("port", s, rows, 2**63, int(r.choice([1, 3])))
        out = res["slices"]
        out["duration"][r.random(len(out)) < zero_dur] = 0
        uc.append(res["counts"])
        us.append(out)
    return uc, us


def test_known_answers():
    # unit 0: interval 0 = {sid 5: 10 + 20, sid 7: 5}, interval 1 empty, interval 2 = {sid 5: 0-length only},
    #         interval 3 = {sid 9: 4};  unit 1: interval 0 = {sid 7: 8}
    u0 = po.make_slices([(0, 10, 5, Y, Y), (10, 5, 7, Y, Y), (15, 20, 5, Y, Y), (50, 0, 5, Y, Y), (60, 4, 9, Y, Y)])
    u1 = po.make_slices([(0, 8, 7, Y, Y)])
    r = po.bin_durations("port", [[3, 0, 1, 1], [1]], [u0, u1], 2, 1)
    f = r["freqs"]
    assert list(f["stack_id"]) == [5, 7, 9]
    assert list(f["duration"]) == [30, 13, 4] and list(f["num_obs"]) == [3, 2, 1] and list(f["num_intervals"]) == [2, 2, 1]
    assert list(f["bid"]) == [1, 2, 0] and r["num_bins"] == 3 and r["num_rows"] == 5
    d = r["durations"]
    assert list(d[0]) == [0, 30, 5 + 8]          # both units start at row 0 (MonData.h:563)
    assert list(d[1]) == [4, 0, 0]               # the 0-length interval did not advance the row (:592-597)
    assert not d[2:].any()


@needs_ref
@pytest.mark.parametrize("min_int,min_dur", [(0, 0), (1, 1), (2, 50), (5, 0), (1, 10**9)])
def test_port_matches_reference(min_int, min_dur):
    r = np.random.default_rng(17 + min_int)
    for trial in range(12):
        uc, us = random_units(r, int(r.integers(1, 6)))
        same_binning(po.bin_durations("reference", uc, us, min_int, min_dur),
                     po.bin_durations("port", uc, us, min_int, min_dur))


@needs_ref
def test_reference_known_answers():
    u0 = po.make_slices([(0, 10, 5, Y, Y), (10, 5, 7, Y, Y), (15, 20, 5, Y, Y), (50, 0, 5, Y, Y), (60, 4, 9, Y, Y)])
    u1 = po.make_slices([(0, 8, 7, Y, Y)])
    same_binning(po.bin_durations("reference", [[3, 0, 1, 1], [1]], [u0, u1], 2, 1),
                 po.bin_durations("port", [[3, 0, 1, 1], [1]], [u0, u1], 2, 1))
