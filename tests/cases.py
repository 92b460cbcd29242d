"""Shared helpers: load golden fixtures, compare decode results."""
import glob
import os

import numpy as np

from oracle import pyoracle as po

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def golden_stream_names():
    return sorted(os.path.basename(p)[len("stream_"):-len(".npz")]
                  for p in glob.glob(os.path.join(GOLDEN, "stream_*.npz")))


def load_golden_stream(name):
    z = np.load(os.path.join(GOLDEN, f"stream_{name}.npz"))
    pv = [int(x) for x in z["params"]]
    params = po.make_params(pv[0], n_buckets=pv[1], tsc_mask=pv[2], time_shift=pv[3],
                            time_mult=pv[4], time_zero=pv[5], t0=pv[6], bucket_ns=pv[7])
    ci = [int(x) for x in z["carry_in"]]
    carry = po.make_carry(valid=bool(ci[0]), tstamp=ci[1], te=ci[2], tr=ci[3], last=ci[4:])
    exp = {
        "counters": dict(zip(po.COUNTER_KEYS, [int(x) for x in z["counters"]])),
        "records": (z["rec_off"], z["rec_type"], z["rec_size"]),
        "rows": z["rows"], "buckets": z["buckets"], "scaled": z["scaled"],
        "carry_out": [int(x) for x in z["carry_out"]],
    }
    return z["data"], params, carry, exp


def carry_list(c, n):
    return [int(c.valid), int(c.tstamp), int(c.te), int(c.tr)] + [int(c.last[i]) for i in range(n)]


def assert_decode_equal(got, exp, n_events, what=""):
    assert got["counters"] == exp["counters"], f"{what}: counters differ"
    for a, b, nm in zip(got["records"], exp["records"], ("off", "type", "size")):
        assert np.array_equal(a, b), f"{what}: record table {nm} differs"
    assert np.array_equal(got["rows"], exp["rows"]), f"{what}: rows differ"
    assert np.array_equal(got["buckets"], exp["buckets"]), f"{what}: bucket sums differ"
    gc = got["carry_out"]
    gc = gc if isinstance(gc, list) else carry_list(gc, n_events)
    ec = exp["carry_out"]
    ec = ec if isinstance(ec, list) else carry_list(ec, n_events)
    if gc[0] == 0 and ec[0] == 0:
        gc[1] = ec[1] = 0  # tstamp is meaningless when the state is invalid
    assert gc == ec, f"{what}: carry-out differs"
