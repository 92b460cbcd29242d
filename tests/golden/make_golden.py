"""Generate tests/golden/*.npz from the REAL reference (oracle/_ref/libhbtref.so).

Run in the build container (where /root/reference exists and `make -C oracle ref`
has been run):   python tests/golden/make_golden.py

Each fixture stores the input ring image (bytes), the parameters/carry, and what
the reference's own consume()/CpuCountSampleGenerator/GroupReadValues/reducers
produced for it.  The fixtures travel to the GPU box; /root/reference does not.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import ringgen as rg  # noqa: E402
from oracle import pyoracle as po  # noqa: E402

STREAM_CASES = {
    # name: (kwargs for random_stream, n_events, n_records, carry_valid, tsc_mask)
    "plain4": (dict(), 4, 600, True, 0),
    "lost_throttle4": (dict(p_lost=0.02, p_throttle=0.01), 4, 800, True, 0),
    "mixed8": (dict(p_lost=0.01, p_throttle=0.01, p_other=0.03, p_read=0.02, mux=True), 8, 900,
               False, 0),
    "forged4": (dict(forge=0.05, p_lost=0.005), 4, 700, True, 0),
    "tsc8": (dict(p_lost=0.01, mux=True), 8, 500, True, 0b00100100),
    "wrap2": (dict(big_values=True, p_lost=0.01), 2, 400, False, 0),
}


def stream_case(name):
    kw, n, nrec, carry_valid, tsc_mask = STREAM_CASES[name]
    seed = 0xD1A0_0000_0000_0000 | (sum(map(ord, name)) << 8)
    data, info = rg.random_stream(seed, nrec, n, **kw)
    t0 = 1_000_000_000
    bucket_ns = 1_000_000
    n_buckets = int((info["t_end"] - t0) // bucket_ns) + 2
    if name == "mixed8":
        n_buckets -= 10  # force some rows out of range
    params = po.make_params(n, n_buckets=n_buckets, t0=t0, bucket_ns=bucket_ns,
                            tsc_mask=tsc_mask, time_shift=10, time_mult=341, time_zero=12345)
    carry = po.make_carry(valid=carry_valid, tstamp=t0 - 5, te=0, tr=0, last=[0] * n)
    return data, params, carry


def main():
    ref = po.load("reference")
    for name in STREAM_CASES:
        data, params, carry = stream_case(name)
        r = ref.decode_stream(params, data, carry)
        scaled = ref.scale_buckets(params, r["buckets"])
        co = r["carry_out"]
        np.savez_compressed(
            os.path.join(HERE, f"stream_{name}.npz"),
            data=data,
            params=np.array([params.n_events, params.n_buckets, params.tsc_mask,
                             params.time_shift, params.time_mult, params.time_zero, params.t0,
                             params.bucket_ns], np.uint64),
            carry_in=np.array([carry.valid, carry.tstamp, carry.te, carry.tr] +
                              list(carry.last)[:params.n_events], np.uint64),
            counters=np.array([r["counters"][k] for k in po.COUNTER_KEYS], np.int64),
            rec_off=r["records"][0], rec_type=r["records"][1], rec_size=r["records"][2],
            rows=r["rows"], buckets=r["buckets"], scaled=scaled,
            carry_out=np.array([co.valid, co.tstamp, co.te, co.tr] +
                               list(co.last)[:params.n_events], np.uint64))
        print(name, len(data), "bytes", r["counters"]["num_records"], "records",
              r["counters"]["num_rows"], "rows")

    # elementwise math (A6, A8, A9, A10, A13)
    g = rg.SplitMix64(0xD1A0_0000_0000_0000 | 0xE1E)
    N = 4096
    cnt = np.array([g.next() >> g.uniform(0, 40) for _ in range(N)], np.uint64)
    te = np.array([g.next() >> g.uniform(16, 60) for _ in range(N)], np.uint64)
    fr = np.array([g.uniform(0, 100) for _ in range(N)], np.uint64)
    tr = (te.astype(np.float64) * fr.astype(np.float64) / 100.0).astype(np.uint64)
    te[:8] = [4, 1, 1000, 0, 5, 0, 1 << 40, 3]
    tr[:8] = [2, 1, 333, 0, 0, 7, 1 << 39, 3]
    cnt[:8] = [200, 100, 12345, 9, 9, 9, (1 << 50) + 1, 0]
    # keep count*te/tr < 2^63 (beyond that the reference's double->u64 cast is UB)
    ok = (cnt.astype(np.float64) * te.astype(np.float64)) < (tr.astype(np.float64) * 2.0**62)
    cnt = np.where(ok | (tr == 0), cnt, cnt >> np.uint64(24))
    ok = (cnt.astype(np.float64) * te.astype(np.float64)) < (tr.astype(np.float64) * 2.0**62)
    cnt = np.where(ok | (tr == 0), cnt, np.uint64(1))
    scaled = ref.get_count(cnt, te, tr)
    num = np.array([g.next() >> g.uniform(0, 50) for _ in range(N)], np.uint64)
    den = np.array([g.next() >> g.uniform(0, 50) for _ in range(N)], np.uint64)
    den[::17] = 0
    rate, rate_valid = ref.rate(num, den)
    addv = np.array([[g.next() >> g.uniform(20, 50) for _ in range(5)] for _ in range(512)],
                    np.uint64)
    addv[0, :2] = [3_000_000_000, 3_000_000_000]
    addv[0, 2:] = 0
    add = ref.add_reduce(addv)
    tp = po.make_params(1, time_shift=10, time_mult=341, time_zero=0x1234_5678_9ABC)
    tsc = np.array([g.next() >> g.uniform(0, 30) for _ in range(N)], np.uint64)
    kt = ref.tsc_to_ktime(tp, tsc)
    back = ref.ktime_to_tsc(tp, kt)
    trn = np.maximum(tr, np.uint64(1))
    psm = ref.per_second_mega(scaled, trn)
    ratio, ratio_i, mux = ref.running_ratio(te, tr)
    np.savez_compressed(os.path.join(HERE, "math.npz"), cnt=cnt, te=te, tr=tr, scaled=scaled,
                        num=num, den=den, rate=rate, rate_valid=rate_valid, addv=addv, add=add,
                        tsc=tsc, kt=kt, back=back, tsc_params=np.array([10, 341, 0x1234_5678_9ABC],
                                                                       np.uint64),
                        trn=trn, psm=psm, ratio=ratio, ratio_i=ratio_i, mux=mux)
    print("math: get_count known answers", scaled[:3])


if __name__ == "__main__":
    main()
