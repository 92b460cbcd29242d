"""Randomised soak of the GPU slicer (both forms) / interval slicer / binner / thread-switch decode with its
device dictionary / sampling decode against the restatements (seeds
beyond the ones the tests pin; not collected by pytest).  python tests/soak_slices.py [seconds] [seed0]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import slicegen
from dynolog_b200 import hbtgpu_slices as gs
from dynolog_b200.hbtgpu import HbtGpuError
from oracle import pyoracle as po
import test_gpu_slices as TS
import test_gpu_slicer as TL
from test_slicer_oracle import random_events

budget = float(sys.argv[1]) if len(sys.argv) > 1 else 120.0
seed0 = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
isl, sl = gs.IntervalSlicer(0), gs.Slicer(0)
t_end = time.time() + budget
n = {"islice": 0, "bin": 0, "slicer": 0, "windows": 0}
seed = seed0
while time.time() < t_end:
    seed += 1
    r = np.random.default_rng(seed)
    try:
        # interval slicer: regular + hostile units mixed
        n_units = int(r.integers(1, 12))
        us, ur, marks = [], [], []
        stride = int(r.integers(2, 11))
        for u in range(n_units):
            if r.random() < 0.25:
                s = slicegen.hostile_slices(r, int(r.integers(1, 600)), str(r.choice(["zero_dur", "overlap", "unsorted", "shuffled", "ok"])))
                rows = slicegen.hostile_rows(r, np.sort(s, order="tstamp"), int(r.integers(1, 900)),
                                             str(r.choice(["overlap", "backwards", "zero_dur", "repeat", "ok"])))
                rows = np.concatenate([rows, r.integers(0, 9, (len(rows), stride - 2)).astype(np.uint64)], axis=1) if stride > 2 else rows
            else:
                s = slicegen.gen_slices(r, int(r.choice([0, 1, 50, 700, 4000, 20000])), hole_prob=float(r.choice([0, 0.01, 0.2])),
                                        max_dur=int(r.choice([2, 40, 3000])))
                rows = slicegen.gen_rows(r, s, int(r.choice([0, 1, 127, 128, 129, 600, 3000])), stride=stride,
                                         align_prob=float(r.choice([0, 0.1, 0.6, 1.0])), gap_prob=float(r.choice([0, 0.1])))
            us.append(s); ur.append(rows); marks.append(int(r.choice([0, 0, 0, 2, 10**6])))
        flags = int(r.integers(0, 4)) | (8 if r.random() < 0.15 else 0)
        allr = np.concatenate(ur)
        stop = 2**63 if r.random() < 0.6 or len(allr) == 0 else int(np.median(allr[:, 0]))
        units = TS.run_and_compare(isl, us, ur, stop, flags, marks)
        n["islice"] += 1
        if (units["status"] == 0).all() and n_units:
            TS.bin_compare(isl, us, ur, flags & 3, int(r.integers(0, 4)), int(r.choice([0, 1, 100, 10**9])),
                           selected=(r.random(n_units) < 0.7).astype(np.uint8) if r.random() < 0.5 else None) if all(
                (po.islice_release("port", us[u], ur[u], 2**63, flags & 3, 0)["rc"] == 0) for u in range(n_units)) else None
            n["bin"] += 1
        # slicer: one window and several
        nu = int(r.integers(1, 8))
        ev = random_events(r, nu, int(r.integers(1, 2500)), int(r.choice([2, 6, 50, 500])),
                           p_create=float(r.choice([0.0, 0.03, 0.3])), p_we=float(r.choice([0.0, 0.01, 0.1])),
                           p_double_in=float(r.choice([0.0, 0.02, 0.4])), p_migrate=float(r.choice([0.0, 0.1, 0.6])),
                           p_tid0=float(r.choice([0.0, 0.2, 0.9])), p_same_ts=float(r.choice([0.0, 0.05, 0.8])))
        uev, merged = TL.split_units(ev, nu)
        if TL.supported(merged, nu):
            exp = po.slicer_run("port", merged, nu)
            TL.compare(sl.run(uev), exp)
            n["slicer"] += 1
            cuts = sorted(set(int(x) for x in r.choice(merged["tstamp"], size=int(r.integers(1, 6)))))
            bounds = [0] + cuts + [2**63]
            acc = [[] for _ in range(nu)]
            got = None
            for w in range(len(bounds) - 1):
                win = [u_[(u_["tstamp"] >= bounds[w]) & (u_["tstamp"] < bounds[w + 1])] for u_ in uev]
                got = sl.run(win, cont=w > 0)
                for u in range(nu):
                    acc[u].append(got["slices"][u])
            got["slices"] = [np.concatenate(x) for x in acc]
            TL.compare(got, exp)
            n["windows"] += 1
        # the Slicer's general form: phases at levels 1..3, one window and several
        nu = int(r.integers(1, 5))
        full = TL.random_phase_events(r, nu, int(r.integers(1, 1500)), int(r.choice([3, 6, 30, 300])),
                                      p_phase=float(r.choice([0.1, 0.35, 0.7])), p_we=float(r.choice([0.0, 0.01, 0.08])),
                                      p_create=float(r.choice([0.0, 0.03, 0.2])), p_bad=float(r.choice([0.0, 0.02, 0.1])))
        gev, gexp, _ = TL.cut_at_fence(full, nu)
        TL.same_general(TL.run_general(sl, gev, nu, general=bool(r.integers(0, 2))), gexp)
        n["general"] = n.get("general", 0) + 1
        if len(gev) > 4:
            cuts = sorted(set(int(x) for x in r.choice(gev["tstamp"], size=int(r.integers(1, 5)))))
            acc, lo = [[] for _ in range(nu)], 0
            for k, cut in enumerate(cuts + [int(gev["tstamp"].max()) + 1]):
                w = gev[(gev["tstamp"] >= lo) & (gev["tstamp"] < cut)]
                got = TL.run_general(sl, w, nu, cont=k > 0, general=True)
                for u in range(nu):
                    acc[u].append(got["slices"][u])
                lo = cut
            got["slices"] = [np.concatenate(x) for x in acc]
            TL.same_general(got, gexp)
            n["general_windows"] = n.get("general_windows", 0) + 1
        # thread-switch decode + the dictionary on the device
        import ringgen as rg
        from dynolog_b200 import hbtgpu_cs
        streams = rg.random_cs_streams(seed, int(r.integers(1, 6)), int(r.integers(50, 4000)), n_tids=int(r.choice([4, 48, 400])),
                                       p_lost=float(r.choice([0.0, 0.004, 0.05])), p_life=float(r.choice([0.0, 0.05, 0.3])),
                                       p_comm=float(r.choice([0.0, 0.03, 0.3])), p_throttle=float(r.choice([0.0, 0.01])),
                                       p_other=float(r.choice([0.0, 0.02])),
                                       tid_base=int(r.choice([5_000_000, (1 << 24) - 7, 1_900_000_000])))
        want = po.cs_run("port", streams, tid_level=2)
        got = hbtgpu_cs.decode(streams, tid_level=2, device_dictionary=True)
        for cpu in range(len(streams)):
            for f in ("tstamp", "type", "level", "comp_unit_id", "tag"):
                assert np.array_equal(got["events"][cpu][f], want["events"][cpu][f]), (cpu, f)
        assert got["threads"] == want["threads"] and got["names"] == want["names"]
        n["cs"] = n.get("cs", 0) + 1
        # the sampling decode: record mixes, group sizes, bucket sizes, one to many images per cpu
        import gpu_util
        from dynolog_b200 import hbtgpu
        nev = int(r.choice([1, 3, 4, 8, 8, 17, 50]))
        ncpu = int(r.integers(1, 5))
        strs, cars = [], []
        for cpu in range(ncpu):
            d, _ = rg.random_stream(seed * 7 + cpu, int(r.integers(1, 6000)), nev, p_lost=float(r.choice([0.0, 0.002, 0.05])),
                                    p_throttle=float(r.choice([0.0, 0.01, 0.2])), p_other=float(r.choice([0.0, 0.02])),
                                    p_read=float(r.choice([0.0, 0.02])), forge=float(r.choice([0.0, 0.0, 0.02])),
                                    mux=bool(r.integers(0, 2)))
            strs.append(d)
            cars.append(po.make_carry(valid=bool(r.integers(0, 2)), tstamp=1_000_000_000 - 5, te=4, tr=3, last=list(range(1, nev + 1))))
        bns = int(r.choice([1_000, 100_000, 1_000_000, 50_000_000, 10**12]))
        nb = int(min(200_000, 700_000_000 // bns + 3))
        params = po.make_params(nev, n_buckets=nb, t0=1_000_000_000, bucket_ns=bns)
        flags = int(r.choice([0, 0, 0, hbtgpu.F_FORCE_GENERAL, hbtgpu.F_FORCE_EXACT]))
        port = po.load("port")
        exps = [port.decode_stream(params, strs[cpu], cars[cpu]) for cpu in range(ncpu)]
        splits = []
        for cpu in range(ncpu):   # one image, or several cut at record boundaries
            offs, k = exps[cpu]["records"][0], int(r.choice([1, 1, 3, 17]))
            idx = sorted(set(int(x) for x in r.integers(0, max(1, len(offs)), size=k - 1))) if k > 1 and len(offs) else []
            cuts = [0] + [int(offs[i]) for i in idx if int(offs[i]) > 0] + [len(strs[cpu])]
            cuts = sorted(set(cuts))
            splits.append([b - a for a, b in zip(cuts[:-1], cuts[1:])] or [len(strs[cpu])])
        got, _ = gpu_util.gpu_decode(params, strs, cars, flags=flags, split=splits)
        for cpu in range(ncpu):
            exps[cpu]["scaled"] = port.scale_buckets(params, exps[cpu]["buckets"])
            gpu_util.assert_matches_oracle(got[cpu], exps[cpu], nev, what=f"seed {seed} cpu {cpu}")
        n["decode"] = n.get("decode", 0) + 1
    except Exception as e:
        print("FAILED at seed", seed, type(e).__name__, str(e)[:500])
        raise
print("soak ok", n, "seeds", seed0 + 1, "..", seed)
