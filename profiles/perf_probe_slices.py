"""Count intervals x slices throughput: python profiles/perf_probe_slices.py [units] [slices_per_unit] [rows_per_unit] [aligned]
Default: 192 comp units, 200k slices (20k context switches/s x 10 s) and 100k count rows
(10 kHz x 10 s, BASELINE config 2's sampling) per unit, TRIM | ABORT_ON_MISSING as
TraceMonitor::releaseCountData_ calls it.  aligned = fraction of row boundaries that fall exactly on a
slice end (default 0.001; those rows depend on their predecessor).
Algorithmic bytes = 32 B per slice read + 32 B per slice written + 16 B per row read + 8 B per
row written (offset, count).  The CPU line times the reference's own code (real IntervalSlicer,
oracle/_ref) and the C restatement on ONE unit, single thread (the reference releases unit by
unit under a mutex, TraceMonitor.cpp:121-146)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from dynolog_b200 import hbtgpu_slices as gs

PEAK = 6547.5


def gen_unit(r, ns, K, align=0.02, hole=0.01, stride=10):
    s = np.zeros(ns, gs.SLICE_DTYPE)
    dur = r.integers(2_000, 100_000, ns).astype(np.uint64)           # 2-100 us slices (ns units)
    gap = np.where(r.random(ns) < hole, r.integers(1, 5_000, ns), 0).astype(np.uint64)
    start = np.uint64(1_000_000) + np.cumsum(gap) + np.concatenate([[0], np.cumsum(dur)[:-1]]).astype(np.uint64)
    s["tstamp"], s["duration"] = start, dur
    s["stack_id"] = r.integers(0, 4096, ns)
    s["swin_type"] = r.integers(2, 4, ns)
    s["swout_type"] = r.integers(2, 4, ns)
    ends = start + dur
    span = int(ends[-1] - start[0])
    P = max(2, span // K)
    b = int(start[0]) + np.cumsum(r.integers(P // 2, P + P // 2, K + 1)).astype(np.int64) - P
    snap = r.random(K + 1) < align
    j = np.minimum(np.searchsorted(ends, b[snap].astype(np.uint64)), ns - 1)
    b[snap] = ends[j].astype(np.int64)
    b = np.unique(b)                                                   # sorted, strictly increasing
    rows = np.zeros((len(b) - 1, stride), np.uint64)
    rows[:, 0] = b[:-1]
    rows[:, 1] = b[1:] - b[:-1]
    rows[:, 2:] = r.integers(0, 1 << 30, (len(b) - 1, stride - 2))
    return s, rows


def main():
    U = int(sys.argv[1]) if len(sys.argv) > 1 else 192
    NS = int(sys.argv[2]) if len(sys.argv) > 2 else 200_000
    K = int(sys.argv[3]) if len(sys.argv) > 3 else 100_000
    ALIGN = float(sys.argv[4]) if len(sys.argv) > 4 else 0.001
    STRIDE = int(sys.argv[5]) if len(sys.argv) > 5 else 10  # words per count row (2: the decode's compact {tstamp, duration})
    import torch
    r = np.random.default_rng(1)
    protos = [gen_unit(r, NS, K, align=ALIGN, stride=STRIDE) for _ in range(4)]
    print(f"units {U}, {NS} slices and {K} rows per unit, {ALIGN} of the row boundaries on a slice end")
    us = [protos[u % 4][0] for u in range(U)]
    ur = [protos[u % 4][1] for u in range(U)]
    sb = np.concatenate([[0], np.cumsum([len(s) for s in us])]).astype(np.uint64)
    rb = np.concatenate([[0], np.cumsum([len(x) for x in ur])]).astype(np.uint64)
    slices = np.concatenate(us)
    rows = np.concatenate(ur)
    stride = rows.shape[1]
    d_s = torch.from_numpy(slices.view(np.uint8)).cuda()
    d_r = torch.from_numpy(rows.view(np.int64)).cuda()
    torch.cuda.synchronize()
    g = gs.IntervalSlicer(0)
    for flags, name in ((3, "trim|abort"), (1, "trim"), (3 | gs.FORCE_EXACT, "trim|abort, serial walk")):
        if flags & gs.FORCE_EXACT and U * NS > 4_000_000:
            Us = max(1, 4_000_000 // NS)
        else:
            Us = U
        ms = []
        for it in range(6 if not flags & gs.FORCE_EXACT else 2):
            marks = np.zeros(Us, np.uint64)
            units = g.release_device(d_s.data_ptr(), sb[:Us + 1], d_r.data_ptr(), stride, rb[:Us + 1], 2**63, flags, marks)
            ms.append(g.last_ms)
        t = float(np.mean(ms[1:]))
        n_out = int(units["n_slices"].sum())
        n_s, n_r = int(sb[Us]), int(rb[Us])
        alg = 32 * n_s + 32 * n_out + 16 * n_r + 8 * n_r
        print(f"{name:26s} units {Us:4d} slices {n_s/1e6:6.2f} M rows {n_r/1e6:6.2f} M out {n_out/1e6:6.2f} M "
              f"exact {int(units['exact_path'].sum())}  kernels {t:8.3f} ms = {n_s/1e3/t:8.1f} M slices/s, "
              f"{n_r/1e3/t:8.1f} M rows/s, {alg/1e6/t:7.1f} GB/s algorithmic = {100*alg/1e6/t/PEAK:4.1f}% of {PEAK}")
    # row f1: frequencies + the interval x bin matrix over a released window (256 stack ids)
    Ub, NSb, Kb = min(U, 192), max(1000, NS // 10), max(500, K // 10)
    pb = [gen_unit(r, NSb, Kb, align=ALIGN) for _ in range(4)]
    for s_, _ in pb:
        s_["stack_id"] %= 256
    bs = np.concatenate([pb[u % 4][0] for u in range(Ub)])
    br = np.concatenate([pb[u % 4][1] for u in range(Ub)])
    bsb = np.concatenate([[0], np.cumsum([len(pb[u % 4][0]) for u in range(Ub)])]).astype(np.uint64)
    brb = np.concatenate([[0], np.cumsum([len(pb[u % 4][1]) for u in range(Ub)])]).astype(np.uint64)
    units = g.release(bs, bsb, br, brb, 2**63, 3, np.zeros(Ub, np.uint64))
    ms = []
    for it in range(4):
        res = g.bin(2, 1000, max_matrix_bytes=16 << 30, fetch=False)
        ms.append(g.last_ms)
    t = float(np.mean(ms[1:]))
    n_out = int(units["n_slices"].sum())
    print(f"bin (f1): units {Ub} intervals {len(br)/1e6:.2f} M slices {n_out/1e6:.2f} M stack ids {res['num_stack_ids']} "
          f"matrix {res['num_rows']} x {res['num_bins']} ({res['num_rows']*res['num_bins']*8/1e9:.2f} GB)  kernels {t:.3f} ms "
          f"= {n_out/1e3/t:.0f} M slices/s (matrix memset not included)")
    from oracle import pyoracle as po_
    t0 = time.time()
    uc = [g.get(u) for u in range(2)]
    for kind in ("reference", "port"):
        if not po_.slices_available(kind):
            continue
        t0 = time.time()
        po_.bin_durations(kind, [uc[0][0], uc[1][0]], [uc[0][1], uc[1][1]], 2, 1000)
        dt = time.time() - t0
        print(f"cpu {kind:9s} bin, 2 units, 1 thread: {dt*1e3:.1f} ms = {(len(uc[0][1]) + len(uc[1][1]))/1e6/dt:.2f} M slices/s"
              f"{' (real MonData + TagStackIdBinner, building the MonData included)' if kind == 'reference' else ''}")
    # end to end from host memory (pinned), counts + slices of every unit back
    pin_s = torch.from_numpy(slices.view(np.uint8)).pin_memory().numpy().view(gs.SLICE_DTYPE)
    pin_r = torch.from_numpy(rows.view(np.int64)).pin_memory().numpy().view(np.uint64).reshape(rows.shape)
    best = None
    for it in range(3):
        marks = np.zeros(U, np.uint64)
        t0 = time.time()
        units = g.release(pin_s, sb, pin_r, rb, 2**63, 3, marks)
        t1 = time.time()
        best = t1 - t0 if best is None else min(best, t1 - t0)
    print(f"host in (pinned), summary out: {best*1e3:8.1f} ms = {int(sb[U])/1e6/best:7.1f} M slices/s "
          f"({(slices.nbytes + 16 * len(rows))/1e9/best:5.1f} GB/s over PCIe)")
    g.close()
    # CPU: the reference on one unit
    from oracle import pyoracle as po
    for kind in ("reference", "port"):
        if not po.slices_available(kind):
            continue
        t0 = time.time()
        res = po.islice_release(kind, us[0], ur[0], 2**63, 3, 0)
        dt = time.time() - t0
        print(f"cpu {kind:9s} 1 unit, 1 thread: {dt*1e3:8.1f} ms = {len(us[0])/1e6/dt:6.2f} M slices/s, "
              f"{len(ur[0])/1e6/dt:6.2f} M rows/s (rc {res['rc']}, {len(res['slices'])} slices out"
              f"{'; includes filling the IntervalSlicer ring' if kind == 'reference' else ''})")


if __name__ == "__main__":
    main()
