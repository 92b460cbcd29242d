"""Slicer throughput (events -> slices): python profiles/perf_probe_slicer.py [cpus] [bytes_per_cpu] [tids]
Thread-switch ring images are decoded on the GPU (events with virtual thread ids, resident), then
hbtgpu_slicer_run turns them into slices.  Algorithmic bytes = 16 B per event read + 32 B per
slice written.  The CPU line is the reference's own Slicer<> (oracle/_ref) on the same merged
events, one thread (it is one state machine by construction)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from dynolog_b200 import hbtgpu_cs, hbtgpu_slices as gs, synth

PEAK = 6547.5


def main():
    ncpu = int(sys.argv[1]) if len(sys.argv) > 1 else 192
    bpc = int(sys.argv[2]) if len(sys.argv) > 2 else 4 << 20
    ntid = int(sys.argv[3]) if len(sys.argv) > 3 else 4096
    import torch
    streams = synth.cs_capture(ncpu, bpc, n_tids=ntid, life_ppm=2000, comm_ppm=0, lost_ppm=20, throttle_ppm=0, seed=2)
    dec = hbtgpu_cs.decode(streams, tid_level=0, device_dictionary=True)
    units = dec["events"]
    n_ev = sum(len(u) for u in units)
    eb = np.concatenate([[0], np.cumsum([len(u) for u in units])]).astype(np.uint64)
    d_ev = torch.from_numpy(np.concatenate(units).view(np.uint8)).cuda()
    torch.cuda.synchronize()
    sl = gs.Slicer(0)
    ms, wall = [], []
    for it in range(5):
        stats = np.zeros(8, np.uint64)
        n_slices = np.zeros(ncpu, np.uint64)
        import ctypes as C
        ns, lt = C.c_uint64(0), C.c_uint64(0)
        t0 = time.time()
        sl._check(sl.L.hbtgpu_slicer_run(sl.h, d_ev.data_ptr(), eb.ctypes.data, ncpu, 1, stats.ctypes.data,
                                         n_slices.ctypes.data, C.byref(ns), C.byref(lt)))
        wall.append(time.time() - t0)
        ms.append(sl.last_ms)
    t = float(np.mean(ms[1:]))
    w = float(np.min(wall[1:])) * 1e3
    n_out = int(n_slices.sum())
    alg = 16 * n_ev + 32 * n_out
    print(f"slicer: {ncpu} units, {n_ev/1e6:.2f} M events -> {n_out/1e6:.2f} M slices, {ns.value} stacks, "
          f"{int(stats[1])} write-error windows, {int(stats[4])} dropped")
    print(f"  device {t:8.3f} ms = {n_ev/1e3/t:8.1f} M events/s, {alg/1e6/t:7.1f} GB/s algorithmic = {100*alg/1e6/t/PEAK:4.1f}% of {PEAK}"
          f"   (call incl. host synchronisations and the stack table: {w:.2f} ms)")
    # the general form (phases, any level): a serial walk on the device
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
    from test_slicer_oracle import random_phase_events
    from oracle import pyoracle as po
    r = np.random.default_rng(5)
    gev = random_phase_events(r, 16, 400_000, 300, p_phase=0.35, p_we=0.0005, p_create=0.01, p_bad=0.0)
    gev = gev[np.lexsort((gev["comp_unit_id"], gev["tstamp"]))]
    k = len(gev)
    while po.slicer_run("port_general", gev[:k], 16)["rc"] == -95:  # (cut in front of the reference's uninitialised-tag reads)
        k = k * 9 // 10
    gev = gev[:k]
    gunits = [gev[gev["comp_unit_id"] == u] for u in range(16)]
    gms = []
    for it in range(3):
        got = sl.run(gunits, general=True)
        gms.append(sl.last_ms)
    t0 = time.time()
    exp = po.slicer_run("port_general", gev, 16)
    dt = time.time() - t0
    same = all(np.array_equal(a, b) for a, b in zip(got["slices"], exp["slices"])) and got["stats"] == exp["stats"]
    print(f"  general form: 16 units, {len(gev)/1e3:.0f} k events with phases at levels 1..3 -> {int(sum(len(x) for x in got['slices']))/1e3:.0f} k slices, "
          f"{len(got['stacks'])} stacks: device {min(gms):.1f} ms = {len(gev)/1e3/min(gms):.2f} M events/s ({1e3*min(gms)/len(gev):.2f} us per event); "
          f"cpu restatement {dt*1e3:.1f} ms; equal: {same}")
    if po.slices_available("reference"):
        n_ref = min(ncpu, 16)
        sub = np.concatenate(units[:n_ref])
        sub = sub[np.lexsort((sub["comp_unit_id"], sub["tstamp"]))]
        t0 = time.time()
        r = po.slicer_run("reference", sub, n_ref)
        dt = time.time() - t0
        print(f"  cpu reference Slicer<>, {n_ref} units merged, 1 thread: {dt*1e3:.1f} ms = {len(sub)/1e6/dt:.2f} M events/s (rc {r['rc']})")
        t0 = time.time()
        r = po.slicer_run("port", sub, n_ref)
        dt = time.time() - t0
        print(f"  cpu restatement: {dt*1e3:.1f} ms = {len(sub)/1e6/dt:.2f} M events/s")


if __name__ == "__main__":
    main()
