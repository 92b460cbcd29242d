"""ContextSwitch-mode decode throughput (resident images): python profiles/perf_probe_cs.py
Algorithmic bytes = the image bytes, read once; what the two passes move is stated besides."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from dynolog_b200 import hbtgpu, hbtgpu_cs, synth

PEAK = 6547.5

def probe(name, ncpu, bytes_per_cpu, flags=0, host=False, **kw):
    streams = synth.cs_capture(ncpu, bytes_per_cpu, **kw)
    total = sum(len(s) for s in streams)
    g = hbtgpu_cs.CsContext(ncpu, total + 256 * ncpu + (1 << 20), flags=flags)
    for cpu, s in enumerate(streams):
        g.submit(cpu, s)
    for _ in range(3):
        g.run(); g.sync()
    cnt, emit, run = [], [], []
    for _ in range(10):
        g.run(); g.sync()
        st = g.stats()
        cnt.append(st["last_count_ms"]); emit.append(st["last_emit_ms"]); run.append(st["last_run_ms"])
    cnt, emit, run = np.mean(cnt), np.mean(emit), np.mean(run)
    nrec = sum(g.counters(i)["num_records"] for i in range(ncpu))
    nev = sum(g.num_events(i)[0] for i in range(ncpu))
    st = g.stats()
    print(f"{name:34s} {total/1e9:5.2f} GB {nrec/1e6:6.1f} Mrec  count {cnt:6.3f} ms ({total/1e6/cnt:6.0f} GB/s, {100*total/1e6/cnt/PEAK:4.1f}%)"
          f"  emit {emit:6.3f} ms ({(total + 20*nev)/1e6/emit:6.0f} GB/s moved)  run {run:6.3f} ms = {nrec/1e6/run:5.1f} Grec/s"
          f" = {100*total/1e6/run/PEAK:4.1f}% of {PEAK}  exact_runs={st['exact_runs']} why={st['exact_reason']} chunks={st['n_chunks']}")
    if host:
        import time
        t0 = time.time()
        data = [(g.events(c), g.ops(c)) for c in range(ncpu)]
        t1 = time.time()
        th = hbtgpu_cs.Threads()
        for (ev, pid), ops in data:
            th.assign(ev, pid, ops)
        t2 = time.time()
        print(f"{'':34s} host: D2H of events+pids+ops (pageable) {t1-t0:.3f} s; thread dictionary (hbtgpu_cs_assign_vids, 1 thread) "
              f"{t2-t1:.3f} s = {nev/1e6/(t2-t1):.0f} M events/s, {th.L.hbtgpu_cs_threads_count(__import__("ctypes").c_void_p(th.h))} vids")
        th.close()
        ms = []
        for rep in range(3):  # (fresh dictionary each time; the first call also pays for its allocations)
            dev = hbtgpu_cs.DeviceThreads()
            g.run(); g.sync()
            dev.assign(g)
            ms.append(dev.last_ms())
            nv = dev.count()
            dev.close()
        print(f"{'':34s} device: thread dictionary (hbtgpu_cs_assign_vids_device) {min(ms):.2f} ms (first call {ms[0]:.2f}) = "
              f"{nev/1e3/min(ms):.0f} M events/s, {nv} vids")
    g.close()

def e2e(ncpu, bytes_per_cpu, ref_cpus=4):
    """Host images in, events with virtual ids out (pinned host memory both ways), against the
    reference's own code on a bounded sample (it is single-threaded by design: one shared
    PerThreadStats)."""
    import time
    import torch
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from oracle import pyoracle
    cap = (bytes_per_cpu + 127) // 128 * 128
    pin_in = torch.empty(ncpu * cap, dtype=torch.uint8, pin_memory=True).numpy()
    streams = synth.cs_capture(ncpu, bytes_per_cpu, out=pin_in)
    total = sum(len(s) for s in streams)
    g = hbtgpu_cs.CsContext(ncpu, total + 256 * ncpu + (1 << 20))
    # at most one event per 20 bytes of input (a 40-byte LOST writes two)
    ev_out = torch.empty((total // 20 + 64) * 16, dtype=torch.uint8, pin_memory=True).numpy().view(hbtgpu_cs.EVENT_DTYPE)
    import ctypes as C
    best = None
    for it in range(3):
        dev = hbtgpu_cs.DeviceThreads()
        t0 = time.time()
        g.reset()
        for cpu, s in enumerate(streams):
            g.submit(cpu, s)
        g.run()
        dev.assign(g)
        off = 0
        for cpu in range(ncpu):
            n, _ = g.num_events(cpu)
            got = C.c_uint64(0)
            g._ck(g.L.hbtgpu_cs_get_events(C.c_void_p(g.h), C.c_uint32(cpu), C.c_void_p(ev_out[off:].ctypes.data),
                                           None, C.c_uint64(n), C.byref(got)))
            off += n
        dt = time.time() - t0
        best = dt if best is None else min(best, dt)
        dev.close()
    nrec = sum(g.counters(i)["num_records"] for i in range(ncpu))
    t0 = time.time()
    ref = pyoracle.cs_run("reference_nolog" if pyoracle.available("reference_nolog") else "port", streams[:ref_cpus], batch=1024)
    tr = time.time() - t0
    nref = sum(c["num_records"] for c in ref["counters"])
    print(f"e2e {ncpu} cpus x {bytes_per_cpu >> 20} MiB: H2D {total/1e9:.2f} GB + decode + GPU dictionary + D2H {off*16/1e9:.2f} GB of events: "
          f"{best*1e3:.1f} ms = {nrec/1e6/best:.0f} M records/s;  reference code ({ref_cpus} cpus, {nref/1e6:.1f} M records, 1 thread): "
          f"{tr:.2f} s = {nref/1e6/tr:.2f} M records/s")
    g.close()

if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "e2e":
        e2e(192, 10 << 20)
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "one":   # for ncu
        probe("192 cpus x 10 MiB", 192, 10 << 20)
        sys.exit(0)
    probe("1 cpu x 4 MiB", 1, 4 << 20)
    probe("192 cpus x 10 MiB", 192, 10 << 20, host=True)
    probe("64 cpus x 32 MiB", 64, 32 << 20)
    probe("64 cpus x 32 MiB, 5% other records", 64, 32 << 20, life_ppm=30000, comm_ppm=20000, lost_ppm=2000)
    probe("64 cpus x 16 MiB forced exact", 64, 16 << 20, flags=hbtgpu.F_FORCE_EXACT)
