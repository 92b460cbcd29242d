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
        dev = hbtgpu_cs.DeviceThreads()
        g.run(); g.sync()
        dev.assign(g)
        print(f"{'':34s} device: thread dictionary (hbtgpu_cs_assign_vids_device) {dev.last_ms():.2f} ms = "
              f"{nev/1e3/dev.last_ms():.0f} M events/s, {dev.count()} vids")
        dev.close()
    g.close()

if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "one":   # for ncu
        probe("192 cpus x 10 MiB", 192, 10 << 20)
        sys.exit(0)
    probe("1 cpu x 4 MiB", 1, 4 << 20)
    probe("192 cpus x 10 MiB", 192, 10 << 20, host=True)
    probe("64 cpus x 32 MiB", 64, 32 << 20)
    probe("64 cpus x 32 MiB, 5% other records", 64, 32 << 20, life_ppm=30000, comm_ppm=20000, lost_ppm=2000)
    probe("64 cpus x 16 MiB forced exact", 64, 16 << 20, flags=hbtgpu.F_FORCE_EXACT)
