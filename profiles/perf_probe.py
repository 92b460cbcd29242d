"""Decode-kernel throughput on the BASELINE configs (resident data): python profiles/perf_probe.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from dynolog_b200 import hbtgpu, synth

def probe(name, cfg, ncpu, n, nb, flags=0, forge=0, fmt=0, series=False, **kw):
    spec, streams = synth.capture(cfg, ncpu, forge_ppm=forge, **kw)
    total = sum(len(s) for s in streams)
    g = hbtgpu.Context(n_events=n, max_cpus=ncpu, n_buckets=nb, t0=synth.T_START, bucket_ns=1_000_000,
                       arena_bytes=total + 256 * ncpu + (1 << 20), flags=flags, sample_format=fmt)
    c = hbtgpu.make_carry(valid=True, tstamp=synth.T_START, last=[0] * n)
    for cpu in range(ncpu):
        g.set_carry(cpu, c)
    g.submit_images(list(range(ncpu)), streams)
    for _ in range(3):
        g.run()
    g.sync()
    s0 = g.stats()
    for _ in range(10):
        g.run()
    g.sync()
    s1 = g.stats()
    dec = (s1["decode_ms_sum"] - s0["decode_ms_sum"]) / (s1["decode_runs"] - s0["decode_runs"])
    nrec = sum(g.counters(i)["num_records"] for i in range(ncpu))
    g.rows(0)  # triggers the emission pass (rows + record table of every cpu)
    emit = g.stats()["last_emit_ms"]
    print(f"{name:34s} {total/1e9:6.2f} GB  decode {dec:7.3f} ms  {total/1e6/dec:7.1f} GB/s "
          f"({100*total/1e6/dec/6547.5:4.1f}% of 6547.5)  step {s1['last_run_ms']:.3f} ms  "
          f"{nrec/1e6/ (s1['last_run_ms']):.1f} Grec/s  emit {emit:.3f} ms  exact_runs={s1['exact_runs']} why={s1['exact_reason']}")
    if series:
        import time
        g.series_reduce(0.95, 1_000_000_000, 10_000_000_000)
        t0 = time.time()
        for _ in range(5):
            g.series_reduce(0.95, 1_000_000_000, 10_000_000_000)
        dt = (time.time() - t0) / 5
        print(f"{'':34s} series reductions (avg, p95, diff, rate) of {ncpu * n} series x {nb} buckets: {dt*1e3:.2f} ms per call")
    g.close()

if __name__ == "__main__":
    probe("cfg1 4MiB 1cpu n=4", 1, 1, 4, 6000, max_bytes_per_cpu=4 << 20)
    probe("cfg2 192cpu n=8 (bench)", 2, 192, 8, 10100, n_samples=100_000, series=True)
    probe("cfg3 64cpu x 64MiB mux+LOST", 3, 64, 8, 70000, max_bytes_per_cpu=64 << 20)
    probe("cfg3 forged 1% (exact path)", 3, 64, 8, 70000, max_bytes_per_cpu=16 << 20, forge=10_000)
    probe("cfg3 force_general", 3, 64, 8, 70000, max_bytes_per_cpu=16 << 20, flags=hbtgpu.F_FORCE_GENERAL)
    probe("cfg3 force_exact", 3, 64, 8, 70000, max_bytes_per_cpu=16 << 20, flags=hbtgpu.F_FORCE_EXACT)
    probe("cfg5 192cpu x 16MiB LOST", 5, 192, 8, 20000, max_bytes_per_cpu=16 << 20)
    probe("n=4 72B records 64cpu x 32MiB", 1, 64, 4, 50000, max_bytes_per_cpu=32 << 20)
    probe("cfg4 IP|TID 88B 64cpu x 32MiB", 4, 64, 4, 50000, fmt=1, max_bytes_per_cpu=32 << 20)
