"""BASELINE config 5 at full size on ONE GPU: 192 CPUs x 21 windows x 16 MiB = 63 GiB, submitted
window by window (21 images per cpu), decoded + aggregated in one run.
    python profiles/probe_cfg5_full.py [n_windows]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from dynolog_b200 import hbtgpu, synth

W = int(sys.argv[1]) if len(sys.argv) > 1 else 21
ncpu, n, win = 192, 8, 16 << 20
cap = synth.Capture(synth.make_spec(5), ncpu)
nb = int(W * (win / 104) * 100_000 / 1_000_000 * 1.1) + 1000
g = hbtgpu.Context(n_events=n, max_cpus=ncpu, n_buckets=nb, t0=synth.T_START, bucket_ns=1_000_000,
                   arena_bytes=W * ncpu * (win + 256) + (1 << 20))
c = hbtgpu.make_carry(valid=True, tstamp=synth.T_START, last=[0] * n)
for cpu in range(ncpu):
    g.set_carry(cpu, c)
buf = np.empty(ncpu * win, np.uint8)
total = 0
t0 = time.time()
for w in range(W):
    streams = cap.window(max_bytes_per_cpu=win, out=buf)
    g.submit_images(list(range(ncpu)), streams)
    total += sum(len(s) for s in streams)
print(f"generated + uploaded {total/2**30:.1f} GiB in {time.time()-t0:.1f} s")
for _ in range(2):
    g.run()
g.sync()
s0 = g.stats()
for _ in range(5):
    g.run()
g.sync()
s1 = g.stats()
dec = (s1["decode_ms_sum"] - s0["decode_ms_sum"]) / (s1["decode_runs"] - s0["decode_runs"])
nrec = sum(g.counters(i)["num_records"] for i in range(ncpu))
lost = sum(g.counters(i)["lost"] for i in range(ncpu))
print(f"cfg5 full: {total/1e9:.2f} GB, {nrec/1e6:.1f} M records, lost={lost}, chunks={s1['n_chunks']}, "
      f"decode {dec:.3f} ms = {total/1e6/dec:.1f} GB/s ({100*total/1e6/dec/6547.5:.1f}% of 6547.5), "
      f"step {s1['last_run_ms']:.3f} ms = {nrec/1e6/s1['last_run_ms']:.1f} G records/s, exact_runs={s1['exact_runs']}")
g.close()
