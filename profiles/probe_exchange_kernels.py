"""The kernels of one sharded step, two ranks emulated on one GPU (for an ncu launch list: the multi-process
run cannot be profiled): python profiles/probe_exchange_kernels.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dynolog_b200 import hbtgpu, synth

ncpu, n, world = 192, 8, 2
cap = synth.Capture(synth.make_spec(2), ncpu)
ctxs = []
for r in range(world):
    streams = cap.window(n_samples=20_000)
    total = sum(len(s) for s in streams)
    g = hbtgpu.Context(n_events=n, max_cpus=ncpu, n_buckets=10_100, t0=synth.T_START, bucket_ns=1_000_000,
                       arena_bytes=total + (1 << 20), flags=hbtgpu.F_EXTERNAL_CARRY,
                       metrics=[(hbtgpu.METRIC_RATE, [0, 1])])
    c = hbtgpu.make_carry(valid=True, tstamp=synth.T_START, last=[0] * n)
    for cpu in range(ncpu):
        g.set_carry(cpu, c)
    g.submit_images(list(range(ncpu)), [s.copy() for s in streams])
    ctxs.append(g)
for r, g in enumerate(ctxs):
    g.peer_attach_local(world, r, ctxs)
for rep in range(4):
    for g in ctxs:
        g.run()
    for g in ctxs:
        g.exchange_boundaries()
    for g in ctxs:
        g.sync()
print("ok")
for g in ctxs:
    g.close()
