import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dynolog_b200 import hbtgpu, synth
N_CPUS, N_EVENTS, NB = 192, 8, 10100
spec, streams = synth.capture(2, N_CPUS, n_samples=100_000)
total = sum(len(s) for s in streams)
for mode in ("plain", "shard"):
    g = hbtgpu.Context(n_events=N_EVENTS, max_cpus=N_CPUS, n_buckets=NB, t0=synth.T_START, bucket_ns=1_000_000,
                       arena_bytes=total + (1 << 20), flags=hbtgpu.F_EXTERNAL_CARRY if mode == "shard" else 0,
                       metrics=[(hbtgpu.METRIC_RATE, [1, 0])])
    c = hbtgpu.make_carry(valid=True, tstamp=synth.T_START, last=[0] * N_EVENTS)
    for cpu in range(N_CPUS):
        g.set_carry(cpu, c)
    g.submit_images(list(range(N_CPUS)), streams)
    sb = g.stub_bytes()
    buf = torch.empty(2 * sb, dtype=torch.uint8, device="cuda")
    ext = torch.cuda.ExternalStream(g.stream())
    def step():
        g.run()
        if mode == "shard":
            g.export_stubs_device(buf.data_ptr())
            g.apply_boundaries_gathered(buf.data_ptr(), 1, 0)
    for _ in range(3): step()
    g.sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(ext)
    for _ in range(20): step()
    e1.record(ext)
    g.sync()
    print(mode, e0.elapsed_time(e1) / 20, "ms/step", g.stats()["kernel_launches"], "launches")
    g.close()
