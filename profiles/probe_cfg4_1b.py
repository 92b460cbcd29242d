"""BASELINE config 4 at its stated size: 1 B PERF_RECORD_SAMPLE records (88 B, IP|TID|TIME|READ) = 88 GB
streamed through ONE context in pieces of 192 CPUs x 44 MiB (8.9 GB) with HBTGPU_F_ACCUMULATE_ATTR; prints
the decode rate per piece, the totals and the head of the top-100.  python profiles/probe_cfg4_1b.py [pieces]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from dynolog_b200 import hbtgpu, synth  # noqa: E402

pieces = int(sys.argv[1]) if len(sys.argv) > 1 else 10
n, ncpu, per_cpu = 4, 192, 46_200_000
cap = synth.Capture(synth.make_spec(4), ncpu)
host = torch.empty(ncpu * ((per_cpu + 127) // 128 * 128), dtype=torch.uint8, pin_memory=True)
t0 = synth.T_START
g = hbtgpu.Context(n_events=n, max_cpus=ncpu, n_buckets=64, t0=t0, bucket_ns=10**12, arena_bytes=host.numel() + (1 << 20),
                   sample_format=hbtgpu.SAMPLE_IP_TID_TIME_READ, flags=hbtgpu.F_ACCUMULATE_ATTR)
c = hbtgpu.make_carry(valid=True, tstamp=t0, last=[0] * n)
for cpu in range(ncpu):
    g.set_carry(cpu, c)
total = nbytes = 0
dec_ms = 0.0
t_all = time.time()
for w in range(pieces):
    streams = cap.window(max_bytes_per_cpu=per_cpu, out=host.numpy())
    if w:
        g.reset(keep_carry=True)
    g.submit_images(list(range(ncpu)), streams)
    g.run()
    g.sync()
    st = g.stats()
    assert st["exact_runs"] == 0, st
    k = sum(g.counters(i)["sample"] for i in range(ncpu))
    total += k
    nbytes += st["image_bytes"]
    dec_ms += st["last_run_ms"]
    print(f"piece {w}: {st['image_bytes'] / 1e9:.2f} GB, {k / 1e6:.1f} M samples, run {st['last_run_ms']:.3f} ms "
          f"({st['image_bytes'] / 1e6 / st['last_run_ms']:.0f} GB/s), decode kernel {st['last_decode_ms']:.3f} ms", flush=True)
hist = g.tid_histogram()
top = g.top_ips(100)
print(f"{total} samples in {nbytes / 1e9:.1f} GB; GPU time {dec_ms:.1f} ms = {nbytes / 1e6 / dec_ms:.0f} GB/s, "
      f"{total / dec_ms / 1e6:.2f} G samples/s; wall {time.time() - t_all:.1f} s (generation + H2D included)")
print(f"histogram: {len(hist)} (pid,tid) pairs, {sum(hist.values())} samples; top IPs: {top[:3]} ... sum of top-100 {sum(c for _, c in top)}")
assert sum(hist.values()) == total
g.close()
