import os, sys
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import numpy as np
from dynolog_b200 import hbtgpu, synth
from oracle import pyoracle as po
port = po.load("port")
n = 8
spec, streams = synth.capture(3, 4, max_bytes_per_cpu=16 << 20, forge_ppm=10_000)
s0 = streams[0]
r = port.decode_stream(po.make_params(n), s0, po.make_carry(), want_rows=False, want_buckets=False)
off = r["records"][0]
def run(img, flags):
    g = hbtgpu.Context(n_events=n, max_cpus=1, n_buckets=70000, t0=synth.T_START, bucket_ns=1_000_000,
                       arena_bytes=len(img) + (1 << 20), flags=flags)
    g.submit_image(0, img); g.run(); g.sync(); st = g.stats(); g.close()
    return st["exact_runs"], st["exact_reason"], st["exact_chunk"]
beg = 3735552 + 32768
i = np.searchsorted(off, beg)
start = int(off[i])
for ln in (16384 * 4, 16384 * 2, 16384):
    j = np.searchsorted(off, start + ln)
    img = s0[start:int(off[j])]
    print(ln, len(img), run(img, hbtgpu.F_FORCE_GENERAL), run(img, 0))
j = np.searchsorted(off, start + 16384 * 2)
img = s0[start:int(off[j])]
np.save("/root/repo/gpurun_out/forged_img.npy", img)
