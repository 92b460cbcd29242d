import os, sys
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import numpy as np
from dynolog_b200 import hbtgpu, synth
from oracle import pyoracle as po
port = po.load("port")
ncpu, n = 4, 8
spec, streams = synth.capture(3, ncpu, max_bytes_per_cpu=16 << 20, forge_ppm=10_000)
for flags in (0, hbtgpu.F_FORCE_GENERAL):
    total = sum(len(s) for s in streams)
    g = hbtgpu.Context(n_events=n, max_cpus=ncpu, n_buckets=70000, t0=synth.T_START, bucket_ns=1_000_000,
                       arena_bytes=total + (1 << 20), flags=flags)
    g.submit_images(list(range(ncpu)), streams)
    g.run(); g.sync()
    st = g.stats()
    print("flags", flags, {k: st[k] for k in ("n_chunks", "exact_runs", "exact_reason", "exact_chunk")})
    if st["exact_runs"]:
        # chunk geometry: chunks are in (cpu, position) order, all streams equal length here
        per = [len(s) for s in streams]
        cb = 16384
        tgt = total // (148 * 12)
        while cb < tgt and cb < (512 << 10): cb <<= 1
        c = st["exact_chunk"]; cpu = 0
        while c >= (per[cpu] + cb - 1) // cb:
            c -= (per[cpu] + cb - 1) // cb; cpu += 1
        beg = c * cb
        print("chunk bytes", cb, "cpu", cpu, "chunk in stream", c, "begin", beg)
        r = port.decode_stream(po.make_params(n), streams[cpu], po.make_carry(), want_rows=False, want_buckets=False)
        off, typ, siz = r["records"]
        i = np.searchsorted(off, beg)
        print("records around chunk begin:", [(int(off[k]) - beg, int(typ[k]), int(siz[k])) for k in range(max(0, i - 4), i + 4)])
        # find first tile (16 KiB) boundary records with odd types in that chunk
        sel = (off >= beg) & (off < beg + cb) & (typ != 9)
        print("non-sample records in chunk:", [(int(o) - beg, int(t), int(z)) for o, t, z in zip(off[sel], typ[sel], siz[sel])][:20])
    g.close()
