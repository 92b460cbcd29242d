"""BASELINE config 4 over N GPUs (last bullet): the capture is sharded by time
window, every rank streams its windows through one context (HBTGPU_F_ACCUMULATE_ATTR), then
hbtgpu_attr_merge gives every rank the (pid,tid) histogram and the sampled-IP counts of the whole
capture over peer memory.  Checked on EVERY rank against the generator's own parameters (each
rank regenerates nothing: the per-rank reference counts are summed with one NCCL all-reduce of
two small integer vectors, which is test infrastructure, not the product path).
  python -m torch.distributed.run --nproc-per-node N profiles/probe_attr_merge.py [windows_per_rank] [MiB_per_cpu]"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from dynolog_b200 import hbtgpu, synth  # noqa: E402
from oracle import pyoracle as po  # noqa: E402  (the checker)

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
W = int(sys.argv[1]) if len(sys.argv) > 1 else 2
per_cpu = (int(sys.argv[2]) if len(sys.argv) > 2 else 8) << 20
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
n, ncpu, t0, nb = 4, 192, synth.T_START, 4096
port = po.load("port")
p = po.make_params(n, n_buckets=nb, t0=t0, bucket_ns=100_000_000, pre=2)
cap = synth.Capture(synth.make_spec(4), ncpu, n_threads=max(2, (os.cpu_count() or 1) // world))
for _ in range(rank * W):  # this rank's windows follow those of the ranks before it
    cap.window(max_bytes_per_cpu=per_cpu, dry=True)
ctx = hbtgpu.Context(n_events=n, max_cpus=ncpu, n_buckets=nb, t0=t0, bucket_ns=100_000_000,
                     arena_bytes=ncpu * per_cpu + (1 << 20), device=lr, sample_format=hbtgpu.SAMPLE_IP_TID_TIME_READ,
                     flags=hbtgpu.F_EXTERNAL_CARRY | hbtgpu.F_ACCUMULATE_ATTR)
handles = [None] * world
dist.all_gather_object(handles, ctx.peer_export(world))
ctx.peer_attach(world, rank, handles)
ext = torch.cuda.ExternalStream(ctx.stream(), device=lr)
tid_counts = np.zeros(4096, np.int64)
ip_counts = np.zeros(1 << 20, np.int64)
buf = np.empty(ncpu * ((per_cpu + 127) // 128 * 128), np.uint8)
samples = 0
for w in range(W):
    streams = cap.window(max_bytes_per_cpu=per_cpu, out=buf)
    ctx.reset(keep_carry=True)
    ctx.submit_images(list(range(ncpu)), streams)
    ctx.run()
    ctx.sync()
    samples += sum(ctx.counters(i)["sample"] for i in range(ncpu))
    for s in streams:  # the checker's counts of this rank's windows
        tids, ips = port.collect_ip_tid(p, s, po.make_carry())
        tid_counts += np.bincount((tids >> np.uint64(32)).astype(np.int64) - 5000, minlength=4096)
        ip_counts += np.bincount(((ips - np.uint64(0xffffffff81000000)) >> np.uint64(4)).astype(np.int64),
                                 minlength=1 << 20)
dist.barrier()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
reps = 10
ctx.attr_merge()  # warm-up
ctx.sync()
dist.barrier()
e0.record(ext)
for _ in range(reps):
    ctx.attr_merge()
e1.record(ext)
ctx.sync()
ms = torch.tensor([e0.elapsed_time(e1) / reps], device="cuda", dtype=torch.float64)
dist.all_reduce(ms, op=dist.ReduceOp.MAX)
hist = ctx.tid_histogram()
top = ctx.top_ips(100)
# the whole capture's reference counts
tc = torch.from_numpy(tid_counts).cuda()
ic = torch.from_numpy(ip_counts).cuda()
ns = torch.tensor([samples], device="cuda", dtype=torch.int64)
dist.all_reduce(tc)
dist.all_reduce(ic)
dist.all_reduce(ns)
tid_all, ip_all = tc.cpu().numpy(), ic.cpu().numpy()
exp_hist = {(1000 + r // 8, 5000 + r): int(k) for r, k in enumerate(tid_all) if k}
order = np.lexsort((np.arange(1 << 20), -ip_all))[:100]
exp_top = [(0xffffffff81000000 + 16 * int(r), int(ip_all[r])) for r in order]
ok = torch.tensor([int(hist == exp_hist and top == exp_top and sum(hist.values()) == int(ns.item()))], device="cuda")
dist.all_reduce(ok, op=dist.ReduceOp.MIN)
if rank == 0:
    print(json.dumps({"n_gpus": world, "windows_per_rank": W, "bytes_per_rank": int(W * ncpu * per_cpu),
                      "samples_total": int(ns.item()), "attr_merge_ms": round(float(ms.item()), 4),
                      "tables_bytes_per_rank": int(8 * (2 * (1 << 16) + 2 * (1 << 22) + 4)),
                      "distinct_tids": len(exp_hist), "distinct_ips": int((ip_all > 0).sum()),
                      "parity_all_ranks": bool(ok.item())}))
ctx.close()
dist.barrier()
dist.destroy_process_group()
