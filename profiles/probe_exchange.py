"""What the exchange and the merge cost per step (run under torch.distributed.run, N >= 2):
cfg2 per GPU resident in HBM; variants of the step timed with CUDA events on the context stream,
max over ranks.  python -m torch.distributed.run --nproc-per-node N profiles/probe_exchange.py"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import bench  # noqa: E402
from dynolog_b200 import hbtgpu  # noqa: E402

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
streams, keep = bench.gen_capture(rank, n_threads=max(2, (os.cpu_count() or 1) // world))
nbytes = sum(len(s) for s in streams)
ctx = hbtgpu.Context(n_events=8, max_cpus=192, n_buckets=bench.N_BUCKETS, t0=bench.T_START + rank * 10_000_000_000,
                     bucket_ns=1_000_000, arena_bytes=nbytes + (1 << 20), device=lr, flags=hbtgpu.F_EXTERNAL_CARRY,
                     metrics=[(hbtgpu.METRIC_RATE, [0, 1])])
c = hbtgpu.make_carry(valid=True, tstamp=bench.T_START, last=[0] * 8)
for cpu in range(192):
    ctx.set_carry(cpu, c)
handles = [None] * world
dist.all_gather_object(handles, ctx.peer_export(world))
ctx.peer_attach(world, rank, handles)
ctx.submit_images(list(range(192)), streams)
ext = torch.cuda.ExternalStream(ctx.stream(), device=lr)
sb = ctx.stub_bytes()
send = torch.empty(sb, dtype=torch.uint8, device="cuda")
recv = torch.empty(world * sb, dtype=torch.uint8, device="cuda")


def nccl_exchange():
    with torch.cuda.stream(ext):
        ctx.export_stubs_device(send.data_ptr())
        dist.all_gather_into_tensor(recv, send)
        ctx.apply_boundaries_gathered(recv.data_ptr(), world, rank)


variants = {
    "run": lambda: ctx.run(),
    "run+derive": lambda: (ctx.run(), ctx.derive()),
    "run+exchange(derive overlapped)": lambda: (ctx.run(), ctx.exchange_boundaries()),
    "run+exchange(no derive)": lambda: (ctx.run(), ctx.exchange_boundaries(no_derive=True)),
    "run+exchange+merge": lambda: (ctx.run(), ctx.exchange_boundaries(no_derive=True), ctx.merge()),
    "run+nccl all_gather exchange": lambda: (ctx.run(), nccl_exchange()),
}
out = {}
for name, f in variants.items():
    for _ in range(5):
        f()
    ctx.sync()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(ext)
    for _ in range(20):
        f()
    e1.record(ext)
    ctx.sync()
    t = torch.tensor([e0.elapsed_time(e1) / 20], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    out[name] = round(float(t.item()), 4)
if rank == 0:
    print(json.dumps({"n_gpus": world, "ms_per_step": out}))
ctx.close()
dist.barrier()
dist.destroy_process_group()
