#!/usr/bin/env python
"""bench.py -- decode + aggregate of synthetic hbt perf_event captures on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A "step" is one pass of the hot path over one capture: decode every ring image,
aggregate into (cpu, event, 1 ms bucket) sums, multiplex-scale, IPC / miss-rate.
Workload at every N: BASELINE.json configs[1] per GPU -- 192 CPUs, 8-event group,
10 kHz for 10 s (19.2 M SAMPLE records of 104 B = 2.0 GB) -- i.e. weak scaling:
rank r holds time window r of the same 192 CPUs.

Prints ONE JSON line (rank 0).  `value` = records/s with the capture resident in
HBM; `e2e` = the same through the C ABI with HOST (pinned) buffers, H2D of the
images and D2H of the results inside the timed region.  `--impl reference` times
the reference's own CPU code (oracle/_ref, the real hbt sources) on the host cores.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_CPUS = 192
N_EVENTS = 8
SAMPLES_PER_CPU = 100_000
BUCKET_NS = 1_000_000
N_BUCKETS = 10_000 + 100  # 10 s of 1 ms buckets (+ slack for period jitter)
RECORD_BYTES = 40 + 8 * N_EVENTS
METRIC = "perf_records_decoded_aggregated_per_s"
UNIT = "records/s"
WORKLOAD = ("cfg2: 192-CPU capture, 8-event group, 10 kHz x 10 s, 1 ms buckets, "
            "IPC + LLC-miss-rate (synthetic, 19.2M SAMPLE records, 2.0 GB per GPU)")


def ncu_traffic():
    """dram read+write bytes per launch of the dominant kernel, from the committed ncu capture."""
    p = os.path.join(ROOT, "profiles", "r01_traffic.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["traffic_bytes_per_launch"])
    return None


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """SM clock and throttle reasons DURING the timed region: NVML polled from a thread (about
    every millisecond; the timed region lasts ~10 ms), nvidia-smi as a fallback."""
    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown",
               0x4: "sw_power_cap"}

    def __init__(self, index):
        self.index = index
        self.sm, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thr = None
        self._nvml = None

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nvml = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM))
            self._thr = threading.Thread(target=self._poll, daemon=True)
            self._thr.start()
        except Exception:
            self._nvml = None

    def _poll(self):
        nv = self._nvml
        while not self._stop.is_set():
            try:
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM)))
                r = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self._h))
                for bit, name in self.REASONS.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                break
            time.sleep(0.001)

    def stop(self):
        if self._nvml is None:
            return self._smi_once()
        self._stop.set()
        self._thr.join(timeout=1.0)
        return {"sm_mhz": statistics.median(self.sm) if self.sm else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.sm), "source": "nvml"}

    def _smi_once(self):
        try:
            out = subprocess.check_output(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=clocks.sm,clocks.max.sm",
                 "--format=csv,noheader,nounits"], text=True, timeout=10).strip().split(",")
            return {"sm_mhz": float(out[0]), "sm_max_mhz": float(out[1]), "reasons": [],
                    "samples": 1, "source": "nvidia-smi after the run"}
        except Exception:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["clock sampling unavailable"],
                    "samples": 0}


def gen_capture(rank, pinned=True):
    """Time window `rank` of the 192-CPU capture, in pinned host memory. Returns (streams, keepalive)."""
    import numpy as np
    import torch
    from dynolog_b200 import synth
    spec = synth.make_spec(2)
    cap = synth.Capture(spec, N_CPUS)
    per_cpu = ((SAMPLES_PER_CPU * RECORD_BYTES + 4096) + 127) // 128 * 128
    if pinned and torch.cuda.is_available():
        host = torch.empty(N_CPUS * per_cpu, dtype=torch.uint8, pin_memory=True)
    else:
        host = torch.empty(N_CPUS * per_cpu, dtype=torch.uint8)
    buf = host.numpy()
    streams = None
    for w in range(rank + 1):  # the generator is sequential in time: skip to our window
        streams = cap.window(n_samples=SAMPLES_PER_CPU, max_bytes_per_cpu=per_cpu, out=buf)
    t_start = synth.T_START + rank * SAMPLES_PER_CPU * 100_000  # nominal window start
    return streams, host, t_start


def cpu_reference_run(streams, t0, n_threads, reps, kind):
    """One timed pass = reference consume() on every stream + row drain + bucket sums."""
    from oracle import pyoracle as po
    o = po.load(kind)
    p = po.make_params(N_EVENTS, n_buckets=N_BUCKETS, t0=t0, bucket_ns=BUCKET_NS)
    carries = [po.make_carry(valid=True, tstamp=t0, last=[0] * N_EVENTS) for _ in streams]
    times = []
    nrec = 0
    for _ in range(reps):
        secs, _, ctrs = o.run_capture(p, streams, carries, n_threads=n_threads, batch=4096,
                                      want_buckets=True)
        times.append(secs)
        nrec = sum(c["sample"] + c["throttle"] + c["unthrottle"] + c["read"] for c in ctrs)
    return times, nrec, o.description


def pick_cpu_kind():
    from oracle import pyoracle as po
    if po.available("reference"):
        return "reference", "reference"
    if not po.available("port"):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "port"],
                              stdout=subprocess.DEVNULL)
    return "port", "port"


def run_reference_arm(args, rank, world):
    """--impl reference: the reference's CPU implementation on the host cores (rank 0 only)."""
    if rank != 0:
        return
    lib_kind, kind = pick_cpu_kind()
    streams, keep, t0 = gen_capture(0, pinned=False)
    cores = os.cpu_count() or 1
    n_threads = min(cores, N_CPUS)
    cpu_reference_run(streams, t0, n_threads, max(1, args.warmup), lib_kind)
    times, nrec, desc = cpu_reference_run(streams, t0, n_threads, args.steps, lib_kind)
    ms = 1e3 * sum(times) / len(times)
    value = nrec / (ms * 1e-3)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "n_cpus": N_CPUS, "n_events": N_EVENTS,
                   "bucket_ns": BUCKET_NS},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": n_threads, "kind": kind,
                         "sample": f"one full 192-stream capture per step ({nrec} records), "
                                   f"{n_threads} threads by CPU partition; {desc}; "
                                   "sums delta[] and n_rows per bucket (rows carry no te/tr)"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--e2e-steps", type=int, default=5)
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        run_reference_arm(args, rank, world)
        return

    import numpy as np
    import torch
    import torch.distributed as dist
    from dynolog_b200 import hbtgpu

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the hot path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    streams, keep, t_win = gen_capture(rank)
    image_bytes = int(sum(len(s) for s in streams))
    t0 = 1_000_000_000 + rank * SAMPLES_PER_CPU * 100_000
    metrics = [(hbtgpu.METRIC_RATE, [0, 1]),  # IPC = instructions / cycles (BuiltinMetrics.cpp:616-666)
               (hbtgpu.METRIC_RATE, [2, 0])]  # LLC misses per instruction (BuiltinMetrics.cpp:792-818)
    # N > 1: time-window shards; every rank decodes from an invalid state and the rows that
    # straddle shard boundaries are added after a small all_gather of per-cpu stubs
    ctx = hbtgpu.Context(n_events=N_EVENTS, max_cpus=N_CPUS, n_buckets=N_BUCKETS, t0=t0,
                         bucket_ns=BUCKET_NS, arena_bytes=image_bytes + (1 << 20), metrics=metrics,
                         device=local_rank, flags=hbtgpu.F_EXTERNAL_CARRY if world > 1 else 0)
    carry = hbtgpu.make_carry(valid=True, tstamp=1_000_000_000, last=[0] * N_EVENTS)
    for cpu in range(N_CPUS):
        ctx.set_carry(cpu, carry)
    ext = torch.cuda.ExternalStream(ctx.stream(), device=local_rank)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    if world > 1:
        sb = ctx.stub_bytes()
        send = torch.empty(sb, dtype=torch.uint8, device="cuda")
        recv = torch.empty(world * sb, dtype=torch.uint8, device="cuda")

    def gpu_step():
        """One pass of the hot path over this rank's resident shard (+ the boundary exchange)."""
        ctx.run()
        if world > 1:
            with torch.cuda.stream(ext):
                ctx.export_stubs_device(send.data_ptr())
                dist.all_gather_into_tensor(recv, send)  # NCCL, ordered on the context stream
                ctx.apply_boundaries_gathered(recv.data_ptr(), world, rank)

    # ---- resident leg: images already in HBM --------------------------------------------
    for cpu, s in enumerate(streams):
        ctx.submit_image(cpu, s)
    for _ in range(args.warmup):
        gpu_step()
    ctx.sync()
    st0 = ctx.stats()
    n_records = sum(ctx.counters(c)["num_records"] for c in range(N_CPUS))
    if st0["exact_runs"]:
        print(f"[bench] note: speculation disproved (reason {st0['exact_reason']})", file=sys.stderr)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sampler = ClockSampler(local_rank)
    sampler.start()
    barrier()
    ev0.record(ext)
    for _ in range(args.steps):
        gpu_step()
    ev1.record(ext)
    ctx.sync()
    barrier()
    clocks = sampler.stop()
    ms_total = ev0.elapsed_time(ev1)
    st1 = ctx.stats()
    dec_runs = st1["decode_runs"] - st0["decode_runs"]
    dec_ms = (st1["decode_ms_sum"] - st0["decode_ms_sum"]) / max(1, dec_runs)
    launches_per_step = st1["kernel_launches"]
    exact_runs = st1["exact_runs"]

    # ---- end-to-end leg: host buffers in, host results out ---------------------------------
    # two sets of pinned result buffers: step k+1 must not overwrite what step k is still copying
    out_scaled = [torch.empty((N_CPUS, N_BUCKETS, N_EVENTS), dtype=torch.int64, pin_memory=True) for _ in range(2)]
    out_metric = [torch.empty((N_CPUS, N_BUCKETS, len(metrics)), dtype=torch.float64, pin_memory=True) for _ in range(2)]
    out_valid = [torch.empty((N_CPUS, N_BUCKETS), dtype=torch.int32, pin_memory=True) for _ in range(2)]
    import ctypes as C

    def e2e_step():
        ctx.reset(keep_carry=False)
        ctx.submit_images(list(range(N_CPUS)), streams)
        gpu_step()
        # D2H of this step's results is queued on the context's second stream: it overlaps the
        # next step's H2D (full-duplex PCIe); ctx.sync() below completes the last one
        k = e2e_step.k = (getattr(e2e_step, "k", 0) + 1) % 2
        ctx._ck(ctx.L.hbtgpu_get_buckets_host_async(ctx.h, C.c_void_p(0),
                                                    C.c_void_p(out_scaled[k].data_ptr()),
                                                    C.c_void_p(out_metric[k].data_ptr()),
                                                    C.c_void_p(out_valid[k].data_ptr())))
    e2e_step()
    barrier()
    t_a = time.perf_counter()
    eva, evb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    eva.record(ext)
    for _ in range(args.e2e_steps):
        e2e_step()
    evb.record(ext)
    ctx.sync()
    barrier()
    e2e_wall_ms = 1e3 * (time.perf_counter() - t_a) / args.e2e_steps
    e2e_ms = max(eva.elapsed_time(evb) / args.e2e_steps, e2e_wall_ms)
    d2h = out_scaled[0].numel() * 8 + out_metric[0].numel() * 8 + out_valid[0].numel() * 4

    # ---- max over ranks, whole-job aggregate ---------------------------------------------------
    t = torch.tensor([ms_total, e2e_ms, dec_ms], dtype=torch.float64, device="cuda")
    tot = torch.tensor([float(n_records), float(image_bytes)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    ms_total, e2e_ms, dec_ms = [float(x) for x in t.tolist()]
    all_records, all_bytes = [float(x) for x in tot.tolist()]
    ms_step = ms_total / args.steps

    line = None
    if rank == 0:
        peak, peak_src = measured_peaks()
        achieved = (image_bytes / 1e9) / (dec_ms * 1e-3)  # per GPU: algorithmic bytes / kernel time
        line = {
            "metric": METRIC, "value": all_records / (ms_step * 1e-3), "unit": UNIT,
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64",
            "data": "synthetic",
            "config": {"workload": WORKLOAD, "n_cpus": N_CPUS, "n_events": N_EVENTS,
                       "bucket_ns": BUCKET_NS, "image_bytes_per_gpu": image_bytes,
                       "l2": "inputs (2.0 GB) larger than L2 (126 MB); no flush needed",
                       "step_gbs": (all_bytes / 1e9) / (ms_step * 1e-3)},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": ncu_traffic(), "peak_source": peak_src,
                         "traffic_source": "profiles/r01_traffic.json (ncu --set full, dram read+write per launch)",
                         "kernel": "k_decode<spec,agg>", "kernel_ms": dec_ms,
                         "algorithmic_bytes_per_launch": image_bytes,
                         "frac_of_nominal_8TBs": achieved / 8000.0},
            "e2e": {"value": all_records / (e2e_ms * 1e-3), "unit": UNIT,
                    "h2d_bytes_per_step": image_bytes, "d2h_bytes_per_step": d2h,
                    "ms_per_step": e2e_ms},
            "gpu_launches": int(launches_per_step * args.steps),
            "clocks": clocks,
        }
        if not args.no_cpu_baseline and world == 1:
            lib_kind, kind = pick_cpu_kind()
            cores = os.cpu_count() or 1
            n_threads = min(cores, N_CPUS)
            times, nrec, desc = cpu_reference_run(streams, t0, n_threads, 3, lib_kind)
            best = min(times)
            t1, n1, _ = cpu_reference_run(streams[:8], t0, 1, 1, lib_kind)
            line["cpu_baseline"] = {
                "value": nrec / best, "unit": UNIT, "cores": n_threads, "kind": kind,
                "sample": f"the full capture ({nrec} records), best of 3 passes, {n_threads} threads "
                          f"by CPU partition; {desc}; single-thread: {n1 / t1[0]:.3e} records/s on 8 streams",
                "single_thread_value": n1 / t1[0]}
        print(json.dumps(line), flush=True)
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
