"""Synthetic perf_event captures (ctypes over dynolog_b200/libhbtsynth.so, host only).

Shapes follow SURVEY.md section 8(d) / BASELINE.json `configs`.  A capture is a
list of per-CPU byte streams (numpy uint8 views into one buffer), deterministic
in (config, cpu).  Used by tests and bench.py for BOTH the CUDA path and the
CPU reference, so they decode identical bytes.
"""
import ctypes as C
import os

import numpy as np

PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(PKG, "libhbtsynth.so")
T_START = 1_000_000_000


class Spec(C.Structure):
    _fields_ = [("cfg_id", C.c_uint32), ("n_events", C.c_uint32), ("layout", C.c_uint32),
                ("forge_ppm", C.c_uint32), ("t_start", C.c_uint64), ("dt_ns", C.c_uint64),
                ("dt_jitter", C.c_uint64), ("lost_lo", C.c_uint32), ("lost_hi", C.c_uint32),
                ("throttle_lo", C.c_uint32), ("throttle_hi", C.c_uint32),
                ("mux_epoch", C.c_uint32), ("sample_format", C.c_uint32), ("n_tids", C.c_uint32),
                ("n_ips", C.c_uint32)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise FileNotFoundError(f"{LIB_PATH} not built: run `make -C dynolog_b200/csrc`")
        _lib = C.CDLL(LIB_PATH)
        _lib.hbtsynth_state_size.restype = C.c_uint64
    return _lib


# BASELINE.json configs -> generator parameters
CONFIGS = {
    # cfg1: single-CPU 4 MiB ring, 4-event group (cycles, instructions, LLC-misses, branch-misses)
    1: dict(n_events=4, layout=0, dt_ns=100_000, dt_jitter=5_000, lost=(4000, 6000),
            throttle=(15000, 25000), mux_epoch=0, bucket_ns=1_000_000),
    # cfg2: 192 CPUs, 8-event group, 10 kHz for 10 s, 1 ms buckets; (instr, cycles, llc, ...)
    2: dict(n_events=8, layout=1, dt_ns=100_000, dt_jitter=5_000, lost=(0, 0), throttle=(0, 0),
            mux_epoch=0, bucket_ns=1_000_000),
    # cfg3: multiplexed counters with interleaved LOST/THROTTLE
    3: dict(n_events=8, layout=1, dt_ns=100_000, dt_jitter=5_000, lost=(200, 2000),
            throttle=(1000, 10000), mux_epoch=1000, bucket_ns=1_000_000),
    # cfg4: IP|TID|TIME|READ samples, 4-event group, 4096 (pid,tid) Zipf(1.1), 2^20 IPs Zipf(1.2)
    4: dict(n_events=4, layout=0, dt_ns=100_000, dt_jitter=5_000, lost=(0, 0), throttle=(0, 0),
            mux_epoch=0, bucket_ns=1_000_000, sample_format=1, n_tids=4096, n_ips=1 << 20),
    # cfg5: windows contiguous in time per CPU, f=1, with LOST
    5: dict(n_events=8, layout=1, dt_ns=100_000, dt_jitter=5_000, lost=(200, 2000),
            throttle=(0, 0), mux_epoch=0, bucket_ns=1_000_000),
}


def make_spec(cfg_id, forge_ppm=0, **over):
    c = dict(CONFIGS[cfg_id])
    c.update(over)
    return Spec(cfg_id=cfg_id, n_events=c["n_events"], layout=c["layout"], forge_ppm=forge_ppm,
                t_start=T_START, dt_ns=c["dt_ns"], dt_jitter=c["dt_jitter"],
                lost_lo=c["lost"][0], lost_hi=c["lost"][1], throttle_lo=c["throttle"][0],
                throttle_hi=c["throttle"][1], mux_epoch=c["mux_epoch"],
                sample_format=c.get("sample_format", 0), n_tids=c.get("n_tids", 0),
                n_ips=c.get("n_ips", 0))


class Capture:
    """Generator of successive time windows for a fixed set of CPUs."""

    def __init__(self, spec, n_cpus, first_cpu=0, n_threads=None):
        self.L = lib()
        self.spec = spec
        self.n_cpus = n_cpus
        self.first_cpu = first_cpu
        self.n_threads = n_threads or min(n_cpus, os.cpu_count() or 1)
        self.state_size = int(self.L.hbtsynth_state_size())
        self.states = np.zeros(n_cpus * self.state_size, np.uint8)

    def window(self, n_samples=None, max_bytes_per_cpu=None, out=None, dry=False):
        """Next window: up to n_samples samples (or max_bytes_per_cpu bytes) per CPU.
        `out`: optional uint8 array of n_cpus*cap bytes (e.g. a view of pinned memory).
        Returns list of per-CPU uint8 views.  dry=True: nothing is written, the generator only
        moves past the window (how a rank skips to its own time windows); returns the lengths."""
        n = self.spec.n_events
        rec = 40 + 8 * n + (16 if self.spec.sample_format == 1 else 0)
        if n_samples is None:
            n_samples = (1 << 62)
            cap = int(max_bytes_per_cpu)
        else:
            cap = int(n_samples) * rec + (n_samples // max(1, self.spec.lost_lo or 10**9) + 2) * 24 \
                + (n_samples // max(1, self.spec.throttle_lo or 10**9) + 2) * 64 + 4096
            if max_bytes_per_cpu is not None:
                cap = min(cap, int(max_bytes_per_cpu))
        cap = (cap + 127) // 128 * 128
        lens = np.zeros(self.n_cpus, np.uint64)
        if dry:
            self.L.hbtsynth_window(C.byref(self.spec), C.c_void_p(self.states.ctypes.data),
                                   C.c_uint32(self.first_cpu), C.c_uint32(self.n_cpus),
                                   C.c_uint64(n_samples), C.c_void_p(0), C.c_uint64(cap),
                                   C.c_void_p(lens.ctypes.data), C.c_uint32(self.n_threads))
            return [int(x) for x in lens]
        if out is None:
            out = np.empty(self.n_cpus * cap, np.uint8)
        assert out.size >= self.n_cpus * cap
        self.L.hbtsynth_window(C.byref(self.spec), C.c_void_p(self.states.ctypes.data),
                               C.c_uint32(self.first_cpu), C.c_uint32(self.n_cpus),
                               C.c_uint64(n_samples), C.c_void_p(out.ctypes.data), C.c_uint64(cap),
                               C.c_void_p(lens.ctypes.data), C.c_uint32(self.n_threads))
        return [out[i * cap:i * cap + int(lens[i])] for i in range(self.n_cpus)]


def capture(cfg_id, n_cpus, n_samples=None, max_bytes_per_cpu=None, forge_ppm=0, out=None, **over):
    """One-window capture of a BASELINE config. Returns (spec, [streams])."""
    spec = make_spec(cfg_id, forge_ppm=forge_ppm, **over)
    cap = Capture(spec, n_cpus)
    return spec, cap.window(n_samples=n_samples, max_bytes_per_cpu=max_bytes_per_cpu, out=out)


# --- ContextSwitch-mode captures ------------------------------------------------
class CsSpec(C.Structure):
    _fields_ = [("n_cpus", C.c_uint32), ("n_tids", C.c_uint32), ("life_ppm", C.c_uint32),
                ("comm_ppm", C.c_uint32), ("lost_ppm", C.c_uint32), ("throttle_ppm", C.c_uint32),
                ("t_start", C.c_uint64), ("dt_ns", C.c_uint64), ("seed", C.c_uint64)]


def cs_capture(n_cpus, bytes_per_cpu, n_tids=4096, life_ppm=2000, comm_ppm=500, lost_ppm=20,
               throttle_ppm=10, seed=1, out=None, n_threads=None):
    """Per-CPU context-switch streams of ~bytes_per_cpu bytes each: pairs of 32-byte
    SWITCH_CPU_WIDE records with FORK / COMM / EXIT / LOST / THROTTLE at the given rates per
    million pairs.  Returns a list of uint8 views."""
    L = lib()
    spec = CsSpec(n_cpus, n_tids, life_ppm, comm_ppm, lost_ppm, throttle_ppm, T_START, 20_000, seed)
    cap = (int(bytes_per_cpu) + 127) // 128 * 128
    if out is None:
        out = np.empty(n_cpus * cap, np.uint8)
    lens = np.zeros(n_cpus, np.uint64)
    L.hbtsynth_cs_capture(C.byref(spec), C.c_void_p(out.ctypes.data), C.c_uint64(cap),
                          C.c_void_p(lens.ctypes.data), C.c_uint32(n_threads or min(n_cpus, os.cpu_count() or 1)))
    return [out[i * cap:i * cap + int(lens[i])] for i in range(n_cpus)]
