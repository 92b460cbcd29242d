"""ctypes binding of the ContextSwitch-mode decode in libhbtgpu.so (include/hbtgpu.h, section
"ContextSwitch-mode decode").  Plumbing for tests and probes; no CPU fallback."""
import ctypes as C

import numpy as np

from . import hbtgpu
from .hbtgpu import Counters, HbtGpuError

EVENT_DTYPE = np.dtype([("tstamp", "<u8"), ("type", "u1"), ("level", "u1"),
                        ("comp_unit_id", "<u2"), ("tag", "<u4")])
OP_DTYPE = np.dtype([("tstamp", "<u8"), ("event_index", "<u8"), ("type", "<u4"), ("misc", "<u4"),
                     ("pid", "<i4"), ("ppid", "<i4"), ("tid", "<i4"), ("ptid", "<i4"),
                     ("name_len", "<u4"), ("pad", "<u4"), ("name", "S16")])
THREAD_DTYPE = np.dtype([("vid", "<u4"), ("ppid", "<i4"), ("ptid", "<i4"), ("pid", "<i4"),
                         ("tid", "<i4"), ("n_names", "<u4"), ("fork_ts", "<u8"), ("end_ts", "<u8")])
NAME_DTYPE = np.dtype([("vid", "<u4"), ("is_exec", "u1"), ("len", "u1"), ("pad", "u1", 2),
                       ("tstamp", "<u8"), ("name", "V16")])
assert EVENT_DTYPE.itemsize == 16 and OP_DTYPE.itemsize == 64
assert THREAD_DTYPE.itemsize == 40 and NAME_DTYPE.itemsize == 32

ABI_SYMBOLS = [
    "hbtgpu_cs_create", "hbtgpu_cs_destroy", "hbtgpu_cs_last_error", "hbtgpu_cs_set_state",
    "hbtgpu_cs_submit_image", "hbtgpu_cs_submit_image_slices", "hbtgpu_cs_reset", "hbtgpu_cs_run", "hbtgpu_cs_sync",
    "hbtgpu_cs_get_counters", "hbtgpu_cs_get_state", "hbtgpu_cs_num_events", "hbtgpu_cs_get_events",
    "hbtgpu_cs_get_ops", "hbtgpu_cs_events_device", "hbtgpu_cs_get_stats",
    "hbtgpu_cs_threads_create", "hbtgpu_cs_threads_destroy", "hbtgpu_cs_assign_vids",
    "hbtgpu_cs_threads_count", "hbtgpu_cs_threads_name_count", "hbtgpu_cs_threads_get",
    "hbtgpu_cs_vids_create", "hbtgpu_cs_vids_destroy", "hbtgpu_cs_vids_last_error",
    "hbtgpu_cs_assign_vids_device", "hbtgpu_cs_vids_count", "hbtgpu_cs_vids_name_count",
    "hbtgpu_cs_vids_last_ms", "hbtgpu_cs_vids_get",
]


class CsConfig(C.Structure):
    _fields_ = [("device", C.c_int32), ("max_cpus", C.c_uint32), ("tid_level", C.c_uint32),
                ("flags", C.c_uint32), ("arena_bytes", C.c_uint64)]


class CsState(C.Structure):
    _fields_ = [("last_tstamp", C.c_uint64), ("in_write_errors", C.c_uint32), ("pad", C.c_uint32)]


class CsStats(C.Structure):
    _fields_ = [("image_bytes", C.c_uint64), ("n_chunks", C.c_uint32), ("exact_runs", C.c_uint32),
                ("exact_reason", C.c_uint32), ("exact_chunk", C.c_uint32),
                ("last_count_ms", C.c_float), ("last_emit_ms", C.c_float), ("last_run_ms", C.c_float),
                ("kernel_launches", C.c_uint32)]


_ready = False


def _lib():
    global _ready
    L = hbtgpu.lib()
    if not _ready:
        L.hbtgpu_cs_create.restype = C.c_void_p
        L.hbtgpu_cs_destroy.restype = None
        L.hbtgpu_cs_destroy.argtypes = [C.c_void_p]
        L.hbtgpu_cs_last_error.restype = C.c_char_p
        L.hbtgpu_cs_last_error.argtypes = [C.c_void_p]
        L.hbtgpu_cs_threads_create.restype = C.c_void_p
        L.hbtgpu_cs_threads_destroy.restype = None
        L.hbtgpu_cs_threads_destroy.argtypes = [C.c_void_p]
        L.hbtgpu_cs_threads_count.restype = C.c_uint64
        L.hbtgpu_cs_threads_count.argtypes = [C.c_void_p]
        L.hbtgpu_cs_threads_name_count.restype = C.c_uint64
        L.hbtgpu_cs_threads_name_count.argtypes = [C.c_void_p]
        L.hbtgpu_cs_vids_create.restype = C.c_void_p
        L.hbtgpu_cs_vids_destroy.restype = None
        L.hbtgpu_cs_vids_destroy.argtypes = [C.c_void_p]
        L.hbtgpu_cs_vids_last_error.restype = C.c_char_p
        L.hbtgpu_cs_vids_last_error.argtypes = [C.c_void_p]
        L.hbtgpu_cs_vids_count.restype = C.c_uint64
        L.hbtgpu_cs_vids_count.argtypes = [C.c_void_p]
        L.hbtgpu_cs_vids_name_count.restype = C.c_uint64
        L.hbtgpu_cs_vids_name_count.argtypes = [C.c_void_p]
        L.hbtgpu_cs_vids_last_ms.restype = C.c_float
        L.hbtgpu_cs_vids_last_ms.argtypes = [C.c_void_p]
        _ready = True
    return L


class CsContext:
    def __init__(self, max_cpus, arena_bytes, tid_level=1, flags=0, device=0):
        self.L = _lib()
        cfg = CsConfig(device, max_cpus, tid_level, flags, arena_bytes)
        self.h = self.L.hbtgpu_cs_create(C.byref(cfg))
        if not self.h:
            raise HbtGpuError(-5, self.L.hbtgpu_cs_last_error(None).decode())
        self.max_cpus = max_cpus

    def _ck(self, rc):
        if rc != 0:
            raise HbtGpuError(rc, self.L.hbtgpu_cs_last_error(C.c_void_p(self.h)).decode())

    def close(self):
        if self.h:
            self.L.hbtgpu_cs_destroy(C.c_void_p(self.h))
            self.h = None

    def __del__(self):
        self.close()

    def set_state(self, cpu, last_tstamp, in_write_errors):
        st = CsState(last_tstamp, 1 if in_write_errors else 0, 0)
        self._ck(self.L.hbtgpu_cs_set_state(C.c_void_p(self.h), C.c_uint32(cpu), C.byref(st)))

    def submit(self, cpu, data):
        data = np.ascontiguousarray(data, dtype=np.uint8)
        self._ck(self.L.hbtgpu_cs_submit_image(C.c_void_p(self.h), C.c_uint32(cpu),
                                               C.c_void_p(data.ctypes.data), C.c_uint64(data.size)))

    def reset(self):
        self._ck(self.L.hbtgpu_cs_reset(C.c_void_p(self.h)))

    def run(self):
        self._ck(self.L.hbtgpu_cs_run(C.c_void_p(self.h)))

    def sync(self):
        self._ck(self.L.hbtgpu_cs_sync(C.c_void_p(self.h)))

    def counters(self, cpu):
        c = Counters()
        self._ck(self.L.hbtgpu_cs_get_counters(C.c_void_p(self.h), C.c_uint32(cpu), C.byref(c)))
        return c.as_dict()

    def state(self, cpu):
        st = CsState()
        self._ck(self.L.hbtgpu_cs_get_state(C.c_void_p(self.h), C.c_uint32(cpu), C.byref(st)))
        return int(st.last_tstamp), int(st.in_write_errors)

    def num_events(self, cpu):
        a, b = C.c_uint64(0), C.c_uint64(0)
        self._ck(self.L.hbtgpu_cs_num_events(C.c_void_p(self.h), C.c_uint32(cpu), C.byref(a), C.byref(b)))
        return a.value, b.value

    def events(self, cpu):
        """(events with tag = raw tid, pids)"""
        n, _ = self.num_events(cpu)
        ev = np.zeros(max(n, 1), EVENT_DTYPE)
        pid = np.zeros(max(n, 1), np.uint32)
        got = C.c_uint64(0)
        self._ck(self.L.hbtgpu_cs_get_events(C.c_void_p(self.h), C.c_uint32(cpu), C.c_void_p(ev.ctypes.data),
                                             C.c_void_p(pid.ctypes.data), C.c_uint64(n), C.byref(got)))
        return ev[:n], pid[:n]

    def ops(self, cpu):
        _, n = self.num_events(cpu)
        op = np.zeros(max(n, 1), OP_DTYPE)
        got = C.c_uint64(0)
        self._ck(self.L.hbtgpu_cs_get_ops(C.c_void_p(self.h), C.c_uint32(cpu), C.c_void_p(op.ctypes.data),
                                          C.c_uint64(n), C.byref(got)))
        return op[:n]

    def stats(self):
        s = CsStats()
        self._ck(self.L.hbtgpu_cs_get_stats(C.c_void_p(self.h), C.byref(s)))
        return {k: getattr(s, k) for k, _ in s._fields_}


class Threads:
    """Host mirror of PerThreadStats (hbtgpu_cs_threads)."""

    def __init__(self):
        self.L = _lib()
        self.h = self.L.hbtgpu_cs_threads_create()

    def close(self):
        if self.h:
            self.L.hbtgpu_cs_threads_destroy(C.c_void_p(self.h))
            self.h = None

    def __del__(self):
        self.close()

    def assign(self, events, pids, ops):
        """Rewrites events['tag'] from tid to vid in place."""
        assert events.dtype == EVENT_DTYPE and events.flags.c_contiguous
        ops = np.ascontiguousarray(ops)
        pids = np.ascontiguousarray(pids, dtype=np.uint32)
        rc = self.L.hbtgpu_cs_assign_vids(C.c_void_p(self.h), C.c_void_p(events.ctypes.data),
                                          C.c_void_p(pids.ctypes.data), C.c_uint64(len(events)),
                                          C.c_void_p(ops.ctypes.data), C.c_uint64(len(ops)))
        if rc != 0:
            raise HbtGpuError(rc, "hbtgpu_cs_assign_vids: input outside the reference's contract")

    def table(self):
        nt = self.L.hbtgpu_cs_threads_count(C.c_void_p(self.h))
        nn = self.L.hbtgpu_cs_threads_name_count(C.c_void_p(self.h))
        th = np.zeros(max(nt, 1), THREAD_DTYPE)
        nm = np.zeros(max(nn, 1), NAME_DTYPE)
        rc = self.L.hbtgpu_cs_threads_get(C.c_void_p(self.h), C.c_void_p(th.ctypes.data), C.c_uint64(nt),
                                          C.c_void_p(nm.ctypes.data), C.c_uint64(nn))
        assert rc == 0
        return _table(th, nm, nt, nn)


def _table(th, nm, nt, nn):
    threads = [tuple(int(x) for x in (t["vid"], t["ppid"], t["ptid"], t["pid"], t["tid"], t["n_names"],
                                      t["fork_ts"], t["end_ts"])) for t in th[:nt]]
    names = sorted((int(n["vid"]), int(n["tstamp"]), bytes(n["name"])[:int(n["len"])], int(n["is_exec"]))
                   for n in nm[:nn])
    return threads, names


class DeviceThreads:
    """The thread dictionary kept on the GPU (hbtgpu_cs_vids): assign() rewrites the tags of a
    whole decoded window on the device."""

    def __init__(self, device=0):
        self.L = _lib()
        self.h = self.L.hbtgpu_cs_vids_create(C.c_int(device))
        if not self.h:
            raise HbtGpuError(-5, "hbtgpu_cs_vids_create failed (no CUDA device?)")

    def close(self):
        if self.h:
            self.L.hbtgpu_cs_vids_destroy(C.c_void_p(self.h))
            self.h = None

    def __del__(self):
        self.close()

    def assign(self, ctx):
        rc = self.L.hbtgpu_cs_assign_vids_device(C.c_void_p(ctx.h), C.c_void_p(self.h))
        if rc != 0:
            raise HbtGpuError(rc, self.L.hbtgpu_cs_vids_last_error(C.c_void_p(self.h)).decode())

    def last_ms(self):
        return float(self.L.hbtgpu_cs_vids_last_ms(C.c_void_p(self.h)))

    def count(self):
        return int(self.L.hbtgpu_cs_vids_count(C.c_void_p(self.h)))

    def table(self):
        nt = self.count()
        nn = int(self.L.hbtgpu_cs_vids_name_count(C.c_void_p(self.h)))
        th = np.zeros(max(nt, 1), THREAD_DTYPE)
        nm = np.zeros(max(nn, 1), NAME_DTYPE)
        rc = self.L.hbtgpu_cs_vids_get(C.c_void_p(self.h), C.c_void_p(th.ctypes.data), C.c_uint64(nt),
                                       C.c_void_p(nm.ctypes.data), C.c_uint64(nn))
        assert rc == 0
        return _table(th, nm, nt, nn)


def decode(streams, tid_level=1, states=None, flags=0, ctx=None, device_dictionary=False):
    """Decode per-CPU ContextSwitch-mode streams on the GPU and map tids to virtual ids on the
    host, cpus in index order.  Returns dict(events, counters, states, threads, names, stats)."""
    ns = len(streams)
    own = ctx is None
    if own:
        total = sum(len(s) for s in streams)
        ctx = CsContext(ns, total + 256 * ns + (1 << 16), tid_level=tid_level, flags=flags)
    try:
        ctx.reset()
        for cpu, s in enumerate(streams):
            if states is not None:
                ctx.set_state(cpu, *states[cpu])
            ctx.submit(cpu, s)
        ctx.run()
        ctx.sync()
        th = DeviceThreads() if device_dictionary else Threads()
        if device_dictionary:
            th.assign(ctx)
        out = {"events": [], "counters": [], "states": [], "stats": ctx.stats()}
        for cpu in range(ns):
            ev, pid = ctx.events(cpu)
            ev = ev.copy()
            if not device_dictionary:
                th.assign(ev, pid, ctx.ops(cpu))
            out["events"].append(ev)
            out["counters"].append(ctx.counters(cpu))
            out["states"].append(ctx.state(cpu))
        out["threads"], out["names"] = th.table()
        th.close()
        return out
    finally:
        if own:
            ctx.close()
