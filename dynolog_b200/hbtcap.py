"""ctypes binding of the capture container and the Logger sink (hbtcap_*, hbtgpu_log_series in
include/hbtgpu.h).  Plumbing for tests and bench.py."""
import ctypes as C

import numpy as np

from . import hbtgpu

NAME_BYTES = 32


class Desc(C.Structure):
    _fields_ = [("n_events", C.c_uint32), ("sample_format", C.c_uint32), ("n_cpus", C.c_uint32),
                ("n_windows", C.c_uint32), ("time_shift", C.c_uint32), ("time_mult", C.c_uint32),
                ("time_zero", C.c_uint64), ("tsc_mask", C.c_uint64),
                ("event_names", (C.c_char * NAME_BYTES) * hbtgpu.MAX_EVENTS)]


def _lib():
    L = hbtgpu.lib()
    L.hbtcap_last_error.restype = C.c_char_p
    L.hbtcap_close.restype = None
    return L


def _ck(rc, what):
    if rc != 0:
        raise hbtgpu.HbtGpuError(rc, what)


def write(path, event_names, windows, sample_format=0, time_shift=0, time_mult=1, time_zero=0, tsc_mask=0,
          wrap_at=None):
    """windows[w][cpu] = uint8 array (or None).  wrap_at: hand every image over as two slices cut
    at this byte (a wrapped ring)."""
    L = _lib()
    d = Desc()
    d.n_events, d.sample_format = len(event_names), sample_format
    d.n_cpus, d.n_windows = len(windows[0]), len(windows)
    d.time_shift, d.time_mult, d.time_zero, d.tsc_mask = time_shift, time_mult, time_zero, tsc_mask
    for i, nm in enumerate(event_names):
        d.event_names[i].value = nm.encode()[:NAME_BYTES - 1]
    h = C.c_void_p()
    _ck(L.hbtcap_create(path.encode(), C.byref(d), C.byref(h)), "hbtcap_create")
    for w, imgs in enumerate(windows):
        for cpu, a in enumerate(imgs):
            if a is None or len(a) == 0:
                continue
            a = np.ascontiguousarray(a, dtype=np.uint8)
            k = min(wrap_at, a.size) if wrap_at else a.size
            _ck(L.hbtcap_append(h, C.c_uint32(w), C.c_uint32(cpu), C.c_void_p(a.ctypes.data), C.c_size_t(k),
                                C.c_void_p(a.ctypes.data + k), C.c_size_t(a.size - k)), "hbtcap_append")
    _ck(L.hbtcap_finish(h), "hbtcap_finish")


class Capture:
    def __init__(self, path):
        self.L = _lib()
        self.h = C.c_void_p()
        _ck(self.L.hbtcap_open(path.encode(), C.byref(self.h)), f"hbtcap_open({path})")
        self.desc = Desc()
        _ck(self.L.hbtcap_describe(self.h, C.byref(self.desc)), "hbtcap_describe")

    @property
    def event_names(self):
        return [self.desc.event_names[i].value.decode() for i in range(self.desc.n_events)]

    def image(self, window, cpu):
        p = C.c_void_p()
        n = C.c_size_t()
        _ck(self.L.hbtcap_image(self.h, C.c_uint32(window), C.c_uint32(cpu), C.byref(p), C.byref(n)), "hbtcap_image")
        if not n.value:
            return np.zeros(0, np.uint8)
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(n.value,))

    def submit(self, ctx, window_lo=0, window_hi=None):
        hi = self.desc.n_windows if window_hi is None else window_hi
        ctx._ck(self.L.hbtcap_submit(self.h, ctx.h, C.c_uint32(window_lo), C.c_uint32(hi)))

    def close(self):
        if self.h:
            self.L.hbtcap_close(self.h)
            self.h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


_LOG_UINT = C.CFUNCTYPE(None, C.c_void_p, C.c_char_p, C.c_uint64)
_LOG_FLOAT = C.CFUNCTYPE(None, C.c_void_p, C.c_char_p, C.c_float)
_FINALIZE = C.CFUNCTYPE(None, C.c_void_p)


class _Logger(C.Structure):
    _fields_ = [("user", C.c_void_p), ("log_uint", _LOG_UINT), ("log_float", _LOG_FLOAT), ("finalize", _FINALIZE)]


def log_series(ctx, event_names, percentage, period_ns, duration_ns, first_bucket=0, count=None):
    """hbtgpu_log_series into a list of dicts, one per finalize() (i.e. per cpu)."""
    entries, cur = [], {}

    def on_uint(_, key, v):
        cur[key.decode()] = int(v)

    def on_float(_, key, v):
        cur[key.decode()] = float(v)

    def on_final(_):
        entries.append(dict(cur))
        cur.clear()

    lg = _Logger(None, _LOG_UINT(on_uint), _LOG_FLOAT(on_float), _FINALIZE(on_final))
    names = (C.c_char_p * len(event_names))(*[n.encode() for n in event_names])
    count = ctx.n_buckets - first_bucket if count is None else count
    ctx._ck(_lib().hbtgpu_log_series(ctx.h, C.byref(lg), names, C.c_uint32(first_bucket), C.c_uint32(count),
                                     C.c_float(percentage), C.c_int64(period_ns), C.c_int64(duration_ns)))
    return entries
