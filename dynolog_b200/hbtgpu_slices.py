"""ctypes binding of the count-interval x slice intersection in libhbtgpu.so (include/hbtgpu.h,
section "Count intervals x tagstack slices").  Plumbing for tests and probes; no CPU fallback."""
import ctypes as C

import numpy as np

from . import hbtgpu
from .hbtgpu import HbtGpuError

SLICE_DTYPE = np.dtype([("tstamp", "<u8"), ("duration", "<u8"), ("stack_id", "<u8"),
                        ("swin_type", "u1"), ("swout_type", "u1"), ("pad", "u1", (6,))])
UNIT_DTYPE = np.dtype([("n_intervals", "<u8"), ("n_slices", "<u8"), ("mark", "<u8"), ("status", "<i8"),
                       ("err_interval", "<u8"), ("exact_path", "<u4"), ("pad", "<u4")])
FREQ_DTYPE = np.dtype([("stack_id", "<u8"), ("duration", "<u8"), ("num_obs", "<u8"), ("num_intervals", "<u8"),
                       ("bid", "<u4"), ("pad", "<u4")])
assert SLICE_DTYPE.itemsize == 32 and UNIT_DTYPE.itemsize == 48 and FREQ_DTYPE.itemsize == 40

TRIM, ABORT_ON_MISSING, DEVICE_INPUT, FORCE_EXACT = 1, 2, 4, 8

EVENT_DTYPE = np.dtype([("tstamp", "<u8"), ("type", "u1"), ("level", "u1"), ("comp_unit_id", "<u2"), ("tag", "<u4")])
STACK_DTYPE = np.dtype([("stack_id", "<u8"), ("parent_stack_id", "<u8"), ("tags", "<u4", (6,)),
                        ("num_set_tags", "u1"), ("known_num_levels", "u1"), ("pad", "u1", (6,)),
                        ("cum_active_time", "<u8"), ("num_swins", "<u8"), ("num_swouts", "<u8")])
SLICER_STAT_KEYS = ("num_processed", "num_error_start", "num_error_end", "num_error_out_sync", "num_dropped",
                    "num_out_of_order", "num_slices_completed", "num_logic_errors")
assert EVENT_DTYPE.itemsize == 16 and STACK_DTYPE.itemsize == 72

ABI_SYMBOLS = [
    "hbtgpu_islice_create", "hbtgpu_islice_destroy", "hbtgpu_islice_last_error",
    "hbtgpu_islice_release", "hbtgpu_islice_get_units", "hbtgpu_islice_get",
    "hbtgpu_islice_device_view", "hbtgpu_islice_last_ms",
    "hbtgpu_islice_bin", "hbtgpu_islice_bin_with", "hbtgpu_islice_bin_get", "hbtgpu_islice_bin_durations_device",
    "hbtgpu_slicer_create", "hbtgpu_slicer_destroy", "hbtgpu_slicer_last_error", "hbtgpu_slicer_run",
    "hbtgpu_slicer_get_slices", "hbtgpu_slicer_get_stacks", "hbtgpu_slicer_slices_device", "hbtgpu_slicer_last_ms",
]


class BinSummary(C.Structure):
    _fields_ = [("num_rows", C.c_uint64), ("num_bins", C.c_uint64), ("num_stack_ids", C.c_uint64),
                ("rows_skipped", C.c_uint64)]


class View(C.Structure):
    _fields_ = [("slices", C.c_void_p), ("unit_base", C.c_void_p), ("interval_offset", C.c_void_p),
                ("interval_count", C.c_void_p), ("n_units", C.c_uint32)]


_ready = False


def _lib():
    global _ready
    L = hbtgpu.lib()
    if not _ready:
        L.hbtgpu_islice_create.restype = C.c_void_p
        L.hbtgpu_islice_create.argtypes = [C.c_int]
        L.hbtgpu_islice_destroy.restype = None
        L.hbtgpu_islice_destroy.argtypes = [C.c_void_p]
        L.hbtgpu_islice_last_error.restype = C.c_char_p
        L.hbtgpu_islice_last_error.argtypes = [C.c_void_p]
        L.hbtgpu_islice_last_ms.restype = C.c_float
        L.hbtgpu_islice_last_ms.argtypes = [C.c_void_p]
        L.hbtgpu_islice_release.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64,
                                            C.c_void_p, C.c_uint32, C.c_uint64, C.c_uint32, C.c_void_p,
                                            C.c_void_p]
        L.hbtgpu_islice_get_units.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32]
        L.hbtgpu_islice_get.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint64, C.c_void_p,
                                        C.c_uint64]
        L.hbtgpu_islice_device_view.argtypes = [C.c_void_p, C.c_void_p]
        L.hbtgpu_islice_bin.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64,
                                        C.c_void_p]
        L.hbtgpu_islice_bin_get.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64]
        L.hbtgpu_islice_bin_with.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64,
                                             C.c_uint64, C.c_uint64, C.c_uint64, C.c_void_p]
        L.hbtgpu_islice_bin_durations_device.restype = C.c_void_p
        L.hbtgpu_islice_bin_durations_device.argtypes = [C.c_void_p]
        L.hbtgpu_slicer_create.restype = C.c_void_p
        L.hbtgpu_slicer_create.argtypes = [C.c_int]
        L.hbtgpu_slicer_destroy.restype = None
        L.hbtgpu_slicer_destroy.argtypes = [C.c_void_p]
        L.hbtgpu_slicer_last_error.restype = C.c_char_p
        L.hbtgpu_slicer_last_error.argtypes = [C.c_void_p]
        L.hbtgpu_slicer_last_ms.restype = C.c_float
        L.hbtgpu_slicer_last_ms.argtypes = [C.c_void_p]
        L.hbtgpu_slicer_run.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p,
                                        C.c_void_p, C.c_void_p, C.c_void_p]
        L.hbtgpu_slicer_get_slices.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint64]
        L.hbtgpu_slicer_get_stacks.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64]
        L.hbtgpu_slicer_slices_device.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        _ready = True
    return L


class IntervalSlicer:
    """Batch form of tagstack::IntervalSlicer + the releaseCountData_ row loop: all comp units of a
    window in one call."""

    def __init__(self, device=0):
        self.L = _lib()
        self.h = self.L.hbtgpu_islice_create(device)
        if not self.h:
            raise HbtGpuError(-19, "hbtgpu_islice_create failed (no CUDA device? there is no CPU fallback)")
        self.n_units = 0

    def close(self):
        if self.h:
            self.L.hbtgpu_islice_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != 0:
            raise HbtGpuError(rc, self.L.hbtgpu_islice_last_error(self.h).decode())

    def release(self, slices, slice_begin, rows, row_begin, stop_ts, flags, marks, row_stride=None):
        """slices: SLICE_DTYPE array (all units), rows: [n, stride] u64; *_begin: [n_units + 1].
        marks: [n_units] u64, updated in place.  Returns the per-unit summary (UNIT_DTYPE)."""
        slices = np.ascontiguousarray(slices, dtype=SLICE_DTYPE)
        rows = np.ascontiguousarray(rows, dtype=np.uint64)
        if rows.ndim == 1:
            rows = rows.reshape(-1, row_stride or 2)
        slice_begin = np.ascontiguousarray(slice_begin, dtype=np.uint64)
        row_begin = np.ascontiguousarray(row_begin, dtype=np.uint64)
        n_units = len(slice_begin) - 1
        assert len(row_begin) == n_units + 1 and len(marks) == n_units and marks.dtype == np.uint64
        units = np.zeros(n_units, UNIT_DTYPE)
        self._check(self.L.hbtgpu_islice_release(
            self.h, slices.ctypes.data, slice_begin.ctypes.data, rows.ctypes.data, rows.shape[1],
            row_begin.ctypes.data, n_units, stop_ts, flags, marks.ctypes.data, units.ctypes.data))
        self.n_units = n_units
        self._units = units
        return units

    def release_device(self, d_slices, slice_begin, d_rows, row_stride, row_begin, stop_ts, flags, marks):
        """Same with slices / rows already on the device (raw pointers)."""
        slice_begin = np.ascontiguousarray(slice_begin, dtype=np.uint64)
        row_begin = np.ascontiguousarray(row_begin, dtype=np.uint64)
        n_units = len(slice_begin) - 1
        units = np.zeros(n_units, UNIT_DTYPE)
        self._check(self.L.hbtgpu_islice_release(
            self.h, d_slices, slice_begin.ctypes.data, d_rows, row_stride, row_begin.ctypes.data, n_units,
            stop_ts, flags | DEVICE_INPUT, marks.ctypes.data, units.ctypes.data))
        self.n_units = n_units
        self._units = units
        return units

    def units(self):
        u = np.zeros(self.n_units, UNIT_DTYPE)
        self._check(self.L.hbtgpu_islice_get_units(self.h, u.ctypes.data, self.n_units))
        return u

    def get(self, unit):
        """(counts per interval, slices in interval order) of one unit."""
        u = self._units[unit]
        counts = np.zeros(max(1, int(u["n_intervals"])), np.uint32)
        out = np.zeros(max(1, int(u["n_slices"])), SLICE_DTYPE)
        self._check(self.L.hbtgpu_islice_get(self.h, unit, counts.ctypes.data, len(counts), out.ctypes.data,
                                             len(out)))
        return counts[:int(u["n_intervals"])], out[:int(u["n_slices"])]

    def device_view(self):
        v = View()
        self._check(self.L.hbtgpu_islice_device_view(self.h, C.byref(v)))
        return v

    def bin(self, min_distinct_intervals, min_duration, selected=None, max_stack_ids=1 << 20,
            max_matrix_bytes=1 << 30, fetch=True, numbering=None):
        """MonData::stackIdFreqs + TagStackIdBinner over the last release.  Returns dict(freqs,
        durations [num_rows, num_bins], num_rows, num_bins, num_stack_ids, rows_skipped).
        numbering = (stack_ids ascending, bids, num_bins, num_rows): bins agreed between ranks
        (hbtgpu_islice_bin_with) instead of the thresholds."""
        summ = BinSummary()
        sel = None if selected is None else np.ascontiguousarray(selected, dtype=np.uint8)
        sel_p = None if sel is None else sel.ctypes.data
        if numbering is None:
            self._check(self.L.hbtgpu_islice_bin(self.h, sel_p, min_distinct_intervals, min_duration, max_stack_ids,
                                                 max_matrix_bytes, C.byref(summ)))
        else:
            ids = np.ascontiguousarray(numbering[0], dtype=np.uint64)
            bids = np.ascontiguousarray(numbering[1], dtype=np.uint32)
            self._check(self.L.hbtgpu_islice_bin_with(self.h, sel_p, ids.ctypes.data, bids.ctypes.data, len(ids),
                                                      int(numbering[2]), int(numbering[3]), max_stack_ids,
                                                      max_matrix_bytes, C.byref(summ)))
        out = {"num_rows": summ.num_rows, "num_bins": summ.num_bins, "num_stack_ids": summ.num_stack_ids,
               "rows_skipped": summ.rows_skipped, "freqs": None, "durations": None}
        if not fetch:  # frequencies and matrix stay on the device
            return out
        freqs = np.zeros(max(1, summ.num_stack_ids), FREQ_DTYPE)
        cells = summ.num_rows * summ.num_bins
        dur = np.zeros(max(1, cells), np.uint64)
        self._check(self.L.hbtgpu_islice_bin_get(self.h, freqs.ctypes.data, len(freqs), dur.ctypes.data, len(dur)))
        out["freqs"] = freqs[:summ.num_stack_ids]
        out["durations"] = dur[:cells].reshape(summ.num_rows, summ.num_bins)
        return out

    def bin_durations_device(self):
        return self.L.hbtgpu_islice_bin_durations_device(self.h)

    @property
    def last_ms(self):
        return float(self.L.hbtgpu_islice_last_ms(self.h))


class Slicer:
    """tagstack::slice::Slicer over one window of events of all comp units."""

    def __init__(self, device=0):
        self.L = _lib()
        self.h = self.L.hbtgpu_slicer_create(device)
        if not self.h:
            raise HbtGpuError(-19, "hbtgpu_slicer_create failed (no CUDA device? there is no CPU fallback)")

    def close(self):
        if self.h:
            self.L.hbtgpu_slicer_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != 0:
            raise HbtGpuError(rc, self.L.hbtgpu_slicer_last_error(self.h).decode())

    def run(self, unit_events=None, device_events=None, event_begin=None, cont=False, general=False):
        """unit_events: list of EVENT_DTYPE arrays (one per comp unit, stream order), or
        device_events (raw pointer) + event_begin.  general: HBTGPU_SLICER_GENERAL (the path for
        phases and any level; a fresh window that needs it takes it by itself).  Returns
        dict(slices=[per unit], stats, stacks, last_ts)."""
        if device_events is None:
            n_units = len(unit_events)
            event_begin = np.concatenate([[0], np.cumsum([len(e) for e in unit_events])]).astype(np.uint64)
            ev = np.ascontiguousarray(np.concatenate(unit_events) if n_units else np.zeros(0, EVENT_DTYPE),
                                      dtype=EVENT_DTYPE)
            ptr, flags = ev.ctypes.data, 0
            if len(ev) == 0:
                ev = np.zeros(1, EVENT_DTYPE)
                ptr = ev.ctypes.data
        else:
            event_begin = np.ascontiguousarray(event_begin, dtype=np.uint64)
            n_units = len(event_begin) - 1
            ptr, flags = device_events, 1
        if cont:
            flags |= 2
        if general:
            flags |= 4
        stats = np.zeros(len(SLICER_STAT_KEYS), np.uint64)
        n_slices = np.zeros(n_units, np.uint64)
        n_stacks, last_ts = C.c_uint64(0), C.c_uint64(0)
        self._check(self.L.hbtgpu_slicer_run(self.h, ptr, event_begin.ctypes.data, n_units, flags, stats.ctypes.data,
                                             n_slices.ctypes.data, C.byref(n_stacks), C.byref(last_ts)))
        slices = []
        for u in range(n_units):
            out = np.zeros(max(1, int(n_slices[u])), SLICE_DTYPE)
            self._check(self.L.hbtgpu_slicer_get_slices(self.h, u, out.ctypes.data, len(out)))
            slices.append(out[:int(n_slices[u])])
        stacks = np.zeros(max(1, n_stacks.value), STACK_DTYPE)
        self._check(self.L.hbtgpu_slicer_get_stacks(self.h, stacks.ctypes.data, len(stacks)))
        return {"slices": slices, "stats": dict(zip(SLICER_STAT_KEYS, (int(x) for x in stats))),
                "stacks": stacks[:n_stacks.value], "last_ts": last_ts.value, "n_slices": n_slices}

    def slices_device(self):
        a, b = C.c_void_p(0), C.c_void_p(0)
        self._check(self.L.hbtgpu_slicer_slices_device(self.h, C.byref(a), C.byref(b)))
        return a.value, b.value

    @property
    def last_ms(self):
        return float(self.L.hbtgpu_slicer_last_ms(self.h))
