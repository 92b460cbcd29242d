"""ctypes bindings for the CPU checkers (TEST INFRASTRUCTURE ONLY).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs may import this module.  The product package
(dynolog_b200) never does.

Two implementations of oracle/hbtref_api.h:
  * ``load("reference")`` -> oracle/_ref/libhbtref.so: the real hbt code
    (PerfEventsGroup::consume, CpuCountSampleGenerator, GroupReadValues,
    getRateReducer/getAddReducer) compiled from /root/reference.
  * ``load("port")`` -> oracle/libhbtoracle.so: plain-C restatement.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
MAX_EVENTS = 64


class Params(C.Structure):
    _fields_ = [
        ("n_events", C.c_uint32),
        ("n_buckets", C.c_uint32),
        ("tsc_mask", C.c_uint64),
        ("time_shift", C.c_uint32),
        ("time_mult", C.c_uint32),
        ("time_zero", C.c_uint64),
        ("t0", C.c_uint64),
        ("bucket_ns", C.c_uint64),
        ("pre", C.c_uint32),
        ("pad", C.c_uint32),
    ]


class Carry(C.Structure):
    _fields_ = [
        ("valid", C.c_uint64),
        ("tstamp", C.c_uint64),
        ("te", C.c_uint64),
        ("tr", C.c_uint64),
        ("last", C.c_uint64 * MAX_EVENTS),
    ]


class Counters(C.Structure):
    _fields_ = [(k, C.c_uint64) for k in (
        "lost", "comm", "exit", "throttle", "unthrottle", "fork", "sample", "read",
        "aux", "itrace_start", "switch_cpu_wide", "num_records",
        "num_unknown_count_interval", "num_rows", "rows_out_of_range")] + [
        ("err", C.c_int64), ("err_offset", C.c_uint64), ("last_tstamp", C.c_uint64)]

    def as_dict(self):
        return {k: int(getattr(self, k)) for k, _ in self._fields_}


COUNTER_KEYS = [k for k, _ in Counters._fields_]


def make_params(n_events, n_buckets=1, t0=0, bucket_ns=1, tsc_mask=0, time_shift=0,
                time_mult=1, time_zero=0, pre=0):
    return Params(n_events, n_buckets, tsc_mask, time_shift, time_mult, time_zero, t0,
                  bucket_ns, pre, 0)


def make_carry(valid=False, tstamp=0, te=0, tr=0, last=()):
    c = Carry()
    c.valid = 1 if valid else 0
    c.tstamp = tstamp
    c.te = te
    c.tr = tr
    for i, v in enumerate(last):
        c.last[i] = int(v)
    return c


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t)) if a is not None else None


def _u64(a):
    return np.ascontiguousarray(a, dtype=np.uint64)


class Oracle:
    def __init__(self, path, prefix):
        self.path = path
        self.lib = C.CDLL(path)
        self.prefix = prefix
        f = lambda name: getattr(self.lib, prefix + "_" + name)
        self._decode = f("decode_stream")
        self._decode.restype = C.c_int
        self._scale = f("scale_buckets")
        self._get_count = f("get_count")
        self._rate = f("rate")
        self._add = f("add_reduce")
        self._t2k = f("tsc_to_ktime")
        self._k2t = f("ktime_to_tsc")
        self._psm = f("per_second_mega")
        self._rr = f("running_ratio")
        self._run = f("run_capture")
        self._run.restype = C.c_double
        d = f("describe")
        d.restype = C.c_char_p
        self.description = d().decode()

    # --- A1-A5, A7: stream walk ------------------------------------------------
    def decode_stream(self, params, data, carry=None, want_records=True, want_rows=True,
                      want_buckets=True):
        data = np.ascontiguousarray(data, dtype=np.uint8)
        n = params.n_events
        carry = carry if carry is not None else make_carry()
        out = Carry()
        ctr = Counters()
        cap = max(1, len(data) // 8)
        rec_off = np.zeros(cap, np.uint64) if want_records else None
        rec_type = np.zeros(cap, np.uint32) if want_records else None
        rec_size = np.zeros(cap, np.uint32) if want_records else None
        rows = np.zeros((cap, 2 + n), np.uint64) if want_rows else None
        acc = np.zeros((params.n_buckets, n + 3), np.uint64) if want_buckets else None
        self._decode(C.byref(params), C.byref(carry), C.byref(out),
                     _p(data, C.c_uint8), C.c_uint64(len(data)), C.byref(ctr),
                     _p(rec_off, C.c_uint64), _p(rec_type, C.c_uint32),
                     _p(rec_size, C.c_uint32), C.c_uint64(cap if want_records else 0),
                     _p(rows, C.c_uint64), C.c_uint64(cap if want_rows else 0),
                     _p(acc, C.c_uint64))
        res = {"counters": ctr.as_dict(), "carry_out": out}
        if want_records:
            k = ctr.num_records
            res["records"] = (rec_off[:k].copy(), rec_type[:k].copy(), rec_size[:k].copy())
        if want_rows:
            res["rows"] = rows[:ctr.num_rows].copy()
        if want_buckets:
            res["buckets"] = acc
        return res

    def scale_buckets(self, params, acc):
        acc = _u64(acc)
        out = np.zeros((params.n_buckets, params.n_events), np.uint64)
        self._scale(C.byref(params), _p(acc, C.c_uint64), _p(out, C.c_uint64))
        return out

    # --- A6, A8, A9, A10, A13: elementwise ------------------------------------------
    def get_count(self, count, te, tr):
        count, te, tr = _u64(count), _u64(te), _u64(tr)
        out = np.zeros_like(count)
        self._get_count(_p(count, C.c_uint64), _p(te, C.c_uint64), _p(tr, C.c_uint64),
                        _p(out, C.c_uint64), C.c_uint64(count.size))
        return out

    def rate(self, num, den):
        num, den = _u64(num), _u64(den)
        out = np.zeros(num.shape, np.float64)
        valid = np.zeros(num.shape, np.uint8)
        self._rate(_p(num, C.c_uint64), _p(den, C.c_uint64), _p(out, C.c_double),
                   _p(valid, C.c_uint8), C.c_uint64(num.size))
        return out, valid.astype(bool)

    def add_reduce(self, vals):
        vals = _u64(vals)
        n, k = vals.shape
        out = np.zeros(n, np.float64)
        self._add(_p(vals, C.c_uint64), C.c_uint64(k), _p(out, C.c_double), C.c_uint64(n))
        return out

    def tsc_to_ktime(self, params, tsc):
        tsc = _u64(tsc)
        out = np.zeros_like(tsc)
        self._t2k(C.byref(params), _p(tsc, C.c_uint64), _p(out, C.c_uint64),
                  C.c_uint64(tsc.size))
        return out

    def ktime_to_tsc(self, params, kt):
        kt = _u64(kt)
        out = np.zeros_like(kt)
        self._k2t(C.byref(params), _p(kt, C.c_uint64), _p(out, C.c_uint64),
                  C.c_uint64(kt.size))
        return out

    def per_second_mega(self, count, tr):
        count, tr = _u64(count), _u64(tr)
        out = np.zeros(count.shape, np.float32)
        self._psm(_p(count, C.c_uint64), _p(tr, C.c_uint64), _p(out, C.c_float),
                  C.c_uint64(count.size))
        return out

    def running_ratio(self, te, tr):
        te, tr = _u64(te), _u64(tr)
        ratio = np.zeros(te.shape, np.float32)
        ri = np.zeros(te.shape, np.uint64)
        mux = np.zeros(te.shape, np.uint8)
        self._rr(_p(te, C.c_uint64), _p(tr, C.c_uint64), _p(ratio, C.c_float),
                 _p(ri, C.c_uint64), _p(mux, C.c_uint8), C.c_uint64(te.size))
        return ratio, ri, mux.astype(bool)

    # --- BASELINE config 4 (spec extension; port only) ---------------------------------------
    def collect_ip_tid(self, params, data, carry=None):
        """(pid|tid<<32) and ip of every SAMPLE the walk counts, in stream order."""
        data = np.ascontiguousarray(data, dtype=np.uint8)
        carry = carry if carry is not None else make_carry()
        cap = max(1, len(data) // 48)
        tid = np.zeros(cap, np.uint64)
        ips = np.zeros(cap, np.uint64)
        f = getattr(self.lib, self.prefix + "_collect_ip_tid")
        f.restype = C.c_uint64
        k = f(C.byref(params), C.byref(carry), _p(data, C.c_uint8), C.c_uint64(len(data)),
              _p(tid, C.c_uint64), _p(ips, C.c_uint64), C.c_uint64(cap))
        return tid[:k].copy(), ips[:k].copy()

    # --- timed whole-capture CPU path ----------------------------------------------
    def run_capture(self, params, streams, carries=None, n_threads=1, batch=1024,
                    want_buckets=True):
        """streams: list of uint8 arrays (one per CPU). Returns (seconds, acc, counters)."""
        ns = len(streams)
        n = params.n_events
        ptrs = (C.c_void_p * ns)(*[s.ctypes.data for s in streams])
        lens = np.array([len(s) for s in streams], np.uint64)
        cin = (Carry * ns)()
        if carries is not None:
            for i, c in enumerate(carries):
                cin[i] = c
        acc = np.zeros((ns, params.n_buckets, n + 3), np.uint64) if want_buckets else None
        ctrs = (Counters * ns)()
        secs = self._run(C.byref(params), C.c_uint32(ns), ptrs, _p(lens, C.c_uint64), cin,
                         C.c_uint32(n_threads), C.c_uint32(batch), _p(acc, C.c_uint64), ctrs)
        return secs, acc, [c.as_dict() for c in ctrs]



# --- ContextSwitch-mode decode (oracle/hbtcs_api.h) ----------------------------------------------
class CsEvent(C.Structure):
    _fields_ = [("tstamp", C.c_uint64), ("type", C.c_uint8), ("level", C.c_uint8),
                ("comp_unit_id", C.c_uint16), ("tag", C.c_uint32)]


class CsThread(C.Structure):
    _fields_ = [("vid", C.c_uint32), ("ppid", C.c_int32), ("ptid", C.c_int32), ("pid", C.c_int32),
                ("tid", C.c_int32), ("n_names", C.c_uint32), ("fork_ts", C.c_uint64),
                ("end_ts", C.c_uint64)]


class CsName(C.Structure):
    _fields_ = [("vid", C.c_uint32), ("is_exec", C.c_uint8), ("len", C.c_uint8),
                ("pad", C.c_uint8 * 2), ("tstamp", C.c_uint64), ("name", C.c_char * 16)]


class CsState(C.Structure):
    _fields_ = [("last_tstamp", C.c_uint64), ("in_write_errors", C.c_uint32), ("pad", C.c_uint32)]


CS_EVENT_DTYPE = np.dtype([("tstamp", "<u8"), ("type", "u1"), ("level", "u1"),
                           ("comp_unit_id", "<u2"), ("tag", "<u4")])


def cs_run(kind, streams, tid_level=1, batch=256, states=None):
    """Decode per-CPU ContextSwitch-mode streams with the reference ("reference") or the C
    restatement ("port").  Returns dict(events=[structured arrays], counters=[dict],
    states=[(last_tstamp, in_write_errors)], threads=[tuples], names=[tuples])."""
    o = load(kind)
    f = getattr(o.lib, o.prefix + "_cs_run")
    f.restype = C.c_int
    ns = len(streams)
    streams = [np.ascontiguousarray(s, dtype=np.uint8) for s in streams]
    ptrs = (C.c_void_p * ns)(*[s.ctypes.data for s in streams])
    lens = np.array([len(s) for s in streams], np.uint64)
    caps = np.array([len(s) // 16 + 8 for s in streams], np.uint64)   # <= 2 events per >= 32 bytes
    offs = np.concatenate([[0], np.cumsum(caps)[:-1]]).astype(np.uint64)
    ev = np.zeros(int(caps.sum()), CS_EVENT_DTYPE)
    nev = np.zeros(ns, np.uint64)
    ctrs = (Counters * ns)()
    sin = None
    if states is not None:
        sin = (CsState * ns)()
        for i, (lt, we) in enumerate(states):
            sin[i].last_tstamp = lt
            sin[i].in_write_errors = we
    sout = (CsState * ns)()
    tcap = sum(len(s) for s in streams) // 32 + 16
    th = (CsThread * tcap)()
    nm = (CsName * tcap)()
    nt = C.c_uint64(0)
    nn = C.c_uint64(0)
    rc = f(C.c_uint32(ns), ptrs, _p(lens, C.c_uint64), C.c_uint32(tid_level), C.c_uint32(batch), sin, sout,
           C.c_void_p(ev.ctypes.data), _p(offs, C.c_uint64), _p(caps, C.c_uint64), _p(nev, C.c_uint64),
           ctrs, th, C.c_uint64(tcap), C.byref(nt), nm, C.c_uint64(tcap), C.byref(nn))
    assert rc == 0
    return {
        "events": [ev[int(offs[i]):int(offs[i]) + int(nev[i])].copy() for i in range(ns)],
        "counters": [c.as_dict() for c in ctrs],
        "states": [(int(s.last_tstamp), int(s.in_write_errors)) for s in sout],
        "threads": [(t.vid, t.ppid, t.ptid, t.pid, t.tid, t.n_names, t.fork_ts, t.end_ts)
                    for t in th[:nt.value]],
        "names": sorted((n.vid, n.tstamp, bytes(n.name[:0]) + C.string_at(C.addressof(n) + 16, n.len), n.is_exec)
                        for n in nm[:nn.value]),
    }


# --- series reductions (oracle/hbtseries_api.h) ---------------------------------------------------
def series_reduce(kind, series, percentage, period_ns, duration_ns):
    """series: [n_series, len] u64.  Returns dict of per-series arrays (MetricSeries<uint64_t>:
    avg<double>, avg<uint64_t>, percentile, diff, rate<double>, rate<uint64_t>)."""
    o = load(kind)
    f = getattr(o.lib, o.prefix + "_series_reduce")
    f.restype = C.c_int
    series = np.ascontiguousarray(series, dtype=np.uint64)
    ns, ln = series.shape
    out = {"avg_f64": np.zeros(ns, np.float64), "avg_u64": np.zeros(ns, np.uint64),
           "percentile": np.zeros(ns, np.uint64), "diff": np.zeros(ns, np.uint64),
           "rate_f64": np.zeros(ns, np.float64), "rate_u64": np.zeros(ns, np.uint64)}
    rc = f(_p(series, C.c_uint64), C.c_uint64(ns), C.c_uint64(ln), C.c_float(percentage),
           C.c_int64(period_ns), C.c_int64(duration_ns), _p(out["avg_f64"], C.c_double),
           _p(out["avg_u64"], C.c_uint64), _p(out["percentile"], C.c_uint64), _p(out["diff"], C.c_uint64),
           _p(out["rate_f64"], C.c_double), _p(out["rate_u64"], C.c_uint64))
    if rc != 0:
        raise ValueError("series_reduce: bad arguments")
    return out


_PATHS = {
    "reference": (os.path.join(HERE, "_ref", "libhbtref.so"), "hbtref"),
    "reference_nolog": (os.path.join(HERE, "_ref", "libhbtref_nolog.so"), "hbtref"),
    "port": (os.path.join(HERE, "libhbtoracle.so"), "hbtoracle"),
}
_cache = {}


def available(kind):
    return os.path.exists(_PATHS[kind][0])


def load(kind="port"):
    if kind not in _cache:
        path, prefix = _PATHS[kind]
        if not os.path.exists(path):
            raise FileNotFoundError(
                f"{path} missing: run `make -C oracle` (or __graft_entry__.build())")
        _cache[kind] = Oracle(path, prefix)
    return _cache[kind]


# --- tagstack slices (oracle/hbtslices_api.h) -----------------------------------------------------
SLICE_DTYPE = np.dtype([("tstamp", "<u8"), ("duration", "<u8"), ("stack_id", "<u8"),
                        ("swin_type", "u1"), ("swout_type", "u1"), ("pad", "u1", (6,))])
EVENT_DTYPE = np.dtype([("tstamp", "<u8"), ("type", "u1"), ("level", "u1"),
                        ("comp_unit_id", "<u2"), ("tag", "<u4")])
STACK_DTYPE = np.dtype([("stack_id", "<u8"), ("parent_stack_id", "<u8"), ("tags", "<u4", (6,)),
                        ("num_set_tags", "u1"), ("known_num_levels", "u1"), ("pad", "u1", (6,)),
                        ("cum_active_time", "<u8"), ("num_swins", "<u8"), ("num_swouts", "<u8")])
SLICER_STAT_KEYS = ("num_processed", "num_error_start", "num_error_end", "num_error_out_sync",
                    "num_dropped", "num_out_of_order", "num_slices_completed", "num_logic_errors")
TRIM, ABORT_ON_MISSING = 1, 2
TT_NA, TT_ANALYSIS, TT_PREEMPTED, TT_YIELD, TT_PHASE = 0, 1, 2, 3, 4
assert SLICE_DTYPE.itemsize == 32 and EVENT_DTYPE.itemsize == 16 and STACK_DTYPE.itemsize == 72

_SLICES_PATHS = {
    "reference": (os.path.join(HERE, "_ref", "libhbtref_slices.so"), "hbtref"),
    "port": (os.path.join(HERE, "libhbtoracle.so"), "hbtoracle"),
}
_slices_cache = {}


def slices_available(kind):
    return os.path.exists(_SLICES_PATHS[kind][0])


def _slices_lib(kind):
    if kind not in _slices_cache:
        path, prefix = _SLICES_PATHS[kind]
        if not os.path.exists(path):
            raise FileNotFoundError(f"{path} missing: run `make -C oracle`")
        _slices_cache[kind] = (C.CDLL(path), prefix)
    return _slices_cache[kind]


def make_slices(rows):
    """rows: iterable of (tstamp, duration, stack_id, swin, swout)."""
    rows = list(rows)
    a = np.zeros(len(rows), SLICE_DTYPE)
    for i, r in enumerate(rows):
        a[i]["tstamp"], a[i]["duration"], a[i]["stack_id"], a[i]["swin_type"], a[i]["swout_type"] = r
    return a


def slices_as_tuples(a):
    return [(int(s["tstamp"]), int(s["duration"]), int(s["stack_id"]), int(s["swin_type"]),
             int(s["swout_type"])) for s in a]


def islice_make(kind, slices, offset, tstamp, duration, flags):
    """One makeIntervalIntersectingSlices call.  Returns (slices or None, consumed);
    raises ValueError where the reference throws."""
    lib, prefix = _slices_lib(kind)
    f = getattr(lib, prefix + "_islice_make")
    f.restype = C.c_int
    slices = np.ascontiguousarray(slices, dtype=SLICE_DTYPE)
    out = np.zeros(max(1, len(slices)), SLICE_DTYPE)
    n_out, consumed = C.c_uint64(0), C.c_uint64(0)
    rc = f(C.c_void_p(slices.ctypes.data), C.c_uint64(len(slices)), C.c_uint64(offset),
           C.c_uint64(tstamp), C.c_uint64(duration), C.c_uint32(flags),
           C.c_void_p(out.ctypes.data), C.c_uint64(len(out)), C.byref(n_out), C.byref(consumed))
    if rc < 0:
        raise ValueError(f"islice_make: reference throws (rc={rc})")
    if rc == 1:
        return None, 0
    return out[:n_out.value].copy(), consumed.value


def islice_release(kind, slices, rows, stop_ts, flags, mark=0):
    """The releaseCountData_ row loop for one comp unit.  rows: [n, stride>=2] u64.
    Returns dict(rc, counts, slices, mark, n_intervals, err_interval)."""
    lib, prefix = _slices_lib(kind)
    f = getattr(lib, prefix + "_islice_release")
    f.restype = C.c_int
    slices = np.ascontiguousarray(slices, dtype=SLICE_DTYPE)
    rows = np.ascontiguousarray(rows, dtype=np.uint64)
    if rows.ndim == 1:
        rows = rows.reshape(-1, 2)
    n_rows, stride = rows.shape
    counts = np.zeros(max(1, n_rows), np.uint32)
    cap = len(slices) + 2 * n_rows + 1
    out = np.zeros(cap, SLICE_DTYPE)
    m = C.c_uint64(mark)
    n_out, n_int, err = C.c_uint64(0), C.c_uint64(0), C.c_uint64(0)
    rc = f(C.c_void_p(slices.ctypes.data), C.c_uint64(len(slices)), _p(rows, C.c_uint64),
           C.c_uint64(stride), C.c_uint64(n_rows), C.c_uint64(stop_ts), C.c_uint32(flags),
           C.byref(m), _p(counts, C.c_uint32), C.c_void_p(out.ctypes.data), C.c_uint64(cap),
           C.byref(n_out), C.byref(n_int), C.byref(err))
    return {"rc": rc, "counts": counts[:n_int.value].copy(), "slices": out[:n_out.value].copy(),
            "mark": m.value, "n_intervals": n_int.value, "err_interval": err.value}


def slicer_run(kind, events, n_units, cap_per_unit=None, stack_cap=None):
    """Slicer::processNext over events in stream order.  Returns dict(rc, slices=[per unit],
    stats, stacks, last_ts).  kind "port_general": the restatement with the general stack
    dictionary (oracle/hbt_slicer_general.c: any level, phase events)."""
    general = kind == "port_general"
    lib, prefix = _slices_lib("port" if general else kind)
    f = getattr(lib, prefix + ("_slicer_general_run" if general else "_slicer_run"))
    f.restype = C.c_int
    events = np.ascontiguousarray(events, dtype=EVENT_DTYPE)
    cap = cap_per_unit if cap_per_unit is not None else len(events) + 1
    stack_cap = stack_cap if stack_cap is not None else 8 * len(events) + 8
    out = np.zeros((n_units, cap), SLICE_DTYPE)
    n_slices = np.zeros(n_units, np.uint64)
    stats = np.zeros(len(SLICER_STAT_KEYS), np.uint64)
    stacks = np.zeros(stack_cap, STACK_DTYPE)
    n_stacks, last_ts = C.c_uint64(0), C.c_uint64(0)
    rc = f(C.c_void_p(events.ctypes.data), C.c_uint64(len(events)), C.c_uint32(n_units),
           C.c_void_p(out.ctypes.data), C.c_uint64(cap), _p(n_slices, C.c_uint64),
           _p(stats, C.c_uint64), C.c_void_p(stacks.ctypes.data), C.c_uint64(stack_cap),
           C.byref(n_stacks), C.byref(last_ts))
    return {"rc": rc, "slices": [out[u, :int(n_slices[u])].copy() for u in range(n_units)],
            "stats": dict(zip(SLICER_STAT_KEYS, (int(x) for x in stats))),
            "stacks": stacks[:n_stacks.value].copy(), "last_ts": last_ts.value}


FREQ_DTYPE = np.dtype([("stack_id", "<u8"), ("duration", "<u8"), ("num_obs", "<u8"),
                       ("num_intervals", "<u8"), ("bid", "<u4"), ("pad", "<u4")])
assert FREQ_DTYPE.itemsize == 40


def bin_durations(kind, unit_counts, unit_slices, min_distinct_intervals, min_duration):
    """TagStackIdBinner over per-unit PerIntervalSlices.  unit_counts[u]: slices per interval,
    unit_slices[u]: the unit's slices in interval order.  Returns dict(freqs, durations
    [num_rows, num_bins], num_rows, num_bins); bins numbered as `kind` numbers them."""
    lib, prefix = _slices_lib(kind)
    f = getattr(lib, prefix + "_bin")
    f.restype = C.c_int
    counts = np.ascontiguousarray(np.concatenate([np.asarray(c, np.uint32) for c in unit_counts])
                                  if len(unit_counts) else np.zeros(0, np.uint32), dtype=np.uint32)
    begin = np.concatenate([[0], np.cumsum([len(c) for c in unit_counts])]).astype(np.uint64)
    slices = np.ascontiguousarray(np.concatenate(unit_slices) if len(unit_slices) else
                                  np.zeros(0, SLICE_DTYPE), dtype=SLICE_DTYPE)
    assert int(counts.sum()) == len(slices)
    freqs = np.zeros(max(1, len(slices)), FREQ_DTYPE)
    n_freqs, num_rows, num_bins = C.c_uint64(0), C.c_uint64(0), C.c_uint64(0)
    dur_cap = (len(counts) + 1) * (len(slices) + 2)
    durations = np.zeros(dur_cap, np.uint64)
    counts_p = counts if len(counts) else np.zeros(1, np.uint32)
    rc = f(_p(counts_p, C.c_uint32), _p(begin, C.c_uint64), C.c_void_p(slices.ctypes.data),
           C.c_uint32(len(unit_counts)), C.c_uint64(min_distinct_intervals), C.c_uint64(min_duration),
           C.c_void_p(freqs.ctypes.data), C.c_uint64(len(freqs)), C.byref(n_freqs),
           _p(durations, C.c_uint64), C.c_uint64(dur_cap), C.byref(num_rows), C.byref(num_bins))
    if rc != 0:
        raise ValueError(f"bin_durations: rc={rc}")
    nr, nb = num_rows.value, num_bins.value
    return {"freqs": freqs[:n_freqs.value].copy(), "durations": durations[:nr * nb].reshape(nr, nb).copy(),
            "num_rows": nr, "num_bins": nb}
