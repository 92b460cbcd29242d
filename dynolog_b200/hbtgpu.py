"""ctypes binding of libhbtgpu.so (C ABI: include/hbtgpu.h).

This is plumbing for tests and bench.py: the product is the shared library.
There is no CPU fallback -- importing works anywhere, but creating a context
without the built library or without a CUDA device raises.
"""
import ctypes as C
import errno
import os

import numpy as np

PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(PKG, "libhbtgpu.so")

MAX_EVENTS = 64
MAX_METRICS = 8
MAX_METRIC_EVENTS = 8

F_FORCE_EXACT = 0x1
F_FORCE_GENERAL = 0x2
F_EXTERNAL_CARRY = 0x4
F_ACCUMULATE_ATTR = 0x8

SAMPLE_TIME_READ = 0
SAMPLE_IP_TID_TIME_READ = 1

METRIC_RATE = 1
METRIC_ADD = 2
METRIC_PER_SECOND_MEGA = 3


class Metric(C.Structure):
    _fields_ = [("kind", C.c_uint32), ("n_idx", C.c_uint32),
                ("idx", C.c_uint32 * MAX_METRIC_EVENTS)]


class Config(C.Structure):
    _fields_ = [
        ("device", C.c_int32), ("n_events", C.c_uint32), ("tsc_mask", C.c_uint64),
        ("time_shift", C.c_uint32), ("time_mult", C.c_uint32), ("time_zero", C.c_uint64),
        ("max_cpus", C.c_uint32), ("n_buckets", C.c_uint32), ("t0", C.c_uint64),
        ("bucket_ns", C.c_uint64), ("arena_bytes", C.c_uint64), ("n_metrics", C.c_uint32),
        ("flags", C.c_uint32), ("metrics", Metric * MAX_METRICS),
        ("sample_format", C.c_uint32), ("tid_table_log2", C.c_uint32),
        ("ip_table_log2", C.c_uint32), ("reserved", C.c_uint32),
    ]


class Carry(C.Structure):
    _fields_ = [("valid", C.c_uint64), ("tstamp", C.c_uint64), ("te", C.c_uint64),
                ("tr", C.c_uint64), ("last", C.c_uint64 * MAX_EVENTS)]


class HeadStub(C.Structure):
    _fields_ = [("has_sample", C.c_uint64), ("lost_before", C.c_uint64), ("offset", C.c_uint64),
                ("tstamp", C.c_uint64), ("te", C.c_uint64), ("tr", C.c_uint64),
                ("value", C.c_uint64 * MAX_EVENTS)]


COUNTER_KEYS = ["lost", "comm", "exit", "throttle", "unthrottle", "fork", "sample", "read", "aux",
                "itrace_start", "switch_cpu_wide", "num_records", "num_unknown_count_interval",
                "num_rows", "rows_out_of_range", "err", "err_offset", "last_tstamp"]


class Counters(C.Structure):
    _fields_ = [(k, C.c_int64 if k == "err" else C.c_uint64) for k in COUNTER_KEYS]

    def as_dict(self):
        return {k: int(getattr(self, k)) for k in COUNTER_KEYS}


class BucketView(C.Structure):
    _fields_ = [("acc", C.c_void_p), ("scaled", C.c_void_p), ("metric", C.c_void_p),
                ("valid", C.c_void_p), ("max_cpus", C.c_uint32), ("n_buckets", C.c_uint32),
                ("n_events", C.c_uint32), ("n_metrics", C.c_uint32)]


class Stats(C.Structure):
    _fields_ = [("image_bytes", C.c_uint64), ("n_chunks", C.c_uint32), ("exact_runs", C.c_uint32),
                ("kernel_launches", C.c_uint32), ("last_run_ms", C.c_float),
                ("last_decode_ms", C.c_float), ("exact_reason", C.c_uint32),
                ("exact_chunk", C.c_uint32), ("decode_ms_sum", C.c_double),
                ("decode_runs", C.c_uint64), ("last_emit_ms", C.c_float), ("pad", C.c_uint32)]


# every symbol include/hbtgpu.h declares (tests check that the library exports them all)
ABI_SYMBOLS = [
    "hbtgpu_create", "hbtgpu_destroy", "hbtgpu_last_error", "hbtgpu_set_carry",
    "hbtgpu_submit_image", "hbtgpu_submit_images", "hbtgpu_submit_image_device", "hbtgpu_reset", "hbtgpu_run",
    "hbtgpu_sync", "hbtgpu_get_counters", "hbtgpu_get_carry", "hbtgpu_get_rows", "hbtgpu_rows_device", "hbtgpu_row_times_device",
    "hbtgpu_get_records", "hbtgpu_get_tid_histogram", "hbtgpu_get_top_ips", "hbtgpu_attr_clear",
    "hbtgpu_get_buckets_device", "hbtgpu_get_buckets_host", "hbtgpu_get_buckets_host_async",
    "hbtgpu_derive",
    "hbtgpu_get_stubs", "hbtgpu_compose_tail", "hbtgpu_apply_boundaries", "hbtgpu_stub_bytes",
    "hbtgpu_export_stubs_device", "hbtgpu_apply_boundaries_gathered", "hbtgpu_acc_device", "hbtgpu_stream",
    "hbtgpu_peer_export", "hbtgpu_peer_attach", "hbtgpu_peer_attach_local", "hbtgpu_peer_detach",
    "hbtgpu_exchange_boundaries", "hbtgpu_merge", "hbtgpu_attr_merge", "hbtgpu_get_prev_tail",
    "hbtgpu_math_get_count", "hbtgpu_math_rate", "hbtgpu_math_add_reduce",
    "hbtgpu_math_tsc_to_ktime", "hbtgpu_math_ktime_to_tsc", "hbtgpu_math_per_second_mega",
    "hbtgpu_math_running_ratio", "hbtgpu_get_stats", "hbtgpu_version",
    "hbtgpu_decode_image_noalloc", "hbtgpu_noalloc_last_error", "hbtgpu_series_reduce",
    "hbtcap_create", "hbtcap_append", "hbtcap_finish", "hbtcap_open", "hbtcap_describe", "hbtcap_config",
    "hbtcap_image", "hbtcap_submit", "hbtcap_close", "hbtcap_last_error", "hbtgpu_log_series",
]

NOALLOC_MAX_IMAGE_BYTES = 64 << 20
PEER_HANDLE_BYTES = 256

_lib = None


class HbtGpuError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"hbtgpu error {code} ({errno.errorcode.get(-code, '?')}): {msg}")
        self.code = code


def lib():
    """Load libhbtgpu.so; fail loudly if it was never built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise FileNotFoundError(
                f"{LIB_PATH} not built: run __graft_entry__.build() or "
                "`make -C dynolog_b200/csrc`. There is no CPU fallback.")
        L = C.CDLL(LIB_PATH)
        L.hbtgpu_last_error.restype = C.c_char_p
        L.hbtgpu_last_error.argtypes = [C.c_void_p]
        L.hbtgpu_version.restype = C.c_char_p
        L.hbtgpu_stream.restype = C.c_void_p
        L.hbtgpu_stream.argtypes = [C.c_void_p]
        L.hbtgpu_destroy.restype = None
        L.hbtgpu_destroy.argtypes = [C.c_void_p]
        L.hbtgpu_noalloc_last_error.restype = C.c_char_p
        L.hbtgpu_compose_tail.restype = None
        L.hbtgpu_stub_bytes.restype = C.c_size_t
        L.hbtgpu_stub_bytes.argtypes = [C.c_void_p]
        _lib = L
    return _lib


def _ptr(a):
    return C.c_void_p(a.ctypes.data) if a is not None else C.c_void_p(0)


def _u64(a):
    return np.ascontiguousarray(a, dtype=np.uint64)


def make_carry(valid=False, tstamp=0, te=0, tr=0, last=()):
    c = Carry()
    c.valid = 1 if valid else 0
    c.tstamp = tstamp
    c.te = te
    c.tr = tr
    for i, v in enumerate(last):
        c.last[i] = int(v)
    return c


class Context:
    """One hbtgpu context (one CUDA device + one stream)."""

    def __init__(self, n_events, max_cpus=1, n_buckets=1, t0=0, bucket_ns=1, arena_bytes=1 << 20,
                 tsc_mask=0, time_shift=0, time_mult=1, time_zero=0, metrics=(), flags=0, device=0,
                 sample_format=0, tid_table_log2=0, ip_table_log2=0):
        self.L = lib()
        cfg = Config()
        cfg.device = device
        cfg.n_events = n_events
        cfg.tsc_mask = tsc_mask
        cfg.time_shift = time_shift
        cfg.time_mult = time_mult
        cfg.time_zero = time_zero
        cfg.max_cpus = max_cpus
        cfg.n_buckets = n_buckets
        cfg.t0 = t0
        cfg.bucket_ns = bucket_ns
        cfg.arena_bytes = arena_bytes
        cfg.flags = flags
        cfg.sample_format = sample_format
        cfg.tid_table_log2 = tid_table_log2
        cfg.ip_table_log2 = ip_table_log2
        cfg.n_metrics = len(metrics)
        for m, (kind, idx) in enumerate(metrics):
            cfg.metrics[m].kind = kind
            cfg.metrics[m].n_idx = len(idx)
            for k, i in enumerate(idx):
                cfg.metrics[m].idx[k] = i
        self.cfg = cfg
        self.h = C.c_void_p()
        rc = self.L.hbtgpu_create(C.byref(cfg), C.byref(self.h))
        if rc != 0:
            raise HbtGpuError(rc, self.L.hbtgpu_last_error(None).decode())
        self.n_events, self.max_cpus, self.n_buckets = n_events, max_cpus, n_buckets
        self.n_metrics = len(metrics)

    def _ck(self, rc):
        if rc != 0:
            raise HbtGpuError(rc, self.L.hbtgpu_last_error(self.h).decode())

    def close(self):
        if self.h:
            self.L.hbtgpu_destroy(self.h)
            self.h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- input
    def set_carry(self, cpu, carry):
        self._ck(self.L.hbtgpu_set_carry(self.h, C.c_uint32(cpu), C.byref(carry)))

    def submit_image(self, cpu, first, second=None):
        first = np.ascontiguousarray(first, dtype=np.uint8)
        n2 = 0
        p2 = C.c_void_p(0)
        if second is not None:
            second = np.ascontiguousarray(second, dtype=np.uint8)
            n2 = second.size
            p2 = _ptr(second)
        self._ck(self.L.hbtgpu_submit_image(self.h, C.c_uint32(cpu), _ptr(first),
                                            C.c_size_t(first.size), p2, C.c_size_t(n2)))

    def submit_images(self, cpus, arrays):
        """One call for many images (array order = submission order)."""
        n = len(cpus)
        arrays = [np.ascontiguousarray(a, dtype=np.uint8) for a in arrays]
        c = (C.c_uint32 * n)(*cpus)
        ptrs = (C.c_void_p * n)(*[a.ctypes.data for a in arrays])
        sizes = (C.c_size_t * n)(*[a.size for a in arrays])
        self._ck(self.L.hbtgpu_submit_images(self.h, C.c_size_t(n), c, ptrs, sizes))

    def submit_image_ptr(self, cpu, host_ptr, nbytes):
        self._ck(self.L.hbtgpu_submit_image(self.h, C.c_uint32(cpu), C.c_void_p(host_ptr),
                                            C.c_size_t(nbytes), C.c_void_p(0), C.c_size_t(0)))

    def submit_image_device(self, cpu, device_ptr, nbytes):
        self._ck(self.L.hbtgpu_submit_image_device(self.h, C.c_uint32(cpu), C.c_void_p(device_ptr),
                                                   C.c_size_t(nbytes)))

    def reset(self, keep_carry=False):
        self._ck(self.L.hbtgpu_reset(self.h, C.c_int(1 if keep_carry else 0)))

    # ---- hot path
    def run(self):
        self._ck(self.L.hbtgpu_run(self.h))

    def sync(self):
        self._ck(self.L.hbtgpu_sync(self.h))

    def derive(self):
        self._ck(self.L.hbtgpu_derive(self.h))

    # ---- results
    def counters(self, cpu):
        c = Counters()
        self._ck(self.L.hbtgpu_get_counters(self.h, C.c_uint32(cpu), C.byref(c)))
        return c.as_dict()

    def carry(self, cpu):
        c = Carry()
        self._ck(self.L.hbtgpu_get_carry(self.h, C.c_uint32(cpu), C.byref(c)))
        return c

    def rows(self, cpu):
        n = C.c_size_t(0)
        self._ck(self.L.hbtgpu_get_rows(self.h, C.c_uint32(cpu), C.c_void_p(0), C.c_size_t(0),
                                        C.byref(n)))
        out = np.zeros((n.value, self.n_events + 2), np.uint64)
        if n.value:
            self._ck(self.L.hbtgpu_get_rows(self.h, C.c_uint32(cpu), _ptr(out),
                                            C.c_size_t(n.value), C.byref(n)))
        return out

    def rows_device(self):
        """(device pointer, host row_begin [max_cpus + 1], words per row) of the row table."""
        ptr = C.c_void_p(0)
        begin = np.zeros(self.max_cpus + 1, np.uint64)
        stride = C.c_uint32(0)
        self._ck(self.L.hbtgpu_rows_device(self.h, C.byref(ptr), C.c_void_p(begin.ctypes.data), C.byref(stride)))
        return ptr.value, begin, stride.value

    def row_times_device(self):
        """device pointer to {tstamp, duration} of every row, two words per row (row_begin as rows_device)."""
        ptr = C.c_void_p(0)
        self._ck(self.L.hbtgpu_row_times_device(self.h, C.byref(ptr), C.c_void_p(0)))
        return ptr.value

    def records(self, cpu):
        n = C.c_size_t(0)
        z = C.c_void_p(0)
        self._ck(self.L.hbtgpu_get_records(self.h, C.c_uint32(cpu), z, z, z, C.c_size_t(0),
                                           C.byref(n)))
        off = np.zeros(n.value, np.uint64)
        typ = np.zeros(n.value, np.uint32)
        siz = np.zeros(n.value, np.uint32)
        if n.value:
            self._ck(self.L.hbtgpu_get_records(self.h, C.c_uint32(cpu), _ptr(off), _ptr(typ),
                                               _ptr(siz), C.c_size_t(n.value), C.byref(n)))
        return off, typ, siz

    def tid_histogram(self):
        """IP|TID format: dict {(pid, tid): samples} over all cpus."""
        n = C.c_size_t(0)
        z = C.c_void_p(0)
        self._ck(self.L.hbtgpu_get_tid_histogram(self.h, z, z, z, C.c_size_t(0), C.byref(n)))
        pid = np.zeros(n.value, np.uint32)
        tid = np.zeros(n.value, np.uint32)
        cnt = np.zeros(n.value, np.uint64)
        if n.value:
            self._ck(self.L.hbtgpu_get_tid_histogram(self.h, _ptr(pid), _ptr(tid), _ptr(cnt),
                                                     C.c_size_t(n.value), C.byref(n)))
        return {(int(a), int(b)): int(c) for a, b, c in zip(pid, tid, cnt)}

    def top_ips(self, k):
        """IP|TID format: [(ip, samples)] of the k most-sampled IPs, (count desc, ip asc)."""
        ip = np.zeros(k, np.uint64)
        cnt = np.zeros(k, np.uint64)
        n = C.c_size_t(0)
        self._ck(self.L.hbtgpu_get_top_ips(self.h, C.c_uint32(k), _ptr(ip), _ptr(cnt), C.byref(n)))
        return [(int(a), int(b)) for a, b in zip(ip[:n.value], cnt[:n.value])]

    def attr_clear(self):
        self._ck(self.L.hbtgpu_attr_clear(self.h))

    def buckets(self, want_acc=True, want_scaled=True, want_metric=True, want_valid=True):
        """Returns dict of host arrays: acc[cpu,bucket,n+3], scaled[cpu,bucket,n], metric, valid."""
        s = (self.max_cpus, self.n_buckets)
        acc = np.zeros(s + (self.n_events + 3,), np.uint64) if want_acc else None
        scaled = np.zeros(s + (self.n_events,), np.uint64) if want_scaled else None
        metric = np.zeros(s + (self.n_metrics,), np.float64) if (
            want_metric and self.n_metrics) else None
        valid = np.zeros(s, np.uint32) if want_valid else None
        self._ck(self.L.hbtgpu_get_buckets_host(self.h, _ptr(acc), _ptr(scaled), _ptr(metric),
                                                _ptr(valid)))
        return {"acc": acc, "scaled": scaled, "metric": metric, "valid": valid}

    def series_reduce(self, percentage, period_ns, duration_ns, first_bucket=0, count=None):
        """MetricSeries reductions of every (cpu, event) series of scaled counts over the buckets
        [first_bucket, first_bucket + count).  Returns dict of [max_cpus, n_events] arrays."""
        count = self.n_buckets - first_bucket if count is None else count
        s = (self.max_cpus, self.n_events)
        out = {"avg_f64": np.zeros(s, np.float64), "avg_u64": np.zeros(s, np.uint64),
               "percentile": np.zeros(s, np.uint64), "diff": np.zeros(s, np.uint64),
               "rate_f64": np.zeros(s, np.float64), "rate_u64": np.zeros(s, np.uint64)}
        self._ck(self.L.hbtgpu_series_reduce(self.h, C.c_uint32(first_bucket), C.c_uint32(count),
                                             C.c_float(percentage), C.c_int64(period_ns), C.c_int64(duration_ns),
                                             _ptr(out["avg_f64"]), _ptr(out["avg_u64"]), _ptr(out["percentile"]),
                                             _ptr(out["diff"]), _ptr(out["rate_f64"]), _ptr(out["rate_u64"])))
        return out

    def bucket_view(self):
        v = BucketView()
        self._ck(self.L.hbtgpu_get_buckets_device(self.h, C.byref(v)))
        return v

    def acc_device(self):
        p = C.c_void_p()
        n = C.c_size_t()
        self._ck(self.L.hbtgpu_acc_device(self.h, C.byref(p), C.byref(n)))
        return p.value, n.value

    def stream(self):
        return self.L.hbtgpu_stream(self.h)

    def stats(self):
        s = Stats()
        self._ck(self.L.hbtgpu_get_stats(self.h, C.byref(s)))
        return {k: getattr(s, k) for k, _ in Stats._fields_}

    def get_stubs(self):
        """Shard mode: (head stubs, tails) as ctypes arrays of max_cpus entries."""
        heads = (HeadStub * self.max_cpus)()
        tails = (Carry * self.max_cpus)()
        self._ck(self.L.hbtgpu_get_stubs(self.h, heads, tails))
        return heads, tails

    def stub_bytes(self):
        return int(self.L.hbtgpu_stub_bytes(self.h))

    def export_stubs_device(self, device_dst):
        self._ck(self.L.hbtgpu_export_stubs_device(self.h, C.c_void_p(device_dst)))

    def apply_boundaries_gathered(self, device_gathered, world, rank):
        self._ck(self.L.hbtgpu_apply_boundaries_gathered(self.h, C.c_void_p(device_gathered),
                                                         C.c_uint32(world), C.c_uint32(rank)))

    # ---- exchange and merge over peer memory (one process per GPU, or contexts of one process)
    def peer_export(self, world):
        """This rank's handle (bytes) for hbtgpu_peer_attach on the other ranks."""
        buf = C.create_string_buffer(PEER_HANDLE_BYTES)
        self._ck(self.L.hbtgpu_peer_export(self.h, C.c_uint32(world), buf))
        return buf.raw

    def peer_attach(self, world, rank, handles):
        """handles: list of `world` byte strings from peer_export, in rank order."""
        blob = b"".join(handles)
        assert len(blob) == world * PEER_HANDLE_BYTES
        self._ck(self.L.hbtgpu_peer_attach(self.h, C.c_uint32(world), C.c_uint32(rank), blob))

    def peer_attach_local(self, world, rank, ctxs):
        """ctxs: the `world` Context objects of this process, in rank order."""
        arr = (C.c_void_p * world)(*[c.h for c in ctxs])
        self._ck(self.L.hbtgpu_peer_attach_local(self.h, C.c_uint32(world), C.c_uint32(rank), arr))

    def peer_detach(self):
        self._ck(self.L.hbtgpu_peer_detach(self.h))

    def exchange_boundaries(self, no_derive=False):
        self._ck(self.L.hbtgpu_exchange_boundaries(self.h, C.c_uint32(1 if no_derive else 0)))

    def prev_tail(self, cpu):
        """Shard mode: the state in front of this rank's shard (after the boundaries were applied)."""
        c = Carry()
        self._ck(self.L.hbtgpu_get_prev_tail(self.h, C.c_uint32(cpu), C.byref(c)))
        return c

    def merge(self):
        self._ck(self.L.hbtgpu_merge(self.h))

    def attr_merge(self):
        """IP|TID format: the (pid,tid) histogram and the IP counts of all ranks (peer memory)."""
        self._ck(self.L.hbtgpu_attr_merge(self.h))

    def apply_boundaries(self, prev_tails):
        """prev_tails: ctypes array (Carry * max_cpus): state before this shard, per cpu."""
        self._ck(self.L.hbtgpu_apply_boundaries(self.h, prev_tails))

    # ---- elementwise arithmetic
    def get_count(self, count, te, tr):
        count, te, tr = _u64(count), _u64(te), _u64(tr)
        out = np.zeros_like(count)
        self._ck(self.L.hbtgpu_math_get_count(self.h, _ptr(count), _ptr(te), _ptr(tr), _ptr(out),
                                              C.c_size_t(count.size)))
        return out

    def rate(self, num, den):
        num, den = _u64(num), _u64(den)
        out = np.zeros(num.shape, np.float64)
        valid = np.zeros(num.shape, np.uint8)
        self._ck(self.L.hbtgpu_math_rate(self.h, _ptr(num), _ptr(den), _ptr(out), _ptr(valid),
                                         C.c_size_t(num.size)))
        return out, valid.astype(bool)

    def add_reduce(self, vals):
        vals = _u64(vals)
        n, k = vals.shape
        out = np.zeros(n, np.float64)
        self._ck(self.L.hbtgpu_math_add_reduce(self.h, _ptr(vals), C.c_size_t(k), _ptr(out),
                                               C.c_size_t(n)))
        return out

    def tsc_to_ktime(self, tsc):
        tsc = _u64(tsc)
        out = np.zeros_like(tsc)
        self._ck(self.L.hbtgpu_math_tsc_to_ktime(self.h, _ptr(tsc), _ptr(out), C.c_size_t(tsc.size)))
        return out

    def ktime_to_tsc(self, kt):
        kt = _u64(kt)
        out = np.zeros_like(kt)
        self._ck(self.L.hbtgpu_math_ktime_to_tsc(self.h, _ptr(kt), _ptr(out), C.c_size_t(kt.size)))
        return out

    def per_second_mega(self, count, tr):
        count, tr = _u64(count), _u64(tr)
        out = np.zeros(count.shape, np.float32)
        self._ck(self.L.hbtgpu_math_per_second_mega(self.h, _ptr(count), _ptr(tr), _ptr(out),
                                                    C.c_size_t(count.size)))
        return out

    def running_ratio(self, te, tr):
        te, tr = _u64(te), _u64(tr)
        ratio = np.zeros(te.shape, np.float32)
        ri = np.zeros(te.shape, np.uint64)
        mux = np.zeros(te.shape, np.uint8)
        self._ck(self.L.hbtgpu_math_running_ratio(self.h, _ptr(te), _ptr(tr), _ptr(ratio), _ptr(ri),
                                                  _ptr(mux), C.c_size_t(te.size)))
        return ratio, ri, mux.astype(bool)


def decode_image_noalloc(cfg, device_image, image_bytes, device_acc, carry_in=None, stream=0):
    """hbtgpu_decode_image_noalloc: for callers that may not allocate (see include/hbtgpu.h).
    cfg: a Config; device_image / device_acc: raw device pointers (ints) the caller owns.
    Returns (counters dict, carry_out)."""
    L = lib()
    ctr = Counters()
    cout = Carry()
    rc = L.hbtgpu_decode_image_noalloc(
        C.byref(cfg), C.byref(carry_in) if carry_in is not None else C.c_void_p(0),
        C.c_void_p(device_image), C.c_size_t(image_bytes), C.c_void_p(device_acc), C.byref(ctr),
        C.byref(cout), C.c_void_p(stream))
    if rc != 0:
        raise HbtGpuError(rc, L.hbtgpu_noalloc_last_error().decode())
    return ctr.as_dict(), cout


def compose_tail(prev, head, own_tail):
    """hbtgpu_compose_tail (host only): state after a shard."""
    out = Carry()
    lib().hbtgpu_compose_tail(C.byref(prev), C.byref(head), C.byref(own_tail), C.byref(out))
    return out
