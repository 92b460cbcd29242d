"""Hand-built perf_event ring images for tests (pure Python/numpy, small sizes).

Record layouts follow the kernel ABI as hbt overlays it
(hbt/src/perf_event/PerfEventsGroup.h:241-298): every record starts with
perf_event_header {u32 type; u16 misc; u16 size}, all fields little-endian u64
after that; sizes are multiples of 8.
"""
import struct

import numpy as np

LOST, COMM, EXIT, THROTTLE, UNTHROTTLE, FORK, READ, SAMPLE = 2, 3, 4, 5, 6, 7, 8, 9
AUX, ITRACE_START, SWITCH_CPU_WIDE = 11, 12, 15


def header(rtype, size, misc=0):
    return struct.pack("<IHH", rtype, misc, size)


def header_u64(rtype, size, misc=0):
    return rtype | (misc << 32) | (size << 48)


def sample(tstamp, te, tr, values, nr=None, size=None, ip_tid=None):
    """PERF_RECORD_SAMPLE for sample_type=TIME|READ, read_format=GROUP|TE|TR (:281-298).
    ip_tid=(ip, pid, tid): the IP|TID|TIME|READ layout of BASELINE config 4 (spec extension)."""
    n = len(values)
    nr = n if nr is None else nr
    extra = b"" if ip_tid is None else struct.pack("<QII", ip_tid[0], ip_tid[1], ip_tid[2])
    size = 40 + len(extra) + 8 * n if size is None else size
    m = (1 << 64) - 1
    return header(SAMPLE, size) + extra + struct.pack(
        "<4Q", tstamp & m, nr, te & m, tr & m) + struct.pack(f"<{n}Q", *[v & m for v in values])


def read_rec(pid, tid, te, tr, values, size=None):
    """PERF_RECORD_READ (:267-278)."""
    n = len(values)
    size = 40 + 8 * n if size is None else size
    return header(READ, size) + struct.pack("<IIQQQ", pid, tid, n, te, tr) + struct.pack(
        f"<{n}Q", *values)


def lost(num_lost, ident=7):
    """PERF_RECORD_LOST, Sampling mode: {hdr, id, num_lost} = 24 B (:245-255)."""
    return header(LOST, 24) + struct.pack("<QQ", ident, num_lost)


def throttle(tstamp, ident=7, stream_id=9):
    return header(THROTTLE, 32) + struct.pack("<QQQ", tstamp, ident, stream_id)


def unthrottle(tstamp, ident=7, stream_id=9):
    return header(UNTHROTTLE, 32) + struct.pack("<QQQ", tstamp, ident, stream_id)


def other(rtype, payload_words):
    """A record of any other type: header + payload u64 words."""
    size = 8 + 8 * len(payload_words)
    return header(rtype, size) + struct.pack(f"<{len(payload_words)}Q", *payload_words)


def to_u8(b):
    return np.frombuffer(bytes(b), dtype=np.uint8).copy()


class SplitMix64:
    def __init__(self, seed):
        self.s = seed & ((1 << 64) - 1)

    def next(self):
        m = (1 << 64) - 1
        self.s = (self.s + 0x9E3779B97F4A7C15) & m
        z = self.s
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & m
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & m
        return z ^ (z >> 31)

    def uniform(self, lo, hi):
        """integer in [lo, hi]"""
        return lo + self.next() % (hi - lo + 1)


def random_stream(seed, n_records, n_events, *, p_lost=0.0, p_throttle=0.0, p_other=0.0,
                  p_read=0.0, mux=False, forge=0.0, t_start=1_000_000_000, dt=100_000,
                  big_values=False, ip_tid=False):
    """A plausible single-CPU Sampling-mode stream. Returns (bytes, info)."""
    r = SplitMix64(seed)
    out = bytearray()
    ts = t_start
    te = tr = 0
    vals = [0] * n_events
    if big_values:
        vals = [(1 << 63) + r.next() % (1 << 62) for _ in range(n_events)]
    f = 1.0
    for i in range(n_records):
        u = r.next() / 2.0**64
        if u < p_lost:
            out += lost(r.uniform(1, 64))
            continue
        u -= p_lost
        if u < p_throttle:
            out += throttle(ts) + unthrottle(ts + 10)
            continue
        u -= p_throttle
        if u < p_other:
            k = r.uniform(0, 6)
            out += other(r.uniform(16, 40), [r.next() for _ in range(k)])
            continue
        u -= p_other
        if u < p_read:
            out += read_rec(1, 2, te, tr, vals)
            continue
        d = dt + r.uniform(0, dt // 10) - dt // 20
        ts += d
        if mux and i % 50 == 0:
            f = [0.0, 0.25, 0.5, 0.75, 1.0][r.uniform(0, 4)]
        te += d
        tr += int(d * f)
        for e in range(n_events):
            vals[e] = (vals[e] + r.uniform(0, 300_000)) & ((1 << 64) - 1)
        v = list(vals)
        if forge > 0:
            for e in range(n_events):
                if r.next() / 2.0**64 < forge:
                    # a counter value that is bit-identical to a valid header
                    kind = r.uniform(0, 3)
                    v[e] = [header_u64(SAMPLE, 40 + 8 * n_events), header_u64(LOST, 24),
                            header_u64(THROTTLE, 32), header_u64(77, 8 * r.uniform(1, 20))][kind]
                    vals[e] = v[e]
        if ip_tid:
            k = r.uniform(0, 40) ** 2 % 37          # skewed thread choice
            ip = 0xffffffff81000000 + 16 * (r.uniform(0, 30) ** 2 % 211)
            out += sample(ts, te, tr, v, ip_tid=(ip, 100 + k // 4, 500 + k))
        else:
            out += sample(ts, te, tr, v)
    return to_u8(out), {"t_end": ts}


# --- ContextSwitch mode (mode::ContextSwitch, PerfEventsGroup.h:105-191): sample_type TIME|TID
# with sample_id_all, so every record ends in sample_id {u32 pid, tid; u64 time} -------------
MISC_SWITCH_OUT = 1 << 13
MISC_SWITCH_OUT_PREEMPT = 1 << 14
EV_CREATE, EV_DESTROY, EV_IN, EV_OUT_PREEMPT, EV_OUT_YIELD, EV_WE_START, EV_WE_END = 2, 3, 4, 5, 6, 7, 8


def sample_id(pid, tid, ts):
    return struct.pack("<IIQ", pid, tid, ts & ((1 << 64) - 1))


def cs_lost(num_lost, ts, pid=0, tid=0, ident=7):
    return header(LOST, 40) + struct.pack("<QQ", ident, num_lost) + sample_id(pid, tid, ts)


def cs_comm(pid, tid, name, ts, misc=0):
    name = bytes(name) + b"\0"
    name += b"\0" * (-len(name) % 8)
    return header(COMM, 16 + len(name) + 16, misc) + struct.pack("<II", pid, tid) + name + sample_id(pid, tid, ts)


def cs_exit(pid, ppid, tid, ptid, ts, rtype=EXIT):
    return header(rtype, 48) + struct.pack("<IIIIQ", pid, ppid, tid, ptid, ts) + sample_id(pid, tid, ts)


def cs_fork(pid, ppid, tid, ptid, ts):
    return cs_exit(pid, ppid, tid, ptid, ts, rtype=FORK)


def cs_switch(ts, pid, tid, out=False, preempt=False, np_pid=0, np_tid=0):
    misc = (MISC_SWITCH_OUT if out else 0) | (MISC_SWITCH_OUT_PREEMPT if out and preempt else 0)
    return header(SWITCH_CPU_WIDE, 32, misc) + struct.pack("<II", np_pid, np_tid) + sample_id(pid, tid, ts)


def cs_throttle(ts, rtype=THROTTLE, pid=0, tid=0):
    return header(rtype, 48) + struct.pack("<QQQ", ts, 7, 9) + sample_id(pid, tid, ts)


def random_cs_streams(seed, n_cpus, n_records, *, p_lost=0.0, p_life=0.0, p_comm=0.0, p_throttle=0.0,
                      p_other=0.0, n_tids=64, t_start=1_000_000_000, dt=20_000, tid_base=5_000_000):
    """Per-CPU ContextSwitch-mode streams that keep the reference's handlers inside their
    contract: lifecycle records (FORK/EXIT/COMM) of a tid only on its home CPU (tid % n_cpus) in
    fork -> comm* -> exit order with increasing time, tids above pid_max so /proc lookups of the
    test host find nothing.  tid 0 (idle) only switches."""
    r = SplitMix64(seed)
    streams = []
    for cpu in range(n_cpus):
        out = bytearray()
        ts = t_start + cpu * 7
        home = [t for t in range(n_tids) if t % n_cpus == cpu]
        alive = {}
        cur = 0
        for _ in range(n_records):
            ts += dt + r.uniform(0, dt // 4)
            u = r.next() / 2.0**64
            if u < p_lost:
                out += cs_lost(r.uniform(1, 90), ts, 1, cur)
                alive.clear()      # the reference forgets which tids are active
                continue
            u -= p_lost
            if u < p_life and home:
                k = home[r.uniform(0, len(home) - 1)]
                tid = tid_base + k
                pid = tid_base + (k // 4) * 4
                if not alive.get(k):
                    out += cs_fork(pid, 1, tid, 1, ts)
                    alive[k] = True
                else:
                    out += cs_exit(pid, 1, tid, 1, ts)
                    alive[k] = False
                continue
            u -= p_life
            if u < p_comm and home:
                k = home[r.uniform(0, len(home) - 1)]
                name = b"t%d_%d" % (k, r.uniform(0, 99999)) + b"x" * r.uniform(0, 6)
                out += cs_comm(tid_base + (k // 4) * 4, tid_base + k, name[:15], ts)
                continue
            u -= p_comm
            if u < p_throttle:
                out += cs_throttle(ts) + cs_throttle(ts + 5, rtype=UNTHROTTLE)
                continue
            u -= p_throttle
            if u < p_other:
                out += other(r.uniform(16, 40), [r.next() for _ in range(r.uniform(0, 5))])
                continue
            # a context switch: current thread out, next thread in
            nxt = r.uniform(0, n_tids - 1)
            nxt_tid = 0 if nxt == 0 else tid_base + nxt
            out += cs_switch(ts, 1, cur, out=True, preempt=r.uniform(0, 1) == 1, np_tid=nxt_tid)
            ts += 800 + r.uniform(0, 400)
            out += cs_switch(ts, 1, nxt_tid, np_tid=cur)
            cur = nxt_tid
        streams.append(to_u8(out))
    return streams
