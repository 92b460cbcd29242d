"""Helpers for the -m gpu tests: run streams through the C ABI and shape the
results like the oracle's so the same comparison code applies."""
import numpy as np

from dynolog_b200 import hbtgpu
from oracle import pyoracle as po


def to_gpu_carry(c):
    g = hbtgpu.Carry()
    g.valid, g.tstamp, g.te, g.tr = c.valid, c.tstamp, c.te, c.tr
    for i in range(hbtgpu.MAX_EVENTS):
        g.last[i] = c.last[i]
    return g


def gpu_decode(params, streams, carries=None, flags=0, split=None, metrics=(), want_tables=True):
    """streams: list of uint8 arrays (one per cpu).  split: optional list of lists of image
    sizes per stream (images are submitted in order).  Returns (list of per-stream dicts, ctx-level)."""
    n = params.n_events
    total = sum(len(s) for s in streams)
    ctx = hbtgpu.Context(n_events=n, max_cpus=len(streams), n_buckets=params.n_buckets,
                         t0=params.t0, bucket_ns=params.bucket_ns,
                         arena_bytes=total + 4096 * (len(streams) + 8) + (1 << 16),
                         tsc_mask=params.tsc_mask, time_shift=params.time_shift,
                         time_mult=params.time_mult, time_zero=params.time_zero, metrics=metrics,
                         flags=flags,
                         sample_format=(hbtgpu.SAMPLE_IP_TID_TIME_READ if params.pre == 2 else 0))
    try:
        for cpu, s in enumerate(streams):
            if carries is not None:
                ctx.set_carry(cpu, to_gpu_carry(carries[cpu]))
        # interleave images of different cpus to exercise the stream bookkeeping
        cuts = []
        for cpu, s in enumerate(streams):
            sizes = split[cpu] if split is not None else [len(s)]
            assert sum(sizes) == len(s)
            o = 0
            parts = []
            for z in sizes:
                parts.append(s[o:o + z])
                o += z
            cuts.append(parts)
        k = 0
        while any(k < len(p) for p in cuts):
            for cpu, parts in enumerate(cuts):
                if k < len(parts):
                    ctx.submit_image(cpu, parts[k])
            k += 1
        ctx.run()
        ctx.sync()
        b = ctx.buckets()
        out = []
        for cpu in range(len(streams)):
            d = {"counters": ctx.counters(cpu), "carry_out": ctx.carry(cpu),
                 "buckets": b["acc"][cpu], "scaled": b["scaled"][cpu]}
            if want_tables:
                d["records"] = ctx.records(cpu)
                d["rows"] = ctx.rows(cpu)
            out.append(d)
        extra = {"stats": ctx.stats(), "metric": b["metric"], "valid": b["valid"]}
        return out, extra
    finally:
        ctx.close()


def assert_matches_oracle(got, exp, n_events, what="", tables=True):
    """got: gpu_decode per-stream dict; exp: oracle decode_stream dict (+ 'scaled')."""
    from cases import carry_list
    ge = dict(got["counters"])
    ee = dict(exp["counters"])
    assert ge == ee, f"{what}: counters differ\n gpu={ge}\n ref={ee}"
    if tables:
        for a, b, nm in zip(got["records"], exp["records"], ("off", "type", "size")):
            assert np.array_equal(a, b), f"{what}: record table {nm} differs"
        assert np.array_equal(got["rows"], exp["rows"]), f"{what}: rows differ"
    assert np.array_equal(got["buckets"], exp["buckets"]), f"{what}: bucket sums differ"
    if "scaled" in exp:
        assert np.array_equal(got["scaled"], exp["scaled"]), f"{what}: scaled counts differ"
    gc = carry_list(got["carry_out"], n_events)
    ec = exp["carry_out"]
    ec = list(ec) if isinstance(ec, list) else carry_list(ec, n_events)
    if gc[0] == 0 and ec[0] == 0:
        gc[1] = ec[1] = 0
    assert gc == ec, f"{what}: carry-out differs\n gpu={gc[:8]}\n ref={ec[:8]}"
