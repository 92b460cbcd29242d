"""The capture container (hbtcap_*): host-side round trip on the CPU, and on
a GPU the decode of a capture file against the decode of the same images submitted directly, plus
the Logger-shaped sink against hbtgpu_series_reduce."""
import os

import numpy as np
import pytest

import ringgen as rg

hbtgpu = pytest.importorskip("dynolog_b200.hbtgpu")
from dynolog_b200 import hbtcap  # noqa: E402

NAMES = ["instructions", "cycles", "llc_misses", "branch_misses"]


def _windows():
    out = []
    for w in range(3):
        row = []
        for cpu in range(4):
            if w == 1 and cpu == 2:
                row.append(None)  # a cpu without an image in this window
                continue
            d, _ = rg.random_stream(100 * w + cpu, 400, 4, p_lost=0.01, t_start=1_000_000_000 + w * 50_000_000)
            row.append(d)
        out.append(row)
    return out


def test_container_round_trip(tmp_path):
    path = str(tmp_path / "a.hbtcap")
    win = _windows()
    hbtcap.write(path, NAMES, win, time_shift=3, time_mult=7, time_zero=11, tsc_mask=2, wrap_at=1000)
    with hbtcap.Capture(path) as cap:
        assert cap.event_names == NAMES
        assert (cap.desc.n_cpus, cap.desc.n_windows, cap.desc.time_shift, cap.desc.time_mult) == (4, 3, 3, 7)
        assert (cap.desc.time_zero, cap.desc.tsc_mask) == (11, 2)
        for w in range(3):
            for cpu in range(4):
                exp = win[w][cpu] if win[w][cpu] is not None else np.zeros(0, np.uint8)
                got = cap.image(w, cpu)
                assert np.array_equal(got, exp), (w, cpu)
                if len(got):
                    assert got.ctypes.data % 128 == 0  # what the arena wants


def test_container_rejects_damaged_files(tmp_path):
    path = str(tmp_path / "a.hbtcap")
    hbtcap.write(path, NAMES, _windows())
    raw = bytearray(open(path, "rb").read())
    for name, patch in (("magic", lambda b: b.__setitem__(0, 0x58)), ("truncated", lambda b: b.__delitem__(slice(-16, None))),
                        ("index offset", lambda b: b.__setitem__(slice(16, 24), (len(b) - 8).to_bytes(8, "little")))):
        bad = bytearray(raw)
        patch(bad)
        p2 = str(tmp_path / f"bad_{name.replace(' ', '_')}.hbtcap")
        open(p2, "wb").write(bad)
        with pytest.raises(hbtgpu.HbtGpuError):
            hbtcap.Capture(p2)
    with pytest.raises(hbtgpu.HbtGpuError):
        hbtcap.Capture(str(tmp_path / "missing.hbtcap"))


@pytest.mark.gpu
def test_decode_from_a_capture_file_and_log_the_series(tmp_path):
    from oracle import pyoracle as po
    path = str(tmp_path / "a.hbtcap")
    win = _windows()
    hbtcap.write(path, NAMES, win, wrap_at=504)
    n, nb, t0, bns = 4, 200, 1_000_000_000, 1_000_000
    outs = []
    for from_file in (False, True):
        with hbtgpu.Context(n_events=n, max_cpus=4, n_buckets=nb, t0=t0, bucket_ns=bns, arena_bytes=1 << 22) as g:
            c = hbtgpu.make_carry(valid=True, tstamp=t0, last=[0] * n)
            for cpu in range(4):
                g.set_carry(cpu, c)
            if from_file:
                with hbtcap.Capture(path) as cap:
                    cap.submit(g)
            else:
                for row in win:
                    for cpu, a in enumerate(row):
                        if a is not None:
                            g.submit_image(cpu, a)
            g.run()
            g.sync()
            b = g.buckets()
            red = g.series_reduce(0.95, 1_000_000_000, 10_000_000_000)
            logged = hbtcap.log_series(g, NAMES, 0.95, 1_000_000_000, 10_000_000_000) if from_file else None
            outs.append((b, [g.counters(i) for i in range(4)], red, logged))
    for k in ("acc", "scaled"):
        assert np.array_equal(outs[0][0][k], outs[1][0][k]), k
    assert outs[0][1] == outs[1][1]
    # and against the oracle: the whole stream of a cpu is its images in window order
    port = po.load("port")
    p = po.make_params(n, n_buckets=nb, t0=t0, bucket_ns=bns)
    for cpu in range(4):
        whole = np.concatenate([row[cpu] for row in win if row[cpu] is not None])
        exp = port.decode_stream(p, whole, po.make_carry(valid=True, tstamp=t0, last=[0] * n), want_records=False, want_rows=False)
        assert np.array_equal(outs[1][0]["acc"][cpu], exp["buckets"])
    # the sink: one entry per cpu, values of hbtgpu_series_reduce
    red, logged = outs[1][2], outs[1][3]
    assert len(logged) == 4
    for cpu in range(4):
        for e, nm in enumerate(NAMES):
            assert logged[cpu][f"{nm}_p95.cpu{cpu}"] == int(red["percentile"][cpu, e])
            assert logged[cpu][f"{nm}_diff.cpu{cpu}"] == int(red["diff"][cpu, e])
            assert logged[cpu][f"{nm}_avg.cpu{cpu}"] == np.float32(red["avg_f64"][cpu, e])
            assert logged[cpu][f"{nm}_rate.cpu{cpu}"] == np.float32(red["rate_f64"][cpu, e])
