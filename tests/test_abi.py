"""CPU-side checks of the boundary: the C-ABI library loads and exports every symbol
include/hbtgpu.h declares; struct layouts in the ctypes mirror match the header."""
import ctypes
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def built():
    from dynolog_b200 import hbtgpu
    if not os.path.exists(hbtgpu.LIB_PATH):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "dynolog_b200", "csrc")])
    return hbtgpu


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "hbtgpu.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(hbtgpu_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(built):
    lib = ctypes.CDLL(built.LIB_PATH)
    decl = declared_symbols()
    assert len(decl) >= 25
    for s in decl:
        assert hasattr(lib, s), f"{s} declared in include/hbtgpu.h but not exported"
    assert sorted(built.ABI_SYMBOLS) == decl, "ctypes binding list out of sync with the header"


def test_struct_sizes_match_header(built, tmp_path):
    """Compile a tiny C program against include/hbtgpu.h and compare sizeof()s."""
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include "hbtgpu.h"\nint main(){printf("%zu %zu %zu %zu %zu %zu %zu\\n",'
                   "sizeof(hbtgpu_config),sizeof(hbtgpu_carry),sizeof(hbtgpu_counters),"
                   "sizeof(hbtgpu_bucket_view),sizeof(hbtgpu_head_stub),sizeof(hbtgpu_stats),"
                   "sizeof(hbtgpu_metric));return 0;}\n")
    exe = tmp_path / "sz"
    subprocess.check_call(["gcc", "-std=c11", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    sizes = [int(x) for x in subprocess.check_output([str(exe)]).split()]
    h = built
    assert sizes == [ctypes.sizeof(h.Config), ctypes.sizeof(h.Carry), ctypes.sizeof(h.Counters),
                     ctypes.sizeof(h.BucketView), ctypes.sizeof(h.HeadStub), ctypes.sizeof(h.Stats),
                     ctypes.sizeof(h.Metric)]


def test_header_is_plain_c(tmp_path):
    src = tmp_path / "c.c"
    src.write_text('#include "hbtgpu.h"\nint main(void){return 0;}\n')
    subprocess.check_call(["gcc", "-std=c99", "-pedantic", "-Werror", "-I", os.path.join(ROOT, "include"),
                           "-c", str(src), "-o", str(tmp_path / "c.o")])


def test_no_cuda_device_fails_loudly(built):
    """Without a GPU (this container) create must fail with -ENODEV: no CPU fallback."""
    import errno
    try:
        import torch
        if torch.cuda.is_available():
            pytest.skip("a CUDA device is present")
    except ImportError:
        pass
    with pytest.raises(built.HbtGpuError) as e:
        built.Context(n_events=4)
    assert e.value.code in (-errno.ENODEV, -errno.EIO)


def test_product_does_not_touch_oracle():
    """The oracle is test infrastructure: nothing under dynolog_b200/ or include/ may mention it."""
    bad = []
    for d in ("dynolog_b200", "include"):
        for dp, _, fs in os.walk(os.path.join(ROOT, d)):
            for f in fs:
                if f.endswith((".py", ".cu", ".cuh", ".c", ".h", ".cpp")):
                    txt = open(os.path.join(dp, f), errors="ignore").read()
                    if re.search(r"\boracle\b|hbtref|pyoracle", txt):
                        bad.append(os.path.join(dp, f))
    assert not bad, f"product sources reference the oracle: {bad}"


def test_synth_matches_oracles(port_oracle):
    """The synthetic generator's bytes decode cleanly and identically on both CPU checkers."""
    import numpy as np
    from dynolog_b200 import synth
    from oracle import pyoracle as po
    for cfg, ncpu, forge in ((1, 1, 0), (2, 2, 0), (3, 2, 0), (3, 1, 20000), (5, 1, 0)):
        spec, streams = synth.capture(cfg, ncpu, max_bytes_per_cpu=256 << 10, forge_ppm=forge)
        n = spec.n_events
        p = po.make_params(n, n_buckets=400, t0=synth.T_START, bucket_ns=1_000_000)
        c = po.make_carry(valid=True, tstamp=synth.T_START, last=[0] * n)
        for s in streams:
            a = port_oracle.decode_stream(p, s, c)
            assert a["counters"]["err"] == 0 and a["counters"]["err_offset"] == len(s)
            assert a["counters"]["num_rows"] > 1000
            if po.available("reference"):
                b = po.load("reference").decode_stream(p, s, c)
                assert a["counters"] == b["counters"]
                assert np.array_equal(a["buckets"], b["buckets"])
                assert np.array_equal(a["rows"], b["rows"])
