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
    return sorted(set(re.findall(r"\b(hbt(?:gpu|cap)_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(built):
    lib = ctypes.CDLL(built.LIB_PATH)
    decl = declared_symbols()
    assert len(decl) >= 25
    for s in decl:
        assert hasattr(lib, s), f"{s} declared in include/hbtgpu.h but not exported"
    from dynolog_b200 import hbtgpu_cs, hbtgpu_slices
    assert sorted(built.ABI_SYMBOLS + hbtgpu_cs.ABI_SYMBOLS + hbtgpu_slices.ABI_SYMBOLS) == decl, "ctypes binding lists out of sync with the header"


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


def test_cs_struct_sizes_match_header(built, tmp_path):
    from dynolog_b200 import hbtgpu_cs as cs
    src = tmp_path / "szcs.c"
    src.write_text('#include <stdio.h>\n#include "hbtgpu.h"\nint main(){printf("%zu %zu %zu %zu %zu %zu %zu\\n",'
                   "sizeof(hbtgpu_cs_event),sizeof(hbtgpu_cs_op),sizeof(hbtgpu_cs_state),"
                   "sizeof(hbtgpu_cs_config),sizeof(hbtgpu_cs_stats),sizeof(hbtgpu_cs_thread),"
                   "sizeof(hbtgpu_cs_name));return 0;}\n")
    exe = tmp_path / "szcs"
    subprocess.check_call(["gcc", "-std=c11", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    sizes = [int(x) for x in subprocess.check_output([str(exe)]).split()]
    assert sizes == [cs.EVENT_DTYPE.itemsize, cs.OP_DTYPE.itemsize, ctypes.sizeof(cs.CsState),
                     ctypes.sizeof(cs.CsConfig), ctypes.sizeof(cs.CsStats), cs.THREAD_DTYPE.itemsize,
                     cs.NAME_DTYPE.itemsize]


def test_cs_thread_dictionary_on_host(built):
    """hbtgpu_cs_assign_vids is host code (mirror of PerThreadStats): hand-made events and ops."""
    import numpy as np
    from dynolog_b200 import hbtgpu_cs as cs
    T = 5_000_000
    ev = np.zeros(6, cs.EVENT_DTYPE)
    ev["type"] = [4, 5, 2, 4, 3, 4]          # in(T+1), out, create(T+2), in(T+2), destroy(T+2), in(0)
    ev["tag"] = [T + 1, 0xFFFFFFFF, T + 2, T + 2, T + 2, 0]
    ev["tstamp"] = [10, 20, 30, 40, 50, 60]
    pids = np.array([7, 0, 8, 8, 8, 0], np.uint32)
    ops = np.zeros(3, cs.OP_DTYPE)
    ops["type"] = [7, 3, 4]                   # FORK at event 2, COMM before event 3, EXIT at event 4
    ops["event_index"] = [2, 3, 4]
    ops["tstamp"] = [30, 35, 50]
    ops["pid"], ops["tid"], ops["ppid"], ops["ptid"] = 8, T + 2, 1, 1
    ops["name_len"][1] = 8
    ops["name"][1] = b"worker"
    th = cs.Threads()
    th.assign(ev, pids, ops)
    assert ev["tag"].tolist() == [1, 0xFFFFFFFF, 2, 2, 2, 0]
    threads, names = th.table()
    assert threads == [(1, -1, -1, 7, T + 1, 0, 2**64 - 1, 2**64 - 1), (2, 1, 1, 8, T + 2, 1, 30, 50)]
    assert names == [(2, 35, b"worker\0\0", 1)]
    # after the exit the tid is new again; a LOST forgets every active tid
    ev2 = np.zeros(3, cs.EVENT_DTYPE)
    ev2["type"] = [4, 7, 4]
    ev2["tag"] = [T + 2, 0xFFFFFFFF, T + 1]
    ops2 = np.zeros(1, cs.OP_DTYPE)
    ops2["type"], ops2["event_index"] = 2, 1
    th.assign(ev2, np.zeros(3, np.uint32), ops2)
    assert ev2["tag"].tolist() == [3, 0xFFFFFFFF, 4]


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


def test_slices_struct_sizes_match_header(built, tmp_path):
    from dynolog_b200 import hbtgpu_slices as sl
    src = tmp_path / "szsl.c"
    src.write_text('#include <stdio.h>\n#include "hbtgpu.h"\nint main(){printf("%zu %zu %zu %zu %zu %zu %zu\\n",'
                   "sizeof(hbtgpu_slice),sizeof(hbtgpu_islice_unit),sizeof(hbtgpu_islice_view),"
                   "sizeof(hbtgpu_bin_freq),sizeof(hbtgpu_bin_summary),sizeof(hbtgpu_stack_stats),"
                   "sizeof(hbtgpu_slicer_stats));return 0;}\n")
    exe = tmp_path / "szsl"
    subprocess.check_call(["gcc", "-std=c11", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    sizes = [int(x) for x in subprocess.check_output([str(exe)]).split()]
    assert sizes == [sl.SLICE_DTYPE.itemsize, sl.UNIT_DTYPE.itemsize, ctypes.sizeof(sl.View),
                     sl.FREQ_DTYPE.itemsize, ctypes.sizeof(sl.BinSummary), sl.STACK_DTYPE.itemsize,
                     8 * len(sl.SLICER_STAT_KEYS)]
