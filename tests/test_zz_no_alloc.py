"""-m gpu: hbtgpu_decode_image_noalloc against the context-based entry points.

The no-alloc entry point exists for callers that cannot call cudaMalloc: an always-on
monitor that shares the GPU with a training job which owns nearly all device memory (an
allocation there may fail at any time, and cudaMalloc/cudaFree stall the whole device), or a
plugin living inside someone else's memory budget.  It decodes one ring image into a
caller-owned accumulator using a static __device__ scratch array inside the library.

This file is named so that pytest collects it last: a run that stops at the first failure has
then already run every other test.  The check is the library's own bar for integer work:
bit-exact equality with what hbtgpu_submit_image + hbtgpu_run produce for the same bytes, on
two calls in a row (the second with a smaller image, so stale scratch contents would show).
"""
import errno

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
hbtgpu = pytest.importorskip("dynolog_b200.hbtgpu")


def _images():
    from dynolog_b200 import synth
    # multiplexed counters with LOST / THROTTLE records (BASELINE config 3 shape), two sizes
    _, a = synth.capture(3, 1, max_bytes_per_cpu=3 << 20)
    spec = synth.make_spec(3, forge_ppm=5000)
    cap = synth.Capture(spec, 1, first_cpu=7)
    b = cap.window(max_bytes_per_cpu=600 << 10)
    return [np.ascontiguousarray(a[0]), np.ascontiguousarray(b[0])]


def _via_context(data, n_events, n_buckets, t0, bucket_ns, carry):
    with hbtgpu.Context(n_events=n_events, max_cpus=1, n_buckets=n_buckets, t0=t0,
                        bucket_ns=bucket_ns, arena_bytes=len(data) + (1 << 16)) as g:
        g.set_carry(0, carry)
        g.submit_image(0, data)
        g.run()
        g.sync()
        return g.buckets(want_scaled=False, want_metric=False, want_valid=False)["acc"][0].copy(), \
            g.counters(0), g.carry(0)


def _carry_tuple(c, n):
    v = int(c.valid)
    return (v, int(c.tstamp) if v else 0, int(c.te), int(c.tr), tuple(int(c.last[i]) for i in range(n)))


def test_noalloc_matches_context_twice_in_a_row():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from dynolog_b200 import synth
    n, nb, t0, bns = 8, 4096, synth.T_START, 1_000_000
    carry = hbtgpu.make_carry(valid=True, tstamp=t0, last=[0] * n)
    cfg = hbtgpu.Config()
    cfg.device, cfg.n_events, cfg.n_buckets, cfg.t0, cfg.bucket_ns = 0, n, nb, t0, bns
    cfg.time_mult, cfg.max_cpus = 1, 1
    acc_dev = torch.empty(nb * (n + 3), dtype=torch.int64, device="cuda:0")
    free_after_first = None
    for k, data in enumerate(_images()):
        exp_acc, exp_ctr, exp_carry = _via_context(data, n, nb, t0, bns, carry)
        img_dev = torch.from_numpy(data).to("cuda:0")
        acc_dev.fill_(-1)  # the entry point must overwrite, not accumulate
        torch.cuda.synchronize()
        if k == 1:
            free_before = torch.cuda.mem_get_info(0)[0]
        ctr, cout = hbtgpu.decode_image_noalloc(cfg, img_dev.data_ptr(), len(data),
                                                acc_dev.data_ptr(), carry_in=carry)
        if k == 1:
            # the second call on a device must not have taken any device memory
            assert torch.cuda.mem_get_info(0)[0] == free_before
        got = acc_dev.cpu().numpy().view(np.uint64).reshape(nb, n + 3)
        assert ctr == exp_ctr, f"call {k}: counters differ\n noalloc={ctr}\n context={exp_ctr}"
        assert ctr["num_rows"] > 1000 and ctr["lost"] > 0
        assert np.array_equal(got, exp_acc), f"call {k}: bucket sums differ"
        assert _carry_tuple(cout, n) == _carry_tuple(exp_carry, n), f"call {k}: carry-out differs"
        del img_dev


@pytest.mark.parametrize("n", [8, 4, 3])
def test_noalloc_matches_the_oracle(n):
    """The same entry point against the CPU oracle directly (not only against the context path),
    for the group sizes that have a specialised kernel build (8, 4) and one that has not (3)."""
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from dynolog_b200 import synth
    from oracle import pyoracle as po
    import ringgen as rg
    nb, t0, bns = 3000, synth.T_START, 1_000_000
    data, _ = rg.random_stream(77 + n, 20_000, n, p_lost=0.002, p_throttle=0.002, mux=True)
    port = po.load("port")
    params = po.make_params(n, n_buckets=nb, t0=t0, bucket_ns=bns)
    exp = port.decode_stream(params, data, po.make_carry(valid=True, tstamp=t0, last=[0] * n),
                             want_records=False, want_rows=False)
    cfg = hbtgpu.Config()
    cfg.device, cfg.n_events, cfg.n_buckets, cfg.t0, cfg.bucket_ns = 0, n, nb, t0, bns
    cfg.time_mult, cfg.max_cpus = 1, 1
    acc_dev = torch.empty(nb * (n + 3), dtype=torch.int64, device="cuda:0")
    img_dev = torch.from_numpy(data).to("cuda:0")
    for _ in range(2):
        acc_dev.fill_(-1)
        torch.cuda.synchronize()
        ctr, cout = hbtgpu.decode_image_noalloc(cfg, img_dev.data_ptr(), len(data), acc_dev.data_ptr(),
                                                carry_in=hbtgpu.make_carry(valid=True, tstamp=t0, last=[0] * n))
        got = acc_dev.cpu().numpy().view(np.uint64).reshape(nb, n + 3)
        assert ctr == exp["counters"]
        assert np.array_equal(got, exp["buckets"])
        assert _carry_tuple(cout, n) == _carry_tuple(exp["carry_out"], n)


def test_noalloc_rejects_oversized_image():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    cfg = hbtgpu.Config()
    cfg.device, cfg.n_events, cfg.n_buckets, cfg.t0, cfg.bucket_ns = 0, 4, 16, 0, 1000
    cfg.time_mult, cfg.max_cpus = 1, 1
    acc_dev = torch.zeros(16 * 7, dtype=torch.int64, device="cuda:0")
    img = torch.zeros(16, dtype=torch.uint8, device="cuda:0")
    with pytest.raises(hbtgpu.HbtGpuError) as e:
        hbtgpu.decode_image_noalloc(cfg, img.data_ptr(), hbtgpu.NOALLOC_MAX_IMAGE_BYTES + 8,
                                    acc_dev.data_ptr())
    assert e.value.code == -errno.E2BIG
    assert "HBTGPU_NOALLOC_MAX_IMAGE_BYTES" in str(e.value)
