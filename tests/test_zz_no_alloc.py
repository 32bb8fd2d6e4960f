"""bgr_bn_stats_static -- the allocation-free variant of bgr_bn_stats (scratch in a static device array of the
library; no counterpart in the reference) -- through the C ABI against
the existing entry point, on two calls in a row.  Collected last on purpose."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _reference_sums(K, x):
    """The existing entry point: partial rows in a caller-provided workspace, added here in fp64."""
    return K.bn_stats(x).double().sum(0)


def _call_static(lib, x, M, C, ld, out):
    return lib.bgr_bn_stats_static(ctypes.c_void_p(x.data_ptr()), ctypes.c_long(M), C, ld, ctypes.c_void_p(out.data_ptr()),
                                   ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))


def test_bn_stats_static_matches_bn_stats_on_consecutive_calls():
    from bias_gan_b200 import _lib, kernels as K
    lib = _lib.lib()
    dev = torch.device("cuda", 0)
    g = torch.Generator(device="cpu").manual_seed(17)
    # first call: the widest case (more partial rows, 728 channels at the padded pixel stride 768); second call:
    # fewer rows and channels, so stale contents of the shared array would show if anything read beyond its rows
    shapes = [(2, 48, 72, 728), (1, 9, 7, 128), (2, 6, 9, 2048)]
    outs = [torch.full((2 * s[3],), -7.0, dtype=torch.float64, device=dev) for s in shapes]   # allocated up front
    xs = []
    for (N, H, W, C) in shapes:
        x = K.act_empty((N, H, W, C), dev)
        x.copy_((torch.randn(N, H, W, C, generator=g) * 1.5 + 0.25).to(dev))
        xs.append(x)
    torch.cuda.synchronize()
    for i, (x, out, (N, H, W, C)) in enumerate(zip(xs, outs, shapes)):
        free_before = torch.cuda.mem_get_info(dev)[0]
        rc = _call_static(lib, x, N * H * W, C, x.stride(2), out)
        assert rc == 0, _lib.last_error()
        torch.cuda.synchronize()
        if i > 0:   # after the first (warm-up) call the device's free memory must not move: nothing is allocated
            assert abs(torch.cuda.mem_get_info(dev)[0] - free_before) <= 2 * 2 ** 20
        want = _reference_sums(K, x)
        # the tolerance of tests/test_kernels_gpu.py::test_conv_fused_bn_stats for the same sums
        assert torch.allclose(out, want, rtol=1e-5, atol=1e-3), (i, float((out - want).abs().max()))
        xd = x.double().reshape(-1, C)
        assert torch.allclose(out, torch.cat((xd.sum(0), (xd * xd).sum(0))), rtol=1e-5, atol=1e-3)
    # the same input twice in a row: bit-identical (fixed-order sums, no state carried between calls)
    again = torch.empty_like(outs[0])
    K.bn_stats_static(xs[0], again)
    torch.cuda.synchronize()
    assert torch.equal(again, outs[0])


def test_bn_stats_static_rejects_sizes_beyond_its_maximum():
    from bias_gan_b200 import _lib, kernels as K
    lib = _lib.lib()
    dev = torch.device("cuda", 0)
    C = K.BN_STATS_STATIC_MAX_C + 8
    x = torch.zeros((4, C), dtype=torch.bfloat16, device=dev)
    out = torch.zeros(2 * C, dtype=torch.float64, device=dev)
    rc = _call_static(lib, x, 4, C, C, out)
    assert rc == _lib.BGR_ERR_ARG and "BGR_BN_STATS_STATIC_MAX_C" in _lib.last_error()
    with pytest.raises(ValueError):
        _lib.check(rc)
    assert float(out.abs().max()) == 0.0     # nothing was launched
