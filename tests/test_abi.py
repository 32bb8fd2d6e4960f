"""The C-ABI library loads on a CPU-only machine and exports every symbol include/bgr.h declares
(no compute calls here), and the product refuses to run without its native library or a GPU."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "bgr.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(bgr_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_path():
    names = _declared()
    for must in ("bgr_reader_create", "bgr_reader_read", "bgr_stage_normalize", "bgr_conv_fprop", "bgr_conv_wgrad",
                 "bgr_bn_backward", "bgr_dw_fwd", "bgr_bilinear_fwd", "bgr_head_fwd", "bgr_l1_loss_grad"):
        assert must in names


def test_library_exports_every_declared_symbol():
    from bias_gan_b200 import _lib
    L = _lib.lib()
    missing = [n for n in _declared() if not hasattr(L, n)]
    assert not missing, missing
    assert L.bgr_version() >= 100
    L.bgr_launch_count.restype = ctypes.c_ulonglong
    assert L.bgr_launch_count() >= 0


def test_no_cpu_fallback():
    import torch
    from bias_gan_b200 import numpy_reader as nr
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(RuntimeError):
        nr.numpy_reader(False, 0)  # device path without a CUDA device must fail loudly


def test_product_does_not_import_the_oracle():
    pkg = os.path.join(ROOT, "bias_gan_b200")
    for f in os.listdir(pkg):
        if f.endswith(".py"):
            assert "oracle" not in open(os.path.join(pkg, f)).read(), f


def test_padded_layout_helpers():
    """Host logic of the 128-byte-line layout rule (kernels.pad_channels / act views / weight pitch)."""
    import torch
    from bias_gan_b200 import kernels as K, layouts as L
    assert [K.pad_channels(c) for c in (16, 48, 64, 128, 304, 728, 1536)] == [16, 48, 64, 128, 320, 768, 1536]
    t = torch.empty(2, 3, 5, 768, dtype=torch.bfloat16)
    assert K.is_act_view(t) and K.is_act_view(t[..., :728]) and not K.is_act_view(t[:, :, ::2, :728])
    assert not K.is_act_view(t.float()[..., :728]) and not K.is_act_view(t.permute(0, 3, 1, 2))
    w = torch.randn(24, 728, 3, 3)
    g = L.gemm_weight(w)
    assert tuple(g.shape) == (24, 9, 728) and g.stride() == (9 * 768, 768, 1)
    assert torch.equal(g, w.permute(0, 2, 3, 1).reshape(24, 9, 728).bfloat16())
    gt = L.gemm_weight_t(w)
    assert tuple(gt.shape) == (728, 9, 24) and gt.stride() == (9 * 64, 64, 1)
    assert torch.equal(gt, w.permute(1, 2, 3, 0).reshape(728, 9, 24).bfloat16())


def test_weight_cache():
    """layouts.cached: reuse until the parameter changes in place, is moved, or is invalidated."""
    import torch
    from bias_gan_b200 import layouts as L
    lin = torch.nn.Linear(4, 3)
    calls = []
    fn = lambda w: (calls.append(1), w.detach().clone())[1]
    a = L.cached(lin.weight, "k", fn)
    assert L.cached(lin.weight, "k", fn) is a and len(calls) == 1
    assert L.cached(lin.bias, "k", fn) is not a and len(calls) == 2
    with torch.no_grad():
        lin.weight.add_(1.0)
    b = L.cached(lin.weight, "k", fn)
    assert b is not a and torch.equal(b, lin.weight) and len(calls) == 3
    L.invalidate_cache(lin)
    assert L.cached(lin.weight, "k", fn) is not b and len(calls) == 4
    opt = torch.optim.AdamW(lin.parameters(), lr=0.1)
    lin(torch.ones(2, 4)).sum().backward()
    c = L.cached(lin.weight, "k", fn)
    opt.step()
    assert L.cached(lin.weight, "k", fn) is not c   # optimizer steps bump the version counter
    n = len(L._cache)
    del lin, opt, a, b, c
    import gc; gc.collect()
    assert len(L._cache) < n
