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
