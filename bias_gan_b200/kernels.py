"""Thin Python wrappers over the C ABI (include/bgr.h): raw pointers in, status out.

Tensors here are bf16 NHWC views `[N, H, W, C]` whose last dim is contiguous; a view may be a
channel slice of a wider buffer (pixel stride = `stride(2)`), which is how concatenations are
written in place.
"""
import ctypes

import torch

from . import _lib
from ._lib import check, int_array, ptr, stream_ptr


def taps_for(kernel_size, dilation=1):
    """(dy, dx) input offsets of a `kernel_size` x `kernel_size` conv with 'same' padding
    (fixed_padding, deeplab.py:66-72, and padding=dilation for the dense 3x3 convs)."""
    if kernel_size == 1:
        return [0], [0]
    r = (kernel_size - 1) // 2
    dy, dx = [], []
    for i in range(kernel_size):
        for j in range(kernel_size):
            dy.append((i - r) * dilation)
            dx.append((j - r) * dilation)
    return dy, dx


def _nhwc_view(t):
    assert t.dim() == 4 and t.dtype == torch.bfloat16 and t.is_cuda, "expected a bf16 NHWC cuda view"
    N, H, W, C = t.shape
    ld = t.stride(2)
    assert t.stride(3) == 1 and t.stride(1) == W * ld and t.stride(0) == H * W * ld, \
        "expected dense pixels with a uniform pixel stride"
    return N, H, W, C, ld


def conv_fprop(x, w, dy, dx, out=None, bias=None):
    """y = conv(x, w).  x: [N,H,W,Cin] bf16 NHWC; w: [Cout, ntaps, Cin] bf16; out: [N,H,W,Cout]."""
    N, H, W, Cin, ldx = _nhwc_view(x)
    Cout, ntaps, wc = w.shape
    assert wc == Cin and w.is_contiguous() and w.dtype == torch.bfloat16 and ntaps == len(dy)
    if out is None:
        out = torch.empty((N, H, W, Cout), dtype=torch.bfloat16, device=x.device)
    No, Ho, Wo, Co, ldy = _nhwc_view(out)
    assert (No, Ho, Wo, Co) == (N, H, W, Cout)
    if bias is not None:
        assert bias.dtype == torch.float32 and bias.numel() == Cout
    check(_lib.lib().bgr_conv_fprop(ptr(x), N, H, W, Cin, ldx, ptr(w), Cout, ntaps, int_array(dy),
                                    int_array(dx), ptr(out), ldy, ptr(bias), stream_ptr()))
    return out


def conv_wgrad(x, gy, dy, dx, dw=None):
    """dw[Cout, ntaps, Cin] (fp32) += sum_pixels gy[p, co] * x[p + tap, ci]."""
    N, H, W, Cin, ldx = _nhwc_view(x)
    Ng, Hg, Wg, Cout, ldy = _nhwc_view(gy)
    assert (Ng, Hg, Wg) == (N, H, W)
    ntaps = len(dy)
    if dw is None:
        dw = torch.zeros((Cout, ntaps, Cin), dtype=torch.float32, device=x.device)
    assert dw.is_contiguous() and dw.dtype == torch.float32 and tuple(dw.shape) == (Cout, ntaps, Cin)
    check(_lib.lib().bgr_conv_wgrad(ptr(x), N, H, W, Cin, ldx, ptr(gy), Cout, ldy, ntaps,
                                    int_array(dy), int_array(dx), ptr(dw), stream_ptr()))
    return dw


_SRC_DTYPES = {torch.int64: 0, torch.int32: 1, torch.float64: 2, torch.float32: 3}


def stage_normalize(src, layout, channels, shift, scale, crop=None, flip=(False, False), s2d=False, out=None):
    """Fused standardise/crop/flip/bf16 staging.  src: [B,Cf,H,W] ('NCHW') or [B,H,W,Cf] ('NHWC')
    device tensor straight from the reader; shift/scale: fp32 device [C]; returns bf16
    [B,h,w,C] or the space-to-depth field layout [B,h/2,w/2,4C]."""
    assert src.is_cuda and src.is_contiguous() and src.dim() == 4
    if layout == "NCHW":
        B, Cf, H, W = src.shape
        lay = 0
    else:
        B, H, W, Cf = src.shape
        lay = 1
    channels = list(range(Cf)) if channels is None else [int(c) for c in channels]
    C = len(channels)
    y0, x0, h, w = crop if crop is not None else (0, 0, H, W)
    shape = (B, h // 2, w // 2, 4 * C) if s2d else (B, h, w, C)
    if out is None:
        out = torch.empty(shape, dtype=torch.bfloat16, device=src.device)
    assert tuple(out.shape) == shape and out.is_contiguous() and out.dtype == torch.bfloat16
    assert shift.dtype == torch.float32 and scale.dtype == torch.float32 and shift.numel() == C == scale.numel()
    check(_lib.lib().bgr_stage_normalize(ptr(src), _SRC_DTYPES[src.dtype], lay, B, Cf, H, W, int_array(channels), C,
                                         ptr(shift), ptr(scale), y0, x0, h, w, int(flip[0]), int(flip[1]),
                                         ptr(out), int(s2d), stream_ptr()))
    return out
