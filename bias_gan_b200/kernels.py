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


def pad_channels(C):
    """Channel stride of an activation with C channels: TMA and the vector kernels move 128-byte
    lines, so wide tensors whose rows would straddle lines (728 channels) are stored with the
    pixel stride rounded up to 64 elements (include/bgr.h, bgr_conv_fprop)."""
    return C if (C <= 64 or C % 64 == 0) else (C + 63) // 64 * 64


def act_empty(shape, device):
    """Uninitialised bf16 [..., C] activation view with the padded pixel stride."""
    C = shape[-1]
    Cp = pad_channels(C)
    t = torch.empty(tuple(shape[:-1]) + (Cp,), dtype=torch.bfloat16, device=device)
    return t if Cp == C else t[..., :C]


def is_act_view(t):
    """bf16 [..., C] view with unit channel stride and one uniform pixel stride."""
    if t.dtype != torch.bfloat16 or t.stride(-1) != 1:
        return False
    if t.is_contiguous():
        return True
    exp = None
    for d in range(t.dim() - 2, -1, -1):
        if t.shape[d] == 1:
            continue
        if exp is None:
            exp = t.stride(d)
            if exp < t.shape[-1]:
                return False
        elif t.stride(d) != exp:
            return False
        exp *= t.shape[d]
    return True


def _nhwc_view(t):
    assert t.dim() == 4 and t.dtype == torch.bfloat16 and t.is_cuda, "expected a bf16 NHWC cuda view"
    N, H, W, C = t.shape
    if t.is_contiguous():
        return N, H, W, C, C
    ld = t.stride(2) if W > 1 else (t.stride(1) if H > 1 else (t.stride(0) if N > 1 else C))
    assert t.stride(3) == 1 and (W == 1 or t.stride(2) == ld) and (H == 1 or t.stride(1) == W * ld) \
        and (N == 1 or t.stride(0) == H * W * ld), "expected dense pixels with a uniform pixel stride"
    return N, H, W, C, ld


BN_MAX_ROWS = 448            # include/bgr.h: capacities of the partial-row workspaces
DW_WGRAD_WS_FLOATS = 1 << 20
GAP_MAX_ROWS, HEAD_MAX_ROWS, LOSS_MAX_ROWS = 64, 296, 2368
_sm_counts = {}


def _conv_stat_rows(device):
    """Upper bound of the partial rows a conv epilogue writes: one per group of thread blocks <= SM count."""
    idx = device.index if device.index is not None else torch.cuda.current_device()
    if idx not in _sm_counts:
        _sm_counts[idx] = min(BN_MAX_ROWS, torch.cuda.get_device_properties(idx).multi_processor_count)
    return _sm_counts[idx]


def conv_fprop(x, w, dy, dx, out=None, bias=None, want_stats=False):
    """y = conv(x, w).  x: [N,H,W,Cin] bf16 NHWC; w: [Cout, ntaps, Cin] bf16; out: [N,H,W,Cout].
    want_stats=True returns (y, part): part = fp32 [rows, 2*Cout] partial rows of the per-channel sum /
    sum of squares of y, written by the conv epilogue, for bn_finalize to add in row order."""
    N, H, W, Cin, ldx = _nhwc_view(x)
    Cout, ntaps, wc = w.shape
    ldw = w.stride(1) if ntaps > 1 else w.stride(0)
    assert wc == Cin and w.stride(2) == 1 and w.stride(0) == ntaps * ldw and w.dtype == torch.bfloat16 \
        and ntaps == len(dy), "expected bf16 [Cout, ntaps, Cin] weights with a uniform row pitch"
    if out is None:
        out = act_empty((N, H, W, Cout), x.device)
    No, Ho, Wo, Co, ldy = _nhwc_view(out)
    assert (No, Ho, Wo, Co) == (N, H, W, Cout)
    if bias is not None:
        assert bias.dtype == torch.float32 and bias.numel() == Cout
    part, rows = None, ctypes.c_int(0)
    if want_stats:
        part = torch.empty((_conv_stat_rows(x.device), 2 * Cout), dtype=torch.float32, device=x.device)
    check(_lib.lib().bgr_conv_fprop(ptr(x), N, H, W, Cin, ldx, ptr(w), ldw, Cout, ntaps, int_array(dy),
                                    int_array(dx), ptr(out), ldy, ptr(bias), ptr(part), ctypes.byref(rows),
                                    stream_ptr()))
    if want_stats:
        assert 1 <= rows.value <= part.shape[0]
        return out, part[:rows.value]
    return out


def weight_to_gemm(w, transpose):
    """fp32 cuda [Cout, Cin, kh, kw] -> bf16 [rows, taps, inner] view of a zero-padded [rows, taps, pitch]."""
    assert w.is_cuda and w.dtype == torch.float32 and w.is_contiguous() and w.dim() == 4
    Cout, Cin, kh, kw = w.shape
    rows, inner = (Cin, Cout) if transpose else (Cout, Cin)
    pitch = inner if inner % 64 == 0 else (inner + 63) // 64 * 64
    out = torch.empty((rows, kh * kw, pitch), dtype=torch.bfloat16, device=w.device)
    check(_lib.lib().bgr_weight_to_gemm(ptr(w), Cout, Cin, kh * kw, int(transpose), ptr(out), pitch, stream_ptr()))
    return out if pitch == inner else out[..., :inner]


def conv_wgrad(x, gy, dy, dx, dw=None):
    """dw[Cout, ntaps, Cin] (fp32) = sum_pixels gy[p, co] * x[p + tap, ci]  (overwritten; the pixel splits
    write slabs of a workspace that are added in split order)."""
    N, H, W, Cin, ldx = _nhwc_view(x)
    Ng, Hg, Wg, Cout, ldy = _nhwc_view(gy)
    assert (Ng, Hg, Wg) == (N, H, W)
    ntaps = len(dy)
    if dw is None:
        dw = torch.empty((Cout, ntaps, Cin), dtype=torch.float32, device=x.device)
    assert dw.is_contiguous() and dw.dtype == torch.float32 and tuple(dw.shape) == (Cout, ntaps, Cin)
    lib = _lib.lib()
    lib.bgr_conv_wgrad_workspace.restype = ctypes.c_size_t
    nbytes = int(lib.bgr_conv_wgrad_workspace(N, H, W, Cin, Cout, ntaps))
    ws = torch.empty(nbytes // 4, dtype=torch.float32, device=x.device) if nbytes else None
    check(lib.bgr_conv_wgrad(ptr(x), N, H, W, Cin, ldx, ptr(gy), Cout, ldy, ntaps, int_array(dy), int_array(dx),
                             ptr(dw), ptr(ws), ctypes.c_size_t(nbytes), stream_ptr()))
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


def channel_stats(src, layout="NCHW"):
    """Per-sample, per-channel (mean, mean-square, min, max) in fp64 on the device: returns a [B,4,Cf]
    fp64 tensor.  src: [B,Cf,H,W] ('NCHW') or [B,H,W,Cf] ('NHWC') device tensor as the reader returns
    it (see bgr_channel_stats in bgr.h)."""
    assert src.is_cuda and src.is_contiguous() and src.dim() == 4
    if layout == "NCHW":
        B, Cf, H, W = src.shape
        lay, rows = 0, B * Cf
    else:
        B, H, W, Cf = src.shape
        lay, rows = 1, B
    # enough chunks for four resident blocks per SM, but at least 16K elements each
    splits = max(1, min((148 * 4 + rows - 1) // rows, (H * W * (1 if lay == 0 else Cf)) // 16384, 4096))
    part = torch.empty(B * Cf * splits * 4, dtype=torch.float64, device=src.device)
    out = torch.empty((B, 4, Cf), dtype=torch.float64, device=src.device)
    check(_lib.lib().bgr_channel_stats(ptr(src), _SRC_DTYPES[src.dtype], lay, B, Cf, H, W, splits, ptr(part), ptr(out),
                                       stream_ptr()))
    return out


# ---------------------------------------------------------------------------------------------
# BatchNorm / activation / depthwise / decoder helpers.  Tensors are bf16 [..., C] whose leading
# dims are dense with a uniform pixel stride (channel slices of wider buffers are fine).
# ---------------------------------------------------------------------------------------------
c_long, c_float = ctypes.c_long, ctypes.c_float


def _rows(t):
    """(M, C, ld) of a bf16 [..., C] view with uniform pixel stride."""
    assert t.dtype == torch.bfloat16 and t.is_cuda and t.stride(-1) == 1
    C = t.shape[-1]
    M = t.numel() // C
    if t.is_contiguous():
        return M, C, C
    ld = C
    for d in range(t.dim() - 2, -1, -1):
        if t.shape[d] > 1:
            ld = t.stride(d)
            break
    exp = ld
    for d in range(t.dim() - 2, -1, -1):
        assert t.shape[d] == 1 or t.stride(d) == exp, "non-uniform pixel stride"
        exp *= t.shape[d]
    return M, C, ld


def bn_stats(x):
    """fp32 [rows, 2C] partial rows of (sum x | sum x^2) per channel, one row per thread block."""
    M, C, ld = _rows(x)
    part = torch.empty((BN_MAX_ROWS, 2 * C), dtype=torch.float32, device=x.device)
    rows = ctypes.c_int(0)
    check(_lib.lib().bgr_bn_stats(ptr(x), c_long(M), C, ld, ptr(part), ctypes.byref(rows), stream_ptr()))
    return part[:rows.value]


def channel_sums(x):
    """fp32 [C]: sum over pixels of a bf16 [..., C] tensor (partial rows added in row order in fp64)."""
    C = x.shape[-1]
    return bn_stats(x)[:, :C].double().sum(0).float()


def bn_finalize(part, M, gamma, beta, eps, momentum, running_mean, running_var, num_batches_tracked=None):
    """part: the [rows, 2C] partial rows of bn_stats or of a conv epilogue (conv_fprop(want_stats=True))."""
    assert part.dtype == torch.float32 and part.dim() == 2 and part.is_contiguous()
    rows, C = part.shape[0], part.shape[1] // 2
    dev = part.device
    mean, invstd, a, b = (torch.empty(C, dtype=torch.float32, device=dev) for _ in range(4))
    check(_lib.lib().bgr_bn_finalize(ptr(part), rows, c_long(M), C, ptr(gamma), ptr(beta), c_float(eps),
                                     c_float(momentum), ptr(mean), ptr(invstd), ptr(a), ptr(b), ptr(running_mean),
                                     ptr(running_var), ptr(num_batches_tracked), stream_ptr()))
    return mean, invstd, a, b


def bn_apply(x, a, b, res=None, act=True, slope=0.2, out=None):
    M, C, ldx = _rows(x)
    if out is None:
        out = act_empty(x.shape, x.device)
    Mo, Co, ldo = _rows(out)
    assert (Mo, Co) == (M, C)
    ldr = 0
    if res is not None:
        Mr, Cr, ldr = _rows(res)
        assert (Mr, Cr) == (M, C)
    check(_lib.lib().bgr_bn_apply(ptr(x), ldx, ptr(a), ptr(b), ptr(res), ldr, ptr(out), ldo, c_long(M), C,
                                  int(act), c_float(slope), stream_ptr()))
    return out


def bn_backward(g, out, x, mean, invstd, gamma, act=True, slope=0.2, g2=None):
    """Returns (dx, dyh, grads) — dyh is also the gradient of the residual input; grads = fp32
    [dL/dbeta (C) | dL/dgamma (C)]."""
    M, C, ldx = _rows(x)
    Mg, Cg, ldg = _rows(g)
    assert (Mg, Cg) == (M, C)
    ldg2 = 0
    if g2 is not None:
        M2, C2, ldg2 = _rows(g2)
        assert (M2, C2) == (M, C)
    ldo = 0
    if act:
        Mo, Co, ldo = _rows(out)
        assert (Mo, Co) == (M, C)
    dyh = act_empty(x.shape, x.device)
    dx = act_empty(x.shape, x.device)
    sums = torch.empty((BN_MAX_ROWS, 2 * C), dtype=torch.float32, device=x.device)   # partial-row workspace
    coef = torch.empty(7 * C, dtype=torch.float32, device=x.device)
    check(_lib.lib().bgr_bn_backward(ptr(g), ldg, ptr(g2), ldg2, ptr(out) if act else ctypes.c_void_p(0), ldo, ptr(x), ldx,
                                     ptr(mean), ptr(invstd), ptr(gamma), int(act), c_float(slope), c_long(M), C,
                                     ptr(dyh), ptr(dx), _rows(dx)[2], ptr(sums), ptr(coef), stream_ptr()))
    return dx, dyh, coef[5 * C:7 * C]   # fp32 [dL/dbeta | dL/dgamma]


def bn_backward_frozen(g, out, x, mean, invstd, gamma, act=True, slope=0.2, g2=None, want_dx=True):
    """Backward with frozen statistics (eval-mode BatchNorm / Identity normaliser): (dx or None, dyh,
    fp32 [dL/dbeta | dL/dgamma])."""
    M, C, ldx = _rows(x)
    Mg, Cg, ldg = _rows(g)
    assert (Mg, Cg) == (M, C)
    ldg2 = 0
    if g2 is not None:
        ldg2 = _rows(g2)[2]
    ldo = _rows(out)[2] if act else 0
    dyh = act_empty(x.shape, x.device)
    dx = act_empty(x.shape, x.device) if want_dx else None
    part = torch.empty((BN_MAX_ROWS, 2 * C), dtype=torch.float32, device=x.device)
    coef = torch.empty(7 * C, dtype=torch.float32, device=x.device)
    check(_lib.lib().bgr_bn_backward_frozen(ptr(g), ldg, ptr(g2), ldg2, ptr(out) if act else ctypes.c_void_p(0), ldo,
                                            ptr(x), ldx, ptr(mean), ptr(invstd), ptr(gamma), int(act), c_float(slope),
                                            c_long(M), C, ptr(dyh), ptr(dx), _rows(dyh)[2], ptr(part), ptr(coef),
                                            stream_ptr()))
    return dx, dyh, coef[5 * C:7 * C]


def dw_fwd(x, w9, stride, dil, out=None):
    N, H, W, C, ldx = _nhwc_view(x)
    Ho, Wo = (H - 1) // stride + 1, (W - 1) // stride + 1
    if out is None:
        out = act_empty((N, Ho, Wo, C), x.device)
    _, _, _, _, ldy = _nhwc_view(out)
    assert w9.dtype == torch.float32 and tuple(w9.shape) == (9, C) and w9.is_contiguous()
    check(_lib.lib().bgr_dw_fwd(ptr(x), N, H, W, C, ldx, ptr(w9), stride, dil, ptr(out), ldy, stream_ptr()))
    return out


def dw_fwd_bn(x, w9, a, b, act, slope=0.2, out=None):
    """Stride-1 dilation-1 depthwise conv of act(a*x + b) with the transform applied on load."""
    N, H, W, C, ldx = _nhwc_view(x)
    if out is None:
        out = act_empty((N, H, W, C), x.device)
    _, _, _, _, ldy = _nhwc_view(out)
    assert w9.dtype == torch.float32 and tuple(w9.shape) == (9, C) and w9.is_contiguous()
    assert a.dtype == b.dtype == torch.float32 and a.numel() == b.numel() == C
    check(_lib.lib().bgr_dw_fwd_bn(ptr(x), N, H, W, C, ldx, ptr(w9), ptr(a), ptr(b), int(act), c_float(slope),
                                   ptr(out), ldy, stream_ptr()))
    return out


def dw_wgrad_bn(x, a, b, act, gy, slope=0.2):
    N, H, W, C, ldx = _nhwc_view(x)
    _, _, _, _, ldy = _nhwc_view(gy)
    dw = torch.empty((C, 9), dtype=torch.float32, device=x.device)   # the parameter's [C,1,3,3] layout
    ws = torch.empty(DW_WGRAD_WS_FLOATS, dtype=torch.float32, device=x.device)
    check(_lib.lib().bgr_dw_wgrad_bn(ptr(x), N, H, W, C, ldx, ptr(a), ptr(b), int(act), c_float(slope), ptr(gy), ldy,
                                     ptr(dw), ptr(ws), stream_ptr()))
    return dw


def bn_backward_recompute(g, a, b, x, mean, invstd, gamma, act=True, slope=0.2, want_dyh=False, g2=None):
    """bn_backward for a BatchNorm(+act) whose output was never stored: act' from sign(a*x + b).
    g2: a second gradient of the same output (dual handle), added on load by both passes."""
    M, C, ldx = _rows(x)
    Mg, Cg, ldg = _rows(g)
    assert (Mg, Cg) == (M, C)
    if g2 is not None:
        M2, C2, ldg2 = _rows(g2)
        assert (M2, C2) == (M, C)
        dyh = act_empty(x.shape, x.device) if want_dyh else None
        dx = act_empty(x.shape, x.device)
        sums = torch.empty((BN_MAX_ROWS, 2 * C), dtype=torch.float32, device=x.device)
        coef = torch.empty(7 * C, dtype=torch.float32, device=x.device)
        check(_lib.lib().bgr_bn_backward_recompute_dual(ptr(g), ldg, ptr(g2), ldg2, ptr(a), ptr(b), ptr(x), ldx, ptr(mean),
                                                        ptr(invstd), ptr(gamma), int(act), c_float(slope), c_long(M), C,
                                                        ptr(dyh), ptr(dx), _rows(dx)[2], ptr(sums), ptr(coef), stream_ptr()))
        return dx, dyh, coef[5 * C:7 * C]
    dyh = act_empty(x.shape, x.device) if want_dyh else None   # NULL: pass 2 recomputes it from g and x
    dx = act_empty(x.shape, x.device)
    sums = torch.empty((BN_MAX_ROWS, 2 * C), dtype=torch.float32, device=x.device)   # partial-row workspace
    coef = torch.empty(7 * C, dtype=torch.float32, device=x.device)
    check(_lib.lib().bgr_bn_backward_recompute(ptr(g), ldg, ptr(a), ptr(b), ptr(x), ldx, ptr(mean), ptr(invstd),
                                               ptr(gamma), int(act), c_float(slope), c_long(M), C, ptr(dyh), ptr(dx),
                                               _rows(dx)[2], ptr(sums), ptr(coef), stream_ptr()))
    return dx, dyh, coef[5 * C:7 * C]   # fp32 [dL/dbeta | dL/dgamma]


def dw_dgrad(gy, w9, stride, dil, in_shape):
    N, H, W, C = in_shape
    _, _, _, _, ldy = _nhwc_view(gy)
    dx = act_empty((N, H, W, C), gy.device)
    check(_lib.lib().bgr_dw_dgrad(ptr(gy), N, H, W, C, ldy, ptr(w9), stride, dil, ptr(dx), _nhwc_view(dx)[4],
                                  stream_ptr()))
    return dx


def dw_wgrad(x, gy, stride, dil):
    N, H, W, C, ldx = _nhwc_view(x)
    _, _, _, _, ldy = _nhwc_view(gy)
    dw = torch.empty((C, 9), dtype=torch.float32, device=x.device)   # the parameter's [C,1,3,3] layout
    ws = torch.empty(DW_WGRAD_WS_FLOATS, dtype=torch.float32, device=x.device)
    check(_lib.lib().bgr_dw_wgrad(ptr(x), N, H, W, C, ldx, ptr(gy), ldy, stride, dil, ptr(dw), ptr(ws), stream_ptr()))
    return dw


def subsample2(x):
    N, H, W, C, ldx = _nhwc_view(x)
    out = act_empty((N, (H + 1) // 2, (W + 1) // 2, C), x.device)
    check(_lib.lib().bgr_subsample2(ptr(x), N, H, W, C, ldx, ptr(out), _nhwc_view(out)[4], stream_ptr()))
    return out


def subsample2_bwd(gy, in_shape):
    N, H, W, C = in_shape
    _, _, _, _, ldy = _nhwc_view(gy)
    dx = act_empty((N, H, W, C), gy.device)
    check(_lib.lib().bgr_subsample2_bwd(ptr(gy), N, H, W, C, ldy, ptr(dx), _nhwc_view(dx)[4], stream_ptr()))
    return dx


def bilinear_fwd(x, H, W, out=None, s2d=False):
    N, h, w, C, ldx = _nhwc_view(x)
    if out is None:
        shape = (N, H // 2, W // 2, 4 * C) if s2d else (N, H, W, C)
        out = torch.empty(shape, dtype=torch.bfloat16, device=x.device)
    ldy = 4 * C if s2d else _nhwc_view(out)[4]
    check(_lib.lib().bgr_bilinear_fwd(ptr(x), N, h, w, C, ldx, ptr(out), H, W, ldy, int(s2d), stream_ptr()))
    return out


def bilinear_bwd(gy, in_shape, H, W, s2d=False):
    N, h, w, C = in_shape
    ldy = 4 * C if s2d else _nhwc_view(gy)[4]
    assert gy.dtype == torch.bfloat16
    gx = torch.empty((N, h, w, C), dtype=torch.float32, device=gy.device)
    check(_lib.lib().bgr_bilinear_bwd(ptr(gy), N, h, w, C, H, W, ldy, int(s2d), ptr(gx), stream_ptr()))
    return gx


def gap_fwd(x):
    N, H, W, C, ld = _nhwc_view(x)
    sums = torch.empty((N, C), dtype=torch.float32, device=x.device)
    ws = torch.empty((GAP_MAX_ROWS, N, C), dtype=torch.float32, device=x.device)
    check(_lib.lib().bgr_gap_fwd(ptr(x), N, c_long(H * W), C, ld, ptr(sums), ptr(ws), stream_ptr()))
    return sums


def gap_bwd(g, N, H, W, C):
    assert g.dtype == torch.bfloat16 and g.is_contiguous()
    dx = torch.empty((N, H, W, C), dtype=torch.bfloat16, device=g.device)
    check(_lib.lib().bgr_gap_bwd(ptr(g), N, c_long(H * W), C, ptr(dx), stream_ptr()))
    return dx


def head_fwd(x, w, logits):
    """logits (fp32 [N], pre-set to the bias) += x.view(N,-1) @ w."""
    N = x.shape[0]
    K = x.numel() // N
    assert x.is_contiguous() and w.is_contiguous() and w.numel() == K and logits.dtype == torch.float32
    ws = torch.empty((HEAD_MAX_ROWS, N), dtype=torch.float32, device=x.device)
    check(_lib.lib().bgr_head_fwd(ptr(x), ptr(w), N, c_long(K), ptr(logits), ptr(ws), stream_ptr()))
    return logits


def head_bwd(g, x, w, need_dx=True, need_dw=True):
    N = x.shape[0]
    K = x.numel() // N
    assert g.dtype == torch.float32 and g.numel() == N and g.is_contiguous()
    dx = torch.empty(x.shape, dtype=torch.bfloat16, device=x.device) if need_dx else None
    dw = torch.empty(K, dtype=torch.float32, device=x.device) if need_dw else None
    check(_lib.lib().bgr_head_bwd(ptr(g), ptr(x), ptr(w), N, c_long(K), ptr(dx), ptr(dw), stream_ptr()))
    return dx, dw


def l1_loss_grad(a, b, gscale, need_grad=True):
    """Returns (sum|a-b| as fp64 0-d tensor, grad bf16 or None)."""
    assert a.shape == b.shape and a.is_contiguous() and b.is_contiguous() and a.dtype == b.dtype == torch.bfloat16
    s = torch.empty((), dtype=torch.float64, device=a.device)
    grad = torch.empty(a.shape, dtype=torch.bfloat16, device=a.device) if need_grad else None
    ws = torch.empty(LOSS_MAX_ROWS, dtype=torch.float32, device=a.device)
    check(_lib.lib().bgr_l1_loss_grad(ptr(a), ptr(b), c_long(a.numel()), c_float(gscale), ptr(grad), ptr(s), ptr(ws),
                                      stream_ptr()))
    return s, grad


def reg_loss_grad(a, b, mode, gscale, weights=None, beta=1.0, gscale_dev=None, need_grad=True):
    """(sum w*l(a-b) as an fp64 0-d tensor, grad bf16 or None); mode 0 L1, 1 SmoothL1(beta), 2 MSE."""
    assert a.shape == b.shape and a.is_contiguous() and b.is_contiguous() and a.dtype == b.dtype == torch.bfloat16
    if weights is not None:
        assert weights.shape == a.shape and weights.is_contiguous() and weights.dtype == torch.bfloat16
    if gscale_dev is not None:
        assert gscale_dev.dtype == torch.float32 and gscale_dev.numel() == 1 and gscale_dev.is_cuda
    s = torch.empty((), dtype=torch.float64, device=a.device)
    grad = torch.empty(a.shape, dtype=torch.bfloat16, device=a.device) if need_grad else None
    ws = torch.empty(LOSS_MAX_ROWS, dtype=torch.float32, device=a.device)
    check(_lib.lib().bgr_reg_loss_grad(ptr(a), ptr(b), ptr(weights), c_long(a.numel()), int(mode), c_float(beta),
                                       c_float(gscale), ptr(gscale_dev), ptr(grad), ptr(s), ptr(ws), stream_ptr()))
    return s, grad


def field_to_nchw(field, C):
    N, H2, W2, C4 = field.shape
    assert C4 == 4 * C and field.is_contiguous() and field.dtype == torch.bfloat16
    out = torch.empty((N, C, 2 * H2, 2 * W2), dtype=torch.float32, device=field.device)
    check(_lib.lib().bgr_field_to_nchw(ptr(field), ptr(out), N, C, 2 * H2, 2 * W2, stream_ptr()))
    return out


def nchw_to_field(x):
    N, C, H, W = x.shape
    assert x.dtype == torch.float32 and x.is_contiguous() and H % 2 == 0 and W % 2 == 0
    out = torch.empty((N, H // 2, W // 2, 4 * C), dtype=torch.bfloat16, device=x.device)
    check(_lib.lib().bgr_nchw_to_field(ptr(x), ptr(out), N, C, H, W, stream_ptr()))
    return out


def field_append_noise(field, noise, Cp):
    """deeplab_gan.py:88-90 in the field layout: [N,H/2,W/2,4C] bf16 + fp32 NCHW noise [N,nd,H,W] ->
    [N,H/2,W/2,4*Cp] (channels per sub-pixel: input, noise, zero padding)."""
    N, H2, W2, C4 = field.shape
    C = C4 // 4
    Nn, nd, H, W = noise.shape
    assert field.dtype == torch.bfloat16 and field.is_contiguous() and noise.dtype == torch.float32 \
        and noise.is_contiguous() and noise.is_cuda and (Nn, H, W) == (N, 2 * H2, 2 * W2)
    out = torch.empty((N, H2, W2, 4 * Cp), dtype=torch.bfloat16, device=field.device)
    check(_lib.lib().bgr_field_append_noise(ptr(field), ptr(noise), ptr(out), N, C, nd, Cp, H, W, stream_ptr()))
    return out


def weights_prepare(descs, first_block, ndesc, total_blocks):
    """One launch for every derived weight copy of a model (layouts.WeightPack owns the tables)."""
    assert descs.is_cuda and descs.dtype == torch.uint8 and descs.numel() == 40 * ndesc
    assert first_block.is_cuda and first_block.dtype == torch.int32 and first_block.numel() == ndesc + 1
    check(_lib.lib().bgr_weights_prepare(ptr(descs), ptr(first_block), ndesc, total_blocks, stream_ptr()))



# ---------------------------------------------------------------------------------------------
# Deconv / Deconv1x decoders and odd-sized, odd-channel-count fields (csrc/deconv.cu)
# ---------------------------------------------------------------------------------------------
def zero_stuff2(x, H, W):
    """[N,h,w,C] -> [N,H,W,C] with y[:, 2i, 2j] = x[:, i, j] and zeros elsewhere (ConvTranspose2d s2 as a conv)."""
    N, h, w, C, ldx = _nhwc_view(x)
    y = act_empty((N, H, W, C), x.device)
    check(_lib.lib().bgr_zero_stuff2(ptr(x), N, h, w, C, ldx, ptr(y), H, W, _nhwc_view(y)[4], stream_ptr()))
    return y


def zero_stuff2_bwd(gy, h, w):
    N, H, W, C, ldy = _nhwc_view(gy)
    gx = act_empty((N, h, w, C), gy.device)
    check(_lib.lib().bgr_zero_stuff2_bwd(ptr(gy), N, h, w, C, H, W, ldy, ptr(gx), _nhwc_view(gx)[4], stream_ptr()))
    return gx


def avgpool2_fwd(x, pad, act, slope=0.2):
    """nn.AvgPool2d(2, stride=1, padding=pad) [+ LeakyReLU]."""
    N, H, W, C, ldx = _nhwc_view(x)
    y = act_empty((N, H + 2 * pad - 1, W + 2 * pad - 1, C), x.device)
    check(_lib.lib().bgr_avgpool2_fwd(ptr(x), N, H, W, C, ldx, int(pad), int(act), c_float(slope), ptr(y),
                                      _nhwc_view(y)[4], stream_ptr()))
    return y


def avgpool2_bwd(gy, yout, in_shape, pad, act, slope=0.2):
    N, H, W, C = in_shape
    ldg = _nhwc_view(gy)[4]
    ldo = _nhwc_view(yout)[4] if act else 0
    gx = act_empty((N, H, W, C), gy.device)
    check(_lib.lib().bgr_avgpool2_bwd(ptr(gy), ldg, ptr(yout) if act else ctypes.c_void_p(0), ldo, N, H, W, C, int(pad),
                                      int(act), c_float(slope), ptr(gx), _nhwc_view(gx)[4], stream_ptr()))
    return gx


def nhwc_to_field(x):
    """bf16 NHWC [N,H,W,C] -> field [N,ceil(H/2),ceil(W/2),4C] (zero row / column for odd H / W)."""
    N, H, W, C, ldx = _nhwc_view(x)
    f = torch.empty((N, (H + 1) // 2, (W + 1) // 2, 4 * C), dtype=torch.bfloat16, device=x.device)
    check(_lib.lib().bgr_nhwc_to_field(ptr(x), N, H, W, C, ldx, ptr(f), stream_ptr()))
    return f


def field_to_nhwc(field, H, W):
    N, H2, W2, C4 = field.shape
    assert field.is_contiguous() and field.dtype == torch.bfloat16 and H2 == (H + 1) // 2 and W2 == (W + 1) // 2
    y = act_empty((N, H, W, C4 // 4), field.device)
    check(_lib.lib().bgr_field_to_nhwc(ptr(field), N, H, W, C4 // 4, ptr(y), _nhwc_view(y)[4], stream_ptr()))
    return y


def stage_normalize_padded(src, channels, shift, scale, Cp):
    """fp32 NCHW [B,Cf,H,W] of any size -> field [B,ceil(H/2),ceil(W/2),4*Cp], zero beyond the selected
    channels and beyond H, W; shift / scale fp32 device [C] or both None (plain cast)."""
    assert src.is_cuda and src.is_contiguous() and src.dim() == 4 and src.dtype == torch.float32
    B, Cf, H, W = src.shape
    channels = list(range(Cf)) if channels is None else [int(c) for c in channels]
    C = len(channels)
    if shift is not None:
        assert shift.dtype == scale.dtype == torch.float32 and shift.numel() == scale.numel() == C
    f = torch.empty((B, (H + 1) // 2, (W + 1) // 2, 4 * Cp), dtype=torch.bfloat16, device=src.device)
    check(_lib.lib().bgr_stage_normalize_padded(ptr(src), B, Cf, H, W, int_array(channels), C, ptr(shift), ptr(scale),
                                                ptr(f), int(Cp), stream_ptr()))
    return f


def field_to_nchw_padded(field, C, H, W):
    N, H2, W2, C4 = field.shape
    assert field.is_contiguous() and field.dtype == torch.bfloat16 and H2 == (H + 1) // 2 and W2 == (W + 1) // 2
    out = torch.empty((N, C, H, W), dtype=torch.float32, device=field.device)
    check(_lib.lib().bgr_field_to_nchw_padded(ptr(field), ptr(out), N, C, C4 // 4, H, W, stream_ptr()))
    return out


BN_STATS_STATIC_MAX_C = 2048   # include/bgr.h


def bn_stats_static(x, out):
    """bgr_bn_stats_static: per-channel (sum x | sum x^2) of a bf16 [..., C] view into the caller's fp64 [2C]
    device tensor `out`; the partial rows live in the library's own static device array, so neither this wrapper
    nor the entry point allocates anything (not added by the reference: see include/bgr.h)."""
    M, C, ld = _rows(x)
    assert out.is_cuda and out.dtype == torch.float64 and out.is_contiguous() and out.numel() == 2 * C
    check(_lib.lib().bgr_bn_stats_static(ptr(x), c_long(M), C, ld, ptr(out), stream_ptr()))
    return out
