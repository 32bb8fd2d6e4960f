"""Autograd bindings of the sm_100a kernels (bf16 NHWC activations, fp32 master parameters).

Each Function's forward/backward is one or a few C-ABI launches (kernels.py); PyTorch's autograd
only orders them.  Parameter tensors keep the reference's shapes ([Cout,Cin,kh,kw] etc.) so that
state dicts interchange; GEMM-layout bf16 copies are derived per call (tiny tensors).
"""
import contextlib
import os

import torch

from . import kernels as K
from . import layouts as L

SLOPE = 0.2  # nn.LeakyReLU(0.2) everywhere in the reference (deeplab.py:100,180,334,365)

# The weight gradient of a layer does not feed the rest of the backward pass, so it is launched on a
# second stream, forked from and joined back into the calling stream inside the same autograd node.
# Under graph capture the fork/join become plain edges: the weight-gradient kernel fills the SMs that
# the data-gradient kernel leaves idle in its last (partial) round of tiles and during its ramps.
SIDE_STREAM = os.environ.get("BGR_NO_SIDE_STREAM") is None
# BatchNorm backward of residual-free layers from (g, y) alone (A/B switch, see _BNAct)
RECOMPUTE_BN_BWD = os.environ.get("BGR_RECOMPUTE_BN_BWD", "1") != "0"
_side_streams = {}
_forked = set()


@contextlib.contextmanager
def _side_stream(ref):
    """Work inside runs on the device's side stream, after everything queued so far; the caller
    joins with _join(ref) before returning the results."""
    if not (SIDE_STREAM and ref.is_cuda):
        yield
        return
    dev = ref.device.index
    side = _side_streams.get(dev)
    if side is None:
        side = _side_streams[dev] = torch.cuda.Stream(device=ref.device)
    side.wait_stream(torch.cuda.current_stream(ref.device))
    _forked.add(dev)
    with torch.cuda.stream(side):
        yield


def _join(ref):
    # only a node that forked joins: waiting on a side stream that holds no work of the current CUDA-graph
    # capture would wait on a non-captured event and invalidate the capture (data-gradient-only backward
    # passes, e.g. deeplab_gan.gradient_penalty)
    if ref.is_cuda and ref.device.index in _forked:
        _forked.discard(ref.device.index)
        torch.cuda.current_stream(ref.device).wait_stream(_side_streams[ref.device.index])


def _bf16c(t):
    """Incoming gradient as a bf16 activation view (padded pixel strides are kept as they are)."""
    return t if K.is_act_view(t) else t.to(torch.bfloat16).contiguous()


class _Conv(torch.autograd.Function):
    """Dense conv (1x1 / 3x3 / atrous, stride 1) as tcgen05 implicit GEMM."""

    @staticmethod
    def forward(ctx, x, weight, bias, dil, want_stats):
        ks = weight.shape[-1]
        dy, dx = K.taps_for(ks, dil)
        ctx.save_for_backward(x, weight)
        ctx.set_materialize_grads(False)   # no zero-filled "gradient" for the BatchNorm partial-sums output
        ctx.taps = (dy, dx)
        ctx.has_bias = bias is not None
        # want_stats: BatchNorm statistics of y from the conv epilogue (no second pass over y)
        y = K.conv_fprop(x, L.cached(weight, "gemm", L.gemm_weight), dy, dx, bias=bias, want_stats=want_stats)
        if want_stats:
            ctx.mark_non_differentiable(y[1])
        return y

    @staticmethod
    def backward(ctx, g, *unused):
        if g is None:
            return None, None, None, None, None
        x, weight = ctx.saved_tensors
        dy, dx = ctx.taps
        g = _bf16c(g)
        gx = gw = gb = None
        if ctx.needs_input_grad[1]:
            with _side_stream(g):
                gw = L.from_gemm_grad(K.conv_wgrad(x, g, dy, dx), weight.shape)
        if ctx.needs_input_grad[0]:
            gx = K.conv_fprop(g, L.cached(weight, "gemm_t", L.gemm_weight_t), [-v for v in dy], [-v for v in dx])
        if ctx.has_bias and ctx.needs_input_grad[2]:
            gb = K.channel_sums(g)
        _join(g)
        return gx, gw, gb, None, None


def conv2d(x, weight, bias=None, dilation=1, want_stats=False):
    """want_stats=True returns (y, part) with the fp32 [rows, 2*Cout] partial rows of the BatchNorm sums of y."""
    return _Conv.apply(x, weight, bias, dilation, want_stats)


def _pad_cin(w, Cp):
    return w if w.shape[1] == Cp else torch.nn.functional.pad(w, (0, 0, 0, 0, 0, Cp - w.shape[1]))


class _ConvS2D(torch.autograd.Function):
    """3x3 stride-2 pad-1 entry conv (deeplab.py:178) on a field stored as 2x2 space-to-depth
    blocks: a 2x2-tap stride-1 implicit GEMM with K = 4 taps x 4*C channels."""

    @staticmethod
    def forward(ctx, field, weight, want_stats):
        ctx.save_for_backward(field, weight)
        ctx.set_materialize_grads(False)
        # a field wider than the weight's Cin carries zero channels (noise concat padded to a 128-byte
        # line, deeplab_gan.Generator): the weight gets zero rows there and their gradient is dropped
        ctx.Cp = Cp = field.shape[-1] // 4
        assert Cp >= weight.shape[1]
        w2 = L.cached(weight, "s2d", lambda w: L.s2d_weight(_pad_cin(w, Cp)).to(torch.bfloat16).contiguous())
        y = K.conv_fprop(field, w2, L.S2D_DY, L.S2D_DX, want_stats=want_stats)
        if want_stats:
            ctx.mark_non_differentiable(y[1])
        return y

    @staticmethod
    def backward(ctx, g, *unused):
        if g is None:
            return None, None, None
        field, weight = ctx.saved_tensors
        g = _bf16c(g)
        gf = gw = None
        if ctx.needs_input_grad[1]:
            with _side_stream(g):
                Cout, Cin = weight.shape[:2]
                gw = L.s2d_weight_grad(K.conv_wgrad(field, g, L.S2D_DY, L.S2D_DX), (Cout, ctx.Cp, 3, 3))[:, :Cin].contiguous()
        if ctx.needs_input_grad[0]:
            Cp = ctx.Cp
            w2t = L.cached(weight, "s2d_t",
                           lambda w: L.s2d_weight(_pad_cin(w, Cp)).permute(2, 1, 0).to(torch.bfloat16).contiguous())
            gf = K.conv_fprop(g, w2t, [-v for v in L.S2D_DY], [-v for v in L.S2D_DX])
        _join(g)
        return gf, gw, None


def conv_s2d(field, weight, want_stats=False):
    return _ConvS2D.apply(field, weight, want_stats)


class _Depthwise(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, stride, dil):
        C = weight.shape[0]
        w9 = L.cached(weight, "dw9", lambda w: w.reshape(C, 9).t().contiguous().float())
        ctx.save_for_backward(x, w9)
        ctx.cfg = (stride, dil, tuple(x.shape))
        return K.dw_fwd(x, w9, stride, dil)

    @staticmethod
    def backward(ctx, g):
        x, w9 = ctx.saved_tensors
        stride, dil, shape = ctx.cfg
        g = _bf16c(g)
        gx = gw = None
        if ctx.needs_input_grad[1]:
            with _side_stream(g):
                gw = K.dw_wgrad(x, g, stride, dil).view(shape[3], 1, 3, 3)
        if ctx.needs_input_grad[0]:
            gx = K.dw_dgrad(g, w9, stride, dil, shape)
        _join(g)
        return gx, gw, None, None


def depthwise3x3(x, weight, stride=1, dilation=1):
    return _Depthwise.apply(x, weight, stride, dilation)


class _BNAct(torch.autograd.Function):
    """out = act(BatchNorm_train(y) [+ res]) with batch statistics (nn.BatchNorm2d training mode)."""

    @staticmethod
    def forward(ctx, y, gamma, beta, res, running_mean, running_var, eps, momentum, act, sums, nbt, dual, slot=None):
        M = y.numel() // y.shape[-1]
        if sums is None:
            sums = K.bn_stats(y)
        mean, invstd, a, b = K.bn_finalize(sums, M, gamma, beta, eps, momentum, running_mean, running_var, nbt)
        out = K.bn_apply(y, a, b, res=res, act=act, slope=SLOPE, out=slot.view if slot is not None else None)
        if slot is not None:
            out = out[...]      # a fresh view object of the concat buffer's channel slice
        ctx.has_res = res is not None
        # without a residual input act' = sign(a*y + b) is recomputed in backward (RECOMPUTE_BN_BWD): the stored
        # output is not read and no dyh tensor makes the round trip between the two passes
        # (not for dual handles: with a second gradient to add the stored path measures faster, 521 vs 718 us
        # on a 453 MB tensor -- the recompute kernels are half issue-bound, profiles/r01_bn_backward_ncu.txt)
        ctx.recompute = RECOMPUTE_BN_BWD and not ctx.has_res and not dual
        ctx.save_for_backward(y, out if (act and not ctx.recompute) else None, mean, invstd, gamma,
                              a if ctx.recompute else None, b if ctx.recompute else None)
        ctx.act = act
        if dual:   # a second handle of the same tensor: its gradient arrives separately (see bn_act)
            return out, out[...]
        return out

    @staticmethod
    def backward(ctx, g, g2=None):
        y, out, mean, invstd, gamma, a, b = ctx.saved_tensors
        if g is None:
            g, g2 = g2, None
        g = _bf16c(g)
        g2 = _bf16c(g2) if g2 is not None else None
        C = y.shape[-1]
        if ctx.recompute:
            dx, dyh, sums = K.bn_backward_recompute(g, a, b, y, mean, invstd, gamma, act=ctx.act, slope=SLOPE, g2=g2)
        else:
            dx, dyh, sums = K.bn_backward(g, out, y, mean, invstd, gamma, act=ctx.act, slope=SLOPE, g2=g2)
        ggamma = sums[C:] if ctx.needs_input_grad[1] else None
        gbeta = sums[:C] if ctx.needs_input_grad[2] else None
        gres = dyh if (ctx.has_res and ctx.needs_input_grad[3]) else None
        return dx, ggamma, gbeta, gres, None, None, None, None, None, None, None, None, None


class _BNActDepthwise(torch.autograd.Function):
    """t = depthwise3x3(act(BatchNorm_train(y))) (stride 1, dilation 1) without storing the activated
    tensor: the statistics come from the producing conv's epilogue, the normalisation and LeakyReLU
    are applied while the depthwise kernel loads y, and backward recomputes them the same way."""

    @staticmethod
    def forward(ctx, y, sums, gamma, beta, running_mean, running_var, eps, momentum, act, dw_weight, nbt):
        M = y.numel() // y.shape[-1]
        C = y.shape[-1]
        if sums is None:
            sums = K.bn_stats(y)
        mean, invstd, a, b = K.bn_finalize(sums, M, gamma, beta, eps, momentum, running_mean, running_var, nbt)
        w9 = L.cached(dw_weight, "dw9", lambda w: w.reshape(C, 9).t().contiguous().float())
        ctx.save_for_backward(y, mean, invstd, a, b, gamma, w9)
        ctx.act = act
        return K.dw_fwd_bn(y, w9, a, b, act, SLOPE)

    @staticmethod
    def backward(ctx, g):
        y, mean, invstd, a, b, gamma, w9 = ctx.saved_tensors
        g = _bf16c(g)
        C = y.shape[-1]
        gw = None
        if ctx.needs_input_grad[9]:
            with _side_stream(g):
                gw = K.dw_wgrad_bn(y, a, b, ctx.act, g, SLOPE).view(C, 1, 3, 3)
        gy = ggamma = gbeta = None
        if ctx.needs_input_grad[0]:
            gx = K.dw_dgrad(g, w9, 1, 1, tuple(y.shape))
            gy, _, bsums = K.bn_backward_recompute(gx, a, b, y, mean, invstd, gamma, act=ctx.act, slope=SLOPE)
            ggamma = bsums[C:] if ctx.needs_input_grad[2] else None
            gbeta = bsums[:C] if ctx.needs_input_grad[3] else None
        _join(g)
        return gy, None, ggamma, gbeta, None, None, None, None, None, gw, None


def bn_act_depthwise(y, bn, dw_weight, act=True):
    """depthwise3x3(bn_act(y, bn, act=act), dw_weight) for a training-mode BatchNorm, fused (see
    _BNActDepthwise).  `y` may be the (y, sums) pair of a conv called with want_stats=True."""
    sums = None
    if isinstance(y, tuple):
        y, sums = y
    if not needs_stats(bn):   # eval mode / Identity normaliser: the unfused pair (both differentiable)
        return depthwise3x3(bn_act(y, bn, act=act), dw_weight)
    return _BNActDepthwise.apply(y, sums, bn.weight, bn.bias, bn.running_mean, bn.running_var, bn.eps, bn.momentum,
                                 act, dw_weight, _nbt(bn))


class _AffineAct(torch.autograd.Function):
    """out = act(a*y + b [+ res]) with a, b that do not depend on the batch: nn.BatchNorm2d in eval mode
    (running statistics; validation train_gan.py:330-398, freeze_bn) or an Identity normaliser
    (--disable_batchnorm, train_gan.py:125: gamma = beta = None, mean 0, invstd 1)."""

    @staticmethod
    def forward(ctx, y, gamma, beta, mean, invstd, res, act, dual, slot=None):
        C = y.shape[-1]
        if gamma is None:
            a, b = _const_vec(1.0, C, y.device), _const_vec(0.0, C, y.device)
        else:
            a = (gamma.detach() * invstd).float().contiguous()
            b = (beta.detach() - mean * a).float().contiguous()
        out = K.bn_apply(y, a, b, res=res, act=act, slope=SLOPE, out=slot.view if slot is not None else None)
        if slot is not None:
            out = out[...]
        ctx.save_for_backward(y, out if act else None, gamma, mean, invstd)
        ctx.act, ctx.has_res = act, res is not None
        if dual:
            return out, out[...]
        return out

    @staticmethod
    def backward(ctx, g, g2=None):
        y, out, gamma, mean, invstd = ctx.saved_tensors
        if g is None:
            g, g2 = g2, None
        g = _bf16c(g)
        g2 = _bf16c(g2) if g2 is not None else None
        C = y.shape[-1]
        identity = gamma is None
        dx, dyh, grads = K.bn_backward_frozen(g, out, y, mean, invstd, gamma, act=ctx.act, slope=SLOPE, g2=g2,
                                              want_dx=not identity)
        ggamma = grads[C:] if (not identity and ctx.needs_input_grad[1]) else None
        gbeta = grads[:C] if (not identity and ctx.needs_input_grad[2]) else None
        gres = dyh if (ctx.has_res and ctx.needs_input_grad[5]) else None
        return (dyh if identity else dx), ggamma, gbeta, None, None, gres, None, None, None


_consts = {}


def _const_vec(value, C, device):
    key = (float(value), int(C), str(device))
    t = _consts.get(key)
    if t is None:
        t = _consts[key] = torch.full((C,), float(value), dtype=torch.float32, device=device)
    return t


def is_batchnorm(m):
    return isinstance(m, torch.nn.modules.batchnorm._BatchNorm)


def needs_stats(m):
    """True when the normaliser after a conv wants that conv's batch statistics (training-mode BatchNorm)."""
    return is_batchnorm(m) and m.training


def _nbt(bn):
    """num_batches_tracked as the device int64 the finalize kernel increments (None if untracked)."""
    t = bn.num_batches_tracked
    if t is None:
        return None
    if not t.is_cuda:   # a module left on the host: keep nn.BatchNorm2d's own bookkeeping
        t += 1
        return None
    return t


class ConcatBuffer:
    """A channel concatenation written in place: producers (bn_act / bilinear with slot=...) store their result
    straight into their channel slice of one NHWC buffer (the TMA-store / vector-store epilogues take a pixel
    stride), `cat()` hands the whole buffer to the consumer, and in backward the consumer's gradient is split
    into slice *views* -- no copy in either direction (torch.cat((x1, ..., x5), 1) of deeplab.py:664 and :377)."""

    class Slot:
        def __init__(self, view):
            self.view = view

    def __init__(self, N, H, W, channels, device):
        self.channels = list(channels)
        self.buf = K.act_empty((N, H, W, sum(self.channels)), device)
        self.offsets = [sum(self.channels[:i]) for i in range(len(self.channels))]

    def slot(self, i):
        a = self.offsets[i]
        return ConcatBuffer.Slot(self.buf[..., a:a + self.channels[i]])

    def cat(self, parts):
        return _CatInPlace.apply(self, *parts)


class _CatInPlace(torch.autograd.Function):
    @staticmethod
    def forward(ctx, cb, *parts):
        for i, p in enumerate(parts):   # every part must already live in its slice
            want = cb.buf[..., cb.offsets[i]:cb.offsets[i] + cb.channels[i]]
            if p.data_ptr() != want.data_ptr() or tuple(p.shape) != tuple(want.shape) or p.stride() != want.stride():
                raise RuntimeError("ConcatBuffer.cat: part %d was not produced into its slot" % i)
        ctx.cb = cb
        return cb.buf[...]

    @staticmethod
    def backward(ctx, g):
        cb = ctx.cb
        g = _bf16c(g)
        return (None,) + tuple(g[..., a:a + c] for a, c in zip(cb.offsets, cb.channels))


def bn_act(y, bn, res=None, act=True, sums=None, dual=False, slot=None):
    """`bn` is a torch.nn.BatchNorm2d-compatible module (weight, bias, running_*, eps, momentum).
    `y` may be the (y, sums) pair of a conv called with want_stats=True.
    dual=True returns two handles (out, out') of the same tensor for a result that feeds two paths
    (a block input: separable convs + skip/residual): autograd then delivers the two gradients
    separately and the BatchNorm backward kernel adds them while loading, instead of a separate
    full-tensor add kernel."""
    if isinstance(y, tuple):
        y, sums = y
    if not is_batchnorm(bn):      # Identity normaliser (train_gan.py:125): activation / residual only
        if not isinstance(bn, torch.nn.Identity):
            raise NotImplementedError("normalizer %s: BatchNorm2d and Identity are built" % type(bn).__name__)
        C = y.shape[-1]
        return _AffineAct.apply(y, None, None, _const_vec(0.0, C, y.device), _const_vec(1.0, C, y.device), res, act, dual, slot)
    if bn.training:
        return _BNAct.apply(y, bn.weight, bn.bias, res, bn.running_mean, bn.running_var, bn.eps,
                            bn.momentum, act, sums, _nbt(bn), dual, slot)
    with torch.no_grad():
        invstd = torch.rsqrt(bn.running_var + bn.eps).float().contiguous()
        mean = bn.running_mean.float().contiguous()
    return _AffineAct.apply(y, bn.weight, bn.bias, mean, invstd, res, act, dual, slot)


class _Subsample2(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        ctx.shape = tuple(x.shape)
        return K.subsample2(x)

    @staticmethod
    def backward(ctx, g):
        return K.subsample2_bwd(_bf16c(g), ctx.shape)


def subsample2(x):
    return _Subsample2.apply(x)


class _Bilinear(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, H, W, s2d, slot=None):
        ctx.cfg = (tuple(x.shape), H, W, s2d)
        if slot is not None:
            return K.bilinear_fwd(x, H, W, out=slot.view, s2d=s2d)[...]
        return K.bilinear_fwd(x, H, W, s2d=s2d)

    @staticmethod
    def backward(ctx, g):
        shape, H, W, s2d = ctx.cfg
        g = _bf16c(g)
        if shape[1] == 1 and shape[2] == 1 and not s2d:   # broadcast of a 1x1 map: the adjoint is the pixel sum
            return K.gap_fwd(g).to(torch.bfloat16).view(shape), None, None, None, None
        return K.bilinear_bwd(g, shape, H, W, s2d=s2d).to(torch.bfloat16), None, None, None, None


def bilinear(x, H, W, s2d=False, slot=None):
    """F.interpolate(x, (H, W), mode='bilinear', align_corners=True) on NHWC; s2d=True writes the
    2x2 space-to-depth field layout [N, H/2, W/2, 4C]; slot = a ConcatBuffer slot to write into."""
    return _Bilinear.apply(x, H, W, s2d, slot)


class _GlobalAvgPool(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        N, H, W, C = x.shape
        ctx.shape = (N, H, W, C)
        return (K.gap_fwd(x) * (1.0 / (H * W))).to(torch.bfloat16).view(N, 1, 1, C)

    @staticmethod
    def backward(ctx, g):
        N, H, W, C = ctx.shape
        return K.gap_bwd(_bf16c(g).view(N, C), N, H, W, C)


def global_avg_pool(x):
    return _GlobalAvgPool.apply(x)


class _Head(torch.autograd.Function):
    """flatten (C,H,W order of the reference) + Linear(K, 1) on an NHWC feature map."""

    @staticmethod
    def forward(ctx, x, weight, bias):
        N, H, W, C = x.shape
        w_hwc = L.cached(weight, "head", lambda w: w.view(C, H, W).permute(1, 2, 0).reshape(-1).to(torch.bfloat16).contiguous())
        logits = bias.detach().float().repeat(N).contiguous()
        K.head_fwd(x, w_hwc, logits)
        ctx.save_for_backward(x, w_hwc)
        return logits.view(N, 1)

    @staticmethod
    def backward(ctx, g):
        x, w_hwc = ctx.saved_tensors
        N, H, W, C = x.shape
        g = g.reshape(N).float().contiguous()
        gx, gw = K.head_bwd(g, x, w_hwc, need_dx=ctx.needs_input_grad[0], need_dw=ctx.needs_input_grad[1])
        if gw is not None:
            gw = gw.view(H, W, C).permute(2, 0, 1).reshape(1, -1)
        gb = g.sum().view(1) if ctx.needs_input_grad[2] else None
        return gx, gw, gb


def linear_head(x, weight, bias):
    return _Head.apply(x, weight, bias)


class _L1(torch.autograd.Function):
    """mean |a - b| with the sign gradient produced in the same pass (train_gan.py:146,281)."""

    @staticmethod
    def forward(ctx, a, b):
        n = a.numel()
        s, grad = K.l1_loss_grad(a, b, 1.0 / n, need_grad=ctx.needs_input_grad[0])
        ctx.save_for_backward(grad)
        return (s / n).float()

    @staticmethod
    def backward(ctx, g):
        (grad,) = ctx.saved_tensors
        return (grad * g.to(grad.dtype)), None


def l1_loss(a, b):
    return _L1.apply(a, b)


class _RegLoss(torch.autograd.Function):
    """The script's regression losses on fields with the gradient produced in the same pass:
    mean(l(a - b) [* w]) or, normalised, sum(l * w) / (sum(w) + eps) (utils/losses.py:101-112)."""

    @staticmethod
    def forward(ctx, a, b, weights, mode, beta, count, normalize, eps):
        n = float(count if count is not None else a.numel())
        if normalize and weights is not None:   # 1 / (sum w + eps) stays on the device: no host sync, capturable
            inv = (1.0 / (weights.float().sum() + eps)).reshape(1).float()
            s, grad = K.reg_loss_grad(a, b, mode, 1.0, weights=weights, beta=beta, gscale_dev=inv,
                                      need_grad=ctx.needs_input_grad[0])
            out = (s * inv.double()[0]).float()
        else:
            s, grad = K.reg_loss_grad(a, b, mode, 1.0 / n, weights=weights, beta=beta, need_grad=ctx.needs_input_grad[0])
            out = (s / n).float()
        ctx.save_for_backward(grad)
        return out

    @staticmethod
    def backward(ctx, g):
        (grad,) = ctx.saved_tensors
        return (grad * g.to(grad.dtype)), None, None, None, None, None, None, None


def regression_loss(a, b, weights=None, mode=0, beta=1.0, count=None, normalize=False, eps=1e-8):
    return _RegLoss.apply(a, b, weights, mode, beta, count, normalize, eps)


class _ToField(torch.autograd.Function):
    """fp32 NCHW -> bf16 field layout [N, H/2, W/2, (py, px, c)] (and back in backward)."""

    @staticmethod
    def forward(ctx, x):
        ctx.C = x.shape[1]
        return K.nchw_to_field(x.float().contiguous())

    @staticmethod
    def backward(ctx, g):
        return K.field_to_nchw(_bf16c(g), ctx.C)


class _FromField(torch.autograd.Function):
    @staticmethod
    def forward(ctx, field, C):
        return K.field_to_nchw(field, C)

    @staticmethod
    def backward(ctx, g):
        return K.nchw_to_field(g.float().contiguous()), None


class _AppendNoise(torch.autograd.Function):
    """cat((input_raw, input_noise), dim=1) of deeplab_gan.py:90 on fields; Cp = padded channel count."""

    @staticmethod
    def forward(ctx, field, noise, Cp):
        ctx.C = field.shape[-1] // 4
        return K.field_append_noise(field, noise, Cp)

    @staticmethod
    def backward(ctx, g):
        N, H2, W2, _ = g.shape
        gf = g.view(N, H2, W2, 4, -1)[..., :ctx.C].reshape(N, H2, W2, 4 * ctx.C) if ctx.needs_input_grad[0] else None
        return gf, None, None   # the noise is a sample, not a parameter: no gradient is asked for


def append_noise(field, noise, Cp):
    return _AppendNoise.apply(field, noise, Cp)


def to_field(x_nchw):
    return _ToField.apply(x_nchw)


def from_field(field, C):
    return _FromField.apply(field, C)


# ---------------------------------------------------------------------------------------------
# Deconv / Deconv1x decoders, odd-sized fields (csrc/deconv.cu)
# ---------------------------------------------------------------------------------------------
class _ConvT(torch.autograd.Function):
    """nn.ConvTranspose2d(kernel 3, stride 2, padding p, output_padding) (deeplab.py:406-431) on the tcgen05
    implicit-GEMM kernel: out[2i - p + k] += x[i] * w[k] is the stride-1 conv of the zero-stuffed input
    (z[2i] = x[i]) with tap offsets p - k, so forward, data gradient and weight gradient are the same three
    launches as every other conv; weight is the module's [Cin, Cout, 3, 3] parameter."""

    @staticmethod
    def forward(ctx, x, weight, pad, out_pad, want_stats):
        N, h, w, Cin = x.shape
        Cout = weight.shape[1]
        H = (h - 1) * 2 - 2 * pad + 3 + out_pad[0]
        W = (w - 1) * 2 - 2 * pad + 3 + out_pad[1]
        dy = [pad - ky for ky in range(3) for _ in range(3)]
        dx = [pad - kx for _ in range(3) for kx in range(3)]
        ctx.save_for_backward(x, weight)
        ctx.set_materialize_grads(False)
        ctx.cfg = (H, W, dy, dx)
        z = K.zero_stuff2(x, H, W)
        wg = L.cached(weight, "convT", lambda t: L._padded_rows(t.permute(1, 2, 3, 0).reshape(Cout, 9, Cin), Cout, 9, Cin))
        y = K.conv_fprop(z, wg, dy, dx, want_stats=want_stats)
        if want_stats:
            ctx.mark_non_differentiable(y[1])
        return y

    @staticmethod
    def backward(ctx, g, *unused):
        if g is None:
            return None, None, None, None, None
        x, weight = ctx.saved_tensors
        H, W, dy, dx = ctx.cfg
        N, h, w, Cin = x.shape
        Cout = weight.shape[1]
        g = _bf16c(g)
        gx = gw = None
        if ctx.needs_input_grad[1]:
            with _side_stream(g):
                gwg = K.conv_wgrad(K.zero_stuff2(x, H, W), g, dy, dx)            # [Cout, 9, Cin]
                gw = gwg.reshape(Cout, 3, 3, Cin).permute(3, 0, 1, 2).contiguous()
        if ctx.needs_input_grad[0]:
            wt = L.cached(weight, "convT_t", lambda t: L._padded_rows(t.permute(0, 2, 3, 1).reshape(Cin, 9, Cout), Cin, 9, Cout))
            gz = K.conv_fprop(g, wt, [-v for v in dy], [-v for v in dx])
            gx = K.zero_stuff2_bwd(gz, h, w)
        _join(g)
        return gx, gw, None, None, None


def conv_transpose2d(x, weight, padding, output_padding, want_stats=False):
    return _ConvT.apply(x, weight, int(padding), tuple(output_padding), want_stats)


class _AvgPool2(torch.autograd.Function):
    """nn.AvgPool2d(2, stride=1, padding=pad) followed (act=True) by LeakyReLU(0.2), one kernel each way."""

    @staticmethod
    def forward(ctx, x, pad, act):
        y = K.avgpool2_fwd(x, pad, act, SLOPE)
        ctx.save_for_backward(y if act else None)
        ctx.cfg = (tuple(x.shape), pad, act)
        return y

    @staticmethod
    def backward(ctx, g):
        (y,) = ctx.saved_tensors
        shape, pad, act = ctx.cfg
        return K.avgpool2_bwd(_bf16c(g), y, shape, pad, act, SLOPE), None, None


def avgpool2(x, pad, act=False):
    return _AvgPool2.apply(x, int(pad), bool(act))


class _NhwcToField(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        ctx.hw = (x.shape[1], x.shape[2])
        return K.nhwc_to_field(x)

    @staticmethod
    def backward(ctx, g):
        return K.field_to_nhwc(_bf16c(g).contiguous(), *ctx.hw)


class _FieldToNhwc(torch.autograd.Function):
    @staticmethod
    def forward(ctx, field, H, W):
        return K.field_to_nhwc(field, H, W)

    @staticmethod
    def backward(ctx, g):
        return K.nhwc_to_field(_bf16c(g)), None, None


def nhwc_to_field(x):
    return _NhwcToField.apply(x)


def field_to_nhwc(field, H, W):
    return _FieldToNhwc.apply(field, H, W)


class _ToFieldPadded(torch.autograd.Function):
    """fp32 NCHW of any size / channel count -> field with Cp channels per sub-pixel, zero padded."""

    @staticmethod
    def forward(ctx, x, Cp):
        ctx.cfg = tuple(x.shape)
        return K.stage_normalize_padded(x.float().contiguous(), None, None, None, Cp)

    @staticmethod
    def backward(ctx, g):
        N, C, H, W = ctx.cfg
        return K.field_to_nchw_padded(_bf16c(g).contiguous(), C, H, W), None


class _FromFieldPadded(torch.autograd.Function):
    @staticmethod
    def forward(ctx, field, C, H, W):
        ctx.Cp = field.shape[-1] // 4
        return K.field_to_nchw_padded(field, C, H, W)

    @staticmethod
    def backward(ctx, g):
        return K.stage_normalize_padded(g.float().contiguous(), None, None, None, ctx.Cp), None, None, None


def to_field_padded(x_nchw, Cp):
    return _ToFieldPadded.apply(x_nchw, Cp)


def from_field_padded(field, C, H, W):
    return _FromFieldPadded.apply(field, C, H, W)


def pad8(C):
    return (C + 7) // 8 * 8
