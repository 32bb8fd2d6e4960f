"""DeepLabv3+ / modified aligned Xception on the sm_100a kernels.

Module and parameter names mirror src/deepCam/architecture/gpsro/deeplab.py so state dicts
interchange with the reference (`xception_features.block4.rep.1.conv1.weight`, ...).  nn.Conv2d /
nn.BatchNorm2d objects are used only as parameter holders with the reference's shapes; the
arithmetic is the C-ABI kernels through functional.py on bf16 NHWC activations.

In-place LeakyReLU semantics of the reference (deeplab.py:100,137-141: the first `rep` element
mutates the block input, so the skip path and `low_level_feat` see the activated tensor) are
reproduced by fusing that activation into the producer of each block input.
"""
import math
import os

import torch
import torch.nn as nn

from . import functional as F_


# debugging switch: False materialises every BatchNorm output (bn_apply) before its depthwise conv
FUSE_BN_INTO_DEPTHWISE = os.environ.get("BGR_NO_BN_DW_FUSION", "0") != "1"


def _hold_conv(cin, cout, k, groups=1, bias=False):
    return nn.Conv2d(cin, cout, k, groups=groups, bias=bias)


class _Act(nn.Identity):
    """Place holder of the reference's nn.LeakyReLU(0.2, inplace=True) modules inside nn.Sequential containers:
    it keeps the `rep.N` / `last_conv.N` indices of the state-dict keys; the activation itself is fused into the
    neighbouring kernel.  A class of its own so that an Identity *normaliser* (--disable_batchnorm,
    train_gan.py:125) in the same container is not mistaken for it."""


def _is_norm(m):
    """BatchNorm2d or the Identity normaliser (anything the `normalizer` argument built), not an activation slot."""
    return F_.is_batchnorm(m) or (isinstance(m, nn.Identity) and not isinstance(m, _Act))


class SeparableConv2d_same(nn.Module):
    """fixed_padding + depthwise 3x3 + pointwise 1x1 (deeplab.py:75-87)."""

    def __init__(self, inplanes, planes, kernel_size=3, stride=1, dilation=1, bias=False):
        super().__init__()
        assert kernel_size == 3 and not bias
        self.conv1 = _hold_conv(inplanes, inplanes, 3, groups=inplanes)
        self.pointwise = _hold_conv(inplanes, planes, 1)
        self.stride, self.dilation = stride, dilation

    def forward(self, x, want_stats=False):
        """want_stats=True returns (y, sums): the BatchNorm sums of y from the 1x1 conv's epilogue."""
        x = F_.depthwise3x3(x, self.conv1.weight, self.stride, self.dilation)
        return F_.conv2d(x, self.pointwise.weight, want_stats=want_stats)


class Block(nn.Module):
    """Xception block (deeplab.py:90-143).  `forward(a)` takes the already-activated block input
    when start_with_relu (the reference's in-place relu) and returns the pre-activation sum's
    fused result: act_out=True applies the *next* block's in-place LeakyReLU."""

    def __init__(self, inplanes, planes, reps, stride=1, dilation=1, start_with_relu=True, grow_first=True,
                 is_last=False, normalizer=nn.BatchNorm2d):
        super().__init__()
        if planes != inplanes or stride != 1:
            self.skip = _hold_conv(inplanes, planes, 1)
            self.skipbn = normalizer(planes)
        else:
            self.skip = None
        self.relu = _Act()
        self.stride = stride
        rep = []
        filters = inplanes
        if grow_first:
            rep += [_Act(), SeparableConv2d_same(inplanes, planes, 3, 1, dilation), normalizer(planes)]
            filters = planes
        for _ in range(reps - 1):
            rep += [_Act(), SeparableConv2d_same(filters, filters, 3, 1, dilation), normalizer(filters)]
        if not grow_first:
            rep += [_Act(), SeparableConv2d_same(inplanes, planes, 3, 1, dilation), normalizer(planes)]
        if not start_with_relu:
            rep = rep[1:]
        if stride != 1:
            rep.append(SeparableConv2d_same(planes, planes, 3, stride=2))
        if stride == 1 and is_last:
            rep.append(SeparableConv2d_same(planes, planes, 3, stride=1))
        self.rep = nn.Sequential(*rep)

    def forward(self, a, act_out, dual_out=False):
        """`a` may be a pair of handles (a_main, a_side) of the same tensor (bn_act(dual=True)): the
        separable convs read a_main, the skip / residual path a_side, so that their gradients reach
        the producing BatchNorm's backward separately and are summed there.  dual_out=True returns
        such a pair for the next block."""
        a, a_side = a if isinstance(a, tuple) else (a, a)
        mods = list(self.rep)
        x = a
        i = 0
        pending = None  # (raw conv output [+ its BatchNorm sums], bn, followed by LeakyReLU?)
        tail_bn = None
        while i < len(mods):
            m = mods[i]
            if isinstance(m, SeparableConv2d_same):
                has_bn = i + 1 < len(mods) and _is_norm(mods[i + 1])
                want = has_bn and F_.needs_stats(mods[i + 1])
                if pending is not None:
                    py, pbn, pact = pending
                    pending = None
                    if m.stride == 1 and m.dilation == 1 and F_.needs_stats(pbn) and FUSE_BN_INTO_DEPTHWISE:
                        # BatchNorm (+LeakyReLU) applied while the depthwise kernel loads: never stored
                        t = F_.bn_act_depthwise(py, pbn, m.conv1.weight, act=pact)
                        y = F_.conv2d(t, m.pointwise.weight, want_stats=want)
                    else:
                        y = m(F_.bn_act(py, pbn, act=pact), want_stats=want)
                else:
                    y = m(x, want_stats=want)
                if has_bn:
                    bn = mods[i + 1]
                    i += 2
                    act = i < len(mods) and isinstance(mods[i], _Act)  # the (place-held) relu
                    if act:
                        i += 1
                    pending = (y, bn, act)
                    continue
                x = y
            i += 1
        if pending is not None:
            tail_bn = pending  # last rep element is a BN: fuse with the residual add
        if self.skip is not None:
            s_in = F_.subsample2(a_side) if self.stride == 2 else a_side
            s = F_.conv2d(s_in, self.skip.weight, want_stats=F_.needs_stats(self.skipbn))
            if tail_bn is not None:
                x = F_.bn_act(tail_bn[0], tail_bn[1], act=False)
            return F_.bn_act(s, self.skipbn, res=x, act=act_out, dual=dual_out)
        if tail_bn is not None:
            return F_.bn_act(tail_bn[0], tail_bn[1], res=a_side, act=act_out, dual=dual_out)
        raise NotImplementedError("identity-skip block ending in a convolution does not occur in Xception")


class Xception(nn.Module):
    """Modified aligned Xception trunk (deeplab.py:146-278), os=16 or 8."""

    def __init__(self, inplanes=3, os=16, pretrained=False, normalizer=nn.BatchNorm2d):
        super().__init__()
        assert not pretrained, "pretrained weights need network access"
        if os == 16:
            b3s, mid_rate, exit_rates = 2, 1, (1, 2)
        elif os == 8:
            b3s, mid_rate, exit_rates = 1, 2, (2, 4)
        else:
            raise NotImplementedError
        self.inplanes = inplanes
        self.conv1 = nn.Conv2d(inplanes, 128, 3, stride=2, padding=1, bias=False)
        self.bn1 = normalizer(128)
        self.relu = _Act()
        self.conv2 = nn.Conv2d(128, 128, 3, stride=1, padding=1, bias=False)
        self.bn2 = normalizer(128)
        self.block1 = Block(128, 128, reps=2, stride=2, start_with_relu=False, normalizer=normalizer)
        self.block2 = Block(128, 256, reps=2, stride=2, normalizer=normalizer)
        self.block3 = Block(256, 728, reps=2, stride=b3s, is_last=True, normalizer=normalizer)
        for i in range(4, 20):
            setattr(self, "block%d" % i, Block(728, 728, reps=3, dilation=mid_rate, normalizer=normalizer))
        self.block20 = Block(728, 1024, reps=2, dilation=exit_rates[0], grow_first=False, is_last=True,
                             normalizer=normalizer)
        self.conv3 = SeparableConv2d_same(1024, 1536, 3, dilation=exit_rates[1])
        self.bn3 = normalizer(1536)
        self.conv4 = SeparableConv2d_same(1536, 1536, 3, dilation=exit_rates[1])
        self.bn4 = normalizer(1536)
        self.conv5 = SeparableConv2d_same(1536, 2048, 3, dilation=exit_rates[1])
        self.bn5 = normalizer(2048)
        self._init_weight()

    def _init_weight(self):
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight)
            elif isinstance(m, nn.BatchNorm2d):
                nn.init.ones_(m.weight)
                nn.init.zeros_(m.bias)

    def forward(self, field):
        """field: bf16 [N, H/2, W/2, 4*inplanes] (2x2 space-to-depth).  Returns (features
        [N,H/16,W/16,2048], low_level [N,H/4,W/4,128]) — both activated, as the reference's
        in-place LeakyReLUs leave them."""
        x = F_.bn_act(F_.conv_s2d(field, self.conv1.weight, F_.needs_stats(self.bn1)), self.bn1, act=True)
        x = F_.bn_act(F_.conv2d(x, self.conv2.weight, want_stats=F_.needs_stats(self.bn2)), self.bn2, act=True, dual=True)
        x = self.block1(x, act_out=True, dual_out=True)
        low = x[0]
        x = self.block2(x, act_out=True, dual_out=True)
        x = self.block3(x, act_out=True, dual_out=True)
        for i in range(4, 20):
            x = getattr(self, "block%d" % i)(x, act_out=True, dual_out=True)
        x = self.block20(x, act_out=False)
        x = F_.bn_act(self.conv3(x, F_.needs_stats(self.bn3)), self.bn3, act=True)
        x = F_.bn_act(self.conv4(x, F_.needs_stats(self.bn4)), self.bn4, act=True)
        x = F_.bn_act(self.conv5(x, F_.needs_stats(self.bn5)), self.bn5, act=True)
        return x, low


def _init_decoder(mod):
    gain = nn.init.calculate_gain("leaky_relu", 0.2)
    for m in mod.modules():
        if isinstance(m, nn.Conv2d):
            n = m.kernel_size[0] * m.kernel_size[1] * m.out_channels
            nn.init.normal_(m.weight, mean=0.0, std=gain / math.sqrt(n))
        elif isinstance(m, nn.BatchNorm2d):
            nn.init.ones_(m.weight)
            nn.init.zeros_(m.bias)


class ASPP_module(nn.Module):
    """deeplab.py:322-355."""

    def __init__(self, inplanes, planes, rate, normalizer=nn.BatchNorm2d):
        super().__init__()
        k = 1 if rate == 1 else 3
        self.atrous_convolution = nn.Conv2d(inplanes, planes, k, padding=0 if rate == 1 else rate,
                                            dilation=rate, bias=False)
        self.bn = normalizer(planes)
        self.relu = _Act()
        self.rate = rate
        nn.init.kaiming_normal_(self.atrous_convolution.weight)

    def forward(self, x, slot=None):
        """slot: the ConcatBuffer slice the activated result is written into (no torch.cat afterwards)."""
        return F_.bn_act(F_.conv2d(x, self.atrous_convolution.weight, dilation=self.rate,
                                   want_stats=F_.needs_stats(self.bn)), self.bn, act=True, slot=slot)


class InterpolationUpsampler(nn.Module):
    """deeplab.py:358-381; returns the output field (2x2 space-to-depth bf16)."""

    def __init__(self, n_output, normalizer=nn.BatchNorm2d):
        super().__init__()
        self.last_conv = nn.Sequential(nn.Conv2d(304, 256, 3, padding=1, bias=False), normalizer(256), _Act(),
                                       nn.Conv2d(256, 256, 3, padding=1, bias=False), normalizer(256), _Act(),
                                       nn.Conv2d(256, n_output, 1))
        _init_decoder(self)

    def forward(self, x, low_conv, low_bn, H, W):
        """low_conv / low_bn: the raw 1x1-reduced low-level features and their normaliser; both halves of the
        304-channel concat (deeplab.py:377) are written straight into one buffer."""
        h4, w4 = int(math.ceil(H / 4)), int(math.ceil(W / 4))
        c_low = (low_conv[0] if isinstance(low_conv, tuple) else low_conv).shape[-1]
        cb = F_.ConcatBuffer(x.shape[0], h4, w4, (x.shape[-1], c_low), x.device)
        x = cb.cat([F_.bilinear(x, h4, w4, slot=cb.slot(0)), F_.bn_act(low_conv, low_bn, act=True, slot=cb.slot(1))])
        lc = self.last_conv
        x = F_.bn_act(F_.conv2d(x, lc[0].weight, want_stats=F_.needs_stats(lc[1])), lc[1], act=True)
        x = F_.bn_act(F_.conv2d(x, lc[3].weight, want_stats=F_.needs_stats(lc[4])), lc[4], act=True)
        x = F_.conv2d(x, lc[6].weight, lc[6].bias)
        return F_.bilinear(x, H, W, s2d=True)


def _pad_cout(w, dim=0):
    """Weight with its output-channel dimension zero-padded to a multiple of 8 (45 -> 48): the kernels move
    128-bit channel groups; the padded outputs are exact zeros and their gradient rows are dropped by autograd."""
    C = w.shape[dim]
    Cp = F_.pad8(C)
    if Cp == C:
        return w
    pad = [0, 0] * (w.dim() - 1 - dim) + [0, Cp - C]
    return torch.nn.functional.pad(w, pad)


def _pad_cin(w, Cp):
    return w if w.shape[1] == Cp else torch.nn.functional.pad(w, (0, 0, 0, 0, 0, Cp - w.shape[1]))


def _init_deconv(mod):
    gain = nn.init.calculate_gain("leaky_relu", 0.2)
    for m in mod.modules():
        if isinstance(m, (nn.Conv2d, nn.ConvTranspose2d)):
            n = m.kernel_size[0] * m.kernel_size[1] * m.out_channels
            nn.init.normal_(m.weight, mean=0.0, std=gain / math.sqrt(n))
        elif isinstance(m, nn.BatchNorm2d):
            nn.init.ones_(m.weight)
            nn.init.zeros_(m.bias)


class _Pool(nn.Identity):
    """Place holder of nn.AvgPool2d(2, stride=1, padding=p) in the decoder's nn.Sequential containers (no
    parameters; keeps the indices of the state-dict keys); the pooling runs fused with the LeakyReLU behind it."""

    def __init__(self, padding):
        super().__init__()
        self.padding = padding


class DeconvUpsampler(nn.Module):
    """deeplab.py:398-444.  Four ConvTranspose2d(k3, s2) whose paddings / output paddings are hard-wired for the
    19 x 37 GPSRO fields (2x3 -> 3x6 -> [pool] 2x5 -> 6x11 -> [pool] 5x10 -> 10x19 -> [pool] 9x18 -> 18x36), each
    followed by normaliser, AvgPool2d(2, 1) and LeakyReLU; state-dict keys deconv1.0.weight, deconv1.1.*, ...
    as the reference's.  The transposed convs run on the implicit-GEMM kernel (functional._ConvT)."""

    def __init__(self, n_output, normalizer=nn.BatchNorm2d, nn_pooling=True):
        super().__init__()
        pool = (lambda: _Pool(0)) if nn_pooling else (lambda: nn.Identity())
        ct = lambda cin, cout, p, op: nn.ConvTranspose2d(cin, cout, 3, stride=2, padding=p, output_padding=op, bias=False)
        self.deconv1 = nn.Sequential(ct(256, 256, 1, (0, 1)), normalizer(256), pool(), _Act())
        self.deconv2 = nn.Sequential(ct(256, 256, 0, (1, 0)), normalizer(256), pool(), _Act())
        self.conv1 = nn.Sequential(nn.Conv2d(304, 256, 3, padding=1, bias=False), normalizer(256), _Act(),
                                   nn.Conv2d(256, 256, 3, padding=1, bias=False), normalizer(256), _Act(),
                                   nn.Conv2d(256, 256, 1))
        self.deconv3 = nn.Sequential(ct(256, 128, 1, (1, 0)), normalizer(128), pool(), _Act())
        self.last_deconv = nn.Sequential(ct(128, n_output, 1, (1, 1)))
        _init_deconv(self)

    @staticmethod
    def _stage(seq, x):
        ct, norm, pool = seq[0], seq[1], seq[2]
        y = F_.conv_transpose2d(x, ct.weight, ct.padding[0], ct.output_padding, want_stats=F_.needs_stats(norm))
        if isinstance(pool, _Pool):   # normaliser -> AvgPool2d(2, 1) -> LeakyReLU
            return F_.avgpool2(F_.bn_act(y, norm, act=False), pool.padding, act=True)
        return F_.bn_act(y, norm, act=True)

    def forward(self, x, low_level_features):
        x = self._stage(self.deconv2, self._stage(self.deconv1, x))
        x = torch.cat((x, low_level_features), dim=3)
        c = self.conv1
        x = F_.bn_act(F_.conv2d(x, c[0].weight, want_stats=F_.needs_stats(c[1])), c[1], act=True)
        x = F_.bn_act(F_.conv2d(x, c[3].weight, want_stats=F_.needs_stats(c[4])), c[4], act=True)
        x = F_.conv2d(x, c[6].weight, c[6].bias)
        x = self._stage(self.deconv3, x)
        ld = self.last_deconv[0]
        return F_.conv_transpose2d(x, _pad_cout(ld.weight, 1), ld.padding[0], ld.output_padding)


class UpsamplerExtension(nn.Module):
    """deeplab.py:466-500: full-resolution two-conv skip from the raw input, joined with the normalised,
    pooled (18x36 -> 19x37) decoder output."""

    def __init__(self, n_input, n_output, normalizer=nn.BatchNorm2d, nn_pooling=True):
        super().__init__()
        self.init_norm = nn.Sequential(normalizer(128), _Pool(1) if nn_pooling else nn.Identity(), _Act())
        self.conv1 = nn.Sequential(nn.Conv2d(n_input, 64, 3, padding=1, bias=False), normalizer(64), _Act(),
                                   nn.Conv2d(64, 128, 3, padding=1, bias=False), normalizer(128), _Act())
        self.conv2 = nn.Sequential(nn.Conv2d(256, 64, 3, padding=1, bias=False), normalizer(64), _Act(),
                                   nn.Conv2d(64, n_output, 3, padding=1, bias=False))
        _init_deconv(self)

    def forward(self, inp, x):
        """inp: the raw generator input as bf16 NHWC [N,H,W,Cp] (channels zero-padded to a multiple of 8)."""
        a, b = self.conv1, self.conv2
        skip = F_.bn_act(F_.conv2d(inp, _pad_cin(a[0].weight, inp.shape[-1]), want_stats=F_.needs_stats(a[1])), a[1], act=True)
        skip = F_.bn_act(F_.conv2d(skip, a[3].weight, want_stats=F_.needs_stats(a[4])), a[4], act=True)
        n = self.init_norm
        if isinstance(n[1], _Pool):
            x = F_.avgpool2(F_.bn_act(x, n[0], act=False), n[1].padding, act=True)
        else:
            x = F_.bn_act(x, n[0], act=True)
        x = torch.cat((x, skip), dim=3)
        x = F_.bn_act(F_.conv2d(x, b[0].weight, want_stats=F_.needs_stats(b[1])), b[1], act=True)
        return F_.conv2d(x, _pad_cout(b[3].weight, 0))


class DeepLabv3_plus(nn.Module):
    """deeplab.py:583-684: upsampler_type 'Interpolate' (shape-generic), 'Deconv' and 'Deconv1x'
    (the reference's default / its GAN run script's choice; 19 x 37 fields only, like the reference's)."""

    def __init__(self, n_input=3, n_output=21, os=16, upsampler_type="Interpolate", pretrained=False, _print=False,
                 normalizer=nn.BatchNorm2d, nn_pooling=True):
        super().__init__()
        if upsampler_type not in ("Interpolate", "Deconv", "Deconv1x"):
            raise NotImplementedError("Error, upsampler {} not implemented.".format(upsampler_type))
        self.upsampler_type = upsampler_type
        self.n_output = n_output
        self.nn_pooling = nn_pooling
        self.xception_features = Xception(n_input, os, pretrained, normalizer)
        rates = [1, 6, 12, 18] if os == 16 else [1, 12, 24, 36]
        self.aspp1 = ASPP_module(2048, 256, rates[0], normalizer)
        self.aspp2 = ASPP_module(2048, 256, rates[1], normalizer)
        self.aspp3 = ASPP_module(2048, 256, rates[2], normalizer)
        self.aspp4 = ASPP_module(2048, 256, rates[3], normalizer)
        self.relu = _Act()
        self.global_avg_pool = nn.Sequential(_Act(), nn.Conv2d(2048, 256, 1, bias=False), normalizer(256),
                                             _Act())
        self.conv1 = nn.Conv2d(1280, 256, 1, bias=False)
        self.bn1 = normalizer(256)
        self.conv2 = nn.Conv2d(128, 48, 1, bias=False)
        self.bn2 = normalizer(48)
        if upsampler_type == "Interpolate":
            self.upsample = InterpolationUpsampler(n_output, normalizer)
        else:
            self.upsample = DeconvUpsampler(128 if upsampler_type == "Deconv1x" else n_output, normalizer, nn_pooling)
            if upsampler_type == "Deconv1x":
                self.upsample_extension = UpsamplerExtension(n_input, n_output, normalizer, nn_pooling)
            else:
                self.final_pool = _Pool(1) if nn_pooling else nn.Identity()

    def forward(self, field, hw=None):
        """field: bf16 [N, ceil(H/2), ceil(W/2), 4*Cp]; hw = the true (H, W) when the field carries a zero row /
        column (odd sizes).  Returns the output field (same padding convention, channels padded to 8)."""
        N, H2, W2, _ = field.shape
        H, W = hw if hw is not None else (2 * H2, 2 * W2)
        x, low = self.xception_features(field)
        h, w = x.shape[1], x.shape[2]
        # the five ASPP branches write their 256 channels each into one 1280-channel buffer (deeplab.py:664)
        cb = F_.ConcatBuffer(N, h, w, (256,) * 5, x.device)
        parts = [m(x, slot=cb.slot(i)) for i, m in enumerate((self.aspp1, self.aspp2, self.aspp3, self.aspp4))]
        gp = self.global_avg_pool
        x5 = F_.bn_act(F_.conv2d(F_.global_avg_pool(x), gp[1].weight, want_stats=F_.needs_stats(gp[2])), gp[2], act=True)
        parts.append(F_.bilinear(x5, h, w, slot=cb.slot(4)))
        x = cb.cat(parts)
        x = F_.bn_act(F_.conv2d(x, self.conv1.weight, want_stats=F_.needs_stats(self.bn1)), self.bn1, act=True)
        low_raw = F_.conv2d(low, self.conv2.weight, want_stats=F_.needs_stats(self.bn2))
        if self.upsampler_type == "Interpolate":
            if (H | W) & 1:
                raise NotImplementedError("the Interpolate upsampler is built for even field sizes")
            return self.upsample(x, low_raw, self.bn2, H, W)
        low = F_.bn_act(low_raw, self.bn2, act=True)
        x = self.upsample(x, low)
        if self.upsampler_type == "Deconv1x":      # deeplab.py:676-677
            x = self.upsample_extension(F_.field_to_nhwc(field, H, W), x)
        elif isinstance(self.final_pool, _Pool):   # deeplab.py:678-679
            x = F_.avgpool2(x, self.final_pool.padding, act=False)
        if (x.shape[1], x.shape[2]) != (H, W):
            raise RuntimeError("Deconv upsampler produced %dx%d for a %dx%d input (its paddings are hard-wired for "
                               "19x37 fields, deeplab.py:406-431)" % (x.shape[1], x.shape[2], H, W))
        return F_.nhwc_to_field(x)

    def freeze_bn(self):
        for m in self.modules():
            if isinstance(m, nn.BatchNorm2d):
                m.eval()
