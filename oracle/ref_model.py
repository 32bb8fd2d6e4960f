"""TEST INFRASTRUCTURE — plain PyTorch fp32 (NCHW, ATen ops) restatement of the reference's GAN
networks and losses, used only as the parity checker by tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs.  Product code never imports it.

Restates (state-dict keys identical, so weights interchange with the reference and the product):
  Sep / Blk / XceptionRef     src/deepCam/architecture/gpsro/deeplab.py:75-87, 90-143, 146-278
  Aspp / DeepLabRef           deeplab.py:322-342, 358-381 (Interpolate upsampler), 583-684
  _Deconv / _Extension        deeplab.py:398-444, 466-500 (Deconv / Deconv1x upsamplers, native 19x37 fields)
  GeneratorRef / DiscriminatorRef   architecture/gpsro/deeplab_gan.py:12-39, 65-94
  GANLossRef                  src/deepCam/utils/losses.py:129-172
  gan_step_ref                src/deepCam/gpsro_train/train_gan.py:244-301 (one G+D step)

Pinned: tests/test_oracle_model.py checks it against tests/golden/model_golden.pt, noise_golden.pt, gp_golden.pt and
deconv_golden.pt -- vectors (outputs, logits, losses, BatchNorm statistics, every per-parameter gradient norm)
produced by the reference's OWN modules imported from /root/reference in the build container by the committed
scripts tests/golden/make_*_golden.py.  The test itself never imports the reference (it must run on the GPU box).
The arithmetic itself lives in PyTorch ATen
(un-vendored third-party code; the reference pins only the NGC 20.12 container).
"""
import math

import torch
import torch.nn as nn
import torch.nn.functional as F


class Sep(nn.Module):
    def __init__(self, cin, cout, stride=1, dilation=1):
        super().__init__()
        self.conv1 = nn.Conv2d(cin, cin, 3, stride, 0, dilation, groups=cin, bias=False)
        self.pointwise = nn.Conv2d(cin, cout, 1, bias=False)
        self.d = dilation

    def forward(self, x):
        return self.pointwise(self.conv1(F.pad(x, (self.d,) * 4)))


class Blk(nn.Module):
    def __init__(self, cin, cout, reps, stride=1, dilation=1, start_with_relu=True, grow_first=True, is_last=False):
        super().__init__()
        self.skip = None
        if cout != cin or stride != 1:
            self.skip = nn.Conv2d(cin, cout, 1, stride=stride, bias=False)
            self.skipbn = nn.BatchNorm2d(cout)
        self.relu = nn.LeakyReLU(0.2, inplace=True)
        seq, c = [], cin
        if grow_first:
            seq += [self.relu, Sep(cin, cout, 1, dilation), nn.BatchNorm2d(cout)]
            c = cout
        for _ in range(reps - 1):
            seq += [self.relu, Sep(c, c, 1, dilation), nn.BatchNorm2d(c)]
        if not grow_first:
            seq += [self.relu, Sep(cin, cout, 1, dilation), nn.BatchNorm2d(cout)]
        if not start_with_relu:
            seq = seq[1:]
        if stride != 1:
            seq.append(Sep(cout, cout, 2))
        elif is_last:
            seq.append(Sep(cout, cout, 1))
        self.rep = nn.Sequential(*seq)

    def forward(self, inp):
        x = self.rep(inp)  # NB: the leading in-place LeakyReLU mutates `inp` (reference behaviour)
        s = inp if self.skip is None else self.skipbn(self.skip(inp))
        x += s
        return x


class XceptionRef(nn.Module):
    def __init__(self, inplanes, os=16):
        super().__init__()
        b3s, mid, ex = (2, 1, (1, 2)) if os == 16 else (1, 2, (2, 4))
        self.conv1 = nn.Conv2d(inplanes, 128, 3, stride=2, padding=1, bias=False)
        self.bn1 = nn.BatchNorm2d(128)
        self.relu = nn.LeakyReLU(0.2, inplace=True)
        self.conv2 = nn.Conv2d(128, 128, 3, padding=1, bias=False)
        self.bn2 = nn.BatchNorm2d(128)
        self.block1 = Blk(128, 128, 2, 2, start_with_relu=False)
        self.block2 = Blk(128, 256, 2, 2)
        self.block3 = Blk(256, 728, 2, b3s, is_last=True)
        for i in range(4, 20):
            setattr(self, "block%d" % i, Blk(728, 728, 3, 1, mid))
        self.block20 = Blk(728, 1024, 2, 1, ex[0], grow_first=False, is_last=True)
        self.conv3, self.bn3 = Sep(1024, 1536, 1, ex[1]), nn.BatchNorm2d(1536)
        self.conv4, self.bn4 = Sep(1536, 1536, 1, ex[1]), nn.BatchNorm2d(1536)
        self.conv5, self.bn5 = Sep(1536, 2048, 1, ex[1]), nn.BatchNorm2d(2048)
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight)

    def forward(self, x):
        x = self.relu(self.bn1(self.conv1(x)))
        x = self.relu(self.bn2(self.conv2(x)))
        x = self.block1(x)
        low = x
        x = self.block3(self.block2(x))
        for i in range(4, 21):
            x = getattr(self, "block%d" % i)(x)
        x = self.relu(self.bn3(self.conv3(x)))
        x = self.relu(self.bn4(self.conv4(x)))
        x = self.relu(self.bn5(self.conv5(x)))
        return x, low


def _decoder_init(mod):
    gain = nn.init.calculate_gain("leaky_relu", 0.2)
    for m in mod.modules():
        if isinstance(m, nn.Conv2d):
            nn.init.normal_(m.weight, 0.0, gain / math.sqrt(m.kernel_size[0] * m.kernel_size[1] * m.out_channels))


class Aspp(nn.Module):
    def __init__(self, cin, cout, rate):
        super().__init__()
        k, p = (1, 0) if rate == 1 else (3, rate)
        self.atrous_convolution = nn.Conv2d(cin, cout, k, 1, p, rate, bias=False)
        self.bn = nn.BatchNorm2d(cout)
        self.relu = nn.LeakyReLU(0.2)
        nn.init.kaiming_normal_(self.atrous_convolution.weight)

    def forward(self, x):
        return self.relu(self.bn(self.atrous_convolution(x)))


class _Upsample(nn.Module):
    def __init__(self, n_out):
        super().__init__()
        self.last_conv = nn.Sequential(nn.Conv2d(304, 256, 3, padding=1, bias=False), nn.BatchNorm2d(256),
                                       nn.LeakyReLU(0.2, inplace=True),
                                       nn.Conv2d(256, 256, 3, padding=1, bias=False), nn.BatchNorm2d(256),
                                       nn.LeakyReLU(0.2, inplace=True), nn.Conv2d(256, n_out, 1))
        _decoder_init(self)

    def forward(self, x, low, size):
        q = (int(math.ceil(size[-2] / 4)), int(math.ceil(size[-1] / 4)))
        x = F.interpolate(x, size=q, mode="bilinear", align_corners=True)
        x = self.last_conv(torch.cat((x, low), 1))
        return F.interpolate(x, size=size[2:], mode="bilinear", align_corners=True)


class _Deconv(nn.Module):
    """DeconvUpsampler, deeplab.py:398-444: four ConvTranspose2d k3 s2 whose paddings / output paddings are
    hard-wired for the 19x37 GPSRO fields (2x3 -> 3x6 -> 6x11 -> 10x19 -> 18x36), each followed by
    normaliser, AvgPool2d(2, stride 1) and LeakyReLU (the last one bare)."""

    def __init__(self, n_out, nn_pooling=True):
        super().__init__()
        pool = (lambda: nn.AvgPool2d(2, stride=1, padding=0)) if nn_pooling else (lambda: nn.Identity())
        act = lambda: nn.LeakyReLU(0.2, inplace=True)
        self.deconv1 = nn.Sequential(nn.ConvTranspose2d(256, 256, 3, 2, 1, output_padding=(0, 1), bias=False),
                                     nn.BatchNorm2d(256), pool(), act())
        self.deconv2 = nn.Sequential(nn.ConvTranspose2d(256, 256, 3, 2, 0, output_padding=(1, 0), bias=False),
                                     nn.BatchNorm2d(256), pool(), act())
        self.conv1 = nn.Sequential(nn.Conv2d(304, 256, 3, 1, 1, bias=False), nn.BatchNorm2d(256), act(),
                                   nn.Conv2d(256, 256, 3, 1, 1, bias=False), nn.BatchNorm2d(256), act(),
                                   nn.Conv2d(256, 256, 1))
        self.deconv3 = nn.Sequential(nn.ConvTranspose2d(256, 128, 3, 2, 1, output_padding=(1, 0), bias=False),
                                     nn.BatchNorm2d(128), pool(), act())
        self.last_deconv = nn.Sequential(nn.ConvTranspose2d(128, n_out, 3, 2, 1, output_padding=(1, 1), bias=False))
        gain = nn.init.calculate_gain("leaky_relu", 0.2)
        for m in self.modules():
            if isinstance(m, (nn.Conv2d, nn.ConvTranspose2d)):
                nn.init.normal_(m.weight, 0.0, gain / math.sqrt(m.kernel_size[0] * m.kernel_size[1] * m.out_channels))

    def forward(self, x, low, size):
        x = self.deconv2(self.deconv1(x))
        x = self.conv1(torch.cat((x, low), 1))
        return self.last_deconv(self.deconv3(x))


class _Extension(nn.Module):
    """UpsamplerExtension, deeplab.py:466-500: full-resolution skip from the raw input."""

    def __init__(self, n_in, n_out, nn_pooling=True):
        super().__init__()
        act = lambda: nn.LeakyReLU(0.2, inplace=True)
        self.init_norm = nn.Sequential(nn.BatchNorm2d(128),
                                       nn.AvgPool2d(2, stride=1, padding=1) if nn_pooling else nn.Identity(), act())
        self.conv1 = nn.Sequential(nn.Conv2d(n_in, 64, 3, 1, 1, bias=False), nn.BatchNorm2d(64), act(),
                                   nn.Conv2d(64, 128, 3, 1, 1, bias=False), nn.BatchNorm2d(128), act())
        self.conv2 = nn.Sequential(nn.Conv2d(256, 64, 3, 1, 1, bias=False), nn.BatchNorm2d(64), act(),
                                   nn.Conv2d(64, n_out, 3, 1, 1, bias=False))
        _decoder_init(self)

    def forward(self, inp, x):
        skip = self.conv1(inp)
        return self.conv2(torch.cat((self.init_norm(x), skip), 1))


class DeepLabRef(nn.Module):
    def __init__(self, n_input, n_output, os=16, upsampler_type="Interpolate"):
        super().__init__()
        self.upsampler_type = upsampler_type
        self.xception_features = XceptionRef(n_input, os)
        r = [1, 6, 12, 18] if os == 16 else [1, 12, 24, 36]
        self.aspp1, self.aspp2, self.aspp3, self.aspp4 = (Aspp(2048, 256, k) for k in r)
        self.relu = nn.LeakyReLU(0.2)
        self.global_avg_pool = nn.Sequential(nn.AdaptiveAvgPool2d((1, 1)), nn.Conv2d(2048, 256, 1, bias=False),
                                             nn.BatchNorm2d(256), nn.LeakyReLU(0.2))
        self.conv1, self.bn1 = nn.Conv2d(1280, 256, 1, bias=False), nn.BatchNorm2d(256)
        self.conv2, self.bn2 = nn.Conv2d(128, 48, 1, bias=False), nn.BatchNorm2d(48)
        if upsampler_type == "Interpolate":
            self.upsample = _Upsample(n_output)
        elif upsampler_type in ("Deconv", "Deconv1x"):
            self.upsample = _Deconv(128 if upsampler_type == "Deconv1x" else n_output)
            if upsampler_type == "Deconv1x":
                self.upsample_extension = _Extension(n_input, n_output)
            else:
                self.final_pool = nn.AvgPool2d(2, stride=1, padding=1)
        else:
            raise NotImplementedError("Error, upsampler {} not implemented.".format(upsampler_type))

    def forward(self, inp):
        x, low = self.xception_features(inp)
        parts = [self.aspp1(x), self.aspp2(x), self.aspp3(x), self.aspp4(x)]
        g = F.interpolate(self.global_avg_pool(x), size=x.shape[2:], mode="bilinear", align_corners=True)
        x = self.relu(self.bn1(self.conv1(torch.cat(parts + [g], 1))))
        low = self.relu(self.bn2(self.conv2(low)))
        x = self.upsample(x, low, inp.size())
        if self.upsampler_type == "Deconv1x":     # deeplab.py:676-679
            x = self.upsample_extension(inp, x)
        elif self.upsampler_type == "Deconv":
            x = self.final_pool(x)
        return x


class GeneratorRef(nn.Module):
    """deeplab_gan.py:65-94.  With noise_dimensions > 0 the noise planes are drawn on the CPU
    generator (`dist.rsample((B, nd, H, W))`), moved to the input's device and concatenated after
    the input channels (deeplab_gan.py:88-90)."""

    def __init__(self, n_input, n_output, os=16, noise_type="Normal", noise_dimensions=0, upsampler_type="Interpolate"):
        super().__init__()
        self.noise_dimensions = noise_dimensions
        if noise_type == "Uniform":
            self.dist = torch.distributions.uniform.Uniform(0.0, 1.0)
        elif noise_type == "Normal":
            self.dist = torch.distributions.normal.Normal(0.0, 1.0)
        else:
            raise NotImplementedError("Error, noise type {} not supported.".format(noise_type))
        self.model = DeepLabRef(n_input + noise_dimensions, n_output, os, upsampler_type)

    def forward(self, x):
        if self.noise_dimensions > 0:
            z = self.dist.rsample((x.shape[0], self.noise_dimensions, x.shape[2], x.shape[3])).to(x.device)
            x = torch.cat((x, z), dim=1)
        return self.model(x)


class DiscriminatorRef(nn.Module):
    def __init__(self, n_input, input_size=None, os=16):
        super().__init__()
        self.xception_features = XceptionRef(n_input, os)
        feat = 12288 if input_size is None else 2048 * math.ceil(input_size[0] / os) * math.ceil(input_size[1] / os)
        self.linear = nn.Linear(feat, 1)   # the reference hard-codes 12288 = 2048*2*3 (19x37 fields, deeplab_gan.py:21)
        self.sigmoid = nn.Sigmoid()
        _decoder_init(self)
        nn.init.xavier_uniform_(self.linear.weight, 1.0)
        nn.init.zeros_(self.linear.bias)

    def forward(self, x):
        f, _ = self.xception_features(x)
        logits = self.linear(f.reshape(f.shape[0], -1))
        return logits, self.sigmoid(logits)


class GANLossRef:
    """losses.py:129-172.  Label noise and the swap decision are drawn on the CPU generator in the
    reference's order: label_fake U(0,.2), label_real U(.8,1), swap U(0,1)."""

    def __init__(self, mode, batch_size, device):
        self.mode, self.device = mode, device
        if mode == "ModifiedMinMax":
            self.criterion = nn.BCEWithLogitsLoss()
            self.ones = torch.ones((batch_size, 1), device=device)
            self.shape = (batch_size, 1)
        elif mode != "Wasserstein":
            raise NotImplementedError(mode)

    def d_loss(self, logits_real, logits_fake):
        if self.mode == "Wasserstein":
            return torch.mean(logits_fake - logits_real)
        lf = torch.distributions.uniform.Uniform(0.0, 0.2).rsample(self.shape).to(self.device)
        lr = torch.distributions.uniform.Uniform(0.8, 1.0).rsample(self.shape).to(self.device)
        if torch.distributions.uniform.Uniform(0.0, 1.0).sample() < 0.05:
            lf, lr = lr, lf
        return 0.5 * (self.criterion(logits_fake, lf) + self.criterion(logits_real, lr))

    def g_loss(self, logits_fake):
        if self.mode == "Wasserstein":
            return -1.0 * torch.mean(logits_fake)
        return self.criterion(logits_fake, self.ones)


def gradient_penalty_ref(critic, images_fake, images_real):
    """deeplab_gan.py:98-114.  eta ~ U(0,1) per sample on the CPU generator; the interpolate is a fresh
    leaf (`Variable(..., requires_grad=True)`), the critic runs in its current (train) mode — so its
    BatchNorm running statistics advance once more — and the input gradient is taken WITHOUT
    create_graph: the penalty is a number added to the reported d_loss, it contributes no parameter
    gradient (the reference has the create_graph arguments commented out)."""
    eta = torch.distributions.uniform.Uniform(0.0, 1.0).rsample((images_fake.shape[0], 1, 1, 1)).to(images_fake.device)
    interpol = (eta * images_fake + (1.0 - eta) * images_real).detach().requires_grad_(True)
    logits, _ = critic(interpol)
    gradients = torch.autograd.grad(outputs=logits, inputs=interpol, grad_outputs=torch.ones_like(logits))[0]
    return ((gradients.norm(2, dim=1) - 1.0) ** 2).mean()


def gan_step_ref(G, D, g_opt, d_opt, crit, inputs, real, w_gan=1.0, w_reg=1.0, w_gp=10.0):
    """One step with both phases active (train_gan.py:250-298, generator_warmup_steps = 0, update
    frequencies 1, amp O0).  Returns (d_loss, g_loss, outputs_fake of the G phase)."""
    fake = G(inputs)  # not detached in the reference
    lr_, _ = D(real)
    lf_, _ = D(fake)
    d_loss = crit.d_loss(lr_, lf_)
    if crit.mode == "Wasserstein":  # train_gan.py:257-258
        d_loss = d_loss + w_gp * gradient_penalty_ref(D, fake, real)
    d_opt.zero_grad()
    d_loss.backward()
    d_opt.step()
    fake = G(inputs)
    lf_, _ = D(fake)
    g_loss = w_gan * crit.g_loss(lf_) + w_reg * F.l1_loss(fake, real)
    g_opt.zero_grad()
    g_loss.backward()
    g_opt.step()
    return d_loss.detach(), g_loss.detach(), fake.detach()
