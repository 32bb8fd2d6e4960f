"""Generator / Discriminator with the reference's constructor signatures and state-dict keys
(src/deepCam/architecture/gpsro/deeplab_gan.py:12-94), running on the sm_100a kernels.

`forward(x)` keeps the reference contract: fp32 NCHW in, fp32 NCHW (generator) or
(logits, sigmoid) (discriminator) out, differentiable with loss.backward().
`forward_field(field)` is the fused path used by the step loop: bf16 2x2 space-to-depth fields
[N, H/2, W/2, 4C] in and out, no layout round trips.
"""
import math

import torch
import torch.nn as nn

from . import functional as F_
from .deeplab import DeepLabv3_plus, Xception


class Discriminator(nn.Module):
    def __init__(self, n_input=3, os=16, pretrained=False, normalizer=nn.BatchNorm2d, input_size=None):
        """input_size=(H, W) sizes the linear head as 2048*ceil(H/16)*ceil(W/16); the reference
        hard-codes 12288 = 2048*2*3 for 19x37 fields (deeplab_gan.py:21)."""
        super().__init__()
        self.n_input = n_input
        self.xception_features = Xception(n_input, os, pretrained, normalizer)
        feat = 12288
        if input_size is not None:
            feat = 2048 * int(math.ceil(input_size[0] / os)) * int(math.ceil(input_size[1] / os))
        self.linear = nn.Linear(feat, 1)
        self.sigmoid = nn.Sigmoid()
        self._init_weight()

    def _init_weight(self):
        gain = nn.init.calculate_gain("leaky_relu", 0.2)
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                n = m.kernel_size[0] * m.kernel_size[1] * m.out_channels
                nn.init.normal_(m.weight, mean=0.0, std=gain / math.sqrt(n))
            elif isinstance(m, nn.Linear):
                nn.init.xavier_uniform_(m.weight, 1.0)
                nn.init.zeros_(m.bias)
            elif isinstance(m, nn.BatchNorm2d):
                nn.init.ones_(m.weight)
                nn.init.zeros_(m.bias)

    def forward_field(self, field):
        x, _ = self.xception_features(field)
        logits = F_.linear_head(x, self.linear.weight, self.linear.bias)
        return logits, torch.sigmoid(logits)

    def forward(self, input):
        N, C, H, W = input.shape
        if (H | W) & 1 or C % 8:   # the reference's native 45 x 19 x 37 fields: zero-padded field
            return self.forward_field(F_.to_field_padded(input, F_.pad8(C)))
        return self.forward_field(F_.to_field(input))

    def freeze_bn(self):
        for m in self.modules():
            if isinstance(m, nn.BatchNorm2d):
                m.eval()


class _NoiseSlot:
    def __init__(self, shape, device):
        self.host = torch.empty(shape, pin_memory=torch.device(device).type == "cuda")
        self.dev = torch.empty(shape, device=device)
        self.copied = None

    def draw(self, dist):
        if self.copied is not None:
            self.copied.synchronize()      # the previous copy out of the pinned buffer has finished
        self.host.copy_(dist.rsample(self.host.shape))
        self.dev.copy_(self.host, non_blocking=True)   # stream-ordered after the previous step's reads
        if self.dev.is_cuda:
            self.copied = torch.cuda.Event()
            self.copied.record()


class Generator(nn.Module):
    def __init__(self, n_input, n_output, upsampler_type, noise_type, noise_dimensions, os=16, pretrained=False,
                 normalizer=nn.BatchNorm2d):
        super().__init__()
        self.noise_dimensions = noise_dimensions
        self.n_output = n_output
        self.n_input = n_input
        # true (H, W) of the fields handed to forward_field when they carry a zero row / column (odd sizes, e.g. the
        # native 19 x 37): a padded field cannot tell its own size.  None = the field is exactly 2*H2 x 2*W2.
        self.field_hw = None
        self._noise_slots, self._noise_cursor = None, 0
        if noise_type == "Uniform":
            self.dist = torch.distributions.uniform.Uniform(0.0, 1.0)
        elif noise_type == "Normal":
            self.dist = torch.distributions.normal.Normal(0.0, 1.0)
        else:
            raise NotImplementedError("Error, noise type {} not supported.".format(noise_type))
        self.model = DeepLabv3_plus(n_input=n_input + noise_dimensions, n_output=n_output, os=os,
                                    upsampler_type=upsampler_type, pretrained=pretrained, normalizer=normalizer)

    def forward_field(self, field, hw=None):
        """deeplab_gan.py:85-94.  hw = the true (H, W) of a field that carries a zero row / column (odd sizes).
        The noise planes are drawn like the reference's — `dist.rsample`
        on the CPU generator, [B, nd, H, W] fp32 — so a seeded run sees the same noise; they join the
        field in one kernel, padded with zero channels to a multiple of 16 so that the 4*(C+nd)
        channels of a block pixel fill whole 128-byte lines for the entry conv's TMA boxes."""
        if hw is None:
            hw = self.field_hw
        nd = self.noise_dimensions
        if nd > 0:
            if hw is not None and hw != (2 * field.shape[1], 2 * field.shape[2]) or (field.shape[-1] // 4) != self.n_input:
                raise NotImplementedError("noise planes on zero-padded fields (odd sizes or channel counts that are "
                                          "not multiples of 8) are not built; the reference's GAN run uses "
                                          "--noise_dimensions 0 there (run_training_gan_coccobello.sh:36)")
            B, H2, W2, C4 = field.shape
            shape = (B, nd, 2 * H2, 2 * W2)
            if self._noise_slots is None:
                noise = self.dist.rsample(shape)
                if field.is_cuda:
                    noise = noise.pin_memory().to(field.device, non_blocking=True)
            else:   # static buffers (CUDA graphs): the k-th forward of a step reads slot k
                if self._noise_cursor == len(self._noise_slots):
                    self._noise_slots.append(_NoiseSlot(shape, field.device))
                    self._noise_slots[-1].draw(self.dist)
                noise = self._noise_slots[self._noise_cursor].dev
                self._noise_cursor += 1
            field = F_.append_noise(field, noise, (C4 // 4 + nd + 15) // 16 * 16)
        return self.model(field, hw)

    def static_noise(self, enable=True):
        """Keep the noise planes in persistent device buffers, one per forward of a step, so that a
        captured CUDA graph (step.GraphedGanStep) contains no host RNG and no host->device copy."""
        self._noise_slots = [] if enable else None
        self._noise_cursor = 0

    def rewind_noise(self):
        self._noise_cursor = 0

    def redraw_noise(self, first=0, last=None):
        """Draw fresh noise into slots [first, last) on the CPU generator, in slot order."""
        for slot in (self._noise_slots or [])[first:last]:
            slot.draw(self.dist)

    def forward(self, input_raw):
        N, C, H, W = input_raw.shape
        if (H | W) & 1 or C % 8 or self.n_output % 8:   # the reference's native 45 x 19 x 37 fields
            out = self.forward_field(F_.to_field_padded(input_raw, F_.pad8(C)), hw=(H, W))
            return F_.from_field_padded(out, self.n_output, H, W)
        return F_.from_field(self.forward_field(F_.to_field(input_raw)), self.n_output)


def gradient_penalty(critic, images_fake, images_real, eta=None):
    """WGAN-GP term of the critic loss (deeplab_gan.py:98-114, train_gan.py:257-258): eta ~ U(0,1) per
    sample drawn on the CPU generator, one more train-mode critic forward on the interpolate, the
    gradient of the logits with respect to it (no create_graph: like the reference's, the result
    carries no parameter gradient), mean over pixels of (||grad||_2 over channels - 1)^2.

    Takes what the caller has: fp32 NCHW tensors (the reference's call) or bf16 fields
    [N, H/2, W/2, 4C] (the step loop); `eta` = a [N,1,1,1] device tensor drawn earlier
    (GANLoss.draw_labels, for CUDA graphs)."""
    if eta is None:
        eta = torch.distributions.uniform.Uniform(0.0, 1.0).rsample((images_fake.shape[0], 1, 1, 1))
        eta = eta.to(images_fake.device)
    field = images_fake.dtype == torch.bfloat16
    interpol = eta * images_fake.detach().float() + (1.0 - eta) * images_real.detach().float()
    interpol = (interpol.bfloat16() if field else interpol).requires_grad_(True)
    params = [p for p in critic.parameters() if p.requires_grad]
    for p in params:   # only the data-gradient kernels run: no weight gradient is asked for
        p.requires_grad_(False)
    try:
        with torch.enable_grad():
            logits, _ = critic.forward_field(interpol) if field else critic(interpol)
            gradients = torch.autograd.grad(outputs=logits, inputs=interpol, grad_outputs=torch.ones_like(logits))[0]
    finally:
        for p in params:
            p.requires_grad_(True)
    if field:   # channels of one pixel = the C entries of one (py, px) sub-block
        N, H2, W2, C4 = gradients.shape
        norms = gradients.view(N, H2, W2, 4, C4 // 4).float().norm(2, dim=-1)
    else:
        norms = gradients.norm(2, dim=1)
    return ((norms - 1.0) ** 2).mean()
