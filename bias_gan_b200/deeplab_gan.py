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
        return self.forward_field(F_.to_field(input))

    def freeze_bn(self):
        for m in self.modules():
            if isinstance(m, nn.BatchNorm2d):
                m.eval()


class Generator(nn.Module):
    def __init__(self, n_input, n_output, upsampler_type, noise_type, noise_dimensions, os=16, pretrained=False,
                 normalizer=nn.BatchNorm2d):
        super().__init__()
        self.noise_dimensions = noise_dimensions
        self.n_output = n_output
        if noise_type == "Uniform":
            self.dist = torch.distributions.uniform.Uniform(0.0, 1.0)
        elif noise_type == "Normal":
            self.dist = torch.distributions.normal.Normal(0.0, 1.0)
        else:
            raise NotImplementedError("Error, noise type {} not supported.".format(noise_type))
        if noise_dimensions > 0:
            raise NotImplementedError("noise_dimensions > 0 needs C+nd channel fields (not built yet)")
        self.model = DeepLabv3_plus(n_input=n_input + noise_dimensions, n_output=n_output, os=os,
                                    upsampler_type=upsampler_type, pretrained=pretrained, normalizer=normalizer)

    def forward_field(self, field):
        return self.model(field)

    def forward(self, input_raw):
        return F_.from_field(self.model(F_.to_field(input_raw)), self.n_output)
