"""TEST INFRASTRUCTURE — deterministic parameter and input generators shared by the golden-vector
script (applied to the *reference* modules) and the parity tests (applied to the oracle and to
the product modules), so that a 100 M-parameter state never has to be stored in a fixture.
"""
import zlib

import torch


def _gen(seed, name):
    return torch.Generator(device="cpu").manual_seed((seed * 1000003 + zlib.crc32(name.encode())) % (2 ** 31))


@torch.no_grad()
def seed_state(module, seed):
    """Overwrite every parameter/buffer of `module` with values that depend only on (seed, key,
    shape): conv/linear weights ~ N(0, 1/fan_in) * 1.4, BN gamma ~ 1 + 0.1 N, beta ~ 0.1 N,
    running stats at their defaults."""
    for name, t in module.state_dict().items():
        g = _gen(seed, name)
        if name.endswith("num_batches_tracked"):
            t.zero_()
        elif name.endswith("running_mean"):
            t.zero_()
        elif name.endswith("running_var"):
            t.fill_(1.0)
        elif t.dim() >= 2:
            fan_in = t[0].numel()
            t.copy_(torch.randn(t.shape, generator=g) * (1.4 / fan_in ** 0.5))
        elif name.endswith("weight"):
            t.copy_(1.0 + 0.1 * torch.randn(t.shape, generator=g))
        else:
            t.copy_(0.1 * torch.randn(t.shape, generator=g))
    return module


def seeded_fields(seed, B, C, H, W):
    """(inputs, real) fp32 NCHW, roughly standardised, already rounded to bf16-representable values
    so that the bf16 product path and the fp32 oracle see identical inputs."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    x = torch.randn(B, C, H, W, generator=g)
    y = x + 0.1 * torch.randn(B, C, H, W, generator=g)
    return x.bfloat16().float(), y.bfloat16().float()
