"""Generates tests/golden/noise_golden.pt: the *reference's own* Generator with noise channels
(/root/reference/src/deepCam/architecture/gpsro/deeplab_gan.py:65-94, noise_dimensions 1 and 2,
"Normal" and "Uniform") on seeded weights / inputs (oracle/seeded.py) with the CPU generator seeded
right before the forward.  Only runnable in the build container.  Stores the output (bf16), an L1
gradient summary of the entry conv weight (the only tensor whose shape changes with the noise
planes) and of one deep weight.

    python tests/golden/make_noise_golden.py
"""
import os
import sys
import types

import torch
import torch.nn as nn
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle.seeded import seed_state, seeded_fields  # noqa: E402

REF = "/root/reference/src/deepCam"
B, C, H, W = 2, 16, 64, 96
CASES = [("Normal", 1), ("Uniform", 2)]
NOISE_SEED = 4242
KEYS = ("model.xception_features.conv1.weight", "model.xception_features.block4.rep.1.pointwise.weight")


def main():
    sys.modules.setdefault("conv2d_local", types.ModuleType("conv2d_local"))  # dead import, SURVEY F4
    sys.path.insert(0, REF)
    from architecture.gpsro import deeplab_gan as dxg
    res = {"meta": dict(B=B, C=C, H=H, W=W, noise_seed=NOISE_SEED, cases=CASES)}
    x, y = seeded_fields(5, B, C, H, W)
    for kind, nd in CASES:
        G = seed_state(dxg.Generator(C, C, "Interpolate", kind, nd, os=16, normalizer=nn.BatchNorm2d), 11).train()
        torch.manual_seed(NOISE_SEED)
        fake = G(x)
        F.l1_loss(fake, y).backward()
        params = dict(G.named_parameters())
        res["%s%d" % (kind, nd)] = dict(
            fake=fake.detach().bfloat16(),
            grads={k: (float(params[k].grad.double().norm()), params[k].grad.flatten()[:8].clone()) for k in KEYS})
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "noise_golden.pt")
    torch.save(res, out)
    print("wrote", out, os.path.getsize(out), "bytes")


if __name__ == "__main__":
    main()
