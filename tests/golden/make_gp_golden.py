"""Generates tests/golden/gp_golden.pt: the *reference's own* `gradient_penalty`
(/root/reference/src/deepCam/architecture/gpsro/deeplab_gan.py:98-114) and Wasserstein GANLoss
(utils/losses.py:129-172) on the reference Discriminator with seeded weights / inputs.  Only
runnable in the build container.

    python tests/golden/make_gp_golden.py
"""
import os
import sys
import types

import torch
import torch.nn as nn

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle.seeded import seed_state, seeded_fields  # noqa: E402

REF = "/root/reference/src/deepCam"
B, C, H, W = 2, 16, 64, 96
SEED = 777


def main():
    sys.modules.setdefault("conv2d_local", types.ModuleType("conv2d_local"))  # dead import, SURVEY F4
    sys.path.insert(0, REF)
    from architecture.gpsro import deeplab_gan as dxg
    from utils.losses import GANLoss
    D = dxg.Discriminator(C, os=16, normalizer=nn.BatchNorm2d)
    D.linear = nn.Linear(2048 * -(-H // 16) * -(-W // 16), 1)
    seed_state(D, 12).train()
    fake, real = seeded_fields(5, B, C, H, W)
    torch.manual_seed(SEED)
    crit = GANLoss("Wasserstein", B, torch.device("cpu"))
    lr_, _ = D(real)
    lf_, _ = D(fake)
    d_loss = crit.d_loss(lr_, lf_)
    gp = dxg.gradient_penalty(D, fake, real)
    total = d_loss + 10.0 * gp
    D.zero_grad()
    total.backward()
    key = "xception_features.conv1.weight"
    g = dict(D.named_parameters())[key].grad
    res = dict(meta=dict(B=B, C=C, H=H, W=W, seed=SEED), d_loss=d_loss.detach(), gp=gp.detach(),
               g_loss=crit.g_loss(lf_).detach(), gp_requires_grad=bool(gp.requires_grad),
               nbt=int(D.xception_features.bn1.num_batches_tracked), rng_after=torch.get_rng_state(),
               grad=(float(g.double().norm()), g.flatten()[:8].clone()))
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "gp_golden.pt")
    torch.save(res, out)
    print("wrote", out, {k: v for k, v in res.items() if k in ("d_loss", "gp", "g_loss", "gp_requires_grad", "nbt")})


if __name__ == "__main__":
    main()
