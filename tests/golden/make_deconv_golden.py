"""Generates tests/golden/deconv_golden.pt by running the *reference's own* modules
(/root/reference/src/deepCam/architecture/gpsro/deeplab_gan.py) UNMODIFIED at the configuration the reference's
only GAN run script uses (run_scripts/run_training_gan_coccobello.sh: 45-channel 19x37 GPSRO fields, local batch
4, `Deconv1x`), and with `Deconv`: generator output, discriminator logits (the hard-wired Linear(12288, 1)), the
L1 + adversarial generator loss and per-parameter gradient summaries, on seeded weights and inputs
(oracle/seeded.py).  Only runnable in the build container.

    python tests/golden/make_deconv_golden.py
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
B, C, H, W = 4, 45, 19, 37


def grad_summary(model):
    return {n: (float(p.grad.double().norm()), p.grad.flatten()[:8].clone()) for n, p in model.named_parameters()
            if p.grad is not None}


def main():
    sys.modules.setdefault("conv2d_local", types.ModuleType("conv2d_local"))
    sys.path.insert(0, REF)
    from architecture.gpsro import deeplab_gan as dxg
    out = {"meta": dict(B=B, C=C, H=H, W=W)}
    x, y = seeded_fields(7, B, C, H, W)
    for up in ("Deconv", "Deconv1x"):
        G = seed_state(dxg.Generator(C, C, up, "Normal", 0, os=16, normalizer=nn.BatchNorm2d), 11).train()
        D = seed_state(dxg.Discriminator(C, os=16, normalizer=nn.BatchNorm2d), 12).train()
        fake = G(x)
        logits, _ = D(fake)
        loss = F.binary_cross_entropy_with_logits(logits, torch.ones_like(logits)) + F.l1_loss(fake, y)
        loss.backward()
        out[up] = dict(fake=fake.detach().clone(), logits=logits.detach().clone(), loss=loss.detach().clone(),
                       g_grads=grad_summary(G), d_keys=sorted(D.state_dict()), g_keys=sorted(G.state_dict()),
                       bn_rm=G.model.upsample.deconv1[1].running_mean.clone())
        print(up, tuple(fake.shape), float(loss), logits.flatten().tolist())
    torch.save(out, os.path.join(ROOT, "tests", "golden", "deconv_golden.pt"))


if __name__ == "__main__":
    main()
