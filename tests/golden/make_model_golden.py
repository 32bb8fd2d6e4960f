"""Generates tests/golden/model_golden.pt by running the *reference's own* modules
(/root/reference/src/deepCam/architecture/gpsro/deeplab_gan.py, utils/losses.py) on seeded weights
and inputs (oracle/seeded.py).  Only runnable in the build container.  The fixture stores outputs,
losses and per-parameter gradient summaries — not weights.

    python tests/golden/make_model_golden.py
"""
import os
import sys
import types

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle.seeded import seed_state, seeded_fields  # noqa: E402

REF = "/root/reference/src/deepCam"
B, C, H, W = 2, 16, 64, 96


def load_reference():
    sys.modules.setdefault("conv2d_local", types.ModuleType("conv2d_local"))  # dead import, SURVEY F4
    sys.path.insert(0, REF)
    from architecture.gpsro import deeplab_gan as dxg
    from utils.losses import GANLoss
    import torch.nn as nn
    G = dxg.Generator(C, C, "Interpolate", "Normal", 0, os=16, normalizer=nn.BatchNorm2d)
    D = dxg.Discriminator(C, os=16, normalizer=nn.BatchNorm2d)
    D.linear = nn.Linear(2048 * -(-H // 16) * -(-W // 16), 1)
    return G, D, GANLoss


def grad_summary(model):
    out = {}
    for n, p in model.named_parameters():
        g = p.grad
        out[n] = None if g is None else (float(g.double().norm()), g.flatten()[:8].clone())
    return out


def run(G, D, GANLoss):
    import torch.nn as nn
    seed_state(G, 11)
    seed_state(D, 12)
    G.train(); D.train()
    x, y = seeded_fields(5, B, C, H, W)
    torch.manual_seed(333)
    crit = GANLoss("ModifiedMinMax", B, torch.device("cpu"))
    res = {}
    # D phase forward/backward (train_gan.py:250-266), fake not detached
    fake = G(x)
    lr_, _ = D(y.clone())
    lf_, _ = D(fake)
    d_loss = crit.d_loss(lr_, lf_)
    G.zero_grad(); D.zero_grad()
    d_loss.backward()
    res.update(fake_d=fake.detach().clone(), logits_real=lr_.detach().clone(), logits_fake=lf_.detach().clone(),
               d_loss=d_loss.detach().clone(), d_grads=grad_summary(D))
    # G phase forward/backward (train_gan.py:273-293) on the same weights
    G.zero_grad(); D.zero_grad()
    fake = G(x)
    lf_, _ = D(fake)
    g_loss = 1.0 * crit.g_loss(lf_) + 1.0 * nn.L1Loss()(fake, y)
    g_loss.backward()
    res.update(fake_g=fake.detach().clone(), logits_fake_g=lf_.detach().clone(), g_loss=g_loss.detach().clone(),
               g_grads=grad_summary(G), bn1_running_mean=G.model.xception_features.bn1.running_mean.clone(),
               bn1_running_var=G.model.xception_features.bn1.running_var.clone())
    return res


if __name__ == "__main__":
    torch.set_num_threads(8)
    G, D, GANLoss = load_reference()
    res = run(G, D, GANLoss)
    res["meta"] = dict(B=B, C=C, H=H, W=W, torch=torch.__version__)
    # keep the fixture small: fields are stored at bf16 precision
    for k in ("fake_d", "fake_g"):
        res[k] = res[k].bfloat16()
    out = os.path.join(ROOT, "tests", "golden", "model_golden.pt")
    torch.save(res, out)
    print("wrote", out, os.path.getsize(out), "bytes", "d_loss", float(res["d_loss"]), "g_loss", float(res["g_loss"]))
