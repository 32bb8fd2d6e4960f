"""The normaliser / mode flag surface of the training script on the sm_100a path, against the fp32 oracle:
  * eval mode (train_gan.py:330-398 validation: generator.eval() / discriminator.eval() under no_grad; and
    freeze_bn): BatchNorm uses its running statistics, folded into a per-channel affine; forward AND backward;
  * --disable_batchnorm (train_gan.py:125,480): an Identity normaliser in place of every BatchNorm2d.
Without batch statistics there is no train-mode noise amplification, so the bounds are absolute.
"""
import copy

import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

B, C, H, W = 2, 16, 128, 192


def rel(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float((a - b).norm() / (b.norm() + 1e-30))


def cos(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float((a @ b) / (a.norm() * b.norm() + 1e-30))


def _strip_norms(module):
    """Replace every BatchNorm2d of the oracle by nn.Identity (what normalizer=Identity builds in the reference)."""
    for parent in list(module.modules()):
        for name, child in list(parent.named_children()):
            if isinstance(child, nn.BatchNorm2d):
                setattr(parent, name, nn.Identity())
    return module


def _randomize_running_stats(*mods):
    """Non-trivial running statistics, identical in product and oracle (same module names)."""
    for m in mods:
        for name, buf in m.named_buffers():
            g = torch.Generator(device="cpu").manual_seed(abs(hash(name)) % (2 ** 31))
            if name.endswith("running_mean"):
                buf.copy_(0.1 * torch.randn(buf.shape, generator=g))
            elif name.endswith("running_var"):
                buf.copy_(0.5 + torch.rand(buf.shape, generator=g))


@pytest.fixture(scope="module")
def env():
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    from oracle.seeded import seeded_fields
    dev = torch.device("cuda", 0)
    x, y = seeded_fields(5, B, C, H, W)
    return dict(dev=dev, x=x.to(dev), y=y.to(dev))


def _flat(m, prefix=""):
    return torch.cat([p.grad.flatten().double() for n, p in m.named_parameters() if n.startswith(prefix) and p.grad is not None])


def test_eval_mode_forward_and_backward(env):
    from bias_gan_b200 import deeplab_gan as P
    from oracle import ref_model as R
    from oracle.seeded import seed_state
    dev, x, y = env["dev"], env["x"], env["y"]
    Gp = seed_state(P.Generator(C, C, "Interpolate", "Normal", 0), 11)
    Gr = seed_state(R.GeneratorRef(C, C), 11)
    Dp = seed_state(P.Discriminator(C, input_size=(H, W)), 12)
    Dr = seed_state(R.DiscriminatorRef(C, (H, W)), 12)
    # seed_state leaves running_mean 0 / running_var 1.  Give both sides the SAME realistic running statistics:
    # one train-mode pass of the oracle with momentum 1 stores the batch statistics of this input, which are
    # then copied into the product's buffers (arbitrary statistics would let activations explode through 70
    # un-normalised layers and make the comparison ill-conditioned).
    Gr, Dr = Gr.to(dev).train(), Dr.to(dev).train()
    bns = [m for net in (Gr, Dr) for m in net.modules() if isinstance(m, nn.BatchNorm2d)]
    for m in bns:
        m.momentum = 1.0
    with torch.no_grad():
        Gr(x)
        Dr(y.clone())
    for m in bns:
        m.momentum = 0.1
    for mp, mr in ((Gp, Gr), (Dp, Dr)):
        sp, sr = mp.state_dict(), mr.state_dict()
        for k in sp:
            if k.endswith("running_mean") or k.endswith("running_var"):
                sp[k].copy_(sr[k].cpu())
    Gp, Gr, Dp, Dr = (m.to(dev).eval() for m in (Gp, Gr, Dp, Dr))
    before = {k: v.clone() for k, v in Gp.state_dict().items() if "running" in k or "num_batches" in k}
    # validation forward (no_grad), as train_gan.py:345-360
    with torch.no_grad():
        fake_r, fake_p = Gr(x), Gp(x)
        lo_r, lo_p = Dr(y.clone())[0], Dp(y)[0]
    for k, v in Gp.state_dict().items():
        if k in before:
            assert torch.equal(v, before[k]), "eval forward changed " + k
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):   # the bf16 noise floor of this function
        fake_a, lo_a = Gr(x).float(), Dr(y.clone())[0].float()
    print("PARITY eval: G out rel product %.4f autocast %.4f  D logits %s / %s / %s"
          % (rel(fake_p, fake_r), rel(fake_a, fake_r), lo_p.flatten().tolist(), lo_a.flatten().tolist(), lo_r.flatten().tolist()))
    # frozen statistics do not re-centre the activations layer by layer, so bf16 rounding accumulates through the
    # 70 layers further than in train mode: the bar is again the oracle's own bf16-autocast deviation
    assert rel(fake_p, fake_r) <= 1.25 * rel(fake_a, fake_r) + 1e-3
    assert float((lo_p - lo_r).abs().max()) <= 1.5 * float((lo_a - lo_r).abs().max()) + 2e-2 * (float(lo_r.abs().max()) + 1.0)
    # frozen-BN training (freeze_bn): gradients flow through the eval-mode normalisers to every parameter
    out_r, out_p = Gr(x), Gp(x)
    F.l1_loss(out_r, y).backward(); F.l1_loss(out_p, y).backward()
    Ga = copy.deepcopy(Gr)
    Ga.zero_grad()
    with torch.autocast("cuda", dtype=torch.bfloat16):
        out_a = Ga(x).float()
    F.l1_loss(out_a, y).backward()
    for pre, cap in (("model.upsample.", 0.99), ("model.xception_features.block10.", 0.9), ("model.xception_features.conv1.", 0.9)):
        c, ca = cos(_flat(Gp, pre), _flat(Gr, pre)), cos(_flat(Ga, pre), _flat(Gr, pre))
        print("PARITY eval: G grad cos %-34s product %.5f autocast %.5f" % (pre, c, ca))
        assert c >= min(cap, ca - 0.03), (pre, c, ca)
    sel = lambda m: torch.cat([p.grad.flatten().double() for n, p in m.named_parameters() if p.dim() == 1 and "xception_features.block1." in n])
    cg, cga = cos(sel(Gp), sel(Gr)), cos(sel(Ga), sel(Gr))
    print("PARITY eval: dL/dgamma,dbeta (block1) cos product %.5f autocast %.5f" % (cg, cga))
    assert cg >= min(0.95, cga - 0.03)      # dL/dgamma, dL/dbeta of the frozen normalisers


def test_identity_normaliser(env):
    """normalizer = Identity: same state-dict keys minus the BatchNorm entries, activation / residual fused
    into one kernel, no statistics anywhere."""
    from bias_gan_b200 import deeplab_gan as P
    from oracle import ref_model as R
    from oracle.seeded import seed_state
    dev, x, y = env["dev"], env["x"], env["y"]
    Gp = P.Generator(C, C, "Interpolate", "Normal", 0, normalizer=nn.Identity)
    Dp = P.Discriminator(C, normalizer=nn.Identity, input_size=(H, W))
    Gr, Dr = _strip_norms(R.GeneratorRef(C, C)), _strip_norms(R.DiscriminatorRef(C, (H, W)))
    assert set(Gp.state_dict()) == set(Gr.state_dict()) and set(Dp.state_dict()) == set(Dr.state_dict())
    # without normalisation activations grow layer by layer: scale the weights down so that bf16 stays in range
    for m, seed in ((Gp, 11), (Gr, 11), (Dp, 12), (Dr, 12)):
        seed_state(m, seed)
        with torch.no_grad():
            for p in m.parameters():
                if p.dim() == 4:
                    p.mul_(0.7)
    Gp, Gr, Dp, Dr = (m.to(dev).train() for m in (Gp, Gr, Dp, Dr))
    out_r, out_p = Gr(x), Gp(x)
    lo_r, lo_p = Dr(y.clone())[0], Dp(y)[0]
    print("PARITY identity: G out rel %.4f (|out| %.3g)  D logits %s vs %s"
          % (rel(out_p, out_r), float(out_r.abs().mean()), lo_p.flatten().tolist(), lo_r.flatten().tolist()))
    assert rel(out_p, out_r) <= 3e-2
    assert float((lo_p - lo_r).abs().max()) <= 3e-2 * (float(lo_r.abs().max()) + 1.0)
    F.l1_loss(out_r, y).backward(); F.l1_loss(out_p, y).backward()
    F.binary_cross_entropy_with_logits(lo_r, torch.full_like(lo_r, 0.9)).backward()
    F.binary_cross_entropy_with_logits(lo_p, torch.full_like(lo_p, 0.9)).backward()
    for m_p, m_r, pre in ((Gp, Gr, "model.upsample."), (Gp, Gr, "model.xception_features.block10."),
                          (Dp, Dr, "xception_features.block10."), (Dp, Dr, "linear.")):
        c = cos(_flat(m_p, pre), _flat(m_r, pre))
        print("PARITY identity: grad cos %-34s %.5f" % (pre, c))
        assert c >= 0.98, (pre, c)
