"""Run-to-run determinism of the product path.

No kernel of libbgr.so combines partial results with floating-point atomics (csrc/reduce.cu): BatchNorm
statistics / backward sums, conv and depthwise weight gradients, pooling, the linear head and the loss
write one partial row per thread block (or split) and the rows are added in index order.  Two runs from
the same state are therefore bit-identical -- which is what allows test_model_gpu.py to hold fixed bounds
instead of bands derived from a run-to-run spread.
"""
import copy

import pytest
import torch

pytestmark = pytest.mark.gpu


def _build(dev, B, C, H, W, nd=0):
    from bias_gan_b200.deeplab_gan import Discriminator, Generator
    from oracle.seeded import seed_state
    G = seed_state(Generator(C, C, "Interpolate", "Normal", nd), 21).to(dev).train()
    D = seed_state(Discriminator(C, input_size=(H, W)), 22).to(dev).train()
    return G, D


def _state(*mods):
    return [t.detach().clone() for m in mods for t in list(m.parameters()) + list(m.buffers())]


@pytest.mark.parametrize("lean", [True, False])
def test_gan_step_is_bit_identical_across_runs(lean):
    from bias_gan_b200 import kernels as K
    from bias_gan_b200.losses import GANLoss, L1LossField
    from bias_gan_b200.step import gan_step
    from oracle.seeded import seeded_fields
    dev = torch.device("cuda", 0)
    B, C, H, W = 4, 16, 128, 192
    x, y = seeded_fields(3, B, C, H, W)
    xf, yf = K.nchw_to_field(x.to(dev)), K.nchw_to_field(y.to(dev))
    opt = lambda m: torch.optim.AdamW(m.parameters(), lr=1e-4, eps=1e-8, weight_decay=1e-5)
    runs = []
    for _ in range(3):
        G, D = _build(dev, B, C, H, W)
        go, do = opt(G), opt(D)
        torch.manual_seed(5)
        losses = []
        for _ in range(2):   # two steps: the second starts from updated weights / Adam moments
            dl, gl = gan_step(G, D, go, do, GANLoss("ModifiedMinMax", B, dev), L1LossField(), xf, yf, lean=lean)
            losses += [dl.clone(), gl.clone()]
        torch.cuda.synchronize()
        runs.append((losses, _state(G, D)))
    for losses, state in runs[1:]:
        for a, b in zip(losses, runs[0][0]):
            assert torch.equal(a, b), (float(a), float(b))
        for a, b in zip(state, runs[0][1]):
            assert torch.equal(a, b)


def test_graphed_step_is_bit_identical_to_eager():
    """One CUDA-graph replay (side-stream weight gradients as parallel branches, cached weight copies)
    produces exactly the bits of the kernel-by-kernel step: scheduling cannot change any sum."""
    from bias_gan_b200 import kernels as K
    from bias_gan_b200.losses import GANLoss, L1LossField
    from bias_gan_b200.step import GraphedGanStep, gan_step
    from oracle.seeded import seeded_fields
    dev = torch.device("cuda", 0)
    B, C, H, W = 4, 16, 64, 96
    x, y = seeded_fields(3, B, C, H, W)
    xf, yf = K.nchw_to_field(x.to(dev)), K.nchw_to_field(y.to(dev))
    opt = lambda m: torch.optim.AdamW(m.parameters(), lr=1e-4, eps=1e-8, weight_decay=1e-5, fused=True, capturable=True)
    G0, D0 = _build(dev, B, C, H, W)
    go0, do0 = opt(G0), opt(D0)
    torch.manual_seed(5)
    ref = [gan_step(G0, D0, go0, do0, GANLoss("ModifiedMinMax", B, dev), L1LossField(), xf, yf) for _ in range(2)]
    G1, D1 = _build(dev, B, C, H, W)
    go1, do1 = opt(G1), opt(D1)
    torch.manual_seed(5)
    gs = GraphedGanStep(G1, D1, go1, do1, GANLoss("ModifiedMinMax", B, dev), L1LossField(), tuple(xf.shape), dev)
    got = []
    for _ in range(2):
        dl, gl = gs(xf, yf)
        got.append((dl.clone(), gl.clone()))
    torch.cuda.synchronize()
    for (a0, a1), (b0, b1) in zip(ref, got):
        assert torch.equal(a0, b0) and torch.equal(a1, b1), (float(a0), float(b0), float(a1), float(b1))
    for a, b in zip(_state(G0, D0), _state(G1, D1)):
        assert torch.equal(a, b)


@pytest.mark.parametrize("C,N,H,W", [(728, 2, 24, 36), (128, 2, 48, 40), (2048, 1, 6, 9), (48, 2, 20, 28)])
def test_reductions_repeat_bit_exactly(C, N, H, W):
    from bias_gan_b200 import kernels as K
    dev = torch.device("cuda", 0)
    g = torch.Generator(device="cpu").manual_seed(C)
    x = K.act_empty((N, H, W, C), dev)
    gy = K.act_empty((N, H, W, C), dev)
    x.copy_(torch.randn(N, H, W, C, generator=g).to(dev))
    gy.copy_(torch.randn(N, H, W, C, generator=g).to(dev))
    w9 = torch.randn(9, C, generator=g).to(dev)
    gam = torch.rand(C, generator=g).to(dev) + 0.5
    for fn in (lambda: K.bn_stats(x),
               lambda: K.dw_wgrad(x, gy, 1, 1),
               lambda: K.dw_wgrad(x, gy, 1, 2),
               lambda: K.gap_fwd(x),
               lambda: K.bn_backward_recompute(gy, gam, gam - 1.0, x, gam, gam, gam, act=True)[2],
               lambda: K.l1_loss_grad(x.contiguous(), gy.contiguous(), 1.0)[0]):
        a = fn().clone()
        for _ in range(3):
            assert torch.equal(fn(), a)
