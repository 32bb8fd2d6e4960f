"""Generator with noise planes (deeplab_gan.py:85-94) on the sm_100a path.
 * the field concat kernel is bit-exact against cat((x, noise), 1) -> bf16 -> field layout;
 * the product generator, seeded like the oracle and fed the same CPU-generator draw, matches the fp32
   oracle (pinned by tests/golden/noise_golden.pt) relative to the bf16-autocast noise floor, the entry
   conv's 17-/18-channel weight gradient included;
 * the graphed step draws on the host in the reference's order: noise, labels, noise.
"""
import os

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

NOISE_GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "noise_golden.pt")


def rel(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float((a - b).norm() / (b.norm() + 1e-30))


def cos(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float((a @ b) / (a.norm() * b.norm() + 1e-30))


@pytest.mark.parametrize("N,C,nd,Cp,H,W", [(2, 16, 1, 32, 8, 12), (1, 16, 3, 32, 64, 96), (3, 8, 8, 16, 6, 10),
                                           (1, 16, 5, 24, 4, 4)])
def test_field_append_noise_bit_exact(N, C, nd, Cp, H, W):
    from bias_gan_b200 import kernels as K
    dev = torch.device("cuda", 0)
    g = torch.Generator().manual_seed(N * 100 + nd)
    x = torch.randn(N, C, H, W, generator=g).to(dev)
    z = torch.randn(N, nd, H, W, generator=g).to(dev)
    got = K.field_append_noise(K.nchw_to_field(x), z, Cp)
    want = K.nchw_to_field(torch.cat((x, z, torch.zeros(N, Cp - C - nd, H, W, device=dev)), 1).contiguous())
    assert got.shape == want.shape == (N, H // 2, W // 2, 4 * Cp)
    assert torch.equal(got.view(torch.int16), want.view(torch.int16))


def test_field_append_noise_rejects_bad_shapes():
    from bias_gan_b200 import kernels as K
    dev = torch.device("cuda", 0)
    f = torch.zeros(1, 2, 2, 64, dtype=torch.bfloat16, device=dev)
    with pytest.raises(ValueError):
        K.field_append_noise(f, torch.zeros(1, 9, 4, 4, device=dev), 24)   # Cp < C + nd


@pytest.mark.parametrize("kind,nd", [("Normal", 1), ("Uniform", 2)])
def test_noise_generator_matches_oracle(kind, nd):
    from bias_gan_b200 import deeplab_gan as P
    from oracle import ref_model as R
    from oracle.seeded import seed_state, seeded_fields
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    dev = torch.device("cuda", 0)
    B, C, H, W = 4, 16, 256, 384
    x, y = seeded_fields(5, B, C, H, W)
    x, y = x.to(dev), y.to(dev)
    Gp = seed_state(P.Generator(C, C, "Interpolate", kind, nd), 11).to(dev).train()
    Gr = seed_state(R.GeneratorRef(C, C, noise_type=kind, noise_dimensions=nd), 11).to(dev).train()
    Ga = seed_state(R.GeneratorRef(C, C, noise_type=kind, noise_dimensions=nd), 11).to(dev).train()
    assert list(Gp.state_dict().keys()) == list(Gr.state_dict().keys())
    assert Gp.model.xception_features.conv1.weight.shape[1] == C + nd
    outs = {}
    for name, G in (("p", Gp), ("r", Gr), ("a", Ga)):
        torch.manual_seed(4242)
        if name == "a":
            with torch.autocast("cuda", dtype=torch.bfloat16):
                out = G(x).float()
        else:
            out = G(x)
        F.l1_loss(out, y).backward()
        outs[name] = out.detach()
        rng_after = torch.get_rng_state()
        if name == "p":
            rng_p = rng_after
    assert torch.equal(rng_p, rng_after)    # the product consumed exactly the reference's draw
    assert rel(outs["p"], outs["r"]) <= 1.5 * rel(outs["a"], outs["r"]) + 1e-3
    key = "model.xception_features.conv1.weight"
    gp, gr, ga = (dict(G.named_parameters())[key].grad for G in (Gp, Gr, Ga))
    assert gp.shape == gr.shape
    assert (1 - cos(gp, gr)) <= 1.5 * (1 - cos(ga, gr)) + 1e-3
    # the noise channels' own weight slice must carry signal, not the zero padding's
    assert float(gp[:, C:].abs().max()) > 0
    assert (1 - cos(gp[:, C:], gr[:, C:])) <= 1.5 * (1 - cos(ga[:, C:], gr[:, C:])) + 2e-2


def test_noise_generator_matches_reference_golden():
    """Against the reference's own module output (64x96, batch 2; tests/golden/make_noise_golden.py)."""
    from bias_gan_b200 import deeplab_gan as P
    from oracle.seeded import seed_state, seeded_fields
    gold = torch.load(NOISE_GOLD, weights_only=False)
    m = gold["meta"]
    dev = torch.device("cuda", 0)
    x, _ = seeded_fields(5, m["B"], m["C"], m["H"], m["W"])
    for kind, nd in m["cases"]:
        G = seed_state(P.Generator(m["C"], m["C"], "Interpolate", kind, nd), 11).to(dev).train()
        torch.manual_seed(m["noise_seed"])
        with torch.no_grad():
            out = G(x.to(dev))
        assert rel(out.cpu(), gold["%s%d" % (kind, nd)]["fake"].float()) <= 0.2   # bf16 floor here ~0.13


def test_graphed_step_with_noise_draw_order():
    from bias_gan_b200 import kernels as K
    from bias_gan_b200.deeplab_gan import Discriminator, Generator
    from bias_gan_b200.losses import GANLoss, L1LossField
    from bias_gan_b200.step import GraphedGanStep, gan_step
    from oracle.seeded import seed_state, seeded_fields
    dev = torch.device("cuda", 0)
    B, C, H, W, nd = 2, 16, 64, 96, 1
    x, y = seeded_fields(3, B, C, H, W)
    xf, yf = K.nchw_to_field(x.to(dev)), K.nchw_to_field(y.to(dev))
    opt = lambda m: torch.optim.AdamW(m.parameters(), lr=1e-4, eps=1e-8, weight_decay=1e-5, fused=True, capturable=True)

    def build():
        G = seed_state(Generator(C, C, "Interpolate", "Normal", nd), 21).to(dev).train()
        D = seed_state(Discriminator(C, input_size=(H, W)), 22).to(dev).train()
        return G, D, opt(G), opt(D)

    G0, D0, go0, do0 = build()
    torch.manual_seed(9)
    dl0, gl0 = gan_step(G0, D0, go0, do0, GANLoss("ModifiedMinMax", B, dev), L1LossField(), xf, yf)
    rng_eager = torch.get_rng_state()
    # the reference's order on the CPU generator: noise, label_fake, label_real, swap, noise
    torch.manual_seed(9)
    z1 = torch.distributions.normal.Normal(0.0, 1.0).rsample((B, nd, H, W))
    torch.distributions.uniform.Uniform(0.0, 0.2).rsample((B, 1))
    torch.distributions.uniform.Uniform(0.8, 1.0).rsample((B, 1))
    torch.distributions.uniform.Uniform(0.0, 1.0).sample()
    z2 = torch.distributions.normal.Normal(0.0, 1.0).rsample((B, nd, H, W))
    assert torch.equal(torch.get_rng_state(), rng_eager)

    G1, D1, go1, do1 = build()
    torch.manual_seed(9)
    gs = GraphedGanStep(G1, D1, go1, do1, GANLoss("ModifiedMinMax", B, dev), L1LossField(), tuple(xf.shape), dev)
    torch.manual_seed(9)
    dl1, gl1 = gs(xf, yf)
    assert torch.equal(torch.get_rng_state(), rng_eager)
    assert len(G1._noise_slots) == 2
    assert torch.equal(G1._noise_slots[0].dev.cpu(), z1) and torch.equal(G1._noise_slots[1].dev.cpu(), z2)
    # same state, same labels, same noise planes and no atomics anywhere: the replayed graph reproduces the
    # kernel-by-kernel step bit for bit
    assert torch.equal(dl0, dl1) and torch.equal(gl0, gl1), (float(dl0), float(dl1), float(gl0), float(gl1))
    dl2, gl2 = gs(xf, yf)
    assert not torch.equal(G1._noise_slots[0].dev.cpu(), z1)    # fresh noise every replay
    assert all(torch.isfinite(t) for t in (dl2, gl2))
