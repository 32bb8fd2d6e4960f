"""Pins oracle/ref_model.py (the fp32 PyTorch restatement used as the model-parity checker):
 * against tests/golden/model_golden.pt, generated from the reference's own modules
   (tests/golden/make_model_golden.py), everywhere;
 * state-dict key/shape identity with the product modules (so weights interchange).
CPU only; a few seconds.
"""
import os

import pytest
import torch
import torch.nn.functional as F

from oracle import ref_model as R
from oracle.seeded import seed_state, seeded_fields

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "model_golden.pt")


@pytest.fixture(scope="module")
def golden():
    return torch.load(GOLD, weights_only=False)


@pytest.fixture(scope="module")
def oracle_run(golden):
    m = golden["meta"]
    B, C, H, W = m["B"], m["C"], m["H"], m["W"]
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    G = seed_state(R.GeneratorRef(C, C), 11).train()
    D = seed_state(R.DiscriminatorRef(C, (H, W)), 12).train()
    x, y = seeded_fields(5, B, C, H, W)
    torch.manual_seed(333)
    crit = R.GANLossRef("ModifiedMinMax", B, torch.device("cpu"))
    out = {}
    fake = G(x)
    lr_, _ = D(y.clone())
    lf_, _ = D(fake)
    d_loss = crit.d_loss(lr_, lf_)
    d_loss.backward()
    out.update(fake_d=fake.detach(), logits_real=lr_.detach(), logits_fake=lf_.detach(), d_loss=d_loss.detach(),
               d_grads={n: p.grad.clone() for n, p in D.named_parameters()})
    G.zero_grad(); D.zero_grad()
    fake = G(x)
    lf_, _ = D(fake)
    g_loss = crit.g_loss(lf_) + F.l1_loss(fake, y)
    g_loss.backward()
    out.update(fake_g=fake.detach(), g_loss=g_loss.detach(), logits_fake_g=lf_.detach(),
               g_grads={n: p.grad.clone() for n, p in G.named_parameters()},
               bn1_running_mean=G.model.xception_features.bn1.running_mean.clone(),
               bn1_running_var=G.model.xception_features.bn1.running_var.clone())
    return out


def test_forward_matches_reference_golden(golden, oracle_run):
    for k in ("logits_real", "logits_fake", "logits_fake_g", "d_loss", "g_loss", "bn1_running_mean",
              "bn1_running_var"):
        torch.testing.assert_close(oracle_run[k], golden[k], rtol=1e-4, atol=1e-5)
    for k in ("fake_d", "fake_g"):  # stored at bf16 precision
        torch.testing.assert_close(oracle_run[k].bfloat16().float(), golden[k].float(), rtol=2e-2, atol=2e-2)


@pytest.mark.parametrize("which", ["d_grads", "g_grads"])
def test_gradients_match_reference_golden(golden, oracle_run, which):
    gold, got = golden[which], oracle_run[which]
    assert set(gold) == set(got)
    for n, (norm, head) in gold.items():
        g = got[n]
        assert abs(float(g.double().norm()) - norm) <= 2e-3 * norm + 1e-7, n
        torch.testing.assert_close(g.flatten()[:8], head, rtol=5e-3, atol=1e-6 + 1e-3 * norm / max(1, g.numel()) ** 0.5)


def test_state_dict_keys_match_product():
    """Product modules (bias_gan_b200) and oracle share key names and shapes."""
    from bias_gan_b200 import deeplab_gan as P
    C, H, W = 16, 64, 96
    for prod, ref in ((P.Generator(C, C, "Interpolate", "Normal", 0), R.GeneratorRef(C, C)),
                      (P.Discriminator(C, input_size=(H, W)), R.DiscriminatorRef(C, (H, W)))):
        a, b = prod.state_dict(), ref.state_dict()
        assert list(a.keys()) == list(b.keys())
        for k in a:
            assert a[k].shape == b[k].shape, k


NOISE_GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "noise_golden.pt")


@pytest.mark.parametrize("kind,nd", [("Normal", 1), ("Uniform", 2)])
def test_noise_generator_matches_reference_golden(kind, nd):
    """GeneratorRef with noise planes (deeplab_gan.py:85-94) against the reference's own module
    (tests/golden/make_noise_golden.py): same CPU-generator draw, same concat position."""
    gold = torch.load(NOISE_GOLD, weights_only=False)
    m = gold["meta"]
    x, y = seeded_fields(5, m["B"], m["C"], m["H"], m["W"])
    G = seed_state(R.GeneratorRef(m["C"], m["C"], noise_type=kind, noise_dimensions=nd), 11).train()
    torch.manual_seed(m["noise_seed"])
    fake = G(x)
    F.l1_loss(fake, y).backward()
    g = gold["%s%d" % (kind, nd)]
    torch.testing.assert_close(fake.detach().bfloat16().float(), g["fake"].float(), rtol=2e-2, atol=2e-2)
    params = dict(G.named_parameters())
    for k, (norm, head) in g["grads"].items():
        gr = params[k].grad
        assert abs(float(gr.double().norm()) - norm) <= 2e-3 * norm + 1e-7, k
        torch.testing.assert_close(gr.flatten()[:8], head, rtol=5e-3, atol=1e-6 + 1e-3 * norm / gr.numel() ** 0.5)


def test_gradient_penalty_matches_reference_golden():
    """gradient_penalty_ref / Wasserstein GANLossRef against the reference's own functions
    (tests/golden/make_gp_golden.py): value, RNG consumption, the extra BatchNorm update, and that
    the penalty carries no parameter gradient (create_graph is commented out in the reference)."""
    gold = torch.load(os.path.join(os.path.dirname(NOISE_GOLD), "gp_golden.pt"), weights_only=False)
    m = gold["meta"]
    D = seed_state(R.DiscriminatorRef(m["C"], (m["H"], m["W"])), 12).train()
    fake, real = seeded_fields(5, m["B"], m["C"], m["H"], m["W"])
    torch.manual_seed(m["seed"])
    crit = R.GANLossRef("Wasserstein", m["B"], torch.device("cpu"))
    lr_, _ = D(real)
    lf_, _ = D(fake)
    d_loss = crit.d_loss(lr_, lf_)
    gp = R.gradient_penalty_ref(D, fake, real)
    assert gp.requires_grad == gold["gp_requires_grad"] is False
    assert torch.equal(torch.get_rng_state(), gold["rng_after"])
    assert int(D.xception_features.bn1.num_batches_tracked) == gold["nbt"] == 3
    torch.testing.assert_close(d_loss.detach(), gold["d_loss"], rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(crit.g_loss(lf_).detach(), gold["g_loss"], rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(gp, gold["gp"], rtol=2e-3, atol=1e-6)
    (d_loss + 10.0 * gp).backward()
    g = D.xception_features.conv1.weight.grad
    norm, head = gold["grad"]
    assert abs(float(g.double().norm()) - norm) <= 2e-3 * norm
    torch.testing.assert_close(g.flatten()[:8], head, rtol=5e-3, atol=1e-3 * norm / g.numel() ** 0.5)


# ---- Deconv / Deconv1x upsamplers at the reference's native 45 x 19 x 37 configuration ------------------
DECONV_GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "deconv_golden.pt")


@pytest.mark.parametrize("up", ["Deconv", "Deconv1x"])
def test_deconv_oracle_matches_reference_golden(up):
    """oracle/ref_model.py's restatement of DeconvUpsampler / UpsamplerExtension (deeplab.py:398-500) against
    the vectors the reference's own, unmodified modules produced (tests/golden/make_deconv_golden.py):
    state-dict keys, generator output, critic logits (Linear(12288,1)), loss and every gradient norm."""
    gold = torch.load(DECONV_GOLD, weights_only=False)
    m = gold["meta"]
    B, C, H, W = m["B"], m["C"], m["H"], m["W"]
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    G = seed_state(R.GeneratorRef(C, C, upsampler_type=up), 11).train()
    D = seed_state(R.DiscriminatorRef(C), 12).train()
    assert sorted(G.state_dict()) == gold[up]["g_keys"] and sorted(D.state_dict()) == gold[up]["d_keys"]
    x, y = seeded_fields(7, B, C, H, W)
    fake = G(x)
    logits, _ = D(fake)
    loss = F.binary_cross_entropy_with_logits(logits, torch.ones_like(logits)) + F.l1_loss(fake, y)
    loss.backward()
    torch.testing.assert_close(fake.detach(), gold[up]["fake"], rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(logits.detach(), gold[up]["logits"], rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(loss.detach(), gold[up]["loss"], rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(G.model.upsample.deconv1[1].running_mean, gold[up]["bn_rm"], rtol=1e-4, atol=1e-6)
    for n, p in G.named_parameters():
        norm, head = gold[up]["g_grads"][n]
        assert abs(float(p.grad.double().norm()) - norm) <= 2e-3 * norm + 1e-9, n
        torch.testing.assert_close(p.grad.flatten()[:8], head, rtol=5e-3, atol=1e-6 + 1e-3 * norm / max(p.numel(), 1) ** 0.5)


@pytest.mark.parametrize("up", ["Deconv", "Deconv1x"])
def test_product_deconv_state_dict_is_the_references(up):
    """The product's Generator / Discriminator at the native configuration expose exactly the reference's
    state-dict keys and shapes (golden g_keys / d_keys come from the reference's own modules), so `.cpt` files
    interchange; construction needs no GPU."""
    from bias_gan_b200 import deeplab_gan as P
    gold = torch.load(DECONV_GOLD, weights_only=False)
    C = gold["meta"]["C"]
    G, D = P.Generator(C, C, up, "Normal", 0), P.Discriminator(C)
    assert sorted(G.state_dict()) == gold[up]["g_keys"] and sorted(D.state_dict()) == gold[up]["d_keys"]
    Gr = R.GeneratorRef(C, C, upsampler_type=up)
    for k, v in Gr.state_dict().items():
        assert tuple(G.state_dict()[k].shape) == tuple(v.shape), k
