"""Host-side RNG bookkeeping of the graphed step (CPU only): the pre-drawn noise planes, noisy labels and
WGAN-GP eta consume the CPU generator exactly like the reference's in-line draws
(deeplab_gan.py:88, losses.py:146-153, deeplab_gan.py:100-101; order per SURVEY 8c)."""
import torch

from bias_gan_b200.deeplab_gan import Generator, _NoiseSlot
from bias_gan_b200.losses import GANLoss


def test_noise_slots_and_labels_follow_the_reference_order():
    B, nd, H, W = 2, 3, 8, 12
    G = Generator(16, 16, "Interpolate", "Uniform", nd)
    assert G.model.xception_features.conv1.weight.shape[1] == 16 + nd     # state-dict shape of the reference
    G.static_noise(True)
    G._noise_slots = [_NoiseSlot((B, nd, H, W), "cpu"), _NoiseSlot((B, nd, H, W), "cpu")]
    crit = GANLoss("ModifiedMinMax", B, torch.device("cpu"))
    torch.manual_seed(123)
    G.redraw_noise(0, 1)
    crit.draw_labels()
    G.redraw_noise(1, None)
    after = torch.get_rng_state()
    label_fake, label_real = (t.clone() for t in crit._static)
    torch.manual_seed(123)
    u = torch.distributions.uniform.Uniform
    z1 = u(0.0, 1.0).rsample((B, nd, H, W))
    lf = u(0.0, 0.2).rsample((B, 1))
    lr = u(0.8, 1.0).rsample((B, 1))
    if u(0.0, 1.0).sample() < 0.05:
        lf, lr = lr, lf
    z2 = u(0.0, 1.0).rsample((B, nd, H, W))
    assert torch.equal(torch.get_rng_state(), after)
    assert torch.equal(G._noise_slots[0].dev, z1) and torch.equal(G._noise_slots[1].dev, z2)
    assert torch.equal(label_fake, lf) and torch.equal(label_real, lr)
    G.static_noise(False)
    assert G._noise_slots is None


def test_wasserstein_eta_is_predrawn_like_gradient_penalty():
    B = 3
    crit = GANLoss("Wasserstein", B, torch.device("cpu"))
    assert crit.static_eta() is None
    torch.manual_seed(7)
    crit.draw_labels()
    after = torch.get_rng_state()
    torch.manual_seed(7)
    eta = torch.distributions.uniform.Uniform(0.0, 1.0).rsample((B, 1, 1, 1))
    assert torch.equal(torch.get_rng_state(), after)
    assert torch.equal(crit.static_eta(), eta)
    lr_, lf_ = torch.randn(B, 1), torch.randn(B, 1)
    assert torch.equal(crit.d_loss(lr_, lf_), torch.mean(lf_ - lr_))     # losses.py:161-162, 171-172
    assert torch.equal(crit.g_loss(lf_), -1.0 * torch.mean(lf_))
