"""layouts.WeightPack / bgr_weights_prepare: every derived weight copy of a network from one launch,
bit-exact against the per-parameter builders it replaces, and kept in step with the optimizer."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _lazy(w, kind):
    from bias_gan_b200 import layouts as L
    if kind == "gemm":
        return L.gemm_weight(w)
    if kind == "gemm_t":
        return L.gemm_weight_t(w)
    C = w.shape[0]
    return w.detach().reshape(C, 9).t().contiguous().float()


@pytest.mark.parametrize("which", ["G", "D"])
def test_pack_matches_per_parameter_builders(which):
    from bias_gan_b200 import layouts as L
    from bias_gan_b200.deeplab_gan import Discriminator, Generator
    from oracle.seeded import seed_state
    dev = torch.device("cuda", 0)
    m = Generator(16, 16, "Interpolate", "Normal", 1) if which == "G" else Discriminator(16, input_size=(64, 96))
    m = seed_state(m, 3).to(dev)
    pack = L.prepare_weights(m)
    kinds = [k for _, k, _, _ in pack.items]
    assert kinds.count("gemm") == kinds.count("gemm_t") > 40 and kinds.count("dw9") > 60
    # the stride-2 entry conv keeps its space-to-depth path
    entry = (m.model if which == "G" else m).xception_features.conv1.weight
    assert all(p is not entry for p, _, _, _ in pack.items)
    for p, kind, view, buf in pack.items:
        want = _lazy(p, kind)
        assert view.shape == want.shape and view.dtype == want.dtype, (kind, tuple(p.shape))
        assert torch.equal(view, want), (kind, tuple(p.shape))
        if buf.shape != view.shape:   # padding columns are zeros (TMA reads whole 128-byte lines)
            assert float(buf[..., view.shape[-1]:].float().abs().max()) == 0.0
        assert L.cached(p, kind, lambda w: None) is view        # functional.py will pick these up
    # parameters change -> one more launch brings every copy up to date
    with torch.no_grad():
        for p in m.parameters():
            p.mul_(1.5)
    assert L.prepare_weights(m) is pack
    for p, kind, view, _ in pack.items[:: max(1, len(pack.items) // 25)]:
        assert torch.equal(view, _lazy(p, kind))


def test_pack_follows_the_optimizer_inside_a_step():
    """gan_step rebuilds D's copies between the critic update and the generator phase, and G's at the
    start of the next step: after a step D's copies equal its (updated) parameters bit for bit, G's are
    the ones its forward used (now stale) until the next prepare.  (Comparing losses or weights of a
    packed and an unpacked run says nothing at this size: AdamW turns the trunk's sign-noisy bf16
    gradients into +-lr moves, measured median |dw| between two identical runs 1.6e-4 = 1.6 lr.)"""
    from bias_gan_b200 import kernels as K, layouts as L, step as S
    from bias_gan_b200.deeplab_gan import Discriminator, Generator
    from bias_gan_b200.losses import GANLoss, L1LossField
    from oracle.seeded import seed_state, seeded_fields
    dev = torch.device("cuda", 0)
    B, C, H, W = 2, 16, 64, 96
    x, y = seeded_fields(3, B, C, H, W)
    xf, yf = K.nchw_to_field(x.to(dev)), K.nchw_to_field(y.to(dev))
    assert S.PACK_WEIGHTS
    G = seed_state(Generator(C, C, "Interpolate", "Normal", 0), 21).to(dev).train()
    D = seed_state(Discriminator(C, input_size=(H, W)), 22).to(dev).train()
    go, do = (torch.optim.AdamW(m.parameters(), lr=1e-4, eps=1e-8, weight_decay=1e-5) for m in (G, D))
    torch.manual_seed(5)
    crit = GANLoss("ModifiedMinMax", B, dev)
    for it in range(2):
        dl, gl = S.gan_step(G, D, go, do, crit, L1LossField(), xf, yf)
        assert torch.isfinite(dl) and torch.isfinite(gl)
        pd, pg = D.__dict__["_bgr_weight_pack"], G.__dict__["_bgr_weight_pack"]
        for p, kind, view, _ in pd.items[::7]:
            assert torch.equal(view, _lazy(p, kind)), ("D copy stale after the step", kind, tuple(p.shape))
        stale = sum(not torch.equal(view, _lazy(p, kind)) for p, kind, view, _ in pg.items[::7])
        assert stale > 0.9 * len(pg.items[::7])      # G was updated after its copies were last built
    L.prepare_weights(G)
    for p, kind, view, _ in pg.items[::7]:
        assert torch.equal(view, _lazy(p, kind))
