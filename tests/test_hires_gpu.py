"""BASELINE.json configs[4]: 16 x 1536 x 2304 fields, batch 4 per GPU (`bench.py --config hires`).

At 4x the pixels the trunk runs 96x144 mid-flow maps, 384x576 quarter-resolution decoder convs and a
768x1152x128 entry layer: (1) the kernels at exactly those shapes against plain PyTorch fp32 on
bf16-rounded inputs (the usual 4e-3 / 1e-4 bars); (2) the whole generator at 1536x2304.  The fp32 oracle
does not fit at this size next to the product (batch 2 -- BatchNorm's minimum -- needs about 170 GB), so
the whole-network check is against the oracle under bf16 autocast: two independent bf16 evaluations of the
same function, each about 0.1 rel-L2 from fp32 (tests/test_fullsize_gpu.py), must agree within 0.2, their L1
terms within 2 %, and the run must repeat bit for bit.
"""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def rel(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float((a - b).norm() / (b.norm() + 1e-30))


def cos(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float((a @ b) / (a.norm() * b.norm() + 1e-30))


def nhwc(x):
    return x.permute(0, 2, 3, 1).contiguous()


def nchw(x):
    return x.permute(0, 3, 1, 2).contiguous()


@pytest.fixture(scope="module")
def dev():
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    return torch.device("cuda", 0)


@pytest.mark.parametrize("N,H,W,Cin,Cout,ks", [(1, 96, 144, 728, 728, 1),      # mid-flow pointwise at hi-res
                                               (1, 384, 576, 304, 256, 3),     # decoder 3x3 on the concat
                                               (1, 768, 1152, 128, 128, 3),    # entry conv2
                                               (1, 96, 144, 2048, 256, 1)])    # ASPP 1x1
def test_conv_at_hires_shapes(dev, N, H, W, Cin, Cout, ks):
    from bias_gan_b200 import functional as F_
    g = torch.Generator(device="cpu").manual_seed(1)
    x = torch.randn(N, Cin, H, W, generator=g).to(dev).bfloat16()
    w = (torch.randn(Cout, Cin, ks, ks, generator=g) / (Cin * ks * ks) ** 0.5).to(dev).bfloat16().float()
    gy = torch.randn(N, Cout, H, W, generator=g).to(dev).bfloat16()
    xr, wr = x.float().requires_grad_(True), w.clone().requires_grad_(True)
    ref = F.conv2d(xr, wr, None, padding=(ks - 1) // 2)
    ref.backward(gy.float())
    xp, wp = nhwc(x).requires_grad_(True), w.clone().requires_grad_(True)
    out = F_.conv2d(xp, wp, None)
    out.backward(nhwc(gy))
    assert rel(nchw(out), ref) < 4e-3
    assert rel(nchw(xp.grad), xr.grad) < 4e-3
    assert rel(wp.grad, wr.grad) < 1e-4


@pytest.mark.parametrize("C,H,W,stride,dil", [(728, 96, 144, 1, 1), (128, 768, 1152, 2, 1), (1536, 96, 144, 1, 2),
                                              (256, 384, 576, 2, 1)])
def test_depthwise_and_bn_at_hires_shapes(dev, C, H, W, stride, dil):
    from bias_gan_b200 import functional as F_
    g = torch.Generator(device="cpu").manual_seed(5)
    x = torch.randn(1, C, H, W, generator=g).to(dev).bfloat16()
    w = (torch.randn(C, 1, 3, 3, generator=g) / 3).to(dev)
    xr, wr = x.float().requires_grad_(True), w.clone().requires_grad_(True)
    ref = F.conv2d(F.pad(xr, (dil,) * 4), wr, None, stride=stride, dilation=dil, groups=C)
    gy = torch.randn(ref.shape, generator=g).to(dev).bfloat16()
    ref.backward(gy.float())
    xp, wp = nhwc(x).requires_grad_(True), w.clone().requires_grad_(True)
    out = F_.depthwise3x3(xp, wp, stride, dil)
    out.backward(nhwc(gy))
    assert rel(nchw(out), ref) < 4e-3
    assert rel(nchw(xp.grad), xr.grad) < 4e-3
    assert rel(wp.grad, wr.grad) < 1e-4
    # BatchNorm + LeakyReLU forward / backward on the same tensor
    bn_r, bn_p = torch.nn.BatchNorm2d(C).to(dev).train(), torch.nn.BatchNorm2d(C).to(dev).train()
    yr = x.float().requires_grad_(True)
    o = F.leaky_relu(bn_r(yr), 0.2)
    go = torch.randn(o.shape, generator=g).to(dev).bfloat16()
    o.backward(go.float())
    yp = nhwc(x).requires_grad_(True)
    op = F_.bn_act(yp, bn_p, act=True)
    op.backward(nhwc(go))
    assert rel(nchw(op), o) < 4e-3
    assert rel(nchw(yp.grad), yr.grad) < 1e-2
    assert rel(bn_p.weight.grad, bn_r.weight.grad) < 5e-3 and rel(bn_p.bias.grad, bn_r.bias.grad) < 5e-3


def test_generator_and_step_at_hires(dev):
    if torch.cuda.get_device_properties(0).total_memory < 100 * 2 ** 30:
        pytest.skip("needs about 90 GB of device memory")
    from bias_gan_b200 import deeplab_gan as P, kernels as K
    from bias_gan_b200.losses import GANLoss, L1LossField
    from bias_gan_b200.step import gan_step
    from oracle import ref_model as R
    from oracle.seeded import seed_state, seeded_fields
    B, C, H, W = 2, 16, 1536, 2304
    x, y = seeded_fields(5, B, C, H, W)
    x, y = x.to(dev), y.to(dev)
    res = {}
    for name in ("autocast", "product", "product-again"):
        if name == "autocast":
            G = seed_state(R.GeneratorRef(C, C), 11).to(dev).train()
            with torch.autocast("cuda", dtype=torch.bfloat16):
                out = G(x).float()
        else:
            G = seed_state(P.Generator(C, C, "Interpolate", "Normal", 0), 11).to(dev).train()
            out = G(x)
        loss = F.l1_loss(out, y)
        loss.backward()
        dec = torch.cat([p.grad.flatten().double() for n, p in G.named_parameters() if n.startswith("model.upsample.")])
        res[name] = (out.detach(), float(loss.detach()), dec)
        del G, out, loss
        torch.cuda.empty_cache()
    o_a, l_a, d_a = res["autocast"]
    o_p, l_p, d_p = res["product"]
    print("PARITY hires G: product vs autocast oracle rel %.4f  L1 %.5f / %.5f  decoder-grad cos %.4f"
          % (rel(o_p, o_a), l_p, l_a, cos(d_p, d_a)))
    assert tuple(o_p.shape) == (B, C, H, W)
    assert rel(o_p, o_a) <= 0.2
    assert abs(l_p - l_a) <= 2e-2 * abs(l_a)
    assert cos(d_p, d_a) >= 0.95
    assert torch.equal(o_p, res["product-again"][0]) and torch.equal(d_p, res["product-again"][2])
    del res, o_a, o_p
    torch.cuda.empty_cache()
    # one full G+D step at this size: finite losses, parameters moved, peak memory recorded
    G = seed_state(P.Generator(C, C, "Interpolate", "Normal", 0), 11).to(dev).train()
    D = seed_state(P.Discriminator(C, input_size=(H, W)), 12).to(dev).train()
    opt = lambda m: torch.optim.AdamW(m.parameters(), lr=1e-4, eps=1e-8, weight_decay=1e-5)
    before = [p.detach().clone() for p in D.parameters()][:4]
    torch.manual_seed(333)
    dl, gl = gan_step(G, D, opt(G), opt(D), GANLoss("ModifiedMinMax", B, dev), L1LossField(), K.nchw_to_field(x), K.nchw_to_field(y))
    assert bool(torch.isfinite(dl)) and bool(torch.isfinite(gl))
    assert any(not torch.equal(a, b) for a, b in zip(before, list(D.parameters())[:4]))
    print("PARITY hires step: d_loss %.4f g_loss %.4f  peak %.1f GiB" % (float(dl), float(gl), torch.cuda.max_memory_allocated() / 2 ** 30))
