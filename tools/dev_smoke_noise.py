"""Losses of one G+D step at the smoke configuration: fp32 oracle, the same oracle under bf16
autocast (the rounding-noise floor of any bf16 implementation) and the product, a few runs each."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bias_gan_b200 import functional as F_
from bias_gan_b200.deeplab_gan import Discriminator, Generator
from bias_gan_b200.losses import GANLoss, L1LossField
from bias_gan_b200.step import gan_step
from oracle import ref_model as R
from oracle.seeded import seed_state

dev = torch.device("cuda", 0)
B, C, H, W = [int(v) for v in (sys.argv[1:5] if len(sys.argv) > 4 else (2, 16, 64, 96))]
g = torch.Generator().manual_seed(0)
x, y = torch.randn(B, C, H, W, generator=g).to(dev), torch.randn(B, C, H, W, generator=g).to(dev)
x, y = x.bfloat16().float(), y.bfloat16().float()
opt = lambda m: torch.optim.AdamW(m.parameters(), lr=1e-4, eps=1e-8, weight_decay=1e-5)


def oracle(autocast):
    Gr = seed_state(R.GeneratorRef(C, C), 11).to(dev).train()
    Dr = seed_state(R.DiscriminatorRef(C, (H, W)), 12).to(dev).train()
    torch.manual_seed(333)
    with torch.autocast("cuda", dtype=torch.bfloat16, enabled=autocast):
        dl, gl, _ = R.gan_step_ref(Gr, Dr, opt(Gr), opt(Dr), R.GANLossRef("ModifiedMinMax", B, dev), x, y.clone())
    return float(dl), float(gl)


def product():
    Gp = seed_state(Generator(C, C, "Interpolate", "Normal", 0), 11).to(dev).train()
    Dp = seed_state(Discriminator(C, input_size=(H, W)), 12).to(dev).train()
    torch.manual_seed(333)
    dl, gl = gan_step(Gp, Dp, opt(Gp), opt(Dp), GANLoss("ModifiedMinMax", B, dev), L1LossField(), F_.to_field(x), F_.to_field(y))
    return float(dl), float(gl)


print("config", (B, C, H, W))
print("fp32 oracle      ", oracle(False))
for _ in range(3):
    print("autocast oracle  ", oracle(True))
for _ in range(4):
    print("product          ", product())

# well-conditioned pieces: G's output and the L1 term, D's logits on real data (train-mode forward, no update)
import torch.nn.functional as TF
Gr = seed_state(R.GeneratorRef(C, C), 11).to(dev).train()
Dr = seed_state(R.DiscriminatorRef(C, (H, W)), 12).to(dev).train()
with torch.no_grad():
    fr = Gr(x); lr_ = Dr(y)[0]
    for _ in range(3):
        Gp = seed_state(Generator(C, C, "Interpolate", "Normal", 0), 11).to(dev).train()
        Dp = seed_state(Discriminator(C, input_size=(H, W)), 12).to(dev).train()
        fp = Gp(x).float(); lp = Dp(y)[0].float()
        print("G out rel-L2 %.4f | L1 term product %.5f oracle %.5f | D logits product %s oracle %s" % (
            float((fp - fr).norm() / fr.norm()), float(TF.l1_loss(fp, y)), float(TF.l1_loss(fr, y)),
            [round(float(v), 3) for v in lp.flatten()], [round(float(v), 3) for v in lr_.flatten()]))
