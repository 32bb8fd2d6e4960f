"""GPU debugging aid: per-parameter gradient cosine of product / autocast-oracle vs the fp32 oracle."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
from bias_gan_b200 import deeplab_gan as P
from oracle import ref_model as R
from oracle.seeded import seed_state, seeded_fields

torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
B, C, H, W = int(os.environ.get("B", 2)), 16, int(os.environ.get("H", 64)), int(os.environ.get("W", 96))
dev = torch.device("cuda")
x, y = seeded_fields(5, B, C, H, W)
x, y = x.to(dev), y.to(dev)


def cos(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float((a @ b) / (a.norm() * b.norm() + 1e-30))


Gr = seed_state(R.GeneratorRef(C, C), 11).to(dev).train()
Ga = seed_state(R.GeneratorRef(C, C), 11).to(dev).train()
Gp = seed_state(P.Generator(C, C, "Interpolate", "Normal", 0), 11).to(dev).train()
F.l1_loss(Gr(x), y).backward()
with torch.autocast("cuda", dtype=torch.bfloat16):
    oa = Ga(x).float()
F.l1_loss(oa, y).backward()
F.l1_loss(Gp(x), y).backward()
names = [n for n, _ in Gr.named_parameters()]
pr, pa, pp = dict(Gr.named_parameters()), dict(Ga.named_parameters()), dict(Gp.named_parameters())
sel = [n for n in names if n.endswith("weight") and ("pointwise" in n or "conv" in n or "skip" in n) and "bn" not in n]
for n in sel[::6] + sel[-8:]:
    print("%-60s prod %.3f  autocast %.3f  |g| %.2e" % (n, cos(pp[n].grad, pr[n].grad), cos(pa[n].grad, pr[n].grad), float(pr[n].grad.norm())))
