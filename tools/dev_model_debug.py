"""GPU debugging aid: per-stage relative error of the product trunk vs the fp32 oracle (and the
oracle under bf16 autocast for scale)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
from bias_gan_b200 import deeplab_gan as P, functional as F_
from oracle import ref_model as R
from oracle.seeded import seed_state, seeded_fields

torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
B, C, H, W = 2, 16, int(os.environ.get("H", 64)), int(os.environ.get("W", 96))
dev = torch.device("cuda")
Gp = seed_state(P.Generator(C, C, "Interpolate", "Normal", 0), 11).to(dev).train()
Gr = seed_state(R.GeneratorRef(C, C), 11).to(dev).train()
x, y = seeded_fields(5, B, C, H, W)
x = x.to(dev)


def rel(a, b):
    return float((a.double() - b.double()).norm() / (b.double().norm() + 1e-30))


ref_out, prod_out = {}, {}
names = ["bn1", "bn2"] + ["block%d" % i for i in range(1, 21)] + ["bn3", "bn4", "bn5"]
xr, xp = Gr.model.xception_features, Gp.model.xception_features
for n in names:
    getattr(xr, n).register_forward_hook(lambda m, i, o, n=n: ref_out.__setitem__(n, o.detach().clone()))
for n in names:
    if n.startswith("block"):
        getattr(xp, n).register_forward_hook(lambda m, i, o, n=n: prod_out.__setitem__(n, (o[0] if isinstance(o, tuple) else o).detach().permute(0, 3, 1, 2).float()))
for tag, mods in (("aspp1", (Gr.model.aspp1, Gp.model.aspp1)), ("aspp2", (Gr.model.aspp2, Gp.model.aspp2)),
                  ("aspp4", (Gr.model.aspp4, Gp.model.aspp4))):
    mods[0].register_forward_hook(lambda m, i, o, n=tag: ref_out.__setitem__(n, o.detach().clone()))
    mods[1].register_forward_hook(lambda m, i, o, n=tag: prod_out.__setitem__(n, (o[0] if isinstance(o, tuple) else o).detach().permute(0, 3, 1, 2).float()))

out_r = Gr(x)
with torch.autocast("cuda", dtype=torch.bfloat16):
    Gr2 = seed_state(R.GeneratorRef(C, C), 11).to(dev).train()
    out_a = Gr2(x).float()
out_p = Gp(x)
print("G out: product vs fp32 %.4f | autocast-bf16 vs fp32 %.4f" % (rel(out_p, out_r), rel(out_a, out_r)))
for n in names:
    if n in prod_out:
        r = ref_out[n]
        if n != "block20":
            r = F.leaky_relu(r, 0.2)
        print("%-8s prod vs ref %.4f  shape %s" % (n, rel(prod_out[n], r), tuple(r.shape)))
for n in ("aspp1", "aspp2", "aspp4"):
    print("%-8s prod vs ref %.4f" % (n, rel(prod_out[n], ref_out[n])))
