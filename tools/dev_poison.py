"""GPU debug: run the same forward/backward twice with the caching allocator poisoned (NaN patterns)
in between; report the first functional op whose output differs => something reads uninitialised
memory or races."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bias_gan_b200 import functional as F_, kernels as K
from bias_gan_b200.deeplab_gan import Discriminator, Generator
from oracle.seeded import seed_state, seeded_fields

dev = torch.device("cuda", 0)
B, C, H, W = 2, 16, 64, 96
x, y = seeded_fields(3, B, C, H, W)
xf = K.nchw_to_field(x.to(dev))
G = seed_state(Generator(C, C, "Interpolate", "Normal", 0), 21).to(dev).train()
D = seed_state(Discriminator(C, input_size=(H, W)), 22).to(dev).train()

names = ["conv2d", "conv_s2d", "depthwise3x3", "bn_act", "subsample2", "bilinear", "global_avg_pool", "linear_head"]
orig = {n: getattr(F_, n) for n in names}
log = []


def wrap(n):
    def f(*a, **kw):
        out = orig[n](*a, **kw)
        t = out[0] if isinstance(out, tuple) else out
        idx = len(log)
        log.append([n, tuple(t.shape), t.detach().float().clone(), None])
        if isinstance(out, tuple):
            log[-1].append(out[1].detach().clone())
        if t.requires_grad:
            t.register_hook(lambda g, idx=idx: log[idx].__setitem__(3, g.detach().float().clone()))
        return out
    return f


for n in names:
    setattr(F_, n, wrap(n))


def poison():
    t = torch.full((6 * 1024 ** 3 // 2,), float("nan"), dtype=torch.bfloat16, device=dev)
    del t
    t = torch.full((2 * 1024 ** 3 // 4,), 1e30, dtype=torch.float32, device=dev)
    del t


def run():
    log.clear()
    for m in (G, D):
        m.zero_grad(set_to_none=True)
    fake = G.forward_field(xf)
    logits, _ = D.forward_field(fake)
    (logits.sum() + fake.float().abs().mean()).backward()
    torch.cuda.synchronize()
    return [list(e) for e in log], {n: p.grad.detach().clone() for m in (G, D) for n, p in m.named_parameters() if p.grad is not None}


a, ga = run()
poison()
b, gb = run()
rows = []
for i, (ea, eb) in enumerate(zip(a, b)):
    r_out = ((ea[2] - eb[2]).norm() / (ea[2].norm() + 1e-30)).item()
    r_g = -1.0 if ea[3] is None or eb[3] is None else ((ea[3] - eb[3]).norm() / (ea[3].norm() + 1e-30)).item()
    nan = bool(torch.isnan(eb[2]).any()) or (eb[3] is not None and bool(torch.isnan(eb[3]).any()))
    rows.append((i, ea[0], ea[1], r_out, r_g, nan))
print("forward order, first 12 with out rel diff > 1e-3:")
k = 0
for r in rows:
    if r[3] > 1e-3 and k < 12:
        print("  op %3d %-14s %-22s out rel %.3e grad rel %.3e nan %s" % r); k += 1
print("backward order (last ops first):")
for r in rows[::-1][:60]:
    print("  op %3d %-14s %-22s out rel %.3e grad rel %.3e nan %s" % r)
