"""Run-to-run repeatability of the product networks at the smoke configuration: two forwards and
backwards from identical state; prints the first module whose output differs and the largest
relative gradient differences (fp32 atomics may move the last bits, nothing more)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bias_gan_b200 import kernels as K
from bias_gan_b200.deeplab_gan import Discriminator, Generator
from oracle.seeded import seed_state, seeded_fields

dev = torch.device("cuda", 0)
B, C, H, W = [int(v) for v in (sys.argv[1:5] if len(sys.argv) > 4 else (2, 16, 64, 96))]
x, y = seeded_fields(3, B, C, H, W)
xf = x.to(dev)


def run(kind):
    torch.manual_seed(0)
    if kind == "G":
        m = seed_state(Generator(C, C, "Interpolate", "Normal", 0), 11).to(dev).train()
    else:
        m = seed_state(Discriminator(C, input_size=(H, W)), 12).to(dev).train()
    outs = {}

    def hook(name):
        def f(mod, inp, out):
            o = out[0] if isinstance(out, tuple) else out
            if torch.is_tensor(o):
                outs[name] = o.detach().float().clone()
        return f
    for n, mod in m.named_modules():
        if n:
            mod.register_forward_hook(hook(n))
    out = m(xf)
    o = out[0] if isinstance(out, tuple) else out
    o.float().square().mean().backward()
    torch.cuda.synchronize()
    grads = {n: p.grad.detach().float().clone() for n, p in m.named_parameters() if p.grad is not None}
    return outs, grads, o.detach().float().clone()


for kind in ("G", "D"):
    a_out, a_g, a_o = run(kind)
    b_out, b_g, b_o = run(kind)
    bad = [(n, float((a_out[n] - b_out[n]).norm() / (a_out[n].norm() + 1e-30))) for n in a_out if n in b_out]
    first = next(((n, d) for n, d in bad if d > 0), None)
    print(kind, "output rel diff %.3e" % float((a_o - b_o).norm() / (a_o.norm() + 1e-30)), "| first differing module:", first)
    worst = sorted(((float((a_g[n] - b_g[n]).norm() / (a_g[n].norm() + 1e-30)), n) for n in a_g), reverse=True)[:5]
    print("   worst gradient rel diffs:", [(n, "%.2e" % d) for d, n in worst])
    big = [(n, "%.2e" % d) for n, d in bad if d > 1e-3][:8]
    print("   modules with forward rel diff > 1e-3:", big)
