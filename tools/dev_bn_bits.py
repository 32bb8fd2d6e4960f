"""Bit-equality harness for the BatchNorm streaming kernels: runs bn_apply / bn_backward (stored, recompute, dual,
frozen) on seeded inputs at several shapes and saves every output; run once per library build (BGR_LIB=...) and
compare the two files with --compare.
    python tools/dev_bn_bits.py out_a.pt ; BGR_LIB=other.so python tools/dev_bn_bits.py out_b.pt
    python tools/dev_bn_bits.py --compare out_a.pt out_b.pt"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

if sys.argv[1] == "--compare":
    a, b = torch.load(sys.argv[2]), torch.load(sys.argv[3])
    assert a.keys() == b.keys()
    bad = [k for k in a if not torch.equal(a[k], b[k])]
    print("compared %d tensors: %s" % (len(a), "ALL BIT-IDENTICAL" if not bad else "DIFFER: %s" % bad))
    sys.exit(1 if bad else 0)

from bias_gan_b200 import kernels as K
dev = torch.device("cuda", 0)
out = {}
for (N, H, W, C) in [(2, 48, 72, 728), (1, 37, 19, 128), (3, 5, 7, 2048), (2, 24, 20, 48), (1, 1, 1, 256)]:
    g = torch.Generator(device="cpu").manual_seed(C + H)
    def act(scale=1.0):
        t = K.act_empty((N, H, W, C), dev)
        t.copy_((torch.randn(N, H, W, C, generator=g) * scale).to(dev))
        return t
    x, go, g2, res = act(1.5), act(), act(), act()
    vec = lambda lo, hi: (lo + (hi - lo) * torch.rand(C, generator=g)).to(dev)
    a, b, mean, invstd, gamma = vec(0.5, 1.5), vec(-0.5, 0.5), vec(-0.3, 0.3), vec(0.5, 2.0), vec(0.8, 1.2)
    tag = "%dx%dx%dx%d" % (N, H, W, C)
    for actv in (True, False):
        o = K.bn_apply(x, a, b, act=actv)
        out[tag + "/apply%d" % actv] = o.clone()
        out[tag + "/apply_res%d" % actv] = K.bn_apply(x, a, b, res=res, act=actv).clone()
        for name, second in (("", None), ("_g2", g2)):
            dx, dyh, gr = K.bn_backward(go, o, x, mean, invstd, gamma, act=actv, g2=second)
            out[tag + "/bwd%s%d" % (name, actv)] = torch.cat((dx.float().flatten(), dyh.float().flatten(), gr))
            dx, _, gr = K.bn_backward_recompute(go, a, b, x, mean, invstd, gamma, act=actv, g2=second)
            out[tag + "/rec%s%d" % (name, actv)] = torch.cat((dx.float().flatten(), gr))
            dx, dyh, gr = K.bn_backward_frozen(go, o, x, mean, invstd, gamma, act=actv, g2=second)
            out[tag + "/frz%s%d" % (name, actv)] = torch.cat((dx.float().flatten(), dyh.float().flatten(), gr))
torch.cuda.synchronize()
torch.save({k: v.cpu() for k, v in out.items()}, sys.argv[1])
print("saved %d tensors to %s" % (len(out), sys.argv[1]))
