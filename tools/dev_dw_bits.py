"""Bit-equality harness for the depthwise marches (plain and BatchNorm-on-load variants): run once per library build
(BGR_LIB=...), then --compare.  See tools/dev_bn_bits.py."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

if sys.argv[1] == "--compare":
    a, b = torch.load(sys.argv[2]), torch.load(sys.argv[3])
    bad = [k for k in a if not torch.equal(a[k], b[k])]
    print("compared %d tensors: %s" % (len(a), "ALL BIT-IDENTICAL" if not bad else "DIFFER: %s" % bad))
    sys.exit(1 if bad else 0)

from bias_gan_b200 import kernels as K
dev = torch.device("cuda", 0)
out = {}
for (N, H, W, C) in [(2, 48, 72, 728), (1, 37, 19, 128), (2, 5, 3, 256), (1, 24, 20, 1536), (3, 9, 70, 48)]:
    g = torch.Generator(device="cpu").manual_seed(C + H)
    def act(scale=1.0):
        t = K.act_empty((N, H, W, C), dev)
        t.copy_((torch.randn(N, H, W, C, generator=g) * scale).to(dev))
        return t
    x, gy = act(1.5), act()
    w9 = torch.randn(9, C, generator=g).to(dev)
    a = (0.5 + torch.rand(C, generator=g)).to(dev)
    b = (torch.rand(C, generator=g) - 0.5).to(dev)
    tag = "%dx%dx%dx%d" % (N, H, W, C)
    for dil in (1, 2):
        out[tag + "/fwd_d%d" % dil] = K.dw_fwd(x, w9, 1, dil).float().clone()
        out[tag + "/dgrad_d%d" % dil] = K.dw_dgrad(gy, w9, 1, dil, tuple(x.shape)).float().clone()
        out[tag + "/wgrad_d%d" % dil] = K.dw_wgrad(x, gy, 1, dil).clone()
    for actv in (True, False):
        out[tag + "/fwd_bn%d" % actv] = K.dw_fwd_bn(x, w9, a, b, actv).float().clone()
        out[tag + "/wgrad_bn%d" % actv] = K.dw_wgrad_bn(x, a, b, actv, gy).clone()
torch.cuda.synchronize()
torch.save({k: v.cpu() for k, v in out.items()}, sys.argv[1])
print("saved %d tensors to %s" % (len(out), sys.argv[1]))
