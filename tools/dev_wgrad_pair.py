"""CTA-pair weight-gradient kernel against the 1-CTA kernel: same values up to the fp32 summation order of another
split partition (both checked against an fp64 reference), bit-identical repeats, and the time per call."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bias_gan_b200 import kernels as K, _lib

dev = torch.device("cuda", 0)
L = _lib.lib()
bad = 0
cases = [(8, 48, 72, 728, 728, 0), (1, 48, 72, 728, 728, 0), (2, 17, 23, 256, 256, 1), (2, 48, 72, 1536, 2048, 0),
         (1, 30, 50, 304, 256, 1), (2, 48, 72, 2048, 256, 6), (3, 9, 20, 264, 1000, 0), (8, 48, 72, 1024, 1536, 0)]
for (N, H, W, Cin, Cout, dil) in cases:
    g = torch.Generator(device="cpu").manual_seed(N * 1000 + Cin)
    x = K.act_empty((N, H, W, Cin), dev)
    x.copy_(torch.randn(N, H, W, Cin, generator=g).to(dev))
    gy = K.act_empty((N, H, W, Cout), dev)
    gy.copy_(torch.randn(N, H, W, Cout, generator=g).to(dev))
    dy = [0] if dil == 0 else [(t // 3 - 1) * dil for t in range(9)]
    dx = [0] if dil == 0 else [(t % 3 - 1) * dil for t in range(9)]
    out = {}
    for ctas in (1, 2):
        L.bgr_debug_conv_ctas(ctas)
        a = K.conv_wgrad(x, gy, dy, dx).clone()
        b = K.conv_wgrad(x, gy, dy, dx).clone()
        out[ctas] = (a, torch.equal(a, b))
    # fp64 reference of the centre tap only (cheap): dw[co, tap, ci] = sum_p gy[p, co] x[p + off, ci]
    tap = len(dy) // 2
    ref = gy.double().reshape(-1, Cout).t() @ x.double().reshape(-1, Cin)
    e1 = float((out[1][0][:, tap].double() - ref).abs().max() / ref.abs().max())
    e2 = float((out[2][0][:, tap].double() - ref).abs().max() / ref.abs().max())
    d12 = float((out[1][0] - out[2][0]).abs().max() / out[1][0].abs().max())
    ok = out[1][1] and out[2][1] and e2 < 2e-5 and d12 < 2e-5
    bad += 0 if ok else 1
    print("%-32s err vs fp64: 1-CTA %.2e pair %.2e   1-CTA vs pair %.2e   repeats identical %s/%s  %s" % (
        (N, H, W, Cin, Cout, dil), e1, e2, d12, out[1][1], out[2][1], "ok" if ok else "FAIL"))
print("FAILED %d" % bad if bad else "all cases ok")


def timed(fn, n=20):
    for _ in range(3):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for i in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / n


if not bad:
    for (N, H, W, Cin, Cout) in ((8, 48, 72, 728, 728), (8, 48, 72, 1536, 2048), (8, 48, 72, 1536, 1536),
                                 (8, 96, 144, 728, 728), (8, 48, 72, 728, 1024), (8, 48, 72, 1024, 1536)):
        x = K.act_empty((N, H, W, Cin), dev).normal_()
        gy = K.act_empty((N, H, W, Cout), dev).normal_()
        line = "%dx%dx%d %4d -> %4d:" % (N, H, W, Cin, Cout)
        for ctas in (1, 2):
            L.bgr_debug_conv_ctas(ctas)
            dw = K.conv_wgrad(x, gy, [0], [0])
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                for _ in range(10):
                    K.conv_wgrad(x, gy, [0], [0], dw=dw)
            line += "   ctas %d %6.1f us" % (ctas, timed(g.replay) / 10)
        print(line + "   (kernel + slab sum, graph of 10)")
L.bgr_debug_conv_ctas(0)
