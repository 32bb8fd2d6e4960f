"""CTA-pair (cta_group::2) forward kernel against the 1-CTA kernel: outputs must be bit-identical, the fused
BatchNorm partial rows must add up to the same sums; then the per-round cost of both (see dev_conv_rounds.py)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bias_gan_b200 import kernels as K, _lib

dev = torch.device("cuda", 0)
L = _lib.lib()


def run(x, w, dy, dx, Cout, ctas, bias=None, stats=True):
    L.bgr_debug_conv_ctas(ctas)
    y = K.act_empty(tuple(x.shape[:3]) + (Cout,), dev)
    y.fill_(7.0)
    if stats and bias is None:
        y, part = K.conv_fprop(x, w, dy, dx, out=y, want_stats=True)
        return y.float().clone(), part.double().sum(0), part.shape[0]
    y = K.conv_fprop(x, w, dy, dx, out=y, bias=bias)
    return y.float().clone(), None, 0


bad = 0
cases = [  # N, H, W, Cin, Cout, taps(dilation or 0 for 1x1), bias
    (8, 48, 72, 728, 728, 0, False), (1, 48, 72, 728, 728, 0, False), (3, 17, 23, 256, 256, 1, False),
    (2, 48, 72, 2048, 256, 6, False), (1, 8, 16, 128, 128, 0, False), (5, 8, 16, 64, 128, 0, True),
    (2, 96, 144, 256, 728, 0, False), (1, 30, 50, 304, 256, 1, True), (7, 8, 16, 1536, 2048, 0, False),
    (2, 37, 41, 128, 128, 1, False), (1, 9, 200, 136, 1000, 0, False)]
for (N, H, W, Cin, Cout, dil, use_bias) in cases:
    g = torch.Generator(device="cpu").manual_seed(N * 1000 + Cin)
    x = K.act_empty((N, H, W, Cin), dev)
    x.copy_(torch.randn(N, H, W, Cin, generator=g).to(dev))
    taps = 1 if dil == 0 else 9
    dy = [0] if dil == 0 else [(t // 3 - 1) * dil for t in range(9)]
    dx = [0] if dil == 0 else [(t % 3 - 1) * dil for t in range(9)]
    w = (torch.randn(Cout, taps, K.pad_channels(Cin), generator=g) * 0.05).to(dev).bfloat16()[..., :Cin]
    bias = torch.randn(Cout, generator=g).to(dev) if use_bias else None
    y1, s1, r1 = run(x, w, dy, dx, Cout, 1, bias)
    y2, s2, r2 = run(x, w, dy, dx, Cout, 2, bias)
    same = torch.equal(y1, y2)
    srel = 0.0 if s1 is None else float(((s1 - s2).abs() / (s1.abs() + 1e-3)).max())
    ok = same and srel < 1e-5
    bad += 0 if ok else 1
    print("%-38s y %s  max|dy| %.3g  stats rel %.2e rows %d/%d  %s" % (
        (N, H, W, Cin, Cout, dil, use_bias), "bit-identical" if same else "DIFFERENT", float((y1 - y2).abs().max()),
        srel, r1, r2, "ok" if ok else "FAIL"))
print("FAILED %d" % bad if bad else "all cases ok")


def timed(fn, n=40):
    for _ in range(5):
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
    for (Cin, Cout, taps) in ((728, 728, 1), (1536, 1536, 1), (128, 128, 9), (256, 256, 1), (128, 128, 1)):
        w = (torch.randn(Cout, taps, K.pad_channels(Cin), device=dev) * 0.05).bfloat16()[..., :Cin]
        dy = [0] if taps == 1 else [t // 3 - 1 for t in range(9)]
        dx = [0] if taps == 1 else [t % 3 - 1 for t in range(9)]
        ntile = (Cout + 255) // 256 if Cout > 128 else 1
        groups = 148 // ntile
        line = "Cin %4d Cout %4d taps %d:" % (Cin, Cout, taps)
        for ctas in (1, 2):
            L.bgr_debug_conv_ctas(ctas)
            r = []
            for k in (4, 12):
                N = (groups // 2 * 2) * k
                x = K.act_empty((N, 8, 16, Cin), dev).normal_()
                y = K.act_empty((N, 8, 16, Cout), dev)
                r.append(timed(lambda: K.conv_fprop(x, w, dy, dx, out=y, want_stats=True)))
            line += "   ctas %d: %7.1f %7.1f us, %6.2f us/round" % (ctas, r[0], r[1], (r[1] - r[0]) / 8)
        print(line)
    N, H, W, C = 8, 48, 72, 728
    x = K.act_empty((N, H, W, C), dev).normal_()
    y = K.act_empty((N, H, W, C), dev)
    w = (torch.randn(C, 1, K.pad_channels(C), device=dev) * 0.05).bfloat16()[..., :C]
    for ctas in (1, 2):
        L.bgr_debug_conv_ctas(ctas)
        g = torch.cuda.CUDAGraph()
        K.conv_fprop(x, w, [0], [0], out=y, want_stats=True)
        with torch.cuda.graph(g):
            for _ in range(20):
                K.conv_fprop(x, w, [0], [0], out=y, want_stats=True)
        print("mid-flow 8x48x72 728->728 with statistics, ctas %d: %.1f us per launch (graph of 20)" % (
            ctas, timed(g.replay, 20) / 20))
L.bgr_debug_conv_ctas(0)
