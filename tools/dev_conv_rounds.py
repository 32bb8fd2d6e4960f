"""Separate the fixed cost of one implicit-GEMM launch from its per-round cost: the 728 -> 728 1x1 signature on
49*k images of 8 x 16 pixels (one 128-pixel tile each; 3 Cout tiles -> 49 CTA groups -> exactly k rounds of tiles),
timed with CUDA events over back-to-back launches, rotating over buffers larger than L2 (cold) or one buffer (warm)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bias_gan_b200 import kernels as K

dev = torch.device("cuda", 0)
C = int(os.environ.get("C", "728"))


def timed(fn, n):
    for _ in range(5):
        fn(0)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for i in range(n):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / n


w = torch.randn(C, 1, K.pad_channels(C), device=dev).bfloat16()[..., :C]
print("C=%d   us per launch (50 back-to-back launches)" % C)
print("%7s %9s %9s %9s %9s %9s" % ("rounds", "f warm", "f cold", "f nostat", "wg warm", "wg cold"))
for k in (1, 2, 3, 4, 5, 6, 8):
    N = 49 * k
    nbuf = max(2, int(300e6 // (N * 128 * C * 2 * 2)) + 1)
    xs = [K.act_empty((N, 8, 16, C), dev).normal_() for _ in range(nbuf)]
    ys = [K.act_empty((N, 8, 16, C), dev).normal_() for _ in range(nbuf)]
    r = [timed(lambda i: K.conv_fprop(xs[0], w, [0], [0], out=ys[0], want_stats=True), 50),
         timed(lambda i: K.conv_fprop(xs[i % nbuf], w, [0], [0], out=ys[i % nbuf], want_stats=True), 50),
         timed(lambda i: K.conv_fprop(xs[0], w, [0], [0], out=ys[0]), 50),
         timed(lambda i: K.conv_wgrad(xs[0], ys[0], [0], [0]), 50),
         timed(lambda i: K.conv_wgrad(xs[i % nbuf], ys[i % nbuf], [0], [0]), 50)]
    print("%7d " % k + " ".join("%9.1f" % v for v in r))
    del xs, ys
