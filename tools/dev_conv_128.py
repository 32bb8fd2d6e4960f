"""Forced pair kernel on the 128-wide signatures at their real shapes (the pick rule leaves them on the 1-CTA kernel)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bias_gan_b200 import kernels as K, _lib
dev = torch.device("cuda", 0)
L = _lib.lib()

def timed(fn, n=10):
    for _ in range(2):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for i in range(n):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / n

for (N, H, W, Cin, Cout, ks) in ((8, 384, 576, 128, 128, 3), (8, 384, 576, 64, 128, 2), (8, 192, 288, 304, 256, 3),
                                 (8, 192, 288, 256, 256, 3), (8, 192, 288, 256, 256, 1), (8, 96, 144, 256, 728, 1),
                                 (8, 48, 72, 2048, 256, 3)):
    x = K.act_empty((N, H, W, Cin), dev).normal_()
    y = K.act_empty((N, H, W, Cout), dev)
    taps = ks * ks
    w = (torch.randn(Cout, taps, K.pad_channels(Cin), device=dev) * 0.05).bfloat16()[..., :Cin]
    dy = [t // ks - ks // 2 for t in range(taps)]
    dx = [t % ks - ks // 2 for t in range(taps)]
    line = "%dx%dx%d %4d -> %4d taps %d:" % (N, H, W, Cin, Cout, taps)
    for ctas in (1, 2):
        L.bgr_debug_conv_ctas(ctas)
        line += "   ctas %d %7.1f us" % (ctas, timed(lambda: K.conv_fprop(x, w, dy, dx, out=y, want_stats=True)))
    print(line)
L.bgr_debug_conv_ctas(0)
