"""Per-round cost of the fprop kernel under the kernel's timing-experiment modes (bgr_debug_conv_mode):
0 normal, 1 no statistics, 2 no epilogue, 3 TMEM reads only, 4 all but the TMA store.  Slope between 4 and 12 rounds."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bias_gan_b200 import kernels as K, _lib

dev = torch.device("cuda", 0)


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


for (Cin, Cout) in ((728, 728), (1536, 1536), (256, 256)):
    w = torch.randn(Cout, 1, K.pad_channels(Cin), device=dev).bfloat16()[..., :Cin]
    ntile = (Cout + 255) // 256
    groups = 148 // ntile
    print("Cin %d Cout %d: %d CTA groups; us per launch at 4 / 12 rounds and the slope per round" % (Cin, Cout, groups))
    for mode in (0, 1, 4, 3, 2):
        _lib.lib().bgr_debug_conv_mode(mode)
        r = []
        for k in (4, 12):
            N = groups * k
            x = K.act_empty((N, 8, 16, Cin), dev).normal_()
            y = K.act_empty((N, 8, 16, Cout), dev)
            r.append(timed(lambda: K.conv_fprop(x, w, [0], [0], out=y, want_stats=True)))
        print("  mode %d  %8.1f %8.1f   slope %6.2f us/round" % (mode, r[0], r[1], (r[1] - r[0]) / 8))
    _lib.lib().bgr_debug_conv_mode(0)
