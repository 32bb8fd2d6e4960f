"""Ad-hoc GPU experiment: does the 128-byte alignment of pixel rows matter to the TMA-fed conv?"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bias_gan_b200 import _lib, kernels as K
dev = torch.device("cuda")
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)

def timeit(fn, n=7, do_flush=True):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(n):
        if do_flush: flush.zero_()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return min(ts) * 1e3

for (H, W, Cin, Cout) in [(48, 72, 728, 728), (48, 72, 768, 768), (48, 72, 704, 704), (48, 72, 1536, 1536), (96, 144, 728, 728), (96, 144, 768, 768)]:
    N = 8
    x = torch.randn(N, H, W, Cin, device=dev).bfloat16()
    w = torch.randn(Cout, 1, Cin, device=dev).bfloat16()
    y = torch.empty(N, H, W, Cout, device=dev, dtype=torch.bfloat16)
    fl = 2.0 * N * H * W * Cin * Cout
    for bn in (128, 256):
        _lib.lib().bgr_debug_conv_block_n(bn)
        t1 = timeit(lambda: K.conv_fprop(x, w, [0], [0], out=y))
        t2 = timeit(lambda: K.conv_fprop(x, w, [0], [0], out=y), do_flush=False)
        print("%dx%d %d->%d bn%d: cold %.1f us %.0f TF/s   warm %.1f us %.0f TF/s" % (H, W, Cin, Cout, bn, t1, fl / t1 / 1e6, t2, fl / t2 / 1e6), flush=True)
_lib.lib().bgr_debug_conv_block_n(0)
