"""GPU micro-benchmark of the HBM-bound kernels at the trunk's dominant shapes (CUDA events,
L2 flushed between iterations)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bias_gan_b200 import kernels as K

dev = torch.device("cuda")
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)


def timeit(fn, n=5):
    fn(); torch.cuda.synchronize()
    if os.environ.get("BGR_ONCE"):   # under ncu: one launch of each kernel is enough
        return 1.0
    ts = []
    for _ in range(n):
        flush.zero_()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return min(ts) * 1e3  # us


SHAPES = [(8, 384, 576, 128), (8, 192, 288, 256), (8, 96, 144, 728), (8, 48, 72, 728), (8, 48, 72, 1536)]
if len(sys.argv) > 2 and sys.argv[1] == "--only":   # a short run for ncu
    SHAPES = [SHAPES[int(sys.argv[2])]]


def act(N, H, W, C):   # padded pixel stride as the product allocates it
    t = K.act_empty((N, H, W, C), dev)
    t.normal_()
    return t


for (N, H, W, C) in SHAPES:
    x = act(N, H, W, C)
    g = act(N, H, W, C)
    w9 = torch.randn(9, C, device=dev)
    nbytes = x.numel() * 2
    rows = []
    rows.append(("dw_fwd s1d1", timeit(lambda: K.dw_fwd(x, w9, 1, 1)), 2 * nbytes))
    rows.append(("dw_fwd s1d2", timeit(lambda: K.dw_fwd(x, w9, 1, 2)), 2 * nbytes))
    rows.append(("dw_fwd s2d1", timeit(lambda: K.dw_fwd(x, w9, 2, 1)), 1.25 * nbytes))
    rows.append(("dw_dgrad s1d1", timeit(lambda: K.dw_dgrad(g, w9, 1, 1, tuple(x.shape))), 2 * nbytes))
    rows.append(("dw_wgrad s1d1", timeit(lambda: K.dw_wgrad(x, g, 1, 1)), 2 * nbytes))
    rows.append(("dw_wgrad s1d2", timeit(lambda: K.dw_wgrad(x, g, 1, 2)), 2 * nbytes))
    sums = torch.empty(2 * C, dtype=torch.float64, device=dev)
    rows.append(("bn_stats", timeit(lambda: K.bn_stats(x)), nbytes))
    a = torch.ones(C, device=dev); b = torch.zeros(C, device=dev)
    out = K.act_empty((N, H, W, C), dev)
    rows.append(("bn_apply act", timeit(lambda: K.bn_apply(x, a, b, act=True, out=out)), 2 * nbytes))
    rows.append(("bn_apply act+res", timeit(lambda: K.bn_apply(x, a, b, res=g, act=True, out=out)), 3 * nbytes))
    mean = torch.zeros(C, device=dev); inv = torch.ones(C, device=dev)
    rows.append(("bn_backward (2 kernels)", timeit(lambda: K.bn_backward(g, out, x, mean, inv, a, act=True)), 7 * nbytes))
    rows.append(("bn_backward recompute", timeit(lambda: K.bn_backward_recompute(g, a, b, x, mean, inv, a, act=True)), 5 * nbytes))
    rows.append(("bn_backward recompute+g2", timeit(lambda: K.bn_backward_recompute(g, a, b, x, mean, inv, a, act=True, g2=out)), 7 * nbytes))
    rows.append(("bn_backward stored+g2", timeit(lambda: K.bn_backward(g, out, x, mean, inv, a, act=True, g2=out)), 8 * nbytes))
    rows.append(("bn_backward recompute+g2+dyh", timeit(lambda: K.bn_backward_recompute(g, a, b, x, mean, inv, a, act=True, g2=out, want_dyh=True)), 7 * nbytes))
    print("shape N%d %dx%d C%d  (%.1f MB per tensor)" % (N, H, W, C, nbytes / 1e6))
    for name, us, by in rows:
        print("   %-26s %9.1f us   %7.0f GB/s" % (name, us, by / us / 1e3))

# dataset statistics (bgr_channel_stats): one read of the fp32 payload, 8 fields of 16 x 768 x 1152
if len(sys.argv) <= 2:
    xs = torch.randn(8, 16, 768, 1152, device=dev)
    print("shape 8 fields 16x768x1152 fp32  (%.1f MB)" % (xs.numel() * 4 / 1e6))
    for lay, t in (("NCHW", xs), ("NHWC", xs.permute(0, 2, 3, 1).contiguous())):
        us = timeit(lambda: K.channel_stats(t, lay))
        print("   %-26s %9.1f us   %7.0f GB/s" % ("channel_stats " + lay, us, xs.numel() * 4 / us / 1e3))
