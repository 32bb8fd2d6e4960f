"""GPU micro-benchmark of the tcgen05 implicit-GEMM kernels at the step's layer shapes (CUDA events,
L2 flushed between iterations).  usage: python tools/dev_conv_bench.py [block_n ...]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bias_gan_b200 import _lib, kernels as K

dev = torch.device("cuda")
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)


def timeit(fn, n=5):
    fn(); torch.cuda.synchronize()
    ts = []
    if os.environ.get("BGR_B2B"):   # ten launches back to back, warm L2: what a graph replay sees
        for _ in range(n):
            e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
            e0.record()
            for _ in range(10):
                fn()
            e1.record(); torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1) / 10)
        return min(ts) * 1e3
    for _ in range(n):
        if not os.environ.get("BGR_NOFLUSH"):
            flush.zero_()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return min(ts) * 1e3  # us


SHAPES = [  # N, H, W, Cin, Cout, ks, dil
    (8, 48, 72, 728, 728, 1, 1), (8, 48, 72, 1024, 1536, 1, 1), (8, 48, 72, 1536, 1536, 1, 1), (8, 48, 72, 1536, 2048, 1, 1),
    (8, 48, 72, 2048, 256, 1, 1), (8, 48, 72, 2048, 256, 3, 6), (8, 48, 72, 1280, 256, 1, 1),
    (8, 96, 144, 256, 728, 1, 1), (8, 96, 144, 728, 728, 1, 1), (8, 192, 288, 128, 256, 1, 1), (8, 192, 288, 256, 256, 1, 1),
    (8, 192, 288, 304, 256, 3, 1), (8, 192, 288, 256, 256, 3, 1), (8, 192, 288, 256, 16, 1, 1),
    (8, 384, 576, 128, 128, 1, 1), (8, 384, 576, 128, 128, 3, 1), (8, 384, 576, 64, 128, 4, 0)]

args = sys.argv[1:]
if args and args[0] == "--only":   # --only <shape index>: a short run for ncu
    SHAPES = [SHAPES[int(args[1])]]
    args = args[2:]
bns = [int(a) for a in args] or [0]
for (N, H, W, Cin, Cout, ks, dil) in SHAPES:
    x = K.act_empty((N, H, W, Cin), dev); x.normal_()        # padded pixel stride, as the product allocates
    gy = K.act_empty((N, H, W, Cout), dev); gy.normal_()
    if ks == 4:  # the 2x2-tap space-to-depth entry conv
        dy, dx = [-1, -1, 0, 0], [-1, 0, -1, 0]
    else:
        dy, dx = K.taps_for(ks, dil)
    nt = len(dy)
    w = torch.randn(Cout, nt, K.pad_channels(Cin) if Cin > 64 else Cin, device=dev).bfloat16()[..., :Cin]
    wt = torch.randn(Cin, nt, K.pad_channels(Cout) if Cout > 64 else Cout, device=dev).bfloat16()[..., :Cout]
    y = K.act_empty((N, H, W, Cout), dev)
    gx = K.act_empty((N, H, W, Cin), dev)
    sums = torch.empty(2 * Cout, dtype=torch.float64, device=dev)
    dw = torch.zeros(Cout, nt, Cin, device=dev)
    fl = 2.0 * N * H * W * Cin * Cout * nt
    line = "N%d %3dx%3d %4d->%4d t%d: " % (N, H, W, Cin, Cout, nt)
    for bn in bns:
        _lib.lib().bgr_debug_conv_block_n(bn)
        t1 = timeit(lambda: K.conv_fprop(x, w, dy, dx, out=y))
        t2 = timeit(lambda: K.conv_fprop(x, w, dy, dx, out=y, want_stats=True))
        if os.environ.get("BGR_DBG"):
            _lib.lib().bgr_debug_conv_mode(int(os.environ["BGR_DBG"]))
            t2b = timeit(lambda: K.conv_fprop(x, w, dy, dx, out=y, want_stats=True))
            _lib.lib().bgr_debug_conv_mode(0)
            line += "(dbg %.1f us) " % t2b
        t3 = timeit(lambda: K.conv_fprop(gy, wt, dy, dx, out=gx))
        line += "[bn%3d] fprop %7.1f us %6.0f TF/s | +stats %7.1f us | dgrad %7.1f us %6.0f TF/s  " % (
            bn, t1, fl / t1 / 1e6, t2, t3, fl / t3 / 1e6)
    _lib.lib().bgr_debug_conv_block_n(0)
    t4 = timeit(lambda: K.conv_wgrad(x, gy, dy, dx, dw=dw))
    line += "| wgrad %7.1f us %6.0f TF/s" % (t4, fl / t4 / 1e6)
    print(line, flush=True)
