"""Where the roles of the forward implicit-GEMM kernel wait (development build with -DBGR_CONV_TIMING, selected with
BGR_LIB=.../libbgr_timing.so).  Cycles per CTA, averaged over the CTAs that ran the role, per launch."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from bias_gan_b200 import kernels as K, _lib

dev = torch.device("cuda", 0)
L = _lib.lib()
buf = np.zeros((512, 8), dtype=np.int64)


def read(reset=1):
    assert L.bgr_debug_conv_timing(buf.ctypes.data_as(ctypes.c_void_p), reset) == 0
    return buf.copy()


def report(tag, launches):
    t = read().astype(np.float64) / launches
    def avg(col):
        v = t[:, col]
        v = v[v > 0]
        return (v.mean(), v.max()) if v.size else (0.0, 0.0)
    names = ["prod wait empty", "prod role", "mma wait tempty", "mma wait full", "mma role", "epi wait tfull",
             "epi wait staging", "epi role"]
    print(tag)
    for i, n in enumerate(names):
        a, m = avg(i)
        print("    %-18s avg %9.0f  max %9.0f cycles" % (n, a, m))


cases = [("mid-flow 8x48x72 728->728", 8, 48, 72, 728, 728), ("12 rounds 728->728", 48 * 12, 8, 16, 728, 728),
         ("8x48x72 1536->1536", 8, 48, 72, 1536, 1536), ("8x192x288 256->256", 8, 192, 288, 256, 256)]
for (name, N, H, W, Cin, Cout) in cases:
    x = K.act_empty((N, H, W, Cin), dev).normal_()
    y = K.act_empty((N, H, W, Cout), dev)
    w = (torch.randn(Cout, 1, K.pad_channels(Cin), device=dev) * 0.05).bfloat16()[..., :Cin]
    for ctas in (1, 2):
        L.bgr_debug_conv_ctas(ctas)
        for _ in range(3):
            K.conv_fprop(x, w, [0], [0], out=y, want_stats=True)
        read()
        n = 10
        for _ in range(n):
            K.conv_fprop(x, w, [0], [0], out=y, want_stats=True)
        report("%s  ctas %d" % (name, ctas), n)
