"""Where the roles of the weight-gradient kernel wait (timing build, BGR_LIB=.../libbgr_timing.so)."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from bias_gan_b200 import kernels as K, _lib
dev = torch.device("cuda", 0)
L = _lib.lib()
buf = np.zeros((512, 8), dtype=np.int64)
for (N, H, W, Cin, Cout) in ((8, 48, 72, 728, 728), (8, 48, 72, 1536, 1536), (8, 96, 144, 728, 728)):
    x = K.act_empty((N, H, W, Cin), dev).normal_()
    gy = K.act_empty((N, H, W, Cout), dev).normal_()
    for ctas in (1, 2):
        L.bgr_debug_conv_ctas(ctas)
        for _ in range(3):
            K.conv_wgrad(x, gy, [0], [0])
        L.bgr_debug_conv_timing(buf.ctypes.data_as(ctypes.c_void_p), 1)
        n = 10
        for _ in range(n):
            K.conv_wgrad(x, gy, [0], [0])
        L.bgr_debug_conv_timing(buf.ctypes.data_as(ctypes.c_void_p), 1)
        t = buf.astype(np.float64) / n
        col = lambda i: (t[:, i][t[:, i] > 0].mean() if (t[:, i] > 0).any() else 0.0)
        print("%dx%dx%d %d->%d ctas %d (%d CTAs): producer wait empty %7.0f role %7.0f | mma wait full %7.0f role %7.0f | "
              "epilogue: wait for accumulator %7.0f, whole role %7.0f cycles" % (
                  N, H, W, Cin, Cout, ctas, int((t[:, 1] > 0).sum()), col(0), col(1), col(3), col(4), col(5), col(7)))
L.bgr_debug_conv_ctas(0)
