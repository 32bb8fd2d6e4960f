"""Pair kernel, 728 -> 728, 12 rounds: ring depth 3 / 4 / 5 (timing build only) -- is the operand feed bound by
latency (bytes in flight) or by rate?"""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from bias_gan_b200 import kernels as K, _lib
dev = torch.device("cuda", 0)
L = _lib.lib()
buf = np.zeros((512, 8), dtype=np.int64)
N, Cin, Cout = 48 * 12, 728, 728
x = K.act_empty((N, 8, 16, Cin), dev).normal_()
y = K.act_empty((N, 8, 16, Cout), dev)
w = (torch.randn(Cout, 1, K.pad_channels(Cin), device=dev) * 0.05).bfloat16()[..., :Cin]
L.bgr_debug_conv_ctas(2)
for stages in (3, 4, 5):
    L.bgr_debug_conv_stages(stages)
    for _ in range(3):
        K.conv_fprop(x, w, [0], [0], out=y, want_stats=True)
    L.bgr_debug_conv_timing(buf.ctypes.data_as(ctypes.c_void_p), 1)
    for _ in range(10):
        K.conv_fprop(x, w, [0], [0], out=y, want_stats=True)
    L.bgr_debug_conv_timing(buf.ctypes.data_as(ctypes.c_void_p), 1)
    t = buf.astype(np.float64) / 10
    col = lambda i: t[:, i][t[:, i] > 0].mean()
    print("stages %d: mma role %8.0f  wait full %8.0f  wait tempty %6.0f   producer wait empty %8.0f cycles" % (
        stages, col(4), col(3), col(2), col(0)))
