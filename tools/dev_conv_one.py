"""One launch signature of the implicit-GEMM kernel for ncu: the mid-flow 728 -> 728 1x1 conv at 8 x 48 x 72 with the
fused BatchNorm statistics, and its weight gradient.   ncu --set full -k regex:conv_ python tools/dev_conv_one.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bias_gan_b200 import kernels as K

dev = torch.device("cuda", 0)
N, H, W, C = 8, 48, 72, 728
x = K.act_empty((N, H, W, C), dev); x.normal_()
gy = K.act_empty((N, H, W, C), dev); gy.normal_()
w = torch.randn(C, 1, K.pad_channels(C), device=dev).bfloat16()[..., :C]
y = K.act_empty((N, H, W, C), dev)
for _ in range(3):
    K.conv_fprop(x, w, [0], [0], out=y, want_stats=True)
    K.conv_wgrad(x, gy, [0], [0])
torch.cuda.synchronize()
