"""GPU development check of the tcgen05 conv kernels against torch fp32 convs (not a test)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
from bias_gan_b200 import kernels as K, layouts as L

torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
dev = "cuda"


def relerr(a, b):
    return ((a.float() - b.float()).norm() / (b.float().norm() + 1e-30)).item()


def case(N, H, W, Cin, Cout, ks, dil, bias=False):
    g = torch.Generator(device="cpu").manual_seed(1)
    x = torch.randn(N, Cin, H, W, generator=g).to(dev).bfloat16()
    w = (torch.randn(Cout, Cin, ks, ks, generator=g) / (Cin * ks * ks) ** 0.5).to(dev).bfloat16()
    b = torch.randn(Cout, generator=g).to(dev) if bias else None
    pad = dil * (ks - 1) // 2
    ref = F.conv2d(x.float(), w.float(), b, padding=pad, dilation=dil)
    dy, dx = K.taps_for(ks, dil)
    xn = x.permute(0, 2, 3, 1).contiguous()
    y = K.conv_fprop(xn, L.gemm_weight(w.float()), dy, dx, bias=b)
    torch.cuda.synchronize()
    e1 = relerr(y.permute(0, 3, 1, 2), ref)
    # dgrad / wgrad
    gy = torch.randn(N, Cout, H, W, generator=g).to(dev).bfloat16()
    xr = x.float().requires_grad_(True); wr = w.float().requires_grad_(True)
    F.conv2d(xr, wr, None, padding=pad, dilation=dil).backward(gy.float())
    gyn = gy.permute(0, 2, 3, 1).contiguous()
    gx = K.conv_fprop(gyn, L.gemm_weight_t(w.float()), [-v for v in dy], [-v for v in dx])
    dw = K.conv_wgrad(xn, gyn, dy, dx)
    torch.cuda.synchronize()
    e2 = relerr(gx.permute(0, 3, 1, 2), xr.grad)
    e3 = relerr(L.from_gemm_grad(dw, w.shape), wr.grad)
    print(f"N{N} {H}x{W} {Cin}->{Cout} k{ks} d{dil} bias={bias}: fprop {e1:.2e} dgrad {e2:.2e} wgrad {e3:.2e}", flush=True)


def case_s2d(N, H, W, Cout=128):
    g = torch.Generator(device="cpu").manual_seed(2)
    x = torch.randn(N, 16, H, W, generator=g).to(dev).bfloat16()
    w = (torch.randn(Cout, 16, 3, 3, generator=g) / 12.0).to(dev).bfloat16()
    ref = F.conv2d(x.float(), w.float(), None, stride=2, padding=1)
    xn = x.permute(0, 2, 3, 1).contiguous()
    w2 = L.s2d_weight(w.float()).bfloat16().contiguous()
    y = K.conv_fprop(xn, w2, L.S2D_DY, L.S2D_DX, s2d=True)
    torch.cuda.synchronize()
    e1 = relerr(y.permute(0, 3, 1, 2), ref)
    gy = torch.randn(N, Cout, H // 2, W // 2, generator=g).to(dev).bfloat16()
    wr = w.float().requires_grad_(True)
    F.conv2d(x.float(), wr, None, stride=2, padding=1).backward(gy.float())
    dw2 = K.conv_wgrad(xn, gy.permute(0, 2, 3, 1).contiguous(), L.S2D_DY, L.S2D_DX, s2d=True)
    torch.cuda.synchronize()
    e3 = relerr(L.s2d_weight_grad(dw2, w.shape), wr.grad)
    print(f"s2d N{N} {H}x{W} 16->{Cout}: fprop {e1:.2e} wgrad {e3:.2e}", flush=True)


if __name__ == "__main__":
    case(1, 16, 8, 64, 128, 1, 1)
    case(2, 48, 72, 728, 728, 1, 1)
    case(2, 24, 40, 128, 48, 1, 1)
    case(1, 16, 24, 256, 16, 1, 1, bias=True)
    case(2, 32, 48, 128, 128, 3, 1)
    case(1, 19, 37, 304, 256, 3, 1)
    case(1, 48, 72, 256, 256, 3, 6)
    case(1, 48, 72, 192, 64, 3, 18)
    case_s2d(2, 64, 96)
    case_s2d(1, 38, 74)
    # timing of the dominant shapes
    for (N, H, W, Cin, Cout, ks) in [(8, 48, 72, 728, 728, 1), (8, 384, 576, 128, 128, 3), (8, 192, 288, 304, 256, 3)]:
        x = torch.randn(N, H, W, Cin, device=dev).bfloat16()
        w = torch.randn(Cout, ks * ks, Cin, device=dev).bfloat16()
        dy, dx = K.taps_for(ks, 1)
        y = K.conv_fprop(x, w, dy, dx)
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record()
        for _ in range(5):
            K.conv_fprop(x, w, dy, dx, out=y)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 5
        fl = 2.0 * N * H * W * Cin * Cout * ks * ks
        print(f"fprop N{N} {H}x{W} {Cin}->{Cout} k{ks}: {ms:.3f} ms {fl / ms / 1e9:.1f} TFLOP/s", flush=True)
        dw = K.conv_wgrad(x, y, dy, dx)
        e0.record()
        for _ in range(5):
            K.conv_wgrad(x, y, dy, dx, dw=dw)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 5
        print(f"wgrad same: {ms:.3f} ms {fl / ms / 1e9:.1f} TFLOP/s", flush=True)
