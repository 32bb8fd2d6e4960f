"""Weight-layout transforms between PyTorch's [Cout, Cin, kh, kw] parameters (the state-dict
contract, deeplab.py) and the GEMM layouts the tcgen05 kernels consume.  Tiny tensors; plain torch.
"""
import torch


def gemm_weight(w):
    """[Cout, Cin, kh, kw] -> bf16 [Cout, kh*kw, Cin] (tap-major, Cin contiguous)."""
    Cout, Cin, kh, kw = w.shape
    return w.permute(0, 2, 3, 1).reshape(Cout, kh * kw, Cin).to(torch.bfloat16).contiguous()


def gemm_weight_t(w):
    """Data-gradient weights: bf16 [Cin, kh*kw, Cout]; used with negated tap offsets."""
    Cout, Cin, kh, kw = w.shape
    return w.permute(1, 2, 3, 0).reshape(Cin, kh * kw, Cout).to(torch.bfloat16).contiguous()


def from_gemm_grad(dw, shape):
    """fp32 [Cout, kh*kw, Cin] -> [Cout, Cin, kh, kw]."""
    Cout, Cin, kh, kw = shape
    return dw.reshape(Cout, kh, kw, Cin).permute(0, 3, 1, 2).contiguous()


# stride-2 3x3 pad-1 conv on C4 channels == 2x2-tap stride-1 conv on 2x2 space-to-depth blocks.
# block tap (ty, tx) in {-1, 0}^2; inside a block (py, px); kernel index k = 2*t + p + 1 in [0, 3).
S2D_DY = [-1, -1, 0, 0]
S2D_DX = [-1, 0, -1, 0]


def s2d_weight(w):
    """[Cout, C4, 3, 3] -> fp32 [Cout, 4, 4*C4]: tap-major, channel = (py, px, c)."""
    Cout, C4, kh, kw = w.shape
    assert kh == 3 and kw == 3
    out = w.new_zeros((Cout, 2, 2, 2, 2, C4))  # [co, ty, tx, py, px, c]
    for ty in range(2):
        for py in range(2):
            ky = 2 * (ty - 1) + py + 1
            if ky < 0:
                continue
            for tx in range(2):
                for px in range(2):
                    kx = 2 * (tx - 1) + px + 1
                    if kx < 0:
                        continue
                    out[:, ty, tx, py, px, :] = w[:, :, ky, kx]
    return out.reshape(Cout, 4, 4 * C4)


def s2d_weight_grad(dw2, shape):
    """Inverse of s2d_weight for gradients: fp32 [Cout, 4, 4*C4] -> [Cout, C4, 3, 3]."""
    Cout, C4, _, _ = shape
    g = dw2.reshape(Cout, 2, 2, 2, 2, C4)
    out = dw2.new_zeros(shape)
    for ty in range(2):
        for py in range(2):
            ky = 2 * (ty - 1) + py + 1
            if ky < 0:
                continue
            for tx in range(2):
                for px in range(2):
                    kx = 2 * (tx - 1) + px + 1
                    if kx < 0:
                        continue
                    out[:, :, ky, kx] = g[:, ty, tx, py, px, :]
    return out
