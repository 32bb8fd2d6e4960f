"""Weight-layout transforms between PyTorch's [Cout, Cin, kh, kw] parameters (the state-dict
contract, deeplab.py) and the GEMM layouts the tcgen05 kernels consume.  Tiny tensors; plain torch.
"""
import weakref

import torch

# Derived GEMM-layout copies are cached per parameter and reused until the parameter changes
# (optimizer steps / load_state_dict / .to() bump `_version` or move the storage).  A CUDA-graph
# replay changes parameters behind Python's back: GraphedGanStep calls invalidate_cache().
_cache = {}   # id(param) -> (weakref to param, {kind: ((data_ptr, version), derived tensor)})


def cached(param, kind, fn):
    """fn(param) memoised until the parameter's storage or version changes."""
    ent = _cache.get(id(param))
    if ent is None or ent[0]() is not param:
        pid = id(param)
        ent = _cache[pid] = (weakref.ref(param, lambda _r, pid=pid: _cache.pop(pid, None)), {})
    key = (param.data_ptr(), param._version)
    hit = ent[1].get(kind)
    if hit is not None and hit[0] == key:
        return hit[1]
    with torch.no_grad():
        val = fn(param)
    ent[1][kind] = (key, val)
    return val


def invalidate_cache(module=None):
    """Drop the derived copies of `module`'s parameters (all parameters if None)."""
    if module is None:
        _cache.clear()
        return
    for p in module.parameters():
        _cache.pop(id(p), None)


def _row_pitch(K):
    """GEMM-K row pitch of a weight matrix: rounded up to 64 elements (128-byte lines for TMA)."""
    return K if K % 64 == 0 else (K + 63) // 64 * 64


def _padded_rows(src, rows, taps, K):
    """bf16 [rows, taps, K] view of a [rows, taps, pitch] buffer filled from `src` (same logical shape)."""
    pitch = _row_pitch(K)
    if pitch == K:
        return src.to(torch.bfloat16).contiguous()
    buf = torch.zeros((rows, taps, pitch), dtype=torch.bfloat16, device=src.device)
    buf[..., :K].copy_(src)
    return buf[..., :K]


def gemm_weight(w):
    """[Cout, Cin, kh, kw] -> bf16 [Cout, kh*kw, Cin] (tap-major, Cin contiguous, padded row pitch)."""
    Cout, Cin, kh, kw = w.shape
    if w.is_cuda and w.dtype == torch.float32 and w.is_contiguous():
        from . import kernels as K
        return K.weight_to_gemm(w.detach(), False)   # one kernel instead of permute + cast + pad
    return _padded_rows(w.permute(0, 2, 3, 1).reshape(Cout, kh * kw, Cin), Cout, kh * kw, Cin)


def gemm_weight_t(w):
    """Data-gradient weights: bf16 [Cin, kh*kw, Cout]; used with negated tap offsets."""
    Cout, Cin, kh, kw = w.shape
    if w.is_cuda and w.dtype == torch.float32 and w.is_contiguous():
        from . import kernels as K
        return K.weight_to_gemm(w.detach(), True)
    return _padded_rows(w.permute(1, 2, 3, 0).reshape(Cin, kh * kw, Cout), Cin, kh * kw, Cout)


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
