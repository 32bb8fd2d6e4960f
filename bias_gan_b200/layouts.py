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


def _store(param, kind, val):
    """Enter a derived copy built elsewhere (WeightPack) under the parameter's current version."""
    ent = _cache.get(id(param))
    if ent is None or ent[0]() is not param:
        pid = id(param)
        ent = _cache[pid] = (weakref.ref(param, lambda _r, pid=pid: _cache.pop(pid, None)), {})
    ent[1][kind] = ((param.data_ptr(), param._version), val)


class WeightPack:
    """All derived weight copies of one network — bf16 [Cout,taps,Cin] / [Cin,taps,Cout] GEMM operands of
    the dense convs, fp32 [9,C] taps of the depthwise convs — rebuilt by ONE kernel launch
    (bgr_weights_prepare) into persistent buffers and entered into the per-parameter cache that
    functional.py reads.  Without it every conv rebuilds its own copies after each optimizer step
    (~680 launches of 3-5 us per G+D step, all on the critical path of the step's CUDA graph).
    The stride-2 entry conv (space-to-depth weight) and the linear head keep their own paths."""

    def __init__(self, module):
        import numpy as np
        import torch.nn as nn
        self.items = []   # (param, kind, view handed to the kernels)
        rows, first, nblk = [], [0], 0
        for m in module.modules():
            if not isinstance(m, nn.Conv2d):
                continue
            w = m.weight
            if not (w.is_cuda and w.dtype == torch.float32 and w.is_contiguous()):
                raise ValueError("WeightPack needs contiguous fp32 CUDA parameters")
            Cout, Cin, kh, kw = w.shape
            taps = kh * kw
            if m.groups == 1 and taps <= 9 and not (taps > 1 and m.stride[0] != 1):
                for kind, tr in (("gemm", False), ("gemm_t", True)):
                    nrows, inner = (Cin, Cout) if tr else (Cout, Cin)
                    pitch = _row_pitch(inner)
                    buf = torch.empty((nrows, taps, pitch), dtype=torch.bfloat16, device=w.device)
                    gx = (Cin * taps + 31) // 32 if tr else (pitch + 255) // 256
                    blocks = gx * ((pitch + 31) // 32) if tr else Cout * gx
                    rows.append((w.data_ptr(), buf.data_ptr(), Cout, Cin, taps, pitch, int(tr), gx))
                    self.items.append((w, kind, buf if pitch == inner else buf[..., :inner], buf))
                    nblk += blocks
                    first.append(nblk)
            elif m.groups == Cout and Cin == 1 and taps == 9:
                buf = torch.empty((9, Cout), dtype=torch.float32, device=w.device)
                rows.append((w.data_ptr(), buf.data_ptr(), Cout, 1, 9, Cout, 2, 1))
                self.items.append((w, "dw9", buf, buf))
                nblk += (9 * Cout + 255) // 256
                first.append(nblk)
        self.ndesc, self.total_blocks = len(rows), nblk
        self._key = None
        if not rows:
            return
        dt = np.dtype([("src", "<u8"), ("dst", "<u8"), ("Cout", "<i4"), ("Cin", "<i4"), ("taps", "<i4"),
                       ("pitch", "<i4"), ("mode", "<i4"), ("gx", "<i4")])
        assert dt.itemsize == 40
        dev = self.items[0][0].device
        self.descs = torch.from_numpy(np.array(rows, dtype=dt).view(np.uint8).copy()).to(dev)
        self.first = torch.tensor(first, dtype=torch.int32).to(dev)

    def ensure(self):
        """Rebuild the copies if any parameter moved or changed since the last rebuild, or if the cache
        entries were dropped (invalidate_cache), and enter them into the cache."""
        if not self.ndesc:
            return
        key = tuple((p.data_ptr(), p._version) for p, kind, _, _ in self.items if kind != "gemm_t")
        live = all((_cache.get(id(p)) or (None, {}))[1].get(kind, (None,))[0] == (p.data_ptr(), p._version)
                   for p, kind, _, _ in self.items)
        if key == self._key and live:
            return
        if self._key is not None and any(k[0] != o[0] for k, o in zip(key, self._key)):
            raise RuntimeError("WeightPack: a parameter's storage moved; build a new pack")
        from . import kernels as K
        K.weights_prepare(self.descs, self.first, self.ndesc, self.total_blocks)
        for p, kind, view, _ in self.items:
            _store(p, kind, view)
        self._key = key


def prepare_weights(module):
    """module's WeightPack (created on first use), brought up to date."""
    pack = module.__dict__.get("_bgr_weight_pack")
    if pack is None or any(p.data_ptr() != q for (p, _, _, _), q in zip(pack.items, pack._ptrs)):
        pack = WeightPack(module)
        pack._ptrs = [p.data_ptr() for p, _, _, _ in pack.items]
        module.__dict__["_bgr_weight_pack"] = pack
    pack.ensure()
    return pack


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
