"""Dataset statistics (reference src/utils/summarize_data_gpsro.py): GPU tokens
against the oracle's numpy restatement, host merge bit-exact, stats.npz end to end."""
import os

import numpy as np
import pytest
import torch

from oracle import npy_oracle as orc


def _rand_token(rng, C, n):
    return (n, rng.standard_normal(C), rng.random(C) + 1.0, -rng.random(C) - 2.0, rng.random(C) + 2.0)


def test_merge_token_is_the_reference_formula():
    from bias_gan_b200 import stats as S
    rng = np.random.default_rng(0)
    a, b, c = _rand_token(rng, 16, 1), _rand_token(rng, 16, 3), _rand_token(rng, 16, 1)
    for t1, t2 in ((a, b), (b, a), (c, S.merge_token(a, b))):
        got, want = S.merge_token(t1, t2), orc.merge_token(t1, t2)
        assert got[0] == want[0]
        for g, w in zip(got[1:], want[1:]):
            assert g.dtype == np.float64 and np.array_equal(g, w)   # bit-exact: same formula, same order


def _check_token(tok, arr, data_format):
    n, mean, sq, lo, hi = tok
    wn, wmean, wsq, wlo, whi = orc.create_token(arr, data_format)
    assert n == wn == 1
    assert np.array_equal(lo, wlo) and np.array_equal(hi, whi)                      # exact
    scale = np.sqrt(wsq)                                                             # the channel's magnitude
    assert np.all(np.abs(mean - wmean) <= 1e-12 * np.maximum(np.abs(wmean), scale))  # fp64, other summation order
    assert np.all(np.abs(sq - wsq) <= 1e-12 * wsq)


@pytest.mark.gpu
@pytest.mark.parametrize("fmt,shape,dtype", [
    ("nchw", (2, 16, 48, 64), np.float32), ("nchw", (1, 3, 7, 9), np.float32), ("nchw", (3, 5, 33, 31), np.float64),
    ("nchw", (2, 4, 16, 20), np.int32), ("nchw", (1, 2, 6, 10), np.int64),
    ("nhwc", (2, 48, 64, 16), np.float32), ("nhwc", (1, 7, 9, 3), np.float32), ("nhwc", (2, 13, 11, 8), np.float64),
    ("nhwc", (1, 40, 24, 128), np.float32), ("nhwc", (1, 9, 5, 2), np.int32)])
def test_tokens_match_oracle(fmt, shape, dtype):
    from bias_gan_b200 import stats as S
    rng = np.random.default_rng(7)
    if np.issubdtype(dtype, np.integer):
        arr = rng.integers(-1000, 1000, size=shape).astype(dtype)
    else:
        C = shape[1] if fmt == "nchw" else shape[3]
        mu = (np.arange(C) - 8.0).reshape((1, C, 1, 1) if fmt == "nchw" else (1, 1, 1, C))
        arr = (rng.standard_normal(shape) * 3.0 + mu).astype(dtype)
    toks = S.create_tokens(torch.from_numpy(arr).cuda(), fmt)
    assert len(toks) == shape[0]
    for b, tok in enumerate(toks):
        _check_token(tok, arr[b], fmt)


@pytest.mark.gpu
def test_full_field_properties():
    """16 x 768 x 1152 (BASELINE configs[1]): size-independent properties."""
    from bias_gan_b200 import stats as S
    C, H, W = 16, 768, 1152
    const = torch.arange(C, dtype=torch.float32, device="cuda").view(1, C, 1, 1).expand(1, C, H, W).contiguous() * 0.37
    n, mean, sq, lo, hi = S.create_tokens(const, "nchw")[0]
    want = (np.arange(C, dtype=np.float32) * np.float32(0.37)).astype(np.float64)
    assert np.array_equal(lo, want) and np.array_equal(hi, want)
    assert np.all(np.abs(mean - want) <= 1e-12 * np.maximum(np.abs(want), 1e-30)) and np.all(np.abs(sq - want * want) <= 1e-12 * want * want)
    g = torch.Generator(device="cuda").manual_seed(1)
    x = torch.randn(2, C, H, W, device="cuda", generator=g)
    t_nchw = S.create_tokens(x, "nchw")
    t_nhwc = S.create_tokens(x.permute(0, 2, 3, 1).contiguous(), "nhwc")    # the same numbers in the other layout
    for a, b in zip(t_nchw, t_nhwc):
        assert np.array_equal(a[3], b[3]) and np.array_equal(a[4], b[4])
        assert np.allclose(a[1], b[1], rtol=0, atol=1e-13) and np.allclose(a[2], b[2], rtol=1e-13, atol=0)
        assert np.all(a[3] <= a[1]) and np.all(a[1] <= a[4]) and np.all(a[2] >= a[1] ** 2)
    # linearity: tokens of 2x are (2 mean, 4 sqmean, 2 min, 2 max) exactly (power-of-two scaling)
    for a, b in zip(t_nchw, S.create_tokens(x * 2.0, "nchw")):
        assert np.array_equal(b[1], 2 * a[1]) and np.array_equal(b[2], 4 * a[2]) and np.array_equal(b[3], 2 * a[3])


@pytest.mark.gpu
def test_summarize_dataset_matches_oracle_loop(tmp_path):
    from bias_gan_b200 import stats as S
    rng = np.random.default_rng(3)
    C, H, W, N = 16, 24, 36, 5
    mu, sig = (np.arange(C) - 8.0).reshape(C, 1, 1), (1.0 + np.arange(C) / 4.0).reshape(C, 1, 1)
    d_tok = l_tok = None
    for i in range(N):
        a = (rng.standard_normal((C, H, W)) * sig + mu).astype(np.float32)
        l = (a + 0.1 * rng.standard_normal((C, H, W))).astype(np.float32)
        np.save(tmp_path / ("data_in_%05d.npy" % i), a)
        np.save(tmp_path / ("data_out_%05d.npy" % i), l)
        ta, tl = orc.create_token(a), orc.create_token(l)     # the reference's loop, :68-82
        d_tok = ta if d_tok is None else orc.merge_token(ta, d_tok)
        l_tok = tl if l_tok is None else orc.merge_token(tl, l_tok)
    out = str(tmp_path / "stats.npz")
    S.summarize_dataset(str(tmp_path), out)
    st = np.load(out)
    assert int(st["count"]) == N
    for prefix, tok in (("data", d_tok), ("label", l_tok)):
        assert np.allclose(st[prefix + "_mean"], tok[1], rtol=1e-12, atol=1e-12)
        assert np.allclose(st[prefix + "_sqmean"], tok[2], rtol=1e-12, atol=0)
        assert np.array_equal(st[prefix + "_minval"], tok[3]) and np.array_equal(st[prefix + "_maxval"], tok[4])
        # what the hot path consumes (gpsro_dataset.py:87-116): identical standardisation constants in fp32
        for kind in ("MeanVariance", "MinMax"):
            want = {k: v for k, v in zip(("mean", "sqmean", "minval", "maxval"), tok[1:])}
            ref_stats = {prefix + "_" + k: v for k, v in want.items()}
            s_got, c_got = orc.norm_constants(st, prefix, list(range(C)), kind)
            s_ref, c_ref = orc.norm_constants(ref_stats, prefix, list(range(C)), kind)
            assert np.allclose(s_got, s_ref, rtol=1e-6, atol=0) and np.allclose(c_got, c_ref, rtol=1e-6, atol=0)
