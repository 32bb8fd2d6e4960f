"""Host-side launch planning of the implicit-GEMM kernels (bgr_conv_plan, no device work): Cout tile width, CTA
pairs (cta_group::2), CTA groups, rounds of tiles, BatchNorm partial rows, weight-gradient splits.  The numbers
asserted here are the ones DESIGN.md section 3.1 quotes for a 148-SM B200 (the SM count the library assumes when
no device is present)."""
import ctypes

import pytest

from bias_gan_b200 import _lib

BN_MAX_ROWS = 448   # include/bgr.h BGR_BN_MAX_ROWS


def plan(kind, N, H, W, Cin, Cout, taps, stats=1):
    out = (ctypes.c_int * 8)()
    rc = _lib.lib().bgr_conv_plan(kind, N, H, W, Cin, Cout, taps, stats, out)
    assert rc == 0, _lib.lib().bgr_last_error()
    return list(out)


def test_mid_flow_forward_runs_as_cta_pairs():
    bn, ctas, n_tiles, groups, m_tiles, rounds, rows, grid = plan(0, 8, 48, 72, 728, 728, 1)
    assert (bn, ctas, n_tiles) == (256, 2, 3)
    assert m_tiles == 216                  # 8 images x 27 tiles of 8 x 16 pixels
    assert groups == 24 and grid == 144    # 74 pairs / 3 Cout tiles
    assert rounds == 5                     # 108 pair tiles over 24 groups: 4.5 rounds run as 5
    assert rows == 48                      # one partial row per CTA of a group


@pytest.mark.parametrize("sig,want_bn,want_ctas", [
    ((8, 384, 576, 128, 128, 9), 128, 1),    # 128-wide tiles stay on the 1-CTA kernel
    ((8, 384, 576, 64, 128, 4), 128, 1),
    ((8, 192, 288, 256, 256, 1), 256, 1),    # 4 k-blocks, one Cout tile: no gain from pairs
    ((8, 96, 144, 256, 728, 1), 256, 2),     # 4 k-blocks, three Cout tiles re-read the pixel tile: pairs
    ((8, 192, 288, 256, 256, 9), 256, 2),
    ((8, 48, 72, 2048, 256, 9), 128, 1),     # 216 tiles: 3 rounds of 128-wide tiles beat 2 rounds of 256-wide ones
    ((8, 48, 72, 1536, 2048, 1), 256, 2),
    ((8, 384, 576, 128, 48, 1), 64, 1)])
def test_forward_tile_width_and_pairing(sig, want_bn, want_ctas):
    p = plan(0, *sig)
    assert (p[0], p[1]) == (want_bn, want_ctas)


def test_forward_plans_respect_capacities():
    """Every signature of the networks (and a few hostile ones): partial rows within BGR_BN_MAX_ROWS, at least one
    group, enough rounds to cover every pixel tile, pairs only in whole clusters."""
    sigs = [(8, 48, 72, 728, 728, 1), (1, 48, 72, 728, 728, 1), (8, 24, 36, 1536, 2048, 1), (8, 384, 576, 128, 128, 9),
            (2, 19, 37, 304, 256, 9), (1, 8, 16, 128, 128, 1), (1, 1, 1, 8, 8, 1), (64, 48, 72, 728, 30000, 1),
            (4, 1536, 2304, 64, 128, 4), (3, 17, 23, 256, 256, 9)]
    for sig in sigs:
        bn, ctas, n_tiles, groups, m_tiles, rounds, rows, grid = plan(0, *sig)
        assert ctas in (1, 2) and groups >= 1 and n_tiles * bn >= sig[4]
        assert rows == groups * ctas and rows <= BN_MAX_ROWS
        assert grid == groups * n_tiles * ctas and grid % ctas == 0
        assert rounds * groups * ctas >= m_tiles
        assert plan(0, *sig, stats=0)[6] == 0


def test_mid_flow_weight_gradient_runs_as_cta_pairs():
    bn, ctas, tiles, splits, cps, chunks, grid, needs_ws = plan(1, 8, 48, 72, 728, 728, 1)
    assert (bn, ctas, tiles) == (256, 2, 9)           # 3 pair tiles of 256 couts x 3 tiles of 256 cins
    assert chunks == 432 and splits == 8 and cps == 54 and grid == 144 and needs_ws == 1
    lib = _lib.lib()
    lib.bgr_conv_wgrad_workspace.restype = ctypes.c_size_t
    assert lib.bgr_conv_wgrad_workspace(8, 48, 72, 728, 728, 1) == splits * 728 * 728 * 4


@pytest.mark.parametrize("sig,want_ctas", [
    ((8, 192, 288, 256, 304, 9), 1),     # 304 couts: 256-cout pair tiles would pad to 512, 128-cout tiles to 384
    ((8, 48, 72, 1536, 1536, 1), 2), ((8, 48, 72, 728, 1024, 1), 2), ((8, 48, 72, 2048, 256, 9), 2),
    ((8, 384, 576, 128, 128, 9), 1),     # 128-wide Cin tile: no pair kernel
    ((8, 96, 144, 728, 128, 1), 1)])     # fewer than 256 couts
def test_weight_gradient_pairing_and_split_coverage(sig, want_ctas):
    bn, ctas, tiles, splits, cps, chunks, grid, needs_ws = plan(1, *sig)
    assert ctas == want_ctas
    assert splits >= 1 and cps * splits >= chunks and cps * (splits - 1) < chunks   # every split has work
    assert grid == tiles * ctas * splits and needs_ws == (1 if splits > 1 else 0)


def test_plan_rejects_bad_arguments():
    out = (ctypes.c_int * 8)()
    lib = _lib.lib()
    lib.bgr_last_error.restype = ctypes.c_char_p
    assert lib.bgr_conv_plan(2, 8, 48, 72, 728, 728, 1, 1, out) != 0
    assert b"conv_plan" in lib.bgr_last_error()
    assert lib.bgr_conv_plan(0, 8, 48, 72, 728, 728, 10, 1, out) != 0
    assert lib.bgr_conv_plan(0, 8, 48, 72, 728, 728, 1, 1, None) != 0
