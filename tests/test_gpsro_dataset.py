"""GPSRODataset / PrefetchLoader against the oracle's restatement of
data/gpsro_dataset.py: file order and sharding, standardisation constants, per-sample tensors
(bit-exact), batched bf16 fields (bit-exact) and the prefetching loader's batch stream."""
import os

import numpy as np
import pytest
import torch

from oracle import npy_oracle as orc

C, H, W = 16, 16, 24


def _make_dir(tmp_path, n, seed=0):
    rng = np.random.default_rng(seed)
    mu, sig = (np.arange(C) - 8.0).reshape(C, 1, 1), (1.0 + np.arange(C) / 4.0).reshape(C, 1, 1)
    d_tok = l_tok = None
    for i in range(n):
        a = (rng.standard_normal((C, H, W)) * sig + mu).astype(np.float32)
        l = (a + 0.1 * rng.standard_normal((C, H, W))).astype(np.float32)
        np.save(tmp_path / ("data_in_%05d.npy" % i), a)
        np.save(tmp_path / ("data_out_%05d.npy" % i), l)
        ta, tl = orc.create_token(a), orc.create_token(l)
        d_tok = ta if d_tok is None else orc.merge_token(ta, d_tok)
        l_tok = tl if l_tok is None else orc.merge_token(tl, l_tok)
    np.savez(tmp_path / "stats.npz", count=n, data_mean=d_tok[1], data_sqmean=d_tok[2], data_minval=d_tok[3],
             data_maxval=d_tok[4], label_mean=l_tok[1], label_sqmean=l_tok[2], label_minval=l_tok[3], label_maxval=l_tok[4])
    return str(tmp_path), str(tmp_path / "stats.npz")


@pytest.mark.parametrize("shuffle,shard_num,shard_idx", [(True, 1, 0), (False, 1, 0), (True, 2, 1), (True, 3, 2), (False, 4, 0)])
@pytest.mark.parametrize("norm", ["MeanVariance", "MinMax"])
def test_files_constants_and_samples_on_host(tmp_path, shuffle, shard_num, shard_idx, norm):
    from bias_gan_b200.gpsro_dataset import GPSRODataset
    root, statsfile = _make_dir(tmp_path, 11)
    ds = GPSRODataset(root, statsfile, list(range(C)), normalization_type=norm, shuffle=shuffle,
                      shard_idx=shard_idx, shard_num=shard_num, num_intra_threads=2)
    want_files = orc.gpsro_file_shard(os.listdir(root), shuffle, 12345, shard_idx, shard_num)
    assert ds.files == want_files and len(ds) == 11 // shard_num
    stats = np.load(statsfile)
    for prefix, shift, scale in (("data", ds.data_shift, ds.data_scale), ("label", ds.label_shift, ds.label_scale)):
        w_shift, w_scale = orc.norm_constants(stats, prefix, list(range(C)), norm)
        assert tuple(shift.shape) == (C, 1, 1) and shift.dtype == torch.float32
        assert np.array_equal(shift.numpy().ravel(), w_shift) and np.array_equal(scale.numpy().ravel(), w_scale)
    for idx in (0, len(ds) - 1):
        data, label, name = ds[idx]
        assert name == want_files[idx]
        for t, prefix in ((data, "data"), (label, "label")):
            raw = np.load(os.path.join(root, ("data_in_" if prefix == "data" else "data_out_") + name))
            sh, sc = orc.norm_constants(stats, prefix, list(range(C)), norm)
            assert np.array_equal(t.numpy(), orc.standardize(raw, sh, sc))     # bit-exact fp32


def test_static_mixing_draws_follow_the_shuffle(tmp_path):
    from bias_gan_b200.gpsro_dataset import GPSRODataset
    root, statsfile = _make_dir(tmp_path, 6)
    ds = GPSRODataset(root, statsfile, list(range(C)), normalization_type="MeanVariance", augmentation_mode="static")
    rng = np.random.RandomState(12345)
    names = sorted(x.replace("data_in_", "") for x in os.listdir(root) if x.startswith("data_in_"))
    rng.shuffle(names)
    assert np.array_equal(ds.mixing, rng.rand(6))      # gpsro_dataset.py:57-58: same generator, after the shuffle


@pytest.mark.gpu
@pytest.mark.parametrize("channels,s2d", [(list(range(C)), True), (list(range(C)), False), ([1, 3, 4, 6, 9, 11, 14, 15], True)])
def test_batch_fields_bit_exact(tmp_path, channels, s2d):
    from bias_gan_b200.gpsro_dataset import GPSRODataset
    root, statsfile = _make_dir(tmp_path, 7)
    dev = torch.device("cuda", 0)
    ds = GPSRODataset(root, statsfile, channels, normalization_type="MeanVariance", shard_num=1,
                      num_intra_threads=2, read_device=dev, send_device=dev)
    stats = np.load(statsfile)
    idx = [4, 0, 6]
    x, y, names = ds.get_batch_fields(idx, s2d=s2d)
    assert names == [ds.files[i] for i in idx]
    for t, prefix, fp in ((x, "data", "data_in_"), (y, "label", "data_out_")):
        raw = np.stack([np.load(os.path.join(root, fp + n)) for n in names])
        sh, sc = orc.norm_constants(stats, prefix, channels, "MeanVariance")
        want = orc.stage_reference(raw, "NCHW", channels, sh, sc, s2d=s2d)
        assert np.array_equal(t.view(torch.int16).cpu().numpy().view(np.uint16), want)


@pytest.mark.gpu
def test_prefetch_loader_stream(tmp_path):
    from bias_gan_b200.gpsro_dataset import GPSRODataset, PrefetchLoader
    root, statsfile = _make_dir(tmp_path, 11)
    dev = torch.device("cuda", 0)
    ds = GPSRODataset(root, statsfile, list(range(C)), normalization_type="MeanVariance",
                      num_intra_threads=2, read_device=dev, send_device=dev)
    ds_sync = GPSRODataset(root, statsfile, list(range(C)), normalization_type="MeanVariance",   # own readers: a reader
                           num_intra_threads=2, read_device=dev, send_device=dev)               # instance is not re-entrant
    loader = PrefetchLoader(ds, 2, depth=2)
    assert len(loader) == 5                                     # drop_last
    for epoch in range(2):
        seen, late = 0, None
        for i, (x, y, names) in enumerate(loader):
            if late is not None:        # the consumer's late read of the previous batch saw it intact
                assert torch.equal(late[0], late[1])
            wx, wy, wn = ds_sync.get_batch_fields([2 * i, 2 * i + 1])  # the synchronous path, on the consumer's stream
            assert names == wn
            assert torch.equal(x.view(torch.int16), wx.view(torch.int16)) and torch.equal(y.view(torch.int16), wy.view(torch.int16))
            # a "training step": long GPU work, then a read of the batch at its very end.  The slot may be
            # rewritten only after this work, although the host asks for the next batch right away.
            torch.cuda._sleep(20_000_000)
            late = (x.view(torch.int16).clone(), wx.view(torch.int16).clone())
            seen += 1
        assert seen == 5 and torch.equal(late[0], late[1])
