"""CamDataset / CamDataloader (reference data/cam_numpy_singlefile_dataset.py) against
numpy restatements of the reference's indexing: sample and batch tensors bit-exact, batch order with
and without shuffle, drop of the ragged tail, and the fused field path against the oracle."""
import numpy as np
import pytest
import torch

from oracle import npy_oracle as orc

N, H, W, CF = 7, 8, 12, 16


def _files(tmp_path):
    rng = np.random.default_rng(5)
    data = rng.standard_normal((N, H, W, CF)).astype(np.float32)
    label = rng.integers(0, 3, size=(N, H, W, 2)).astype(np.float32)
    names = np.array(["sample_%03d.h5" % i for i in range(N)])
    np.save(tmp_path / "data.npy", data); np.save(tmp_path / "label.npy", label); np.save(tmp_path / "filenames.npy", names)
    return str(tmp_path / "filenames.npy"), str(tmp_path / "data.npy"), str(tmp_path / "label.npy"), data, label, names


def test_cam_dataset_host(tmp_path):
    from bias_gan_b200.cam_numpy_singlefile_dataset import CamDataset
    fn, fd, fl, data, label, names = _files(tmp_path)
    ch = [0, 1, 2, 10]
    ds = CamDataset(fn, fd, fl, ch)
    assert len(ds) == N and ds.shapes == ([H, W, CF], [H, W, 2])
    for idx in (0, 3, N - 1):
        X, Y, name = ds[idx]
        assert name == names[idx]
        assert np.array_equal(X.numpy(), orc.cam_select(data[idx:idx + 1], ch)[0]) and tuple(X.shape) == (len(ch), H, W)
        assert np.array_equal(Y.numpy(), label[idx, ..., 0])
    raw = CamDataset(fn, fd, fl, ch, preprocess=False)
    X, Y, _ = raw[2]
    assert np.array_equal(X.numpy(), data[2]) and np.array_equal(Y.numpy(), label[2])


@pytest.mark.parametrize("batchsize,shuffle", [(2, False), (3, True), (1, False)])
def test_cam_dataloader_host(tmp_path, batchsize, shuffle):
    from bias_gan_b200.cam_numpy_singlefile_dataset import CamDataloader
    fn, fd, fl, data, label, names = _files(tmp_path)
    ch = list(range(8))
    np.random.seed(11)
    dl = CamDataloader(fn, fd, fl, ch, batchsize, shuffle=shuffle, num_intra_threads=2)
    np.random.seed(11)
    order = np.random.permutation(range(N)) if shuffle else np.arange(N)
    batches = list(dl)
    assert len(batches) == N // batchsize                      # the ragged tail is dropped (:187-188)
    for b, (x, y, f) in enumerate(batches):
        idx = order[b * batchsize:(b + 1) * batchsize]
        assert list(f) == list(names[idx])
        assert tuple(x.shape) == (batchsize, len(ch), H, W) and np.array_equal(x.numpy(), orc.cam_select(data[idx], ch))
        assert tuple(y.shape) == (batchsize, H, W) and np.array_equal(y.numpy(), label[idx][..., 0])


@pytest.mark.gpu
@pytest.mark.parametrize("batchsize", [2, 1])
def test_cam_dataloader_fields(tmp_path, batchsize):
    from bias_gan_b200.cam_numpy_singlefile_dataset import CamDataloader
    fn, fd, fl, data, label, names = _files(tmp_path)
    ch = [0, 1, 2, 10, 4, 8, 15, 7]
    dl = CamDataloader(fn, fd, fl, ch, batchsize, device=0, num_intra_threads=2, fields=True)
    ref = CamDataloader(fn, fd, fl, ch, batchsize, device=0, num_intra_threads=2)
    zero, one = np.zeros(len(ch), np.float32), np.ones(len(ch), np.float32)
    for b, ((x, y, f), (xr, yr, fr)) in enumerate(zip(dl, ref)):
        idx = np.arange(b * batchsize, (b + 1) * batchsize)
        want = orc.stage_reference(data[idx], "NHWC", ch, zero, one, s2d=True)
        assert np.array_equal(x.view(torch.int16).cpu().numpy().view(np.uint16), want)       # bit-exact bf16 field
        assert np.array_equal(xr.cpu().numpy(), orc.cam_select(data[idx], ch))              # the reference-style view
        assert torch.equal(y, yr) and list(f) == list(fr) == list(names[idx])
