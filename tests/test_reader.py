"""Reader parity tests, modelled on the reference's src/numpy_reader/tests/reader_test.py:
every non-failing fixture must equal np.load over the inter x intra thread grid, and the six
failure modes must raise the reference's exception types/messages (including the two the reference
itself fails).  The oracle (oracle/npy_oracle.py) is checked against the same vectors.
"""
import glob
import os
import sys

import numpy as np
import pytest
import torch

from bias_gan_b200 import numpy_reader as nr
from oracle import npy_oracle as orc

HERE = os.path.dirname(os.path.abspath(__file__))
DIRS = [os.path.join(HERE, "golden", "reader")]
_REF_DATA = "/root/reference/src/numpy_reader/tests/data"  # only exists in the build container
if os.path.isdir(_REF_DATA):
    DIRS.append(_REF_DATA)


def _files(pred):
    out = []
    for d in DIRS:
        out += [f for f in sorted(glob.glob(os.path.join(d, "*.npy"))) if pred(os.path.basename(f))]
    return out


ALL = _files(lambda n: "_fail" not in n)
ROW_MAJOR = _files(lambda n: "_rm" in n and "_fail" not in n)
COL_MAJOR = _files(lambda n: "_cm" in n and "_fail" not in n)
DEVICES = [-1, pytest.param(0, marks=pytest.mark.gpu)]


def _get_sample(reader, filename, sample_id=0):
    reader.parse(filename)
    reader.init_file(filename)
    sample = reader.get_sample(sample_id)
    reader.finalize_file()
    return sample


def _equal(array, tensor):
    got = tensor.cpu().numpy()
    assert got.dtype == array.dtype
    np.testing.assert_array_equal(np.asarray(array).reshape(got.shape), got)  # bit-exact


@pytest.mark.parametrize("dev", DEVICES)
@pytest.mark.parametrize("ninter", [1, 2, 4, 8])
@pytest.mark.parametrize("nintra", [1, 2, 4, 8])
@pytest.mark.parametrize("filename", ALL)
def test_single_sample_loads(filename, ninter, nintra, dev):
    npr = nr.numpy_reader(False, dev)
    npr.num_inter_threads = ninter
    npr.num_intra_threads = nintra
    assert (npr.num_inter_threads, npr.num_intra_threads) == (ninter, nintra)
    _equal(np.load(filename), _get_sample(npr, filename))


@pytest.mark.parametrize("dev", DEVICES)
@pytest.mark.parametrize("ninter,nintra", [(1, 1), (2, 4), (8, 8)])
@pytest.mark.parametrize("filename", ROW_MAJOR)
def test_multi_sample_loads_row_major(filename, ninter, nintra, dev):
    npr = nr.numpy_reader(True, dev)
    npr.num_inter_threads = ninter
    npr.num_intra_threads = nintra
    npr.parse(filename)
    npr.init_file(filename)
    array = np.load(filename)
    if not array.shape:
        array = np.array([array])
    assert npr.num_samples == array.shape[0]
    for sample in range(array.shape[0]):
        ref = array[sample, ...]
        _equal(ref if ref.shape else ref.reshape(1), npr.get_sample(sample))
    # negative ids wrap (numpy_reader.cpp:438)
    _equal(array[-1, ...] if array[-1, ...].shape else array[-1, ...].reshape(1), npr.get_sample(-1))
    npr.finalize_file()


@pytest.mark.parametrize("dev", DEVICES)
def test_get_batch(dev):
    filename = os.path.join(DIRS[0], "arr_4d_fp32_rm.npy")
    array = np.load(filename)
    npr = nr.numpy_reader(True, dev)
    npr.num_intra_threads = 4
    npr.parse(filename)
    npr.set_batchsize(2)
    npr.init_file(filename)
    t = npr.get_batch([2, 0])
    assert list(t.shape) == [2] + list(array.shape[1:])
    _equal(array[[2, 0]], t)
    with pytest.raises(IndexError):
        npr.get_batch([0])  # len(ids) != batchsize (numpy_reader.cpp:481-483)
    with pytest.raises(IndexError):
        npr.get_sample(0)  # needs batchsize 1 (numpy_reader.cpp:461-463)
    with pytest.raises(IndexError):
        npr.set_batchsize(4)  # > num_samples
    with pytest.raises(IndexError):
        npr.set_batchsize(0)
    npr.finalize_file()
    single = nr.numpy_reader(False, dev)
    single.parse(filename)
    with pytest.raises(IndexError):
        single.set_batchsize(2)  # batching needs split_axis (numpy_reader.cpp:233-236)


@pytest.mark.parametrize("filename", COL_MAJOR)
def test_multi_sample_loads_column_major(filename):
    npr = nr.numpy_reader(True, -1)
    with pytest.raises(RuntimeError) as e:
        npr.parse(filename)
    assert "reading column-major arrays" in str(e.value)


@pytest.mark.parametrize("filename", COL_MAJOR)
def test_column_major_strides(filename):
    npr = nr.numpy_reader(False, -1)
    npr.parse(filename)
    assert npr.shape == [2, 6, 3] and npr.strides == [1, 2, 12]


@pytest.mark.parametrize("d", DIRS)
def test_failure_modes(d):
    npr = nr.numpy_reader(False, -1)
    with pytest.raises(RuntimeError) as e:
        npr.parse(os.path.join(d, "arr_1d_dtype_fail.npy"))
    assert "unsupported datatype" in str(e.value)
    with pytest.raises(RuntimeError) as e:
        npr.parse(os.path.join(d, "does_not_exist_and_never_did.npy"))
    assert "failed to open file" in str(e.value)
    with pytest.raises(RuntimeError) as e:
        npr.parse(os.path.join(d, "arr_1d_header_fail.npy"))
    assert "ill formatted or corrupt" in str(e.value)
    with pytest.raises(RuntimeError) as e:
        npr.parse(os.path.join(d, "arr_1d_format_fail.npy"))
    assert "not a numpy file" in str(e.value)


@pytest.mark.parametrize("ninter", [1, 4])
@pytest.mark.parametrize("nintra", [1, 2, 8])
@pytest.mark.parametrize("d", DIRS)
def test_corrupted_payload(d, ninter, nintra):
    npr = nr.numpy_reader(False, -1)
    npr.num_inter_threads = ninter
    npr.num_intra_threads = nintra
    with pytest.raises(IndexError) as e:
        _get_sample(npr, os.path.join(d, "arr_1d_corruption_fail.npy"))
    assert "file corruption" in str(e.value)


def test_big_endian_rejected():
    npr = nr.numpy_reader(False, -1)
    with pytest.raises(RuntimeError) as e:
        npr.parse(os.path.join(DIRS[0], "arr_1d_bigendian_fail.npy"))
    assert "big endian" in str(e.value)


def test_reparse_changes_shape():
    npr = nr.numpy_reader(False, -1)
    a, b = os.path.join(DIRS[0], "arr_2d_fp64_rm.npy"), os.path.join(DIRS[0], "arr_3d_int_rm.npy")
    _equal(np.load(a), _get_sample(npr, a))
    _equal(np.load(b), _get_sample(npr, b))
    assert npr.shape == [2, 6, 3]


# ---- the oracle itself against the same golden vectors --------------------------------------
@pytest.mark.parametrize("filename", ALL)
def test_oracle_matches_np_load(filename):
    a = np.load(filename)
    got = orc.read_samples(filename, [0], split_axis=False)[0]
    np.testing.assert_array_equal(got.reshape(a.shape) if a.shape else got.reshape(()), a)
    h = orc.parse_header(filename)
    npr = nr.numpy_reader(False, -1)
    npr.parse(filename)
    assert (npr.shape, npr.strides) == (h["shape"], h["strides"])


def test_oracle_failure_modes():
    d = DIRS[0]
    for name, exc, msg in [("arr_1d_dtype_fail.npy", RuntimeError, "unsupported datatype"),
                           ("nope.npy", RuntimeError, "failed to open file"),
                           ("arr_1d_header_fail.npy", RuntimeError, "ill formatted or corrupt"),
                           ("arr_1d_format_fail.npy", RuntimeError, "not a numpy file")]:
        with pytest.raises(exc) as e:
            orc.parse_header(os.path.join(d, name))
        assert msg in str(e.value)
    with pytest.raises(IndexError):
        orc.read_samples(os.path.join(d, "arr_1d_corruption_fail.npy"), [0])


_REF_SO = os.path.join(os.path.dirname(HERE), "oracle", "_ref", "numpy_reader.so")


@pytest.mark.skipif(not os.path.exists(_REF_SO), reason="reference extension not built (oracle/build_ref.py)")
@pytest.mark.parametrize("filename", ALL)
def test_against_compiled_reference(filename):
    """Our reader and the oracle against the unmodified reference extension (CPU path)."""
    sys.path.insert(0, os.path.dirname(_REF_SO))
    try:
        import numpy_reader as ref_nr
    finally:
        sys.path.pop(0)
    for split in (False, True):
        if split and "_cm" in filename:
            continue
        ref = ref_nr.numpy_reader(split, -1)
        ref.num_intra_threads = 2
        ours = nr.numpy_reader(split, -1)
        ours.num_intra_threads = 2
        ref.parse(filename); ours.parse(filename)
        assert (ref.shape, ref.strides, ref.num_samples) == (ours.shape, ours.strides, ours.num_samples)
        ref.init_file(filename); ours.init_file(filename)
        for i in range(ref.num_samples):
            a, b = ref.get_sample(i), ours.get_sample(i)
            assert a.dtype == b.dtype and list(a.shape) == list(b.shape)
            assert torch.equal(a, b)
            np.testing.assert_array_equal(orc.read_samples(filename, [i], split)[0].reshape(a.shape), a.numpy())
        ref.finalize_file(); ours.finalize_file()


@pytest.mark.gpu
def test_enable_disable_p2p(tmp_path):
    """numpy_reader.enable_p2p / disable_p2p (numpy_reader.cpp:17-56): all-pairs peer access, idempotent, leaves the
    current device alone, and the reader still delivers afterwards.  With one visible GPU there are no pairs and
    both calls succeed; with several, a peer copy between device 0 and 1 must work once enabled."""
    import torch
    from bias_gan_b200 import numpy_reader as nr
    a = np.arange(2 * 3 * 5, dtype=np.float32).reshape(2, 3, 5)
    f = str(tmp_path / "p2p.npy")
    np.save(f, a)
    cur = torch.cuda.current_device()
    r = nr.numpy_reader(False, 0)
    r.enable_p2p()
    r.enable_p2p()          # already enabled pairs are not an error
    assert torch.cuda.current_device() == cur
    r.parse(f); r.init_file(f)
    t = r.get_sample(0)
    r.finalize_file()
    assert np.array_equal(t.cpu().numpy(), a)
    if torch.cuda.device_count() >= 2 and torch.cuda.can_device_access_peer(0, 1):
        u = t.to("cuda:1")
        assert np.array_equal(u.cpu().numpy(), a)
    r.disable_p2p()
    r.disable_p2p()
    assert torch.cuda.current_device() == cur


def test_init_file_refuses_a_file_of_another_shape(tmp_path):
    """The geometry of the parsed file is re-used for every file opened afterwards (gpsro_dataset.py:133-148): a
    larger file cannot be one of the same shape and is refused instead of being delivered mis-shaped; `out=` must
    match the parsed geometry (dtype, size) or the read is refused before the C side writes."""
    a, b = np.zeros((4, 8, 8), np.float32), np.zeros((4, 64, 64), np.float32)
    fa, fb = str(tmp_path / "a.npy"), str(tmp_path / "b.npy")
    np.save(fa, a); np.save(fb, b)
    r = nr.numpy_reader(False, -1)
    r.parse(fa)
    r.init_file(fa)
    with pytest.raises(ValueError):
        r.get_sample(0, out=torch.empty(10))                       # too small
    with pytest.raises(ValueError):
        r.get_sample(0, out=torch.empty(4, 8, 8, dtype=torch.float64))   # wrong dtype
    assert np.array_equal(r.get_sample(0, out=torch.empty(4, 8, 8)).numpy(), a)
    r.finalize_file()
    with pytest.raises(RuntimeError, match="more than the parsed header"):
        r.init_file(fb)
    r.parse(fb); r.init_file(fb)
    assert tuple(r.get_sample(0).shape) == (4, 64, 64)
    r.finalize_file()
