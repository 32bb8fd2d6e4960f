"""TEST INFRASTRUCTURE — CPU restatement (numpy) of the reference's ingest path.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module; product code never does.

Pinned against: the reference's own reader fixtures/tests (src/numpy_reader/tests/reader_test.py,
equality with np.load) and, in the build container, against the unmodified reference extension
compiled by oracle/build_ref.py.

Functions and the reference lines they restate:
  parse_header      src/numpy_reader/cpp/numpy_reader.cpp:58-185   (ParseFile)
  read_samples      src/numpy_reader/cpp/numpy_reader.cpp:434-497  (readSample / getSample / getBatch)
  norm_constants    src/deepCam/data/gpsro_dataset.py:87-116       (shift / scale from stats.npz)
  standardize       src/deepCam/data/gpsro_dataset.py:169-170      (scale * (x - shift))
  cam_select        src/deepCam/data/cam_numpy_singlefile_dataset.py:195-200 (channel select + NHWC->NCHW)
  stage_reference   the staged tensor the CUDA stage kernel must reproduce: standardise, crop, flip,
                    round to bf16 (RNE), lay out NHWC or 2x2 space-to-depth blocks
  create/merge_token src/utils/summarize_data_gpsro.py:5-50        (per-channel statistics tokens)
"""
import re

import numpy as np

_DTYPES = {"i8": np.int64, "i4": np.int32, "f8": np.float64, "f4": np.float32}
_HEADER_RE = re.compile(r"^\{'descr': '(.*?)', 'fortran_order': (.*?), 'shape': \((.*?)\), \}")


def parse_header(filename, split_axis=False):
    """Returns dict(dtype, fortran, shape, strides (elements), num_samples, offset)."""
    try:
        f = open(filename, "rb")
    except OSError as e:
        raise RuntimeError("NumpyReader: failed to open file %s: %s" % (filename, e))
    with f:
        token = f.read(10)
        if len(token) != 10:
            raise RuntimeError("NumpyReader: could not read header.")
        if b"NUMPY" not in token:
            raise RuntimeError("NumpyReader: not a numpy file.")
        header_len = int.from_bytes(token[8:10], "little")
        header = f.read(header_len).decode("latin-1")
    m = _HEADER_RE.search(header)
    if m is None:
        raise RuntimeError("NumpyReader: cannot parse header, ill formatted or corrupt.")
    descr, order, shapestr = m.group(1), m.group(2), m.group(3)
    if descr[0] not in "<|=":
        raise RuntimeError("NumpyReader: the specified file is in big endian. This is currently not supported.")
    if descr[1:] not in _DTYPES:
        raise RuntimeError("NumpyReader: unsupported datatype.")
    fortran = order != "False"
    shape = [int(s) for s in shapestr.replace(" ", "").split(",") if s != ""]
    if not shape:
        shape = [1]
    num_samples = 1
    if split_axis:
        if fortran:
            raise RuntimeError("NumpyReader: reading column-major arrays (Fortran order) is currently "
                               "only supported if the split_axis option is false.")
        num_samples = shape[0]
        shape = shape[1:] or [1]
    strides = [1] * len(shape)
    if not fortran:
        for i in range(len(shape) - 2, -1, -1):
            strides[i] = strides[i + 1] * shape[i + 1]
    else:
        for i in range(1, len(shape)):
            strides[i] = strides[i - 1] * shape[i - 1]
    return dict(dtype=_DTYPES[descr[1:]], fortran=fortran, shape=shape, strides=strides,
                num_samples=num_samples, offset=10 + header_len)


def read_samples(filename, ids, split_axis=False):
    """Bytes of samples `ids` (wrapping like (i + n) % n) as an ndarray [len(ids)] + shape."""
    h = parse_header(filename, split_axis)
    numel = int(np.prod(h["shape"]))
    isz = np.dtype(h["dtype"]).itemsize
    out = []
    with open(filename, "rb") as f:
        for i in ids:
            i = (i % h["num_samples"] + h["num_samples"]) % h["num_samples"]
            f.seek(h["offset"] + i * numel * isz)
            raw = f.read(numel * isz)
            if len(raw) != numel * isz:
                raise IndexError("NumpyReader: file corruption")
            a = np.frombuffer(raw, dtype=h["dtype"])
            out.append(a.reshape(h["shape"], order="F" if h["fortran"] else "C"))
    return np.stack(out, axis=0)


def norm_constants(stats, prefix, channels, normalization_type):
    """shift, scale (fp32, [C]) as gpsro_dataset.py:87-116 derives them from stats.npz."""
    ch = list(channels)
    if normalization_type == "MinMax":
        shift = stats[prefix + "_minval"][ch]
        scale = 1.0 / (stats[prefix + "_maxval"][ch] - shift)
    else:
        shift = stats[prefix + "_mean"][ch]
        scale = 1.0 / np.sqrt(stats[prefix + "_sqmean"][ch] - np.square(shift))
    return shift.astype(np.float32), scale.astype(np.float32)


def standardize(x_nchw, shift, scale):
    """fp32 scale * (x - shift), broadcast over [C,1,1] (gpsro_dataset.py:169-170)."""
    return (scale.reshape(-1, 1, 1) * (x_nchw.astype(np.float32) - shift.reshape(-1, 1, 1))).astype(np.float32)


def cam_select(batch_nhwc, channels):
    """[B,H,W,Cfile] -> [B,C,H,W] (cam_numpy_singlefile_dataset.py:200)."""
    return np.transpose(batch_nhwc[..., list(channels)], (0, 3, 1, 2))


def bf16_round(x):
    """fp32 -> bf16 round-to-nearest-even, returned as the uint16 bit pattern."""
    u = np.ascontiguousarray(x, dtype=np.float32).view(np.uint32)
    r = (u + (np.uint32(0x7FFF) + ((u >> np.uint32(16)) & np.uint32(1)))) >> np.uint32(16)
    return r.astype(np.uint16)


def bf16_to_f32(bits):
    return (bits.astype(np.uint32) << np.uint32(16)).view(np.float32)


def stage_reference(src, layout, channels, shift, scale, crop=None, flip=(False, False), s2d=False):
    """Staged bf16 bits the CUDA stage kernel must reproduce exactly.

    src: [B,Cfile,H,W] (layout 'NCHW') or [B,H,W,Cfile] ('NHWC'), any reader dtype.
    crop = (y0, x0, h, w) in source pixels (default: whole field); flip = (flip_h, flip_v):
    flip_h mirrors x inside the crop, flip_v mirrors y.  Output uint16 [B,h,w,C] or, with s2d,
    [B,h/2,w/2,4C] where channel = (py*2 + px)*C + c.
    """
    x = src if layout == "NCHW" else np.transpose(src, (0, 3, 1, 2))
    x = x[:, list(channels)].astype(np.float32)
    B, C, H, W = x.shape
    y0, x0, h, w = crop if crop is not None else (0, 0, H, W)
    x = x[:, :, y0:y0 + h, x0:x0 + w]
    if flip[0]:
        x = x[:, :, :, ::-1]
    if flip[1]:
        x = x[:, :, ::-1, :]
    # the kernel computes scale * (x - shift) with separate fp32 sub and mul (no FMA contraction)
    v = (x - shift.reshape(1, -1, 1, 1).astype(np.float32)).astype(np.float32)
    v = (v * scale.reshape(1, -1, 1, 1).astype(np.float32)).astype(np.float32)
    bits = bf16_round(np.transpose(v, (0, 2, 3, 1)))  # [B,h,w,C]
    if s2d:
        bits = bits.reshape(B, h // 2, 2, w // 2, 2, C).transpose(0, 1, 3, 2, 4, 5).reshape(B, h // 2, w // 2, 4 * C)
    return np.ascontiguousarray(bits)


def create_token(arr, data_format="nchw"):
    """summarize_data_gpsro.py:31-50 — one file = one token (n = 1): per-channel mean, mean-square,
    min, max over the spatial axes in fp64."""
    a = np.asarray(arr).astype(np.float64)
    axis = (1, 2) if data_format == "nchw" else (0, 1)
    return (1, np.average(a, axis=axis), np.average(np.square(a), axis=axis),
            np.amin(a, axis=axis), np.amax(a, axis=axis))


def merge_token(t1, t2):
    """summarize_data_gpsro.py:5-27 — pairwise merge weighted by counts."""
    n1, m1, s1, lo1, hi1 = t1
    n2, m2, s2, lo2, hi2 = t2
    n = n1 + n2
    return (n, (n1 / n) * m1 + (n2 / n) * m2, (n1 / n) * s1 + (n2 / n) * s2,
            np.minimum(lo1, lo2), np.maximum(hi1, hi2))


def gpsro_file_shard(names, shuffle=True, seed=12345, shard_idx=0, shard_num=1):
    """gpsro_dataset.py:19-33 (init_files): `data_in_*.npy` names with the prefix stripped, sorted,
    shuffled by np.random.RandomState(seed) (the dataset's only RNG, first use), contiguous shard
    [idx*size, (idx+1)*size) with size = n // shard_num (remainder dropped)."""
    allfiles = sorted(x.replace("data_in_", "") for x in names if x.endswith(".npy") and x.startswith("data_in_"))
    if shuffle:
        np.random.RandomState(seed).shuffle(allfiles)
    size = len(allfiles) // shard_num
    return allfiles[size * shard_idx: size * shard_idx + size]
