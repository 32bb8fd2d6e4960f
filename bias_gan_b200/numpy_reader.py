"""Drop-in for the reference's `numpy_reader` extension module.

Same class and surface as the pybind class at src/numpy_reader/cpp/numpy_reader.cpp:501-536:
`numpy_reader(split_axis=False, device=-1)` with `parse`, `init_file`, `finalize_file`,
`get_sample`, `get_batch`, `set_batchsize`, `enable_p2p`, `disable_p2p`, `print_file_info` and the
properties `num_samples`, `shape`, `strides`, `num_inter_threads`, `num_intra_threads`.
Usage: `from bias_gan_b200 import numpy_reader as nr; nr.numpy_reader(False, 0)`.

The native side (csrc/reader.cpp behind the C ABI) reads straight into the tensor returned to the
caller, so there is no IO-buffer clone; error types/messages follow the reference's pybind
translation (RuntimeError / IndexError, message substrings as asserted by its tests).
"""
import ctypes

import torch

from . import _lib
from ._lib import check

_DTYPES = {0: torch.int64, 1: torch.int32, 2: torch.float64, 3: torch.float32}


class numpy_reader(object):
    def __init__(self, split_axis=False, device=-1):
        self._h = ctypes.c_void_p(0)
        self._split_axis = bool(split_axis)
        self._device = int(device)
        L = _lib.lib()
        L.bgr_reader_bytes_read.restype = ctypes.c_int64
        check(L.bgr_reader_create(int(self._split_axis), self._device, ctypes.byref(self._h)))

    def __del__(self):
        h = getattr(self, "_h", None)
        if h is not None and h.value and _lib is not None and ctypes is not None:   # modules may be gone at interpreter exit
            _lib.lib().bgr_reader_destroy(h)
            self._h = ctypes.c_void_p(0)

    # ---- info -----------------------------------------------------------------------------
    def _info(self):
        ndim = ctypes.c_int(0)
        shape = (ctypes.c_int64 * 8)()
        stride = (ctypes.c_int64 * 8)()
        dtype = ctypes.c_int(-1)
        ns = ctypes.c_int64(0)
        fo = ctypes.c_int(0)
        off = ctypes.c_int64(0)
        bs = ctypes.c_int64(0)
        inter = ctypes.c_int(0)
        intra = ctypes.c_int(0)
        check(_lib.lib().bgr_reader_info(self._h, ctypes.byref(ndim), shape, stride, ctypes.byref(dtype),
                                         ctypes.byref(ns), ctypes.byref(fo), ctypes.byref(off),
                                         ctypes.byref(bs), ctypes.byref(inter), ctypes.byref(intra)))
        n = ndim.value
        return dict(shape=[int(shape[i]) for i in range(n)], strides=[int(stride[i]) for i in range(n)],
                    dtype=dtype.value, num_samples=int(ns.value), fortran=bool(fo.value),
                    offset=int(off.value), batchsize=int(bs.value), inter=inter.value, intra=intra.value)

    @property
    def num_samples(self):
        return self._info()["num_samples"]

    @property
    def shape(self):
        return self._info()["shape"]

    @property
    def strides(self):
        return self._info()["strides"]

    @property
    def num_inter_threads(self):
        return self._info()["inter"]

    @num_inter_threads.setter
    def num_inter_threads(self, n):
        check(_lib.lib().bgr_reader_set_threads(self._h, int(n), self._info()["intra"]))

    @property
    def num_intra_threads(self):
        return self._info()["intra"]

    @num_intra_threads.setter
    def num_intra_threads(self, n):
        check(_lib.lib().bgr_reader_set_threads(self._h, self._info()["inter"], int(n)))

    @property
    def bytes_read(self):
        """Extension: payload bytes delivered so far (reader GB/s accounting)."""
        return int(_lib.lib().bgr_reader_bytes_read(self._h))

    def print_file_info(self):
        i = self._info()
        print("Fortran Order: %s" % ("Yes" if i["fortran"] else "No"))
        print("Endianess: little endian")
        print("Number of Samples: %d" % i["num_samples"])
        print("Shape: (%s)" % ", ".join(str(s) for s in i["shape"]))
        print("Stride: (%s)" % ", ".join(str(s) for s in i["strides"]))

    # ---- file handling --------------------------------------------------------------------
    def parse(self, filename):
        check(_lib.lib().bgr_reader_parse(self._h, str(filename).encode()))

    def init_file(self, filename):
        check(_lib.lib().bgr_reader_init_file(self._h, str(filename).encode()))

    def finalize_file(self):
        check(_lib.lib().bgr_reader_finalize_file(self._h))

    def set_batchsize(self, batch_size):
        check(_lib.lib().bgr_reader_set_batchsize(self._h, int(batch_size)))

    def enable_p2p(self):
        check(_lib.lib().bgr_reader_enable_p2p())

    def disable_p2p(self):
        check(_lib.lib().bgr_reader_disable_p2p())

    # ---- reads ----------------------------------------------------------------------------
    def _read(self, ids, out=None):
        i = self._info()
        if i["dtype"] < 0:
            raise RuntimeError("NumpyReader: parse() has to be called before reading.")
        n = len(ids)
        shape, strides = list(i["shape"]), list(i["strides"])
        if i["batchsize"] > 1:  # numpy_reader.cpp:287-296: batch dim only when batchsize > 1
            numel = 1
            for s in shape:
                numel *= s
            shape, strides = [i["batchsize"]] + shape, [numel] + strides
        dev = torch.device("cpu") if self._device < 0 else torch.device("cuda", self._device)
        if out is None:
            out = torch.empty_strided(shape, strides, dtype=_DTYPES[i["dtype"]], device=dev)
        else:
            # the C side writes numel * itemsize * n bytes unconditionally: a wrong-sized / wrong-typed
            # destination (or a file whose shape differs from the parsed one) must fail here, not overrun
            need = 1
            for v in shape:
                need *= v
            if out.dtype != _DTYPES[i["dtype"]]:
                raise ValueError("numpy_reader: out has dtype %s, the file holds %s" % (out.dtype, _DTYPES[i["dtype"]]))
            if out.device != dev and not (out.device.type == dev.type == "cuda" and (out.device.index or 0) == dev.index):
                raise ValueError("numpy_reader: out lives on %s, the reader delivers to %s" % (out.device, dev))
            if not out.is_contiguous() or out.numel() < need:
                raise ValueError("numpy_reader: out must be contiguous with at least %d elements (has %d)"
                                 % (need, out.numel()))
        arr = (ctypes.c_int64 * n)(*[int(v) for v in ids])
        stream = ctypes.c_void_p(0)
        if self._device >= 0:
            stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        check(_lib.lib().bgr_reader_read(self._h, arr, n, ctypes.c_void_p(out.data_ptr()),
                                         self._device, stream))
        return out

    def get_sample(self, element_id, out=None):
        if self._info()["batchsize"] != 1:
            raise IndexError("NumpyReader: get_sample requires batch size 1, use get_batch instead.")
        return self._read([int(element_id)], out)

    def get_batch(self, element_ids, out=None):
        return self._read(list(element_ids), out)
