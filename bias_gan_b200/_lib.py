"""ctypes loader for libbgr.so (the C ABI declared in include/bgr.h).

There is no fallback: if the library is missing the import of any product module fails loudly.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# BGR_LIB = another build of the same C ABI (A/B measurements of one kernel on the same box)
LIB_PATH = os.environ.get("BGR_LIB") or os.path.join(_HERE, "lib", "libbgr.so")

BGR_OK = 0
BGR_ERR_RUNTIME = -1
BGR_ERR_INDEX = -2
BGR_ERR_CUDA = -3
BGR_ERR_ARG = -4

_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                "bias_gan_b200: %s not found — build it with `python -m bias_gan_b200._build` "
                "(there is no CPU or PyTorch fallback)" % LIB_PATH)
        _lib = ctypes.CDLL(LIB_PATH, mode=ctypes.RTLD_GLOBAL)
        _lib.bgr_last_error.restype = ctypes.c_char_p
    return _lib


def last_error():
    return lib().bgr_last_error().decode("utf-8", "replace")


def check(rc):
    """Map a C-ABI status to the Python exception type the reference's pybind layer raises."""
    if rc == BGR_OK:
        return
    msg = last_error()
    if rc == BGR_ERR_INDEX:
        raise IndexError(msg)
    if rc == BGR_ERR_ARG:
        raise ValueError(msg)
    raise RuntimeError(msg)


def ptr(t):
    """Device/host pointer of a torch tensor (or None) as c_void_p."""
    if t is None:
        return ctypes.c_void_p(0)
    return ctypes.c_void_p(t.data_ptr())


def stream_ptr():
    import torch
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def int_array(vals):
    return (ctypes.c_int * len(vals))(*[int(v) for v in vals])
