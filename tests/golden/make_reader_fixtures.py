"""Generates tests/golden/reader/*.npy — the same twelve cases as the reference's reader fixtures
(src/numpy_reader/tests/data, described in SURVEY.md §4), re-created with numpy so no reference
file is copied.  Naming convention as the reference's conftest.py:21-35: `_rm` row-major,
`_cm` column-major, `_fail` failure cases.  Run: python tests/golden/make_reader_fixtures.py
"""
import os

import numpy as np

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "reader")


def main():
    os.makedirs(OUT, exist_ok=True)
    s = lambda name, a: np.save(os.path.join(OUT, name), a)
    s("arr_1d_atomic_fp32_rm.npy", np.array([1.0], dtype="<f4"))
    s("arr_1d_fp32_rm.npy", np.arange(1, 11, dtype="<f4"))
    s("arr_1d_fp64_rm.npy", np.arange(1, 11, dtype="<f8"))
    s("arr_2d_fp64_rm.npy", np.arange(1, 21, dtype="<f8").reshape(4, 5))
    s("arr_2d_int_rm.npy", np.arange(1, 16, dtype="<i8").reshape(3, 5))
    s("arr_2d_int32_rm.npy", np.arange(1, 16, dtype="<i4").reshape(3, 5))
    s("arr_3d_int_rm.npy", np.arange(1, 37, dtype="<i8").reshape(2, 6, 3))
    s("arr_3d_int_cm.npy", np.asfortranarray(np.arange(1, 37, dtype="<i8").reshape(2, 6, 3)))
    s("arr_4d_fp32_rm.npy", np.random.default_rng(7).standard_normal((3, 4, 6, 10)).astype("<f4"))
    s("singleton_1d_fp32_rm.npy", np.array(42.0, dtype="<f8"))
    s("arr_1d_dtype_fail.npy", np.arange(10).astype("<c16"))
    # header with the two version bytes removed
    good = open(os.path.join(OUT, "arr_1d_fp64_rm.npy"), "rb").read()
    open(os.path.join(OUT, "arr_1d_header_fail.npy"), "wb").write(good[:6] + good[8:])
    open(os.path.join(OUT, "arr_1d_format_fail.npy"), "wb").write(
        b"this is a dummy file with a wrong format to cause the reader to crash")
    # payload two elements short of what the header promises
    open(os.path.join(OUT, "arr_1d_corruption_fail.npy"), "wb").write(good[:-16])
    s("arr_1d_bigendian_fail.npy", np.arange(1, 11, dtype=">f8"))


if __name__ == "__main__":
    main()
