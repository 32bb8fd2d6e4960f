"""TEST INFRASTRUCTURE — builds the *unmodified* reference numpy_reader extension from the sources
where they lie under /root/reference into oracle/_ref/ (git-ignored, travels with gpurun).

Used only to validate oracle/npy_oracle.py and, optionally, as the CPU-baseline reader
(`cpu_baseline.kind = "reference"`).  Never imported by product code.  No reference source is
copied: g++ compiles the files in place; `-include stdexcept` works around gcc 13 strictness in
thread_pool.cpp.
"""
import os
import subprocess
import sys
import sysconfig

REF = "/root/reference/src/numpy_reader/cpp"
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")


def build():
    if not os.path.isdir(REF):
        return None
    import torch
    from torch.utils import cpp_extension as ce
    os.makedirs(OUT, exist_ok=True)
    so = os.path.join(OUT, "numpy_reader.so")
    if os.path.exists(so):
        return so
    inc = ce.include_paths(device_type="cuda") if "device_type" in ce.include_paths.__code__.co_varnames else ce.include_paths(cuda=True)
    cmd = ["g++", "-O2", "-fPIC", "-shared", "-std=c++17", "-include", "stdexcept",
           "-DTORCH_EXTENSION_NAME=numpy_reader", "-DTORCH_API_INCLUDE_EXTENSION_H",
           "-D_GLIBCXX_USE_CXX11_ABI=%d" % int(torch._C._GLIBCXX_USE_CXX11_ABI)]
    for i in inc + [sysconfig.get_paths()["include"], "/usr/local/cuda/include"]:
        cmd += ["-I", i]
    cmd += [os.path.join(REF, "thread_pool.cpp"), os.path.join(REF, "numpy_reader.cpp"), "-o", so]
    tl = os.path.join(os.path.dirname(torch.__file__), "lib")
    cmd += ["-L", tl, "-Wl,-rpath," + tl, "-ltorch", "-ltorch_cpu", "-ltorch_python", "-lc10",
            "-ltorch_cuda", "-lc10_cuda", "-L", "/usr/local/cuda/lib64", "-Wl,-rpath,/usr/local/cuda/lib64",
            "-lcudart", "-lcufile"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stderr[-4000:])
        return None
    return so


if __name__ == "__main__":
    print(build())
