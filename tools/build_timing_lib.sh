#!/bin/sh
# Development build of the library with the forward implicit-GEMM kernel's role timers (-DBGR_CONV_TIMING):
# lib/libbgr_timing.so, selected with BGR_LIB=...; never shipped (git-ignored like every .so).
set -e
cd "$(dirname "$0")/.."
python -m bias_gan_b200._build > /dev/null
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -Xcompiler -pthread \
  --expt-relaxed-constexpr -I include -I bias_gan_b200/csrc -DBGR_CONV_TIMING -x cu \
  -c bias_gan_b200/csrc/conv_gemm.cu -o bias_gan_b200/build/conv_gemm_timing.o
OBJS=$(ls bias_gan_b200/build/*.o | grep -v "conv_gemm.o\|conv_gemm_timing.o")
nvcc -Wno-deprecated-gpu-targets -shared -o bias_gan_b200/lib/libbgr_timing.so $OBJS bias_gan_b200/build/conv_gemm_timing.o \
  -lcudart -ldl -lpthread
echo bias_gan_b200/lib/libbgr_timing.so
