// Internal helpers shared by the translation units of libbgr.so (not part of the C ABI).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <string>

#include "../../include/bgr.h"

namespace bgr {

// thread-local last error string, returned by bgr_last_error()
void set_error(const std::string& msg);
int fail(int code, const std::string& msg);

#define BGR_CUDA_OK(expr)                                                              \
  do {                                                                                 \
    cudaError_t _e = (expr);                                                           \
    if (_e != cudaSuccess)                                                             \
      return ::bgr::fail(BGR_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e)); \
  } while (0)

// every kernel launch of the library is counted (bench.py reports it as gpu_launches)
extern unsigned long long g_launch_count;
#define BGR_LAUNCH_OK()                          \
  do {                                           \
    ++::bgr::g_launch_count;                     \
    BGR_CUDA_OK(cudaGetLastError());             \
  } while (0)

// cuTensorMapEncodeTiled through the runtime's driver entry point (no -lcuda at link time).
// dims/strides innermost first; strides has rank-1 entries (bytes) for dims 1..rank-1.
int encode_tmap_bf16(CUtensorMap* map, const void* base, int rank, const uint64_t* dims,
                     const uint64_t* strides_bytes, const uint32_t* box);

// fixed-order reductions of per-block partial rows (reduce.cu)
//   out[i] = sum_r ws[r*n + i], rows added in index order (fp32)
int launch_sum_rows(const float* ws, int rows, long n, float* out, cudaStream_t st);
//   few outputs, many rows: out_f32_add[i] += sum / out_f64[i] = sum (fp64 accumulation, fixed tree)
int launch_sum_rows_block(const float* ws, int rows, int n, float* out_f32_add, double* out_f64, cudaStream_t st);

// SM count of the current device (cached per device ordinal)
int sm_count();
// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) once per (kernel, device ordinal); thread-safe
int ensure_dyn_smem_impl(const void* kern, int bytes);
template <typename K>
inline int ensure_dyn_smem(K kern, int bytes) { return ensure_dyn_smem_impl(reinterpret_cast<const void*>(kern), bytes); }

}  // namespace bgr
