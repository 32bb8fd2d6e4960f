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

}  // namespace bgr
