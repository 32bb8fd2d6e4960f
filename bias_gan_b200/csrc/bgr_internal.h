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

// ---- programmatic dependent launch (PDL) ---------------------------------------------------------------
// A step is ~3400 kernels of 20-40 us; each pays its launch latency and prologue (barrier init, TMEM allocation,
// descriptor prefetch, weight / coefficient loads, index arithmetic) after the previous kernel has drained.  With
// the programmatic-stream-serialization launch attribute the next kernel's CTAs are scheduled as soon as every CTA
// of the running kernel has passed pdl_trigger() (or exited), run their prologue on the SMs the tail leaves idle,
// and block in pdl_wait() until the previous kernel has completed and flushed its memory.  Rules kept by every
// kernel launched through launch_pdl(): (1) no global-memory write and no read of anything an earlier kernel of
// the stream may have produced before pdl_wait(); (2) pdl_wait() is executed by every thread that touches global
// memory afterwards.  Without the attribute (BGR_PDL=0, or a predecessor that is not a kernel) both are no-ops.
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// BGR_PDL: 0 / unset = off, 1 = every kernel, 2 = the implicit-GEMM kernels only (their prologue -- barrier init,
// TMEM allocation, tensor-map prefetch -- is the longest), 3 = every kernel except those.
int pdl_mode();
inline bool pdl_enabled(bool gemm = false) {
  const int m = pdl_mode();
  return m == 1 || (m == 2 && gemm) || (m == 3 && !gemm);
}
template <bool GEMM = false, typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = pdl_enabled(GEMM) ? 1 : 0;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}
#endif

}  // namespace bgr
