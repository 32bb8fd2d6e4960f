// Fixed-order final reductions of per-block partial results.
//
// No kernel of this library combines partial sums with floating-point atomics: every reduction that
// spans thread blocks (BatchNorm statistics and backward sums, conv / depthwise weight gradients,
// pooling, the linear head, the L1 loss) writes one partial row per block (or per split) into a
// workspace, and one of the kernels below adds the rows in index order.  The result therefore does not
// depend on block scheduling: two runs from the same state are bit-identical, which is what lets the
// parity tests hold fixed bounds (tests/test_determinism_gpu.py).
#include "bgr_internal.h"

namespace bgr {

// out[i] = sum_{r < rows} ws[r * n + i]   (thread per output, rows in order, coalesced across i)
__global__ void __launch_bounds__(128)
sum_rows_kernel(const float* __restrict__ ws, int rows, long n4, float4* __restrict__ out) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  const long i = long(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n4) return;
  const float4* p = reinterpret_cast<const float4*>(ws) + i;
  float4 acc = p[0];
  int r = 1;
  for (; r + 3 < rows; r += 4) {   // four loads in flight, added in row order
    const float4 a = p[long(r) * n4], b = p[long(r + 1) * n4], c = p[long(r + 2) * n4], d = p[long(r + 3) * n4];
    acc.x += a.x; acc.y += a.y; acc.z += a.z; acc.w += a.w;
    acc.x += b.x; acc.y += b.y; acc.z += b.z; acc.w += b.w;
    acc.x += c.x; acc.y += c.y; acc.z += c.z; acc.w += c.w;
    acc.x += d.x; acc.y += d.y; acc.z += d.z; acc.w += d.w;
  }
  for (; r < rows; ++r) {
    const float4 a = p[long(r) * n4];
    acc.x += a.x; acc.y += a.y; acc.z += a.z; acc.w += a.w;
  }
  out[i] = acc;
}

// Few outputs, many rows (depthwise weight gradients: 9*C outputs, ~70 rows): 32 outputs x 8 row lanes per
// block; lane w adds rows w, w+8, ... in order, then the lane sums are added in lane order.
__global__ void __launch_bounds__(256)
sum_rows_lanes_kernel(const float* __restrict__ ws, int rows, long n4, float4* __restrict__ out) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  __shared__ float4 sh[8][32];
  const int ol = threadIdx.x & 31, rl = threadIdx.x >> 5;
  const long i = long(blockIdx.x) * 32 + ol;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  if (i < n4) {
    const float4* p = reinterpret_cast<const float4*>(ws) + i;
    int r = rl;
    for (; r + 8 < rows; r += 16) {
      const float4 a = p[long(r) * n4], b = p[long(r + 8) * n4];
      acc.x += a.x; acc.y += a.y; acc.z += a.z; acc.w += a.w;
      acc.x += b.x; acc.y += b.y; acc.z += b.z; acc.w += b.w;
    }
    if (r < rows) {
      const float4 a = p[long(r) * n4];
      acc.x += a.x; acc.y += a.y; acc.z += a.z; acc.w += a.w;
    }
  }
  sh[rl][ol] = acc;
  __syncthreads();
  if (rl == 0 && i < n4) {
#pragma unroll
    for (int w = 1; w < 8; ++w) {
      const float4 a = sh[w][ol];
      acc.x += a.x; acc.y += a.y; acc.z += a.z; acc.w += a.w;
    }
    out[i] = acc;
  }
}

__global__ void __launch_bounds__(256)
sum_rows_scalar_kernel(const float* __restrict__ ws, int rows, long n, float* __restrict__ out) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  const long i = long(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float acc = 0.f;
  for (int r = 0; r < rows; ++r) acc += ws[long(r) * n + i];
  out[i] = acc;
}

// Few outputs, many rows: one block per output i; thread t adds rows t, t+256, ... in fp64, the 256
// thread sums are combined by a fixed binary tree.  out_f32[i] += sum (single writer) and/or
// out_f64[i] = sum.
__global__ void __launch_bounds__(256)
sum_rows_block_kernel(const float* __restrict__ ws, int rows, int n, float* __restrict__ out_f32_add,
                      double* __restrict__ out_f64) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  __shared__ double sh[256];
  const int i = blockIdx.x;
  double acc = 0.0;
  for (int r = threadIdx.x; r < rows; r += 256) acc += double(ws[long(r) * n + i]);
  sh[threadIdx.x] = acc;
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) {
    if (threadIdx.x < s) sh[threadIdx.x] += sh[threadIdx.x + s];
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    if (out_f32_add) out_f32_add[i] += float(sh[0]);
    if (out_f64) out_f64[i] = sh[0];
  }
}

int launch_sum_rows(const float* ws, int rows, long n, float* out, cudaStream_t st) {
  if (rows < 1 || n < 1) return fail(BGR_ERR_ARG, "sum_rows: empty reduction");
  if ((n % 4) == 0 && ((uintptr_t(ws) | uintptr_t(out)) & 15) == 0) {
    const long n4 = n / 4;
    if (rows >= 16 && n4 <= 8192)
      BGR_CUDA_OK(launch_pdl(sum_rows_lanes_kernel, dim3(unsigned((n4 + 31) / 32)), dim3(256), 0, st, ws, rows, n4, reinterpret_cast<float4*>(out)));
    else
      BGR_CUDA_OK(launch_pdl(sum_rows_kernel, dim3(unsigned((n4 + 127) / 128)), dim3(128), 0, st, ws, rows, n4, reinterpret_cast<float4*>(out)));
  } else {
    BGR_CUDA_OK(launch_pdl(sum_rows_scalar_kernel, dim3(unsigned((n + 255) / 256)), dim3(256), 0, st, ws, rows, n, out));
  }
  BGR_LAUNCH_OK();
  return BGR_OK;
}

int launch_sum_rows_block(const float* ws, int rows, int n, float* out_f32_add, double* out_f64, cudaStream_t st) {
  if (rows < 1 || n < 1) return fail(BGR_ERR_ARG, "sum_rows_block: empty reduction");
  BGR_CUDA_OK(launch_pdl(sum_rows_block_kernel, dim3(n), dim3(256), 0, st, ws, rows, n, out_f32_add, out_f64));
  BGR_LAUNCH_OK();
  return BGR_OK;
}

}  // namespace bgr
