// Per-channel dataset statistics on the device: mean, mean-square, min, max over the spatial axes
// of every sample, in fp64 — the "token" of the reference's summarize_data_gpsro.py:31-50
// (np.load(...).astype(float64); np.average / np.amin / np.amax over axis (1,2) for NCHW files or
// (0,1) for NHWC files).  The tokens of a dataset are then merged on the host exactly as the
// reference does (summarize_data_gpsro.py:5-27): they are a few hundred bytes each.
//
// HBM-bound: every payload byte is read once (4 B/element for fp32 files; measured 4.6 TB/s NCHW,
// 3.8 TB/s NHWC on 453 MB -- read-only streams top out near 4.6 TB/s on this part, cf. bn_stats;
// two loads in flight per thread and eight blocks per SM were measured and are slower), nothing is written but
// B*Cf*splits*4 partials.  Pass 1: a block reduces one chunk of one (sample, channel) plane (NCHW)
// or of one sample's pixels for all channels (NHWC, a thread keeps 4 channels) with 128-bit loads,
// fp64 accumulators per thread, warp shuffles, one shared-memory round.  Pass 2: one thread per
// (sample, channel) adds the partials in a fixed order, so the result does not depend on scheduling.
#include "bgr_internal.h"

namespace bgr {

__device__ __forceinline__ double pos_inf() { return __longlong_as_double(0x7ff0000000000000LL); }

// sums in fp64 (exact products of fp32 inputs); min / max in the file's own type T -- converting to
// fp64 is monotonic, so min(double(x)) == double(min(x)), and fp64 has no single min/max instruction
template <typename T>
struct AccT {
  double s, q;
  T lo, hi;
};
template <typename T> __device__ __forceinline__ T type_max();
template <> __device__ __forceinline__ float type_max<float>() { return __int_as_float(0x7f800000); }
template <> __device__ __forceinline__ double type_max<double>() { return pos_inf(); }
template <> __device__ __forceinline__ int type_max<int>() { return 0x7fffffff; }
template <> __device__ __forceinline__ long long type_max<long long>() { return 0x7fffffffffffffffLL; }
template <typename T> __device__ __forceinline__ T type_lowest() { return -type_max<T>(); }
template <> __device__ __forceinline__ int type_lowest<int>() { return -0x7fffffff - 1; }
template <> __device__ __forceinline__ long long type_lowest<long long>() { return -0x7fffffffffffffffLL - 1; }

template <typename T>
__device__ __forceinline__ void acc_init(AccT<T>& a) {
  a.s = 0.0; a.q = 0.0; a.lo = type_max<T>(); a.hi = type_lowest<T>();
}
template <typename T>
__device__ __forceinline__ void acc_add(AccT<T>& a, T x) {
  const double v = double(x);
  a.s += v;
  a.q = fma(v, v, a.q);
  a.lo = x < a.lo ? x : a.lo;
  a.hi = x > a.hi ? x : a.hi;
}
template <typename T>
__device__ __forceinline__ void acc_merge(AccT<T>& a, const AccT<T>& b) {
  a.s += b.s; a.q += b.q;
  a.lo = b.lo < a.lo ? b.lo : a.lo;
  a.hi = b.hi > a.hi ? b.hi : a.hi;
}
template <typename T>
__device__ __forceinline__ AccT<T> acc_shfl_xor(const AccT<T>& a, int m) {
  AccT<T> r;
  r.s = __shfl_xor_sync(0xffffffffu, a.s, m);
  r.q = __shfl_xor_sync(0xffffffffu, a.q, m);
  r.lo = __shfl_xor_sync(0xffffffffu, a.lo, m);
  r.hi = __shfl_xor_sync(0xffffffffu, a.hi, m);
  return r;
}
template <typename T>
__device__ __forceinline__ void acc_store(const AccT<T>& a, double* o) {
  o[0] = a.s; o[1] = a.q; o[2] = double(a.lo); o[3] = double(a.hi);
}

// ---- NCHW: grid = (splits, B*Cf); block (s, row) reduces elements [s*chunk, (s+1)*chunk) of plane `row`
template <typename T>
__global__ void __launch_bounds__(256)
stats_plane_kernel(const T* __restrict__ src, long P, long chunk, int splits, double* __restrict__ part) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  const long row = blockIdx.y;
  const T* plane = src + row * P;
  const long e0 = long(blockIdx.x) * chunk, e1 = min(e0 + chunk, P);
  AccT<T> a;
  acc_init(a);
  constexpr int V = 16 / sizeof(T);   // elements per 128-bit load
  // chunk is a multiple of V and the plane base is 16-byte aligned when P % V == 0 (checked on the host)
  const bool vec = (P % V == 0) && ((reinterpret_cast<uintptr_t>(src) & 15) == 0);
  if (vec) {
    const long v1 = e0 + ((e1 - e0) / V) * V;
    for (long i = e0 + long(threadIdx.x) * V; i < v1; i += 256L * V) {
      const uint4 u = *reinterpret_cast<const uint4*>(plane + i);
      const T* t = reinterpret_cast<const T*>(&u);
#pragma unroll
      for (int j = 0; j < V; ++j) acc_add(a, t[j]);
    }
    for (long i = v1 + threadIdx.x; i < e1; i += 256) acc_add(a, plane[i]);
  } else {
    for (long i = e0 + threadIdx.x; i < e1; i += 256) acc_add(a, plane[i]);
  }
#pragma unroll
  for (int m = 16; m >= 1; m >>= 1) acc_merge(a, acc_shfl_xor(a, m));
  __shared__ AccT<T> sh[8];
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = a;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < 8; ++w) acc_merge(a, sh[w]);
    double* o = part + (row * splits + blockIdx.x) * 4;
    acc_store(a, o);
  }
}

// ---- NHWC with Cf % V == 0 and (Cf / V) a power of two <= 32: grid = (splits, B); a thread keeps the V
// channels of its column of 128-bit vectors; lanes with equal (lane % G) hold the same channels.
template <typename T>
__global__ void __launch_bounds__(256)
stats_pixels_kernel(const T* __restrict__ src, long P, int Cf, long chunk, int splits, double* __restrict__ part) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  constexpr int V = 16 / sizeof(T);
  const int G = Cf / V;                                  // vectors per pixel
  const long b = blockIdx.y;
  const T* base = src + b * P * Cf;
  const long p0 = long(blockIdx.x) * chunk, p1 = min(p0 + chunk, P);
  AccT<T> a[V];
#pragma unroll
  for (int j = 0; j < V; ++j) acc_init(a[j]);
  const long v0 = p0 * G, v1 = p1 * G;                   // vector range of this chunk; 256 % G == 0
  for (long v = v0 + threadIdx.x; v < v1; v += 256) {
    const uint4 u = *reinterpret_cast<const uint4*>(base + v * V);
    const T* t = reinterpret_cast<const T*>(&u);
#pragma unroll
    for (int j = 0; j < V; ++j) acc_add(a[j], t[j]);
  }
  // v0 % G == 0, so thread t holds vector column t % G = lane % G: combine lanes that differ in the higher bits
  for (int m = 16; m >= G; m >>= 1) {
#pragma unroll
    for (int j = 0; j < V; ++j) acc_merge(a[j], acc_shfl_xor(a[j], m));
  }
  __shared__ AccT<T> sh[8][32][V];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane < G) {
#pragma unroll
    for (int j = 0; j < V; ++j) sh[warp][lane][j] = a[j];
  }
  __syncthreads();
  if (threadIdx.x < G * V) {
    const int g = threadIdx.x / V, j = threadIdx.x % V;
    AccT<T> r = sh[0][g][j];
    for (int w = 1; w < 8; ++w) acc_merge(r, sh[w][g][j]);
    const int c = g * V + j;
    double* o = part + ((b * Cf + c) * splits + blockIdx.x) * 4;
    acc_store(r, o);
  }
}

// ---- NHWC, any Cf: grid = (splits, B*Cf), strided scalar loads (slow, correct)
template <typename T>
__global__ void __launch_bounds__(256)
stats_pixels_scalar_kernel(const T* __restrict__ src, long P, int Cf, long chunk, int splits, double* __restrict__ part) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  const long row = blockIdx.y, b = row / Cf;
  const int c = int(row % Cf);
  const T* base = src + b * P * Cf + c;
  const long p0 = long(blockIdx.x) * chunk, p1 = min(p0 + chunk, P);
  AccT<T> a;
  acc_init(a);
  for (long p = p0 + threadIdx.x; p < p1; p += 256) acc_add(a, base[p * Cf]);
#pragma unroll
  for (int m = 16; m >= 1; m >>= 1) acc_merge(a, acc_shfl_xor(a, m));
  __shared__ AccT<T> sh[8];
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = a;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < 8; ++w) acc_merge(a, sh[w]);
    double* o = part + (row * splits + blockIdx.x) * 4;
    acc_store(a, o);
  }
}

// out[b][k][c], k = mean, mean-square, min, max
__global__ void stats_finalize_kernel(const double* __restrict__ part, int B, int Cf, int splits, double n,
                                      double* __restrict__ out) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * Cf) return;
  const double* p = part + long(i) * splits * 4;
  double s = 0.0, q = 0.0, lo = pos_inf(), hi = -pos_inf();
  for (int k = 0; k < splits; ++k) {
    s += p[k * 4 + 0]; q += p[k * 4 + 1];
    lo = fmin(lo, p[k * 4 + 2]); hi = fmax(hi, p[k * 4 + 3]);
  }
  const int b = i / Cf, c = i % Cf;
  double* o = out + long(b) * 4 * Cf;
  o[0 * Cf + c] = s / n;
  o[1 * Cf + c] = q / n;
  o[2 * Cf + c] = lo;
  o[3 * Cf + c] = hi;
}

template <typename T>
static int channel_stats_impl(const T* src, int layout, int B, int Cf, long P, int splits, double* part,
                              double* out, cudaStream_t st) {
  constexpr int V = 16 / sizeof(T);
  if (layout == BGR_NCHW) {
    long chunk = (P + splits - 1) / splits;
    chunk = (chunk + V - 1) / V * V;
    BGR_CUDA_OK(launch_pdl(stats_plane_kernel<T>, dim3(splits, B * Cf), dim3(256), 0, st, src, P, chunk, splits, part));
  } else {
    const long chunk = (P + splits - 1) / splits;
    const int G = Cf / V;
    const bool vec = Cf % V == 0 && G >= 1 && G <= 32 && (G & (G - 1)) == 0 &&
                     (reinterpret_cast<uintptr_t>(src) & 15) == 0;
    if (vec) BGR_CUDA_OK(launch_pdl(stats_pixels_kernel<T>, dim3(splits, B), dim3(256), 0, st, src, P, Cf, chunk, splits, part));
    else BGR_CUDA_OK(launch_pdl(stats_pixels_scalar_kernel<T>, dim3(splits, B * Cf), dim3(256), 0, st, src, P, Cf, chunk, splits, part));
  }
  BGR_LAUNCH_OK();
  BGR_CUDA_OK(launch_pdl(stats_finalize_kernel, dim3((B * Cf + 127) / 128), dim3(128), 0, st, part, B, Cf, splits, double(P), out));
  BGR_LAUNCH_OK();
  return BGR_OK;
}

}  // namespace bgr

using namespace bgr;

extern "C" int bgr_channel_stats(const void* src, int src_dtype, int src_layout, int B, int Cf, int H, int W,
                                 int splits, double* partials, double* out, void* stream) {
  if (B < 1 || Cf < 1 || H < 1 || W < 1 || splits < 1 || splits > 65535)
    return fail(BGR_ERR_ARG, "channel_stats: bad geometry");
  if (src_layout != BGR_NCHW && src_layout != BGR_NHWC) return fail(BGR_ERR_ARG, "channel_stats: unknown layout");
  if (long(B) * Cf > 65535) return fail(BGR_ERR_ARG, "channel_stats: B*Cf must be <= 65535");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const long P = long(H) * W;
  switch (src_dtype) {
    case BGR_F32: return channel_stats_impl(static_cast<const float*>(src), src_layout, B, Cf, P, splits, partials, out, st);
    case BGR_F64: return channel_stats_impl(static_cast<const double*>(src), src_layout, B, Cf, P, splits, partials, out, st);
    case BGR_I32: return channel_stats_impl(static_cast<const int*>(src), src_layout, B, Cf, P, splits, partials, out, st);
    case BGR_I64: return channel_stats_impl(static_cast<const long long*>(src), src_layout, B, Cf, P, splits, partials, out, st);
  }
  return fail(BGR_ERR_ARG, "channel_stats: unsupported datatype");
}
