// Depthwise 3x3 convolution with the reference's "fixed padding" (pad = dilation on every side,
// src/deepCam/architecture/gpsro/deeplab.py:66-72,75-87), stride 1 or 2, bf16 NHWC.
//   fwd   : y[n,oy,ox,c] = sum_k x[n, oy*s + (ky-1)*d, ox*s + (kx-1)*d, c] * w[k][c]
//   dgrad : adjoint gather over the taps
//   wgrad : dw[k][c] = sum_pixels dy * x_shifted   (per-channel reduction)
// plus the stride-2 pixel subsample used in front of the 1x1 stride-2 skip convs (deeplab.py:95)
// and its adjoint.  CUDA-core, HBM/L2-bound: 8 channels (128 bit) per thread.
#include <cstdlib>
#include <cuda_bf16.h>

#include "bgr_internal.h"

#ifndef BGR_DW_FFMA2
#define BGR_DW_FFMA2 1   // packed fp32 FMAs (FFMA2): -16 % instructions in the issue-bound marches; 1-2 % per kernel (r02)
#endif

namespace bgr {

__device__ __forceinline__ void dw_unpack8(const uint4& u, float (&f)[8]) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 t = __bfloat1622float2(h[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}
__device__ __forceinline__ uint4 dw_pack8(const float (&f)[8]) {
  uint4 u;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
  return u;
}

struct DwGeom {
  int N, H, W, C;     // input
  int Ho, Wo;         // output
  int stride, dil;
  int ldx, ldy;       // pixel strides (elements)
};

// w: fp32 [9][C]
__global__ void __launch_bounds__(256)
dw_fwd_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ w,
              __nv_bfloat16* __restrict__ y, const DwGeom g) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  const int G = g.C >> 3;
  const long total = long(g.N) * g.Ho * g.Wo * G;
  for (long i = long(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += long(gridDim.x) * blockDim.x) {
    const int cg = int(i % G);
    long r = i / G;
    const int ox = int(r % g.Wo); r /= g.Wo;
    const int oy = int(r % g.Ho);
    const int n = int(r / g.Ho);
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
    for (int ky = 0; ky < 3; ++ky) {
      const int iy = oy * g.stride + (ky - 1) * g.dil;
      if (iy < 0 || iy >= g.H) continue;
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) {
        const int ix = ox * g.stride + (kx - 1) * g.dil;
        if (ix < 0 || ix >= g.W) continue;
        float xv[8];
        dw_unpack8(*reinterpret_cast<const uint4*>(x + (size_t(size_t(n) * g.H + iy) * g.W + ix) * g.ldx + cg * 8), xv);
        const float4 w0 = *reinterpret_cast<const float4*>(w + (ky * 3 + kx) * g.C + cg * 8);
        const float4 w1 = *reinterpret_cast<const float4*>(w + (ky * 3 + kx) * g.C + cg * 8 + 4);
        acc[0] = fmaf(xv[0], w0.x, acc[0]); acc[1] = fmaf(xv[1], w0.y, acc[1]);
        acc[2] = fmaf(xv[2], w0.z, acc[2]); acc[3] = fmaf(xv[3], w0.w, acc[3]);
        acc[4] = fmaf(xv[4], w1.x, acc[4]); acc[5] = fmaf(xv[5], w1.y, acc[5]);
        acc[6] = fmaf(xv[6], w1.z, acc[6]); acc[7] = fmaf(xv[7], w1.w, acc[7]);
      }
    }
    *reinterpret_cast<uint4*>(y + (size_t(size_t(n) * g.Ho + oy) * g.Wo + ox) * g.ldy + cg * 8) = dw_pack8(acc);
  }
}

// dx[n,iy,ix,c] = sum_k dy[n,(iy-(ky-1)d)/s,(ix-(kx-1)d)/s,c] * w[k][c]  where divisible/in range
__global__ void __launch_bounds__(256)
dw_dgrad_kernel(const __nv_bfloat16* __restrict__ dy, const float* __restrict__ w,
                __nv_bfloat16* __restrict__ dx, const DwGeom g) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  const int G = g.C >> 3;
  const long total = long(g.N) * g.H * g.W * G;
  for (long i = long(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += long(gridDim.x) * blockDim.x) {
    const int cg = int(i % G);
    long r = i / G;
    const int ix = int(r % g.W); r /= g.W;
    const int iy = int(r % g.H);
    const int n = int(r / g.H);
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
    for (int ky = 0; ky < 3; ++ky) {
      const int ty = iy - (ky - 1) * g.dil;
      if (ty < 0 || (ty % g.stride)) continue;
      const int oy = ty / g.stride;
      if (oy >= g.Ho) continue;
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) {
        const int tx = ix - (kx - 1) * g.dil;
        if (tx < 0 || (tx % g.stride)) continue;
        const int ox = tx / g.stride;
        if (ox >= g.Wo) continue;
        float gv[8];
        dw_unpack8(*reinterpret_cast<const uint4*>(dy + (size_t(size_t(n) * g.Ho + oy) * g.Wo + ox) * g.ldy + cg * 8), gv);
        const float4 w0 = *reinterpret_cast<const float4*>(w + (ky * 3 + kx) * g.C + cg * 8);
        const float4 w1 = *reinterpret_cast<const float4*>(w + (ky * 3 + kx) * g.C + cg * 8 + 4);
        acc[0] = fmaf(gv[0], w0.x, acc[0]); acc[1] = fmaf(gv[1], w0.y, acc[1]);
        acc[2] = fmaf(gv[2], w0.z, acc[2]); acc[3] = fmaf(gv[3], w0.w, acc[3]);
        acc[4] = fmaf(gv[4], w1.x, acc[4]); acc[5] = fmaf(gv[5], w1.y, acc[5]);
        acc[6] = fmaf(gv[6], w1.z, acc[6]); acc[7] = fmaf(gv[7], w1.w, acc[7]);
      }
    }
    *reinterpret_cast<uint4*>(dx + (size_t(size_t(n) * g.H + iy) * g.W + ix) * g.ldx + cg * 8) = dw_pack8(acc);
  }
}

// dw[c][k] = sum over output pixels.  Thread = (pixel lane, 8-channel group); each block walks a
// contiguous range of output pixels; 72 fp32 partials per thread are added into shared memory one
// pixel lane at a time (a fixed order) and the block writes one partial row ws[blockIdx.x][C*9];
// the rows are added in block order by sum_rows (reduce.cu).
__global__ void __launch_bounds__(256)
dw_wgrad_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ dy,
                float* __restrict__ ws, const DwGeom g, long pix_per_block) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  extern __shared__ float sh[];  // [9][C]
  for (int i = threadIdx.x; i < 9 * g.C; i += blockDim.x) sh[i] = 0.f;
  __syncthreads();
  const int G = g.C >> 3;
  const int R = G >= 256 ? 1 : 256 / G;
  const long P = long(g.N) * g.Ho * g.Wo;
  const long p0 = long(blockIdx.x) * pix_per_block, p1 = min(p0 + pix_per_block, P);
  const int r = G >= 256 ? 0 : threadIdx.x / G;
  for (int cg = threadIdx.x % G; cg < G; cg += (G >= 256 ? 256 : G)) {
    float acc[9][8];
#pragma unroll
    for (int k = 0; k < 9; ++k)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[k][j] = 0.f;
    if (r < R) {
      for (long p = p0 + r; p < p1; p += R) {
        const int ox = int(p % g.Wo);
        const int oy = int((p / g.Wo) % g.Ho);
        const int n = int(p / (long(g.Wo) * g.Ho));
        float gv[8];
        dw_unpack8(*reinterpret_cast<const uint4*>(dy + size_t(p) * g.ldy + cg * 8), gv);
#pragma unroll
        for (int ky = 0; ky < 3; ++ky) {
          const int iy = oy * g.stride + (ky - 1) * g.dil;
          if (iy < 0 || iy >= g.H) continue;
#pragma unroll
          for (int kx = 0; kx < 3; ++kx) {
            const int ix = ox * g.stride + (kx - 1) * g.dil;
            if (ix < 0 || ix >= g.W) continue;
            float xv[8];
            dw_unpack8(*reinterpret_cast<const uint4*>(x + (size_t(size_t(n) * g.H + iy) * g.W + ix) * g.ldx + cg * 8), xv);
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[ky * 3 + kx][j] = fmaf(gv[j], xv[j], acc[ky * 3 + kx][j]);
          }
        }
      }
    }
    if (R == 1) {   // one thread per channel group: no sharing
      if (r == 0) {
#pragma unroll
        for (int k = 0; k < 9; ++k)
#pragma unroll
          for (int j = 0; j < 8; ++j) sh[k * g.C + cg * 8 + j] += acc[k][j];
      }
    } else {        // pixel lanes add in lane order: schedule-independent (one pass per thread: G < 256)
      for (int turn = 0; turn < R; ++turn) {
        if (turn == r) {
#pragma unroll
          for (int k = 0; k < 9; ++k)
#pragma unroll
            for (int j = 0; j < 8; ++j) sh[k * g.C + cg * 8 + j] += acc[k][j];
        }
        __syncthreads();
      }
    }
    if (G < 256) break;
  }
  __syncthreads();
  // sh is [9][C]; the partial row is stored [C][9] = the parameter's own [C,1,3,3] layout
  float* row = ws + size_t(blockIdx.x) * 9 * g.C;
  for (int i = threadIdx.x; i < 9 * g.C; i += blockDim.x) row[i] = sh[(i % 9) * g.C + i / 9];
}

// y[n,oy,ox,:] = x[n,2oy,2ox,:]  /  adjoint: dx = 0 except dx[n,2oy,2ox,:] = dy[n,oy,ox,:]
__global__ void __launch_bounds__(256)
subsample2_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y, const DwGeom g) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  const int G = g.C >> 3;
  const long total = long(g.N) * g.Ho * g.Wo * G;
  for (long i = long(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += long(gridDim.x) * blockDim.x) {
    const int cg = int(i % G);
    long r = i / G;
    const int ox = int(r % g.Wo); r /= g.Wo;
    const int oy = int(r % g.Ho);
    const int n = int(r / g.Ho);
    *reinterpret_cast<uint4*>(y + (size_t(size_t(n) * g.Ho + oy) * g.Wo + ox) * g.ldy + cg * 8) =
        *reinterpret_cast<const uint4*>(x + (size_t(size_t(n) * g.H + 2 * oy) * g.W + 2 * ox) * g.ldx + cg * 8);
  }
}
__global__ void __launch_bounds__(256)
subsample2_bwd_kernel(const __nv_bfloat16* __restrict__ dy, __nv_bfloat16* __restrict__ dx, const DwGeom g) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  const int G = g.C >> 3;
  const long total = long(g.N) * g.H * g.W * G;
  for (long i = long(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += long(gridDim.x) * blockDim.x) {
    const int cg = int(i % G);
    long r = i / G;
    const int ix = int(r % g.W); r /= g.W;
    const int iy = int(r % g.H);
    const int n = int(r / g.H);
    uint4 v = make_uint4(0, 0, 0, 0);
    if (!(ix & 1) && !(iy & 1))
      v = *reinterpret_cast<const uint4*>(dy + (size_t(size_t(n) * g.Ho + (iy >> 1)) * g.Wo + (ix >> 1)) * g.ldy + cg * 8);
    *reinterpret_cast<uint4*>(dx + (size_t(size_t(n) * g.H + iy) * g.W + ix) * g.ldx + cg * 8) = v;
  }
}

// ---------------------------------------------------------------- strip-mined variants
// One thread = 8 channels x T consecutive output columns: the 72 weights live in registers and
// every input column of the window is loaded once per row (18 loads per 4 outputs instead of 36
// for stride 1), which moves these kernels from LSU-bound to HBM/L2-bound.
// REVERSE applies the taps mirrored (w[8-k]): the stride-1 data gradient.
template <int S, int D, int T, bool REVERSE>
__global__ void __launch_bounds__(256)
dw_fwd_strip_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ w,
                    __nv_bfloat16* __restrict__ y, const DwGeom g) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  constexpr int SPAN = (T - 1) * S + 2 * D + 1;
  const int G = g.C >> 3;
  const int strips = (g.Wo + T - 1) / T;
  const long total = long(g.N) * g.Ho * strips * G;
  for (long i = long(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += long(gridDim.x) * blockDim.x) {
    const int cg = int(i % G);
    long r = i / G;
    const int sx = int(r % strips); r /= strips;
    const int oy = int(r % g.Ho);
    const int n = int(r / g.Ho);
    const int ox0 = sx * T;
    float wr[9][8];
#pragma unroll
    for (int k = 0; k < 9; ++k) {
      const int kk = REVERSE ? 8 - k : k;
      const float4 w0 = *reinterpret_cast<const float4*>(w + kk * g.C + cg * 8);
      const float4 w1 = *reinterpret_cast<const float4*>(w + kk * g.C + cg * 8 + 4);
      wr[k][0] = w0.x; wr[k][1] = w0.y; wr[k][2] = w0.z; wr[k][3] = w0.w;
      wr[k][4] = w1.x; wr[k][5] = w1.y; wr[k][6] = w1.z; wr[k][7] = w1.w;
    }
    float acc[T][8];
#pragma unroll
    for (int j = 0; j < T; ++j)
#pragma unroll
      for (int c = 0; c < 8; ++c) acc[j][c] = 0.f;
#pragma unroll
    for (int ky = 0; ky < 3; ++ky) {
      const int iy = oy * S + (ky - 1) * D;
      if (iy < 0 || iy >= g.H) continue;
      const __nv_bfloat16* row = x + (size_t(size_t(n) * g.H + iy) * g.W) * g.ldx + cg * 8;
#pragma unroll
      for (int c = 0; c < SPAN; ++c) {
        const int ix = ox0 * S - D + c;
        if (ix < 0 || ix >= g.W) continue;
        float xv[8];
        dw_unpack8(*reinterpret_cast<const uint4*>(row + size_t(ix) * g.ldx), xv);
#pragma unroll
        for (int j = 0; j < T; ++j)
#pragma unroll
          for (int kx = 0; kx < 3; ++kx)
            if (j * S + kx * D == c) {
#pragma unroll
              for (int e = 0; e < 8; ++e) acc[j][e] = fmaf(xv[e], wr[ky * 3 + kx][e], acc[j][e]);
            }
      }
    }
#pragma unroll
    for (int j = 0; j < T; ++j)
      if (ox0 + j < g.Wo)
        *reinterpret_cast<uint4*>(y + (size_t(size_t(n) * g.Ho + oy) * g.Wo + ox0 + j) * g.ldy + cg * 8) = dw_pack8(acc[j]);
  }
}

// ---------------------------------------------------------------- column-marching variants
// stride 1, dilation 1 (57 of the 63 depthwise layers of a trunk).  One thread = 4 channels x 4
// output columns, marching down a segment of rows.  Every input row is loaded once (6 column
// vectors, issued together and one row ahead of the arithmetic so that DRAM latency overlaps
// the FMAs) and scattered into three rolling accumulator rows: 1.5 loads per output vector
// instead of 9, weights in registers for the whole march, index arithmetic once per thread.
constexpr int kCpt = 4;  // channels per thread

__device__ __forceinline__ void dw_unpack4(const uint2& u, float (&f)[4]) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
  const float2 a = __bfloat1622float2(h[0]), b = __bfloat1622float2(h[1]);
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y;
}
__device__ __forceinline__ uint2 dw_pack4(const float (&f)[4]) {
  uint2 u;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
  h[0] = __floats2bfloat162_rn(f[0], f[1]);
  h[1] = __floats2bfloat162_rn(f[2], f[3]);
  return u;
}

struct MarchItem {
  int ox0, oy0, oy1, n;
  int H, W;      // extent of the (sub-)grid this item marches over
  int off;       // pixel offset of the sub-grid origin inside the image (dilation parity)
};
// Work decomposition of the marching kernels.  A block is 4 warps that share one block of 32 channel
// groups (128 channels, lane = group); each warp loops over items = (image, dilation parity, row
// segment, 4-column strip).  A dilation-d conv is d*d independent dilation-1 convs on the parity
// sub-grids (rows py, py+d, ...; columns px, px+d, ...), so the same march serves every dilation
// with the row / column strides multiplied by d.  Weights / reduction registers live for the whole
// loop, and the weight-gradient kernel flushes 9 x 128 values per block instead of 9 x C.
struct MarchSched {
  int segs, RY, cblocks, items, dil;
};
__device__ __forceinline__ void march_decode(const DwGeom& g, const MarchSched& s, int item, MarchItem* it) {
  const int Ws = (g.W + s.dil - 1) / s.dil;
  const int strips = (Ws + 3) >> 2;
  it->ox0 = (item % strips) * 4;
  int t = item / strips;
  it->oy0 = (t % s.segs) * s.RY;
  t /= s.segs;
  const int par = t % (s.dil * s.dil);
  it->n = t / (s.dil * s.dil);
  const int py = par / s.dil, px = par - py * s.dil;
  it->H = (g.H - py + s.dil - 1) / s.dil;
  it->W = (g.W - px + s.dil - 1) / s.dil;
  it->off = py * g.W + px;
  it->oy1 = min(it->oy0 + s.RY, it->H);
}

// 4 bf16 channels -> two packed fp32 pairs (bf16 -> fp32 is a 16-bit shift)
__device__ __forceinline__ void dw_unpack22(const uint2& u, float2 (&f)[2]) {
  f[0] = make_float2(__uint_as_float(u.x << 16), __uint_as_float(u.x & 0xffff0000u));
  f[1] = make_float2(__uint_as_float(u.y << 16), __uint_as_float(u.y & 0xffff0000u));
}
__device__ __forceinline__ uint2 dw_pack22(const float2 (&f)[2]) {
  uint2 u;
  __nv_bfloat162 lo = __floats2bfloat162_rn(f[0].x, f[0].y), hi = __floats2bfloat162_rn(f[1].x, f[1].y);
  u.x = *reinterpret_cast<uint32_t*>(&lo);
  u.y = *reinterpret_cast<uint32_t*>(&hi);
  return u;
}

// Input transform fused into the load: XF 0 = none, 1 = a*x + b (BatchNorm with batch statistics
// folded into a, b), 2 = LeakyReLU(a*x + b).  Pixels outside the image stay exactly zero (the
// reference pads the *activated* tensor, deeplab.py:66-72).
struct XfCoef {
  float2 a[2], b[2];
  float slope;
};
template <int XF>
__device__ __forceinline__ void dw_transform(float2 (&v)[2], const XfCoef& t, bool inside) {
  if (XF == 0) return;
#pragma unroll
  for (int p = 0; p < 2; ++p) {
    float2 z = __ffma2_rn(t.a[p], v[p], t.b[p]);
    if (XF == 2) {
      const float2 sz = __fmul2_rn(make_float2(t.slope, t.slope), z);
      z.x = fmaxf(z.x, sz.x);
      z.y = fmaxf(z.y, sz.y);
    }
    v[p] = inside ? z : make_float2(0.f, 0.f);
  }
}
template <int XF>
__device__ __forceinline__ void dw_load_coef(XfCoef* t, const float* xa, const float* xb, float slope, int c0) {
  if (XF == 0) return;
  const float4 a = *reinterpret_cast<const float4*>(xa + c0), b = *reinterpret_cast<const float4*>(xb + c0);
  t->a[0] = make_float2(a.x, a.y); t->a[1] = make_float2(a.z, a.w);
  t->b[0] = make_float2(b.x, b.y); t->b[1] = make_float2(b.z, b.w);
  t->slope = slope;
}

// Packed FFMA2 (same rounding as two scalar FMAs: results are bit-identical).  Round 1 measured it as no gain while
// the marches were bound by their ramp; with the cost-model planner they are issue-bound at the mid-flow shape (ncu:
// 49 % issue-active at 12 warps per SM, profiles/r02_dw_march_mid_flow_ncu.txt) and 864 FFMA -> 432 FFMA2 removes 16 %
// of the kernel's instructions: 1-2 % per kernel at every trunk shape, 0.3 ms of the step
// (profiles/r02_dw_ffma2_ab.txt).  -DBGR_DW_FFMA2=0 restores the scalar form.
__device__ __forceinline__ float2 fma2s(const float2& a, const float2& b, const float2& c) {
#if BGR_DW_FFMA2
  return __ffma2_rn(a, b, c);
#else
  return make_float2(fmaf(a.x, b.x, c.x), fmaf(a.y, b.y, c.y));
#endif
}

// ---- asynchronous row ring.  The march is bound by memory-level parallelism (measured: 168
// registers -> 12 warps per SM x one prefetched row = 18 KB in flight per SM = 2.7 TB/s), so rows are
// fetched kRing ahead with cp.async into a per-warp shared-memory ring: no registers are held by
// loads in flight and every lane reads back only the 8-byte slots it copied itself (no barriers).
constexpr int kRing = 4;
__device__ __forceinline__ void cp_async8(void* smem, const void* gmem, bool valid) {
  const unsigned sa = static_cast<unsigned>(__cvta_generic_to_shared(smem));
  const int sz = valid ? 8 : 0;   // 0 source bytes = zero fill
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" ::"r"(sa), "l"(gmem), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// issue the six column vectors of input row iy into ring slot `slot` ([col][lane] x 8 bytes).
// 32-bit element offsets from the item's base pointer (one image is < 2^31 elements); INTERIOR =
// all six columns of the strip lie inside the image, so only the row needs a validity test.
template <bool INTERIOR>
__device__ __forceinline__ void march_issue_row(uint2* slot, const __nv_bfloat16* base, int H, int rowstride,
                                                int colbase, int ld, int iy, unsigned cmask, bool in_range) {
  const bool rv = in_range && iy >= 0 && iy < H;
  const int ro = (rv ? iy : 0) * rowstride + colbase;
#pragma unroll
  for (int c = 0; c < 6; ++c) {
    const bool ok = INTERIOR ? rv : (rv && ((cmask >> c) & 1u));
    const int off = INTERIOR ? ro + c * ld : (ok ? ro + c * ld : 0);
    cp_async8(slot + c * 32, base + off, ok);
  }
}
// column validity bits of a strip (bit c <-> column ox0 - 1 + c)
__device__ __forceinline__ unsigned march_col_mask(int W, int ox0) {
  unsigned m = 0;
#pragma unroll
  for (int c = 0; c < 6; ++c) {
    const int ix = ox0 - 1 + c;
    if (ix >= 0 && ix < W) m |= 1u << c;
  }
  return m;
}

// one item (image, row segment, 4-column strip) of the forward / data-gradient march
template <bool INTERIOR, int XF>
__device__ __forceinline__ void march_fwd_item(const __nv_bfloat16* __restrict__ xin, __nv_bfloat16* __restrict__ yout,
                                               const DwGeom& g, int dil, const MarchItem& it,
                                               const float2 (&wr)[9][2], const XfCoef& tc, uint2* ring) {
  float2 a0[4][2], a1[4][2], a2[4][2];
#pragma unroll
  for (int j = 0; j < 4; ++j)
#pragma unroll
    for (int e = 0; e < 2; ++e) a0[j][e] = a1[j][e] = a2[j][e] = make_float2(0.f, 0.f);
  const unsigned cmask = INTERIOR ? 0x3fu : march_col_mask(it.W, it.ox0);
  const int first = it.oy0 - 1;   // rows first .. oy1 are consumed, one per phase
  const int cstep = g.ldx * dil, rowstride = g.W * cstep, colbase = (it.ox0 - 1) * cstep;
  const int ostep = g.ldy * dil, orowstride = g.W * ostep, ocol = it.ox0 * ostep;
#pragma unroll
  for (int r = 0; r < kRing; ++r) {
    march_issue_row<INTERIOR>(ring + r * (6 * 32), xin, it.H, rowstride, colbase, cstep, first + r, cmask,
                              first + r <= it.oy1);
    cp_async_commit();
  }
  // input row iy adds to output rows iy+1 (ky=0), iy (ky=1), iy-1 (ky=2); then row iy-1 is complete
  auto phase = [&](int iy, float2 (&prev)[4][2], float2 (&cur)[4][2], float2 (&next)[4][2]) {
    cp_async_wait<kRing - 1>();   // the oldest outstanding group is row iy
    uint2* slot = ring + ((iy - first) & (kRing - 1)) * (6 * 32);
    uint2 rc[6];
#pragma unroll
    for (int c = 0; c < 6; ++c) rc[c] = slot[c * 32];
    march_issue_row<INTERIOR>(slot, xin, it.H, rowstride, colbase, cstep, iy + kRing, cmask,
                              iy + kRing <= it.oy1);   // refill the slot
    cp_async_commit();
    // Padding is zero AFTER the transform (the reference pads the activated tensor).  For an interior strip that is a
    // per-row condition: a row outside the image contributes nothing and is skipped, every other row needs no
    // per-element select at all (the marches are issue-bound: the selects were a quarter of the transform).
    const bool rowok = iy >= 0 && iy < it.H;
    const unsigned mc = rowok ? cmask : 0u;
    if (!(INTERIOR && XF != 0) || rowok) {
#pragma unroll
    for (int c = 0; c < 6; ++c) {
      float2 xv[2];
      dw_unpack22(rc[c], xv);
      dw_transform<XF>(xv, tc, (INTERIOR && XF != 0) ? true : bool((mc >> c) & 1u));
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int kx = c - j;
        if (kx < 0 || kx > 2) continue;
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          next[j][e] = fma2s(xv[e], wr[0 + kx][e], next[j][e]);
          cur[j][e] = fma2s(xv[e], wr[3 + kx][e], cur[j][e]);
          prev[j][e] = fma2s(xv[e], wr[6 + kx][e], prev[j][e]);
        }
      }
    }
    }
    const int oy = iy - 1;
    if (oy >= it.oy0 && oy < it.oy1) {
      __nv_bfloat16* orow = yout + (oy * orowstride + ocol);
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (INTERIOR || it.ox0 + j < it.W) *reinterpret_cast<uint2*>(orow + j * ostep) = dw_pack22(prev[j]);
    }
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int e = 0; e < 2; ++e) prev[j][e] = make_float2(0.f, 0.f);
  };
  for (int iy = first; iy <= it.oy1; iy += 3) {
    phase(iy, a0, a1, a2);
    if (iy + 1 <= it.oy1) phase(iy + 1, a1, a2, a0);
    if (iy + 2 <= it.oy1) phase(iy + 2, a2, a0, a1);
  }
  cp_async_wait<0>();   // drain the (empty) tail groups before the ring is reused by the next item
}

template <bool REVERSE, int XF>
__global__ void __launch_bounds__(128, 3)
dw_march_fwd_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ w,
                    __nv_bfloat16* __restrict__ y, const DwGeom g, const MarchSched sc,
                    const float* __restrict__ xa, const float* __restrict__ xb, float slope) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  __shared__ uint2 ring_all[4][kRing][6 * 32];   // [warp][slot][col][lane]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c0 = ((blockIdx.x % sc.cblocks) * 32 + lane) * kCpt;
  if (c0 >= g.C) return;
  uint2* ring = &ring_all[warp][0][lane];
  float2 wr[9][2];
#pragma unroll
  for (int k = 0; k < 9; ++k) {
    const float4 w0 = *reinterpret_cast<const float4*>(w + (REVERSE ? 8 - k : k) * g.C + c0);
    wr[k][0] = make_float2(w0.x, w0.y);
    wr[k][1] = make_float2(w0.z, w0.w);
  }
  XfCoef tc;
  dw_load_coef<XF>(&tc, xa, xb, slope, c0);
  const int item_step = (gridDim.x / sc.cblocks) * 4;
  for (int item = (blockIdx.x / sc.cblocks) * 4 + warp; item < sc.items; item += item_step) {
    MarchItem it;
    march_decode(g, sc, item, &it);
    const __nv_bfloat16* xin = x + (size_t(it.n) * g.H * g.W + it.off) * g.ldx + c0;
    __nv_bfloat16* yout = y + (size_t(it.n) * g.H * g.W + it.off) * g.ldy + c0;
    if (it.ox0 >= 1 && it.ox0 + 5 <= it.W)   // warp-uniform: columns ox0-1 .. ox0+4 all inside
      march_fwd_item<true, XF>(xin, yout, g, sc.dil, it, wr, tc, ring);
    else
      march_fwd_item<false, XF>(xin, yout, g, sc.dil, it, wr, tc, ring);
  }
}

// dw[ky][kx][c] += sum dy[oy][ox] * x[oy+ky-1][ox+kx-1]; marching over input rows with the three
// neighbouring dy rows kept (packed) in registers.  Rows of dy outside the segment count as zero.
// XF: x is LeakyReLU/affine of the stored tensor, recomputed on load (see dw_transform).
// ---------------------------------------------------------------- stride-2 marches (dilation 1)
// y[oy][ox] = sum w[ky][kx] x[2oy+ky-1][2ox+kx-1].  Lane = 4 channels, warp = 128 channels x a strip of
// 4 output columns (9 input columns 2ox0-1 .. 2ox0+7) marching down the input rows 2oy0-1 .. 2oy1-1:
// an even row 2o is the middle tap row of output row o, an odd row 2o+1 is the bottom tap row of o
// and the top tap row of o+1.  Items = (image, output-row segment, output strip); same cp.async ring.
__device__ __forceinline__ void march_decode_s2(const DwGeom& g, const MarchSched& s, int item, MarchItem* it) {
  const int strips = (g.Wo + 3) >> 2;
  it->ox0 = (item % strips) * 4;
  const int t = item / strips;
  it->oy0 = (t % s.segs) * s.RY;
  it->n = t / s.segs;
  it->H = g.H; it->W = g.W; it->off = 0;
  it->oy1 = min(it->oy0 + s.RY, g.Ho);
}
// nine column vectors of input row iy into ring slot `slot` ([col][lane] x 8 bytes)
template <bool INTERIOR>
__device__ __forceinline__ void march_issue_row9(uint2* slot, const __nv_bfloat16* base, int H, int W, int rowstride,
                                                 int ix0, int ld, int iy, bool in_range) {
  const bool rv = in_range && iy >= 0 && iy < H;
  const int ro = (rv ? iy : 0) * rowstride + ix0 * ld;
#pragma unroll
  for (int c = 0; c < 9; ++c) {
    const bool ok = INTERIOR ? rv : (rv && ix0 + c >= 0 && ix0 + c < W);
    cp_async8(slot + c * 32, base + (INTERIOR || ok ? ro + c * ld : 0), ok);
  }
}

template <bool INTERIOR>
__device__ __forceinline__ void march_fwd_s2_item(const __nv_bfloat16* __restrict__ xin, __nv_bfloat16* __restrict__ yout,
                                                  const DwGeom& g, const MarchItem& it, const float2 (&wr)[9][2],
                                                  uint2* ring) {
  const int first = 2 * it.oy0 - 1, last = 2 * it.oy1 - 1;   // input rows consumed, one per phase
  const int rowstride = g.W * g.ldx, ix0 = 2 * it.ox0 - 1;
  const int orowstride = g.Wo * g.ldy, ocol = it.ox0 * g.ldy;
  float2 cur[4][2];
#pragma unroll
  for (int j = 0; j < 4; ++j) cur[j][0] = cur[j][1] = make_float2(0.f, 0.f);
#pragma unroll
  for (int r = 0; r < kRing; ++r) {
    march_issue_row9<INTERIOR>(ring + r * (9 * 32), xin, g.H, g.W, rowstride, ix0, g.ldx, first + r, first + r <= last);
    cp_async_commit();
  }
  auto fetch = [&](int iy, uint2 (&rc)[9]) {
    cp_async_wait<kRing - 1>();
    uint2* slot = ring + ((iy - first) & (kRing - 1)) * (9 * 32);
#pragma unroll
    for (int c = 0; c < 9; ++c) rc[c] = slot[c * 32];
    march_issue_row9<INTERIOR>(slot, xin, g.H, g.W, rowstride, ix0, g.ldx, iy + kRing, iy + kRing <= last);
    cp_async_commit();
  };
  // odd row 2o+1: finishes output row o (bottom taps), starts output row o+1 (top taps)
  auto phase_odd = [&](int iy, bool store) {
    uint2 rc[9];
    fetch(iy, rc);
    float2 nxt[4][2];
#pragma unroll
    for (int j = 0; j < 4; ++j) nxt[j][0] = nxt[j][1] = make_float2(0.f, 0.f);
#pragma unroll
    for (int c = 0; c < 9; ++c) {
      float2 xv[2];
      dw_unpack22(rc[c], xv);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int kx = c - 2 * j;
        if (kx < 0 || kx > 2) continue;
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          cur[j][e] = fma2s(xv[e], wr[6 + kx][e], cur[j][e]);
          nxt[j][e] = fma2s(xv[e], wr[0 + kx][e], nxt[j][e]);
        }
      }
    }
    if (store) {
      __nv_bfloat16* orow = yout + (((iy - 1) >> 1) * orowstride + ocol);
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (INTERIOR || it.ox0 + j < g.Wo) *reinterpret_cast<uint2*>(orow + j * g.ldy) = dw_pack22(cur[j]);
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) { cur[j][0] = nxt[j][0]; cur[j][1] = nxt[j][1]; }
  };
  auto phase_even = [&](int iy) {
    uint2 rc[9];
    fetch(iy, rc);
#pragma unroll
    for (int c = 0; c < 9; ++c) {
      float2 xv[2];
      dw_unpack22(rc[c], xv);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int kx = c - 2 * j;
        if (kx < 0 || kx > 2) continue;
#pragma unroll
        for (int e = 0; e < 2; ++e) cur[j][e] = fma2s(xv[e], wr[3 + kx][e], cur[j][e]);
      }
    }
  };
  phase_odd(first, false);
  for (int o = it.oy0; o < it.oy1; ++o) {
    phase_even(2 * o);
    phase_odd(2 * o + 1, true);
  }
  cp_async_wait<0>();
}

__global__ void __launch_bounds__(128, 3)
dw_march_fwd_s2_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ w,
                       __nv_bfloat16* __restrict__ y, const DwGeom g, const MarchSched sc) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  __shared__ uint2 ring_all[4][kRing][9 * 32];   // [warp][slot][col][lane]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c0 = ((blockIdx.x % sc.cblocks) * 32 + lane) * kCpt;
  if (c0 >= g.C) return;
  uint2* ring = &ring_all[warp][0][lane];
  float2 wr[9][2];
#pragma unroll
  for (int k = 0; k < 9; ++k) {
    const float4 w0 = *reinterpret_cast<const float4*>(w + k * g.C + c0);
    wr[k][0] = make_float2(w0.x, w0.y);
    wr[k][1] = make_float2(w0.z, w0.w);
  }
  const int item_step = (gridDim.x / sc.cblocks) * 4;
  for (int item = (blockIdx.x / sc.cblocks) * 4 + warp; item < sc.items; item += item_step) {
    MarchItem it;
    march_decode_s2(g, sc, item, &it);
    const __nv_bfloat16* xin = x + size_t(it.n) * g.H * g.W * g.ldx + c0;
    __nv_bfloat16* yout = y + size_t(it.n) * g.Ho * g.Wo * g.ldy + c0;
    // warp-uniform: input columns 2ox0-1 .. 2ox0+7 and output columns ox0 .. ox0+3 all inside
    if (it.ox0 >= 1 && 2 * it.ox0 + 7 < g.W && it.ox0 + 3 < g.Wo)
      march_fwd_s2_item<true>(xin, yout, g, it, wr, ring);
    else
      march_fwd_s2_item<false>(xin, yout, g, it, wr, ring);
  }
}

// Stride-2 data gradient.  x[r][s] feeds y[oy][ox] with ky = r - 2oy + 1, kx = s - 2ox + 1, so the 2x2
// block dx[2i+p][2j+q] depends on gy[i..i+1][j..j+1] only:
//   (0,0) w11 g00                         (0,1) w12 g00 + w10 g01
//   (1,0) w21 g00 + w01 g10               (1,1) w22 g00 + w20 g01 + w02 g10 + w00 g11
// Thread = 8 channels x one low-resolution position: 4 loads (3 of them shared with neighbours through
// L1/L2), 4 stores, 9 FMAs per channel; straight-line body so that the loads of a thread pipeline.
__global__ void __launch_bounds__(256)
dw_dgrad_s2_kernel(const __nv_bfloat16* __restrict__ gy, const float* __restrict__ w,
                   __nv_bfloat16* __restrict__ dx, const DwGeom g) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  const int G = g.C >> 3;
  const int Hb = (g.H + 1) >> 1, Wb = (g.W + 1) >> 1;   // 2x2 blocks of dx (= Ho x Wo)
  const long total = long(g.N) * Hb * Wb * G;
  for (long t = long(blockIdx.x) * blockDim.x + threadIdx.x; t < total; t += long(gridDim.x) * blockDim.x) {
    const int cg = int(t % G);
    long r = t / G;
    const int j = int(r % Wb); r /= Wb;
    const int i = int(r % Hb);
    const int n = int(r / Hb);
    const bool vi = i + 1 < g.Ho, vj = j + 1 < g.Wo;
    const __nv_bfloat16* gb = gy + ((size_t(n) * g.Ho + i) * g.Wo + j) * g.ldy + cg * 8;
    const uint4 z = make_uint4(0, 0, 0, 0);
    const uint4 u00 = *reinterpret_cast<const uint4*>(gb);
    const uint4 u01 = vj ? *reinterpret_cast<const uint4*>(gb + g.ldy) : z;
    const uint4 u10 = vi ? *reinterpret_cast<const uint4*>(gb + size_t(g.Wo) * g.ldy) : z;
    const uint4 u11 = (vi && vj) ? *reinterpret_cast<const uint4*>(gb + size_t(g.Wo + 1) * g.ldy) : z;
    float g00[8], g01[8], g10[8], g11[8], wk[9][8];
    dw_unpack8(u00, g00); dw_unpack8(u01, g01); dw_unpack8(u10, g10); dw_unpack8(u11, g11);
#pragma unroll
    for (int k = 0; k < 9; ++k) {
      const float4 a = *reinterpret_cast<const float4*>(w + k * g.C + cg * 8);
      const float4 b = *reinterpret_cast<const float4*>(w + k * g.C + cg * 8 + 4);
      wk[k][0] = a.x; wk[k][1] = a.y; wk[k][2] = a.z; wk[k][3] = a.w;
      wk[k][4] = b.x; wk[k][5] = b.y; wk[k][6] = b.z; wk[k][7] = b.w;
    }
    float o00[8], o01[8], o10[8], o11[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      o00[e] = wk[4][e] * g00[e];
      o01[e] = fmaf(wk[5][e], g00[e], wk[3][e] * g01[e]);
      o10[e] = fmaf(wk[7][e], g00[e], wk[1][e] * g10[e]);
      o11[e] = fmaf(wk[8][e], g00[e], fmaf(wk[6][e], g01[e], fmaf(wk[2][e], g10[e], wk[0][e] * g11[e])));
    }
    const int y0 = 2 * i, x0 = 2 * j;
    const bool vy = y0 + 1 < g.H, vx = x0 + 1 < g.W;
    __nv_bfloat16* ob = dx + ((size_t(n) * g.H + y0) * g.W + x0) * g.ldx + cg * 8;
    *reinterpret_cast<uint4*>(ob) = dw_pack8(o00);
    if (vx) *reinterpret_cast<uint4*>(ob + g.ldx) = dw_pack8(o01);
    if (vy) *reinterpret_cast<uint4*>(ob + size_t(g.W) * g.ldx) = dw_pack8(o10);
    if (vx && vy) *reinterpret_cast<uint4*>(ob + size_t(g.W + 1) * g.ldx) = dw_pack8(o11);
  }
}

// Stride-2 weight gradient: acc[ky][kx] += gy[o][j] * x[2o+ky-1][2j+kx-1], marching down the input
// rows like march_fwd_s2_item; the gy row o+1 travels in the ring group of input row 2o+1 (where it is
// first needed, as the row whose top taps that input row feeds).
template <bool INTERIOR>
__device__ __forceinline__ void march_wgrad_s2_item(const __nv_bfloat16* __restrict__ xin,
                                                    const __nv_bfloat16* __restrict__ gin, const DwGeom& g,
                                                    const MarchItem& it, float2 (&acc)[9][2], uint2* ring) {
  const int first = 2 * it.oy0 - 1, last = 2 * it.oy1 - 1;
  const int rowstride = g.W * g.ldx, ix0 = 2 * it.ox0 - 1;
  const int grow = g.Wo * g.ldy, gcol = it.ox0 * g.ldy;
  constexpr int kSlot = 13 * 32;   // 9 x columns + 4 gy columns
  auto issue = [&](uint2* slot, int r) {
    const bool in_range = r <= last;
    march_issue_row9<INTERIOR>(slot, xin, g.H, g.W, rowstride, ix0, g.ldx, r, in_range);
    const int o = (r + 1) >> 1;                       // odd r = 2o'-1 carries gy row o' = (r+1)/2
    const bool ok = in_range && (r & 1) && o >= it.oy0 && o < it.oy1;
    const int go = (ok ? o : 0) * grow + gcol;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const bool okj = INTERIOR ? ok : (ok && it.ox0 + j < g.Wo);
      cp_async8(slot + (9 + j) * 32, gin + (INTERIOR || okj ? go + j * g.ldy : 0), okj);
    }
    cp_async_commit();
  };
#pragma unroll
  for (int r = 0; r < kRing; ++r) issue(ring + r * kSlot, first + r);
  uint2 gc[4], gn[4];   // gy rows o (current) and o+1 (next), packed
#pragma unroll
  for (int j = 0; j < 4; ++j) gc[j] = gn[j] = make_uint2(0, 0);
  auto fetch = [&](int iy, uint2 (&rc)[9], bool take_g) {
    cp_async_wait<kRing - 1>();
    uint2* slot = ring + ((iy - first) & (kRing - 1)) * kSlot;
#pragma unroll
    for (int c = 0; c < 9; ++c) rc[c] = slot[c * 32];
    if (take_g) {
#pragma unroll
      for (int j = 0; j < 4; ++j) gn[j] = slot[(9 + j) * 32];
    }
    issue(slot, iy + kRing);
  };
  auto accumulate = [&](const uint2 (&rc)[9], const uint2 (&grow_)[4], int ky) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float2 gv[2];
      dw_unpack22(grow_[j], gv);
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) {
        float2 xv[2];
        dw_unpack22(rc[2 * j + kx], xv);
#pragma unroll
        for (int e = 0; e < 2; ++e) acc[ky * 3 + kx][e] = fma2s(gv[e], xv[e], acc[ky * 3 + kx][e]);
      }
    }
  };
  // odd row 2o+1: bottom taps (ky=2) of gy row o = gc, top taps (ky=0) of gy row o+1 = gn (just fetched)
  {
    uint2 rc[9];
    fetch(first, rc, true);          // row 2oy0-1: only the top taps of gy row oy0
    accumulate(rc, gn, 0);
#pragma unroll
    for (int j = 0; j < 4; ++j) gc[j] = gn[j];
  }
  for (int o = it.oy0; o < it.oy1; ++o) {
    uint2 rc[9];
    fetch(2 * o, rc, false);
    accumulate(rc, gc, 1);
    fetch(2 * o + 1, rc, true);
    accumulate(rc, gc, 2);
    accumulate(rc, gn, 0);
#pragma unroll
    for (int j = 0; j < 4; ++j) gc[j] = gn[j];
  }
  cp_async_wait<0>();
}

__global__ void __launch_bounds__(128, 2)
dw_march_wgrad_s2_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ dy,
                         float* __restrict__ ws, const DwGeom g, const MarchSched sc) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  __shared__ float sh[9 * 128];
  extern __shared__ uint2 ring_dyn[];               // [4 warps][kRing][13 * 32]
  for (int i = threadIdx.x; i < 9 * 128; i += blockDim.x) sh[i] = 0.f;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int cb0 = (blockIdx.x % sc.cblocks) * 128;
  const int lc = lane * kCpt;
  const int c0 = cb0 + lc;
  float2 acc[9][2];
#pragma unroll
  for (int k = 0; k < 9; ++k) acc[k][0] = acc[k][1] = make_float2(0.f, 0.f);
  if (c0 < g.C) {
    uint2* ring = ring_dyn + warp * (kRing * 13 * 32) + lane;
    const int item_step = (gridDim.x / sc.cblocks) * 4;
    for (int item = (blockIdx.x / sc.cblocks) * 4 + warp; item < sc.items; item += item_step) {
      MarchItem it;
      march_decode_s2(g, sc, item, &it);
      const __nv_bfloat16* xin = x + size_t(it.n) * g.H * g.W * g.ldx + c0;
      const __nv_bfloat16* gin = dy + size_t(it.n) * g.Ho * g.Wo * g.ldy + c0;
      if (it.ox0 >= 1 && 2 * it.ox0 + 7 < g.W && it.ox0 + 3 < g.Wo)
        march_wgrad_s2_item<true>(xin, gin, g, it, acc, ring);
      else
        march_wgrad_s2_item<false>(xin, gin, g, it, acc, ring);
    }
  }
  for (int turn = 0; turn < 4; ++turn) {   // the four warps add in warp order: schedule-independent
    if (turn == warp && c0 < g.C) {
#pragma unroll
      for (int k = 0; k < 9; ++k) {
        sh[k * 128 + lc + 0] += acc[k][0].x;
        sh[k * 128 + lc + 1] += acc[k][0].y;
        sh[k * 128 + lc + 2] += acc[k][1].x;
        sh[k * 128 + lc + 3] += acc[k][1].y;
      }
    }
    __syncthreads();
  }
  // partial row ws[blockIdx.x / cblocks][C][9] (the parameter's own layout): this block owns channels
  // [cb0, cb0 + 128); sum_rows (reduce.cu) adds the rows in block order
  const int nch = min(128, g.C - cb0);
  float* row = ws + size_t(blockIdx.x / sc.cblocks) * 9 * g.C + size_t(cb0) * 9;
  for (int i = threadIdx.x; i < 9 * nch; i += blockDim.x) row[i] = sh[(i % 9) * 128 + i / 9];
}

// one item of the weight-gradient march: acc[ky*3+kx] += dy[oy][ox] * x[oy+ky-1][ox+kx-1]
template <bool INTERIOR, int XF>
__device__ __forceinline__ void march_wgrad_item(const __nv_bfloat16* __restrict__ xin,
                                                 const __nv_bfloat16* __restrict__ gin, const DwGeom& g,
                                                 int dil, const MarchItem& it, float2 (&acc)[9][2],
                                                 const XfCoef& tc, uint2* ring) {
  const unsigned cmask = INTERIOR ? 0x3fu : march_col_mask(it.W, it.ox0);
  const int first = it.oy0 - 1;
  const int cstep = g.ldx * dil, rowstride = g.W * cstep, colbase = (it.ox0 - 1) * cstep;
  const int gstep = g.ldy * dil, grow = g.W * gstep, gcol = it.ox0 * gstep;
  // group r = input row r and the dy row r + 1 (first needed, as "next", in phase r); dy rows
  // outside the segment count as zero
  auto issue = [&](uint2* slot, int r) {
    const bool in_range = r <= it.oy1;
    march_issue_row<INTERIOR>(slot, xin, it.H, rowstride, colbase, cstep, r, cmask, in_range);
    const int oy = r + 1;
    const bool ok = in_range && oy >= it.oy0 && oy < it.oy1;
    const int go = (ok ? oy : 0) * grow + gcol;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const bool okj = INTERIOR ? ok : (ok && it.ox0 + j < it.W);
      cp_async8(slot + (6 + j) * 32, gin + (INTERIOR || okj ? go + j * gstep : 0), okj);
    }
    cp_async_commit();
  };
#pragma unroll
  for (int r = 0; r < kRing; ++r) issue(ring + r * (10 * 32), first + r);
  // at input row iy: prev / cur / next = dy rows iy-1 / iy / iy+1
  auto phase = [&](int iy, const uint2 (&prev)[4], const uint2 (&cur)[4], uint2 (&next)[4]) {
    cp_async_wait<kRing - 1>();
    uint2* slot = ring + ((iy - first) & (kRing - 1)) * (10 * 32);
    uint2 rc[6];
#pragma unroll
    for (int c = 0; c < 6; ++c) rc[c] = slot[c * 32];
#pragma unroll
    for (int j = 0; j < 4; ++j) next[j] = slot[(6 + j) * 32];
    issue(slot, iy + kRing);
    // Padding is zero AFTER the transform (the reference pads the activated tensor).  For an interior strip that is a
    // per-row condition: a row outside the image contributes nothing and is skipped, every other row needs no
    // per-element select at all (the marches are issue-bound: the selects were a quarter of the transform).
    const bool rowok = iy >= 0 && iy < it.H;
    const unsigned mc = rowok ? cmask : 0u;
    if (!(INTERIOR && XF != 0) || rowok) {
#pragma unroll
    for (int c = 0; c < 6; ++c) {
      float2 xv[2];
      dw_unpack22(rc[c], xv);
      dw_transform<XF>(xv, tc, (INTERIOR && XF != 0) ? true : bool((mc >> c) & 1u));
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int kx = c - j;
        if (kx < 0 || kx > 2) continue;
        float2 gp[2], gc[2], gn[2];
        dw_unpack22(prev[j], gp);  // oy = iy-1 -> ky = 2
        dw_unpack22(cur[j], gc);   // oy = iy   -> ky = 1
        dw_unpack22(next[j], gn);  // oy = iy+1 -> ky = 0
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          acc[0 + kx][e] = fma2s(gn[e], xv[e], acc[0 + kx][e]);
          acc[3 + kx][e] = fma2s(gc[e], xv[e], acc[3 + kx][e]);
          acc[6 + kx][e] = fma2s(gp[e], xv[e], acc[6 + kx][e]);
        }
      }
    }
    }
  };
  uint2 d0[4], d1[4], d2[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) d0[j] = d1[j] = d2[j] = make_uint2(0, 0);
  for (int iy = first; iy <= it.oy1; iy += 3) {
    phase(iy, d0, d1, d2);
    if (iy + 1 <= it.oy1) phase(iy + 1, d1, d2, d0);
    if (iy + 2 <= it.oy1) phase(iy + 2, d2, d0, d1);
  }
  cp_async_wait<0>();
}

template <int XF>
__global__ void __launch_bounds__(128, 3)
dw_march_wgrad_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ dy,
                      float* __restrict__ ws, const DwGeom g, const MarchSched sc,
                      const float* __restrict__ xa, const float* __restrict__ xb, float slope) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  __shared__ float sh[9 * 128];                       // [9][128 channels of this block]
  __shared__ uint2 ring_all[4][kRing][10 * 32];       // [warp][slot][6 x cols + 4 dy cols][lane]
  for (int i = threadIdx.x; i < 9 * 128; i += blockDim.x) sh[i] = 0.f;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int cb0 = (blockIdx.x % sc.cblocks) * 128;
  const int lc = lane * kCpt;      // channel offset inside the block's 128
  const int c0 = cb0 + lc;
  float2 acc[9][2];
#pragma unroll
  for (int k = 0; k < 9; ++k) acc[k][0] = acc[k][1] = make_float2(0.f, 0.f);
  if (c0 < g.C) {
    uint2* ring = &ring_all[warp][0][lane];
    XfCoef tc;
    dw_load_coef<XF>(&tc, xa, xb, slope, c0);
    const int item_step = (gridDim.x / sc.cblocks) * 4;
    for (int item = (blockIdx.x / sc.cblocks) * 4 + warp; item < sc.items; item += item_step) {
      MarchItem it;
      march_decode(g, sc, item, &it);
      const __nv_bfloat16* xin = x + (size_t(it.n) * g.H * g.W + it.off) * g.ldx + c0;
      const __nv_bfloat16* gin = dy + (size_t(it.n) * g.H * g.W + it.off) * g.ldy + c0;
      if (it.ox0 >= 1 && it.ox0 + 5 <= it.W)
        march_wgrad_item<true, XF>(xin, gin, g, sc.dil, it, acc, tc, ring);
      else
        march_wgrad_item<false, XF>(xin, gin, g, sc.dil, it, acc, tc, ring);
    }
  }
  for (int turn = 0; turn < 4; ++turn) {   // the four warps add in warp order: schedule-independent
    if (turn == warp && c0 < g.C) {
#pragma unroll
      for (int k = 0; k < 9; ++k) {
        sh[k * 128 + lc + 0] += acc[k][0].x;
        sh[k * 128 + lc + 1] += acc[k][0].y;
        sh[k * 128 + lc + 2] += acc[k][1].x;
        sh[k * 128 + lc + 3] += acc[k][1].y;
      }
    }
    __syncthreads();
  }
  // partial row ws[blockIdx.x / cblocks][C][9] (the parameter's own layout): this block owns channels
  // [cb0, cb0 + 128); sum_rows (reduce.cu) adds the rows in block order
  const int nch = min(128, g.C - cb0);
  float* row = ws + size_t(blockIdx.x / sc.cblocks) * 9 * g.C + size_t(cb0) * 9;
  for (int i = threadIdx.x; i < 9 * nch; i += blockDim.x) row[i] = sh[(i % 9) * 128 + i / 9];
}

// Row segmentation by a cost model.  The marches are issue-bound (ncu at the 48x72x728 mid-flow shape:
// 68 % issue-active, long-scoreboard stalls ~0), so a launch costs about
//   rounds x (RY + 2 halo rows + ~3 rows of ring fill / drain),   rounds = ceil(items / resident warps),
// and the segment count that minimises it wins: long segments when they still fill every warp once
// (mid flow: 2 segments of 24 rows, one item per warp, instead of 8 x 6 rows with 33 % halo work),
// many short rounds only when the image is large.  One wave of blocks: 3 resident blocks per SM.
// (Swept on B200 over forced segment lengths and fill fractions: the model is within ~6 % of the best
// forced plan at every trunk shape; the HBM-bound 384x576 shape is insensitive between 20 and 128 rows.)
static void march_plan(const DwGeom& g, MarchSched* sc, unsigned* blocks, int blocks_per_sm = 3) {
  const int sms = sm_count();
  const int dil = g.dil < 1 ? 1 : g.dil;
  const int Hs = (g.H + dil - 1) / dil, Ws = (g.W + dil - 1) / dil, NP = g.N * dil * dil;
  const int G = g.C / kCpt, strips = (Ws + 3) / 4;
  sc->dil = dil;
  sc->cblocks = (G + 31) / 32;
  int pb_max = sms * blocks_per_sm / sc->cblocks;
  if (pb_max < 1) pb_max = 1;
  const long warps = 4L * pb_max, per_seg = long(NP) * strips;
  long best_cost = -1;
  int best_ry = Hs;
  for (int segs = 1; segs <= Hs; ++segs) {
    const int ry = (Hs + segs - 1) / segs;
    if (segs > 1 && ry == (Hs + segs - 2) / (segs - 1)) continue;   // same RY as the previous candidate
    const long items = per_seg * ((Hs + ry - 1) / ry);
    const long cost = ((items + warps - 1) / warps) * (ry + 5);
    if (best_cost < 0 || cost < best_cost) { best_cost = cost; best_ry = ry; }
    if (ry <= 2) break;
  }
  sc->RY = best_ry;
  sc->segs = (Hs + sc->RY - 1) / sc->RY;
  sc->items = NP * sc->segs * strips;
  const int ipw = int((sc->items + warps - 1) / warps);   // items per warp
  const int pb = (sc->items + 4 * ipw - 1) / (4 * ipw);
  *blocks = unsigned(pb) * sc->cblocks;
}

// plan of the stride-2 marches: items over the *output* grid
static void march_plan_s2(const DwGeom& g, MarchSched* sc, unsigned* blocks, int blocks_per_sm = 3) {
  DwGeom o = g;
  o.H = g.Ho; o.W = g.Wo; o.dil = 1;
  march_plan(o, sc, blocks, blocks_per_sm);
}

static int ew_grid(long total) {
  long blocks = (total + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  return int(blocks < 1 ? 1 : blocks);
}

static int make_geom(DwGeom* g, int N, int H, int W, int C, int stride, int dil, int ldx, int ldy) {
  if (C % 8 || C < 8) return fail(BGR_ERR_ARG, "depthwise: channel count must be a multiple of 8");
  if (stride != 1 && stride != 2) return fail(BGR_ERR_ARG, "depthwise: stride must be 1 or 2");
  g->N = N; g->H = H; g->W = W; g->C = C;
  g->Ho = (H - 1) / stride + 1;
  g->Wo = (W - 1) / stride + 1;
  g->stride = stride; g->dil = dil; g->ldx = ldx; g->ldy = ldy;
  return BGR_OK;
}

}  // namespace bgr

using namespace bgr;

constexpr int kStripT = 4;

// strip-mined forward for the (stride, dilation) pairs of the os=16 trunk; generic kernel otherwise
template <bool REVERSE>
static bool launch_fwd_strip(const __nv_bfloat16* x, const float* w, __nv_bfloat16* y, const DwGeom& g,
                             cudaStream_t st) {
  const long total = long(g.N) * g.Ho * ((g.Wo + kStripT - 1) / kStripT) * (g.C / 8);
  const int grid = ew_grid(total);
  if (g.stride == 1 && g.dil == 1) BGR_CUDA_OK(launch_pdl(dw_fwd_strip_kernel<1, 1, kStripT, REVERSE>, dim3(grid), dim3(256), 0, st, x, w, y, g));
  else if (g.stride == 1 && g.dil == 2) BGR_CUDA_OK(launch_pdl(dw_fwd_strip_kernel<1, 2, kStripT, REVERSE>, dim3(grid), dim3(256), 0, st, x, w, y, g));
  else if (g.stride == 2 && g.dil == 1 && !REVERSE) BGR_CUDA_OK(launch_pdl(dw_fwd_strip_kernel<2, 1, kStripT, false>, dim3(grid), dim3(256), 0, st, x, w, y, g));
  else return false;
  return true;
}

extern "C" int bgr_dw_fwd(const void* x, int N, int H, int W, int C, int ldx, const float* w, int stride,
                          int dil, void* y, int ldy, void* stream) {
  DwGeom g;
  if (int rc = make_geom(&g, N, H, W, C, stride, dil, ldx, ldy)) return rc;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  auto X = static_cast<const __nv_bfloat16*>(x);
  auto Y = static_cast<__nv_bfloat16*>(y);
  if (stride == 1) {   // any dilation: parity sub-grids
    MarchSched sc;
    unsigned blocks;
    march_plan(g, &sc, &blocks);
    BGR_CUDA_OK(launch_pdl(dw_march_fwd_kernel<false, 0>, dim3(blocks), dim3(128), 0, st, X, w, Y, g, sc, nullptr, nullptr, 0.f));
  } else if (stride == 2 && dil == 1) {
    MarchSched sc;
    unsigned blocks;
    march_plan_s2(g, &sc, &blocks);
    BGR_CUDA_OK(launch_pdl(dw_march_fwd_s2_kernel, dim3(blocks), dim3(128), 0, st, X, w, Y, g, sc));
  } else if (!launch_fwd_strip<false>(X, w, Y, g, st))
    BGR_CUDA_OK(launch_pdl(dw_fwd_kernel, dim3(ew_grid(long(N) * g.Ho * g.Wo * (C / 8))), dim3(256), 0, st, X, w, Y, g));
  BGR_LAUNCH_OK();
  return BGR_OK;
}

extern "C" int bgr_dw_dgrad(const void* dy, int N, int H, int W, int C, int ldy, const float* w, int stride,
                            int dil, void* dx, int ldx, void* stream) {
  DwGeom g;
  if (int rc = make_geom(&g, N, H, W, C, stride, dil, ldx, ldy)) return rc;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  auto DY = static_cast<const __nv_bfloat16*>(dy);
  auto DX = static_cast<__nv_bfloat16*>(dx);
  bool done = false;
  if (stride == 1) {
    // stride-1 data gradient = the same stencil with mirrored taps, reading dy and writing dx
    DwGeom r = g;
    r.ldx = ldy;
    r.ldy = ldx;
    MarchSched sc;
    unsigned blocks;
    march_plan(r, &sc, &blocks);
    BGR_CUDA_OK(launch_pdl(dw_march_fwd_kernel<true, 0>, dim3(blocks), dim3(128), 0, st, DY, w, DX, r, sc, nullptr, nullptr, 0.f));
    done = true;
  }
  if (!done && stride == 2 && dil == 1) {
    BGR_CUDA_OK(launch_pdl(dw_dgrad_s2_kernel, dim3(ew_grid(long(N) * ((H + 1) / 2) * ((W + 1) / 2) * (C / 8))), dim3(256), 0, st, DY, w, DX, g));
    done = true;
  }
  if (!done) BGR_CUDA_OK(launch_pdl(dw_dgrad_kernel, dim3(ew_grid(long(N) * H * W * (C / 8))), dim3(256), 0, st, DY, w, DX, g));
  BGR_LAUNCH_OK();
  return BGR_OK;
}

// partial rows of a weight-gradient launch -> dw[C][9], added in row order
static int dw_wgrad_finish(const float* ws, int rows, int C, float* dw, cudaStream_t st) {
  return launch_sum_rows(ws, rows, 9L * C, dw, st);
}
static int dw_wgrad_ws_check(long rows, int C) {
  if (rows * 9L * C > long(BGR_DW_WGRAD_WS_FLOATS))
    return fail(BGR_ERR_ARG, "depthwise wgrad: workspace of BGR_DW_WGRAD_WS_FLOATS floats is too small for this launch");
  return BGR_OK;
}

template <int XF>
static int launch_march_wgrad(const __nv_bfloat16* x, const __nv_bfloat16* dy, float* dw, float* ws, const DwGeom& g,
                              const float* xa, const float* xb, float slope, cudaStream_t st) {
  MarchSched sc;
  unsigned blocks;
  march_plan(g, &sc, &blocks);
  const int rows = int(blocks) / sc.cblocks;
  if (int rc = dw_wgrad_ws_check(rows, g.C)) return rc;
  BGR_CUDA_OK(launch_pdl(dw_march_wgrad_kernel<XF>, dim3(blocks), dim3(128), 0, st, x, dy, ws, g, sc, xa, xb, slope));
  BGR_LAUNCH_OK();
  return dw_wgrad_finish(ws, rows, g.C, dw, st);
}

// Stride-1, dilation-1 depthwise conv of act(a*x + b) without materialising it: the BatchNorm
// (+LeakyReLU) that precedes the depthwise conv of every SeparableConv2d_same inside a block
// (deeplab.py:100-118) is applied while loading.  act: 0 = affine only, 1 = LeakyReLU(slope).
extern "C" int bgr_dw_fwd_bn(const void* x, int N, int H, int W, int C, int ldx, const float* w,
                             const float* a, const float* b, int act, float slope, void* y, int ldy,
                             void* stream) {
  DwGeom g;
  if (int rc = make_geom(&g, N, H, W, C, 1, 1, ldx, ldy)) return rc;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  auto X = static_cast<const __nv_bfloat16*>(x);
  auto Y = static_cast<__nv_bfloat16*>(y);
  MarchSched sc;
  unsigned blocks;
  march_plan(g, &sc, &blocks);
  if (act) BGR_CUDA_OK(launch_pdl(dw_march_fwd_kernel<false, 2>, dim3(blocks), dim3(128), 0, st, X, w, Y, g, sc, a, b, slope));
  else BGR_CUDA_OK(launch_pdl(dw_march_fwd_kernel<false, 1>, dim3(blocks), dim3(128), 0, st, X, w, Y, g, sc, a, b, slope));
  BGR_LAUNCH_OK();
  return BGR_OK;
}

// Weight gradient of bgr_dw_fwd_bn: x is the stored pre-BatchNorm tensor, transformed on load.
extern "C" int bgr_dw_wgrad_bn(const void* x, int N, int H, int W, int C, int ldx, const float* a,
                               const float* b, int act, float slope, const void* dy, int ldy, float* dw,
                               float* ws, void* stream) {
  DwGeom g;
  if (int rc = make_geom(&g, N, H, W, C, 1, 1, ldx, ldy)) return rc;
  if (ws == nullptr) return fail(BGR_ERR_ARG, "depthwise wgrad: workspace missing");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  auto X = static_cast<const __nv_bfloat16*>(x);
  auto DY = static_cast<const __nv_bfloat16*>(dy);
  return act ? launch_march_wgrad<2>(X, DY, dw, ws, g, a, b, slope, st)
             : launch_march_wgrad<1>(X, DY, dw, ws, g, a, b, slope, st);
}

extern "C" int bgr_dw_wgrad(const void* x, int N, int H, int W, int C, int ldx, const void* dy, int ldy,
                            int stride, int dil, float* dw, float* ws, void* stream) {
  DwGeom g;
  if (int rc = make_geom(&g, N, H, W, C, stride, dil, ldx, ldy)) return rc;
  if (ws == nullptr) return fail(BGR_ERR_ARG, "depthwise wgrad: workspace missing");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  auto X = static_cast<const __nv_bfloat16*>(x);
  auto DY = static_cast<const __nv_bfloat16*>(dy);
  if (stride == 1) return launch_march_wgrad<0>(X, DY, dw, ws, g, nullptr, nullptr, 0.f, st);   // any dilation
  if (stride == 2 && dil == 1) {
    constexpr int ring_bytes = 4 * kRing * 13 * 32 * 8;   // 53 KB: above the 48 KB static limit
    if (int rc = ensure_dyn_smem(dw_march_wgrad_s2_kernel, ring_bytes)) return rc;
    MarchSched sc;
    unsigned blocks;
    march_plan_s2(g, &sc, &blocks, 2);   // 192 registers: two resident blocks per SM
    const int rows = int(blocks) / sc.cblocks;
    if (int rc = dw_wgrad_ws_check(rows, C)) return rc;
    BGR_CUDA_OK(launch_pdl(dw_march_wgrad_s2_kernel, dim3(blocks), dim3(128), ring_bytes, st, X, DY, ws, g, sc));
    BGR_LAUNCH_OK();
    return dw_wgrad_finish(ws, rows, C, dw, st);
  }
  if (9 * C * sizeof(float) > 160 * 1024) return fail(BGR_ERR_ARG, "depthwise wgrad: too many channels");
  const long P = long(N) * g.Ho * g.Wo;
  const int G = C / 8, R = G >= 256 ? 1 : 256 / G;
  long blocks = (P + long(R) * 4 - 1) / (long(R) * 4);
  if (blocks > 148 * 4) blocks = 148 * 4;
  if (blocks > long(BGR_DW_WGRAD_WS_FLOATS) / (9L * C)) blocks = long(BGR_DW_WGRAD_WS_FLOATS) / (9L * C);
  if (blocks < 1) blocks = 1;
  const long ppb = (P + blocks - 1) / blocks;
  blocks = (P + ppb - 1) / ppb;
  if (int rc = dw_wgrad_ws_check(blocks, C)) return rc;
  if (int rc = ensure_dyn_smem(dw_wgrad_kernel, 160 * 1024)) return rc;
  BGR_CUDA_OK(launch_pdl(dw_wgrad_kernel, dim3(int(blocks)), dim3(256), 9 * C * sizeof(float), st, X, DY, ws, g, ppb));
  BGR_LAUNCH_OK();
  return dw_wgrad_finish(ws, int(blocks), C, dw, st);
}

extern "C" int bgr_subsample2(const void* x, int N, int H, int W, int C, int ldx, void* y, int ldy,
                              void* stream) {
  DwGeom g;
  if (int rc = make_geom(&g, N, H, W, C, 2, 1, ldx, ldy)) return rc;
  BGR_CUDA_OK(launch_pdl(subsample2_kernel, dim3(ew_grid(long(N) * g.Ho * g.Wo * (C / 8))), dim3(256), 0, reinterpret_cast<cudaStream_t>(stream), 
      static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(y), g));
  BGR_LAUNCH_OK();
  return BGR_OK;
}

extern "C" int bgr_subsample2_bwd(const void* dy, int N, int H, int W, int C, int ldy, void* dx, int ldx,
                                  void* stream) {
  DwGeom g;
  if (int rc = make_geom(&g, N, H, W, C, 2, 1, ldx, ldy)) return rc;
  BGR_CUDA_OK(launch_pdl(subsample2_bwd_kernel, dim3(ew_grid(long(N) * H * W * (C / 8))), dim3(256), 0, reinterpret_cast<cudaStream_t>(stream), 
      static_cast<const __nv_bfloat16*>(dy), static_cast<__nv_bfloat16*>(dx), g));
  BGR_LAUNCH_OK();
  return BGR_OK;
}
