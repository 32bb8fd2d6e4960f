// Spatial helpers of the DeepLabv3+ decoder / discriminator head on bf16 NHWC tensors:
//   bilinear align_corners=True up-sampling + adjoint   (deeplab.py:375-379,663)
//   global average pool + adjoint                         (deeplab.py:621)
//   flatten + Linear(->1) head fwd / bwd                   (deeplab_gan.py:32-36)
//   fused L1 loss + gradient                               (train_gan.py:146,281)
//   field layout conversions (2x2 space-to-depth bf16 <-> NCHW fp32)
// All HBM-bound, 128-bit accesses where the layout allows.
#include <cuda_bf16.h>

#include "bgr_internal.h"

namespace bgr {

__device__ __forceinline__ void sp_unpack8(const uint4& u, float (&f)[8]) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 t = __bfloat1622float2(h[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}
__device__ __forceinline__ uint4 sp_pack8(const float (&f)[8]) {
  uint4 u;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
  return u;
}

static int sp_grid(long total) {
  long blocks = (total + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  return int(blocks < 1 ? 1 : blocks);
}

// Source index/weights as ATen's upsample_bilinear2d with align_corners=True:
// scale = (in-1)/(out-1) (0 if out == 1), src = scale*dst, i0 = floor(src), l1 = src - i0.
__device__ __forceinline__ void bil_coord(int o, float scale, int in, int* i0, int* i1, float* l1) {
  const float s = scale * float(o);
  int a = int(s);
  if (a > in - 1) a = in - 1;
  *i0 = a;
  *i1 = a + (a < in - 1 ? 1 : 0);
  *l1 = s - float(a);
}

struct BilGeom {
  int N, h, w, H, W, C;
  int ldx, ldy;
  int s2d;           // output (or incoming gradient) stored as 2x2 space-to-depth blocks [N,H/2,W/2,4C]
  float sy, sx;
};

__device__ __forceinline__ size_t bil_out_index(const BilGeom& g, int n, int y, int x) {
  if (!g.s2d) return (size_t(size_t(n) * g.H + y) * g.W + x) * g.ldy;
  return (size_t(size_t(n) * (g.H >> 1) + (y >> 1)) * (g.W >> 1) + (x >> 1)) * (4 * g.C) +
         size_t(((y & 1) * 2 + (x & 1)) * g.C);
}

__global__ void __launch_bounds__(256)
bilinear_fwd_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y, const BilGeom g) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  const int G = g.C >> 3;
  const long total = long(g.N) * g.H * g.W * G;
  for (long i = long(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += long(gridDim.x) * blockDim.x) {
    const int cg = int(i % G);
    long r = i / G;
    const int ox = int(r % g.W); r /= g.W;
    const int oy = int(r % g.H);
    const int n = int(r / g.H);
    int y0, y1, x0, x1;
    float ly, lx;
    bil_coord(oy, g.sy, g.h, &y0, &y1, &ly);
    bil_coord(ox, g.sx, g.w, &x0, &x1, &lx);
    const __nv_bfloat16* base = x + size_t(n) * g.h * g.w * g.ldx + cg * 8;
    float a[8], b[8], c[8], d[8], o[8];
    sp_unpack8(*reinterpret_cast<const uint4*>(base + (size_t(y0) * g.w + x0) * g.ldx), a);
    sp_unpack8(*reinterpret_cast<const uint4*>(base + (size_t(y0) * g.w + x1) * g.ldx), b);
    sp_unpack8(*reinterpret_cast<const uint4*>(base + (size_t(y1) * g.w + x0) * g.ldx), c);
    sp_unpack8(*reinterpret_cast<const uint4*>(base + (size_t(y1) * g.w + x1) * g.ldx), d);
    const float hy = 1.f - ly, hx = 1.f - lx;
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = hy * (hx * a[j] + lx * b[j]) + ly * (hx * c[j] + lx * d[j]);
    *reinterpret_cast<uint4*>(y + bil_out_index(g, n, oy, ox) + cg * 8) = sp_pack8(o);
  }
}

// adjoint by gathering (no atomics): one thread per low-resolution pixel and 8 channels collects
// every output pixel whose two source rows / columns include it.  The candidate range is padded by
// one and each candidate re-runs bil_coord, so the weights are exactly the forward's.
__global__ void __launch_bounds__(256)
bilinear_bwd_gather_kernel(const __nv_bfloat16* __restrict__ gy, float* __restrict__ gx, const BilGeom g) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  const int G = g.C >> 3;
  const long total = long(g.N) * g.h * g.w * G;
  const float ry = g.sy > 0.f ? 1.f / g.sy : 0.f, rx = g.sx > 0.f ? 1.f / g.sx : 0.f;   // 0: a single source row / column
  for (long i = long(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += long(gridDim.x) * blockDim.x) {
    const int cg = int(i % G);
    long r = i / G;
    const int ix = int(r % g.w); r /= g.w;
    const int iy = int(r % g.h);
    const int n = int(r / g.h);
    const int oy_lo = g.sy > 0.f ? max(0, int(float(iy - 1) * ry) - 1) : 0;
    const int oy_hi = g.sy > 0.f ? int(fminf(float(g.H - 1), float(iy + 1) * ry + 1.f)) : g.H - 1;
    const int ox_lo = g.sx > 0.f ? max(0, int(float(ix - 1) * rx) - 1) : 0;
    const int ox_hi = g.sx > 0.f ? int(fminf(float(g.W - 1), float(ix + 1) * rx + 1.f)) : g.W - 1;
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int oy = oy_lo; oy <= oy_hi; ++oy) {
      int y0, y1;
      float ly;
      bil_coord(oy, g.sy, g.h, &y0, &y1, &ly);
      float wy = 0.f;
      if (y0 == iy) wy += 1.f - ly;
      if (y1 == iy && y1 != y0) wy += ly;
      if (wy == 0.f) continue;
      for (int ox = ox_lo; ox <= ox_hi; ++ox) {
        int x0, x1;
        float lx;
        bil_coord(ox, g.sx, g.w, &x0, &x1, &lx);
        float wx = 0.f;
        if (x0 == ix) wx += 1.f - lx;
        if (x1 == ix && x1 != x0) wx += lx;
        if (wx == 0.f) continue;
        float v[8];
        sp_unpack8(__ldg(reinterpret_cast<const uint4*>(gy + bil_out_index(g, n, oy, ox) + cg * 8)), v);
        const float wgt = wy * wx;
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = fmaf(wgt, v[j], acc[j]);
      }
    }
    float* o = gx + ((size_t(n) * g.h + iy) * g.w + ix) * g.C + cg * 8;
    *reinterpret_cast<float4*>(o) = make_float4(acc[0], acc[1], acc[2], acc[3]);
    *reinterpret_cast<float4*>(o + 4) = make_float4(acc[4], acc[5], acc[6], acc[7]);
  }
}

// ---------------------------------------------------------------- global average pool
// x [N][P][C] (ld) -> partial sums ws[blockIdx.x][N][C]: the pixel lanes of a block are added in lane
// order through shared memory, the blocks' rows in block order by sum_rows; caller scales by 1/P.
__global__ void __launch_bounds__(256)
gap_fwd_kernel(const __nv_bfloat16* __restrict__ x, int ld, long P, int C, long pix_per_block,
               float* __restrict__ ws) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  extern __shared__ float sh[];   // [R][C]
  const int n = blockIdx.y;
  const int G = C >> 3;
  const int R = G >= 256 ? 1 : 256 / G;
  const long p0 = long(blockIdx.x) * pix_per_block, p1 = min(p0 + pix_per_block, P);
  for (int cg = threadIdx.x % G; cg < G; cg += (G >= 256 ? 256 : G)) {
    const int r = G >= 256 ? 0 : threadIdx.x / G;
    if (r >= R) break;
    float s[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (long p = p0 + r; p < p1; p += R) {
      float f[8];
      sp_unpack8(*reinterpret_cast<const uint4*>(x + (size_t(n) * P + p) * ld + cg * 8), f);
#pragma unroll
      for (int j = 0; j < 8; ++j) s[j] += f[j];
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) sh[size_t(r) * C + cg * 8 + j] = s[j];
    if (G < 256) break;
  }
  __syncthreads();
  float* row = ws + (size_t(blockIdx.x) * gridDim.y + n) * C;
  for (int i = threadIdx.x; i < C; i += blockDim.x) {
    float t = sh[i];
    for (int r = 1; r < R; ++r) t += sh[size_t(r) * C + i];
    row[i] = t;
  }
}

// dx[n,p,c] = g[n,c] * inv_p   (g bf16 [N][C])
__global__ void __launch_bounds__(256)
gap_bwd_kernel(const __nv_bfloat16* __restrict__ g, __nv_bfloat16* __restrict__ dx, long P, int C, int N,
               float inv_p) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  const int G = C >> 3;
  const long total = long(N) * P * G;
  for (long i = long(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += long(gridDim.x) * blockDim.x) {
    const int cg = int(i % G);
    const long row = i / G;
    const int n = int(row / P);
    float f[8];
    sp_unpack8(*reinterpret_cast<const uint4*>(g + size_t(n) * C + cg * 8), f);
#pragma unroll
    for (int j = 0; j < 8; ++j) f[j] *= inv_p;
    *reinterpret_cast<uint4*>(dx + row * C + cg * 8) = sp_pack8(f);
  }
}

// ---------------------------------------------------------------- linear head (K -> 1)
// ws[blockIdx.x][n] = this block's share of sum_i x[n,i] * w[i]     x bf16 [N][K], w bf16 [K];
// sum_rows_block adds the blocks' rows in a fixed order onto the logits (pre-set to the bias).
__global__ void __launch_bounds__(256)
head_fwd_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ w, long K,
                float* __restrict__ ws) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  const int n = blockIdx.y;
  const long K8 = K >> 3;
  float acc = 0.f;
  for (long i = long(blockIdx.x) * blockDim.x + threadIdx.x; i < K8; i += long(gridDim.x) * blockDim.x) {
    float a[8], b[8];
    sp_unpack8(*reinterpret_cast<const uint4*>(x + size_t(n) * K + i * 8), a);
    sp_unpack8(*reinterpret_cast<const uint4*>(w + i * 8), b);
#pragma unroll
    for (int j = 0; j < 8; ++j) acc = fmaf(a[j], b[j], acc);
  }
  __shared__ float red[8];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int i = 0; i < 8; ++i) t += red[i];
    ws[size_t(blockIdx.x) * gridDim.y + n] = t;
  }
}

// dx[n,i] = g[n] * w[i] (bf16, optional);  dw[i] = sum_n g[n] * x[n,i] (fp32, optional)
__global__ void __launch_bounds__(256)
head_bwd_kernel(const float* __restrict__ g, const __nv_bfloat16* __restrict__ x,
                const __nv_bfloat16* __restrict__ w, long K, int N, __nv_bfloat16* __restrict__ dx,
                float* __restrict__ dw) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  const long K8 = K >> 3;
  for (long i = long(blockIdx.x) * blockDim.x + threadIdx.x; i < K8; i += long(gridDim.x) * blockDim.x) {
    float wv[8], acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    sp_unpack8(*reinterpret_cast<const uint4*>(w + i * 8), wv);
    for (int n = 0; n < N; ++n) {
      const float gn = g[n];
      if (dw) {
        float xv[8];
        sp_unpack8(*reinterpret_cast<const uint4*>(x + size_t(n) * K + i * 8), xv);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = fmaf(gn, xv[j], acc[j]);
      }
      if (dx) {
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = gn * wv[j];
        *reinterpret_cast<uint4*>(dx + size_t(n) * K + i * 8) = sp_pack8(o);
      }
    }
    if (dw) {
      *reinterpret_cast<float4*>(dw + i * 8) = make_float4(acc[0], acc[1], acc[2], acc[3]);
      *reinterpret_cast<float4*>(dw + i * 8 + 4) = make_float4(acc[4], acc[5], acc[6], acc[7]);
    }
  }
}

// ---------------------------------------------------------------- fused L1 loss + gradient
// ws[blockIdx.x] = this block's share of sum |a - b|;  grad = sign(a - b) * gscale (bf16)
__global__ void __launch_bounds__(256)
l1_loss_grad_kernel(const __nv_bfloat16* __restrict__ a, const __nv_bfloat16* __restrict__ b, long n8,
                    float gscale, __nv_bfloat16* __restrict__ grad, float* __restrict__ ws) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  float acc = 0.f;
  for (long i = long(blockIdx.x) * blockDim.x + threadIdx.x; i < n8; i += long(gridDim.x) * blockDim.x) {
    float x[8], y[8], o[8];
    sp_unpack8(*reinterpret_cast<const uint4*>(a + i * 8), x);
    sp_unpack8(*reinterpret_cast<const uint4*>(b + i * 8), y);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float d = x[j] - y[j];
      acc += fabsf(d);
      o[j] = d > 0.f ? gscale : (d < 0.f ? -gscale : 0.f);
    }
    if (grad) *reinterpret_cast<uint4*>(grad + i * 8) = sp_pack8(o);
  }
  __shared__ float red[8];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int i = 0; i < 8; ++i) t += red[i];
    ws[blockIdx.x] = t;
  }
}

// Regression losses of the script's flag surface (train_gan.py:141-152,278-281; utils/losses.py:101-112) in one
// kernel: MODE 0 = L1 |d|, 1 = SmoothL1 (beta), 2 = MSE d^2; optional per-element weights (the masks of
// --enable_masks, L1LossWeighted): loss = sum w * l(d), grad = w * l'(d) * gscale [* *gscale_dev].
// ws[blockIdx.x] = this block's share of the sum.
template <int MODE, bool WEIGHTED>
__global__ void __launch_bounds__(256)
reg_loss_grad_kernel(const __nv_bfloat16* __restrict__ a, const __nv_bfloat16* __restrict__ b,
                     const __nv_bfloat16* __restrict__ w, long n8, float gscale, const float* __restrict__ gscale_dev,
                     float beta, __nv_bfloat16* __restrict__ grad, float* __restrict__ ws) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  float acc = 0.f;
  const float gs = gscale_dev ? gscale * __ldg(gscale_dev) : gscale;
  const float inv_beta = beta > 0.f ? 1.f / beta : 0.f;
  for (long i = long(blockIdx.x) * blockDim.x + threadIdx.x; i < n8; i += long(gridDim.x) * blockDim.x) {
    float x[8], y[8], wt[8], o[8];
    sp_unpack8(*reinterpret_cast<const uint4*>(a + i * 8), x);
    sp_unpack8(*reinterpret_cast<const uint4*>(b + i * 8), y);
    if (WEIGHTED) sp_unpack8(*reinterpret_cast<const uint4*>(w + i * 8), wt);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float d = x[j] - y[j], ad = fabsf(d);
      float l, g;
      if (MODE == 0) {
        l = ad;
        g = d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f);
      } else if (MODE == 1) {   // nn.SmoothL1Loss: 0.5 d^2 / beta below beta, |d| - 0.5 beta above
        const bool quad = ad < beta;
        l = quad ? 0.5f * d * d * inv_beta : ad - 0.5f * beta;
        g = quad ? d * inv_beta : (d > 0.f ? 1.f : -1.f);
      } else {
        l = d * d;
        g = 2.f * d;
      }
      if (WEIGHTED) { l *= wt[j]; g *= wt[j]; }
      acc += l;
      o[j] = g * gs;
    }
    if (grad) *reinterpret_cast<uint4*>(grad + i * 8) = sp_pack8(o);
  }
  __shared__ float red[8];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int i = 0; i < 8; ++i) t += red[i];
    ws[blockIdx.x] = t;
  }
}

// ---------------------------------------------------------------- field layout conversions
// field = bf16 [N, H/2, W/2, (py, px, c)];  nchw = fp32 [N, C, H, W]
template <bool TO_NCHW>
__global__ void __launch_bounds__(256)
field_convert_kernel(__nv_bfloat16* __restrict__ field, float* __restrict__ nchw, int N, int C, int H, int W) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  const long total = long(N) * C * H * W;
  for (long i = long(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += long(gridDim.x) * blockDim.x) {
    const int x = int(i % W);
    long r = i / W;
    const int y = int(r % H); r /= H;
    const int c = int(r % C);
    const int n = int(r / C);
    const size_t f = (size_t(size_t(n) * (H >> 1) + (y >> 1)) * (W >> 1) + (x >> 1)) * (4 * C) +
                     size_t(((y & 1) * 2 + (x & 1)) * C + c);
    if (TO_NCHW) nchw[i] = __bfloat162float(field[f]);
    else field[f] = __float2bfloat16_rn(nchw[i]);
  }
}

// Generator input with noise planes (deeplab_gan.py:88-90: cat((input_raw, input_noise), dim=1)) in the
// field layout: out[n, by, bx, q, 0:C] = field[n, by, bx, q, :], out[.., C:C+nd] = bf16(noise[n, k, 2by+py, 2bx+px]),
// out[.., C+nd:Cp] = 0 (the entry conv's weight rows for those channels are zero as well).
__global__ void __launch_bounds__(256)
field_append_noise_kernel(const uint4* __restrict__ field, const float* __restrict__ noise, uint4* __restrict__ out,
                          int N, int C, int nd, int Cp, int H, int W) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  const int gin = C >> 3, gout = Cp >> 3;
  const long total = long(N) * (H >> 1) * (W >> 1) * 4 * gout;
  for (long i = long(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += long(gridDim.x) * blockDim.x) {
    const int g = int(i % gout);
    const long pq = i / gout;  // (n, by, bx, q)
    uint4 v;
    if (g < gin) {
      v = field[pq * gin + g];
    } else {
      const int q = int(pq & 3);
      long r = pq >> 2;
      const int bx = int(r % (W >> 1)); r /= (W >> 1);
      const int by = int(r % (H >> 1));
      const int n = int(r / (H >> 1));
      const int y = 2 * by + (q >> 1), x = 2 * bx + (q & 1);
      __align__(16) __nv_bfloat16 e[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int k = g * 8 + j - C;
        e[j] = __float2bfloat16_rn(k < nd ? noise[(size_t(size_t(n) * nd + k) * H + y) * W + x] : 0.f);
      }
      v = *reinterpret_cast<uint4*>(e);
    }
    out[i] = v;
  }
}

}  // namespace bgr

using namespace bgr;

extern "C" int bgr_field_append_noise(const void* field, const float* noise, void* out, int N, int C, int nd, int Cp,
                                      int H, int W, void* stream) {
  if (C % 8 || Cp % 8 || Cp < C + nd || nd < 1 || (H & 1) || (W & 1))
    return fail(BGR_ERR_ARG, "field_append_noise: need C, Cp multiples of 8, Cp >= C + nd, nd >= 1, even H and W");
  BGR_CUDA_OK(launch_pdl(field_append_noise_kernel, dim3(sp_grid(long(N) * (H / 2) * (W / 2) * 4 * (Cp / 8))), dim3(256), 0, reinterpret_cast<cudaStream_t>(stream), 
      static_cast<const uint4*>(field), noise, static_cast<uint4*>(out), N, C, nd, Cp, H, W));
  BGR_LAUNCH_OK();
  return BGR_OK;
}

static int make_bil(BilGeom* g, int N, int h, int w, int H, int W, int C, int ldx, int ldy, int s2d) {
  if (C % 8 || C < 8) return fail(BGR_ERR_ARG, "bilinear: channel count must be a multiple of 8");
  if (s2d && ((H | W) & 1)) return fail(BGR_ERR_ARG, "bilinear: space-to-depth output needs even size");
  g->N = N; g->h = h; g->w = w; g->H = H; g->W = W; g->C = C; g->ldx = ldx; g->ldy = ldy; g->s2d = s2d;
  g->sy = H > 1 ? float(h - 1) / float(H - 1) : 0.f;
  g->sx = W > 1 ? float(w - 1) / float(W - 1) : 0.f;
  return BGR_OK;
}

extern "C" int bgr_bilinear_fwd(const void* x, int N, int h, int w, int C, int ldx, void* y, int H, int W,
                                int ldy, int s2d, void* stream) {
  BilGeom g;
  if (int rc = make_bil(&g, N, h, w, H, W, C, ldx, ldy, s2d)) return rc;
  BGR_CUDA_OK(launch_pdl(bilinear_fwd_kernel, dim3(sp_grid(long(N) * H * W * (C / 8))), dim3(256), 0, reinterpret_cast<cudaStream_t>(stream), 
      static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(y), g));
  BGR_LAUNCH_OK();
  return BGR_OK;
}

extern "C" int bgr_bilinear_bwd(const void* gy, int N, int h, int w, int C, int H, int W, int ldy, int s2d,
                                float* gx, void* stream) {
  BilGeom g;
  if (int rc = make_bil(&g, N, h, w, H, W, C, C, ldy, s2d)) return rc;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  // always the gather form (no atomics, schedule-independent).  Broadcasts of a 1x1 map (the image-pool
  // branch, deeplab.py:623-624) are summed by the caller with bgr_gap_fwd instead: one thread per source
  // pixel would walk the whole image.
  BGR_CUDA_OK(launch_pdl(bilinear_bwd_gather_kernel, dim3(sp_grid(long(N) * h * w * (C / 8))), dim3(256), 0, st, 
      static_cast<const __nv_bfloat16*>(gy), gx, g));
  BGR_LAUNCH_OK();
  return BGR_OK;
}

extern "C" int bgr_gap_fwd(const void* x, int N, long P, int C, int ld, float* out_sums, float* ws, void* stream) {
  if (C % 8) return fail(BGR_ERR_ARG, "gap: channel count must be a multiple of 8");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int G = C / 8, R = G >= 256 ? 1 : 256 / G;
  long blocks = (P + long(R) * 4 - 1) / (long(R) * 4);
  if (blocks > BGR_GAP_MAX_ROWS) blocks = BGR_GAP_MAX_ROWS;
  if (blocks < 1) blocks = 1;
  const long ppb = (P + blocks - 1) / blocks;
  blocks = (P + ppb - 1) / ppb;
  BGR_CUDA_OK(launch_pdl(gap_fwd_kernel, dim3(unsigned(blocks), unsigned(N)), dim3(256), size_t(R) * C * sizeof(float), st, 
      static_cast<const __nv_bfloat16*>(x), ld, P, C, ppb, ws));
  BGR_LAUNCH_OK();
  return launch_sum_rows(ws, int(blocks), long(N) * C, out_sums, st);
}

extern "C" int bgr_gap_bwd(const void* g, int N, long P, int C, void* dx, void* stream) {
  if (C % 8) return fail(BGR_ERR_ARG, "gap: channel count must be a multiple of 8");
  BGR_CUDA_OK(launch_pdl(gap_bwd_kernel, dim3(sp_grid(long(N) * P * (C / 8))), dim3(256), 0, reinterpret_cast<cudaStream_t>(stream), 
      static_cast<const __nv_bfloat16*>(g), static_cast<__nv_bfloat16*>(dx), P, C, N, 1.f / float(P)));
  BGR_LAUNCH_OK();
  return BGR_OK;
}

extern "C" int bgr_head_fwd(const void* x, const void* w, int N, long K, float* logits, float* ws, void* stream) {
  if (K % 8) return fail(BGR_ERR_ARG, "head: K must be a multiple of 8");
  long blocks = (K / 8 + 255) / 256;
  if (blocks > BGR_HEAD_MAX_ROWS) blocks = BGR_HEAD_MAX_ROWS;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  BGR_CUDA_OK(launch_pdl(head_fwd_kernel, dim3(unsigned(blocks), unsigned(N)), dim3(256), 0, st, 
      static_cast<const __nv_bfloat16*>(x), static_cast<const __nv_bfloat16*>(w), K, ws));
  BGR_LAUNCH_OK();
  return launch_sum_rows_block(ws, int(blocks), N, logits, nullptr, st);
}

extern "C" int bgr_head_bwd(const float* g, const void* x, const void* w, int N, long K, void* dx, float* dw,
                            void* stream) {
  if (K % 8) return fail(BGR_ERR_ARG, "head: K must be a multiple of 8");
  BGR_CUDA_OK(launch_pdl(head_bwd_kernel, dim3(sp_grid(K / 8)), dim3(256), 0, reinterpret_cast<cudaStream_t>(stream), 
      g, static_cast<const __nv_bfloat16*>(x), static_cast<const __nv_bfloat16*>(w), K, N,
      static_cast<__nv_bfloat16*>(dx), dw));
  BGR_LAUNCH_OK();
  return BGR_OK;
}

extern "C" int bgr_l1_loss_grad(const void* a, const void* b, long n, float gscale, void* grad, double* sum,
                                float* ws, void* stream) {
  if (n % 8) return fail(BGR_ERR_ARG, "l1: element count must be a multiple of 8");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int blocks = sp_grid(n / 8);   // <= BGR_LOSS_MAX_ROWS
  BGR_CUDA_OK(launch_pdl(l1_loss_grad_kernel, dim3(blocks), dim3(256), 0, st, static_cast<const __nv_bfloat16*>(a),
                                              static_cast<const __nv_bfloat16*>(b), n / 8, gscale,
                                              static_cast<__nv_bfloat16*>(grad), ws));
  BGR_LAUNCH_OK();
  return launch_sum_rows_block(ws, blocks, 1, nullptr, sum, st);
}

extern "C" int bgr_reg_loss_grad(const void* a, const void* b, const void* w, long n, int mode, float beta, float gscale,
                                 const float* gscale_dev, void* grad, double* sum, float* ws, void* stream) {
  if (n % 8) return fail(BGR_ERR_ARG, "reg_loss: element count must be a multiple of 8");
  if (mode < 0 || mode > 2) return fail(BGR_ERR_ARG, "reg_loss: mode must be 0 (L1), 1 (SmoothL1) or 2 (MSE)");
  if (mode == 1 && !(beta > 0.f)) return fail(BGR_ERR_ARG, "reg_loss: SmoothL1 needs beta > 0");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int blocks = sp_grid(n / 8);   // <= BGR_LOSS_MAX_ROWS
  auto A = static_cast<const __nv_bfloat16*>(a);
  auto B = static_cast<const __nv_bfloat16*>(b);
  auto W = static_cast<const __nv_bfloat16*>(w);
  auto G = static_cast<__nv_bfloat16*>(grad);
#define BGR_REG_LAUNCH(M)                                                                                               \
  if (W) { BGR_CUDA_OK(launch_pdl(reg_loss_grad_kernel<M, true>, dim3(blocks), dim3(256), 0, st, A, B, W, n / 8, gscale,   \
                                  gscale_dev, beta, G, ws)); }                                                           \
  else { BGR_CUDA_OK(launch_pdl(reg_loss_grad_kernel<M, false>, dim3(blocks), dim3(256), 0, st, A, B, W, n / 8, gscale,    \
                                gscale_dev, beta, G, ws)); }
  if (mode == 0) { BGR_REG_LAUNCH(0) }
  else if (mode == 1) { BGR_REG_LAUNCH(1) }
  else { BGR_REG_LAUNCH(2) }
#undef BGR_REG_LAUNCH
  BGR_LAUNCH_OK();
  return launch_sum_rows_block(ws, blocks, 1, nullptr, sum, st);
}

extern "C" int bgr_field_to_nchw(const void* field, float* nchw, int N, int C, int H, int W, void* stream) {
  BGR_CUDA_OK(launch_pdl(field_convert_kernel<true>, dim3(sp_grid(long(N) * C * H * W)), dim3(256), 0, reinterpret_cast<cudaStream_t>(stream), 
      const_cast<__nv_bfloat16*>(static_cast<const __nv_bfloat16*>(field)), nchw, N, C, H, W));
  BGR_LAUNCH_OK();
  return BGR_OK;
}

extern "C" int bgr_nchw_to_field(const float* nchw, void* field, int N, int C, int H, int W, void* stream) {
  BGR_CUDA_OK(launch_pdl(field_convert_kernel<false>, dim3(sp_grid(long(N) * C * H * W)), dim3(256), 0, reinterpret_cast<cudaStream_t>(stream), 
      static_cast<__nv_bfloat16*>(field), const_cast<float*>(nchw), N, C, H, W));
  BGR_LAUNCH_OK();
  return BGR_OK;
}

// ---------------------------------------------------------------- weight layouts
// fp32 master weights [Cout][Cin][taps] (PyTorch's [Cout,Cin,kh,kw]) -> bf16 GEMM operand
// [rows][taps][pitch]: rows = Cout, inner = Cin (fprop) or, transposed, rows = Cin, inner = Cout
// (dgrad).  The padding up to `pitch` is written as zeros.  One launch per parameter instead of a
// permute + cast + pad chain.
namespace bgr {
// fprop layout: out[co][t][ci] = w[co][ci][t].  One block = one cout x 256 cins: the contiguous
// [256][taps] fp32 chunk is read coalesced into shared memory and written tap-major.
__global__ void __launch_bounds__(256)
weight_to_gemm_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ out, int Cin, int taps, int pitch) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  extern __shared__ float sh[];   // [256][taps]
  const int co = blockIdx.x, ci0 = blockIdx.y * 256;
  const int n = min(256, pitch - ci0);            // columns to write (zeros beyond Cin)
  const int nv = max(0, min(256, Cin - ci0));     // columns that exist in w
  const float* src = w + (size_t(co) * Cin + ci0) * taps;
  for (int i = threadIdx.x; i < nv * taps; i += 256) sh[i] = src[i];
  __syncthreads();
  for (int i = threadIdx.x; i < n * taps; i += 256) {
    const int t = i / n, k = i - t * n;
    out[(size_t(co) * taps + t) * pitch + ci0 + k] = __float2bfloat16_rn(k < nv ? sh[k * taps + t] : 0.f);
  }
}
// dgrad layout: out[j][co] = w2d[co][j] with j = ci*taps + t -> (ci, t) rows: a 2-D transpose in
// 32x32 shared-memory tiles (coalesced on both sides); co in [Cout, pitch) is written as zero.
__global__ void __launch_bounds__(256)
weight_to_gemm_t_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ out, int Cout, int J, int pitch) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  __shared__ float tile[32][33];
  const int j0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;   // 32 x 8
#pragma unroll
  for (int r = ty; r < 32; r += 8) {
    const int co = c0 + r, j = j0 + tx;
    tile[r][tx] = (co < Cout && j < J) ? w[size_t(co) * J + j] : 0.f;
  }
  __syncthreads();
#pragma unroll
  for (int r = ty; r < 32; r += 8) {
    const int j = j0 + r, co = c0 + tx;
    if (j < J && co < pitch) out[size_t(j) * pitch + co] = __float2bfloat16_rn(tile[tx][r]);
  }
}
}  // namespace bgr

extern "C" int bgr_weight_to_gemm(const float* w, int Cout, int Cin, int taps, int transpose, void* out, int pitch,
                                  void* stream) {
  const int inner = transpose ? Cout : Cin;
  if (pitch < inner) return bgr::fail(BGR_ERR_ARG, "weight_to_gemm: pitch smaller than the inner dimension");
  if (taps < 1 || taps > 16) return bgr::fail(BGR_ERR_ARG, "weight_to_gemm: taps must be in [1,16]");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  auto O = static_cast<__nv_bfloat16*>(out);
  if (transpose) {
    // [Cout][J = Cin*taps] -> [J][pitch]: row j = ci*taps + t is exactly out[ci][t][:]
    const int J = Cin * taps;
    dim3 grid((J + 31) / 32, (pitch + 31) / 32);
    BGR_CUDA_OK(launch_pdl(bgr::weight_to_gemm_t_kernel, dim3(grid), dim3(256), 0, st, w, O, Cout, J, pitch));
  } else {
    dim3 grid(Cout, (pitch + 255) / 256);
    BGR_CUDA_OK(launch_pdl(bgr::weight_to_gemm_kernel, dim3(grid), dim3(256), 256 * taps * sizeof(float), st, w, O, Cin, taps, pitch));
  }
  BGR_LAUNCH_OK();
  return BGR_OK;
}

// ---------------------------------------------------------------- all derived weight copies of a model, one launch
// After an optimizer step every GEMM-layout copy is stale: ~430 weight_to_gemm launches plus ~250 depthwise
// transposes per G+D step, each a 3-5 us node on the critical path of the step's CUDA graph.  This kernel
// rebuilds all of them from a descriptor table: block -> descriptor by binary search over the prefix sums
// of the per-descriptor block counts, then the same tile code as the single-tensor kernels above.
namespace bgr {
__global__ void __launch_bounds__(256)
weights_prepare_kernel(const BgrWeightDesc* __restrict__ descs, const int* __restrict__ first_block, int ndesc) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  __shared__ float sh[256 * 9];   // mode 0: [256][taps <= 9]; mode 1: a 32x33 tile
  int lo = 0, hi = ndesc - 1;
  while (lo < hi) {               // last d with first_block[d] <= blockIdx.x
    const int mid = (lo + hi + 1) >> 1;
    if (first_block[mid] <= int(blockIdx.x)) lo = mid; else hi = mid - 1;
  }
  const BgrWeightDesc d = descs[lo];
  const int b = int(blockIdx.x) - first_block[lo];
  const float* __restrict__ w = d.src;
  if (d.mode == 0) {              // out[co][t][ci] = w[co][ci][t]
    __nv_bfloat16* out = static_cast<__nv_bfloat16*>(d.dst);
    const int co = b / d.gx, ci0 = (b % d.gx) * 256, taps = d.taps;
    const int n = min(256, d.pitch - ci0);
    const int nv = max(0, min(256, d.Cin - ci0));
    const float* src = w + (size_t(co) * d.Cin + ci0) * taps;
    for (int i = threadIdx.x; i < nv * taps; i += 256) sh[i] = src[i];
    __syncthreads();
    for (int i = threadIdx.x; i < n * taps; i += 256) {
      const int t = i / n, k = i - t * n;
      out[(size_t(co) * taps + t) * d.pitch + ci0 + k] = __float2bfloat16_rn(k < nv ? sh[k * taps + t] : 0.f);
    }
  } else if (d.mode == 1) {       // out[j = ci*taps + t][co] = w[co][j]
    __nv_bfloat16* out = static_cast<__nv_bfloat16*>(d.dst);
    float (*tile)[33] = reinterpret_cast<float (*)[33]>(sh);
    const int J = d.Cin * d.taps;
    const int j0 = (b % d.gx) * 32, c0 = (b / d.gx) * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
#pragma unroll
    for (int r = ty; r < 32; r += 8) {
      const int co = c0 + r, j = j0 + tx;
      tile[r][tx] = (co < d.Cout && j < J) ? w[size_t(co) * J + j] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int r = ty; r < 32; r += 8) {
      const int j = j0 + r, co = c0 + tx;
      if (j < J && co < d.pitch) out[size_t(j) * d.pitch + co] = __float2bfloat16_rn(tile[tx][r]);
    }
  } else {                        // depthwise: fp32 out[t][c] = w[c][t], t < 9
    float* out = static_cast<float*>(d.dst);
    const int i = b * 256 + threadIdx.x;
    if (i < 9 * d.Cout) {
      const int t = i / d.Cout, c = i - t * d.Cout;
      out[i] = w[c * 9 + t];
    }
  }
}
}  // namespace bgr

extern "C" int bgr_weights_prepare(const BgrWeightDesc* descs, const int* first_block, int ndesc, int total_blocks,
                                   void* stream) {
  if (ndesc < 1 || total_blocks < 1) return bgr::fail(BGR_ERR_ARG, "weights_prepare: empty descriptor table");
  BGR_CUDA_OK(launch_pdl(bgr::weights_prepare_kernel, dim3(total_blocks), dim3(256), 0, reinterpret_cast<cudaStream_t>(stream), descs, first_block, ndesc));
  BGR_LAUNCH_OK();
  return BGR_OK;
}

