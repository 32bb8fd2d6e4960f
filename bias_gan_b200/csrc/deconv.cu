// Layout / resampling glue of the Deconv and Deconv1x decoders (DeconvUpsampler / UpsamplerExtension,
// src/deepCam/architecture/gpsro/deeplab.py:398-444,466-500) and of odd-sized, odd-channel-count fields
// (the reference's native GPSRO data: 45 channels x 19 x 37):
//   zero-stuffing + adjoint : ConvTranspose2d(k3, s2, p, output_padding) runs as the implicit-GEMM conv
//                             kernel over the zero-stuffed input with tap offsets p - k (bias_gan_b200/
//                             functional.py _ConvT) -- the only dense contraction stays on tcgen05;
//   AvgPool2d(2, stride 1, padding 0|1) (+ LeakyReLU) and its adjoint (deeplab.py:409,414,428,478,658);
//   NHWC / fp32 NCHW  <->  2x2 space-to-depth field with zero padding to even H, W and to a multiple of
//                             8 channels, optionally standardising on the way in (gpsro_dataset.py:169-170).
// All HBM-bound and tiny at the sizes the reference runs them; 128-bit accesses, one thread per
// (pixel, 8 channels).
#include <cuda_bf16.h>

#include "bgr_internal.h"

namespace bgr {

__device__ __forceinline__ void dc_unpack8(const uint4& u, float (&f)[8]) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 t = __bfloat1622float2(h[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}
__device__ __forceinline__ uint4 dc_pack8(const float (&f)[8]) {
  uint4 u;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
  return u;
}
static int dc_grid(long total) {
  long blocks = (total + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  return int(blocks < 1 ? 1 : blocks);
}

// y[n, Y, X, :] = x[n, Y/2, X/2, :] where Y, X even and inside the source, else 0
__global__ void __launch_bounds__(256)
zero_stuff2_kernel(const __nv_bfloat16* __restrict__ x, int h, int w, int ldx, __nv_bfloat16* __restrict__ y, int N,
                   int H, int W, int C, int ldy) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  const int G = C >> 3;
  const long total = long(N) * H * W * G;
  for (long i = long(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += long(gridDim.x) * blockDim.x) {
    const int g = int(i % G);
    long r = i / G;
    const int X = int(r % W); r /= W;
    const int Y = int(r % H);
    const int n = int(r / H);
    uint4 v = make_uint4(0, 0, 0, 0);
    if (!(Y & 1) && !(X & 1) && (Y >> 1) < h && (X >> 1) < w)
      v = *reinterpret_cast<const uint4*>(x + (size_t(size_t(n) * h + (Y >> 1)) * w + (X >> 1)) * ldx + g * 8);
    *reinterpret_cast<uint4*>(y + (size_t(size_t(n) * H + Y) * W + X) * ldy + g * 8) = v;
  }
}
// adjoint: gx[n, y, x, :] = gy[n, 2y, 2x, :]
__global__ void __launch_bounds__(256)
zero_stuff2_bwd_kernel(const __nv_bfloat16* __restrict__ gy, int H, int W, int ldy, __nv_bfloat16* __restrict__ gx,
                       int N, int h, int w, int C, int ldx) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  const int G = C >> 3;
  const long total = long(N) * h * w * G;
  for (long i = long(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += long(gridDim.x) * blockDim.x) {
    const int g = int(i % G);
    long r = i / G;
    const int x = int(r % w); r /= w;
    const int y = int(r % h);
    const int n = int(r / h);
    *reinterpret_cast<uint4*>(gx + (size_t(size_t(n) * h + y) * w + x) * ldx + g * 8) =
        *reinterpret_cast<const uint4*>(gy + (size_t(size_t(n) * H + 2 * y) * W + 2 * x) * ldy + g * 8);
  }
}

// nn.AvgPool2d(2, stride=1, padding=pad) with count_include_pad (always / 4), then optional LeakyReLU.
// Output (H + 2 pad - 1) x (W + 2 pad - 1).
__global__ void __launch_bounds__(256)
avgpool2_fwd_kernel(const __nv_bfloat16* __restrict__ x, int N, int H, int W, int C, int ldx, int pad, int act,
                    float slope, __nv_bfloat16* __restrict__ y, int ldy) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  const int G = C >> 3, Ho = H + 2 * pad - 1, Wo = W + 2 * pad - 1;
  const long total = long(N) * Ho * Wo * G;
  for (long i = long(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += long(gridDim.x) * blockDim.x) {
    const int g = int(i % G);
    long r = i / G;
    const int ox = int(r % Wo); r /= Wo;
    const int oy = int(r % Ho);
    const int n = int(r / Ho);
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
    for (int dy = 0; dy < 2; ++dy) {
      const int iy = oy - pad + dy;
      if (iy < 0 || iy >= H) continue;
#pragma unroll
      for (int dx = 0; dx < 2; ++dx) {
        const int ix = ox - pad + dx;
        if (ix < 0 || ix >= W) continue;
        float f[8];
        dc_unpack8(*reinterpret_cast<const uint4*>(x + (size_t(size_t(n) * H + iy) * W + ix) * ldx + g * 8), f);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += f[j];
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float v = 0.25f * acc[j];
      if (act) v = v > 0.f ? v : slope * v;
      acc[j] = v;
    }
    *reinterpret_cast<uint4*>(y + (size_t(size_t(n) * Ho + oy) * Wo + ox) * ldy + g * 8) = dc_pack8(acc);
  }
}
// adjoint (gather): gx[iy, ix] = 0.25 * sum_{dy,dx} gy'[iy + pad - dy, ix + pad - dx], gy' = gy * act'(yout)
__global__ void __launch_bounds__(256)
avgpool2_bwd_kernel(const __nv_bfloat16* __restrict__ gy, int ldg, const __nv_bfloat16* __restrict__ yout, int ldo,
                    int N, int H, int W, int C, int pad, int act, float slope, __nv_bfloat16* __restrict__ gx, int ldx) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  const int G = C >> 3, Ho = H + 2 * pad - 1, Wo = W + 2 * pad - 1;
  const long total = long(N) * H * W * G;
  for (long i = long(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += long(gridDim.x) * blockDim.x) {
    const int g = int(i % G);
    long r = i / G;
    const int ix = int(r % W); r /= W;
    const int iy = int(r % H);
    const int n = int(r / H);
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
    for (int dy = 0; dy < 2; ++dy) {
      const int oy = iy + pad - dy;
      if (oy < 0 || oy >= Ho) continue;
#pragma unroll
      for (int dx = 0; dx < 2; ++dx) {
        const int ox = ix + pad - dx;
        if (ox < 0 || ox >= Wo) continue;
        const size_t o = size_t(size_t(n) * Ho + oy) * Wo + ox;
        float f[8];
        dc_unpack8(*reinterpret_cast<const uint4*>(gy + o * ldg + g * 8), f);
        if (act) {
          float s[8];
          dc_unpack8(*reinterpret_cast<const uint4*>(yout + o * ldo + g * 8), s);
#pragma unroll
          for (int j = 0; j < 8; ++j) f[j] = s[j] > 0.f ? f[j] : slope * f[j];
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += f[j];
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] *= 0.25f;
    *reinterpret_cast<uint4*>(gx + (size_t(size_t(n) * H + iy) * W + ix) * ldx + g * 8) = dc_pack8(acc);
  }
}

// bf16 NHWC [N,H,W,C] (pixel stride ldx) <-> field [N, Hp/2, Wp/2, (py, px, c)], Hp = H rounded up to even
// (the extra row / column is zero: a stride-2 pad-1 conv over it equals the conv over the odd-sized image).
template <bool TO_FIELD>
__global__ void __launch_bounds__(256)
nhwc_field_kernel(__nv_bfloat16* __restrict__ nhwc, int ld, __nv_bfloat16* __restrict__ field, int N, int H, int W, int C) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  const int G = C >> 3, Hp = (H + 1) & ~1, Wp = (W + 1) & ~1;
  const long total = long(N) * (TO_FIELD ? Hp : H) * (TO_FIELD ? Wp : W) * G;
  for (long i = long(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += long(gridDim.x) * blockDim.x) {
    const int g = int(i % G);
    long r = i / G;
    const int Wl = TO_FIELD ? Wp : W, Hl = TO_FIELD ? Hp : H;
    const int x = int(r % Wl); r /= Wl;
    const int y = int(r % Hl);
    const int n = int(r / Hl);
    const size_t f = (size_t(size_t(n) * (Hp >> 1) + (y >> 1)) * (Wp >> 1) + (x >> 1)) * (4 * C) +
                     size_t(((y & 1) * 2 + (x & 1)) * C) + g * 8;
    const size_t p = (size_t(size_t(n) * H + y) * W + x) * ld + g * 8;
    if (TO_FIELD) {
      uint4 v = make_uint4(0, 0, 0, 0);
      if (y < H && x < W) v = *reinterpret_cast<const uint4*>(nhwc + p);
      *reinterpret_cast<uint4*>(field + f) = v;
    } else {
      *reinterpret_cast<uint4*>(nhwc + p) = *reinterpret_cast<const uint4*>(field + f);
    }
  }
}

constexpr int kMaxPadChannels = 64;
struct PadStage {
  int B, Cf, H, W;      // source fp32 NCHW
  int C, Cp;            // selected channels, padded channel count of the field (multiple of 8, <= 64)
  int channels[kMaxPadChannels];
};
// field[b, y/2, x/2, (py, px, i)] = bf16(scale[i] * (src[b, channels[i], y, x] - shift[i])) for i < C, y < H, x < W; 0 elsewhere
__global__ void __launch_bounds__(256)
stage_padded_kernel(const float* __restrict__ src, const float* __restrict__ shift, const float* __restrict__ scale,
                    __nv_bfloat16* __restrict__ field, const __grid_constant__ PadStage p) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  const int Hp = (p.H + 1) & ~1, Wp = (p.W + 1) & ~1;
  const long total = long(p.B) * Hp * Wp * p.Cp;
  for (long i = long(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += long(gridDim.x) * blockDim.x) {
    const int c = int(i % p.Cp);
    long r = i / p.Cp;
    const int x = int(r % Wp); r /= Wp;
    const int y = int(r % Hp);
    const int b = int(r / Hp);
    float v = 0.f;
    if (c < p.C && y < p.H && x < p.W) {
      v = src[(size_t(size_t(b) * p.Cf + p.channels[c]) * p.H + y) * p.W + x];
      if (shift) v = __fmul_rn(scale[c], __fsub_rn(v, shift[c]));   // exactly scale * (x - shift), like the stage kernel
    }
    field[(size_t(size_t(b) * (Hp >> 1) + (y >> 1)) * (Wp >> 1) + (x >> 1)) * (4 * p.Cp) +
          size_t(((y & 1) * 2 + (x & 1)) * p.Cp + c)] = __float2bfloat16_rn(v);
  }
}
// nchw[b, c, y, x] = float(field[...]) for c < C
__global__ void __launch_bounds__(256)
field_to_nchw_padded_kernel(const __nv_bfloat16* __restrict__ field, float* __restrict__ nchw, int B, int C, int Cp, int H, int W) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  const int Hp = (H + 1) & ~1, Wp = (W + 1) & ~1;
  const long total = long(B) * C * H * W;
  for (long i = long(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += long(gridDim.x) * blockDim.x) {
    const int x = int(i % W);
    long r = i / W;
    const int y = int(r % H); r /= H;
    const int c = int(r % C);
    const int b = int(r / C);
    nchw[i] = __bfloat162float(field[(size_t(size_t(b) * (Hp >> 1) + (y >> 1)) * (Wp >> 1) + (x >> 1)) * (4 * Cp) +
                                     size_t(((y & 1) * 2 + (x & 1)) * Cp + c)]);
  }
}

}  // namespace bgr

using namespace bgr;

#define DC_CHECK_C8(C) \
  if ((C) % 8 || (C) < 8) return fail(BGR_ERR_ARG, "channel count must be a multiple of 8")

extern "C" int bgr_zero_stuff2(const void* x, int N, int h, int w, int C, int ldx, void* y, int H, int W, int ldy,
                               void* stream) {
  DC_CHECK_C8(C);
  if (H < 2 * h - 1 || W < 2 * w - 1) return fail(BGR_ERR_ARG, "zero_stuff2: output smaller than 2*input - 1");
  BGR_CUDA_OK(launch_pdl(zero_stuff2_kernel, dim3(dc_grid(long(N) * H * W * (C / 8))), dim3(256), 0, reinterpret_cast<cudaStream_t>(stream), 
      static_cast<const __nv_bfloat16*>(x), h, w, ldx, static_cast<__nv_bfloat16*>(y), N, H, W, C, ldy));
  BGR_LAUNCH_OK();
  return BGR_OK;
}

extern "C" int bgr_zero_stuff2_bwd(const void* gy, int N, int h, int w, int C, int H, int W, int ldy, void* gx, int ldx,
                                   void* stream) {
  DC_CHECK_C8(C);
  if (H < 2 * h - 1 || W < 2 * w - 1) return fail(BGR_ERR_ARG, "zero_stuff2: output smaller than 2*input - 1");
  BGR_CUDA_OK(launch_pdl(zero_stuff2_bwd_kernel, dim3(dc_grid(long(N) * h * w * (C / 8))), dim3(256), 0, reinterpret_cast<cudaStream_t>(stream), 
      static_cast<const __nv_bfloat16*>(gy), H, W, ldy, static_cast<__nv_bfloat16*>(gx), N, h, w, C, ldx));
  BGR_LAUNCH_OK();
  return BGR_OK;
}

extern "C" int bgr_avgpool2_fwd(const void* x, int N, int H, int W, int C, int ldx, int pad, int act, float slope,
                                void* y, int ldy, void* stream) {
  DC_CHECK_C8(C);
  if (pad < 0 || pad > 1 || H + 2 * pad < 2 || W + 2 * pad < 2) return fail(BGR_ERR_ARG, "avgpool2: padding 0 or 1, window must fit");
  BGR_CUDA_OK(launch_pdl(avgpool2_fwd_kernel, dim3(dc_grid(long(N) * (H + 2 * pad - 1) * (W + 2 * pad - 1) * (C / 8))), dim3(256), 0, reinterpret_cast<cudaStream_t>(stream), static_cast<const __nv_bfloat16*>(x), N, H, W, C, ldx,
                                                                  pad, act, slope, static_cast<__nv_bfloat16*>(y), ldy));
  BGR_LAUNCH_OK();
  return BGR_OK;
}

extern "C" int bgr_avgpool2_bwd(const void* gy, int ldg, const void* yout, int ldo, int N, int H, int W, int C, int pad,
                                int act, float slope, void* gx, int ldx, void* stream) {
  DC_CHECK_C8(C);
  if (pad < 0 || pad > 1) return fail(BGR_ERR_ARG, "avgpool2: padding 0 or 1");
  if (act && yout == nullptr) return fail(BGR_ERR_ARG, "avgpool2_bwd: the stored output is needed for the activation");
  BGR_CUDA_OK(launch_pdl(avgpool2_bwd_kernel, dim3(dc_grid(long(N) * H * W * (C / 8))), dim3(256), 0, reinterpret_cast<cudaStream_t>(stream), 
      static_cast<const __nv_bfloat16*>(gy), ldg, static_cast<const __nv_bfloat16*>(yout), ldo, N, H, W, C, pad, act,
      slope, static_cast<__nv_bfloat16*>(gx), ldx));
  BGR_LAUNCH_OK();
  return BGR_OK;
}

extern "C" int bgr_nhwc_to_field(const void* x, int N, int H, int W, int C, int ldx, void* field, void* stream) {
  DC_CHECK_C8(C);
  const int Hp = (H + 1) & ~1, Wp = (W + 1) & ~1;
  BGR_CUDA_OK(launch_pdl(nhwc_field_kernel<true>, dim3(dc_grid(long(N) * Hp * Wp * (C / 8))), dim3(256), 0, reinterpret_cast<cudaStream_t>(stream), 
      const_cast<__nv_bfloat16*>(static_cast<const __nv_bfloat16*>(x)), ldx, static_cast<__nv_bfloat16*>(field), N, H, W, C));
  BGR_LAUNCH_OK();
  return BGR_OK;
}

extern "C" int bgr_field_to_nhwc(const void* field, int N, int H, int W, int C, void* y, int ldy, void* stream) {
  DC_CHECK_C8(C);
  BGR_CUDA_OK(launch_pdl(nhwc_field_kernel<false>, dim3(dc_grid(long(N) * H * W * (C / 8))), dim3(256), 0, reinterpret_cast<cudaStream_t>(stream), 
      static_cast<__nv_bfloat16*>(y), ldy, const_cast<__nv_bfloat16*>(static_cast<const __nv_bfloat16*>(field)), N, H, W, C));
  BGR_LAUNCH_OK();
  return BGR_OK;
}

extern "C" int bgr_stage_normalize_padded(const float* src, int B, int Cf, int H, int W, const int* channels, int C,
                                          const float* shift, const float* scale, void* field, int Cp, void* stream) {
  if (Cp % 8 || Cp < 8 || Cp > kMaxPadChannels || C < 1 || C > Cp)
    return fail(BGR_ERR_ARG, "stage_padded: Cp must be a multiple of 8 in [8,64] and 1 <= C <= Cp");
  if ((shift == nullptr) != (scale == nullptr)) return fail(BGR_ERR_ARG, "stage_padded: shift and scale go together");
  PadStage p{};
  p.B = B; p.Cf = Cf; p.H = H; p.W = W; p.C = C; p.Cp = Cp;
  for (int i = 0; i < C; ++i) {
    const int c = channels ? channels[i] : i;
    if (c < 0 || c >= Cf) return fail(BGR_ERR_ARG, "stage_padded: channel index out of range");
    p.channels[i] = c;
  }
  const long total = long(B) * ((H + 1) & ~1) * ((W + 1) & ~1) * Cp;
  BGR_CUDA_OK(launch_pdl(stage_padded_kernel, dim3(dc_grid(total)), dim3(256), 0, reinterpret_cast<cudaStream_t>(stream), 
      src, shift, scale, static_cast<__nv_bfloat16*>(field), p));
  BGR_LAUNCH_OK();
  return BGR_OK;
}

extern "C" int bgr_field_to_nchw_padded(const void* field, float* nchw, int B, int C, int Cp, int H, int W, void* stream) {
  if (Cp % 8 || C < 1 || C > Cp) return fail(BGR_ERR_ARG, "field_to_nchw_padded: 1 <= C <= Cp, Cp multiple of 8");
  BGR_CUDA_OK(launch_pdl(field_to_nchw_padded_kernel, dim3(dc_grid(long(B) * C * H * W)), dim3(256), 0, reinterpret_cast<cudaStream_t>(stream), 
      static_cast<const __nv_bfloat16*>(field), nchw, B, C, Cp, H, W));
  BGR_LAUNCH_OK();
  return BGR_OK;
}
