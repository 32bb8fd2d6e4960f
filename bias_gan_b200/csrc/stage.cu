// Fused decode / per-channel standardise / crop / flip / bf16 staging kernel.
//
// Replaces, in one pass over HBM, the reference's  data_scale * (data - data_shift)
// (src/deepCam/data/gpsro_dataset.py:169-170), the channel select + NHWC->NCHW permute of the CAM
// loader (data/cam_numpy_singlefile_dataset.py:200) and the collate copy.  Reads the reader's
// payload (fp32/fp64/int, NCHW per-file layout or NHWC single-file layout), writes bf16 NHWC or
// bf16 2x2 space-to-depth blocks [B, h/2, w/2, (py, px, c)] — the layout the stride-2 entry conv
// consumes as a plain 2x2-tap implicit GEMM.
//
// HBM-bound: 4 B read + 2 B written per element.  Each thread owns 4 consecutive output pixels x 8
// channels: 128-bit loads along x from 8 channel planes (NCHW) or 2 x 128-bit loads per pixel
// (NHWC), four 128-bit stores.  Arithmetic is a separate fp32 subtract and multiply (no FMA
// contraction) so results are bit-identical to the oracle's scale * (x - shift) rounded RNE.
#include <cuda_bf16.h>

#include "bgr_internal.h"

namespace bgr {

constexpr int kMaxStageChannels = 64;

struct StageParams {
  int B, Cf, H, W;          // source dims
  int C;                    // selected channels (multiple of 8)
  int y0, x0, h, w;         // crop window in source pixels
  int flip_h, flip_v;
  int s2d;
  int identity_channels;    // channels[i] == i for all i
  int channels[kMaxStageChannels];
};

__device__ __forceinline__ uint32_t pack2(float a, float b) {
  __nv_bfloat162 t = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&t);
}

__device__ __forceinline__ size_t out_index(const StageParams& p, int b, int y, int x) {
  // element offset of channel 0 of output pixel (y, x)
  if (!p.s2d) return (size_t(size_t(b) * p.h + y) * p.w + x) * p.C;
  return (size_t(size_t(b) * (p.h >> 1) + (y >> 1)) * (p.w >> 1) + (x >> 1)) * (4 * p.C) +
         size_t(((y & 1) * 2 + (x & 1)) * p.C);
}

template <typename T, int LAYOUT, bool VEC>
__global__ void __launch_bounds__(256)
stage_kernel(const T* __restrict__ src, const float* __restrict__ shift,
             const float* __restrict__ scale, __nv_bfloat16* __restrict__ dst,
             const __grid_constant__ StageParams p) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  const int groups = p.C >> 3;                 // 8-channel groups
  const int wq = p.w >> 2;                     // 4-pixel quads per row (w % 4 == 0 required)
  const size_t total = size_t(p.B) * p.h * wq * groups;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += size_t(gridDim.x) * blockDim.x) {
    // NCHW: quads fastest (coalesced plane reads); NHWC: channel groups fastest
    int g, q;
    size_t r;
    if (LAYOUT == BGR_NCHW) { q = int(i % wq); r = i / wq; g = int(r % groups); r /= groups; }
    else { g = int(i % groups); r = i / groups; q = int(r % wq); r /= wq; }
    const int y = int(r % p.h);
    const int b = int(r / p.h);
    const int sy = p.y0 + (p.flip_v ? (p.h - 1 - y) : y);
    // output pixels x = 4q .. 4q+3 read source columns sx(j)
    const int xo = q * 4;
    const int sx_first = p.x0 + (p.flip_h ? (p.w - 1 - xo) : xo);  // source column of output xo
    const int dir = p.flip_h ? -1 : 1;
    float v[4][8];
    if (LAYOUT == BGR_NCHW) {
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        const int cs = p.identity_channels ? (g * 8 + c) : p.channels[g * 8 + c];
        const T* row = src + (size_t(size_t(b) * p.Cf + cs) * p.H + sy) * p.W;
        if (VEC) {
          const int lo = p.flip_h ? (sx_first - 3) : sx_first;
          const float4 f = *reinterpret_cast<const float4*>(row + lo);
          if (p.flip_h) { v[0][c] = f.w; v[1][c] = f.z; v[2][c] = f.y; v[3][c] = f.x; }
          else { v[0][c] = f.x; v[1][c] = f.y; v[2][c] = f.z; v[3][c] = f.w; }
        } else {
#pragma unroll
          for (int j = 0; j < 4; ++j) v[j][c] = static_cast<float>(row[sx_first + dir * j]);
        }
      }
    } else {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const T* px = src + (size_t(size_t(b) * p.H + sy) * p.W + (sx_first + dir * j)) * p.Cf;
        if (VEC) {
          const float4 f0 = *reinterpret_cast<const float4*>(px + g * 8);
          const float4 f1 = *reinterpret_cast<const float4*>(px + g * 8 + 4);
          v[j][0] = f0.x; v[j][1] = f0.y; v[j][2] = f0.z; v[j][3] = f0.w;
          v[j][4] = f1.x; v[j][5] = f1.y; v[j][6] = f1.z; v[j][7] = f1.w;
        } else {
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            const int cs = p.identity_channels ? (g * 8 + c) : p.channels[g * 8 + c];
            v[j][c] = static_cast<float>(px[cs]);
          }
        }
      }
    }
    float sh[8], sc[8];
#pragma unroll
    for (int c = 0; c < 8; ++c) { sh[c] = __ldg(shift + g * 8 + c); sc[c] = __ldg(scale + g * 8 + c); }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float o[8];
#pragma unroll
      for (int c = 0; c < 8; ++c) o[c] = __fmul_rn(sc[c], __fsub_rn(v[j][c], sh[c]));
      uint4 u;
      u.x = pack2(o[0], o[1]); u.y = pack2(o[2], o[3]); u.z = pack2(o[4], o[5]); u.w = pack2(o[6], o[7]);
      *reinterpret_cast<uint4*>(dst + out_index(p, b, y, xo + j) + g * 8) = u;
    }
  }
}

template <typename T>
static int launch_stage(const void* src, int layout, bool vec, const float* shift, const float* scale,
                        void* dst, const StageParams& p, cudaStream_t st) {
  const size_t total = size_t(p.B) * p.h * (p.w / 4) * (p.C / 8);
  int blocks = int(std::min<size_t>((total + 255) / 256, size_t(148) * 16));
  if (blocks < 1) blocks = 1;
  const T* s = static_cast<const T*>(src);
  __nv_bfloat16* d = static_cast<__nv_bfloat16*>(dst);
  if (layout == BGR_NCHW) {
    if (vec) BGR_CUDA_OK(launch_pdl(stage_kernel<T, BGR_NCHW, true>, dim3(blocks), dim3(256), 0, st, s, shift, scale, d, p));
    else BGR_CUDA_OK(launch_pdl(stage_kernel<T, BGR_NCHW, false>, dim3(blocks), dim3(256), 0, st, s, shift, scale, d, p));
  } else {
    if (vec) BGR_CUDA_OK(launch_pdl(stage_kernel<T, BGR_NHWC, true>, dim3(blocks), dim3(256), 0, st, s, shift, scale, d, p));
    else BGR_CUDA_OK(launch_pdl(stage_kernel<T, BGR_NHWC, false>, dim3(blocks), dim3(256), 0, st, s, shift, scale, d, p));
  }
  BGR_LAUNCH_OK();
  return BGR_OK;
}

}  // namespace bgr

using namespace bgr;

extern "C" int bgr_stage_normalize(const void* src, int src_dtype, int src_layout, int B, int Cf,
                                   int H, int W, const int* channels, int C, const float* shift,
                                   const float* scale, int y0, int x0, int h, int w, int flip_h,
                                   int flip_v, void* dst, int s2d, void* stream) {
  if (C < 8 || C % 8 || C > kMaxStageChannels)
    return fail(BGR_ERR_ARG, "stage: number of selected channels must be a multiple of 8 in [8,64]");
  if (w % 4 || h < 1) return fail(BGR_ERR_ARG, "stage: crop width must be a multiple of 4");
  if (s2d && ((h | w) & 1)) return fail(BGR_ERR_ARG, "stage: space-to-depth needs even crop size");
  if (y0 < 0 || x0 < 0 || y0 + h > H || x0 + w > W)
    return fail(BGR_ERR_ARG, "stage: crop window outside the field");
  if (uintptr_t(dst) & 15) return fail(BGR_ERR_ARG, "stage: dst must be 16-byte aligned");
  StageParams p{};
  p.B = B; p.Cf = Cf; p.H = H; p.W = W; p.C = C;
  p.y0 = y0; p.x0 = x0; p.h = h; p.w = w;
  p.flip_h = flip_h != 0; p.flip_v = flip_v != 0; p.s2d = s2d != 0;
  p.identity_channels = 1;
  for (int i = 0; i < C; ++i) {
    const int c = channels ? channels[i] : i;
    if (c < 0 || c >= Cf) return fail(BGR_ERR_ARG, "stage: channel index out of range");
    p.channels[i] = c;
    if (c != i) p.identity_channels = 0;
  }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  bool vec = src_dtype == BGR_F32 && !(uintptr_t(src) & 15);
  if (src_layout == BGR_NCHW) vec = vec && (W % 4 == 0) && (x0 % 4 == 0);
  else if (src_layout == BGR_NHWC) vec = vec && (Cf % 4 == 0) && p.identity_channels;
  else return fail(BGR_ERR_ARG, "stage: unknown source layout");
  switch (src_dtype) {
    case BGR_F32: return launch_stage<float>(src, src_layout, vec, shift, scale, dst, p, st);
    case BGR_F64: return launch_stage<double>(src, src_layout, false, shift, scale, dst, p, st);
    case BGR_I32: return launch_stage<int32_t>(src, src_layout, false, shift, scale, dst, p, st);
    case BGR_I64: return launch_stage<int64_t>(src, src_layout, false, shift, scale, dst, p, st);
  }
  return fail(BGR_ERR_ARG, "stage: unknown source dtype");
}
