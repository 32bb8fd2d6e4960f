// Implicit-GEMM convolution kernels for sm_100a: tcgen05.mma with TMEM accumulators, operands
// staged by TMA (cp.async.bulk.tensor) straight from bf16 NHWC activations.
//
//   fprop / dgrad :  Y[p, co] = sum_{t, ci} X[p + off(t), ci] * Wg[co, t, ci]
//                    A = activation tile (K-major, 128 pixels x 64 channels per stage, the pixel
//                    tile is a BW x BH rectangle so one 5-D TMA box with a per-tap coordinate
//                    offset fetches it; out-of-image taps are zero-filled by TMA = padding),
//                    B = weights [Cout][taps][Cin] (K-major).
//   wgrad         :  dW[co, t, ci] = sum_p dY[p, co] * X[p + off(t), ci]
//                    both operands MN-major (channels contiguous, pixels = GEMM-K), split over
//                    pixel chunks, fp32 atomics into dW.
//
// Replaces the cuDNN/cuBLAS call sites behind nn.Conv2d in the reference
// (src/deepCam/architecture/gpsro/deeplab.py:58,81,95,178,182,331,363,366,622,626,630).
#include "bgr_internal.h"
#include "tc_common.cuh"

namespace bgr {

constexpr int kBlockM = 128;   // pixels per fprop tile / Cout per wgrad tile
constexpr int kBlockK = 64;    // channels (fprop) or pixels (wgrad) per pipeline stage
constexpr int kMaxTaps = 9;

struct ConvParams {
  int N, H, W;              // output grid (pixels) = A-map spatial grid
  int BW, BH;               // pixel rectangle of one tile / chunk
  int tiles_x, tiles_y;
  int ntaps, kblocks;       // taps and 64-channel blocks per tap
  int Cout, Cin, ld_out;    // valid channels, output row stride (elements)
  int chunks_per_split;     // wgrad only
  int total_chunks;         // wgrad only
  int dy[kMaxTaps], dx[kMaxTaps];
};

// ------------------------------------------------------------------------------------------
// fprop: grid = (pixel tiles, Cout tiles), 192 threads: warp0 = TMA, warp1 = MMA, warps2-5 = epilogue
// ------------------------------------------------------------------------------------------
template <int BLOCK_N, int STAGES>
__global__ void __launch_bounds__(192, 1)
conv_fprop_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                  __nv_bfloat16* __restrict__ out, const float* __restrict__ bias,
                  const __grid_constant__ ConvParams p) {
  constexpr uint32_t A_BYTES = kBlockM * kBlockK * 2;
  constexpr uint32_t B_BYTES = BLOCK_N * kBlockK * 2;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + STAGES * A_BYTES;
  uint64_t* full = reinterpret_cast<uint64_t*>(sB + STAGES * B_BYTES);
  uint64_t* empty = full + STAGES;
  uint64_t* tmem_full = empty + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tile = blockIdx.x;
  const int tx = tile % p.tiles_x;
  const int ty = (tile / p.tiles_x) % p.tiles_y;
  const int n = tile / (p.tiles_x * p.tiles_y);
  const int n0 = blockIdx.y * BLOCK_N;
  const int x0 = tx * p.BW, y0 = ty * p.BH;
  const int iters = p.ntaps * p.kblocks;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    mbar_init(tmem_full, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<BLOCK_N>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      for (int it = 0; it < iters; ++it) {
        const int s = it % STAGES;
        const uint32_t ph = (it / STAGES) & 1;
        mbar_wait(&empty[s], ph ^ 1);
        const int t = it / p.kblocks, kb = it - t * p.kblocks;
        mbar_expect_tx(&full[s], A_BYTES + B_BYTES);
        tma_load_5d(sA + s * A_BYTES, &tmA, &full[s], kb * kBlockK, 0, x0 + p.dx[t],
                    y0 + p.dy[t], n);
        tma_load_3d(sB + s * B_BYTES, &tmB, &full[s], kb * kBlockK, t, n0);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(kBlockM, BLOCK_N, 0, 0);
      for (int it = 0; it < iters; ++it) {
        const int s = it % STAGES;
        const uint32_t ph = (it / STAGES) & 1;
        mbar_wait(&full[s], ph);
        tc_fence_after();
        const uint64_t da = umma_desc_sw128(smem_u32(sA + s * A_BYTES), 16, 1024);
        const uint64_t db = umma_desc_sw128(smem_u32(sB + s * B_BYTES), 16, 1024);
#pragma unroll
        for (int k = 0; k < kBlockK / 16; ++k)
          umma_bf16(tmem, da + uint64_t(k * 2), db + uint64_t(k * 2), idesc, (it | k) != 0);
        umma_commit(&empty[s]);
      }
      umma_commit(tmem_full);
    }
  } else {
    const int q = warp & 3;  // TMEM lane quadrant this warp may access
    mbar_wait(tmem_full, 0);
    tc_fence_after();
    const int row = q * 32 + lane;
    const int y = y0 + row / p.BW, x = x0 + row % p.BW;
    const bool row_ok = (y < p.H) && (x < p.W);
    __nv_bfloat16* orow = out + (size_t(size_t(n) * p.H + y) * p.W + x) * p.ld_out + n0;
#pragma unroll 1
    for (int c0 = 0; c0 < BLOCK_N; c0 += 16) {
      uint32_t v[16];
      tmem_ld16(tmem + (uint32_t(q * 32) << 16) + c0, v);
      tmem_ld_wait();
      if (row_ok) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int c = c0 + h * 8;
          if (n0 + c < p.Cout) {
            float f[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) f[j] = __uint_as_float(v[h * 8 + j]);
            if (bias != nullptr) {
#pragma unroll
              for (int j = 0; j < 8; ++j) f[j] += __ldg(bias + n0 + c + j);
            }
            uint4 o;
            o.x = pack_bf16(f[0], f[1]);
            o.y = pack_bf16(f[2], f[3]);
            o.z = pack_bf16(f[4], f[5]);
            o.w = pack_bf16(f[6], f[7]);
            *reinterpret_cast<uint4*>(orow + c) = o;
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<BLOCK_N>(tmem);
}

// ------------------------------------------------------------------------------------------
// wgrad: grid = (Cout tiles * Cin tiles * taps, splits).  Each CTA reduces its share of 64-pixel
// chunks into one 128 x BLOCK_N fp32 tile in TMEM and adds it to dW with fp32 atomics.
// ------------------------------------------------------------------------------------------
template <int BLOCK_N, int STAGES>
__global__ void __launch_bounds__(192, 1)
conv_wgrad_kernel(const __grid_constant__ CUtensorMap tmDY, const __grid_constant__ CUtensorMap tmX,
                  float* __restrict__ dw, const __grid_constant__ ConvParams p) {
  constexpr uint32_t BOX_BYTES = 64 * kBlockK * 2;           // one 64-channel x 64-pixel box
  constexpr uint32_t A_BYTES = (kBlockM / 64) * BOX_BYTES;   // dY^T : 128 couts
  constexpr uint32_t B_BYTES = (BLOCK_N / 64) * BOX_BYTES;   // X^T  : BLOCK_N cins
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + STAGES * A_BYTES;
  uint64_t* full = reinterpret_cast<uint64_t*>(sB + STAGES * B_BYTES);
  uint64_t* empty = full + STAGES;
  uint64_t* tmem_full = empty + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_tiles = (p.Cin + BLOCK_N - 1) / BLOCK_N;
  const int t = blockIdx.x % p.ntaps;
  const int nt = (blockIdx.x / p.ntaps) % n_tiles;
  const int mt = blockIdx.x / (p.ntaps * n_tiles);
  const int m0 = mt * kBlockM, n0 = nt * BLOCK_N;
  const int c_begin = blockIdx.y * p.chunks_per_split;
  const int c_end = min(c_begin + p.chunks_per_split, p.total_chunks);
  const int iters = max(c_end - c_begin, 0);

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmDY);
    tma_prefetch_desc(&tmX);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    mbar_init(tmem_full, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<BLOCK_N>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (iters > 0) {
    if (warp == 0) {
      if (lane == 0) {
        for (int it = 0; it < iters; ++it) {
          const int s = it % STAGES;
          const uint32_t ph = (it / STAGES) & 1;
          mbar_wait(&empty[s], ph ^ 1);
          const int c = c_begin + it;
          const int cx = c % p.tiles_x;
          const int cy = (c / p.tiles_x) % p.tiles_y;
          const int n = c / (p.tiles_x * p.tiles_y);
          const int x0 = cx * p.BW, y0 = cy * p.BH;
          mbar_expect_tx(&full[s], A_BYTES + B_BYTES);
#pragma unroll
          for (int b = 0; b < kBlockM / 64; ++b)
            tma_load_5d(sA + s * A_BYTES + b * BOX_BYTES, &tmDY, &full[s], m0 + b * 64, 0, x0, y0,
                        n);
#pragma unroll
          for (int b = 0; b < BLOCK_N / 64; ++b)
            tma_load_5d(sB + s * B_BYTES + b * BOX_BYTES, &tmX, &full[s], n0 + b * 64, 0,
                        x0 + p.dx[t], y0 + p.dy[t], n);
        }
      }
    } else if (warp == 1) {
      if (lane == 0) {
        constexpr uint32_t idesc = umma_idesc_bf16(kBlockM, BLOCK_N, 1, 1);
        for (int it = 0; it < iters; ++it) {
          const int s = it % STAGES;
          const uint32_t ph = (it / STAGES) & 1;
          mbar_wait(&full[s], ph);
          tc_fence_after();
          const uint64_t da = umma_desc_sw128(smem_u32(sA + s * A_BYTES), BOX_BYTES, 1024);
          const uint64_t db = umma_desc_sw128(smem_u32(sB + s * B_BYTES), BOX_BYTES, 1024);
#pragma unroll
          for (int k = 0; k < kBlockK / 16; ++k)  // 16 pixel rows = 2048 B per MMA
            umma_bf16(tmem, da + uint64_t(k * 128), db + uint64_t(k * 128), idesc, (it | k) != 0);
          umma_commit(&empty[s]);
        }
        umma_commit(tmem_full);
      }
    } else {
      const int q = warp & 3;
      mbar_wait(tmem_full, 0);
      tc_fence_after();
      const int co = m0 + q * 32 + lane;
      float* drow = dw + (size_t(co) * p.ntaps + t) * p.Cin + n0;
#pragma unroll 1
      for (int c0 = 0; c0 < BLOCK_N; c0 += 16) {
        uint32_t v[16];
        tmem_ld16(tmem + (uint32_t(q * 32) << 16) + c0, v);
        tmem_ld_wait();
        if (co < p.Cout) {
#pragma unroll
          for (int j = 0; j < 16; ++j)
            if (n0 + c0 + j < p.Cin) atomicAdd(drow + c0 + j, __uint_as_float(v[j]));
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<BLOCK_N>(tmem);
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
// Choose the BW x BH pixel rectangle (BW*BH == pixels) that wastes the fewest out-of-image pixels;
// ties go to the widest rectangle (longest contiguous runs in NHWC).
static void pick_rect(int H, int W, int pixels, int* BW, int* BH) {
  long best = -1;
  for (int bw = pixels; bw >= 1; bw >>= 1) {
    const int bh = pixels / bw;
    if (bw > 256 || bh > 256) continue;
    const long cover = long((W + bw - 1) / bw) * bw * long((H + bh - 1) / bh) * bh;
    if (best < 0 || cover < best) {
      best = cover;
      *BW = bw;
      *BH = bh;
    }
  }
}

// 5-D activation map (c, 1, x, y, n) over a bf16 NHWC view with pixel stride ld.
static int make_act_map(CUtensorMap* m, const void* base, int N, int H, int W, int C, int ld,
                        int BW, int BH) {
  uint64_t dims[5] = {uint64_t(C), 1, uint64_t(W), uint64_t(H), uint64_t(N)};
  uint64_t strides[4] = {uint64_t(ld) * 2, uint64_t(ld) * 2, uint64_t(W) * ld * 2,
                         uint64_t(H) * W * ld * 2};
  uint32_t box[5] = {64, 1, uint32_t(BW), uint32_t(BH), 1};
  return encode_tmap_bf16(m, base, 5, dims, strides, box);
}

static int check_conv_args(const void* x, const void* w, const void* y, int Cin, int ldx, int Cout,
                           int ldy, int ntaps) {
  if (ntaps < 1 || ntaps > kMaxTaps) return fail(BGR_ERR_ARG, "conv: ntaps must be in [1,9]");
  if ((Cin % 8) || (Cout % 8) || (ldx % 8) || (ldy % 8))
    return fail(BGR_ERR_ARG, "conv: channel counts and pixel strides must be multiples of 8");
  if ((uintptr_t(x) | uintptr_t(w) | uintptr_t(y)) & 15)
    return fail(BGR_ERR_ARG, "conv: pointers must be 16-byte aligned");
  return BGR_OK;
}

template <int BLOCK_N, int STAGES>
static int launch_fprop(const CUtensorMap& tmA, const CUtensorMap& tmB, void* y, const float* bias,
                        const ConvParams& p, cudaStream_t st) {
  constexpr size_t smem = size_t(STAGES) * (kBlockM * kBlockK * 2 + BLOCK_N * kBlockK * 2) +
                          (2 * STAGES + 1) * 8 + 16 + 1024;
  auto kern = conv_fprop_kernel<BLOCK_N, STAGES>;
  static bool configured = false;
  if (!configured) {
    BGR_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    configured = true;
  }
  dim3 grid(p.N * p.tiles_x * p.tiles_y, (p.Cout + BLOCK_N - 1) / BLOCK_N);
  kern<<<grid, 192, smem, st>>>(tmA, tmB, reinterpret_cast<__nv_bfloat16*>(y), bias, p);
  BGR_LAUNCH_OK();
  return BGR_OK;
}

template <int BLOCK_N, int STAGES>
static int launch_wgrad(const CUtensorMap& tmDY, const CUtensorMap& tmX, float* dw,
                        ConvParams& p, cudaStream_t st) {
  constexpr size_t smem = size_t(STAGES) * ((kBlockM + BLOCK_N) * kBlockK * 2) +
                          (2 * STAGES + 1) * 8 + 16 + 1024;
  auto kern = conv_wgrad_kernel<BLOCK_N, STAGES>;
  static bool configured = false;
  if (!configured) {
    BGR_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    configured = true;
  }
  const int tiles = ((p.Cout + kBlockM - 1) / kBlockM) * ((p.Cin + BLOCK_N - 1) / BLOCK_N) * p.ntaps;
  // split the pixel reduction so that the grid covers the 148 SMs about twice
  int splits = (2 * 148 + tiles - 1) / tiles;
  if (splits > p.total_chunks) splits = p.total_chunks;
  if (splits < 1) splits = 1;
  p.chunks_per_split = (p.total_chunks + splits - 1) / splits;
  splits = (p.total_chunks + p.chunks_per_split - 1) / p.chunks_per_split;
  dim3 grid(tiles, splits);
  kern<<<grid, 192, smem, st>>>(tmDY, tmX, dw, p);
  BGR_LAUNCH_OK();
  return BGR_OK;
}

}  // namespace bgr

using namespace bgr;

extern "C" int bgr_conv_fprop(const void* x, int N, int H, int W, int Cin, int ldx, const void* w,
                              int Cout, int ntaps, const int* dy, const int* dx, void* y, int ldy,
                              const float* bias, void* stream) {
  if (int rc = check_conv_args(x, w, y, Cin, ldx, Cout, ldy, ntaps)) return rc;
  ConvParams p{};
  p.N = N; p.H = H; p.W = W;
  pick_rect(H, W, kBlockM, &p.BW, &p.BH);
  p.tiles_x = (W + p.BW - 1) / p.BW;
  p.tiles_y = (H + p.BH - 1) / p.BH;
  p.ntaps = ntaps;
  p.kblocks = (Cin + kBlockK - 1) / kBlockK;
  p.Cout = Cout; p.Cin = Cin; p.ld_out = ldy;
  for (int t = 0; t < ntaps; ++t) { p.dy[t] = dy[t]; p.dx[t] = dx[t]; }
  CUtensorMap tmA, tmB;
  if (int rc = make_act_map(&tmA, x, N, H, W, Cin, ldx, p.BW, p.BH)) return rc;
  const int bn = Cout <= 64 ? 64 : 128;
  {
    uint64_t dims[3] = {uint64_t(Cin), uint64_t(ntaps), uint64_t(Cout)};
    uint64_t strides[2] = {uint64_t(Cin) * 2, uint64_t(Cin) * ntaps * 2};
    uint32_t box[3] = {64, 1, uint32_t(bn)};
    if (int rc = encode_tmap_bf16(&tmB, w, 3, dims, strides, box)) return rc;
  }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (bn == 64) return launch_fprop<64, 6>(tmA, tmB, y, bias, p, st);
  return launch_fprop<128, 6>(tmA, tmB, y, bias, p, st);
}

extern "C" int bgr_conv_wgrad(const void* x, int N, int H, int W, int Cin, int ldx,
                              const void* dy_, int Cout, int ldy, int ntaps, const int* dy,
                              const int* dx, float* dw, void* stream) {
  if (int rc = check_conv_args(x, dw, dy_, Cin, ldx, Cout, ldy, ntaps)) return rc;
  ConvParams p{};
  p.N = N; p.H = H; p.W = W;
  pick_rect(H, W, kBlockK, &p.BW, &p.BH);
  p.tiles_x = (W + p.BW - 1) / p.BW;
  p.tiles_y = (H + p.BH - 1) / p.BH;
  p.total_chunks = N * p.tiles_x * p.tiles_y;
  p.ntaps = ntaps;
  p.Cout = Cout; p.Cin = Cin;
  for (int t = 0; t < ntaps; ++t) { p.dy[t] = dy[t]; p.dx[t] = dx[t]; }
  CUtensorMap tmDY, tmX;
  if (int rc = make_act_map(&tmDY, dy_, N, H, W, Cout, ldy, p.BW, p.BH)) return rc;
  if (int rc = make_act_map(&tmX, x, N, H, W, Cin, ldx, p.BW, p.BH)) return rc;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (Cin <= 64) return launch_wgrad<64, 6>(tmDY, tmX, dw, p, st);
  return launch_wgrad<128, 6>(tmDY, tmX, dw, p, st);
}
