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
//                    pixel chunks; every split writes its own fp32 slab and the slabs are added in
//                    split order (no atomics: the result does not depend on scheduling).
//
// Replaces the cuDNN/cuBLAS call sites behind nn.Conv2d in the reference
// (src/deepCam/architecture/gpsro/deeplab.py:58,81,95,178,182,331,363,366,622,626,630).
#include "bgr_internal.h"
#include "tc_common.cuh"

static int g_force_block_n = 0;   // development switches (bgr_debug_conv_*)
static int g_conv_dbg = 0;
static int g_force_ctas = -1;     // -1 = not read yet (BGR_CONV_CTAS), 0 = pick, 1 / 2 = forced
static int g_force_ctas_read() {
  if (g_force_ctas < 0) {
    const char* e = getenv("BGR_CONV_CTAS");
    g_force_ctas = e ? atoi(e) : 0;
  }
  return g_force_ctas;
}

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
  int m_tiles, n_tiles;     // fprop only: pixel tiles, Cout tiles
  int dbg;                  // development switch (0 in production)
  int chunks_per_split;     // wgrad only
  int total_chunks;         // wgrad only
  long slab_elems;          // wgrad only: Cout*ntaps*Cin (distance between the slabs of two splits)
  int dy[kMaxTaps], dx[kMaxTaps];
};

// ------------------------------------------------------------------------------------------
// fprop: persistent, one CTA per SM, tiles (pixel tile, Cout tile) round-robin with the Cout
// tile fastest (neighbouring CTAs share the activation tile through L2).  192 threads:
//   warp 0    TMA producer   (smem ring of STAGES x {A 128x64, B BLOCK_Nx64}, runs across tiles)
//   warp 1    tcgen05.mma issuer, accumulators double-buffered in TMEM (2 x BLOCK_N columns)
//   warps 2-5 epilogue: tcgen05.ld -> (+bias) -> bf16 -> 128B-swizzled smem -> TMA store in
//             64-channel chunks (two staging buffers), overlapped with the next tile's MMAs.
//             Optional fused BatchNorm statistics: per-channel sum / sum-of-squares of the
//             bf16-rounded outputs, read back from the staged tile (column sums), kept in registers
//             for the CTA's life; at the end the four warps are added in warp order and the CTA writes
//             one partial row (its tiles are a fixed function of blockIdx, so the row is reproducible).
// ------------------------------------------------------------------------------------------
// Development build only (-DBGR_CONV_TIMING, tools/dev_conv_timing.py): where each role of the forward kernel waits.
#ifdef BGR_CONV_TIMING
__device__ long long g_conv_timing[512][8];
#define BGR_T0(v) const long long v = clock64()
#define BGR_TACC(slot, v) g_conv_timing[blockIdx.y * gridDim.x + blockIdx.x][slot] += clock64() - (v)
#else
#define BGR_T0(v)
#define BGR_TACC(slot, v)
#endif

// CTAS = 2: the CTA pair variant (cta_group::2).  The two CTAs of a cluster take two neighbouring pixel tiles and
// the same Cout tile; each stages its own 128 pixels and HALF of the Cout tile's weights (BLOCK_N/2 rows), the
// leader issues one 256 x BLOCK_N MMA over both shared memories, and each CTA drains its own 128 accumulator rows.
// Per k-block a CTA then receives 16 KB + BLOCK_N/2 x 128 B instead of 16 KB + BLOCK_N x 128 B: the 1-CTA kernel is
// bound by what one SM can pull from L2 (profiles/r02_conv_roles.txt), not by the tensor pipe.
// The pair variant also runs two sets of four epilogue warps (320 threads): one warp per scheduler drains a
// 128 x 256 tile in ~10 k cycles (instruction latency; measured with -DBGR_CONV_TIMING), longer than the tile's
// 6144 tensor-pipe cycles at K = 768; set h takes the 64-channel chunks c with c % 2 == h, with its own pair of
// staging buffers and its own bulk-store groups.
template <int BLOCK_N, int STAGES, bool STATS, int CTAS = 1>
__global__ void __launch_bounds__(64 + 128 * CTAS, 1)
conv_fprop_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                  const __grid_constant__ CUtensorMap tmO, const float* __restrict__ bias,
                  float* __restrict__ stats /* [CTA groups][2][Cout] */, const __grid_constant__ ConvParams p) {
  constexpr uint32_t A_BYTES = kBlockM * kBlockK * 2;
  constexpr uint32_t B_BYTES = (BLOCK_N / CTAS) * kBlockK * 2;   // this CTA's share of the weight tile
  constexpr uint32_t O_BYTES = kBlockM * 64 * 2;       // one 64-channel output chunk
  constexpr uint32_t TMEM_COLS = BLOCK_N > 128 ? 512 : 2 * BLOCK_N;   // allocations are powers of two
  constexpr int SETS = CTAS;                           // epilogue warp sets (4 warps each)
  constexpr int NCHUNK = BLOCK_N / 64;
  constexpr int MYCH = (NCHUNK + SETS - 1) / SETS;     // chunks one set handles per tile
  // staging buffers per set: two (write one while the other's store drains), or one for the pair variant, which
  // spends the shared memory on ring depth instead (its operand feed is latency-bound: 3/4/5 stages wait 58/50/40 %
  // of the MMA role, profiles/r02_conv_roles.txt) and waits for the previous chunk's store after the TMEM read
  constexpr int OBUFS = CTAS == 2 ? 1 : 2;
  constexpr uint32_t NEEDED = STAGES * (A_BYTES + B_BYTES) + SETS * OBUFS * O_BYTES + (2 * STAGES + 4) * 8 + 16;
  static_assert(4 * 2 * BLOCK_N * 4 <= SETS * OBUFS * O_BYTES, "the statistics combine reuses the output staging buffers");
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~uintptr_t(1023));
  if (threadIdx.x == 0 && (smem - smem_raw) + NEEDED > dynamic_smem_size()) {
    printf("bgr: conv_fprop shared memory window too small\n");
    __trap();
  }
  uint8_t* sA = smem;
  uint8_t* sB = sA + STAGES * A_BYTES;
  uint8_t* sO = sB + STAGES * B_BYTES;
  uint64_t* full = reinterpret_cast<uint64_t*>(sO + SETS * OBUFS * O_BYTES);
  uint64_t* empty = full + STAGES;
  uint64_t* tfull = empty + STAGES;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int iters = p.ntaps * p.kblocks;
  // gridDim.x is a multiple of n_tiles: this CTA keeps one Cout tile (so its BatchNorm partial sums
  // stay in shared memory for the whole kernel) and strides over the pixel tiles; the n_tiles CTAs
  // of one group work on the same activation tile at the same time (shared through L2).
  // CTA pairs: unit = the pair, which takes pixel tiles (2k, 2k+1); both CTAs walk the same number of tiles (the
  // odd one out is a phantom tile past the last image: its loads are zero-filled, its results dropped).
  const uint32_t rank = CTAS == 2 ? cluster_ctarank() : 0u;
  const int unit = blockIdx.x / CTAS, units = gridDim.x / CTAS;
  const int nt = unit % p.n_tiles;
  const int mt_first = (unit / p.n_tiles) * CTAS + int(rank), mt_step = (units / p.n_tiles) * CTAS;
  const int mt_end = p.m_tiles + int(rank);   // loop bound: the pair's first tile is inside
  const int n0 = nt * BLOCK_N;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmO);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&tfull[b], 1);
      mbar_init(&tempty[b], 4 * SETS * CTAS);   // the leader's barrier collects the epilogue warps of both CTAs
    }
    fence_barrier_init();
  }
  if (warp == 1) {
    if (CTAS == 2) tmem_alloc2<TMEM_COLS>(tmem_slot);
    else tmem_alloc<TMEM_COLS>(tmem_slot);
  }
  tc_fence_before();
  if (CTAS == 2) cluster_sync_all();   // the peer's barriers exist before anything signals them
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  // everything above (barrier init, descriptor prefetch, TMEM allocation) touched no global memory: it may
  // overlap the tail of the previous kernel (programmatic dependent launch, bgr_internal.h)
  pdl_wait();
  pdl_trigger();

  if (warp == 0) {
    if (lane == 0) {
      uint32_t it = 0;
      BGR_T0(t_role);
      const uint32_t full0 = CTAS == 2 ? mapa_shared(smem_u32(&full[0]), 0) : 0u;   // the leader's full[] barriers
      for (int mt = mt_first; mt < mt_end; mt += mt_step) {
        const int tx = mt % p.tiles_x;
        const int ty = (mt / p.tiles_x) % p.tiles_y;
        const int n = mt / (p.tiles_x * p.tiles_y);
        const int x0 = tx * p.BW, y0 = ty * p.BH;
        for (int t = 0; t < p.ntaps; ++t) {
          for (int kb = 0; kb < p.kblocks; ++kb, ++it) {
            const uint32_t s = it % STAGES, ph = (it / STAGES) & 1;
            {
              BGR_T0(t_w);
              mbar_wait(&empty[s], ph ^ 1);
              BGR_TACC(0, t_w);
            }
            if (CTAS == 2) {
              if (rank == 0) mbar_expect_tx(&full[s], 2 * (A_BYTES + B_BYTES));
              tma_load_5d_2sm(sA + s * A_BYTES, &tmA, full0 + s * 8, kb * kBlockK, 0, x0 + p.dx[t],
                              y0 + p.dy[t], n);
              tma_load_3d_2sm(sB + s * B_BYTES, &tmB, full0 + s * 8, kb * kBlockK, t,
                              n0 + int(rank) * (BLOCK_N / 2));
              continue;
            }
            mbar_expect_tx(&full[s], A_BYTES + B_BYTES);
            tma_load_5d(sA + s * A_BYTES, &tmA, &full[s], kb * kBlockK, 0, x0 + p.dx[t],
                        y0 + p.dy[t], n);
            tma_load_3d(sB + s * B_BYTES, &tmB, &full[s], kb * kBlockK, t, n0);
          }
        }
      }
      BGR_TACC(1, t_role);
    }
  } else if (warp == 1) {
    if (CTAS == 1 || rank == 0) {
    // The whole warp walks the loop (uniform control flow); one elected lane issues.  Per k-block
    // the issuing thread has ~512 tensor-pipe cycles of work to hide behind, so the loop carries
    // running (stage, phase) counters and descriptor low words instead of divisions and rebuilds.
    constexpr uint32_t idesc = umma_idesc_bf16(kBlockM * CTAS, BLOCK_N, 0, 0);
    const uint64_t da0 = umma_desc_sw128(smem_u32(sA), 16, 1024), db0 = umma_desc_sw128(smem_u32(sB), 16, 1024);
    const uint32_t a_hi = uint32_t(da0 >> 32), b_hi = uint32_t(db0 >> 32);
    const uint32_t a_lo0 = uint32_t(da0), b_lo0 = uint32_t(db0);
    uint32_t s = 0, ph = 0, lt = 0;
    BGR_T0(t_role);
    for (int mt = mt_first; mt < mt_end; mt += mt_step, ++lt) {
      const uint32_t acc = lt & 1, aph = (lt >> 1) & 1;
      {
        BGR_T0(t_w);
        if (CTAS == 2) mbar_wait_cluster(&tempty[acc], aph ^ 1);
        else mbar_wait(&tempty[acc], aph ^ 1);
        if (lane == 0) BGR_TACC(2, t_w);
      }
      tc_fence_after();
      const uint32_t d = tmem + acc * BLOCK_N;
      for (int i = 0; i < iters; ++i) {
        {
          BGR_T0(t_w);
          mbar_wait(&full[s], ph);
          if (lane == 0) BGR_TACC(3, t_w);
        }
        tc_fence_after();
        if (elect_one_sync()) {
          const uint32_t al = a_lo0 + s * (A_BYTES >> 4), bl = b_lo0 + s * (B_BYTES >> 4);
#pragma unroll
          for (int k = 0; k < kBlockK / 16; ++k) {
            if (CTAS == 2) umma_bf16_w2(d, al + k * 2, a_hi, bl + k * 2, b_hi, idesc, k ? 1u : uint32_t(i != 0));
            else umma_bf16_w(d, al + k * 2, a_hi, bl + k * 2, b_hi, idesc, k ? 1u : uint32_t(i != 0));
          }
          if (CTAS == 2) umma_commit2(&empty[s]);
          else umma_commit(&empty[s]);
        }
        __syncwarp();
        if (++s == STAGES) { s = 0; ph ^= 1; }
      }
      if (elect_one_sync()) {
        if (CTAS == 2) umma_commit2(&tfull[acc]);
        else umma_commit(&tfull[acc]);
      }
      __syncwarp();
    }
    if (lane == 0) BGR_TACC(4, t_role);
    }
  } else {
    const int q = warp & 3;                  // TMEM lane quadrant this warp may access
    const int row = q * 32 + lane;           // tile row = pixel (row % BW, row / BW) of the rectangle
    const int set = (warp - 2) >> 2;         // epilogue warp set (0 when there is one)
    const int epi_tid = (threadIdx.x - 64) & 127;   // 0..127 within the set
    uint8_t* const sOset = sO + set * OBUFS * O_BYTES;
    uint8_t* my_row = nullptr;
    uint32_t lt = 0, chunk_ctr = 0;
    // BatchNorm partial sums of this CTA's Cout tile over this warp's rows, for the chunks of this warp's set:
    // st_*[i*2 + e] <-> channel (i*SETS + set)*64 + 2*lane + e
    float st_sum[2 * MYCH], st_sq[2 * MYCH];
#pragma unroll
    for (int i = 0; i < 2 * MYCH; ++i) st_sum[i] = st_sq[i] = 0.f;
    const uint32_t tempty0 = CTAS == 2 ? mapa_shared(smem_u32(&tempty[0]), 0) : 0u;
    BGR_T0(t_role);
    for (int mt = mt_first; mt < mt_end; mt += mt_step, ++lt) {
      const int tx = mt % p.tiles_x;
      const int ty = (mt / p.tiles_x) % p.tiles_y;
      const int n = mt / (p.tiles_x * p.tiles_y);
      const int x0 = tx * p.BW, y0 = ty * p.BH;
      const bool tile_full = (y0 + p.BH <= p.H) && (x0 + p.BW <= p.W);
      const uint32_t acc = lt & 1, aph = (lt >> 1) & 1;
      {
        BGR_T0(t_w);
        mbar_wait(&tfull[acc], aph);
        if (threadIdx.x == 64) BGR_TACC(5, t_w);
      }
      tc_fence_after();
#pragma unroll
      for (int ci = 0; ci < MYCH; ++ci) {
        const int c = ci * SETS + set;
        if (c >= NCHUNK) break;
        const int ch0 = n0 + c * 64;
        if (ch0 >= p.Cout) break;
        if (CTAS == 2 && mt >= p.m_tiles) break;   // phantom tile of an odd tile count
        if (p.dbg == 2) break;   // timing experiments only (bgr_debug_conv_mode): no epilogue at all
        uint8_t* buf = sOset + (OBUFS == 2 ? (chunk_ctr & 1) * O_BYTES : 0);
        if (OBUFS == 2) {
          BGR_T0(t_w);
          if (epi_tid == 0) bulk_wait_group_read<1>();   // the store issued two chunks ago has left smem
          named_bar_sync(1 + set, 128);
          if (threadIdx.x == 64) BGR_TACC(6, t_w);
        }
        uint32_t v[64];
        const uint32_t taddr = tmem + (uint32_t(q * 32) << 16) + acc * BLOCK_N + c * 64;
        tmem_ld32(taddr, v);
        tmem_ld32(taddr + 32, v + 32);
        tmem_ld_wait();
        if (OBUFS == 1) {   // the previous chunk's store has left the buffer (and every warp its statistics pass)
          BGR_T0(t_w);
          if (epi_tid == 0) bulk_wait_group_read<0>();
          named_bar_sync(1 + set, 128);
          if (threadIdx.x == 64) BGR_TACC(6, t_w);
        }
        if (p.dbg == 3) { ++chunk_ctr; continue; }   // TMEM reads only
        if (bias != nullptr) {
#pragma unroll
          for (int j = 0; j < 64; ++j)
            if (ch0 + j < p.Cout) v[j] = __float_as_uint(__uint_as_float(v[j]) + __ldg(bias + ch0 + j));
        }
        my_row = buf + row * 128;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          uint4 o;
          o.x = pack_bf16(__uint_as_float(v[j * 8 + 0]), __uint_as_float(v[j * 8 + 1]));
          o.y = pack_bf16(__uint_as_float(v[j * 8 + 2]), __uint_as_float(v[j * 8 + 3]));
          o.z = pack_bf16(__uint_as_float(v[j * 8 + 4]), __uint_as_float(v[j * 8 + 5]));
          o.w = pack_bf16(__uint_as_float(v[j * 8 + 6]), __uint_as_float(v[j * 8 + 7]));
          *reinterpret_cast<uint4*>(my_row + ((j ^ (row & 7)) << 4)) = o;
        }
        fence_proxy_async();
        named_bar_sync(1 + set, 128);
        if (epi_tid == 0 && p.dbg != 4) {   // dbg 4: everything but the TMA store
          tma_store_5d(&tmO, buf, ch0, 0, x0, y0, n);
          bulk_commit_group();
        }
        if (STATS && p.dbg != 1) {
          // Column sums straight from the staged bf16 tile (already the rounded values): warp q sums
          // its 32 rows, lane l the channel pair (2l, 2l+1) = one 4-byte word per row; within a row the
          // 32 lanes read 32 distinct banks.  (A register butterfly over the fp32 accumulators cost 3x
          // the instructions: FSEL 26 %, SHFL 10 % of the kernel.)
          const uint8_t* col = buf + ((lane & 3) << 2);
          const int chunk16 = lane >> 2;
          float s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;
          if (tile_full) {   // straight-line body: the 32 loads pipeline
#pragma unroll
            for (int rr = 0; rr < 32; ++rr) {
              const int r = q * 32 + rr;
              const uint32_t wv = *reinterpret_cast<const uint32_t*>(col + r * 128 + ((chunk16 ^ (r & 7)) << 4));
              const float f0 = __uint_as_float(wv << 16), f1 = __uint_as_float(wv & 0xffff0000u);
              s0 += f0; s1 += f1;
              q0 = fmaf(f0, f0, q0); q1 = fmaf(f1, f1, q1);
            }
          } else {           // ragged tile: rows outside the image do not count
            for (int rr = 0; rr < 32; ++rr) {
              const int r = q * 32 + rr;
              const int yy = r / p.BW, xx = r - yy * p.BW;
              if (y0 + yy >= p.H || x0 + xx >= p.W) continue;
              const uint32_t wv = *reinterpret_cast<const uint32_t*>(col + r * 128 + ((chunk16 ^ (r & 7)) << 4));
              const float f0 = __uint_as_float(wv << 16), f1 = __uint_as_float(wv & 0xffff0000u);
              s0 += f0; s1 += f1;
              q0 = fmaf(f0, f0, q0); q1 = fmaf(f1, f1, q1);
            }
          }
          st_sum[ci * 2 + 0] += s0; st_sum[ci * 2 + 1] += s1;
          st_sq[ci * 2 + 0] += q0;  st_sq[ci * 2 + 1] += q1;
        }
        ++chunk_ctr;
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (CTAS == 2) mbar_arrive_cluster(tempty0 + acc * 8);
        else mbar_arrive(&tempty[acc]);
      }
    }
    if (epi_tid == 0) bulk_wait_group_read<0>();
    if (threadIdx.x == 64) BGR_TACC(7, t_role);
    if (STATS) {
      // The four warps' sums are added in warp order through the (now idle) output staging buffers and the
      // CTA writes partial row mt_first of the [groups][2][Cout] workspace: no atomics, a fixed summation order.
      // (every set's stores have left its staging buffers: all sets meet here before set 0's buffers are reused)
      named_bar_sync(3, 128 * SETS);
      float* comb = reinterpret_cast<float*>(sO);   // [4 warps][2][BLOCK_N]
#pragma unroll
      for (int i = 0; i < 2 * MYCH; ++i) {
        const int c = (i >> 1) * SETS + set;
        if (c >= NCHUNK) break;
        const int ch = c * 64 + 2 * lane + (i & 1);
        comb[q * 2 * BLOCK_N + ch] = st_sum[i];
        comb[q * 2 * BLOCK_N + BLOCK_N + ch] = st_sq[i];
      }
      named_bar_sync(3, 128 * SETS);
      float* prow = stats + size_t(mt_first) * 2 * p.Cout;
      for (int i = threadIdx.x - 64; i < 2 * BLOCK_N; i += 128 * SETS) {
        const int half = i / BLOCK_N, ch = i - half * BLOCK_N;
        if (n0 + ch < p.Cout)
          prow[half * p.Cout + n0 + ch] = ((comb[i] + comb[2 * BLOCK_N + i]) + comb[4 * BLOCK_N + i]) + comb[6 * BLOCK_N + i];
      }
    }
  }
  tc_fence_before();
  if (CTAS == 2) {
    cluster_sync_all();   // neither CTA leaves while the pair's MMAs or remote arrivals may still touch it
    if (warp == 1) tmem_dealloc2<TMEM_COLS>(tmem);
    return;
  }
  __syncthreads();
  if (warp == 1) tmem_dealloc<TMEM_COLS>(tmem);
}

// ------------------------------------------------------------------------------------------
// wgrad: grid = (Cout tiles * Cin tiles * taps, splits).  Each CTA reduces its share of 64-pixel
// chunks into one 128 x BLOCK_N fp32 tile in TMEM and stores it into the slab of its split
// (slab 0 = dW itself when there is one split); sum_rows adds the slabs in split order.
// ------------------------------------------------------------------------------------------
// CTAS = 2 (CTA pair, see the forward kernel): the pair owns 256 couts x BLOCK_N cins; each CTA stages its 128
// couts of dY^T and half of the cins of X^T, and drains its own 128 accumulator rows.
template <int BLOCK_N, int STAGES, int CTAS = 1>
__global__ void __launch_bounds__(192, 1)
conv_wgrad_kernel(const __grid_constant__ CUtensorMap tmDY, const __grid_constant__ CUtensorMap tmX,
                  float* __restrict__ dw, const __grid_constant__ ConvParams p) {
  constexpr uint32_t BOX_BYTES = 64 * kBlockK * 2;           // one 64-channel x 64-pixel box
  constexpr uint32_t A_BYTES = (kBlockM / 64) * BOX_BYTES;   // dY^T : 128 couts
  constexpr uint32_t B_BYTES = (BLOCK_N / CTAS / 64) * BOX_BYTES;   // X^T  : this CTA's share of BLOCK_N cins
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
  const uint32_t rank = CTAS == 2 ? cluster_ctarank() : 0u;
  const int unit = blockIdx.x / CTAS;   // the tile (of a CTA or of a pair)
  const int t = unit % p.ntaps;
  const int nt = (unit / p.ntaps) % n_tiles;
  const int mt = unit / (p.ntaps * n_tiles);
  const int m0 = (mt * CTAS + int(rank)) * kBlockM, n0 = nt * BLOCK_N;   // this CTA's couts; the tile's cins
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
  if (warp == 1) {
    if (CTAS == 2) tmem_alloc2<BLOCK_N>(tmem_slot);
    else tmem_alloc<BLOCK_N>(tmem_slot);
  }
  tc_fence_before();
  if (CTAS == 2) cluster_sync_all();
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  // everything above (barrier init, descriptor prefetch, TMEM allocation) touched no global memory: it may
  // overlap the tail of the previous kernel (programmatic dependent launch, bgr_internal.h)
  pdl_wait();
  pdl_trigger();

  if (iters > 0) {
    if (warp == 0) {
      if (lane == 0) {
        const uint32_t full0 = CTAS == 2 ? mapa_shared(smem_u32(&full[0]), 0) : 0u;   // the leader's full[] barriers
        BGR_T0(t_role);
        for (int it = 0; it < iters; ++it) {
          const int s = it % STAGES;
          const uint32_t ph = (it / STAGES) & 1;
          {
            BGR_T0(t_w);
            mbar_wait(&empty[s], ph ^ 1);
            BGR_TACC(0, t_w);
          }
          const int c = c_begin + it;
          const int cx = c % p.tiles_x;
          const int cy = (c / p.tiles_x) % p.tiles_y;
          const int n = c / (p.tiles_x * p.tiles_y);
          const int x0 = cx * p.BW, y0 = cy * p.BH;
          if (CTAS == 2) {
            if (rank == 0) mbar_expect_tx(&full[s], 2 * (A_BYTES + B_BYTES));
#pragma unroll
            for (int b = 0; b < kBlockM / 64; ++b)
              tma_load_5d_2sm(sA + s * A_BYTES + b * BOX_BYTES, &tmDY, full0 + s * 8, m0 + b * 64, 0, x0, y0, n);
#pragma unroll
            for (int b = 0; b < BLOCK_N / 2 / 64; ++b)
              tma_load_5d_2sm(sB + s * B_BYTES + b * BOX_BYTES, &tmX, full0 + s * 8,
                              n0 + int(rank) * (BLOCK_N / 2) + b * 64, 0, x0 + p.dx[t], y0 + p.dy[t], n);
            continue;
          }
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
        BGR_TACC(1, t_role);
      }
    } else if (warp == 1) {
      if (CTAS == 1 || rank == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(kBlockM * CTAS, BLOCK_N, 1, 1);
      const uint64_t da0 = umma_desc_sw128(smem_u32(sA), BOX_BYTES, 1024), db0 = umma_desc_sw128(smem_u32(sB), BOX_BYTES, 1024);
      const uint32_t a_hi = uint32_t(da0 >> 32), b_hi = uint32_t(db0 >> 32);
      const uint32_t a_lo0 = uint32_t(da0), b_lo0 = uint32_t(db0);
      uint32_t s = 0, ph = 0;
      BGR_T0(t_role);
      for (int it = 0; it < iters; ++it) {   // warp-uniform; one elected lane issues (see fprop)
        {
          BGR_T0(t_w);
          mbar_wait(&full[s], ph);
          if (lane == 0) BGR_TACC(3, t_w);
        }
        tc_fence_after();
        if (elect_one_sync()) {
          const uint32_t al = a_lo0 + s * (A_BYTES >> 4), bl = b_lo0 + s * (B_BYTES >> 4);
#pragma unroll
          for (int k = 0; k < kBlockK / 16; ++k) {  // 16 pixel rows = 2048 B per MMA
            if (CTAS == 2) umma_bf16_w2(tmem, al + k * 128, a_hi, bl + k * 128, b_hi, idesc, k ? 1u : uint32_t(it != 0));
            else umma_bf16_w(tmem, al + k * 128, a_hi, bl + k * 128, b_hi, idesc, k ? 1u : uint32_t(it != 0));
          }
          if (CTAS == 2) umma_commit2(&empty[s]);
          else umma_commit(&empty[s]);
        }
        __syncwarp();
        if (++s == STAGES) { s = 0; ph ^= 1; }
      }
      if (elect_one_sync()) {
        if (CTAS == 2) umma_commit2(tmem_full);
        else umma_commit(tmem_full);
      }
      __syncwarp();
      if (lane == 0) BGR_TACC(4, t_role);
      }
    } else {
      const int q = warp & 3;
      BGR_T0(t_role);
      mbar_wait(tmem_full, 0);
      if (threadIdx.x == 64) BGR_TACC(5, t_role);
      tc_fence_after();
      const int co = m0 + q * 32 + lane;
      float* drow = dw + size_t(blockIdx.y) * p.slab_elems + (size_t(co) * p.ntaps + t) * p.Cin + n0;
#pragma unroll 1
      for (int c0 = 0; c0 < BLOCK_N; c0 += 16) {
        uint32_t v[16];
        tmem_ld16(tmem + (uint32_t(q * 32) << 16) + c0, v);
        tmem_ld_wait();
        if (co < p.Cout) {
#pragma unroll
          for (int j = 0; j < 16; j += 4)   // Cin % 8 == 0: a group of 4 is entirely in or out
            if (n0 + c0 + j < p.Cin)
              *reinterpret_cast<float4*>(drow + c0 + j) = make_float4(__uint_as_float(v[j]), __uint_as_float(v[j + 1]),
                                                                      __uint_as_float(v[j + 2]), __uint_as_float(v[j + 3]));
        }
      }
      if (threadIdx.x == 64) BGR_TACC(7, t_role);
    }
  }
  tc_fence_before();
  if (CTAS == 2) {
    cluster_sync_all();
    if (warp == 1) tmem_dealloc2<BLOCK_N>(tmem);
    return;
  }
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

constexpr size_t kMaxDynSmem = 232448;  // 227 KB per CTA on sm_100

// Launch with clusters of two CTAs along x (CTA pairs); gemm-class programmatic dependent launch as launch_pdl<true>.
template <typename... KArgs, typename... Args>
static cudaError_t launch_pairs(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = pdl_enabled(true) ? 1 : 0;
  cfg.attrs = attr;
  cfg.numAttrs = 2;
  return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

template <int BLOCK_N, int STAGES, bool STATS, int CTAS = 1>
static int launch_fprop(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmO,
                        const float* bias, float* stats, const ConvParams& p, int groups, cudaStream_t st) {
  constexpr size_t needed = size_t(STAGES) * (kBlockM * kBlockK * 2 + (BLOCK_N / CTAS) * kBlockK * 2) +
                            2 * (kBlockM * 64 * 2) + (2 * STAGES + 4) * 8 + 16;   // two staging buffers either way
  // the kernel aligns its window to 1024 B itself and traps if the slack does not suffice
  constexpr size_t smem = needed + 1024 > kMaxDynSmem ? kMaxDynSmem : needed + 1024;
  static_assert(needed <= kMaxDynSmem, "fprop stage ring does not fit");
  auto kern = conv_fprop_kernel<BLOCK_N, STAGES, STATS, CTAS>;
  if (int rc = ensure_dyn_smem(kern, int(smem))) return rc;
  dim3 grid(groups * p.n_tiles * CTAS);
  if (CTAS == 2) {
    BGR_CUDA_OK(launch_pairs(kern, dim3(grid), dim3(64 + 128 * CTAS), smem, st, tmA, tmB, tmO, bias, stats, p));
  } else {
    BGR_CUDA_OK(launch_pdl<true>(kern, dim3(grid), dim3(192), smem, st, tmA, tmB, tmO, bias, stats, p));
  }
  BGR_LAUNCH_OK();
  return BGR_OK;
}

// Cin tile of the weight-gradient GEMM: 256 halves the shared-memory operand traffic per FLOP; weigh it
// against the padding
static int wgrad_block_n(int Cin) {
  if (Cin <= 64) return 64;
  const int pad128 = (Cin + 127) / 128 * 128, pad256 = (Cin + 255) / 256 * 256;
  const bool wide = g_force_block_n ? g_force_block_n == 256 : (Cin > 128 && pad256 * 13 <= pad128 * 20);
  return wide ? 256 : 128;
}

// Split the pixel reduction over `splits` CTAs per tile.  Every split costs a prologue and a
// Cout x Cin pass of fp32 stores + one re-read through L2 (about 32 chunks' worth of MMA time), and the
// grid runs in rounds of one CTA per SM: take the split count with the cheapest schedule.
// CTA pairs for the weight gradient: 256 couts per pair, so only where that pads no worse than 128-cout tiles, with
// the wide Cin tile (the feed-bound one).  BGR_CONV_CTAS / bgr_debug_conv_ctas force 1 or 2 where the shape allows.
static int wgrad_ctas(int Cout, int Cin) {
  const int forced = g_force_ctas_read();
  if (wgrad_block_n(Cin) != 256 || Cout < 256) return 1;
  if (forced == 1 || forced == 2) return forced;
  return (Cout + 255) / 256 * 2 == (Cout + 127) / 128 ? 2 : 1;
}

static int wgrad_splits(int Cout, int Cin, int ntaps, int total_chunks, int* chunks_per_split) {
  const int bn = wgrad_block_n(Cin);
  const int ctas = wgrad_ctas(Cout, Cin);
  const int tiles = ((Cout + kBlockM * ctas - 1) / (kBlockM * ctas)) * ((Cin + bn - 1) / bn) * ntaps;
  const int sms = sm_count() / ctas;   // CTAs or CTA pairs that run at once
  int splits = 1;
  long best = -1;
  for (int sp = 1; sp <= 64 && sp <= total_chunks; ++sp) {
    const long rounds = (long(tiles) * sp + sms - 1) / sms;
    const long cost = rounds * ((total_chunks + sp - 1) / sp + 32);
    if (best < 0 || cost < best) {
      best = cost;
      splits = sp;
    }
  }
  *chunks_per_split = (total_chunks + splits - 1) / splits;
  return (total_chunks + *chunks_per_split - 1) / *chunks_per_split;
}

template <int BLOCK_N, int STAGES, int CTAS = 1>
static int launch_wgrad(const CUtensorMap& tmDY, const CUtensorMap& tmX, float* dst, int splits,
                        const ConvParams& p, cudaStream_t st) {
  constexpr size_t smem = size_t(STAGES) * ((kBlockM + BLOCK_N / CTAS) * kBlockK * 2) +
                          (2 * STAGES + 1) * 8 + 16 + 1024;
  static_assert(smem <= kMaxDynSmem, "wgrad stage ring does not fit");
  auto kern = conv_wgrad_kernel<BLOCK_N, STAGES, CTAS>;
  if (int rc = ensure_dyn_smem(kern, int(smem))) return rc;
  const int tiles = ((p.Cout + kBlockM * CTAS - 1) / (kBlockM * CTAS)) * ((p.Cin + BLOCK_N - 1) / BLOCK_N) * p.ntaps;
  dim3 grid(tiles * CTAS, splits);
  if (CTAS == 2) BGR_CUDA_OK(launch_pairs(kern, dim3(grid), dim3(192), smem, st, tmDY, tmX, dst, p));
  else BGR_CUDA_OK(launch_pdl<true>(kern, dim3(grid), dim3(192), smem, st, tmDY, tmX, dst, p));
  BGR_LAUNCH_OK();
  return BGR_OK;
}

}  // namespace bgr

using namespace bgr;

extern "C" void bgr_debug_conv_block_n(int bn) { g_force_block_n = bn; }
extern "C" void bgr_debug_conv_mode(int m) { g_conv_dbg = m; }
extern "C" void bgr_debug_conv_ctas(int c) { g_force_ctas = c; }
#ifdef BGR_CONV_TIMING
static int g_force_stages = 0;
extern "C" void bgr_debug_conv_stages(int s) { g_force_stages = s; }
// out[512][8] cycles per CTA: producer {wait empty, role}, MMA {wait tempty, wait full, role}, epilogue {wait tfull,
// wait staging buffer, role}; reset != 0 clears the counters afterwards
extern "C" int bgr_debug_conv_timing(long long* out, int reset) {
  BGR_CUDA_OK(cudaDeviceSynchronize());
  BGR_CUDA_OK(cudaMemcpyFromSymbol(out, g_conv_timing, sizeof(long long) * 512 * 8));
  if (reset) {
    void* ptr = nullptr;
    BGR_CUDA_OK(cudaGetSymbolAddress(&ptr, g_conv_timing));
    BGR_CUDA_OK(cudaMemset(ptr, 0, sizeof(long long) * 512 * 8));
  }
  return BGR_OK;
}
#endif

// CTA pairs (cta_group::2) for the forward kernel: 1 = 1-CTA tiles, 2 = pairs.  Pairs pay when the tile is bound
// by its operand feed, i.e. a 256-wide Cout tile with a long reduction (measured per round of tiles, 1-CTA -> pair:
// 12 k-blocks 6.37 -> 6.06 us, 24 k-blocks 11.16 -> 10.23 us; 4 k-blocks 4.36 -> 4.34 us; 128-wide tiles never
// gain and lose at 2 k-blocks, 1.28 -> 2.35 us: there the two CTAs only wait for each other; at 4 k-blocks pairs
// gain only when several Cout tiles re-read the pixel tile, 256 -> 728 at 96 x 144: 92.9 -> 82.1 us).
// BGR_CONV_CTAS=1|2 (or bgr_debug_conv_ctas) forces one wherever the pair kernel exists.
static int pick_ctas(int bn, int m_tiles, int iters, int n_tiles) {
  g_force_ctas_read();
  if ((bn != 128 && bn != 256) || m_tiles < 2) return 1;
  if (g_force_ctas == 1 || g_force_ctas == 2) return g_force_ctas;
  return bn == 256 && (iters >= 8 || (iters >= 4 && n_tiles >= 2)) ? 2 : 1;
}

// Cout tile width: 256 halves the activation re-reads and the shared-memory operand traffic per
// FLOP, 128 quantises better onto the SMs, 192 sits between: 728 (= 768 padded) output channels at
// the 48x72 mid-flow shape are 3 x 256 -> 49 CTA groups -> 216 pixel tiles in 5 rounds (4.41 needed),
// but 4 x 192 -> 37 groups (all 148 SMs) -> 6 rounds of 0.75-size tiles = 4.5.  Compare the
// schedules' costs: rounds x tile width / relative efficiency of that width.  Measured: the 192-wide
// tile runs at 0.84-0.87 of the 256-wide one per FLOP (1203 vs 1386 TFLOP/s at 1536->1536), which
// cancels its better quantisation at the mid-flow shape to the microsecond (38.9 us both), so with
// the measured weight it is never picked today; it stays selectable (bgr_debug_conv_block_n, tests).
static int pick_block_n(int m_tiles, int Cout) {
  if (Cout <= 64) return 64;
  if (Cout <= 128) return 128;
  if (g_force_block_n == 128 || g_force_block_n == 192 || g_force_block_n == 256) return g_force_block_n;
  const int sms = sm_count();
  auto rounds = [&](int bn) {
    const int n_tiles = (Cout + bn - 1) / bn;
    const int groups = sms / n_tiles > 0 ? sms / n_tiles : 1;
    return long((m_tiles + groups - 1) / groups);
  };
  const long c128 = rounds(128) * 128 * 100 / 90, c192 = rounds(192) * 192 * 100 / 86, c256 = rounds(256) * 256;
  if (c192 < c256 && c192 < c128) return 192;
  return c256 <= c128 ? 256 : 128;
}

extern "C" int bgr_conv_fprop(const void* x, int N, int H, int W, int Cin, int ldx, const void* w,
                              int ldw, int Cout, int ntaps, const int* dy, const int* dx, void* y,
                              int ldy, const float* bias, float* bn_part, int* bn_rows, void* stream) {
  if (int rc = check_conv_args(x, w, y, Cin, ldx, Cout, ldy, ntaps)) return rc;
  if (ldw < Cin || (ldw % 8)) return fail(BGR_ERR_ARG, "conv: ldw must be >= Cin and a multiple of 8");
  ConvParams p{};
  p.N = N; p.H = H; p.W = W;
  pick_rect(H, W, kBlockM, &p.BW, &p.BH);
  p.tiles_x = (W + p.BW - 1) / p.BW;
  p.tiles_y = (H + p.BH - 1) / p.BH;
  p.m_tiles = N * p.tiles_x * p.tiles_y;
  p.ntaps = ntaps;
  p.kblocks = (Cin + kBlockK - 1) / kBlockK;
  p.Cout = Cout; p.Cin = Cin; p.ld_out = ldy;
  for (int t = 0; t < ntaps; ++t) { p.dy[t] = dy[t]; p.dx[t] = dx[t]; }
  p.dbg = g_conv_dbg;
  const int bn = pick_block_n(p.m_tiles, Cout);
  p.n_tiles = (Cout + bn - 1) / bn;
  const int ctas = pick_ctas(bn, p.m_tiles, p.ntaps * p.kblocks, p.n_tiles);
  CUtensorMap tmA, tmB, tmO;
  if (int rc = make_act_map(&tmA, x, N, H, W, Cin, ldx, p.BW, p.BH)) return rc;
  if (int rc = make_act_map(&tmO, y, N, H, W, Cout, ldy, p.BW, p.BH)) return rc;
  {
    uint64_t dims[3] = {uint64_t(Cin), uint64_t(ntaps), uint64_t(Cout)};
    uint64_t strides[2] = {uint64_t(ldw) * 2, uint64_t(ldw) * ntaps * 2};
    uint32_t box[3] = {64, 1, uint32_t(bn / ctas)};   // a CTA of a pair stages half of the Cout tile
    if (int rc = encode_tmap_bf16(&tmB, w, 3, dims, strides, box)) return rc;
  }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  int groups = (sm_count() / ctas) / p.n_tiles;   // CTAs (or CTA pairs) per Cout tile
  if (groups < 1) groups = 1;                     // more Cout tiles than SMs: the surplus CTAs queue
  if (groups > (p.m_tiles + ctas - 1) / ctas) groups = (p.m_tiles + ctas - 1) / ctas;
  if (bn_part != nullptr) {
    if (bias != nullptr) return fail(BGR_ERR_ARG, "conv: fused BatchNorm statistics exclude a bias");
    if (bn_rows == nullptr) return fail(BGR_ERR_ARG, "conv: bn_rows must be given with bn_part");
    if (groups * ctas > BGR_BN_MAX_ROWS) groups = BGR_BN_MAX_ROWS / ctas;
    *bn_rows = groups * ctas;   // one partial row per CTA of a group
#ifdef BGR_CONV_TIMING
    if (ctas == 2 && bn == 256 && g_force_stages == 3) return launch_fprop<256, 3, true, 2>(tmA, tmB, tmO, bias, bn_part, p, groups, st);
    if (ctas == 2 && bn == 256 && g_force_stages == 4) return launch_fprop<256, 4, true, 2>(tmA, tmB, tmO, bias, bn_part, p, groups, st);
#endif
    if (ctas == 2 && bn == 128) return launch_fprop<128, 8, true, 2>(tmA, tmB, tmO, bias, bn_part, p, groups, st);
    if (ctas == 2) return launch_fprop<256, 6, true, 2>(tmA, tmB, tmO, bias, bn_part, p, groups, st);
    if (bn == 64) return launch_fprop<64, 6, true>(tmA, tmB, tmO, bias, bn_part, p, groups, st);
    if (bn == 128) return launch_fprop<128, 6, true>(tmA, tmB, tmO, bias, bn_part, p, groups, st);
    if (bn == 192) return launch_fprop<192, 4, true>(tmA, tmB, tmO, bias, bn_part, p, groups, st);
    return launch_fprop<256, 4, true>(tmA, tmB, tmO, bias, bn_part, p, groups, st);
  }
  if (ctas == 2 && bn == 128) return launch_fprop<128, 8, false, 2>(tmA, tmB, tmO, bias, nullptr, p, groups, st);
  if (ctas == 2) return launch_fprop<256, 6, false, 2>(tmA, tmB, tmO, bias, nullptr, p, groups, st);
  if (bn == 64) return launch_fprop<64, 6, false>(tmA, tmB, tmO, bias, nullptr, p, groups, st);
  if (bn == 128) return launch_fprop<128, 6, false>(tmA, tmB, tmO, bias, nullptr, p, groups, st);
  if (bn == 192) return launch_fprop<192, 4, false>(tmA, tmB, tmO, bias, nullptr, p, groups, st);
  return launch_fprop<256, 4, false>(tmA, tmB, tmO, bias, nullptr, p, groups, st);
}

static void wgrad_geometry(ConvParams* p, int N, int H, int W, int Cin, int Cout, int ntaps) {
  p->N = N; p->H = H; p->W = W;
  pick_rect(H, W, kBlockK, &p->BW, &p->BH);
  p->tiles_x = (W + p->BW - 1) / p->BW;
  p->tiles_y = (H + p->BH - 1) / p->BH;
  p->total_chunks = N * p->tiles_x * p->tiles_y;
  p->ntaps = ntaps;
  p->Cout = Cout; p->Cin = Cin;
  p->slab_elems = long(Cout) * ntaps * Cin;
}

// Launch plan of the implicit-GEMM kernels for one layer signature, from the same pick functions the launches use
// (no device work; with no device the SM count falls back to 148).  kind 0 = bgr_conv_fprop, 1 = bgr_conv_wgrad.
extern "C" int bgr_conv_plan(int kind, int N, int H, int W, int Cin, int Cout, int ntaps, int with_stats, int* out8) {
  if (out8 == nullptr || ntaps < 1 || ntaps > kMaxTaps || N < 1 || H < 1 || W < 1 || Cin < 8 || Cout < 8)
    return fail(BGR_ERR_ARG, "conv_plan: bad arguments");
  for (int i = 0; i < 8; ++i) out8[i] = 0;
  if (kind == 0) {
    int BW, BH;
    pick_rect(H, W, kBlockM, &BW, &BH);
    const int m_tiles = N * ((W + BW - 1) / BW) * ((H + BH - 1) / BH);
    const int kblocks = (Cin + kBlockK - 1) / kBlockK;
    const int bn = pick_block_n(m_tiles, Cout);
    const int n_tiles = (Cout + bn - 1) / bn;
    const int ctas = pick_ctas(bn, m_tiles, ntaps * kblocks, n_tiles);
    int groups = (sm_count() / ctas) / n_tiles;   // as in bgr_conv_fprop
    if (groups < 1) groups = 1;
    if (groups > (m_tiles + ctas - 1) / ctas) groups = (m_tiles + ctas - 1) / ctas;
    if (with_stats && groups * ctas > BGR_BN_MAX_ROWS) groups = BGR_BN_MAX_ROWS / ctas;
    out8[0] = bn; out8[1] = ctas; out8[2] = n_tiles; out8[3] = groups; out8[4] = m_tiles;
    out8[5] = ((m_tiles + ctas - 1) / ctas + groups - 1) / groups;   // rounds of tiles
    out8[6] = with_stats ? groups * ctas : 0;                        // BatchNorm partial rows
    out8[7] = groups * n_tiles * ctas;                               // grid
    return BGR_OK;
  }
  if (kind == 1) {
    ConvParams p{};
    wgrad_geometry(&p, N, H, W, Cin, Cout, ntaps);
    int cps = 0;
    const int splits = wgrad_splits(Cout, Cin, ntaps, p.total_chunks, &cps);
    const int bn = wgrad_block_n(Cin), ctas = wgrad_ctas(Cout, Cin);
    const int tiles = ((Cout + kBlockM * ctas - 1) / (kBlockM * ctas)) * ((Cin + bn - 1) / bn) * ntaps;
    out8[0] = bn; out8[1] = ctas; out8[2] = tiles; out8[3] = splits; out8[4] = cps; out8[5] = p.total_chunks;
    out8[6] = tiles * ctas * splits;   // grid (CTAs)
    out8[7] = splits > 1 ? 1 : 0;      // needs the slab workspace
    return BGR_OK;
  }
  return fail(BGR_ERR_ARG, "conv_plan: kind must be 0 (fprop) or 1 (wgrad)");
}

extern "C" size_t bgr_conv_wgrad_workspace(int N, int H, int W, int Cin, int Cout, int ntaps) {
  if (ntaps < 1 || ntaps > kMaxTaps) return 0;
  ConvParams p{};
  wgrad_geometry(&p, N, H, W, Cin, Cout, ntaps);
  int cps;
  const int splits = wgrad_splits(Cout, Cin, ntaps, p.total_chunks, &cps);
  return splits > 1 ? size_t(splits) * size_t(p.slab_elems) * sizeof(float) : 0;
}

extern "C" int bgr_conv_wgrad(const void* x, int N, int H, int W, int Cin, int ldx,
                              const void* dy_, int Cout, int ldy, int ntaps, const int* dy,
                              const int* dx, float* dw, float* ws, size_t ws_bytes, void* stream) {
  if (int rc = check_conv_args(x, dw, dy_, Cin, ldx, Cout, ldy, ntaps)) return rc;
  ConvParams p{};
  wgrad_geometry(&p, N, H, W, Cin, Cout, ntaps);
  for (int t = 0; t < ntaps; ++t) { p.dy[t] = dy[t]; p.dx[t] = dx[t]; }
  int splits = wgrad_splits(Cout, Cin, ntaps, p.total_chunks, &p.chunks_per_split);
  if (splits > 1 && (ws == nullptr || ws_bytes < size_t(splits) * size_t(p.slab_elems) * sizeof(float)))
    return fail(BGR_ERR_ARG, "conv_wgrad: workspace smaller than bgr_conv_wgrad_workspace() asks for");
  if ((uintptr_t(ws) & 15) != 0) return fail(BGR_ERR_ARG, "conv_wgrad: workspace must be 16-byte aligned");
  CUtensorMap tmDY, tmX;
  if (int rc = make_act_map(&tmDY, dy_, N, H, W, Cout, ldy, p.BW, p.BH)) return rc;
  if (int rc = make_act_map(&tmX, x, N, H, W, Cin, ldx, p.BW, p.BH)) return rc;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  float* dst = splits > 1 ? ws : dw;
  const int bn = wgrad_block_n(Cin);
  int rc;
  if (bn == 64) rc = launch_wgrad<64, 6>(tmDY, tmX, dst, splits, p, st);
  else if (bn == 256 && wgrad_ctas(Cout, Cin) == 2) rc = launch_wgrad<256, 6, 2>(tmDY, tmX, dst, splits, p, st);
  else if (bn == 256) rc = launch_wgrad<256, 4>(tmDY, tmX, dst, splits, p, st);
  else rc = launch_wgrad<128, 6>(tmDY, tmX, dst, splits, p, st);
  if (rc != BGR_OK || splits == 1) return rc;
  return launch_sum_rows(ws, splits, p.slab_elems, dw, st);
}
