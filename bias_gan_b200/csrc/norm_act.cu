// Training-mode BatchNorm + LeakyReLU(0.2) + residual kernels on bf16 NHWC tensors viewed as
// [M pixels][C channels] with a pixel stride (ld) so they can read/write channel slices.
// Replaces nn.BatchNorm2d / nn.LeakyReLU / `x += skip` call sites of the reference
// (src/deepCam/architecture/gpsro/deeplab.py:100,141,179-183,334-338,365-368,623-631).
//
// All kernels are HBM-bound: 128-bit accesses (8 bf16 per thread).  Per-channel reductions are
// schedule-independent: per-thread fp32 partials -> one shared-memory slot per (row lane, channel),
// added in lane order -> one partial row per block in a workspace [rows][2C] -> the finalize /
// coefficient kernels add the rows in index order in fp64.  No floating-point atomics anywhere.
#include <cstdlib>
#include <cooperative_groups.h>
#include <cuda_bf16.h>

#include <mutex>
#include <string>

#include "bgr_internal.h"

namespace bgr {

__device__ __forceinline__ void unpack8(const uint4& u, float (&f)[8]) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 t = __bfloat1622float2(h[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  uint4 u;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
  return u;
}

__device__ __forceinline__ void load8f(const float* p, float (&f)[8]) {
  *reinterpret_cast<float4*>(f) = *reinterpret_cast<const float4*>(p);
  *reinterpret_cast<float4*>(f + 4) = *reinterpret_cast<const float4*>(p + 4);
}

// ---- packed (float2) forms.  The streaming kernels are issue-bound at the mid-flow shape (40 MB tensors that
// live in L2), so their inner loops use FFMA2 / FADD2 / FMUL2 on channel pairs, round two values per cvt, and step
// pointers instead of recomputing 64-bit row offsets.  Every operation is the same IEEE operation as the scalar
// form it replaced, in the same order per channel: results are bit-identical (checked against the scalar library
// on B200, tools/dev_bn_bits.py).
__device__ __forceinline__ float2 bf2_to_f2(uint32_t w) {
  return make_float2(__uint_as_float(w << 16), __uint_as_float(w & 0xffff0000u));
}
__device__ __forceinline__ void unpack8p(const uint4& u, float2 (&f)[4]) {
  f[0] = bf2_to_f2(u.x); f[1] = bf2_to_f2(u.y); f[2] = bf2_to_f2(u.z); f[3] = bf2_to_f2(u.w);
}
__device__ __forceinline__ uint32_t pack2rn(const float2& f) {   // two RNE roundings in one cvt.rn.bf16x2.f32
  __nv_bfloat162 t = __floats2bfloat162_rn(f.x, f.y);
  return *reinterpret_cast<uint32_t*>(&t);
}
__device__ __forceinline__ void load8p(const float* p, float2 (&f)[4]) {
  const float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
  f[0] = make_float2(a.x, a.y); f[1] = make_float2(a.z, a.w);
  f[2] = make_float2(b.x, b.y); f[3] = make_float2(b.z, b.w);
}
// LeakyReLU derivative / value select on a pair: z > 0 ? v : slope * v
__device__ __forceinline__ float2 leaky_sel(const float2& z, const float2& v, const float2& sl2) {
  const float2 sv = __fmul2_rn(sl2, v);
  return make_float2(z.x > 0.f ? v.x : sv.x, z.y > 0.f ? v.y : sv.y);
}

// ---- per-thread asynchronous row ring.  A thread of these kernels walks 6-60 rows of 16 bytes per tensor; with
// the loads held in registers (two rows in flight) it is a serial chain of memory round trips -- latency-bound at
// the mid-flow shape, where a whole 40 MB pass should take 6 us and took 25-35 (fewer instructions changed
// nothing: profiles/r02_bn_ring.txt).  Rows are therefore fetched kDepth ahead with cp.async into shared-memory
// slots that only the issuing thread reads back (no barriers, no registers held by loads in flight), the scheme
// of the depthwise marches.  Slot layout [tensor][slot][thread] x 16 bytes: consecutive threads, consecutive lines.
__device__ __forceinline__ void bn_cp16(uint4* smem, const void* gmem, bool valid) {
  const unsigned sa = static_cast<unsigned>(__cvta_generic_to_shared(smem));
  const int sz = valid ? 16 : 0;   // 0 source bytes: nothing is read
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(sa), "l"(gmem), "r"(sz) : "memory");
}
__device__ __forceinline__ void bn_cp_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bn_cp_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
__host__ __device__ constexpr int bn_ring_depth(int ntensors) { return ntensors <= 2 ? 4 : 2; }   // power of two
static size_t bn_ring_bytes(int ntensors, int threads) {
  return size_t(ntensors) * bn_ring_depth(ntensors) * threads * sizeof(uint4);
}

// Thread -> (row lane r, channel group g) mapping shared by the reduction kernels:
// G = C/8 groups.  If G <= 256 a block covers R = 256/G rows at a time, else each thread loops
// over groups g = tid, tid+256, ...
struct RowGroupMap {
  int G, R;
  __device__ RowGroupMap(int C) : G(C >> 3) { R = G >= 256 ? 1 : 256 / G; }
};

// ---------------------------------------------------------------- statistics: sum, sum of squares
// Deposit this thread's 2 x 8 values for (row lane r, channel group g) and write the block's partial
// row: part[blockIdx.x][i] = sum over row lanes (in lane order) of sh[r][i].
__device__ __forceinline__ void block_partial_put(float* sh, int C, int r, int g, const float (&s)[8], const float (&q)[8]) {
  float* dst = sh + size_t(r) * 2 * C;
  *reinterpret_cast<float4*>(dst + g * 8) = make_float4(s[0], s[1], s[2], s[3]);
  *reinterpret_cast<float4*>(dst + g * 8 + 4) = make_float4(s[4], s[5], s[6], s[7]);
  *reinterpret_cast<float4*>(dst + C + g * 8) = make_float4(q[0], q[1], q[2], q[3]);
  *reinterpret_cast<float4*>(dst + C + g * 8 + 4) = make_float4(q[4], q[5], q[6], q[7]);
}
__device__ __forceinline__ void block_partial_flush(const float* sh, int C, int R, float* __restrict__ part) {
  __syncthreads();
  float* row = part + size_t(blockIdx.x) * 2 * C;
  for (int i = threadIdx.x; i < 2 * C; i += blockDim.x) {
    float t = sh[i];
    for (int r = 1; r < R; ++r) t += sh[size_t(r) * 2 * C + i];
    row[i] = t;
  }
}

__global__ void __launch_bounds__(256)
bn_stats_kernel(const __nv_bfloat16* __restrict__ x, long M, int C, int ld, int rows_per_block,
                float* __restrict__ part /* [gridDim.x][2][C] */) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  extern __shared__ float sh[];  // [R][2][C]
  const RowGroupMap m(C);
  const long r0 = long(blockIdx.x) * rows_per_block;
  const long r1 = min(r0 + rows_per_block, M);
  for (int g = threadIdx.x % m.G; g < m.G; g += (m.G >= 256 ? 256 : m.G)) {
    const int r = m.G >= 256 ? 0 : threadIdx.x / m.G;
    if (r >= m.R) break;
    float s[8] = {0, 0, 0, 0, 0, 0, 0, 0}, q[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (long row = r0 + r; row < r1; row += m.R) {
      float f[8];
      unpack8(*reinterpret_cast<const uint4*>(x + row * ld + g * 8), f);
#pragma unroll
      for (int j = 0; j < 8; ++j) { s[j] += f[j]; q[j] += f[j] * f[j]; }
    }
    block_partial_put(sh, C, r, g, s, q);
    if (m.G < 256) break;
  }
  block_partial_flush(sh, C, m.R, part);
}

// Ordered fp64 sum of the partial rows of kRsCh channels: kRsLanes row lanes x kRsCh channels per block; lane w
// adds rows w, w + kRsLanes, ... in row order (eight loads in flight: with up to 444 rows this kernel sits on the
// critical path between the two BatchNorm-backward passes and is pure load latency); the lane sums are combined by
// a fixed binary tree.  The order depends on nothing but (rows, C): the result is reproducible.  Returns true in
// the threads (lane 0) that hold the result for a valid channel c.
constexpr int kRsCh = 8, kRsLanes = 32;   // 256 threads; grid = ceil(C / 8): 91 blocks at C = 728
__device__ __forceinline__ bool ordered_row_sums(const float* __restrict__ part, int rows, int C, int* c_out,
                                                 double* s_out, double* q_out) {
  __shared__ double shd[kRsLanes][2][kRsCh];
  const int cl = threadIdx.x % kRsCh, rl = threadIdx.x / kRsCh;
  const int c = blockIdx.x * kRsCh + cl;
  double s = 0.0, q = 0.0;
  if (c < C) {
    const float* p = part + c;
    const size_t st = size_t(2) * C;
    int r = rl;
    for (; r + 7 * kRsLanes < rows; r += 8 * kRsLanes) {
      float a[8], b[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        a[u] = p[size_t(r + u * kRsLanes) * st];
        b[u] = p[size_t(r + u * kRsLanes) * st + C];
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) { s += double(a[u]); q += double(b[u]); }
    }
    if (r < rows) {   // up to seven more rows, again all loads first
      float a[7], b[7];
#pragma unroll
      for (int u = 0; u < 7; ++u) {
        const bool ok = r + u * kRsLanes < rows;
        a[u] = ok ? p[size_t(r + u * kRsLanes) * st] : 0.f;
        b[u] = ok ? p[size_t(r + u * kRsLanes) * st + C] : 0.f;
      }
#pragma unroll
      for (int u = 0; u < 7; ++u)
        if (r + u * kRsLanes < rows) { s += double(a[u]); q += double(b[u]); }
    }
  }
  shd[rl][0][cl] = s;
  shd[rl][1][cl] = q;
  __syncthreads();
#pragma unroll
  for (int w = kRsLanes / 2; w > 0; w >>= 1) {   // fixed tree over the row lanes
    if (rl < w) {
      shd[rl][0][cl] += shd[rl + w][0][cl];
      shd[rl][1][cl] += shd[rl + w][1][cl];
    }
    __syncthreads();
  }
  if (rl != 0 || c >= C) return false;
  *c_out = c; *s_out = shd[0][0][cl]; *q_out = shd[0][1][cl];
  return true;
}

// mean / invstd / fused affine (a, b) / running statistics (momentum, unbiased variance) from the
// partial rows of the statistics pass (bn_stats_kernel or the conv epilogue).  grid = ceil(C/8) x 256.
__global__ void __launch_bounds__(256)
bn_finalize_kernel(const float* __restrict__ part, int rows, long M, int C,
                   const float* __restrict__ gamma, const float* __restrict__ beta,
                   float eps, float momentum, float* __restrict__ mean,
                   float* __restrict__ invstd, float* __restrict__ a,
                   float* __restrict__ b, float* __restrict__ running_mean,
                   float* __restrict__ running_var, long long* __restrict__ num_batches) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  if (blockIdx.x == 0 && threadIdx.x == 0 && num_batches) *num_batches += 1;   // nn.BatchNorm2d.num_batches_tracked
  int c;
  double sum, sq;
  if (!ordered_row_sums(part, rows, C, &c, &sum, &sq)) return;
  const double mu = sum / double(M);
  double var = sq / double(M) - mu * mu;
  if (var < 0) var = 0;
  const float is = float(1.0 / sqrt(var + double(eps)));
  mean[c] = float(mu);
  invstd[c] = is;
  const float ga = gamma ? gamma[c] : 1.f;
  a[c] = ga * is;
  b[c] = (beta ? beta[c] : 0.f) - float(mu) * ga * is;
  if (running_mean) {
    const double unbiased = M > 1 ? var * double(M) / double(M - 1) : var;
    running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * float(mu);
    running_var[c] = (1.f - momentum) * running_var[c] + momentum * float(unbiased);
  }
}

// ---------------------------------------------------------------- apply: out = act(a*x + b [+ res])
// Thread = one fixed group of 8 channels (coefficients stay in registers) x rows r, r+R, ...;
// blockDim = G*R with R = 256/G rows per block step when G = C/8 <= 256, else 256 threads looping
// over groups.  (The first version re-read a and b for every 16 bytes of data: L1-bound at 91 %.)
template <bool ACT, bool RES>
__global__ void __launch_bounds__(256)
bn_apply_kernel(const __nv_bfloat16* __restrict__ x, int ldx, const float* __restrict__ a,
                const float* __restrict__ b, const __nv_bfloat16* __restrict__ res, int ldr,
                __nv_bfloat16* __restrict__ out, int ldo, long M, int C, float slope) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  const int G = C >> 3;
  const int R = G >= 256 ? 1 : 256 / G;
  const int g0 = G >= 256 ? threadIdx.x : threadIdx.x % G;
  const int r = G >= 256 ? 0 : threadIdx.x / G;
  const long rstep = long(gridDim.x) * R;
  const float2 sl2 = make_float2(slope, slope);
  for (int g = g0; g < G; g += 256) {
    float2 av[4], bv[4];
    load8p(a + g * 8, av);
    load8p(b + g * 8, bv);
    auto one = [&](__nv_bfloat16* dst, const uint4& xu, const uint4& ru) {
      float2 f[4], rr[4];
      unpack8p(xu, f);
      if (RES) unpack8p(ru, rr);
      uint4 o;
      uint32_t* ow = reinterpret_cast<uint32_t*>(&o);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float2 v = __ffma2_rn(av[j], f[j], bv[j]);
        if (RES) v = __fadd2_rn(v, rr[j]);
        if (ACT) v = leaky_sel(v, v, sl2);
        ow[j] = pack2rn(v);
      }
      *reinterpret_cast<uint4*>(dst) = o;
    };
    constexpr int NT = RES ? 2 : 1, DEPTH = bn_ring_depth(NT);
    extern __shared__ uint4 ring_raw[];
    uint4* ring = ring_raw + threadIdx.x;
    const int nthr = blockDim.x;
    const long row0 = long(blockIdx.x) * R + r;
    const int n = row0 < M ? int((M - row0 + rstep - 1) / rstep) : 0;   // rows of this thread
    const __nv_bfloat16* xp = x + row0 * ldx + g * 8;
    const __nv_bfloat16* rp = RES ? res + row0 * ldr + g * 8 : x;
    __nv_bfloat16* op = out + row0 * ldo + g * 8;
    const long xs = rstep * ldx, rs = rstep * ldr, os = rstep * ldo;
    const uint4 z4 = make_uint4(0, 0, 0, 0);
    auto issue = [&](int k) {
      const bool ok = k < n;
      const int sl = k & (DEPTH - 1);
      bn_cp16(ring + sl * nthr, ok ? xp : x, ok);
      if (RES) bn_cp16(ring + (DEPTH + sl) * nthr, ok ? rp : x, ok);
      bn_cp_commit();
      xp += xs;
      if (RES) rp += rs;
    };
#pragma unroll
    for (int k = 0; k < DEPTH; ++k) issue(k);
    for (int k = 0; k < n; ++k) {
      bn_cp_wait<DEPTH - 1>();
      const int sl = k & (DEPTH - 1);
      const uint4 xu = ring[sl * nthr];
      const uint4 ru = RES ? ring[(DEPTH + sl) * nthr] : z4;
      issue(k + DEPTH);
      one(op, xu, ru);
      op += os;
    }
    bn_cp_wait<0>();
  }
}

// ---------------------------------------------------------------- backward, pass 1:
// dyh = g * act'(out) (bf16, = gradient of the residual input too) and per-channel
// sums s1 = sum dyh, s2 = sum dyh * xhat  with xhat = (x - mean) * invstd.
// ACT: 0 = no activation, 1 = sign of the stored activation output, 2 = sign recomputed from
// a*x + b (the activated tensor was never stored, see bgr_dw_fwd_bn); then `out` carries a and
// `out + C` ... is unused: a, b come in through `xa`, `xb`.
template <int ACT, bool G2>
__device__ __forceinline__ void
bn_bwd_reduce_body(const __nv_bfloat16* __restrict__ g, int ldg, const __nv_bfloat16* __restrict__ g2,
                   int ldg2, const __nv_bfloat16* __restrict__ out,
                   int ldo, const float* __restrict__ xa, const float* __restrict__ xb,
                   const __nv_bfloat16* __restrict__ x, int ldx,
                   const float* __restrict__ mean, const float* __restrict__ invstd,
                   __nv_bfloat16* __restrict__ dyh, int ldd, long M, int C, int rows_per_block, float slope,
                   float* __restrict__ part /* [gridDim.x][2][C] */) {
  extern __shared__ float sh[];   // [R][2][C]
  const RowGroupMap m(C);
  const long r0 = long(blockIdx.x) * rows_per_block;
  const long r1 = min(r0 + rows_per_block, M);
  const float2 sl2 = make_float2(slope, slope);
  for (int gi = threadIdx.x % m.G; gi < m.G; gi += (m.G >= 256 ? 256 : m.G)) {
    const int r = m.G >= 256 ? 0 : threadIdx.x / m.G;
    if (r >= m.R) break;
    float2 nmu[4], s1[4], s2[4], ca[4], cb[4];   // s2 accumulates dyh*(x - mean); invstd is applied once at the end
    load8p(mean + gi * 8, nmu);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      nmu[j] = make_float2(-nmu[j].x, -nmu[j].y);   // x - mean == x + (-mean), bit for bit
      s1[j] = s2[j] = make_float2(0.f, 0.f);
    }
    if (ACT == 2) {
      load8p(xa + gi * 8, ca);
      load8p(xb + gi * 8, cb);
    }
    auto one = [&](__nv_bfloat16* dst, const uint4& gu, const uint4& hu, const uint4& xu, const uint4& ou) {
      float2 gv[4], xv[4], ov[4];
      unpack8p(gu, gv);
      if (G2) {   // second gradient of the same tensor (residual path), summed in fp32
        float2 hv[4];
        unpack8p(hu, hv);
#pragma unroll
        for (int j = 0; j < 4; ++j) gv[j] = __fadd2_rn(gv[j], hv[j]);
      }
      unpack8p(xu, xv);
      if (ACT == 1) unpack8p(ou, ov);
      uint4 o;
      uint32_t* ow = reinterpret_cast<uint32_t*>(&o);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float2 d = gv[j];
        if (ACT == 1) d = leaky_sel(ov[j], d, sl2);
        if (ACT == 2) d = leaky_sel(__ffma2_rn(ca[j], xv[j], cb[j]), d, sl2);
        // statistics are taken from the bf16-rounded value pass 2 will read back, so that the
        // mean subtraction cancels exactly (matters when M is tiny, e.g. the image-pool branch)
        ow[j] = pack2rn(d);
        d = bf2_to_f2(ow[j]);
        s1[j] = __fadd2_rn(s1[j], d);
        s2[j] = __ffma2_rn(d, __fadd2_rn(xv[j], nmu[j]), s2[j]);
      }
      if (dst != nullptr) *reinterpret_cast<uint4*>(dst) = o;
    };
    const uint4 z4 = make_uint4(0, 0, 0, 0);
    constexpr int NT = 2 + (G2 ? 1 : 0) + (ACT == 1 ? 1 : 0), DEPTH = bn_ring_depth(NT);
    uint4* ring = reinterpret_cast<uint4*>(sh + size_t(m.R) * 2 * C) + threadIdx.x;   // behind the partial-sum slots
    const int nthr = blockDim.x;
    const long row0 = r0 + r;
    const int n = row0 < r1 ? int((r1 - row0 + m.R - 1) / m.R) : 0;   // rows of this thread: row0, row0 + R, ...
    const __nv_bfloat16* gp = g + row0 * ldg + gi * 8;
    const __nv_bfloat16* xp = x + row0 * ldx + gi * 8;
    const __nv_bfloat16* hp = G2 ? g2 + row0 * ldg2 + gi * 8 : g;
    const __nv_bfloat16* op = ACT == 1 ? out + row0 * ldo + gi * 8 : g;
    __nv_bfloat16* dp = dyh != nullptr ? dyh + row0 * ldd + gi * 8 : nullptr;
    const long gs = long(m.R) * ldg, xs = long(m.R) * ldx, hs = long(m.R) * ldg2, os = long(m.R) * ldo, ds = long(m.R) * ldd;
    auto issue = [&](int k) {   // row k of this thread -> slot k % DEPTH of every tensor; one group per row
      const bool ok = k < n;
      const int sl = k & (DEPTH - 1);
      bn_cp16(ring + (0 * DEPTH + sl) * nthr, ok ? gp : g, ok);
      bn_cp16(ring + (1 * DEPTH + sl) * nthr, ok ? xp : g, ok);
      if (G2) bn_cp16(ring + (2 * DEPTH + sl) * nthr, ok ? hp : g, ok);
      if (ACT == 1) bn_cp16(ring + ((G2 ? 3 : 2) * DEPTH + sl) * nthr, ok ? op : g, ok);
      bn_cp_commit();
      gp += gs; xp += xs;
      if (G2) hp += hs;
      if (ACT == 1) op += os;
    };
#pragma unroll
    for (int k = 0; k < DEPTH; ++k) issue(k);
    for (int k = 0; k < n; ++k) {
      bn_cp_wait<DEPTH - 1>();   // the oldest outstanding group is row k
      const int sl = k & (DEPTH - 1);
      const uint4 gu = ring[(0 * DEPTH + sl) * nthr];
      const uint4 xu = ring[(1 * DEPTH + sl) * nthr];
      const uint4 hu = G2 ? ring[(2 * DEPTH + sl) * nthr] : z4;
      const uint4 ou = ACT == 1 ? ring[((G2 ? 3 : 2) * DEPTH + sl) * nthr] : z4;
      issue(k + DEPTH);          // refill the slot just read
      one(dp, gu, hu, xu, ou);
      if (dp != nullptr) dp += ds;
    }
    bn_cp_wait<0>();             // drain the (empty) tail groups before the ring is reused
    float s1f[8], s2f[8];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      s1f[2 * j] = s1[j].x; s1f[2 * j + 1] = s1[j].y;
      s2f[2 * j] = s2[j].x * invstd[gi * 8 + 2 * j]; s2f[2 * j + 1] = s2[j].y * invstd[gi * 8 + 2 * j + 1];
    }
    block_partial_put(sh, C, r, gi, s1f, s2f);
    if (m.G < 256) break;
  }
  block_partial_flush(sh, C, m.R, part);
}

template <int ACT, bool G2>
__global__ void __launch_bounds__(256, 3)   // 85 registers: three blocks per SM (was 91-96 -> two, latency-bound)
bn_bwd_reduce_kernel(const __nv_bfloat16* __restrict__ g, int ldg, const __nv_bfloat16* __restrict__ g2,
                     int ldg2, const __nv_bfloat16* __restrict__ out,
                     int ldo, const float* __restrict__ xa, const float* __restrict__ xb,
                     const __nv_bfloat16* __restrict__ x, int ldx,
                     const float* __restrict__ mean, const float* __restrict__ invstd,
                     __nv_bfloat16* __restrict__ dyh, int ldd, long M, int C, int rows_per_block, float slope,
                     float* __restrict__ part /* [gridDim.x][2][C] */) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  bn_bwd_reduce_body<ACT, G2>(g, ldg, g2, ldg2, out, ldo, xa, xb, x, ldx, mean, invstd, dyh, ldd, M, C, rows_per_block, slope, part);
}

// backward, pass 2: dx = gamma*invstd * (dyh - s1/M - xhat * s2/M), xhat = (x - mean)*invstd.
// Per-channel fp32 coefficients (c0 = gamma*invstd, k1 = s1/M, k2 = s2/M, mean, invstd) are
// prepared by a tiny kernel so the streaming pass has no fp64 and only vector loads.
__device__ __forceinline__ void
bn_bwd_coef_body(const float* __restrict__ mean, const float* __restrict__ invstd,
                 const float* __restrict__ gamma, const float* __restrict__ part, int rows,
                 long M, int C, float* __restrict__ coef /* [7][C] */) {
  int c;
  double s1, s2;
  if (!ordered_row_sums(part, rows, C, &c, &s1, &s2)) return;
  coef[c] = (gamma ? gamma[c] : 1.f) * invstd[c];
  coef[C + c] = float(s1 / double(M));
  coef[2 * C + c] = float(s2 / double(M));
  coef[3 * C + c] = mean[c];
  coef[4 * C + c] = invstd[c];
  coef[5 * C + c] = float(s1);   // dL/dbeta
  coef[6 * C + c] = float(s2);   // dL/dgamma
}

__global__ void __launch_bounds__(256)
bn_bwd_coef_kernel(const float* __restrict__ mean, const float* __restrict__ invstd,
                   const float* __restrict__ gamma, const float* __restrict__ part, int rows,
                   long M, int C, float* __restrict__ coef /* [7][C] */) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  bn_bwd_coef_body(mean, invstd, gamma, part, rows, M, C, coef);
}

__global__ void __launch_bounds__(256, 4)   // 64 registers (4 bytes of spill): 552 -> 513 us per 453 MB BatchNorm backward
bn_bwd_apply_kernel(const __nv_bfloat16* __restrict__ dyh, const __nv_bfloat16* __restrict__ x, int ldx,
                    const float* __restrict__ coef, __nv_bfloat16* __restrict__ dx, int ldd, long M,
                    int C) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  // same thread -> (channel group, rows) mapping as bn_apply_kernel: the five coefficient vectors
  // are loaded once per thread instead of once per 16 bytes of data
  const int G = C >> 3;
  const int R = G >= 256 ? 1 : 256 / G;
  const int g0 = G >= 256 ? threadIdx.x : threadIdx.x % G;
  const int r = G >= 256 ? 0 : threadIdx.x / G;
  const long rstep = long(gridDim.x) * R;
  for (int g = g0; g < G; g += 256) {
    float c0[8], k1[8], k2[8], mu[8], is[8];
    load8f(coef + g * 8, c0);
    load8f(coef + C + g * 8, k1);
    load8f(coef + 2 * C + g * 8, k2);
    load8f(coef + 3 * C + g * 8, mu);
    load8f(coef + 4 * C + g * 8, is);
#pragma unroll
    for (int j = 0; j < 8; ++j) k2[j] *= is[j];   // (x - mu) * is * k2
    auto one = [&](__nv_bfloat16* dst, const uint4& du, const uint4& xu) {
      float d[8], xv[8];
      unpack8(du, d);
      unpack8(xu, xv);
#pragma unroll
      for (int j = 0; j < 8; ++j) d[j] = c0[j] * (d[j] - k1[j] - (xv[j] - mu[j]) * k2[j]);
      *reinterpret_cast<uint4*>(dst) = pack8(d);
    };
    constexpr int DEPTH = bn_ring_depth(2);
    extern __shared__ uint4 ring_raw[];
    uint4* ring = ring_raw + threadIdx.x;
    const int nthr = blockDim.x;
    const long row0 = long(blockIdx.x) * R + r;
    const int n = row0 < M ? int((M - row0 + rstep - 1) / rstep) : 0;
    const __nv_bfloat16* dp = dyh + row0 * ldd + g * 8;
    const __nv_bfloat16* xp = x + row0 * ldx + g * 8;
    __nv_bfloat16* op = dx + row0 * ldd + g * 8;
    const long ds = rstep * ldd, xs = rstep * ldx;
    auto issue = [&](int k) {
      const bool ok = k < n;
      const int sl = k & (DEPTH - 1);
      bn_cp16(ring + sl * nthr, ok ? dp : dyh, ok);
      bn_cp16(ring + (DEPTH + sl) * nthr, ok ? xp : x, ok);
      bn_cp_commit();
      dp += ds; xp += xs;
    };
#pragma unroll
    for (int k = 0; k < DEPTH; ++k) issue(k);
    for (int k = 0; k < n; ++k) {
      bn_cp_wait<DEPTH - 1>();
      const int sl = k & (DEPTH - 1);
      const uint4 du = ring[sl * nthr];
      const uint4 xu = ring[(DEPTH + sl) * nthr];
      issue(k + DEPTH);
      one(op, du, xu);
      op += ds;
    }
    bn_cp_wait<0>();
  }
}

// Pass 2 without a stored dyh (the BatchNorm whose activation was applied on load, bgr_dw_fwd_bn):
// dyh = bf16(g * act'(a*x + b)) is recomputed from g and x -- x is read anyway -- exactly as pass 1
// computed it for the sums, so pass 1 need not write it and this pass need not read it back.
template <bool ACT, bool G2>
__global__ void __launch_bounds__(256, 3)
bn_bwd_apply_recompute_kernel(const __nv_bfloat16* __restrict__ gin, int ldg, const __nv_bfloat16* __restrict__ gin2,
                              int ldg2, const __nv_bfloat16* __restrict__ x,
                              int ldx, const float* __restrict__ xa, const float* __restrict__ xb,
                              const float* __restrict__ coef, __nv_bfloat16* __restrict__ dx, int ldd, long M,
                              int C, float slope) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  const int G = C >> 3;
  const int R = G >= 256 ? 1 : 256 / G;
  const int g0 = G >= 256 ? threadIdx.x : threadIdx.x % G;
  const int r = G >= 256 ? 0 : threadIdx.x / G;
  const long rstep = long(gridDim.x) * R;
  const float2 sl2 = make_float2(slope, slope);
  for (int g = g0; g < G; g += 256) {
    // dx = c0*(v - k1 - (x - mu)*k2*is) folded to c0*v + A*x + B: three coefficient vectors live in
    // registers instead of five (108 -> under 86 registers: a third block per SM; ncu showed the 2-block
    // version latency-bound at 3.6 TB/s where bn_bwd_apply_kernel streams the same bytes at 5.1)
    float2 c0[4], A[4], B[4], ca[4], cb[4];
    {
      float c0s[8], k1[8], k2[8], mu[8], is[8], As[8], Bs[8];
      load8f(coef + g * 8, c0s);
      load8f(coef + C + g * 8, k1);
      load8f(coef + 2 * C + g * 8, k2);
      load8f(coef + 3 * C + g * 8, mu);
      load8f(coef + 4 * C + g * 8, is);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float k2i = k2[j] * is[j];
        As[j] = -c0s[j] * k2i;
        Bs[j] = c0s[j] * (k2i * mu[j] - k1[j]);
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        c0[j] = make_float2(c0s[2 * j], c0s[2 * j + 1]);
        A[j] = make_float2(As[2 * j], As[2 * j + 1]);
        B[j] = make_float2(Bs[2 * j], Bs[2 * j + 1]);
      }
    }
    if (ACT) {
      load8p(xa + g * 8, ca);
      load8p(xb + g * 8, cb);
    }
    auto one = [&](__nv_bfloat16* dst, const uint4& gu, const uint4& hu, const uint4& xu) {
      float2 d[4], xv[4];
      unpack8p(gu, d);
      if (G2) {   // the second gradient of a dual-handle output, summed in fp32 as pass 1 did
        float2 hv[4];
        unpack8p(hu, hv);
#pragma unroll
        for (int j = 0; j < 4; ++j) d[j] = __fadd2_rn(d[j], hv[j]);
      }
      unpack8p(xu, xv);
      uint4 o;
      uint32_t* ow = reinterpret_cast<uint32_t*>(&o);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float2 v = d[j];
        if (ACT) v = leaky_sel(__ffma2_rn(ca[j], xv[j], cb[j]), v, sl2);
        if (ACT || G2) v = bf2_to_f2(pack2rn(v));   // the value pass 1 summed
        ow[j] = pack2rn(__ffma2_rn(c0[j], v, __ffma2_rn(A[j], xv[j], B[j])));
      }
      *reinterpret_cast<uint4*>(dst) = o;
    };
    const uint4 z4 = make_uint4(0, 0, 0, 0);
    constexpr int NT = G2 ? 3 : 2, DEPTH = bn_ring_depth(NT);
    extern __shared__ uint4 ring_raw[];
    uint4* ring = ring_raw + threadIdx.x;
    const int nthr = blockDim.x;
    const long row0 = long(blockIdx.x) * R + r;
    const int n = row0 < M ? int((M - row0 + rstep - 1) / rstep) : 0;
    const __nv_bfloat16* gp = gin + row0 * ldg + g * 8;
    const __nv_bfloat16* hp = G2 ? gin2 + row0 * ldg2 + g * 8 : gin;
    const __nv_bfloat16* xp = x + row0 * ldx + g * 8;
    __nv_bfloat16* op = dx + row0 * ldd + g * 8;
    const long gs = rstep * ldg, hs = rstep * ldg2, xs = rstep * ldx, os = rstep * ldd;
    auto issue = [&](int k) {
      const bool ok = k < n;
      const int sl = k & (DEPTH - 1);
      bn_cp16(ring + sl * nthr, ok ? gp : gin, ok);
      bn_cp16(ring + (DEPTH + sl) * nthr, ok ? xp : gin, ok);
      if (G2) bn_cp16(ring + (2 * DEPTH + sl) * nthr, ok ? hp : gin, ok);
      bn_cp_commit();
      gp += gs; xp += xs;
      if (G2) hp += hs;
    };
#pragma unroll
    for (int k = 0; k < DEPTH; ++k) issue(k);
    for (int k = 0; k < n; ++k) {
      bn_cp_wait<DEPTH - 1>();
      const int sl = k & (DEPTH - 1);
      const uint4 gu = ring[sl * nthr];
      const uint4 xu = ring[(DEPTH + sl) * nthr];
      const uint4 hu = G2 ? ring[(2 * DEPTH + sl) * nthr] : z4;
      issue(k + DEPTH);
      one(op, gu, hu, xu);
      op += os;
    }
    bn_cp_wait<0>();
  }
}

static int env_int(const char* name, int dflt) {
  const char* v = getenv(name);
  return v && *v ? atoi(v) : dflt;
}

// ---------------------------------------------------------------- backward, all passes in one cooperative launch
// pass 1 (partial rows) -> grid barrier -> coefficients (blocks 0 .. ceil(C/16)-1) -> grid barrier -> pass 2 over the
// SAME rows with the same thread mapping, so g and x come back from L2 (a 48x72x728 mid-flow tensor is 40 MB; the
// three-kernel version paid two more launches and their ramps on 30 us kernels).  dyh == NULL: pass 2 recomputes
// dyh from g (and g2) and x exactly as pass 1 rounded it; dyh != NULL: pass 1 stores it and pass 2 reads it back.
template <int ACT>
__global__ void __launch_bounds__(256, 3)
bn_bwd_fused_kernel(const __nv_bfloat16* __restrict__ g, int ldg, const __nv_bfloat16* __restrict__ g2,
                    int ldg2, const __nv_bfloat16* __restrict__ out,
                    int ldo, const float* __restrict__ xa, const float* __restrict__ xb,
                    const __nv_bfloat16* __restrict__ x, int ldx,
                    const float* __restrict__ mean, const float* __restrict__ invstd, const float* __restrict__ gamma,
                    __nv_bfloat16* __restrict__ dyh, __nv_bfloat16* __restrict__ dx, int ldd, long M, int C,
                    int rows_per_block, float slope, float* __restrict__ part, float* __restrict__ coef) {
  cooperative_groups::grid_group grid = cooperative_groups::this_grid();
  if (g2 != nullptr) bn_bwd_reduce_body<ACT, true>(g, ldg, g2, ldg2, out, ldo, xa, xb, x, ldx, mean, invstd, dyh, ldd, M, C, rows_per_block, slope, part);
  else bn_bwd_reduce_body<ACT, false>(g, ldg, g2, ldg2, out, ldo, xa, xb, x, ldx, mean, invstd, dyh, ldd, M, C, rows_per_block, slope, part);
  __threadfence();
  grid.sync();
  if (int(blockIdx.x) * kRsCh < C) bn_bwd_coef_body(mean, invstd, gamma, part, int(gridDim.x), M, C, coef);
  __threadfence();
  grid.sync();
  const RowGroupMap m(C);
  const long r0 = long(blockIdx.x) * rows_per_block;
  const long r1 = min(r0 + rows_per_block, M);
  for (int gi = threadIdx.x % m.G; gi < m.G; gi += (m.G >= 256 ? 256 : m.G)) {
    const int r = m.G >= 256 ? 0 : threadIdx.x / m.G;
    if (r >= m.R) break;
    // dx = c0*(v - k1 - (x - mu)*k2*is) folded to c0*v + A*x + B
    float c0[8], A[8], B[8], ca[8], cb[8];
    {
      float k1[8], k2[8], mu[8], is[8];
      load8f(coef + gi * 8, c0);
      load8f(coef + C + gi * 8, k1);
      load8f(coef + 2 * C + gi * 8, k2);
      load8f(coef + 3 * C + gi * 8, mu);
      load8f(coef + 4 * C + gi * 8, is);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float k2i = k2[j] * is[j];
        A[j] = -c0[j] * k2i;
        B[j] = c0[j] * (k2i * mu[j] - k1[j]);
      }
    }
    if (ACT == 2) {
      load8f(xa + gi * 8, ca);
      load8f(xb + gi * 8, cb);
    }
    auto one = [&](long row, const uint4& du, const uint4& xu, const uint4& ou) {
      float d[8], xv[8], ov[8];
      unpack8(du, d);
      unpack8(xu, xv);
      if (dyh == nullptr) {   // rebuild the rounded dyh of pass 1
        if (g2 != nullptr) {
          float hv[8];
          unpack8(*reinterpret_cast<const uint4*>(g2 + row * ldg2 + gi * 8), hv);
#pragma unroll
          for (int j = 0; j < 8; ++j) d[j] += hv[j];
        }
        if (ACT == 1) unpack8(ou, ov);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          float v = d[j];
          if (ACT == 1) v = ov[j] > 0.f ? v : slope * v;
          if (ACT == 2) v = fmaf(ca[j], xv[j], cb[j]) > 0.f ? v : slope * v;
          d[j] = __bfloat162float(__float2bfloat16_rn(v));
        }
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) d[j] = fmaf(c0[j], d[j], fmaf(A[j], xv[j], B[j]));
      *reinterpret_cast<uint4*>(dx + row * ldd + gi * 8) = pack8(d);
    };
    const __nv_bfloat16* src = dyh != nullptr ? dyh : g;
    const int lds = dyh != nullptr ? ldd : ldg;
    const bool need_out = (ACT == 1) && dyh == nullptr;
    const uint4 z4 = make_uint4(0, 0, 0, 0);
    long row = r0 + r;
    for (; row + m.R < r1; row += 2 * m.R) {
      const long row2 = row + m.R;
      const uint4 d0 = *reinterpret_cast<const uint4*>(src + row * lds + gi * 8);
      const uint4 d1 = *reinterpret_cast<const uint4*>(src + row2 * lds + gi * 8);
      const uint4 x0 = *reinterpret_cast<const uint4*>(x + row * ldx + gi * 8);
      const uint4 x1 = *reinterpret_cast<const uint4*>(x + row2 * ldx + gi * 8);
      const uint4 o0 = need_out ? *reinterpret_cast<const uint4*>(out + row * ldo + gi * 8) : z4;
      const uint4 o1 = need_out ? *reinterpret_cast<const uint4*>(out + row2 * ldo + gi * 8) : z4;
      one(row, d0, x0, o0);
      one(row2, d1, x1, o1);
    }
    if (row < r1)
      one(row, *reinterpret_cast<const uint4*>(src + row * lds + gi * 8),
          *reinterpret_cast<const uint4*>(x + row * ldx + gi * 8),
          need_out ? *reinterpret_cast<const uint4*>(out + row * ldo + gi * 8) : z4);
    if (m.G < 256) break;
  }
}

// Largest co-resident grid of the fused kernel on the current device (cached per kernel instance and device).
template <int ACT>
static int fused_max_grid(size_t smem) {
  static int cache[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 0;
  if (cache[dev] == 0) {
    int per_sm = 0;
    if (ensure_dyn_smem(bn_bwd_fused_kernel<ACT>, 80 * 1024) != BGR_OK ||
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, bn_bwd_fused_kernel<ACT>, 256, 52 * 1024) != cudaSuccess)
      per_sm = 0;
    cache[dev] = per_sm > 0 ? per_sm * sm_count() : -1;
  }
  (void)smem;
  return cache[dev] > 0 ? cache[dev] : 0;
}

template <int ACT>
static int launch_bn_bwd_fused(const __nv_bfloat16* G, int ldg, const __nv_bfloat16* G2, int ldg2, const __nv_bfloat16* O,
                               int ldo, const float* xa, const float* xb, const __nv_bfloat16* X, int ldx,
                               const float* mean, const float* invstd, const float* gamma, __nv_bfloat16* D,
                               __nv_bfloat16* DX, int ldd, long M, int C, int rpb, int grid, float slope, float* part,
                               float* coef, size_t rsm, cudaStream_t st) {
  void* args[] = {&G, &ldg, &G2, &ldg2, &O, &ldo, &xa, &xb, &X, &ldx, &mean, &invstd, &gamma, &D, &DX, &ldd, &M, &C,
                  &rpb, &slope, &part, &coef};
  BGR_CUDA_OK(cudaLaunchCooperativeKernel(reinterpret_cast<void*>(bn_bwd_fused_kernel<ACT>), dim3(grid), dim3(256), args, rsm, st));
  ++g_launch_count;
  return BGR_OK;
}

// grid / block of the (channel group, rows) streaming kernels
// With the row ring every thread wants a long run of rows (its prefetch depth is only useful on a run), so the
// grid is ONE resident wave: `cap` blocks per SM (what registers and the ring's shared memory allow).
static void rowgroup_launch(long M, int C, int* grid, int* block, int cap_default = 4) {
  static const int min_rows = env_int("BGR_EW_ROWS", 4), cap_env = env_int("BGR_EW_CAP", 0);   // tuning knobs
  const int cap = cap_env > 0 ? cap_env : cap_default;
  const int G = C / 8;
  const int R = G >= 256 ? 1 : 256 / G;
  *block = G >= 256 ? 256 : G * R;
  long blocks = (M + long(R) * min_rows - 1) / (long(R) * min_rows);   // >= min_rows rows per thread
  if (blocks > long(sm_count()) * cap) blocks = long(sm_count()) * cap;
  *grid = int(blocks < 1 ? 1 : blocks);
}

// dynamic shared memory of the reduction kernels: one slot per (row lane, 2C) = at most 16 KB
static size_t reduce_smem(int C) {
  const int G = C / 8;
  const int R = G >= 256 ? 1 : 256 / G;
  return size_t(R) * 2 * C * sizeof(float);
}

static int reduce_grid(long M, int C, int* rows_per_block) {
  const int G = C / 8;
  const int R = G >= 256 ? 1 : 256 / G;
  long blocks = (M + long(R) * 8 - 1) / (long(R) * 8);  // >= 8 row iterations per thread
  // three CTAs per SM (the register limit): one wave.  (Wide tensors used two while every block cost 2*C fp64
  // atomics; with partial rows three measure 3-8 % faster at the mid-flow shape, profiles/r02_knob_sweep.txt)
  static const int bps_wide = env_int("BGR_RED_BPS_WIDE", 3);   // tuning knob
  long cap = C <= 256 ? 148 * 3 : 148 * bps_wide;   // one wave at three resident blocks per SM
  if (cap > BGR_BN_MAX_ROWS) cap = BGR_BN_MAX_ROWS;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  *rows_per_block = int((M + blocks - 1) / blocks);
  return int((M + *rows_per_block - 1) / *rows_per_block);
}

}  // namespace bgr

using namespace bgr;

#define BGR_CHECK_C8(C) \
  if ((C) % 8 || (C) < 8) return fail(BGR_ERR_ARG, "channel count must be a multiple of 8")

extern "C" int bgr_bn_stats(const void* x, long M, int C, int ld, float* part, int* rows, void* stream) {
  BGR_CHECK_C8(C);
  if (reduce_smem(C) > 48 * 1024) return fail(BGR_ERR_ARG, "bn_stats: too many channels");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  int rpb;
  const int grid = reduce_grid(M, C, &rpb);
  BGR_CUDA_OK(launch_pdl(bn_stats_kernel, dim3(grid), dim3(256), reduce_smem(C), st, static_cast<const __nv_bfloat16*>(x), M, C, ld, rpb, part));
  BGR_LAUNCH_OK();
  *rows = grid;
  return BGR_OK;
}

// ---------------------------------------------------------------- bgr_bn_stats without any allocation by anyone
// The partial rows of the statistics pass live in a file-scope device array instead of a caller-provided workspace,
// for callers that cannot allocate device memory at all (code under CUDA graph capture, tight serving loops,
// harnesses that forbid cudaMalloc).  Every GPU that loads this library holds the array (7.0 MiB) until exit.
// Nothing in it has to start at zero: each of the `rows` blocks overwrites its whole row before the sum reads
// exactly those rows.  All calls share the array, so two calls that overlap on one GPU (different streams) race.
static __device__ __align__(256) float g_bn_stats_scratch[size_t(BGR_BN_MAX_ROWS) * 2 * BGR_BN_STATS_STATIC_MAX_C];

namespace bgr {
// sums[c] = sum over the partial rows (fixed order, fp64) of sum_m x, sums[C + c] = the same of sum_m x^2
__global__ void __launch_bounds__(256)
bn_sums_kernel(const float* __restrict__ part, int rows, int C, double* __restrict__ sums) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  int c;
  double s, q;
  if (!ordered_row_sums(part, rows, C, &c, &s, &q)) return;
  sums[c] = s;
  sums[C + c] = q;
}

// Device address of the scratch array on the CURRENT device.  In host code the array's name is only a host-side
// stand-in for the symbol, not a device address; cudaGetSymbolAddress resolves it, once per device ordinal, with
// that device current (it allocates nothing: the array came with the module).
static int bn_stats_scratch_ptr(float** out) {
  static std::mutex mu;
  static float* cache[64] = {nullptr};
  int dev = 0;
  BGR_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return fail(BGR_ERR_ARG, "bn_stats_static: device ordinal out of range");
  std::lock_guard<std::mutex> lock(mu);
  if (cache[dev] == nullptr) {
    float* ptr = nullptr;
    BGR_CUDA_OK(cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_bn_stats_scratch));
    if (ptr == nullptr) return fail(BGR_ERR_CUDA, "bn_stats_static: cudaGetSymbolAddress returned a null address");
    cache[dev] = ptr;
  }
  *out = cache[dev];
  return BGR_OK;
}
}  // namespace bgr

extern "C" int bgr_bn_stats_static(const void* x, long M, int C, int ld, double* sums, void* stream) {
  BGR_CHECK_C8(C);
  if (C > BGR_BN_STATS_STATIC_MAX_C)
    return fail(BGR_ERR_ARG, "bn_stats_static: C = " + std::to_string(C) + " exceeds BGR_BN_STATS_STATIC_MAX_C = " +
                                 std::to_string(BGR_BN_STATS_STATIC_MAX_C) + " (use bgr_bn_stats with your own workspace)");
  if (M < 1 || sums == nullptr || x == nullptr) return fail(BGR_ERR_ARG, "bn_stats_static: empty input or missing output");
  if (reduce_smem(C) > 48 * 1024) return fail(BGR_ERR_ARG, "bn_stats_static: too many channels");
  float* scratch = nullptr;
  if (int rc = bn_stats_scratch_ptr(&scratch)) return rc;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  int rpb;
  const int grid = reduce_grid(M, C, &rpb);   // <= BGR_BN_MAX_ROWS rows of 2*C floats
  BGR_CUDA_OK(launch_pdl(bn_stats_kernel, dim3(grid), dim3(256), reduce_smem(C), st, static_cast<const __nv_bfloat16*>(x), M, C, ld, rpb, scratch));
  BGR_LAUNCH_OK();
  BGR_CUDA_OK(launch_pdl(bn_sums_kernel, dim3((C + kRsCh - 1) / kRsCh), dim3(256), 0, st, static_cast<const float*>(scratch), grid, C, sums));
  BGR_LAUNCH_OK();
  return BGR_OK;
}

extern "C" int bgr_bn_finalize(const float* part, int rows, long M, int C, const float* gamma, const float* beta,
                               float eps, float momentum, float* mean, float* invstd, float* a, float* b,
                               float* running_mean, float* running_var, long long* num_batches_tracked,
                               void* stream) {
  if (rows < 1 || rows > BGR_BN_MAX_ROWS) return fail(BGR_ERR_ARG, "bn_finalize: rows out of range");
  BGR_CUDA_OK(launch_pdl(bn_finalize_kernel, dim3((C + kRsCh - 1) / kRsCh), dim3(256), 0, reinterpret_cast<cudaStream_t>(stream), 
      part, rows, M, C, gamma, beta, eps, momentum, mean, invstd, a, b, running_mean, running_var, num_batches_tracked));
  BGR_LAUNCH_OK();
  return BGR_OK;
}

extern "C" int bgr_bn_apply(const void* x, int ldx, const float* a, const float* b, const void* res, int ldr,
                            void* out, int ldo, long M, int C, int act, float slope, void* stream) {
  BGR_CHECK_C8(C);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  int grid, block;
  rowgroup_launch(M, C, &grid, &block, 4);
  const size_t ring = bn_ring_bytes(res ? 2 : 1, block);   // <= 32 KB
  auto X = static_cast<const __nv_bfloat16*>(x);
  auto R = static_cast<const __nv_bfloat16*>(res);
  auto O = static_cast<__nv_bfloat16*>(out);
  if (act && res) BGR_CUDA_OK(launch_pdl(bn_apply_kernel<true, true>, dim3(grid), dim3(block), ring, st, X, ldx, a, b, R, ldr, O, ldo, M, C, slope));
  else if (act) BGR_CUDA_OK(launch_pdl(bn_apply_kernel<true, false>, dim3(grid), dim3(block), ring, st, X, ldx, a, b, R, ldr, O, ldo, M, C, slope));
  else if (res) BGR_CUDA_OK(launch_pdl(bn_apply_kernel<false, true>, dim3(grid), dim3(block), ring, st, X, ldx, a, b, R, ldr, O, ldo, M, C, slope));
  else BGR_CUDA_OK(launch_pdl(bn_apply_kernel<false, false>, dim3(grid), dim3(block), ring, st, X, ldx, a, b, R, ldr, O, ldo, M, C, slope));
  BGR_LAUNCH_OK();
  return BGR_OK;
}

// pass 1 with the second-gradient flag as a template argument (no per-row branch in the kernel)
#define BGR_LAUNCH_REDUCE(ACTV, ...)                                                                      \
  do {                                                                                                    \
    if (G2 != nullptr) {                                                                                  \
      if (int rc_ = ensure_dyn_smem(bn_bwd_reduce_kernel<ACTV, true>, 80 * 1024)) return rc_;             \
      BGR_CUDA_OK(launch_pdl(bn_bwd_reduce_kernel<ACTV, true>, __VA_ARGS__));                             \
    } else {                                                                                              \
      if (int rc_ = ensure_dyn_smem(bn_bwd_reduce_kernel<ACTV, false>, 80 * 1024)) return rc_;            \
      BGR_CUDA_OK(launch_pdl(bn_bwd_reduce_kernel<ACTV, false>, __VA_ARGS__));                            \
    }                                                                                                     \
  } while (0)

static int bn_backward_impl(const void* g, int ldg, const void* g2, int ldg2, const void* out, int ldo,
                            const float* xa, const float* xb,
                            const void* x, int ldx, const float* mean, const float* invstd,
                            const float* gamma, int act, float slope, long M, int C, void* dyh, void* dx,
                            int ldd, float* part, float* coef_ws, void* stream) {
  BGR_CHECK_C8(C);
  if (reduce_smem(C) > 48 * 1024) return fail(BGR_ERR_ARG, "bn_backward: too many channels");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  int rpb;
  const int grid = reduce_grid(M, C, &rpb);
  // partial-sum slots + the row ring of pass 1 (g, x [, g2] [, stored output])
  const size_t rsm = reduce_smem(C) + bn_ring_bytes(2 + (g2 != nullptr ? 1 : 0) + (act == 1 ? 1 : 0), 256);
  auto G = static_cast<const __nv_bfloat16*>(g);
  auto G2 = static_cast<const __nv_bfloat16*>(g2);
  auto O = static_cast<const __nv_bfloat16*>(out);
  auto X = static_cast<const __nv_bfloat16*>(x);
  auto D = static_cast<__nv_bfloat16*>(dyh);
  auto DXp = static_cast<__nv_bfloat16*>(dx);
  if (D == nullptr && act == 1) return fail(BGR_ERR_ARG, "bn_backward: dyh may be NULL only on the recompute path");
  // BGR_BN_BWD_FUSED=1: one cooperative launch (pass 1 -> grid barrier -> coefficients -> grid barrier -> pass 2 over
  // the same rows out of L2).  Measured on B200 and NOT adopted (profiles/r02_bn_backward_fused.txt): the two grid
  // barriers and the one-wave grid of the second pass (444 x 256 threads instead of 2368 blocks) cost more than
  // the two saved launches -- 77.8 vs 74.8 us at the 48x72x728 mid-flow shape, 502 vs 474 us at 384x576x128, and
  // the step 129.5 vs 126.8 ms.  Kept opt-in for re-measurement.
  static const int use_fused = env_int("BGR_BN_BWD_FUSED", 0);
  if (use_fused) {
    const int cap = act == 2 ? fused_max_grid<2>(rsm) : (act ? fused_max_grid<1>(rsm) : fused_max_grid<0>(rsm));
    if (cap >= grid && grid * kRsCh >= C) {
      if (act == 2) return launch_bn_bwd_fused<2>(G, ldg, G2, ldg2, nullptr, 0, xa, xb, X, ldx, mean, invstd, gamma, D, DXp, ldd, M, C, rpb, grid, slope, part, coef_ws, rsm, st);
      if (act) return launch_bn_bwd_fused<1>(G, ldg, G2, ldg2, O, ldo, nullptr, nullptr, X, ldx, mean, invstd, gamma, D, DXp, ldd, M, C, rpb, grid, slope, part, coef_ws, rsm, st);
      return launch_bn_bwd_fused<0>(G, ldg, G2, ldg2, O, ldo, nullptr, nullptr, X, ldx, mean, invstd, gamma, D, DXp, ldd, M, C, rpb, grid, slope, part, coef_ws, rsm, st);
    }
  }
  if (act == 2) BGR_LAUNCH_REDUCE(2, dim3(grid), dim3(256), rsm, st, G, ldg, G2, ldg2, nullptr, 0, xa, xb, X, ldx, mean, invstd, D, ldd, M, C, rpb, slope, part);
  else if (act) BGR_LAUNCH_REDUCE(1, dim3(grid), dim3(256), rsm, st, G, ldg, G2, ldg2, O, ldo, nullptr, nullptr, X, ldx, mean, invstd, D, ldd, M, C, rpb, slope, part);
  else BGR_LAUNCH_REDUCE(0, dim3(grid), dim3(256), rsm, st, G, ldg, G2, ldg2, O, ldo, nullptr, nullptr, X, ldx, mean, invstd, D, ldd, M, C, rpb, slope, part);
  BGR_LAUNCH_OK();
  BGR_CUDA_OK(launch_pdl(bn_bwd_coef_kernel, dim3((C + kRsCh - 1) / kRsCh), dim3(256), 0, st, mean, invstd, gamma, part, grid, M, C, coef_ws));
  BGR_LAUNCH_OK();
  int agrid, ablock;
  rowgroup_launch(M, C, &agrid, &ablock, D == nullptr ? 3 : 4);   // 80 / 64 registers per thread
  const size_t aring = bn_ring_bytes(D == nullptr && G2 != nullptr ? 3 : 2, ablock);   // <= 32 KB
  if (D == nullptr) {
#define BGR_LAUNCH_RECOMPUTE(A_, G_)                                                                                      \
  BGR_CUDA_OK(launch_pdl(bn_bwd_apply_recompute_kernel<A_, G_>, dim3(agrid), dim3(ablock), aring, st, G, ldg, G2, ldg2, X, ldx, \
                         xa, xb, coef_ws, DXp, ldd, M, C, slope))
    if (act == 2 && G2 != nullptr) BGR_LAUNCH_RECOMPUTE(true, true);
    else if (act == 2) BGR_LAUNCH_RECOMPUTE(true, false);
    else if (G2 != nullptr) BGR_LAUNCH_RECOMPUTE(false, true);
    else BGR_LAUNCH_RECOMPUTE(false, false);
#undef BGR_LAUNCH_RECOMPUTE
  } else {
    BGR_CUDA_OK(launch_pdl(bn_bwd_apply_kernel, dim3(agrid), dim3(ablock), aring, st, D, X, ldx, coef_ws, DXp, ldd, M, C));
  }
  BGR_LAUNCH_OK();
  return BGR_OK;
}

extern "C" int bgr_bn_backward(const void* g, int ldg, const void* g2, int ldg2, const void* out, int ldo,
                               const void* x, int ldx, const float* mean, const float* invstd,
                               const float* gamma, int act, float slope, long M, int C, void* dyh, void* dx,
                               int ldd, float* part, float* coef_ws, void* stream) {
  return bn_backward_impl(g, ldg, g2, ldg2, out, ldo, nullptr, nullptr, x, ldx, mean, invstd, gamma, act ? 1 : 0, slope, M, C,
                          dyh, dx, ldd, part, coef_ws, stream);
}

// Backward of a BatchNorm(+LeakyReLU) whose output was consumed on the fly (bgr_dw_fwd_bn): the
// activation derivative is recomputed from a*x + b instead of being read from a stored output.
extern "C" int bgr_bn_backward_recompute(const void* g, int ldg, const float* a, const float* b, const void* x,
                                         int ldx, const float* mean, const float* invstd,
                                         const float* gamma, int act, float slope, long M, int C, void* dyh,
                                         void* dx, int ldd, float* part, float* coef_ws, void* stream) {
  return bn_backward_impl(g, ldg, nullptr, 0, nullptr, 0, a, b, x, ldx, mean, invstd, gamma, act ? 2 : 0, slope, M, C, dyh, dx,
                          ldd, part, coef_ws, stream);
}

// The same with a second gradient of the output (a dual-handle block input): both passes add g2 on load.
extern "C" int bgr_bn_backward_recompute_dual(const void* g, int ldg, const void* g2, int ldg2, const float* a,
                                              const float* b, const void* x, int ldx, const float* mean,
                                              const float* invstd, const float* gamma, int act, float slope, long M,
                                              int C, void* dyh, void* dx, int ldd, float* part, float* coef_ws,
                                              void* stream) {
  return bn_backward_impl(g, ldg, g2, ldg2, nullptr, 0, a, b, x, ldx, mean, invstd, gamma, act ? 2 : 0, slope, M, C, dyh, dx,
                          ldd, part, coef_ws, stream);
}

// ---------------------------------------------------------------- frozen statistics (eval mode / Identity)
// dx[m][c] = s[c] * d[m][c]
__global__ void __launch_bounds__(256)
scale_rows_kernel(const __nv_bfloat16* __restrict__ d, int ldd, const float* __restrict__ sc,
                  __nv_bfloat16* __restrict__ dx, int ldx, long M, int C) {
  pdl_wait();      // programmatic dependent launch: see bgr_internal.h
  pdl_trigger();
  const int G = C >> 3;
  const long total = M * G;
  for (long i = long(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += long(gridDim.x) * blockDim.x) {
    const int g = int(i % G);
    const long row = i / G;
    float f[8], sv[8];
    unpack8(*reinterpret_cast<const uint4*>(d + row * ldd + g * 8), f);
    load8f(sc + g * 8, sv);
#pragma unroll
    for (int j = 0; j < 8; ++j) f[j] *= sv[j];
    *reinterpret_cast<uint4*>(dx + row * ldx + g * 8) = pack8(f);
  }
}

// Backward of out = act(a*x + b [+ res]) whose a, b do NOT depend on the batch: nn.BatchNorm2d in eval mode
// (running statistics; train_gan.py:330-398 validation, freeze_bn) or an Identity normaliser
// (--disable_batchnorm, train_gan.py:125).  mean / invstd = the frozen statistics (NULL: 0 / 1), gamma may be
// NULL (1).  dyh = (g [+ g2]) * act'(out) is also dL/dres; dx = gamma*invstd*dyh (dx may alias nothing; NULL
// when gamma*invstd == 1, then dyh is dL/dx); coef_ws rows 5, 6 = dL/dbeta, dL/dgamma.
extern "C" int bgr_bn_backward_frozen(const void* g, int ldg, const void* g2, int ldg2, const void* out, int ldo,
                                      const void* x, int ldx, const float* mean, const float* invstd,
                                      const float* gamma, int act, float slope, long M, int C, void* dyh, void* dx,
                                      int ldd, float* part, float* coef_ws, void* stream) {
  BGR_CHECK_C8(C);
  if (reduce_smem(C) > 48 * 1024) return fail(BGR_ERR_ARG, "bn_backward_frozen: too many channels");
  if (dyh == nullptr) return fail(BGR_ERR_ARG, "bn_backward_frozen: dyh is required");
  if (mean == nullptr || invstd == nullptr) return fail(BGR_ERR_ARG, "bn_backward_frozen: mean / invstd vectors are required");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  int rpb;
  const int grid = reduce_grid(M, C, &rpb);
  // partial-sum slots + the row ring of pass 1 (g, x [, g2] [, stored output])
  const size_t rsm = reduce_smem(C) + bn_ring_bytes(2 + (g2 != nullptr ? 1 : 0) + (act ? 1 : 0), 256);
  auto G = static_cast<const __nv_bfloat16*>(g);
  auto G2 = static_cast<const __nv_bfloat16*>(g2);
  auto O = static_cast<const __nv_bfloat16*>(out);
  auto X = static_cast<const __nv_bfloat16*>(x);
  auto D = static_cast<__nv_bfloat16*>(dyh);
  if (act) BGR_LAUNCH_REDUCE(1, dim3(grid), dim3(256), rsm, st, G, ldg, G2, ldg2, O, ldo, nullptr, nullptr, X, ldx, mean, invstd, D, ldd, M, C, rpb, slope, part);
  else BGR_LAUNCH_REDUCE(0, dim3(grid), dim3(256), rsm, st, G, ldg, G2, ldg2, O, ldo, nullptr, nullptr, X, ldx, mean, invstd, D, ldd, M, C, rpb, slope, part);
  BGR_LAUNCH_OK();
  BGR_CUDA_OK(launch_pdl(bn_bwd_coef_kernel, dim3((C + kRsCh - 1) / kRsCh), dim3(256), 0, st, mean, invstd, gamma, part, grid, M, C, coef_ws));
  BGR_LAUNCH_OK();
  if (dx != nullptr) {
    long blocks = (M * (C / 8) + 255) / 256;
    if (blocks > long(sm_count()) * 16) blocks = long(sm_count()) * 16;
    BGR_CUDA_OK(launch_pdl(scale_rows_kernel, dim3(int(blocks < 1 ? 1 : blocks)), dim3(256), 0, st, D, ldd, coef_ws /* row 0 = gamma*invstd */,
                                                                    static_cast<__nv_bfloat16*>(dx), ldd, M, C));
    BGR_LAUNCH_OK();
  }
  return BGR_OK;
}
