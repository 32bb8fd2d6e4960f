// Development microbenchmark: how many bytes per clock one SM can take from L2 through the bulk-copy (TMA) path
// when every SM streams, unicast against cluster multicast.  The implicit-GEMM forward kernel waits on its operand
// loads 40 % of the time (profiles/r02_conv_roles.txt); this decides whether multicast would lift that bound.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/bin/mc_bench tools/mc_bench.cu
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

constexpr int kMaxStages = 6;
constexpr uint32_t kChunk = 32768;

__device__ __forceinline__ uint32_t s32(const void* p) { return uint32_t(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ uint32_t mapa(uint32_t a, uint32_t r) {
  uint32_t o;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(o) : "r"(a), "r"(r));
  return o;
}
__device__ __forceinline__ void wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok = 0, spins = 0;
  while (!ok) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(s32(bar)), "r"(parity)
        : "memory");
    if (++spins > (1u << 24)) { printf("timeout block %d thread %d\n", blockIdx.x, threadIdx.x); __trap(); }
  }
}

// mode 0: every CTA of a cluster copies the same chunk itself; 1: each copies 1/CS of it and multicasts to all;
// 2: every CTA copies its own chunk; 3: tensor-map (tiled, 128-byte swizzle) boxes of 64 bf16 x 256 rows, one per
// stage; 4: two boxes of 64 x 128 rows per stage (what a CTA of the implicit-GEMM pair loads per k-block)
__global__ void __launch_bounds__(64, 1) stream_kernel(const uint8_t* src, uint32_t nchunks, int iters, int mode,
                                                       long long* cycles, int kStages,
                                                       const __grid_constant__ CUtensorMap tm256,
                                                       const __grid_constant__ CUtensorMap tm128, uint32_t kblocks,
                                                       uint32_t rowblocks) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* stage = smem;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + kMaxStages * kChunk);
  uint64_t* empty = full + kMaxStages;
  uint32_t rank, cs, cid;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
  asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(cs));
  asm volatile("mov.u32 %0, %%clusterid.x;" : "=r"(cid));
  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(&full[s])), "r"(1));
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(&empty[s])), "r"(mode == 1 ? cs : 1u));
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  if (threadIdx.x == 0) {
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      const uint32_t s = it % kStages, ph = (it / kStages) & 1;
      wait(&empty[s], ph ^ 1);
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(&full[s])), "r"(kChunk) : "memory");
      const uint32_t owner = mode == 2 ? blockIdx.x : cid;
      const uint8_t* g = src + size_t((owner * 131u + uint32_t(it) * 17u) % nchunks) * kChunk;
      if (mode >= 3) {
        const int c0 = int(((owner * 5u + uint32_t(it)) % kblocks) * 64u);
        const int r0 = int(((owner * 131u + uint32_t(it) * 17u) % rowblocks) * 256u);
        if (mode == 3) {
          asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::
                           "r"(s32(stage + s * kChunk)), "l"(&tm256), "r"(s32(&full[s])), "r"(c0), "r"(r0)
                       : "memory");
        } else {
          asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::
                           "r"(s32(stage + s * kChunk)), "l"(&tm128), "r"(s32(&full[s])), "r"(c0), "r"(r0)
                       : "memory");
          asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::
                           "r"(s32(stage + s * kChunk + kChunk / 2)), "l"(&tm128), "r"(s32(&full[s])), "r"(c0), "r"(r0 + 128)
                       : "memory");
        }
      } else if (mode == 1) {
        const uint32_t part = kChunk / cs;
        const uint16_t mask = uint16_t((1u << cs) - 1);
        asm volatile(
            "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1], %2, [%3], %4;" ::
                "r"(s32(stage + s * kChunk + rank * part)), "l"(g + rank * part), "r"(part), "r"(s32(&full[s])), "h"(mask)
            : "memory");
      } else {
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
                         "r"(s32(stage + s * kChunk)), "l"(g), "r"(kChunk), "r"(s32(&full[s]))
                     : "memory");
      }
    }
    cycles[blockIdx.x] = clock64() - t0;
  } else if (threadIdx.x == 32) {
    for (int it = 0; it < iters; ++it) {
      const uint32_t s = it % kStages, ph = (it / kStages) & 1;
      wait(&full[s], ph);
      if (mode == 1) {
        for (uint32_t r = 0; r < cs; ++r)
          asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(mapa(s32(&empty[s]), r)) : "memory");
      } else {
        asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(s32(&empty[s])) : "memory");
      }
    }
  }
  __syncwarp();
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

#define OK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static void make_map(CUtensorMap* m, void* base, uint64_t cols, uint64_t rows, uint64_t pitch_bytes, uint32_t box_rows) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  OK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
  cuuint64_t gdim[2] = {cols, rows}, gstr[1] = {pitch_bytes};
  cuuint32_t box[2] = {64, box_rows}, estr[2] = {1, 1};
  CUresult r = reinterpret_cast<EncodeFn>(p)(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, base, gdim, gstr, box, estr,
                                             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                             CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { printf("encode failed %d\n", int(r)); exit(1); }
}

int main() {
  const size_t smem = kMaxStages * kChunk + 2 * kMaxStages * 8;
  OK(cudaFuncSetAttribute(stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
  uint8_t* buf;
  long long* cyc;
  const size_t bytes = size_t(96) << 20;
  OK(cudaMalloc(&buf, bytes));
  OK(cudaMemset(buf, 1, bytes));
  OK(cudaMalloc(&cyc, 148 * sizeof(long long)));
  const int iters = 4000;
  printf("%8s %6s %5s %3s %7s %10s %12s %10s\n", "buffer", "pitch", "mode", "cs", "stages", "us", "GB/s total", "B/clk/SM");
  for (size_t mb : {size_t(1), size_t(64)}) {
    for (uint32_t pitch : {1536u, 1456u, 128u}) {
      for (int mode : {2, 3, 4}) {
        if (mode == 2 && pitch != 1536u) continue;
        for (int stages : {2, 4, 6}) {
          const uint32_t nchunks = uint32_t((mb << 20) / kChunk);
          const uint64_t rows = (mb << 20) / pitch / 256 * 256;
          CUtensorMap tm256, tm128;
          make_map(&tm256, buf, pitch / 2 / 64 * 64, rows, pitch, 256);
          make_map(&tm128, buf, pitch / 2 / 64 * 64, rows, pitch, 128);
          cudaLaunchConfig_t cfg = {};
          cfg.gridDim = dim3(148);
          cfg.blockDim = dim3(64);
          cfg.dynamicSmemBytes = smem;
          cudaLaunchAttribute attr[1];
          attr[0].id = cudaLaunchAttributeClusterDimension;
          attr[0].val.clusterDim.x = 1;
          attr[0].val.clusterDim.y = 1;
          attr[0].val.clusterDim.z = 1;
          cfg.attrs = attr;
          cfg.numAttrs = 1;
          cudaEvent_t e0, e1;
          OK(cudaEventCreate(&e0));
          OK(cudaEventCreate(&e1));
          for (int rep = 0; rep < 2; ++rep) {
            OK(cudaEventRecord(e0));
            OK(cudaLaunchKernelEx(&cfg, stream_kernel, (const uint8_t*)buf, nchunks, iters, mode, cyc, stages, tm256, tm128,
                                  uint32_t(pitch / 128), uint32_t(rows / 256)));
            OK(cudaEventRecord(e1));
            OK(cudaDeviceSynchronize());
          }
          float ms = 0;
          OK(cudaEventElapsedTime(&ms, e0, e1));
          long long h[148];
          OK(cudaMemcpy(h, cyc, sizeof(long long) * 148, cudaMemcpyDeviceToHost));
          double avg = 0;
          for (unsigned i = 0; i < 148; ++i) avg += double(h[i]);
          avg /= 148;
          const double total = 148.0 * iters * kChunk;
          printf("%6zuMB %6u %5d %3d %7d %10.1f %12.0f %10.1f\n", mb, pitch, mode, 1, stages, ms * 1e3,
                 total / (ms * 1e-3) / 1e9, double(iters) * kChunk / avg);
        }
      }
    }
  }
  return 0;
}
