// Microbenchmark: issue rate of tcgen05.mma (kind::f16, bf16 -> fp32) with both operands resident in
// shared memory, no loads in the loop.  One CTA per SM (cta_group::1) or one CTA pair per TPC
// (cta_group::2).  Prints cycles per MMA and the implied TFLOP/s at the measured clock.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I../../distributed-machine-learning-cluster_b200/csrc umma_rate.cu -o umma_rate
#include <cstdio>
#include <cuda_runtime.h>
#include "ptx.cuh"
using namespace b2i;

__device__ __forceinline__ void mma1(uint32_t d, uint64_t a, uint64_t b, uint32_t id, uint32_t acc) { umma_bf16(d, a, b, id, acc); }
__device__ __forceinline__ void mma2(uint32_t d, uint64_t a, uint64_t b, uint32_t id, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(a), "l"(b), "r"(id), "r"(acc) : "memory");
}

// A-descriptor variants (cta_group::1, N = 64): start offset / SBO / layout, to see what unaligned
// "shifted view" operands cost.
__global__ void __launch_bounds__(128, 1) adesc_kernel(int iters, uint32_t a_off, uint32_t sbo, uint32_t lbo, int swizzled, long long* cycles) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t slot;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) { mbar_init(smem_u32(&bar), 1); mbar_fence_init(); }
  for (uint32_t i = threadIdx.x; i < 200 * 1024 / 16; i += blockDim.x)
    asm volatile("st.shared.v4.u32 [%0], {%1,%1,%1,%1};" ::"r"(base + i * 16), "r"(0u));
  fence_proxy_async_smem();
  if (warp == 0) tmem_alloc(smem_u32(&slot), 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = slot;
  if (warp == 1 && lane == 0) {
    const uint32_t idesc = umma_idesc_bf16(128, 64);
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      const uint32_t a_s = base + (it & 3) * 24576 + a_off;
      uint64_t ad = static_cast<uint64_t>((a_s & 0x3FFFFu) >> 4) | (static_cast<uint64_t>(lbo >> 4) << 16) |
                    (static_cast<uint64_t>(sbo >> 4) << 32) | (1ull << 46) | (swizzled ? (2ull << 61) : 0ull);
      const uint64_t bd = umma_smem_desc_sw128(base + 100 * 1024 + (it & 3) * 8192);
#pragma unroll
      for (int k = 0; k < 4; ++k) umma_bf16(tm, ad + (swizzled ? 2u * k : 0u), bd + 2u * k, idesc, 1u);
    }
    umma_commit(smem_u32(&bar));
    mbar_wait(smem_u32(&bar), 0);
    long long t1 = clock64();
    if (blockIdx.x == 0) cycles[0] = t1 - t0;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tm, 512);
}

template <int CG>
__global__ void __launch_bounds__(128, 1) rate_kernel(int N, int iters, int same_stage, long long* cycles) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t slot;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint32_t rank = 0;
  if (CG == 2) asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
  if (threadIdx.x == 0) { mbar_init(smem_u32(&bar), 1); mbar_fence_init(); }
  // zero the operand area (finite values)
  for (uint32_t i = threadIdx.x; i < 200 * 1024 / 16; i += blockDim.x)
    asm volatile("st.shared.v4.u32 [%0], {%1,%1,%1,%1};" ::"r"(base + i * 16), "r"(0u));
  fence_proxy_async_smem();
  if (warp == 0) {
    if (CG == 1) tmem_alloc(smem_u32(&slot), 512);
    else {
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&slot)), "r"(512u) : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
  }
  tc_fence_before();
  __syncthreads();
  if (CG == 2) { asm volatile("barrier.cluster.arrive.release.aligned;"); asm volatile("barrier.cluster.wait.acquire.aligned;"); }
  tc_fence_after();
  const uint32_t tm = slot;
  if (warp == 1 && lane == 0 && rank == 0) {
    const uint32_t idesc = umma_idesc_bf16(CG == 2 ? 256 : 128, N);
    const uint32_t stage_bytes = 16384 + (CG == 2 ? N / 2 : N) * 128;
    const int stages = same_stage ? 1 : 4;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      const uint32_t a_s = base + (it % stages) * stage_bytes;
      const uint64_t ad = umma_smem_desc_sw128(a_s), bd = umma_smem_desc_sw128(a_s + 16384);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        if (CG == 1) mma1(tm, ad + 2u * k, bd + 2u * k, idesc, 1u);
        else mma2(tm, ad + 2u * k, bd + 2u * k, idesc, 1u);
      }
    }
    if (CG == 1) umma_commit(smem_u32(&bar));
    else asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(smem_u32(&bar)), "h"((uint16_t)1) : "memory");
    mbar_wait(smem_u32(&bar), 0);
    long long t1 = clock64();
    if (blockIdx.x == 0) cycles[0] = t1 - t0;
  }
  tc_fence_before();
  __syncthreads();
  if (CG == 2) { asm volatile("barrier.cluster.arrive.release.aligned;"); asm volatile("barrier.cluster.wait.acquire.aligned;"); }
  if (warp == 0) {
    if (CG == 1) tmem_dealloc(tm, 512);
    else asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tm), "r"(512u) : "memory");
  }
}

// A operand in TENSOR MEMORY (tcgen05.mma [d], [a], bdesc): cycles per MMA by N and by the number of independent
// accumulate chains (different D columns) the issuing thread rotates through, and with two issuing warps.
__global__ void __launch_bounds__(128, 1) ts_kernel(int N, int chains, int issuers, int iters, long long* cycles) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar[2];
  __shared__ uint32_t slot;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) { mbar_init(smem_u32(&bar[0]), 1); mbar_init(smem_u32(&bar[1]), 1); mbar_fence_init(); }
  for (uint32_t i = threadIdx.x; i < 200 * 1024 / 16; i += blockDim.x)
    asm volatile("st.shared.v4.u32 [%0], {%1,%1,%1,%1};" ::"r"(base + i * 16), "r"(0u));
  fence_proxy_async_smem();
  if (warp == 0) tmem_alloc(smem_u32(&slot), 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = slot;
  {  // zero the A region (columns 384..447) through tcgen05.st
    const uint32_t row = tm + (static_cast<uint32_t>(warp * 32) << 16) + 384;
    for (int c = 0; c < 64; c += 8)
      asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%1,%1,%1,%1,%1,%1,%1};" ::"r"(row + c), "r"(0u) : "memory");
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  // The issuing loop runs WARP-UNIFORM (every lane executes it, one elected lane issues): a loop under `lane == 0`
  // costs ~15-20 instructions per MMA of descriptor traffic into uniform registers, which is what an earlier
  // version of this benchmark measured (161 cycles per MMA whatever N) instead of the tensor pipe.
  const int uwarp = __shfl_sync(0xffffffffu, warp, 0);
  const uint32_t utm = __shfl_sync(0xffffffffu, tm, 0);
  if (uwarp == 1 || (uwarp == 2 && issuers == 2)) {
    const int me = uwarp - 1;
    const uint32_t idesc = umma_idesc_bf16(128, N);
    const uint32_t d0 = utm + static_cast<uint32_t>((me * chains * N) % 384);
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      const uint64_t bd = umma_smem_desc_sw128(base + (it & 3) * 32768);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const uint32_t d = d0 + static_cast<uint32_t>(((it * 4 + k) & (chains - 1)) * N);   // chains is a power of two
        asm volatile("{\n\t.reg .pred p, pe;\n\telect.sync _|pe, 0xffffffff;\n\tsetp.ne.b32 p, %4, 0;\n\t@pe tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d),
                     "r"(utm + 384 + 8 * k), "l"(bd + 2u * k), "r"(idesc), "r"(1u) : "memory");
      }
    }
    umma_commit_elect(smem_u32(&bar[me]));
    mbar_wait(smem_u32(&bar[me]), 0);
    long long t1 = clock64();
    if (blockIdx.x == 0 && me == 0 && lane == 0) cycles[0] = t1 - t0;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tm, 512);
}

int main() {
  long long* d_cyc;
  cudaMalloc(&d_cyc, 8);
  int clk_khz = 0;
  cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
  const int iters = 2000;
  const size_t smem = 210 * 1024;
  cudaFuncSetAttribute(rate_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaFuncSetAttribute(rate_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  for (int cg = 1; cg <= 2; ++cg)
    for (int same = 0; same <= 1; ++same)
      for (int N : {64, 128, 256}) {
        for (int rep = 0; rep < 2; ++rep) {
          cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
          cudaEventRecord(e0);
          if (cg == 1) rate_kernel<1><<<148, 128, smem>>>(N, iters, same, d_cyc);
          else {
            cudaLaunchConfig_t cfg = {}; cfg.gridDim = dim3(148); cfg.blockDim = dim3(128); cfg.dynamicSmemBytes = smem;
            cudaLaunchAttribute at[1]; at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
            cfg.attrs = at; cfg.numAttrs = 1;
            cudaLaunchKernelEx(&cfg, rate_kernel<2>, N, iters, same, d_cyc);
          }
          cudaEventRecord(e1);
          cudaError_t err = cudaDeviceSynchronize();
          float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
          long long cyc = 0; cudaMemcpy(&cyc, d_cyc, 8, cudaMemcpyDeviceToHost);
          if (rep == 1) {
            const double per_mma = (double)cyc / (iters * 4.0);
            const double M = cg == 2 ? 256 : 128;
            const double units = cg == 2 ? 74 : 148;
            const double flops = 2.0 * M * N * 16 * iters * 4.0 * units;
            printf("cta_group %d N %3d %s: %7.1f cycles/MMA (ideal %3d) kernel %.3f ms -> %.0f TFLOP/s chip  [%s]\n", cg, N,
                   same ? "one stage   " : "four stages ", per_mma, N / 2, ms, flops / (ms * 1e-3) / 1e12, cudaGetErrorString(err));
          }
        }
      }
  cudaFuncSetAttribute(ts_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  for (int N : {32, 64, 128, 256})
    for (int issuers : {1, 2})
      for (int chains : {1, 2, 4}) {
        if (N * chains * issuers > 384) continue;
        for (int rep = 0; rep < 2; ++rep) {
          ts_kernel<<<148, 128, smem>>>(N, chains, issuers, iters, d_cyc);
          cudaError_t err = cudaDeviceSynchronize();
          long long cyc = 0; cudaMemcpy(&cyc, d_cyc, 8, cudaMemcpyDeviceToHost);
          if (rep == 1) printf("A in TMEM, N %3d, %d issuing warp(s), %d chain(s) each: %6.1f cycles per MMA per SM (floor %3d) [%s]\n", N, issuers, chains,
                               (double)cyc / (iters * 4.0 * issuers), N / 2, cudaGetErrorString(err));
        }
      }
  cudaFuncSetAttribute(adesc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  struct V { const char* name; uint32_t off, sbo, lbo; int sw; } vs[] = {
      {"SW128 aligned           off    0 SBO 1024", 0, 1024, 16, 1},
      {"SW128 shifted view      off  128 SBO 1024", 128, 1024, 16, 1},
      {"SW128 shifted view      off  384 SBO 1280", 384, 1280, 16, 1},
      {"SW128 aligned pitch     off 1024 SBO 2048", 1024, 2048, 16, 1},
      {"none  aligned           off    0 SBO  640 LBO 16", 0, 640, 16, 0},
      {"none  +32B              off   32 SBO  640 LBO 16", 32, 640, 16, 0},
      {"none  +64B              off   64 SBO  640 LBO 16", 64, 640, 16, 0},
      {"none  aligned pitch768  off    0 SBO  768 LBO 16", 0, 768, 16, 0},
      {"none  canonical         off    0 SBO  128 LBO 2048", 0, 128, 2048, 0},
  };
  for (auto& v : vs) {
    for (int rep = 0; rep < 2; ++rep) {
      adesc_kernel<<<148, 128, smem>>>(iters, v.off, v.sbo, v.lbo, v.sw, d_cyc);
      cudaError_t err = cudaDeviceSynchronize();
      long long cyc = 0; cudaMemcpy(&cyc, d_cyc, 8, cudaMemcpyDeviceToHost);
      if (rep == 1) printf("A desc %-48s: %6.1f cycles/MMA (N=64, floor 48) [%s]\n", v.name, (double)cyc / (iters * 4.0), cudaGetErrorString(err));
    }
  }
  printf("clock attr %d kHz\n", clk_khz);
  return 0;
}
