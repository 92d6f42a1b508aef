// Do TMA loads into shared memory slow the tensor pipe down?  CTA pairs (cta_group::2, M = 256, N = 256, bf16, operands
// in shared memory, three 32 KB stages reused over and over) issue MMAs back to back while `np` warps per CTA stream
// 16 KB tiles (TMA tiled, 128-byte swizzle, L2-resident source) into a separate region of the same shared memory.
// Prints cycles per MMA (floor 128) and the TMA intake in bytes per clock per SM, for np = 0 .. 3.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I../../distributed-machine-learning-cluster_b200/csrc \
//        mma_tma_mix.cu ../../distributed-machine-learning-cluster_b200/csrc/conv_umma.cu -o mma_tma_mix -lcuda
#include <cstdio>
#include <cuda_runtime.h>
#include "conv.h"
#include "ptx.cuh"
using namespace b2i;

__device__ __forceinline__ void mma2(uint32_t d, uint64_t a, uint64_t b, uint32_t id, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(a), "l"(b), "r"(id), "r"(acc) : "memory");
}

constexpr int MMA_STAGES = 3, TMA_NST = 2;
constexpr uint32_t TILE = 16 * 1024;

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(192, 1)
mix_kernel(const __grid_constant__ CUtensorMap tm, int iters, int np, int N, int tiles_total, int kblocks, int commit_each, int wait_each, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bars[1 + 4 * TMA_NST];
  __shared__ __align__(8) uint64_t kb_bar[MMA_STAGES];
  __shared__ uint32_t slot;
  __shared__ volatile int stop;
  __shared__ long long tma_bytes[4];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint32_t rank;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
  if (threadIdx.x == 0) {
    for (int i = 0; i < 1 + 4 * TMA_NST; ++i) mbar_init(smem_u32(&bars[i]), 1);
    for (int i = 0; i < MMA_STAGES; ++i) mbar_init(smem_u32(&kb_bar[i]), 1);
    mbar_fence_init();
    stop = 0;
  }
  for (uint32_t i = threadIdx.x; i < 200 * 1024 / 16; i += blockDim.x)
    asm volatile("st.shared.v4.u32 [%0], {%1,%1,%1,%1};" ::"r"(base + i * 16), "r"(0u));
  fence_proxy_async_smem();
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&slot)), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  asm volatile("barrier.cluster.arrive.release.aligned;");
  asm volatile("barrier.cluster.wait.acquire.aligned;");
  tc_fence_after();
  const uint32_t tmem = slot;
  const uint32_t stage_bytes = 16384 + (N / 2) * 128;
  if (warp == 1) {
    if (lane == 0 && rank == 0) {
      const uint32_t idesc = umma_idesc_bf16(256, N);
      const long long t0 = clock64();
      for (int it = 0; it < iters; ++it) {
        const uint32_t a_s = base + (it % MMA_STAGES) * stage_bytes;
        const uint64_t ad = umma_smem_desc_sw128(a_s), bd = umma_smem_desc_sw128(a_s + 16384);
#pragma unroll
        for (int k = 0; k < 4; ++k) mma2(tmem, ad + 2u * k, bd + 2u * k, idesc, 1u);
        // a k-block's stage handed back the way the conv kernels do it: one multicast commit per four MMAs; with
        // wait_each the issuer also waits for the commit of the stage it is about to reuse (as a full / empty ring would)
        if (commit_each)
          asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(smem_u32(&kb_bar[it % MMA_STAGES])), "h"((uint16_t)3) : "memory");
        if (wait_each && it + 1 >= MMA_STAGES) {
          const int j = it + 1 - MMA_STAGES;
          mbar_wait(smem_u32(&kb_bar[j % MMA_STAGES]), (j / MMA_STAGES) & 1);
        }
      }
      asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(smem_u32(&bars[0])), "h"((uint16_t)1) : "memory");
      mbar_wait(smem_u32(&bars[0]), 0);
      const long long t1 = clock64();
      out[blockIdx.x * 2] = t1 - t0;
    }
    if (lane == 0) {
      if (rank != 0) {   // the peer's loaders run for the same time: it watches its own copy of the clock
        const long long t0 = clock64();
        while (clock64() - t0 < static_cast<long long>(iters) * 4 * (N / 2)) {}
      }
      stop = 1;
    }
  } else if (warp >= 2 && warp - 2 < np) {
    const int me = warp - 2;
    uint64_t* b = bars + 1 + me * TMA_NST;
    const uint32_t dst0 = base + MMA_STAGES * stage_bytes + me * TMA_NST * TILE;
    long long bytes = 0;
    const long long t0 = clock64();
    int it = 0;
    if (lane == 0) {
      for (; !stop; ++it) {
        const int s = it % TMA_NST;
        if (it >= TMA_NST) mbar_wait(smem_u32(&b[s]), ((it / TMA_NST) - 1) & 1);
        const int tile = ((blockIdx.x * 4 + me) * 97 + it) % tiles_total;
        mbar_expect_tx(smem_u32(&b[s]), TILE);
        tma_load_2d(dst0 + s * TILE, &tm, smem_u32(&b[s]), (tile % kblocks) * 64, (tile / kblocks) * 128);
        bytes += TILE;
      }
      for (int j = it < TMA_NST ? 0 : it - TMA_NST; j < it; ++j) mbar_wait(smem_u32(&b[j % TMA_NST]), (j / TMA_NST) & 1);
      const long long t1 = clock64();
      tma_bytes[me] = bytes;
      out[blockIdx.x * 2 + 1] = np > 0 ? (me == 0 ? t1 - t0 : out[blockIdx.x * 2 + 1]) : 0;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x == 0 && np > 0) {
    long long tot = 0;
    for (int i = 0; i < np; ++i) tot += tma_bytes[i];
    out[296 + blockIdx.x] = tot;
  }
  asm volatile("barrier.cluster.arrive.release.aligned;");
  asm volatile("barrier.cluster.wait.acquire.aligned;");
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
}

int main() {
  const int K = 4096, kblocks = K / 64, rows = 8192;      // 64 MB, L2-resident after the first pass
  const int tiles_total = (rows / 128) * kblocks;
  void* d = nullptr;
  cudaMalloc(&d, static_cast<size_t>(rows) * K * 2);
  cudaMemset(d, 0, static_cast<size_t>(rows) * K * 2);
  CUtensorMap tm;
  const char* err = "";
  if (make_tmap_rows(&tm, d, rows, K, 128, &err)) { printf("tmap: %s\n", err); return 1; }
  long long* d_out;
  cudaMalloc(&d_out, 444 * sizeof(long long));
  cudaFuncSetAttribute(mix_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
  const int iters = 3000;
  for (int mode = 0; mode <= 2; ++mode)
  for (int N : {256, 128}) {
    for (int np = 0; np <= 3; np += 3) {
      long long h[444];
      for (int rep = 0; rep < 2; ++rep) {
        cudaMemset(d_out, 0, 444 * sizeof(long long));
        mix_kernel<<<148, 192, 210 * 1024>>>(tm, iters, np, N, tiles_total, kblocks, mode >= 1, mode >= 2, d_out);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("error: %s\n", cudaGetErrorString(e)); return 1; }
        cudaMemcpy(h, d_out, 444 * sizeof(long long), cudaMemcpyDeviceToHost);
      }
      long long mma_cycles = 0;
      for (int c = 0; c < 148; c += 2) mma_cycles = h[c * 2] > mma_cycles ? h[c * 2] : mma_cycles;
      double tma = 0;
      for (int c = 0; c < 148; ++c) tma += static_cast<double>(h[296 + c]);
      printf("%s N %3d, %d loading warp(s) per CTA: %6.1f cycles per MMA (floor %d); TMA intake %5.1f B/clk per SM; shared-memory traffic %5.1f B/clk per SM\n",
             mode == 0 ? "one commit at the end;   " : (mode == 1 ? "commit per 4 MMAs;       " : "commit + ring wait per 4;"), N, np, static_cast<double>(mma_cycles) / (iters * 4.0), N / 2, tma / 148.0 / mma_cycles,
             tma / 148.0 / mma_cycles + (4096.0 + N / 2 * 32.0) * iters * 4.0 / mma_cycles);
    }
  }
  return 0;
}
