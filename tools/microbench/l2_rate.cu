// What bounds operand delivery from L2 to shared memory: the SM's own intake, or the L2's output for the whole chip -
// and do CTAs that fetch the SAME tile at about the same time share one L2 read?
// `active` CTAs (one per SM), three issuing warps each, every warp keeps four 16 KB stages in flight and streams
// 128-row x 128-byte tiles (TMA tiled, 128-byte swizzle) out of a 96 MB bf16 matrix that was read once before the
// timed loop (L2-resident).  `share` consecutive CTAs walk the same tile sequence in lockstep order; share = 1 gives
// every CTA its own tiles.  Prints bytes per clock per SM and for the chip.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I../../distributed-machine-learning-cluster_b200/csrc \
//        l2_rate.cu ../../distributed-machine-learning-cluster_b200/csrc/conv_umma.cu -o l2_rate -lcuda
#include <cstdio>
#include <cuda_runtime.h>
#include "conv.h"
#include "ptx.cuh"
using namespace b2i;

constexpr int NW = 3, NST = 4;
constexpr uint32_t STAGE = 16 * 1024;

__global__ void __launch_bounds__(32 * NW, 1) l2_kernel(const __grid_constant__ CUtensorMap tm, int iters, int share, int tiles_total,
                                                        int kblocks, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bars_all[NW * NST];
  const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint64_t* bars = bars_all + NST * wid;
  if (lane == 0) { for (int s = 0; s < NST; ++s) mbar_init(smem_u32(&bars[s]), 1); mbar_fence_init(); }
  __syncthreads();
  const uint32_t base = ((smem_u32(smem_raw) + 1023u) & ~1023u) + wid * NST * STAGE;
  const int group = blockIdx.x / share;                   // CTAs of a group read the same tiles
  const int groups = (gridDim.x + share - 1) / share;
  const int per_group = tiles_total / groups;             // a private range of tiles per group
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    const int s = it % NST;
    if (it >= NST) mbar_wait(smem_u32(&bars[s]), ((it / NST) - 1) & 1);
    const int tile = group * per_group + (it * NW + wid) % per_group;
    const int row = (tile / kblocks) * 128, col = (tile % kblocks) * 64;
    if (lane == 0) {
      mbar_expect_tx(smem_u32(&bars[s]), STAGE);
      tma_load_2d(base + s * STAGE, &tm, smem_u32(&bars[s]), col, row);
    }
    __syncwarp();
  }
  for (int s = 0; s < NST; ++s) {
    const int it_last = iters - 1 - ((iters - 1 - s) % NST + NST) % NST;
    mbar_wait(smem_u32(&bars[s]), (it_last / NST) & 1);
  }
  const long long t1 = clock64();
  if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
}

int main() {
  const int K = 4096, kblocks = K / 64;                    // 8 KB per row
  const int rows = 12288;                                  // 96 MB
  const int tiles_total = (rows / 128) * kblocks;          // 6144 tiles of 16 KB
  void* d = nullptr;
  cudaMalloc(&d, static_cast<size_t>(rows) * K * 2);
  cudaMemset(d, 1, static_cast<size_t>(rows) * K * 2);
  CUtensorMap tm;
  const char* err = "";
  if (make_tmap_rows(&tm, d, rows, K, 128, &err)) { printf("tmap: %s\n", err); return 1; }
  long long* d_out;
  cudaMalloc(&d_out, 148 * sizeof(long long));
  cudaFuncSetAttribute(l2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  const int iters = 600;
  for (int active : {148, 74, 37}) {
    for (int share : {1, 2, 4, 8, 148}) {
      if (share > active) continue;
      long long h[148];
      double best = 0;
      for (int rep = 0; rep < 3; ++rep) {                  // first repetition warms L2 for this pattern
        l2_kernel<<<active, 32 * NW, NW * NST * STAGE + 2048>>>(tm, iters, share, tiles_total, kblocks, d_out);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("error: %s\n", cudaGetErrorString(e)); return 1; }
        cudaMemcpy(h, d_out, active * sizeof(long long), cudaMemcpyDeviceToHost);
        long long mx = 0;
        for (int i = 0; i < active; ++i) mx = h[i] > mx ? h[i] : mx;
        const double per_sm = static_cast<double>(iters) * NW * STAGE / mx;
        if (rep) best = per_sm > best ? per_sm : best;
      }
      printf("active SMs %3d, %3d CTAs per tile sequence: %6.1f B/clk per SM, %7.0f B/clk for the chip\n", active, share, best, best * active);
    }
  }
  return 0;
}
