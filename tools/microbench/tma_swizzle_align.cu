// Does a SWIZZLE_128B tiled TMA load to a shared-memory destination that is only 128-byte aligned place
// its 16-byte chunks by ABSOLUTE address bits (chunk ^= (addr >> 7) & 7) or relative to the box start?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I../../distributed-machine-learning-cluster_b200/csrc \
//        tma_swizzle_align.cu ../../distributed-machine-learning-cluster_b200/csrc/conv_umma.cu -o tma_swizzle_align -lcuda
#include <cstdio>
#include <vector>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include "conv.h"
#include "ptx.cuh"
using namespace b2i;

__device__ __forceinline__ void tma4d(uint32_t dst, const CUtensorMap* tm, uint32_t bar, int c0, int c1, int c2, int c3) {
  asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(dst), "l"(tm), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}

__global__ void k(const __grid_constant__ CUtensorMap tm, int shift, uint32_t* out) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  for (int i = threadIdx.x; i < 1024; i += blockDim.x) reinterpret_cast<uint32_t*>(smem_raw + (base - smem_u32(smem_raw)))[i] = 0xFFFFFFFFu;
  if (threadIdx.x == 0) { mbar_init(smem_u32(&bar), 1); mbar_fence_init(); }
  fence_proxy_async_smem();
  __syncthreads();
  if (threadIdx.x == 0) {
    mbar_expect_tx(smem_u32(&bar), 8 * 128);
    tma4d(base + shift * 128, &tm, smem_u32(&bar), 0, 0, 0, 0);
  }
  mbar_wait(smem_u32(&bar), 0);
  __syncthreads();
  for (int i = threadIdx.x; i < 1024; i += blockDim.x) out[i] = reinterpret_cast<uint32_t*>(smem_raw + (base - smem_u32(smem_raw)))[i];
}

int main() {
  std::vector<__nv_bfloat16> h(16 * 64);
  for (int p = 0; p < 16; ++p) for (int c = 0; c < 64; ++c) h[p * 64 + c] = __float2bfloat16((float)(p * 8 + c / 8));   // value = pixel*8 + chunk
  __nv_bfloat16* d; cudaMalloc(&d, h.size() * 2); cudaMemcpy(d, h.data(), h.size() * 2, cudaMemcpyHostToDevice);
  CUtensorMap tm; const char* err = nullptr;
  if (make_tmap_nhwc64(&tm, d, 1, 1, 16, 8, 1, &err)) { printf("tmap: %s\n", err); return 1; }
  uint32_t* d_out; cudaMalloc(&d_out, 4096);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 8192);
  for (int shift = 0; shift < 8; ++shift) {
    k<<<1, 128, 8192>>>(tm, shift, d_out);
    cudaError_t e = cudaDeviceSynchronize();
    std::vector<uint16_t> o(2048);
    cudaMemcpy(o.data(), d_out, 4096, cudaMemcpyDeviceToHost);
    int abs_ok = 1, rel_ok = 1;
    for (int p = 0; p < 8; ++p)
      for (int c = 0; c < 8; ++c) {
        const int row = shift + p;                    // 128-byte row index from the 1024-aligned base
        auto at = [&](int chunk) { return __bfloat162float(*reinterpret_cast<__nv_bfloat16*>(&o[(row * 128 + chunk * 16) / 2])); };
        if (at(c ^ (row & 7)) != (float)(p * 8 + c)) abs_ok = 0;
        if (at(c ^ (p & 7)) != (float)(p * 8 + c)) rel_ok = 0;
      }
    printf("dst = base + %d*128: absolute-address swizzle %s, box-relative swizzle %s [%s]\n", shift, abs_ok ? "YES" : "no", rel_ok ? "YES" : "no", cudaGetErrorString(e));
  }
  return 0;
}
