// How fast does one SM's TMA unit fill shared memory, by load mode?  148 CTAs, one issuing thread each,
// an 8-deep ring of 16 KB stages, no consumer: bytes per clock per SM for
//   im2col-mode loads (128 pixels x 64 channels, 3x3 window, stride 1 and 2),
//   tiled pixel-row loads (one output row's input pixels per load, stride 1 and 2),
//   tiled 2-D weight loads (128 rows x 128 bytes).
// The source tensors fit in L2.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I../../distributed-machine-learning-cluster_b200/csrc \
//        tma_rate.cu ../../distributed-machine-learning-cluster_b200/csrc/conv_umma.cu -o tma_rate -lcuda
#include <cstdio>
#include <cuda_runtime.h>
#include "conv.h"
#include "ptx.cuh"
using namespace b2i;

__device__ __forceinline__ void tma4d(uint32_t dst, const CUtensorMap* tm, uint32_t bar, int c0, int c1, int c2, int c3) {
  asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(dst), "l"(tm), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}

// mode 0/1: im2col (tm), 2/3: pixel rows (tm, rows_per_tile loads of row_px pixels), 4: 2-D weights
// nst_force > 0: ring depth per issuing warp (several warps per CTA, or two CTAs per SM, share the shared memory)
__global__ void __launch_bounds__(64, 1) rate_kernel(const __grid_constant__ CUtensorMap tm, int mode, int iters, int H, int W, int OH, int OW,
                                                      int stride, int cslices, int nimg, int rows_per_tile, int px, long long* cycles,
                                                      int nst_force) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bars_all[16];
  const int wid = threadIdx.x >> 5;
  uint64_t* bars = bars_all + 8 * wid;
  const int lane = threadIdx.x & 31;
  if (lane == 0) { for (int s = 0; s < 8; ++s) mbar_init(smem_u32(&bars[s]), 1); mbar_fence_init(); }
  __syncthreads();
  const uint32_t stage_bytes = px > 128 ? px * 128 + 4096 : 20 * 1024;
  const int nst = nst_force > 0 ? nst_force : (px > 128 ? (px > 256 ? 2 : 4) : 8);
  const uint32_t base = ((smem_u32(smem_raw) + 1023u) & ~1023u) + wid * nst * stage_bytes;
  long long t0 = clock64();
  uint32_t bytes_per = 0;
  for (int it = 0; it < iters; ++it) {
    const int s = it % nst;
    if (it >= nst) mbar_wait(smem_u32(&bars[s]), ((it / nst) - 1) & 1);
    const uint32_t dst = base + s * stage_bytes + 2048;
    const uint32_t bar = smem_u32(&bars[s]);
    const int tile = ((blockIdx.x * 2 + wid) * 131 + it / 9) % ((nimg * OH * OW) / 128);
    const int tap = it % 9, r = tap / 3, sx = tap % 3, cb = (it / 9) % cslices;
    const int m0 = tile * 128;
    if (mode <= 1) {
      bytes_per = px * 128;
      if (lane == 0) {
        mbar_expect_tx(bar, bytes_per);
        const int n = m0 / (OH * OW), rem = m0 % (OH * OW), oh = rem / OW, ow = rem % OW;
        tma_load_im2col_4d(dst, &tm, bar, cb * 64, ow * stride - 1, oh * stride - 1, n, (uint16_t)sx, (uint16_t)r);
      }
    } else if (mode <= 3) {
      bytes_per = rows_per_tile * OW * 128;
      if (lane == 0) mbar_expect_tx(bar, bytes_per);
      if (lane < rows_per_tile) {
        const int R = m0 / OW + lane;
        const int n = R / OH, oh = R % OH;
        tma4d(dst + (R * OW - m0) * 128, &tm, bar, cb * 64, sx - 1, oh * stride - 1 + r, n);
      }
    } else {
      bytes_per = 128 * 128;
      if (lane == 0) { mbar_expect_tx(bar, bytes_per); tma_load_2d(dst, &tm, bar, ((it * 7) % 36) * 64, (blockIdx.x & 1) * 128); }
    }
    __syncwarp();
  }
  for (int s = 0; s < nst; ++s) {               // the last fill of stage s was the largest it < iters with it % nst == s
    const int it_last = iters - 1 - ((iters - 1 - s) % nst + nst) % nst;
    mbar_wait(smem_u32(&bars[s]), (it_last / nst) & 1);
  }
  long long t1 = clock64();
  if (blockIdx.x == 0 && threadIdx.x == 0) { cycles[0] = t1 - t0; cycles[1] = (long long)bytes_per * iters; }
}

int main() {
  const int nimg = 64;
  long long* d_c; cudaMalloc(&d_c, 16);
  cudaFuncSetAttribute(rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  struct Cfg { const char* name; int mode, H, W, C, stride, px; } cfgs[] = {
      {"im2col  s1  14x14x256, 256 pixels per load", 0, 14, 14, 256, 1, 256},
      {"im2col  s1  14x14x256, 512 pixels per load", 0, 14, 14, 256, 1, 512},
      {"im2col  s1  14x14x256, 64 pixels per load ", 0, 14, 14, 256, 1, 64},
      {"im2col  s1  14x14x256, 32 pixels per load ", 0, 14, 14, 256, 1, 32},
      {"im2col  s1  14x14x256 (layer3 convs)      ", 0, 14, 14, 256, 1},
      {"im2col  s2  28x28x128 -> 14x14 (layer3.0) ", 1, 28, 28, 128, 2},
      {"im2col  s2  56x56x64  -> 28x28 (layer2.0) ", 1, 56, 56, 64, 2},
      {"pixrows s1  14x14x256                     ", 2, 14, 14, 256, 1},
      {"pixrows s2  28x28x128 -> 14x14            ", 3, 28, 28, 128, 2},
      {"pixrows s2  56x56x64  -> 28x28            ", 3, 56, 56, 64, 2},
      {"pixrows s1  7x7x512   (layer4 convs)      ", 2, 7, 7, 512, 1},
      {"pixrows s1  28x28x128 (layer2 convs)      ", 2, 28, 28, 128, 1},
      {"weights 2-D 128 rows x 128 B of [256][2304]", 4, 0, 0, 0, 1},
  };
  for (auto& c : cfgs) {
    void* d_x = nullptr;
    CUtensorMap tm; const char* err = "";
    int OH = 0, OW = 0, rows = 0;
    if (c.mode == 4) {
      cudaMalloc(&d_x, 256 * 2304 * 2); cudaMemset(d_x, 0, 256 * 2304 * 2);
      if (make_tmap_weights(&tm, d_x, 256, 2304, 128, &err)) { printf("tmap %s\n", err); return 1; }
    } else {
      const size_t bytes = (size_t)nimg * c.H * c.W * c.C * 2;
      cudaMalloc(&d_x, bytes); cudaMemset(d_x, 0, bytes);
      OH = c.H / c.stride; OW = c.W / c.stride;
      rows = (127 + OW - 1) / OW + 1;
      if (c.mode <= 1) {
        ConvGeom g{}; g.N = nimg; g.H = c.H; g.W = c.W; g.C = c.C; g.OH = OH; g.OW = OW; g.R = g.S = g.Sp = 3; g.stride = c.stride; g.pad = 1;
        g.Cout = g.Cout_pad = 256; g.K = 9 * c.C; g.M = nimg * OH * OW;
        if (make_tmap_im2col_px(&tm, d_x, g, nimg, c.px ? c.px : 128, &err)) { printf("tmap %s\n", err); return 1; }
      } else if (make_tmap_pixrow(&tm, d_x, nimg, c.H, c.W, c.C, OW, c.stride, &err)) { printf("tmap %s\n", err); return 1; }
    }
    const int iters = 4000;
    for (int rep = 0; rep < 2; ++rep) {
      rate_kernel<<<148, 32, 200 * 1024>>>(tm, c.mode, iters, c.H, c.W, OH, OW, c.stride, c.C / 64, nimg, rows, c.px ? c.px : 128, d_c, 0);
      cudaError_t e = cudaDeviceSynchronize();
      long long h[2]; cudaMemcpy(h, d_c, 16, cudaMemcpyDeviceToHost);
      if (rep) printf("%s: %6.1f B/clk/SM (%lld cycles per load of %lld B) [%s]\n", c.name, (double)h[1] / h[0], h[0] / iters, h[1] / iters, cudaGetErrorString(e));
    }
    // Do im2col requests from DIFFERENT issuers overlap?  (a) two issuing warps in one CTA, each with its own ring;
    // (b) two CTAs per SM.  Rates are per issuer: if requests serialise in the SM's TMA unit, cycles per load doubles.
    if (c.mode <= 1 && (c.px == 0 || c.px == 128 || c.px == 256)) {
      const int px = c.px ? c.px : 128;
      for (int variant = 0; variant < 2; ++variant) {
        for (int rep = 0; rep < 2; ++rep) {
          if (variant == 0) rate_kernel<<<148, 64, 200 * 1024>>>(tm, c.mode, iters, c.H, c.W, OH, OW, c.stride, c.C / 64, nimg, rows, px, d_c, px > 128 ? 2 : 4);
          else rate_kernel<<<296, 32, 100 * 1024>>>(tm, c.mode, iters, c.H, c.W, OH, OW, c.stride, c.C / 64, nimg, rows, px, d_c, px > 128 ? 2 : 4);
          cudaError_t e = cudaDeviceSynchronize();
          long long h[2]; cudaMemcpy(h, d_c, 16, cudaMemcpyDeviceToHost);
          if (rep) printf("    %s: %6.1f B/clk per issuer, %lld cycles per load per issuer -> %.1f B/clk/SM [%s]\n",
                          variant == 0 ? "two issuing warps in one CTA" : "two CTAs per SM            ", (double)h[1] / h[0], h[0] / iters,
                          2.0 * h[1] / h[0], cudaGetErrorString(e));
        }
      }
    }
    cudaFree(d_x);
  }
  return 0;
}
