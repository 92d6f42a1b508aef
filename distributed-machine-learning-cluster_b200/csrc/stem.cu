// Fused ResNet stem for sm_100a: u8 image -> normalise -> conv 7x7 s2 p3 (3->64, BN folded) -> ReLU
// -> max-pool 3x3 s2 p1, one kernel, bf16 out [N][56][56][64].
//
// Replaces, for the bf16 path, the first three stages the reference runs through tch
// (imagenet normalise, resnet `conv1`+`bn1`+relu, max_pool2d; reference src/services.rs:492-493,
// SURVEY.md Appendix B.1/B.2).  Unfused, those stages move 150 KB + 2 x 401 KB + 2 x 1.6 MB + 401 KB
// per image through HBM and were 43 % of the step; fused, the kernel reads the 150 KB u8 image and
// writes the 401 KB pooled map.
//
// One CTA (two co-resident per SM) produces a 7x7 pooled tile = a 15x15 patch of conv outputs
// (1 halo row/column, recomputed), held as M = 256 GEMM rows (15 rows x 16 slots):
//   P1  256 threads stage the 36x40-pixel input patch in shared memory as normalised bf16x4 pixels
//   P2  each thread builds its im2col row for one 64-wide K block (2 filter rows x 8 taps x 4 ch)
//       straight into the 128B-swizzled UMMA layout; 2-stage ring, mbarrier hand-off to the MMA warp
//   MMA warp 8: 4 K blocks x 4 x (two M=128 halves) tcgen05.mma, N = 64, fp32 accumulators in TMEM
//   P3  TMEM -> registers -> +bias, ReLU, bf16 -> shared staging tile
//   P4  3x3/s2 max over the staging tile -> coalesced 16-byte stores of the pooled tile
// Weights ([64][256] bf16, K order (r, s<8, c<4)) are loaded once per CTA by TMA.
#include "conv.h"
#include "ops.h"
#include "ptx.cuh"

namespace b2i {
namespace {

constexpr int ST_BUILDERS = 256;
constexpr int ST_THREADS = ST_BUILDERS + 32;
constexpr int PATCH_H = 36, PATCH_W = 40;          // pixels of 8 bytes
constexpr uint32_t A_STAGE_BYTES = 256 * 128;      // 256 rows x 64 bf16
constexpr uint32_t A_RING_BYTES = 2 * A_STAGE_BYTES;
constexpr uint32_t B_BYTES = 4 * 64 * 128;         // 4 K blocks x 64 out channels x 64 bf16
constexpr uint32_t PATCH_BYTES = PATCH_H * PATCH_W * 8;
constexpr uint32_t STEM_SMEM = A_RING_BYTES + B_BYTES + PATCH_BYTES + 1024;
constexpr uint32_t TMEM_COLS = 128;                // two 128x64 fp32 accumulators

struct StemParams {
  const uint8_t* in;
  __nv_bfloat16* out;
  const float* bias;
  int total_tiles;  // N * 64
};

__global__ void __launch_bounds__(ST_THREADS, 2)
stem_kernel(const __grid_constant__ CUtensorMap tmW, const StemParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bars[6];  // full[2], empty[2], acc_full, w_full
  __shared__ uint32_t tmem_slot;
  __shared__ float lut[3][256];               // (u/255 - mean)/std per channel, exact fp32 ops

  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const int lane = tid & 31;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t a_ring = base;
  const uint32_t b_smem = base + A_RING_BYTES;
  const uint32_t patch = b_smem + B_BYTES;
  uint8_t* const gen_base = smem_raw + (base - smem_u32(smem_raw));  // generic pointer to `base`
  const uint32_t bar0 = smem_u32(bars);
  auto full_bar = [&](int s) { return bar0 + 8u * s; };
  auto empty_bar = [&](int s) { return bar0 + 8u * (2 + s); };
  const uint32_t acc_full = bar0 + 8u * 4;
  const uint32_t w_full = bar0 + 8u * 5;

  if (tid == 0) {
    mbar_init(full_bar(0), ST_BUILDERS);
    mbar_init(full_bar(1), ST_BUILDERS);
    mbar_init(empty_bar(0), 1);
    mbar_init(empty_bar(1), 1);
    mbar_init(acc_full, 1);
    mbar_init(w_full, 1);
    mbar_fence_init();
  }
  for (int i = tid; i < 768; i += ST_THREADS) {
    const int c = i >> 8, u = i & 255;
    const float mean = c == 0 ? 0.485f : (c == 1 ? 0.456f : 0.406f);
    const float sd = c == 0 ? 0.229f : (c == 1 ? 0.224f : 0.225f);
    lut[c][u] = __fdiv_rn(__fsub_rn(__fdiv_rn(static_cast<float>(u), 255.0f), mean), sd);
  }
  if (warp == 8) tmem_alloc(smem_u32(&tmem_slot), TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;

  if (warp == 8) {
    // ------------------------------------------------------------------ weights + MMA issue
    if (lane == 0) {
      tma_prefetch_desc(&tmW);
      mbar_expect_tx(w_full, B_BYTES);
      for (int kb = 0; kb < 4; ++kb) tma_load_2d(b_smem + kb * 8192u, &tmW, w_full, kb * 64, 0);
      mbar_wait(w_full, 0);
    }
    __syncwarp();
    const uint32_t idesc = umma_idesc_bf16(128, 64);
    for (int t = blockIdx.x; t < p.total_tiles; t += gridDim.x) {
      for (int kb = 0; kb < 4; ++kb) {
        const int s = kb & 1;
        mbar_wait(full_bar(s), (kb >> 1) & 1);
        tc_fence_after();
        if (lane == 0) {
          const uint32_t a_s = a_ring + s * A_STAGE_BYTES;
          const uint64_t bdesc = umma_smem_desc_sw128(b_smem + kb * 8192u);
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            const uint64_t adesc = umma_smem_desc_sw128(a_s + h * 16384u);
#pragma unroll
            for (int k = 0; k < 4; ++k)
              umma_bf16(tmem_base + h * 64u, adesc + 2u * k, bdesc + 2u * k, idesc, (kb | k) != 0 ? 1u : 0u);
          }
          umma_commit(empty_bar(s));
          if (kb == 3) umma_commit(acc_full);
        }
        __syncwarp();
      }
    }
  } else {
    // ------------------------------------------------------------------ builders / epilogue
    const int m = tid;                 // GEMM row: conv pixel (ly, lx) = (m / 16, m % 16)
    const int ly = (m >> 4) < 15 ? (m >> 4) : 14;  // slots 240..255 are padding rows: read row 14 again
    const int lx = m & 15;
    const uint32_t swz_row = static_cast<uint32_t>(m & 7);
    const int h = warp >> 2, q = warp & 3;
    int i = 0;
    for (int t = blockIdx.x; t < p.total_tiles; t += gridDim.x, ++i) {
      const int n = t >> 6;
      const int pty = (t >> 3) & 7, ptx = t & 7;
      const int oh0 = 14 * pty - 1, ow0 = 14 * ptx - 1;   // first conv row / col of the patch
      const int ih0 = 2 * oh0 - 3, iw0 = 2 * ow0 - 3;     // first input row / col
      // P1: input patch -> normalised bf16x4 pixels (zero outside the image = conv zero padding)
      const uint8_t* img = p.in + static_cast<size_t>(n) * (224 * 224 * 3);
      for (int e = tid; e < PATCH_H * PATCH_W; e += ST_BUILDERS) {
        const int py = e / PATCH_W, px = e - py * PATCH_W;
        const int ih = ih0 + py, iw = iw0 + px;
        uint2 v = make_uint2(0u, 0u);
        if (ih >= 0 && ih < 224 && iw >= 0 && iw < 224) {
          const uint8_t* s3 = img + (ih * 224 + iw) * 3;
          v.x = pack_bf16x2(lut[0][__ldg(s3)], lut[1][__ldg(s3 + 1)]);
          v.y = pack_bf16x2(lut[2][__ldg(s3 + 2)], 0.f);
        }
        *reinterpret_cast<uint2*>(gen_base + A_RING_BYTES + B_BYTES + e * 8) = v;
      }
      asm volatile("bar.sync 1, %0;" ::"n"(ST_BUILDERS) : "memory");
      // P2: im2col rows, one 64-wide K block (filter rows 2kb, 2kb+1) per ring stage
      for (int kb = 0; kb < 4; ++kb) {
        const int s = kb & 1;
        mbar_wait(empty_bar(s), ((kb >> 1) & 1) ^ 1);
        const uint32_t dst_row = a_ring + s * A_STAGE_BYTES + static_cast<uint32_t>(m) * 128u;
        const int nchunks = kb == 3 ? 4 : 8;  // filter row 7 does not exist: its weights are zero
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          if (j < nchunks) {
            const int r = 2 * kb + (j >> 2);
            const uint32_t src = patch + static_cast<uint32_t>(((2 * ly + r) * PATCH_W + 2 * lx + 2 * (j & 3)) * 8);
            uint32_t x0, x1, x2, x3;
            asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(x0), "=r"(x1), "=r"(x2), "=r"(x3) : "r"(src));
            asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(dst_row + ((static_cast<uint32_t>(j) ^ swz_row) << 4)),
                         "r"(x0), "r"(x1), "r"(x2), "r"(x3) : "memory");
          }
        }
        fence_proxy_async_smem();
        mbar_arrive(full_bar(s));
      }
      // P3: accumulators -> +bias, ReLU, bf16 -> staging tile (aliases the A ring; all MMAs are done)
      mbar_wait(acc_full, i & 1);
      tc_fence_after();
      {
        const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(h * 64);
        const int row = h * 128 + q * 32 + lane;
        const uint32_t st_row = a_ring + static_cast<uint32_t>(row) * 128u;
        const uint32_t sw = static_cast<uint32_t>(row & 7);
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          uint32_t v[32];
          tmem_ld32(t_row + static_cast<uint32_t>(c * 32), v);
          tmem_ld_wait();
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            float f[8];
#pragma unroll
            for (int e = 0; e < 8; ++e)
              f[e] = fmaxf(__uint_as_float(v[g * 8 + e]) + __ldg(p.bias + c * 32 + g * 8 + e), 0.f);
            const uint32_t chunk = static_cast<uint32_t>(c * 4 + g);
            asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(st_row + ((chunk ^ sw) << 4)),
                         "r"(pack_bf16x2(f[0], f[1])), "r"(pack_bf16x2(f[2], f[3])),
                         "r"(pack_bf16x2(f[4], f[5])), "r"(pack_bf16x2(f[6], f[7])) : "memory");
          }
        }
      }
      tc_fence_before();
      asm volatile("bar.sync 1, %0;" ::"n"(ST_BUILDERS) : "memory");
      // P4: 3x3 stride-2 max over the staging tile -> pooled 7x7x64
      for (int item = tid; item < 49 * 8; item += ST_BUILDERS) {
        const int pp = item >> 3, c8 = item & 7;
        const int py = pp / 7, px = pp - py * 7;
        float mx[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) mx[e] = 0.f;  // post-ReLU values are >= 0 and the centre tap is always valid
#pragma unroll
        for (int dy = 0; dy < 3; ++dy) {
          const int cy = 2 * py + dy;
          const int oh = oh0 + cy;
          if (oh < 0 || oh >= 112) continue;
#pragma unroll
          for (int dx = 0; dx < 3; ++dx) {
            const int cx = 2 * px + dx;
            const int ow = ow0 + cx;
            if (ow < 0 || ow >= 112) continue;
            const int row = cy * 16 + cx;
            const uint32_t a = a_ring + static_cast<uint32_t>(row) * 128u + ((static_cast<uint32_t>(c8) ^ (row & 7)) << 4);
            uint32_t x0, x1, x2, x3;
            asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(x0), "=r"(x1), "=r"(x2), "=r"(x3) : "r"(a));
            const uint32_t xs[4] = {x0, x1, x2, x3};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              mx[2 * e] = fmaxf(mx[2 * e], bf16_lo(xs[e]));
              mx[2 * e + 1] = fmaxf(mx[2 * e + 1], bf16_hi(xs[e]));
            }
          }
        }
        const int ph = pty * 7 + py, pw = ptx * 7 + px;
        uint4 o;
        o.x = pack_bf16x2(mx[0], mx[1]);
        o.y = pack_bf16x2(mx[2], mx[3]);
        o.z = pack_bf16x2(mx[4], mx[5]);
        o.w = pack_bf16x2(mx[6], mx[7]);
        *reinterpret_cast<uint4*>(p.out + ((static_cast<size_t>(n) * 56 + ph) * 56 + pw) * 64 + c8 * 8) = o;
      }
      // the staging tile (= ring stage 0/1) and the patch are rewritten by the next tile
      asm volatile("bar.sync 1, %0;" ::"n"(ST_BUILDERS) : "memory");
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 8) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

}  // namespace

cudaError_t launch_stem(const uint8_t* in_u8, void* out_bf16, const CUtensorMap* tmW,
                        const float* bias, int n_images, cudaStream_t st) {
  static bool attr_set[16] = {};
  static int sm_count[16] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 16) return cudaErrorInvalidDevice;
  if (!attr_set[dev]) {
    cudaError_t e = cudaFuncSetAttribute(stem_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, STEM_SMEM);
    if (e != cudaSuccess) return e;
    e = cudaDeviceGetAttribute(&sm_count[dev], cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess) return e;
    attr_set[dev] = true;
  }
  if (n_images < 1 || !tmW) return cudaErrorInvalidValue;
  StemParams p;
  p.in = in_u8;
  p.out = static_cast<__nv_bfloat16*>(out_bf16);
  p.bias = bias;
  p.total_tiles = n_images * 64;
  const int grid = p.total_tiles < 2 * sm_count[dev] ? p.total_tiles : 2 * sm_count[dev];
  stem_kernel<<<grid, ST_THREADS, STEM_SMEM, st>>>(*tmW, p);
  return cudaGetLastError();
}

}  // namespace b2i
