// Fused AlexNet stem for sm_100a: u8 image -> normalise -> conv 11x11 s4 p2 (3->64, +bias) -> ReLU, one
// kernel, bf16 out [N][55][55][64].
//
// Replaces, for the bf16 path, the first two stages the reference runs through tch (imagenet normalise,
// alexnet `features.0` + ReLU; reference src/services.rs:492-493).  The general
// small-channel kernel gathers this layer's im2col rows with 8-byte cp.async copies, 16 per thread and
// k-block, and spends 0.25 ms on it at batch 256 - 42 % of the AlexNet step.
//
// Same idea as the ResNet stem (stem.cu): the A-operand descriptors of the MMAs point STRAIGHT INTO the input
// patch, no im2col matrix exists.  The canonical no-swizzle K-major core matrix is 8 rows x 16 bytes with the
// rows 16 bytes apart; with stride 4 the windows of neighbouring output pixels start 4 input pixels apart, so
// a pixel has to be 4 bytes = two bf16 channels.  The patch is therefore kept as TWO planes, (R, G) and
// (B, 0).  One 16-byte K chunk is four taps of one plane, a K = 16 MMA step eight taps (LBO = 16 bytes), the
// next output pixel 16 bytes further, the next output row four patch rows further (SBO).  A filter row is
// 11 taps, padded to 16 (two steps) per plane: K = 11 rows x 2 planes x 32 = 704, 52 % of it real.
//
// One CTA per SM produces tiles of 16 (y) x 8 (x) output pixels = one M = 128 MMA tile, 44 MMAs (N = 64) each.
//   loaders (8 warps)  stage the 71 x 44-pixel input patch of tile i+1 as two planes of normalised bf16x2
//                      pixels, double buffered; aligned 32-bit loads, one FMA per byte (see stem.cu)
//   MMA (1 warp)       issues the 44 tcgen05.mma of a tile; weights [64][704] resident (11 k-blocks, TMA)
//   drain (8 warps)    TMEM -> registers -> +bias (kernel parameters), ReLU, bf16 -> 32-byte global stores
#include "conv.h"
#include "ops.h"
#include "ptx.cuh"

namespace b2i {
namespace {

constexpr int AS_DRAIN = 256;                       // warps 0-7
constexpr int AS_LOADERS = 256;                     // warps 9-16
constexpr int AS_THREADS = AS_DRAIN + 32 + AS_LOADERS;
constexpr int AP_H = 71, AP_W = 44;                 // patch rows / pixels per row (4-byte pixels)
constexpr uint32_t AP_ROW_BYTES = AP_W * 4;         // 176
constexpr uint32_t AP_PLANE_BYTES = ((AP_H * AP_ROW_BYTES + 127) / 128) * 128;   // 12544
constexpr uint32_t AP_BYTES = 2 * AP_PLANE_BYTES;   // both planes of one patch
constexpr uint32_t AW_BYTES = 11 * 64 * 128;        // 11 k-blocks x 64 out channels x 64 bf16, 128B swizzle
constexpr uint32_t AS_SMEM = AW_BYTES + 2 * AP_BYTES + 1024;
constexpr uint32_t AS_TMEM_COLS = 128;              // two tiles in flight x one 128x64 fp32 accumulator
constexpr int OUT_HW = 55;
constexpr int TILES_X = 7, TILES_Y = 4;             // 8-wide x 16-tall tiles over 55 x 55

struct AlexStemParams {
  const uint8_t* in;
  __nv_bfloat16* out;
  float bias[64];
  int total_tiles;  // N * 28
  int f16;          // activations / weights are IEEE half instead of bf16
};

__device__ __forceinline__ uint64_t plane_desc(uint32_t saddr) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((saddr & 0x3FFFFu) >> 4);
  d |= static_cast<uint64_t>(16 >> 4) << 16;                     // leading (K) byte offset: the next four taps
  d |= static_cast<uint64_t>((4 * AP_ROW_BYTES) >> 4) << 32;     // stride (M group) byte offset: the next output row
  d |= static_cast<uint64_t>(1) << 46;
  return d;                                                       // layout type 0: no swizzle
}

template <bool F16>
__global__ void __launch_bounds__(AS_THREADS, 1)
alex_stem_kernel(const __grid_constant__ CUtensorMap tmW, const AlexStemParams p) {
  griddep_launch();
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bars[9];  // patch_full[2], patch_empty[2], acc_full[2], acc_empty[2], w_full
  __shared__ uint32_t tmem_slot;

  const int tid = threadIdx.x;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);   // warp-uniform, and the compiler knows it (MMA operands stay in uniform registers)
  const int lane = tid & 31;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t w_smem = base;
  const uint32_t patch0 = base + AW_BYTES;
  const uint32_t bar0 = smem_u32(bars);
  auto patch_full = [&](int b) { return bar0 + 8u * b; };
  auto patch_empty = [&](int b) { return bar0 + 8u * (2 + b); };
  auto acc_full = [&](int b) { return bar0 + 8u * (4 + b); };
  auto acc_empty = [&](int b) { return bar0 + 8u * (6 + b); };
  const uint32_t w_full = bar0 + 8u * 8;

  if (tid == 0) {
    for (int b = 0; b < 2; ++b) {
      mbar_init(patch_full(b), AS_LOADERS);
      mbar_init(patch_empty(b), 1);
      mbar_init(acc_full(b), 1);
      mbar_init(acc_empty(b), 8);
    }
    mbar_init(w_full, 1);
    mbar_fence_init();
  }
  if (warp == 8) tmem_alloc(smem_u32(&tmem_slot), AS_TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, tmem_slot, 0);
  if (warp == 8 && lane == 0) {   // the weights do not depend on the previous kernel
    tma_prefetch_desc(&tmW);
    mbar_expect_tx(w_full, AW_BYTES);
    for (int kb = 0; kb < 11; ++kb) tma_load_2d(w_smem + kb * 8192u, &tmW, w_full, kb * 64, 0);
  }
  griddep_wait();

  if (warp == 8) {
    // ------------------------------------------------------------------ MMA issue
    mbar_wait(w_full, 0);
    const uint32_t idesc = umma_idesc_bf16(128, 64, F16 ? 1 : 0);
    int i = 0;
    for (int t = blockIdx.x; t < p.total_tiles; t += gridDim.x, ++i) {
      const int b = i & 1;
      mbar_wait(acc_empty(b), ((i >> 1) & 1) ^ 1);
      mbar_wait(patch_full(b), (i >> 1) & 1);
      tc_fence_after();
      const uint32_t pbase = patch0 + b * AP_BYTES;
      const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(b * 64);
#pragma unroll 1
      for (int r = 0; r < 11; ++r) {           // filter row = weight k-block r
        const uint64_t bdesc = umma_smem_desc_sw128(w_smem + r * 8192u);
#pragma unroll
        for (int j = 0; j < 4; ++j) {           // plane j/2, taps 8*(j&1) .. +7
          const uint32_t a_addr = pbase + static_cast<uint32_t>((j >> 1) * AP_PLANE_BYTES + r * AP_ROW_BYTES + (j & 1) * 32);
          umma_bf16_elect(d_tmem, plane_desc(a_addr), bdesc + 2u * j, idesc, (r | j) != 0 ? 1u : 0u);
        }
      }
      umma_commit_elect(patch_empty(b));
      umma_commit_elect(acc_full(b));
    }
  } else if (warp > 8) {
    // ------------------------------------------------------------------ patch loaders
    const int lt = tid - (AS_DRAIN + 32);
    uint8_t* const gen_base = smem_raw + (base - smem_u32(smem_raw));
    // One task = four neighbouring pixels of one patch row = 12 image bytes.  The patch starts at image column
    // 32*tx - 2, i.e. at byte 96*tx - 6 of the row: 2 past a word boundary (rows are 672 bytes), so the four
    // pixels lie in four aligned words starting at byte 96*tx - 8 + 12*g.
    constexpr int GROUPS = AP_W / 4;                                          // 11 per patch row
    constexpr int TASKS = AP_H * GROUPS;                                      // 781
    constexpr int ROUNDS = (TASKS + AS_LOADERS - 1) / AS_LOADERS;             // 4
    const float na[3] = {0x1.18926cp-6f, 0x1.1ed5b2p-6f, 0x1.1d8f56p-6f};     // fp32(1 / (255 * std))
    const float nb[3] = {-0x1.0f177ap+1f, -0x1.04924ap+1f, -0x1.cdf012p+0f};  // fp32(-mean / std)
    int i = 0;
    for (int t = blockIdx.x; t < p.total_tiles; t += gridDim.x, ++i) {
      const int b = i & 1;
      const int n = t / (TILES_X * TILES_Y);
      const int rem = t - n * (TILES_X * TILES_Y);
      const int ty = rem / TILES_X, tx = rem - ty * TILES_X;
      const int ih0 = 64 * ty - 2, iw0 = 32 * tx - 2;       // input row / col of patch pixel (0, 0)
      const uint8_t* img = p.in + static_cast<size_t>(n) * (224 * 224 * 3);
      uint8_t* dst = gen_base + AW_BYTES + b * AP_BYTES;
#pragma unroll 1
      for (int k0 = 0; k0 < ROUNDS; k0 += 4) {               // four tasks' loads in flight per thread
        uint32_t w[4][4];
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
          const int e = lt + (k0 + kk) * AS_LOADERS;
          const int py = e / GROUPS, g = e - py * GROUPS;
          const int ih = ih0 + py;
          const int wo = 96 * tx - 8 + 12 * g;
          const bool row_ok = (k0 + kk) < ROUNDS && e < TASKS && ih >= 0 && ih < 224;
          const uint32_t* src = reinterpret_cast<const uint32_t*>(img + ih * 672 + wo);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int o = wo + 4 * j;
            w[kk][j] = (row_ok && o >= 0 && o < 672) ? __ldg(src + j) : 0u;
          }
        }
        if (k0 == 0) mbar_wait(patch_empty(b), ((i >> 1) & 1) ^ 1);   // the loads above are already in flight
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
          const int e = lt + (k0 + kk) * AS_LOADERS;
          if ((k0 + kk) >= ROUNDS || e >= TASKS) continue;
          const int py = e / GROUPS, g = e - py * GROUPS;
          const int ih = ih0 + py;
          const bool row_ok = ih >= 0 && ih < 224;
          auto byte_f = [&](int q) {   // byte q of the 16 loaded ones as a float: 0x4B0000uu is 2^23 + u
            return __uint_as_float(__byte_perm(w[kk][q >> 2], 0x4B000000u, 0x7440u | (q & 3))) - 8388608.0f;
          };
          uint32_t rg[4], bz[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) {                      // pixel j = bytes 2+3j .. 4+3j
            const int iw = iw0 + 4 * g + j;
            const bool ok = row_ok && iw >= 0 && iw < 224;   // outside the image: the conv's zero padding
            const float r = fmaf(byte_f(2 + 3 * j), na[0], nb[0]);
            const float gg = fmaf(byte_f(3 + 3 * j), na[1], nb[1]);
            const float bl = fmaf(byte_f(4 + 3 * j), na[2], nb[2]);
            rg[j] = ok ? pack16x2_t<F16>(r, gg) : 0u;
            bz[j] = ok ? pack16x2_t<F16>(bl, 0.f) : 0u;
          }
          const uint32_t off = static_cast<uint32_t>(py) * AP_ROW_BYTES + static_cast<uint32_t>(g) * 16u;
          *reinterpret_cast<uint4*>(dst + off) = make_uint4(rg[0], rg[1], rg[2], rg[3]);
          *reinterpret_cast<uint4*>(dst + AP_PLANE_BYTES + off) = make_uint4(bz[0], bz[1], bz[2], bz[3]);
        }
      }
      fence_proxy_async_smem();
      mbar_arrive(patch_full(b));
    }
  } else {
    // ------------------------------------------------------------------ drain (warps 0-7)
    const int q = warp & 3, half = warp >> 2;   // TMEM lane quarter, 32-column half of the 64 output channels
    int i = 0;
    for (int t = blockIdx.x; t < p.total_tiles; t += gridDim.x, ++i) {
      const int b = i & 1;
      const int n = t / (TILES_X * TILES_Y);
      const int rem = t - n * (TILES_X * TILES_Y);
      const int ty = rem / TILES_X, tx = rem - ty * TILES_X;
      const int mrow = q * 32 + lane;
      const int oy = 16 * ty + (mrow >> 3), ox = 8 * tx + (mrow & 7);
      mbar_wait(acc_full(b), (i >> 1) & 1);
      tc_fence_after();
      uint32_t v[32];
      tmem_ld32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(b * 64 + half * 32), v);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(acc_empty(b));
      if (oy < OUT_HW && ox < OUT_HW) {
        __nv_bfloat16* o = p.out + ((static_cast<size_t>(n) * OUT_HW + oy) * OUT_HW + ox) * 64 + half * 32;
#pragma unroll
        for (int h2 = 0; h2 < 2; ++h2) {
          uint32_t ov[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            const int c = h2 * 16 + 2 * e;
            const float f0 = fmaxf(__uint_as_float(v[c]) + p.bias[half * 32 + c], 0.f);
            const float f1 = fmaxf(__uint_as_float(v[c + 1]) + p.bias[half * 32 + c + 1], 0.f);
            ov[e] = pack16x2_t<F16>(f0, f1);
          }
          stg256(o + h2 * 16, ov);
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 8) {
    tc_fence_after();
    tmem_dealloc(tmem_base, AS_TMEM_COLS);
  }
}

}  // namespace

cudaError_t launch_alex_stem(const uint8_t* in_u8, void* out_bf16, const CUtensorMap* tmW,
                             const float* bias_host, int n_images, int f16, cudaStream_t st) {
  static bool attr_set[16] = {};
  static int sm_count[16] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 16) return cudaErrorInvalidDevice;
  if (!attr_set[dev]) {
    cudaError_t e = cudaFuncSetAttribute(alex_stem_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, AS_SMEM);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(alex_stem_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, AS_SMEM);
    if (e != cudaSuccess) return e;
    e = cudaDeviceGetAttribute(&sm_count[dev], cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess) return e;
    attr_set[dev] = true;
  }
  if (n_images < 1 || !tmW || !bias_host) return cudaErrorInvalidValue;
  if (reinterpret_cast<uintptr_t>(in_u8) & 3) return cudaErrorMisalignedAddress;   // the loaders fetch aligned words
  AlexStemParams p;
  p.in = in_u8;
  p.out = static_cast<__nv_bfloat16*>(out_bf16);
  for (int c = 0; c < 64; ++c) p.bias[c] = bias_host[c];
  p.total_tiles = n_images * TILES_X * TILES_Y;
  p.f16 = f16;
  const int grid = p.total_tiles < sm_count[dev] ? p.total_tiles : sm_count[dev];
  if (f16) return launch_pdl(alex_stem_kernel<true>, dim3(grid), dim3(AS_THREADS), AS_SMEM, st, *tmW, p);
  return launch_pdl(alex_stem_kernel<false>, dim3(grid), dim3(AS_THREADS), AS_SMEM, st, *tmW, p);
}

}  // namespace b2i
