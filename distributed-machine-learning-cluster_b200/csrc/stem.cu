// Fused ResNet stem for sm_100a: u8 image -> normalise -> conv 7x7 s2 p3 (3->64, BN folded) -> ReLU
// -> max-pool 3x3 s2 p1, one kernel, bf16 out [N][56][56][64].
//
// Replaces, for the bf16 path, the first three stages the reference runs through tch
// (imagenet normalise, resnet `conv1`+`bn1`+relu, max_pool2d; reference src/services.rs:492-493,
// SURVEY.md Appendix B.1/B.2).  Unfused, those stages move 150 KB + 2 x 401 KB + 2 x 1.6 MB + 401 KB
// per image through HBM and were 43 % of the step; fused, the kernel reads the 150 KB u8 image and
// writes the 401 KB pooled map.
//
// One CTA (two co-resident per SM) produces 7x7 pooled tiles = 15x15 patches of conv outputs
// (1 halo row/column, recomputed), each computed as two M=128 MMAs (16 conv rows x 8 conv columns).
// Three groups of warps run concurrently, connected by mbarriers, each on its own tile:
//   loaders (4 warps)  stage the 37x40-pixel input patch of tile i+1 in shared memory as normalised
//                      bf16x4 pixels (8 bytes), double buffered
//   MMA (1 warp)       issues 2 x 14 tcgen05.mma (K = 16 = four taps each) whose A descriptors point
//                      STRAIGHT INTO THE PATCH: with 8-byte pixels and stride 2, the tap window of the
//                      next conv pixel starts 16 bytes further, which is exactly the canonical
//                      no-swizzle K-major core matrix (8 rows x 16 B); the next conv row is 640 B
//                      further (SBO), the next two taps 16 B (LBO), the next filter row 320 B (new
//                      descriptor).  No im2col matrix is ever materialised.  TMEM double buffered.
//   drain (8 warps)    TMEM -> registers -> +bias, ReLU, bf16 -> shared staging tile, then the
//                      3x3/s2 max over the staging tile -> 16-byte stores of the pooled tile
// Weights ([64][256] bf16, K order (r, s<8, c<4), rows r = 7 / taps s = 7 / channel 3 are zero) are
// loaded once per CTA by TMA into the 128B-swizzled layout.
#include "conv.h"
#include "ops.h"
#include "ptx.cuh"

namespace b2i {
namespace {

constexpr int ST_DRAIN = 256;                      // warps 0-7
constexpr int ST_LOADERS = 128;                    // warps 9-12
constexpr int ST_THREADS = ST_DRAIN + 32 + ST_LOADERS;
constexpr int PATCH_H = 37, PATCH_W = 40;          // pixels of 8 bytes (bf16 r, g, b, 0)
constexpr uint32_t PATCH_ROW_BYTES = PATCH_W * 8;  // 320
constexpr uint32_t PATCH_BYTES = ((PATCH_H * PATCH_W * 8 + 127) / 128) * 128;
constexpr uint32_t B_BYTES = 4 * 64 * 128;         // 4 K blocks x 64 out channels x 64 bf16, 128B swizzle
constexpr uint32_t STAGE_BYTES = 256 * 128;        // conv outputs of one tile, bf16 [15x16 slots][64]
constexpr uint32_t STEM_SMEM = B_BYTES + STAGE_BYTES + 2 * PATCH_BYTES + 1024;
constexpr uint32_t TMEM_COLS = 256;                // 2 tiles in flight x two 128x64 fp32 accumulators

#ifdef B2I_STEM_TRACE
// tools/microbench/stem_trace.cu: clock64 stamps of CTA 0's three roles for its first 64 tiles
__device__ long long g_stem_trace[64][12];
__device__ long long g_stem_cta[512][2];   // globaltimer at entry / exit of every CTA
__device__ __forceinline__ long long stem_now_ns() { unsigned long long ns; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ns)); return (long long)ns; }
#define STEM_STAMP(tile_i, slot, cond) do { if (blockIdx.x == 0 && (tile_i) < 64 && (cond)) { g_stem_trace[tile_i][slot] = clock64(); \
    if ((slot) == 10) { unsigned long long ns; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ns)); g_stem_trace[tile_i][11] = (long long)ns; } } } while (0)
#else
#define STEM_STAMP(tile_i, slot, cond) do { } while (0)
#endif

struct StemParams {
  const uint8_t* in;
  __nv_bfloat16* out;
  float bias[64];      // folded BN bias, by value: the drain warps read it as constant-bank operands
  int total_tiles;  // N * 64
  int f16;          // activations / weights are IEEE half instead of bf16
  unsigned int* sched;  // [0] tiles handed out beyond the first gridDim.x, [1] CTAs done; both zero between launches
};

// A-operand descriptor over the RAW patch (no swizzle, K-major): row i of an 8-row core matrix is
// the next conv pixel along x = the next 16 bytes; 8-row groups (next conv row) are 2 patch rows
// = 640 bytes apart; the next 16-byte K chunk (two more taps) is 16 bytes further.
__device__ __forceinline__ uint64_t patch_desc(uint32_t saddr) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((saddr & 0x3FFFFu) >> 4);
  d |= static_cast<uint64_t>(16 >> 4) << 16;                      // leading (K) byte offset
  d |= static_cast<uint64_t>((2 * PATCH_ROW_BYTES) >> 4) << 32;   // stride (M group) byte offset
  d |= static_cast<uint64_t>(1) << 46;
  return d;                                                        // layout type 0: no swizzle
}

template <bool F16>
__global__ void __launch_bounds__(ST_THREADS, 2)
stem_kernel(const __grid_constant__ CUtensorMap tmW, const StemParams p) {
  griddep_launch();
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bars[9];  // patch_full[2], patch_empty[2], acc_full[2], acc_empty[2], w_full
  __shared__ uint32_t tmem_slot;
  __shared__ int tile_ring[8];   // tile of iteration i (or -1 = no more), written by the loaders, read by the other roles
  __shared__ int next_tile[2];

  const int tid = threadIdx.x;
#ifdef B2I_STEM_TRACE
  if (tid == 0 && blockIdx.x < 512) g_stem_cta[blockIdx.x][0] = stem_now_ns();
#endif
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);   // warp-uniform, and the compiler knows it (MMA operands stay in uniform registers)
  const int lane = tid & 31;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t b_smem = base;
  const uint32_t stage = base + B_BYTES;
  const uint32_t patch0 = stage + STAGE_BYTES;
  const uint32_t bar0 = smem_u32(bars);
  auto patch_full = [&](int b) { return bar0 + 8u * b; };
  auto patch_empty = [&](int b) { return bar0 + 8u * (2 + b); };
  auto acc_full = [&](int b) { return bar0 + 8u * (4 + b); };
  auto acc_empty = [&](int b) { return bar0 + 8u * (6 + b); };
  const uint32_t w_full = bar0 + 8u * 8;

  if (tid == 0) {
    for (int b = 0; b < 2; ++b) {
      mbar_init(patch_full(b), ST_LOADERS);
      mbar_init(patch_empty(b), 1);
      mbar_init(acc_full(b), 1);
      mbar_init(acc_empty(b), 8);
    }
    mbar_init(w_full, 1);
    mbar_fence_init();
  }
  if (warp == 8) tmem_alloc(smem_u32(&tmem_slot), TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, tmem_slot, 0);
  griddep_wait();   // the pooled map written below may still be read by the previous forward's kernels

  if (warp == 8) {
    // ------------------------------------------------------------------ weights + MMA issue
    if (lane == 0) {
      tma_prefetch_desc(&tmW);
      mbar_expect_tx(w_full, B_BYTES);
      for (int kb = 0; kb < 4; ++kb) tma_load_2d(b_smem + kb * 8192u, &tmW, w_full, kb * 64, 0);
    }
    mbar_wait(w_full, 0);
    const uint32_t idesc = umma_idesc_bf16(128, 64, F16 ? 1 : 0);
    for (int i = 0;; ++i) {
      const int b = i & 1;
      mbar_wait(acc_empty(b), ((i >> 1) & 1) ^ 1);
      mbar_wait(patch_full(b), (i >> 1) & 1);
      if (*reinterpret_cast<volatile int*>(&tile_ring[i & 7]) < 0) {   // out of tiles: pass the news on to the drain warps
        umma_commit_elect(acc_full(b));   // arrives once every earlier MMA has completed, i.e. after their own commits
        break;
      }
      tc_fence_after();
      STEM_STAMP(i, 3, lane == 0);
      const uint32_t pbase = patch0 + b * PATCH_BYTES;
#pragma unroll
      for (int h = 0; h < 2; ++h) {          // conv columns 0-7 / 8-15 of the 16x16 patch
        const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(b * 128 + h * 64);
#pragma unroll
        for (int j = 0; j < 14; ++j) {        // K step j: filter row j/2, taps 4*(j&1) .. +3
          const uint32_t a_addr = pbase + static_cast<uint32_t>(h * 128 + (j >> 1) * PATCH_ROW_BYTES + (j & 1) * 32);
          const uint64_t bdesc = umma_smem_desc_sw128(b_smem + (j >> 2) * 8192u) + 2u * (j & 3);
          umma_bf16_elect(d_tmem, patch_desc(a_addr), bdesc, idesc, j != 0 ? 1u : 0u);
        }
      }
      umma_commit_elect(patch_empty(b));   // the loaders may overwrite this patch buffer
      umma_commit_elect(acc_full(b));
      STEM_STAMP(i, 4, lane == 0);
    }
  } else if (warp > 8) {
    // ------------------------------------------------------------------ patch loaders
    const int lt = tid - (ST_DRAIN + 32);
    uint8_t* const gen_base = smem_raw + (base - smem_u32(smem_raw));  // generic pointer to `base`
    // One task = four neighbouring pixels of one patch row = 12 image bytes, fetched as four aligned
    // 32-bit words: the patch starts at image column 28*ptx - 5, i.e. at byte 84*ptx - 15 of the row,
    // which is always 1 past a word boundary (rows are 672 bytes, images 150528).
    constexpr int GROUPS = PATCH_W / 4;                                       // 10 per patch row
    constexpr int TASKS = PATCH_H * GROUPS;                                   // 370
    constexpr int ROUNDS = (TASKS + ST_LOADERS - 1) / ST_LOADERS;             // 3
    // (u/255 - mean)/std as one FMA per byte: a = 1/(255*std), b = -mean/std.  After rounding to bf16
    // all 3 x 256 values equal the exact divide-subtract-divide sequence (checked exhaustively by the CPU tests: test_stem_fma_normalise_is_exact_in_bf16).
    const float na[3] = {0x1.18926cp-6f, 0x1.1ed5b2p-6f, 0x1.1d8f56p-6f};     // fp32(1 / (255 * std))
    const float nb[3] = {-0x1.0f177ap+1f, -0x1.04924ap+1f, -0x1.cdf012p+0f};  // fp32(-mean / std)
    // Tiles are handed out dynamically (the first one is blockIdx.x, the rest come from a global
    // counter): the two CTAs of an SM do not run at the same speed, a static split leaves a 20 % tail.
    int t = blockIdx.x;
    for (int i = 0;; ++i) {
      const int b = i & 1;
      if (lt == 0) tile_ring[i & 7] = t < p.total_tiles ? t : -1;
      if (t >= p.total_tiles) {
        // not before the MMA warp has consumed this barrier's previous phase (two completed phases
        // look like none to a parity wait)
        mbar_wait(patch_empty(b), ((i >> 1) & 1) ^ 1);
        mbar_arrive(patch_full(b));
        break;
      }
      unsigned int ticket = 0;
      if (lt == 0) ticket = atomicAdd(p.sched, 1u);          // needed only at the end of this tile
      const int n = t >> 6;
      const int pty = (t >> 3) & 7, ptx = t & 7;
      const int ih0 = 28 * pty - 5, iw0 = 28 * ptx - 5;   // input row / col of patch pixel (0, 0)
      const uint8_t* img = p.in + static_cast<size_t>(n) * (224 * 224 * 3);
      uint8_t* dst = gen_base + B_BYTES + STAGE_BYTES + b * PATCH_BYTES;
      STEM_STAMP(i, 0, lt == 0);
      uint32_t w[ROUNDS][4];
#pragma unroll
      for (int k = 0; k < ROUNDS; ++k) {
        const int e = lt + k * ST_LOADERS;
        const int py = e / GROUPS, g = e - py * GROUPS;
        const int ih = ih0 + py;
        const int wo = 84 * ptx - 16 + 12 * g;             // byte offset of the first word in the image row
        const bool row_ok = e < TASKS && ih >= 0 && ih < 224;
        const uint32_t* src = reinterpret_cast<const uint32_t*>(img + ih * 672 + wo);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int o = wo + 4 * j;
          w[k][j] = (row_ok && o >= 0 && o < 672) ? __ldg(src + j) : 0u;
        }
      }
      mbar_wait(patch_empty(b), ((i >> 1) & 1) ^ 1);
      STEM_STAMP(i, 1, lt == 0);
#pragma unroll
      for (int k = 0; k < ROUNDS; ++k) {
        const int e = lt + k * ST_LOADERS;
        if (e >= TASKS) continue;
        const int py = e / GROUPS, g = e - py * GROUPS;
        const int ih = ih0 + py;
        const bool row_ok = ih >= 0 && ih < 224;
        // byte q of the 16 loaded ones as a float: 0x4B0000uu is 2^23 + u
        auto byte_f = [&](int q) {
          return __uint_as_float(__byte_perm(w[k][q >> 2], 0x4B000000u, 0x7440u | (q & 3))) - 8388608.0f;
        };
        uint32_t o[8];
#pragma unroll
        for (int j = 0; j < 4; ++j) {                      // pixel j = bytes 1+3j .. 3+3j
          const int iw = iw0 + 4 * g + j;
          const bool ok = row_ok && iw >= 0 && iw < 224;   // outside the image: the conv's zero padding
          const float r = fmaf(byte_f(1 + 3 * j), na[0], nb[0]);
          const float gg = fmaf(byte_f(2 + 3 * j), na[1], nb[1]);
          const float bl = fmaf(byte_f(3 + 3 * j), na[2], nb[2]);
          o[2 * j] = ok ? pack16x2_t<F16>(r, gg) : 0u;
          o[2 * j + 1] = ok ? pack16x2_t<F16>(bl, 0.f) : 0u;
        }
        // two 16-byte stores; odd groups of four lanes write their halves in the other order so that
        // each store instruction covers all 32 banks once per eight lanes
        uint4* d4 = reinterpret_cast<uint4*>(dst + e * 32);
        const int first = (lt >> 2) & 1;
        const uint4 lo = make_uint4(o[0], o[1], o[2], o[3]), hi = make_uint4(o[4], o[5], o[6], o[7]);
        d4[first] = first ? hi : lo;
        d4[first ^ 1] = first ? lo : hi;
      }
      fence_proxy_async_smem();                // generic-proxy stores -> visible to the tensor core
      mbar_arrive(patch_full(b));
      STEM_STAMP(i, 2, lt == 0);
      if (lt == 0) next_tile[b] = static_cast<int>(gridDim.x + ticket);
      asm volatile("bar.sync 2, %0;" ::"n"(ST_LOADERS) : "memory");
      t = *reinterpret_cast<volatile int*>(&next_tile[b]);
    }
  } else {
    // ------------------------------------------------------------------ drain + pool (warps 0-7)
    const int h = warp >> 2, q = warp & 3;
    const uint8_t* const stage_gen = smem_raw + (stage - smem_u32(smem_raw));  // generic pointer to the staging tile
    for (int i = 0;; ++i) {
      const int b = i & 1;
      mbar_wait(acc_full(b), (i >> 1) & 1);
      const int t = *reinterpret_cast<volatile int*>(&tile_ring[i & 7]);
      if (t < 0) break;
      const int n = t >> 6;
      const int pty = (t >> 3) & 7, ptx = t & 7;

      // accumulators -> +bias, ReLU, bf16 -> staging tile [ly*16 + lx][64], 16-byte chunks XOR-swizzled
      tc_fence_after();
      STEM_STAMP(i, 5, tid == 0);
      {
        const int mrow = q * 32 + lane;                    // row inside this half's 128-row MMA
        const int srow = (mrow >> 3) * 16 + h * 8 + (mrow & 7);
        const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(b * 128 + h * 64);
        const uint32_t st_row = stage + static_cast<uint32_t>(srow) * 128u;
        const uint32_t sw = static_cast<uint32_t>(srow & 7);
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          uint32_t v[32];
          tmem_ld32(t_row + static_cast<uint32_t>(c * 32), v);
          tmem_ld_wait();
          if (c == 1) {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(acc_empty(b));      // TMEM buffer free for tile i+2
            STEM_STAMP(i, 6, tid == 0);
          }
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            float f[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) f[e] = fmaxf(__uint_as_float(v[g * 8 + e]) + p.bias[c * 32 + g * 8 + e], 0.f);
            const uint32_t chunk = static_cast<uint32_t>(c * 4 + g);
            asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(st_row + ((chunk ^ sw) << 4)),
                         "r"(pack16x2_t<F16>(f[0], f[1])), "r"(pack16x2_t<F16>(f[2], f[3])),
                         "r"(pack16x2_t<F16>(f[4], f[5])), "r"(pack16x2_t<F16>(f[6], f[7])) : "memory");
          }
        }
      }
      STEM_STAMP(i, 7, tid == 0);
      asm volatile("bar.sync 1, %0;" ::"n"(ST_DRAIN) : "memory");
      STEM_STAMP(i, 8, tid == 0);
      // 3x3 stride-2 max over the staging tile -> pooled 7x7x64.  One task = one pooled pixel x one
      // 16-byte chunk; eight neighbouring lanes read one whole 128-byte staging row per tap, which
      // is bank-conflict free whatever the row's XOR swizzle.
      const int cy_min = pty == 0 ? 1 : 0, cx_min = ptx == 0 ? 1 : 0;   // patch slot 0 is conv row / column -1 there
      for (int task = tid; task < 49 * 8; task += ST_DRAIN) {
        const int pp = task >> 3;
        const uint32_t ch = static_cast<uint32_t>(task & 7);
        const int py = pp / 7, px = pp - py * 7;
        // Window taps that fall outside the image (only possible at its top / left edge) are clamped
        // onto the first valid row / column: a duplicate tap leaves the max unchanged, and the nine
        // loads carry no branches, so they are all in flight together.
        uint4 tap[9];
#pragma unroll
        for (int dy = 0; dy < 3; ++dy) {
          const int cy = max(2 * py + dy, cy_min);
#pragma unroll
          for (int dx = 0; dx < 3; ++dx) {
            const int cx = max(2 * px + dx, cx_min);
            const uint32_t row = static_cast<uint32_t>(cy * 16 + cx);
            tap[dy * 3 + dx] = *reinterpret_cast<const uint4*>(stage_gen + row * 128u + ((ch ^ (row & 7u)) << 4));
          }
        }
        // packed 16-bit max in the model's storage format
        uint32_t mx[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          mx[e] = reinterpret_cast<const uint32_t*>(&tap[0])[e];
#pragma unroll
          for (int k = 1; k < 9; ++k) mx[e] = max16x2_t<F16>(mx[e], reinterpret_cast<const uint32_t*>(&tap[k])[e]);
        }
        const int ph = pty * 7 + py, pw = ptx * 7 + px;
        const uint32_t* mw = mx;
        *reinterpret_cast<uint4*>(p.out + ((static_cast<size_t>(n) * 56 + ph) * 56 + pw) * 64 + ch * 8) =
            make_uint4(mw[0], mw[1], mw[2], mw[3]);
      }
      STEM_STAMP(i, 9, tid == 0);
      asm volatile("bar.sync 1, %0;" ::"n"(ST_DRAIN) : "memory");  // staging tile is rewritten next iteration
      STEM_STAMP(i, 10, tid == 0);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (tid == 0) {   // the last CTA out re-arms the tile counter for the next launch
    __threadfence();
    if (atomicAdd(p.sched + 1, 1u) == gridDim.x - 1) {
      p.sched[0] = 0;
      p.sched[1] = 0;
      __threadfence();
    }
  }
  if (warp == 8) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
#ifdef B2I_STEM_TRACE
    if (lane == 0 && blockIdx.x < 512) g_stem_cta[blockIdx.x][1] = stem_now_ns();
#endif
  }
}

}  // namespace

#ifdef B2I_STEM_TRACE
cudaError_t stem_trace_read(long long* host) { return cudaMemcpyFromSymbol(host, g_stem_trace, sizeof(g_stem_trace)); }
cudaError_t stem_cta_read(long long* host) { return cudaMemcpyFromSymbol(host, g_stem_cta, sizeof(g_stem_cta)); }
#endif

cudaError_t launch_stem(const uint8_t* in_u8, void* out_bf16, const CUtensorMap* tmW,
                        const float* bias_host, int n_images, unsigned int* sched, int f16, cudaStream_t st) {
  static bool attr_set[16] = {};
  static int sm_count[16] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 16) return cudaErrorInvalidDevice;
  if (!attr_set[dev]) {
    cudaError_t e = cudaFuncSetAttribute(stem_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, STEM_SMEM);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(stem_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, STEM_SMEM);
    if (e != cudaSuccess) return e;
    e = cudaDeviceGetAttribute(&sm_count[dev], cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess) return e;
    attr_set[dev] = true;
  }
  if (n_images < 1 || !tmW || !sched || !bias_host) return cudaErrorInvalidValue;
  if (reinterpret_cast<uintptr_t>(in_u8) & 3) return cudaErrorMisalignedAddress;   // the loaders fetch aligned words
  StemParams p;
  p.in = in_u8;
  p.out = static_cast<__nv_bfloat16*>(out_bf16);
  for (int c = 0; c < 64; ++c) p.bias[c] = bias_host[c];
  p.total_tiles = n_images * 64;
  p.sched = sched;
  p.f16 = f16;
  const int grid = p.total_tiles < 2 * sm_count[dev] ? p.total_tiles : 2 * sm_count[dev];
  if (f16) return launch_pdl(stem_kernel<true>, dim3(grid), dim3(ST_THREADS), STEM_SMEM, st, *tmW, p);
  return launch_pdl(stem_kernel<false>, dim3(grid), dim3(ST_THREADS), STEM_SMEM, st, *tmW, p);
}

}  // namespace b2i
