// Fused ResNet stem, second form ("channels on the accumulator lanes"): u8 image -> normalise -> conv 7x7 s2 p3
// (3 -> 64, BN folded) -> ReLU -> max-pool 3x3 s2 p1 in one kernel, 16-bit out [N][56][56][64].  Same stages of the
// reference as stem.cu (imagenet normalise, resnet conv1 + bn1 + relu, max_pool2d: reference src/services.rs:492-493,
// SURVEY.md Appendix B.1 / B.2) and the same input patch in shared memory; what changes is which operand is which.
//
// stem.cu computes D[pixel][channel]: the MMA's A operand is the pixel patch (descriptor-direct), B the weights, and
// both are read from shared memory for every K step (6 KB per 128 x 64 x 16 MMA: the unit reads 128 B/clk, so an
// MMA that could take 32 cycles takes 48), the accumulator tile then goes through a 32 KB staging tile so that the
// 3 x 3 max can see neighbouring pixels, which sit on other TMEM lanes (other threads).  Shared-memory wavefronts,
// MMA reads plus staging traffic, were that kernel's limit (2607 per tile against 2395 cycles measured).
//
// Here D[(row parity, channel)][pixel]:
//   * A = the weights, 128 rows = 2 x 64 channels, kept in TENSOR MEMORY for the life of the CTA (144 columns):
//     the MMA reads no weights from shared memory at all.  Rows 0-63 hold filter row rho of the 7 x 7 window,
//     rows 64-127 hold filter row rho - 2: the same nine input rows rho = 0..8 under an even conv row serve that row
//     (filter rows 0-6) and the odd conv row below it (stride 2: two input rows further down).  M = 128 is needed
//     for the full MMA rate (M = 64 runs at half), and this pairing fills it with 14 / 18 useful rows.
//   * B = the pixel patch, N = 64 = 8 row pairs x 8 conv columns, descriptor-direct as before (8-byte pixels, stride
//     2: the next conv pixel is 16 bytes on = the rows of a no-swizzle K-major core matrix; next row PAIR 1280 B).
//     2 KB per MMA from shared memory instead of 6 KB; 36 MMAs of 32 tensor-pipe cycles per 16 x 16-pixel tile.
//   * the epilogue thread owns one (parity, channel) lane and receives PIXELS in its registers: the horizontal
//     3-tap max, and the vertical max of the two same-parity rows, are register operations on raw accumulators
//     (+bias, ReLU and the rounding are monotone, so they commute with max and are applied once per POOLED value).
//     Only the pooled values cross threads: 2 x 6 KB per tile through shared memory instead of 32 KB + 56 KB.
// One CTA per SM (all 512 TMEM columns: two accumulator sets of 128 + 144 of weights), 26 warps; tiles alternate
// between two loader groups and two epilogue groups (even / odd tile of the CTA's sequence = patch buffer =
// accumulator set), because both are latency chains (global loads; TMEM loads feeding dependent max chains)
// that take longer than the MMAs of a tile:
//   warps 0-15  two epilogue groups of eight warps: two sets (pooled rows 0-3 / 4-6) x four TMEM lane quarters
//   warps 16-17 MMA issue, one per half of the tile's columns
//   warps 18-25 two loader groups of four warps
#include <cstdio>
#include <cstdlib>
#include "conv.h"
#include "ops.h"
#include "ptx.cuh"

namespace b2i {
namespace {

constexpr int S2_DRAIN = 512;
constexpr int S2_GROUP = 192;                          // threads of one loader group
constexpr int S2_MMA_WARPS = 2;                         // one per half of the tile's columns: see the kernel
constexpr int S2_THREADS = S2_DRAIN + 32 * S2_MMA_WARPS + 2 * S2_GROUP;
constexpr int PATCH_H = 37, PATCH_W = 40;              // pixels of 8 bytes (r, g, b, 0 in the 16-bit storage format)
constexpr uint32_t PATCH_ROW_BYTES = PATCH_W * 8;      // 320
constexpr uint32_t PATCH_BYTES = ((PATCH_H * PATCH_W * 8 + 127) / 128) * 128;
constexpr uint32_t X_BYTES = 2 * 49 * 128;             // pooled candidates of one tile: [parity][7 x 7][64] 16-bit
constexpr int S2_PBUF = 2;                              // patch buffers per loader group: a group converts its next tile while the MMAs read its last one
constexpr uint32_t S2_SMEM = 2 * S2_PBUF * PATCH_BYTES + 2 * X_BYTES + 1024;
constexpr uint32_t TMEM_COLS = 512;
constexpr uint32_t A_COL = 256;                        // weights: columns 256 .. 399 (18 K chunks x 8 columns)
constexpr int KCHUNKS = 18;                            // 9 input rows x 2 halves of 4 taps x 4 channels

struct Stem2Params {
  const uint8_t* in;
  uint16_t* out;
  const uint32_t* w2;   // [128][144] packed pairs: row = parity * 64 + channel, K chunk t = rho * 2 + half
  float bias[64];
  int total_tiles;      // N * 64
  unsigned int* sched;  // [0] tiles handed out beyond the first 2 per CTA, [1] CTAs done; both zero between launches
  long long* trace;     // -DB2I_STEM2_TRACE + B2I_STEM2_TRACE=1: clock stamps of CTA 0's roles, else null
};
#ifdef B2I_STEM2_TRACE
constexpr bool kS2Trace = true;
#else
constexpr bool kS2Trace = false;
#endif

// B-operand descriptor over the raw patch (no swizzle, K-major): row x of an 8-row core matrix is the next conv
// pixel = 16 bytes on; the next 8-row group is the next row PAIR = 4 patch rows = 1280 bytes; the second core
// matrix of a K = 16 step (two more taps) is 16 bytes on.
__device__ __forceinline__ uint64_t patch_desc_pairs(uint32_t saddr) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((saddr & 0x3FFFFu) >> 4);
  d |= static_cast<uint64_t>(16 >> 4) << 16;
  d |= static_cast<uint64_t>((4 * PATCH_ROW_BYTES) >> 4) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  return d;
}

// D[tmem] (+)= A[tmem] * B[smem desc]^T; issued by one elected lane of a converged warp.
__device__ __forceinline__ void umma_ts_elect(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred pe, pa;\n\t"
      "elect.sync _|pe, 0xffffffff;\n\t"
      "setp.ne.b32 pa, %4, 0;\n\t"
      "@pe tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, pa;\n\t}" ::"r"(tmem_d),
      "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tmem_st8(uint32_t taddr, const uint32_t (&r)[8]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(taddr), "r"(r[0]), "r"(r[1]),
               "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
               : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, float (&r)[8]) {
  uint32_t v[8];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
               : "r"(taddr)
               : "memory");
#pragma unroll
  for (int i = 0; i < 8; ++i) r[i] = __uint_as_float(v[i]);
}
template <bool F16> __device__ __forceinline__ uint16_t to16(float v) {
  if constexpr (F16) return __half_as_ushort(__float2half_rn(v));
  else return __bfloat16_as_ushort(__float2bfloat16_rn(v));
}

template <bool F16>
__global__ void __launch_bounds__(S2_THREADS, 1)
stem2_kernel(const Stem2Params p) {
  griddep_launch();
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bars[4 * S2_PBUF + 4];   // patch_full[2 * S2_PBUF], patch_empty[2 * S2_PBUF], acc_full[2], acc_empty[2]
  __shared__ uint32_t tmem_slot;
  __shared__ int tile_ring[32];                // tile of iteration i (or -1 = this loader group is out of tiles)

  const int tid = threadIdx.x;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);   // warp-uniform, and the compiler knows it (MMA operands stay in uniform registers)
  const int lane = tid & 31;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t patch0 = base;
  const uint32_t x0 = base + 2 * S2_PBUF * PATCH_BYTES;
  const uint32_t bar0 = smem_u32(bars);
  // patch buffer of group g's k-th tile: slot g * S2_PBUF + k % S2_PBUF, its barriers flip every S2_PBUF tiles of the group
  auto patch_full = [&](int s) { return bar0 + 8u * s; };
  auto patch_empty = [&](int s) { return bar0 + 8u * (2 * S2_PBUF + s); };
  auto acc_full = [&](int b) { return bar0 + 8u * (4 * S2_PBUF + b); };
  auto acc_empty = [&](int b) { return bar0 + 8u * (4 * S2_PBUF + 2 + b); };

  if (tid == 0) {
    for (int s = 0; s < 2 * S2_PBUF; ++s) {
      mbar_init(patch_full(s), S2_GROUP);
      mbar_init(patch_empty(s), S2_MMA_WARPS);   // one commit per MMA-issuing warp
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(acc_full(b), S2_MMA_WARPS);
      mbar_init(acc_empty(b), 8);
    }
    mbar_fence_init();
  }
  constexpr int MMA_WARP = S2_DRAIN / 32;
  if (warp == MMA_WARP) tmem_alloc(smem_u32(&tmem_slot), TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, tmem_slot, 0);
  if (warp < 4) {
    // weights -> tensor memory, once: thread = row (parity, channel) of its warp's lane quarter
    const uint32_t* src = p.w2 + static_cast<size_t>(warp * 32 + lane) * (KCHUNKS * 8);
    const uint32_t t_row = tmem_base + (static_cast<uint32_t>(warp * 32) << 16) + A_COL;
#pragma unroll 2
    for (int t = 0; t < KCHUNKS; ++t) {
      const uint4 a = __ldg(reinterpret_cast<const uint4*>(src + t * 8));
      const uint4 b = __ldg(reinterpret_cast<const uint4*>(src + t * 8 + 4));
      const uint32_t r[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
      tmem_st8(t_row + static_cast<uint32_t>(t * 8), r);
    }
    tmem_st_wait();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  griddep_wait();   // the pooled map written below may still be read by the previous forward's kernels

  if (warp >= MMA_WARP && warp < MMA_WARP + S2_MMA_WARPS) {
    // ------------------------------------------------------------------ MMA issue
    // Two issuing warps, one per half of the tile's conv columns (its own 64 accumulator columns, the same weights).
    // History: an earlier reading of tools/microbench/umma_rate.cu had a TMEM-A MMA complete ~160 cycles after the same
    // warp's previous one whatever its N; that was the issuing loop of the benchmark (and of this kernel) ferrying
    // descriptors into uniform registers under `lane == 0`, not the tensor pipe.  Issued warp-uniformly the TMEM-A form
    // runs at the same floor as the shared-memory form (profiles/r03_microbench_umma_tmem_a_lean_issue.txt), and the
    // 36 MMAs of a tile are ~1150 cycles of pipe time.  What the clock stamps of one CTA show now (-DB2I_STEM2_TRACE):
    // ~1700 cycles per tile, with the loader groups (~2000 cycles of conversion per tile and group), the accumulator
    // hand-over (MMAs ~1100 + epilogue reads ~1200 per set) and the epilogue groups all near that rate.
    // tcgen05.commit tracks the MMAs of the committing thread, so both warps commit and the barriers count two arrivals.
    const int hx = warp - MMA_WARP;
    const uint32_t idesc = umma_idesc_bf16(128, 64, F16 ? 1 : 0);
    bool done[2] = {false, false};
    for (int i = 0;; ++i) {
      const int b = i & 1;
      if (done[b]) {
        if (done[b ^ 1]) break;
        continue;
      }
      const uint32_t ph = static_cast<uint32_t>((i >> 1) & 1);
      const bool tr = kS2Trace && p.trace != nullptr && blockIdx.x == 0 && hx == 0 && lane == 0 && i < 200;
      if (tr) p.trace[i * 12 + 0] = clock64();
      const int kk = i >> 1;                                   // the group's tile number
      const int ps = b * S2_PBUF + kk % S2_PBUF;               // its patch slot
      mbar_wait(acc_empty(b), ph ^ 1);
      if (tr) p.trace[i * 12 + 1] = clock64();
      mbar_wait(patch_full(ps), static_cast<uint32_t>((kk / S2_PBUF) & 1));
      if (tr) p.trace[i * 12 + 2] = clock64();
      if (*reinterpret_cast<volatile int*>(&tile_ring[i & 31]) < 0) {   // this group is out of tiles: tell the epilogue
        umma_commit_elect(acc_full(b));
        done[b] = true;
        continue;
      }
      tc_fence_after();
      const uint32_t pbase = patch0 + ps * PATCH_BYTES;
      {                                         // conv columns 8 * hx .. 8 * hx + 7 of the tile
        const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(b * 128 + hx * 64);
#pragma unroll
        for (int t = 0; t < KCHUNKS; ++t) {      // input row rho = t / 2 under the pair's even conv row, taps 4 * (t & 1) .. + 3
          const uint32_t b_addr = pbase + static_cast<uint32_t>((t >> 1) * PATCH_ROW_BYTES + hx * 128 + (t & 1) * 32);
          umma_ts_elect(d_tmem, tmem_base + A_COL + static_cast<uint32_t>(t * 8), patch_desc_pairs(b_addr), idesc, t != 0 ? 1u : 0u);
        }
      }
      umma_commit_elect(patch_empty(ps));   // the loaders may overwrite this patch buffer
      umma_commit_elect(acc_full(b));
      if (tr) p.trace[i * 12 + 3] = clock64();
    }
  } else if (warp >= MMA_WARP + S2_MMA_WARPS) {
    // ------------------------------------------------------------------ patch loaders: group g fills buffer g
    const int g = (warp - MMA_WARP - S2_MMA_WARPS) / (S2_GROUP / 32);
    const int lt = tid - (S2_DRAIN + 32 * S2_MMA_WARPS) - g * S2_GROUP;
    uint8_t* const gen_base = smem_raw + (base - smem_u32(smem_raw));
    __shared__ int next_tile[2];
    constexpr int GROUPS = PATCH_W / 4;                                       // 10 four-pixel tasks per patch row
    constexpr int TASKS = PATCH_H * GROUPS;                                   // 370
    constexpr int ROUNDS = (TASKS + S2_GROUP - 1) / S2_GROUP;                 // 3
    // (u/255 - mean)/std as one FMA per byte (bit-identical after rounding to 16 bits: checked exhaustively by the CPU tests)
    const float na[3] = {0x1.18926cp-6f, 0x1.1ed5b2p-6f, 0x1.1d8f56p-6f};
    const float nb[3] = {-0x1.0f177ap+1f, -0x1.04924ap+1f, -0x1.cdf012p+0f};
    int t = 2 * blockIdx.x + g;              // first tile of this group; the rest come from the global counter
    for (int k = 0;; ++k) {
      const int i = 2 * k + g;               // iteration number in the CTA's tile sequence
      if (lt == 0) tile_ring[i & 31] = t < p.total_tiles ? t : -1;
      const int ps = g * S2_PBUF + k % S2_PBUF;
      const uint32_t pph = static_cast<uint32_t>((k / S2_PBUF) & 1);
      if (t >= p.total_tiles) {
        mbar_wait(patch_empty(ps), pph ^ 1);
        mbar_arrive(patch_full(ps));
        break;
      }
      unsigned int ticket = 0;
      if (lt == 0) ticket = atomicAdd(p.sched, 1u);
      const bool trl = kS2Trace && p.trace != nullptr && blockIdx.x == 0 && lt == 0 && i < 200;
      if (trl) p.trace[i * 12 + 4] = clock64();
      const int n = t >> 6;
      const int pty = (t >> 3) & 7, ptx = t & 7;
      const int ih0 = 28 * pty - 5, iw0 = 28 * ptx - 5;
      const uint8_t* img = p.in + static_cast<size_t>(n) * (224 * 224 * 3);
      uint8_t* dst = gen_base + ps * PATCH_BYTES;
      uint32_t w[ROUNDS][4];
#pragma unroll
      for (int r = 0; r < ROUNDS; ++r) {
        const int e = lt + r * S2_GROUP;
        const int py = e / GROUPS, gg = e - py * GROUPS;
        const int ih = ih0 + py;
        const int wo = 84 * ptx - 16 + 12 * gg;            // byte offset of the first aligned word in the image row
        const bool row_ok = e < TASKS && ih >= 0 && ih < 224;
        const uint32_t* src = reinterpret_cast<const uint32_t*>(img + ih * 672 + wo);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int o = wo + 4 * j;
          w[r][j] = (row_ok && o >= 0 && o < 672) ? __ldg(src + j) : 0u;
        }
      }
      if (trl) p.trace[i * 12 + 5] = clock64() + (w[0][0] & 0);
      mbar_wait(patch_empty(ps), pph ^ 1);
      if (trl) p.trace[i * 12 + 6] = clock64();
#pragma unroll
      for (int r = 0; r < ROUNDS; ++r) {
        const int e = lt + r * S2_GROUP;
        if (e >= TASKS) continue;
        const int py = e / GROUPS, gg = e - py * GROUPS;
        const int ih = ih0 + py;
        const bool row_ok = ih >= 0 && ih < 224;
        auto byte_f = [&](int q) {
          return __uint_as_float(__byte_perm(w[r][q >> 2], 0x4B000000u, 0x7440u | (q & 3))) - 8388608.0f;
        };
        uint32_t o[8];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int iw = iw0 + 4 * gg + j;
          const bool ok = row_ok && iw >= 0 && iw < 224;
          const float rr = fmaf(byte_f(1 + 3 * j), na[0], nb[0]);
          const float gr = fmaf(byte_f(2 + 3 * j), na[1], nb[1]);
          const float bl = fmaf(byte_f(3 + 3 * j), na[2], nb[2]);
          o[2 * j] = ok ? pack16x2_t<F16>(rr, gr) : 0u;
          o[2 * j + 1] = ok ? pack16x2_t<F16>(bl, 0.f) : 0u;
        }
        uint4* d4 = reinterpret_cast<uint4*>(dst + e * 32);
        const int first = (lt >> 2) & 1;
        const uint4 lo = make_uint4(o[0], o[1], o[2], o[3]), hi = make_uint4(o[4], o[5], o[6], o[7]);
        d4[first] = first ? hi : lo;
        d4[first ^ 1] = first ? lo : hi;
      }
      fence_proxy_async_smem();
      mbar_arrive(patch_full(ps));
      if (trl) p.trace[i * 12 + 7] = clock64();
      if (lt == 0) next_tile[g] = static_cast<int>(2 * gridDim.x + ticket);
      if (g == 0) asm volatile("bar.sync 2, %0;" ::"n"(S2_GROUP) : "memory");
      else asm volatile("bar.sync 3, %0;" ::"n"(S2_GROUP) : "memory");
      t = *reinterpret_cast<volatile int*>(&next_tile[g]);
    }
  } else {
    // ------------------------------------------------------------------ epilogue: pool in registers, +bias, ReLU, store
    const int b = warp >> 3;                   // epilogue group = accumulator set = tiles 2k + b of the CTA's sequence
    const int set = (warp >> 2) & 1;           // 0: pooled rows 0-3, 1: pooled rows 4-6
    const int q = warp & 3;                    // TMEM lane quarter
    const int par = q >> 1;                    // 0: even conv rows of the tile (slots 2i), 1: odd rows (slots 2i + 1)
    const int ch = (q & 1) * 32 + lane;
    const float bias = p.bias[ch];
    const int py0 = set * 4, npy = set == 0 ? 4 : 3;
    const int st = tid & 127;                  // index inside the set
    uint8_t* const x_gen = smem_raw + (x0 - smem_u32(smem_raw));
    for (int k = 0;; ++k) {
      const int i = 2 * k + b;
      const bool tre = kS2Trace && p.trace != nullptr && blockIdx.x == 0 && (warp & 7) == 0 && lane == 0 && i < 200;
      if (tre) p.trace[i * 12 + 8] = clock64();
      mbar_wait(acc_full(b), static_cast<uint32_t>(k & 1));
      if (tre) p.trace[i * 12 + 9] = clock64();
      const int t = *reinterpret_cast<volatile int*>(&tile_ring[i & 31]);
      if (t < 0) break;
      const int n = t >> 6;
      const int pty = (t >> 3) & 7, ptx = t & 7;
      tc_fence_after();
      // this thread's accumulator row: lanes (quarter q), columns [b*128 + hx*64 + pair*8 + x]
      const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(b * 128);
      uint16_t* xw = reinterpret_cast<uint16_t*>(x_gen + b * X_BYTES + par * (49 * 128)) + ch;   // [pooled pixel][64]
      const float NEG = -3.0e38f;
      // horizontal 3-tap / stride-2 max of one conv row (raw accumulators): h[px] = max(v[2px], v[2px+1], v[2px+2]);
      // tile slot 0 is conv column -1 in the left-most tiles and must not take part
      auto hrow = [&](int pair, float (&h)[7]) {
        float v0[8], v1[8];
        tmem_ld8(t_row + static_cast<uint32_t>(pair * 8), v0);
        tmem_ld8(t_row + static_cast<uint32_t>(64 + pair * 8), v1);
        tmem_ld_wait();
        if (ptx == 0) v0[0] = NEG;
        h[0] = fmaxf(fmaxf(v0[0], v0[1]), v0[2]);
        h[1] = fmaxf(fmaxf(v0[2], v0[3]), v0[4]);
        h[2] = fmaxf(fmaxf(v0[4], v0[5]), v0[6]);
        h[3] = fmaxf(fmaxf(v0[6], v0[7]), v1[0]);
        h[4] = fmaxf(fmaxf(v1[0], v1[1]), v1[2]);
        h[5] = fmaxf(fmaxf(v1[2], v1[3]), v1[4]);
        h[6] = fmaxf(fmaxf(v1[4], v1[5]), v1[6]);
      };
      if (par == 0) {
        // pooled row py needs conv slots 2py and 2py + 2 = pairs py and py + 1 of this parity
        float prev[7], cur[7];
        hrow(py0, prev);
        if (pty == 0 && py0 == 0) {            // slot row 0 is conv row -1 in the top tiles
#pragma unroll
          for (int x = 0; x < 7; ++x) prev[x] = NEG;
        }
        for (int k = 0; k < npy; ++k) {
          hrow(py0 + k + 1, cur);
          const int pp = (py0 + k) * 7;
#pragma unroll
          for (int x = 0; x < 7; ++x) {
            const float m = fmaxf(prev[x], cur[x]);
            xw[(pp + x) * 64] = to16<F16>(fmaxf(m + bias, 0.f));
            prev[x] = cur[x];
          }
        }
      } else {
        float cur[7];
        for (int k = 0; k < npy; ++k) {
          hrow(py0 + k, cur);
          const int pp = (py0 + k) * 7;
#pragma unroll
          for (int x = 0; x < 7; ++x) xw[(pp + x) * 64] = to16<F16>(fmaxf(cur[x] + bias, 0.f));
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(acc_empty(b));   // this warp has read what it needs of accumulator set b
      if (tre) p.trace[i * 12 + 10] = clock64();
      asm volatile("bar.sync %0, %1;" ::"r"(4 + 2 * b + set), "n"(128) : "memory");   // the set's 128 threads
      // the set's pooled rows: max of the two parities' candidates, 16 bytes (8 channels) per task
      const uint8_t* xe = x_gen + b * X_BYTES;
      const int ntask = npy * 7 * 8;
      for (int task = st; task < ntask; task += 128) {
        const int pp = py0 * 7 + (task >> 3);
        const int c8 = task & 7;
        const uint4 e = *reinterpret_cast<const uint4*>(xe + pp * 128 + c8 * 16);
        const uint4 o = *reinterpret_cast<const uint4*>(xe + 49 * 128 + pp * 128 + c8 * 16);
        uint4 m;
        m.x = max16x2_t<F16>(e.x, o.x); m.y = max16x2_t<F16>(e.y, o.y);
        m.z = max16x2_t<F16>(e.z, o.z); m.w = max16x2_t<F16>(e.w, o.w);
        const int py = pp / 7, px = pp - py * 7;
        const int ph = pty * 7 + py, pw = ptx * 7 + px;
        *reinterpret_cast<uint4*>(p.out + ((static_cast<size_t>(n) * 56 + ph) * 56 + pw) * 64 + c8 * 8) = m;
      }
      // this group's candidates buffer is rewritten by its next tile: order these reads before those writes
      asm volatile("bar.sync %0, %1;" ::"r"(4 + 2 * b + set), "n"(128) : "memory");
      if (tre) p.trace[i * 12 + 11] = clock64();
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
  if (warp == MMA_WARP) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

}  // namespace

// Host side: the [128][144] packed weight matrix of the kernel from conv1's folded weights in the layout stem.cu
// uses ([64][256], K order (r < 8, s < 8, c < 4), zeros at r = 7, s = 7, c = 3).  16-bit values, two per word.
void stem2_pack_weights(const uint16_t* w64x256, uint32_t* out) {
  for (int par = 0; par < 2; ++par)
    for (int c = 0; c < 64; ++c)
      for (int t = 0; t < KCHUNKS; ++t) {
        const int rho = t >> 1, half = t & 1;
        const int r = rho - 2 * par;
        for (int e = 0; e < 16; e += 2) {
          uint16_t v[2] = {0, 0};
          for (int u = 0; u < 2; ++u) {
            const int s = 4 * half + ((e + u) >> 2), chn = (e + u) & 3;
            if (r >= 0 && r <= 6) v[u] = w64x256[c * 256 + (r * 8 + s) * 4 + chn];
          }
          out[static_cast<size_t>(par * 64 + c) * (KCHUNKS * 8) + t * 8 + (e >> 1)] = static_cast<uint32_t>(v[0]) | (static_cast<uint32_t>(v[1]) << 16);
        }
      }
}

cudaError_t launch_stem2(const uint8_t* in_u8, void* out16, const uint32_t* w2_dev, const float* bias_host, int n_images,
                         unsigned int* sched, int f16, cudaStream_t st) {
  static bool attr_set[16] = {};
  static int sm_count[16] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 16) return cudaErrorInvalidDevice;
  if (!attr_set[dev]) {
    cudaError_t e = cudaFuncSetAttribute(stem2_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, S2_SMEM);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(stem2_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, S2_SMEM);
    if (e != cudaSuccess) return e;
    e = cudaDeviceGetAttribute(&sm_count[dev], cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess) return e;
    attr_set[dev] = true;
  }
  if (n_images < 1 || !w2_dev || !sched || !bias_host) return cudaErrorInvalidValue;
  if (reinterpret_cast<uintptr_t>(in_u8) & 3) return cudaErrorMisalignedAddress;   // the loaders fetch aligned words
  Stem2Params p;
  p.in = in_u8;
  p.out = static_cast<uint16_t*>(out16);
  p.w2 = w2_dev;
  for (int c = 0; c < 64; ++c) p.bias[c] = bias_host[c];
  p.total_tiles = n_images * 64;
  p.sched = sched;
  p.trace = nullptr;
  static const bool trace_on = [] { const char* e = getenv("B2I_STEM2_TRACE"); return e && e[0] == '1'; }();
  if (kS2Trace && trace_on) {
    static long long* d_trace = nullptr;
    if (!d_trace) cudaMalloc(&d_trace, 2400 * sizeof(long long));
    cudaMemsetAsync(d_trace, 0, 2400 * sizeof(long long), st);
    p.trace = d_trace;
    const int grid_t = (p.total_tiles + 1) / 2 < sm_count[dev] ? (p.total_tiles + 1) / 2 : sm_count[dev];
    cudaError_t e = f16 ? launch_pdl(stem2_kernel<true>, dim3(grid_t), dim3(S2_THREADS), S2_SMEM, st, p)
                        : launch_pdl(stem2_kernel<false>, dim3(grid_t), dim3(S2_THREADS), S2_SMEM, st, p);
    if (e != cudaSuccess) return e;
    e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) return e;
    static long long h[2400];
    cudaMemcpy(h, d_trace, sizeof h, cudaMemcpyDeviceToHost);
    const long long t0 = h[4] ? h[4] : h[0];
    fprintf(stderr, "[stem2 trace] CTA 0, tile by tile (cycles since the first loader started): loader: loads issued / data in registers / buffer free / patch written | MMA warp: starts waiting / accumulators free / patch full / MMAs issued | epilogue: starts waiting / accumulators full / accumulators read / tile stored\n");
    for (int i = 0; i < 200 && (h[i * 12 + 4] || h[i * 12]); ++i) {
      if (i < 24 || (i % 20) == 0)
        fprintf(stderr, "  tile %3d: %7lld %7lld %7lld %7lld | %7lld %7lld %7lld %7lld | %7lld %7lld %7lld %7lld\n", i, h[i * 12 + 4] - t0, h[i * 12 + 5] - t0, h[i * 12 + 6] - t0,
                h[i * 12 + 7] - t0, h[i * 12 + 0] - t0, h[i * 12 + 1] - t0, h[i * 12 + 2] - t0, h[i * 12 + 3] - t0, h[i * 12 + 8] - t0, h[i * 12 + 9] - t0, h[i * 12 + 10] - t0,
                h[i * 12 + 11] - t0);
    }
    return cudaSuccess;
  }
  const int grid = (p.total_tiles + 1) / 2 < sm_count[dev] ? (p.total_tiles + 1) / 2 : sm_count[dev];
  if (f16) return launch_pdl(stem2_kernel<true>, dim3(grid), dim3(S2_THREADS), S2_SMEM, st, p);
  return launch_pdl(stem2_kernel<false>, dim3(grid), dim3(S2_THREADS), S2_SMEM, st, p);
}

}  // namespace b2i
