// bf16 implicit-GEMM convolution / FC on the 5th-gen tensor cores (tcgen05 + TMEM), sm_100a only.
//
// Replaces the libtorch-CPU conv2d / linear calls the reference reaches through
// `forward_t` (reference src/services.rs:493, model graphs in tch::vision::{resnet,alexnet},
// SURVEY.md 2.3 / Appendix C).  Eval-mode BatchNorm is folded into the weights at load time, so
// the epilogue is  out = relu?( acc + bias (+ residual) )  with one bf16 rounding per layer.
//
// One CTA computes one 128 x BN output tile (128 output pixels x BN output channels):
//   warps 0-3  producer (gather modes: cp.async im2col gather into 128B-swizzled smem) and,
//              after the main loop, epilogue (TMEM -> registers -> bias/residual/ReLU -> global)
//   warp  4    TMA producer: im2col-mode / tiled tensor maps for A, tiled map for the weights
//   warp  5    TMEM allocation + single-thread tcgen05.mma issue, tcgen05.commit -> mbarriers
// Operand tiles are K-major rows of 64 bf16 (128 bytes) with the 128-byte swizzle, which is what
// both TMA (CU_TENSOR_MAP_SWIZZLE_128B) and the hand gather produce and what the UMMA shared-memory
// descriptors in ptx.cuh declare.
#include <algorithm>
#include <cstdio>

#include "conv.h"
#include "ptx.cuh"

namespace b2i {

namespace {

constexpr int BM = 128;
constexpr int BK = 64;
constexpr int A_TILE_BYTES = BM * BK * 2;
constexpr int MAX_STAGES = 8;
constexpr int NTHREADS = 192;
constexpr int LAG = 2;  // cp.async groups kept in flight per gather thread (needs stages >= 3)

struct KP {
  int H, W, C, OH, OW, R, S, Sp, stride, pad;
  int Cout, K, M, BN, stages, num_kb, tiles_n, tiles_m, ldo, relu, out_fp32;
  uint32_t tmem_cols;
  const __nv_bfloat16* x;
  const __nv_bfloat16* w;
  const float* bias;
  const __nv_bfloat16* residual;
  void* out;
  // optional second problem sharing the A operand (a 1x1/s2 downsample riding on its sibling 3x3/s2
  // conv): `tiles_n2` extra N-tiles that run over k-blocks [kb_first2, kb_first2 + num_kb2) of the
  // same im2col stream with their own weights (tmB2), bias and output; no residual, bf16 out.
  int tiles_n2, kb_first2, num_kb2, ldo2, relu2, Cout2;
  const float* bias2;
  void* out2;
  int resb;   // pair kernel: this CTA's half of ALL weight k-blocks stays resident in shared memory
  int msub;   // pair kernel: 128-row sub-tiles per CTA (1 or 2)
  int f16;    // 16-bit tensors are IEEE half instead of bf16
  int rev;    // pair kernel: walk the tiles from the last to the first
  int nprod;  // pair kernel: warps that issue TMA requests (1..4), k-blocks dealt round-robin among them
  // persistent kernel, split-K (FC layers with few rows: all SMs stream a share of the weights): `splits` CTAs
  // compute one output tile over disjoint k-block ranges, each stores its accumulators in its own slice of `ws`
  // (fp32 [splits][tiles_m * 128][ldw]) and the one that arrives last on `ctr[tile]` (all zero between launches)
  // sums the slices in split order, runs the epilogue and re-zeroes the counter.
  int splits, ldw;
  float* ws;
  int* ctr;
  int a_rows;   // persistent kernel, A_TILED: rows of the A box (the tensor map's), 128 unless the layer has fewer rows
  long long* trace;   // pair kernel, debugging (B2I_PAIR_TRACE=1): clock stamps of one CTA's producers and MMA warp, else null
  int trace_cta;
};

// Where one tile's results go.
struct Epi {
  const float* bias;
  uint32_t bias_smem;   // shared-memory copy of `bias` (address; 0 = none): the epilogue reads it from there instead of L2
  const __nv_bfloat16* residual;
  void* out;
  int ldo, relu, out_fp32, Cout, f16;
};
__device__ __forceinline__ Epi epi_of(const KP& p, bool second) {
  Epi e;
  e.bias = second ? p.bias2 : p.bias;
  e.bias_smem = 0;
  e.residual = second ? nullptr : p.residual;
  e.out = second ? p.out2 : p.out;
  e.ldo = second ? p.ldo2 : p.ldo;
  e.relu = second ? p.relu2 : p.relu;
  e.out_fp32 = second ? 0 : p.out_fp32;
  e.Cout = second ? p.Cout2 : p.Cout;
  e.f16 = p.f16;
  return e;
}


// 16-bit tail of epilogue_chunk for one storage format: (+ residual) -> ReLU -> pack -> two 32-byte stores.
template <bool F16>
__device__ __forceinline__ void epilogue_store16(const Epi& p, size_t off, float (&f)[32], const uint32_t (*rpre)[16]) {
  if (p.residual) {
    // Written by an earlier kernel that may have overlapped this one's start: read it from L2, never L1,
    // one whole 32-byte sector per load (row stride and channel offset are multiples of 32 channels).
    if (rpre) {
#pragma unroll
      for (int e = 0; e < 16; ++e) {
        f[e * 2] += lo16_t<F16>((*rpre)[e]);
        f[e * 2 + 1] += hi16_t<F16>((*rpre)[e]);
      }
    } else {
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        uint32_t rr[8];
        ldg256_cg(p.residual + off + q * 16, rr);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          f[q * 16 + e * 2] += lo16_t<F16>(rr[e]);
          f[q * 16 + e * 2 + 1] += hi16_t<F16>(rr[e]);
        }
      }
    }
  }
  if (p.relu) {
#pragma unroll
    for (int j = 0; j < 32; ++j) f[j] = fmaxf(f[j], 0.f);
  }
  __nv_bfloat16* op = reinterpret_cast<__nv_bfloat16*>(p.out) + off;
#pragma unroll
  for (int q = 0; q < 2; ++q) {
    uint32_t ov[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) ov[e] = pack16x2_t<F16>(f[q * 16 + 2 * e], f[q * 16 + 2 * e + 1]);
    stg256(op + q * 16, ov);
  }
}

// One output row x 32 consecutive channels: acc + bias (+ residual) -> ReLU -> store.
// `rpre`: the residual's 32 channels already in registers (see prefetch_residual), or null to load them here.
template <bool F16>
__device__ __forceinline__ void epilogue_chunk(const Epi& p, int m, int nb, const uint32_t (&v)[32],
                                               const uint32_t (*rpre)[16] = nullptr) {
  float f[32];
  if (p.bias_smem) {   // staged at kernel start: no L2 round trip between the accumulator and the store
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float4 b;
      asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w) : "r"(p.bias_smem + static_cast<uint32_t>(nb + 4 * j) * 4u));
      f[4 * j + 0] = __uint_as_float(v[4 * j + 0]) + b.x;
      f[4 * j + 1] = __uint_as_float(v[4 * j + 1]) + b.y;
      f[4 * j + 2] = __uint_as_float(v[4 * j + 2]) + b.z;
      f[4 * j + 3] = __uint_as_float(v[4 * j + 3]) + b.w;
    }
  } else {
    const float4* b4 = reinterpret_cast<const float4*>(p.bias + nb);   // bias is padded to Cout_pad: always in range
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float4 b = __ldg(b4 + j);
      f[4 * j + 0] = __uint_as_float(v[4 * j + 0]) + b.x;
      f[4 * j + 1] = __uint_as_float(v[4 * j + 1]) + b.y;
      f[4 * j + 2] = __uint_as_float(v[4 * j + 2]) + b.z;
      f[4 * j + 3] = __uint_as_float(v[4 * j + 3]) + b.w;
    }
  }
  if (p.out_fp32) {
    float* o = reinterpret_cast<float*>(p.out) + static_cast<size_t>(m) * p.ldo + nb;
    if (p.relu) {
#pragma unroll
      for (int j = 0; j < 32; ++j) f[j] = fmaxf(f[j], 0.f);
    }
    if (nb + 32 <= p.Cout && (p.ldo & 3) == 0) {
#pragma unroll
      for (int j = 0; j < 32; j += 4)
        *reinterpret_cast<float4*>(o + j) = make_float4(f[j], f[j + 1], f[j + 2], f[j + 3]);
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (nb + j < p.Cout) o[j] = f[j];
    }
    return;
  }
  if (nb >= p.Cout) return;
  const size_t off = static_cast<size_t>(m) * p.ldo + nb;
  epilogue_store16<F16>(p, off, f, rpre);
}

// Issue the residual loads of one output row x 32 channels early (before the accumulator is ready), so
// their L2 latency overlaps the wait instead of following it.
__device__ __forceinline__ void prefetch_residual(const Epi& p, int m, int nb, uint32_t (&r)[16]) {
  const __nv_bfloat16* src = p.residual + static_cast<size_t>(m) * p.ldo + nb;
  uint32_t lo[8], hi[8];
  ldg256_cg(src, lo);
  ldg256_cg(src + 16, hi);
#pragma unroll
  for (int e = 0; e < 8; ++e) { r[e] = lo[e]; r[8 + e] = hi[e]; }
}

template <int AMODE, bool B_TMA, bool F16>
__global__ void __launch_bounds__(NTHREADS)
conv_umma_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const KP p) {
  constexpr bool A_IS_TMA = (AMODE == A_IM2COL || AMODE == A_TILED);
  static_assert(!(A_IS_TMA && !B_TMA), "TMA A with gathered B is not a supported combination");

  griddep_launch();
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bars[2 * MAX_STAGES + 1];
  __shared__ uint32_t tmem_slot;

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);   // warp-uniform, and the compiler knows it (MMA operands stay in uniform registers)
  const int lane = threadIdx.x & 31;
  const int tile_n = blockIdx.x % p.tiles_n;
  const int tile_m = blockIdx.x / p.tiles_n;
  const int m0 = tile_m * BM;
  const int n0 = tile_n * p.BN;

  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t b_tile_bytes = static_cast<uint32_t>(p.BN) * 128u;
  const uint32_t stage_bytes = A_TILE_BYTES + b_tile_bytes;
  const uint32_t bar0 = smem_u32(bars);
  auto full_bar = [&](int s) { return bar0 + 8u * s; };
  auto empty_bar = [&](int s) { return bar0 + 8u * (MAX_STAGES + s); };
  const uint32_t tmem_full_bar = bar0 + 8u * (2 * MAX_STAGES);

  if (threadIdx.x == 0) {
    const uint32_t full_count = (A_IS_TMA ? 0u : 128u) + (B_TMA ? 1u : 0u);
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(full_bar(s), full_count);
      mbar_init(empty_bar(s), 1);
    }
    mbar_init(tmem_full_bar, 1);
    mbar_fence_init();
  }
  if (warp == 4 && lane == 0) {
    if (A_IS_TMA) tma_prefetch_desc(&tmA);
    if (B_TMA) tma_prefetch_desc(&tmB);
  }
  if (warp == 5) tmem_alloc(smem_u32(&tmem_slot), p.tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, tmem_slot, 0);
  griddep_wait();   // everything above overlaps the previous kernel's tail; activations are read below

  if (warp < 4) {
    // ------------------------------------------------------------------ gather producer
    if (!A_IS_TMA) {
      const int t = threadIdx.x;
      const int chunk = t & 7;   // 16-byte chunk inside the 128-byte row
      const int r0 = t >> 3;     // rows r0 + 16*i
      const uint32_t swz = static_cast<uint32_t>(chunk ^ (r0 & 7)) << 4;
      int ih0[8], iw0[8];
      const __nv_bfloat16* rowbase[8];
      const int ohw = p.OH * p.OW;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int m = m0 + r0 + 16 * i;
        if (m < p.M) {
          const int n = m / ohw;
          const int rem = m - n * ohw;
          const int oh = rem / p.OW;
          const int ow = rem - oh * p.OW;
          ih0[i] = oh * p.stride - p.pad;
          iw0[i] = ow * p.stride - p.pad;
          rowbase[i] = p.x + static_cast<size_t>(n) * p.H * p.W * p.C;
        } else {
          ih0[i] = -(1 << 28);  // every tap lands out of bounds -> zero fill
          iw0[i] = 0;
          rowbase[i] = p.x;
        }
      }
      const int cpb = (AMODE == A_GATHER) ? (p.C >> 6) : 1;  // 64-channel blocks per tap
      const int b_passes = p.BN >> 4;
      int s_w = 0, ph_w = 0;  // stage / phase of the k-block being issued
      int s_a = 0;            // stage of the k-block being published
      for (int it = 0; it < p.num_kb + LAG; ++it) {
        if (it < p.num_kb) {
          mbar_wait(empty_bar(s_w), ph_w ^ 1);
          const uint32_t a_s = smem_base + s_w * stage_bytes;
          if (AMODE == A_GATHER) {
            const int tap = it / cpb;
            const int c0 = (it - tap * cpb) << 6;
            const int r = tap / p.S;
            const int s = tap - r * p.S;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const int ih = ih0[i] + r;
              const int iw = iw0[i] + s;
              const bool ok = (ih >= 0) && (ih < p.H) && (iw >= 0) && (iw < p.W);
              const __nv_bfloat16* src =
                  ok ? rowbase[i] + (static_cast<size_t>(ih) * p.W + iw) * p.C + c0 + chunk * 8
                     : p.x;
              cp_async16(a_s + static_cast<uint32_t>(r0 + 16 * i) * 128u + swz, src, ok ? 16u : 0u);
            }
          } else {  // A_SMALLC: C == 4, one 8-byte pixel per tap, two taps per 16-byte chunk
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              const int tap = it * 16 + chunk * 2 + h;
              const int r = tap / p.Sp;
              const int s = tap - r * p.Sp;
              const bool tap_ok = (r < p.R) && (s < p.S);
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                const int ih = ih0[i] + r;
                const int iw = iw0[i] + s;
                const bool ok = tap_ok && (ih >= 0) && (ih < p.H) && (iw >= 0) && (iw < p.W);
                const __nv_bfloat16* src =
                    ok ? rowbase[i] + (static_cast<size_t>(ih) * p.W + iw) * 4 : p.x;
                cp_async8(a_s + static_cast<uint32_t>(r0 + 16 * i) * 128u + swz + 8u * h, src,
                          ok ? 8u : 0u);
              }
            }
          }
          if (!B_TMA) {
            const uint32_t b_s = a_s + A_TILE_BYTES;
            const __nv_bfloat16* wsrc =
                p.w + static_cast<size_t>(n0 + r0) * p.K + static_cast<size_t>(it) * BK + chunk * 8;
            for (int i = 0; i < b_passes; ++i) {
              cp_async16(b_s + static_cast<uint32_t>(r0 + 16 * i) * 128u + swz,
                         wsrc + static_cast<size_t>(16 * i) * p.K, 16u);
            }
          }
          if (++s_w == p.stages) { s_w = 0; ph_w ^= 1; }
        }
        cp_async_commit();
        if (it >= LAG) {
          cp_async_wait<LAG>();       // k-block (it - LAG) has landed for this thread
          fence_proxy_async_smem();   // generic-proxy writes -> visible to the tensor-core proxy
          mbar_arrive(full_bar(s_a));
          if (++s_a == p.stages) s_a = 0;
        }
      }
    }

    // ------------------------------------------------------------------ epilogue
    mbar_wait(tmem_full_bar, 0);
    tc_fence_after();
    const int m = m0 + warp * 32 + lane;
    const bool row_ok = m < p.M;
    const uint32_t t_row = tmem_base + (static_cast<uint32_t>(warp * 32) << 16);
    const int n_chunks = p.BN >> 5;
    for (int c = 0; c < n_chunks; ++c) {
      uint32_t v[32];
      tmem_ld32(t_row + static_cast<uint32_t>(c * 32), v);
      tmem_ld_wait();
      if (row_ok) epilogue_chunk<F16>(epi_of(p, false), m, n0 + c * 32, v);
    }
  } else if (warp == 4) {
    // ------------------------------------------------------------------ TMA producer
    if (B_TMA && lane == 0) {
      int iw_base = 0, ih_base = 0, n_base = 0;
      if (AMODE == A_IM2COL) {
        const int ohw = p.OH * p.OW;
        n_base = m0 / ohw;
        const int rem = m0 - n_base * ohw;
        const int oh = rem / p.OW;
        const int ow = rem - oh * p.OW;
        ih_base = oh * p.stride - p.pad;
        iw_base = ow * p.stride - p.pad;
      }
      const int cpb = (AMODE == A_IM2COL) ? (p.C >> 6) : 1;
      const uint32_t tx = b_tile_bytes + (A_IS_TMA ? A_TILE_BYTES : 0);
      int s = 0, ph = 0;
      for (int kb = 0; kb < p.num_kb; ++kb) {
        mbar_wait(empty_bar(s), ph ^ 1);
        const uint32_t a_s = smem_base + s * stage_bytes;
        mbar_expect_tx(full_bar(s), tx);
        if (AMODE == A_IM2COL) {
          const int tap = kb / cpb;
          const int c0 = (kb - tap * cpb) << 6;
          const int r = tap / p.S;
          const int sx = tap - r * p.S;
          tma_load_im2col_4d(a_s, &tmA, full_bar(s), c0, iw_base, ih_base, n_base,
                             static_cast<uint16_t>(sx), static_cast<uint16_t>(r));
        } else if (AMODE == A_TILED) {
          tma_load_2d(a_s, &tmA, full_bar(s), kb * BK, m0);
        }
        tma_load_2d(a_s + A_TILE_BYTES, &tmB, full_bar(s), kb * BK, n0);
        if (++s == p.stages) { s = 0; ph ^= 1; }
      }
    }
  } else {
    // ------------------------------------------------------------------ MMA issuer (warp 5)
    const uint32_t idesc = umma_idesc_bf16(BM, p.BN, F16 ? 1 : 0);
    int s = 0, ph = 0;
    for (int kb = 0; kb < p.num_kb; ++kb) {
      mbar_wait(full_bar(s), ph);
      tc_fence_after();
      {
        const uint32_t a_s = smem_base + s * stage_bytes;
        const uint64_t adesc = umma_smem_desc_sw128(a_s);
        const uint64_t bdesc = umma_smem_desc_sw128(a_s + A_TILE_BYTES);
#pragma unroll
        for (int k = 0; k < BK / 16; ++k) {
          // advancing 16 bf16 (32 bytes) along K inside the swizzle atom: +2 in 16-byte units
          umma_bf16_elect(tmem_base, adesc + 2u * k, bdesc + 2u * k, idesc, (kb | k) != 0 ? 1u : 0u);
        }
        umma_commit_elect(empty_bar(s));  // smem slot reusable once these MMAs have read it
        if (kb == p.num_kb - 1) umma_commit_elect(tmem_full_bar);
      }
      if (++s == p.stages) { s = 0; ph ^= 1; }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 5) {
    tc_fence_after();
    tmem_dealloc(tmem_base, p.tmem_cols);
  }
}

// ------------------------------------------------------------------------------------------------
// Persistent variant for the TMA-fed modes: one CTA per SM walks a static round-robin list of
// output tiles.  The TMA warp streams k-blocks of successive tiles through one shared-memory ring
// without draining between tiles, the MMA thread alternates between two TMEM accumulators, and the
// four epilogue warps drain accumulator (i & 1) while the tensor core fills the other one.
//   warp 0: TMA producer   warp 1: TMEM alloc + MMA issue   warps 2-9: epilogue
constexpr int PTHREADS = 64 + 256;   // TMA warp, MMA warp, 8 epilogue warps (two per TMEM lane quarter)

// As in the pair kernel below, the k-blocks are dealt round-robin to up to three TMA-issuing warps (warp 0 and the
// two warps after the epilogue warps): the requests of one warp complete one after the other, those of
// different warps overlap.  An FC layer on a small batch is a pure weight stream and was bound by exactly that
// (517 cycles per 8 KB weight tile with one issuing warp, whatever the tile width or the A box).
constexpr int PERS_MAX_PROD = 3;
constexpr int PERS_THREADS = PTHREADS + 32 * (PERS_MAX_PROD - 1);
template <int AMODE, bool F16>
__global__ void __launch_bounds__(PERS_THREADS, 1)
conv_umma_persistent_kernel(const __grid_constant__ CUtensorMap tmA,
                            const __grid_constant__ CUtensorMap tmB,
                            const __grid_constant__ CUtensorMap tmB2, const KP p) {
  griddep_launch();
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bars[2 * MAX_STAGES + 4];
  __shared__ uint32_t tmem_slot;
  __shared__ int split_last;

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);   // warp-uniform, and the compiler knows it (MMA operands stay in uniform registers)
  const int lane = threadIdx.x & 31;
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t b_tile_bytes = static_cast<uint32_t>(p.BN) * 128u;
  const uint32_t stage_bytes = A_TILE_BYTES + b_tile_bytes;
  const int splits = p.splits;                                   // 1 unless the launcher chose split-K
  const int kb_per = (p.num_kb + splits - 1) / splits;           // k-blocks of one split (the last may be shorter, never empty)
  const uint32_t bar0 = smem_u32(bars);
  auto full_bar = [&](int s) { return bar0 + 8u * s; };
  auto empty_bar = [&](int s) { return bar0 + 8u * (MAX_STAGES + s); };
  auto acc_full_bar = [&](int b) { return bar0 + 8u * (2 * MAX_STAGES + b); };
  auto acc_empty_bar = [&](int b) { return bar0 + 8u * (2 * MAX_STAGES + 2 + b); };
  const int tn_all = p.tiles_n + p.tiles_n2;
  const int total_tiles = p.tiles_m * tn_all * splits;           // work items: (output tile, split), splits of a tile adjacent

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(acc_full_bar(b), 1);
      mbar_init(acc_empty_bar(b), 8);  // one arrival per epilogue warp
    }
    mbar_fence_init();
  }
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1) tmem_alloc(smem_u32(&tmem_slot), p.tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, tmem_slot, 0);
  griddep_wait();   // everything above overlaps the previous kernel's tail; activations are read below

  const int prod_id = warp == 0 ? 0 : (warp >= PTHREADS / 32 ? warp - PTHREADS / 32 + 1 : -1);
  if (prod_id >= 0) {
    if (lane == 0 && prod_id < p.nprod) {
      int turn = 0;   // whose k-block the next one is
      const int cpb = (AMODE == A_IM2COL) ? (p.C >> 6) : 1;
      const int ohw = p.OH * p.OW;
      // A_TILED with fewer than 128 rows in the layer: the map's box may hold only those rows; the MMA still reads 128
      // rows, the rest are whatever the ring held and are never stored
      const uint32_t tx = b_tile_bytes + (AMODE == A_TILED ? static_cast<uint32_t>(p.a_rows) * 128u : static_cast<uint32_t>(A_TILE_BYTES));
      int s = 0, ph = 0;
      for (int w = blockIdx.x; w < total_tiles; w += gridDim.x) {
        const int t = w / splits, sp = w - t * splits;
        const int tm = t / tn_all;
        const int nn = t - tm * tn_all;
        const bool second = nn >= p.tiles_n;
        const int m0 = tm * BM;
        const int n0 = (second ? nn - p.tiles_n : nn) * p.BN;
        const int kb_begin = second ? p.kb_first2 : sp * kb_per;
        const int kb_count = second ? p.num_kb2 : min(kb_per, p.num_kb - sp * kb_per);
        int iw_base = 0, ih_base = 0, n_base = 0;
        if (AMODE == A_IM2COL) {
          n_base = m0 / ohw;
          const int rem = m0 - n_base * ohw;
          const int oh = rem / p.OW;
          const int ow = rem - oh * p.OW;
          ih_base = oh * p.stride - p.pad;
          iw_base = ow * p.stride - p.pad;
        }
        const int tap0 = kb_begin / cpb;
        int cblk = kb_begin - tap0 * cpb, r = tap0 / p.S, sx = tap0 - (tap0 / p.S) * p.S;
        for (int j = 0; j < kb_count; ++j) {
          const int kb = kb_begin + j;
          const bool mine = turn == prod_id;
          if (++turn == p.nprod) turn = 0;
          if (mine) {
            mbar_wait(empty_bar(s), ph ^ 1);
            const uint32_t a_s = smem_base + s * stage_bytes;
            mbar_expect_tx(full_bar(s), tx);
            if (AMODE == A_IM2COL) {
              tma_load_im2col_4d(a_s, &tmA, full_bar(s), cblk << 6, iw_base, ih_base, n_base,
                                 static_cast<uint16_t>(sx), static_cast<uint16_t>(r));
            } else {
              tma_load_2d(a_s, &tmA, full_bar(s), kb * BK, m0);
            }
            if (second) tma_load_2d(a_s + A_TILE_BYTES, &tmB2, full_bar(s), j * BK, n0);
            else tma_load_2d(a_s + A_TILE_BYTES, &tmB, full_bar(s), kb * BK, n0);
          }
          if (AMODE == A_IM2COL && ++cblk == cpb) {
            cblk = 0;
            if (++sx == p.S) { sx = 0; ++r; }
          }
          if (++s == p.stages) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    const uint32_t idesc = umma_idesc_bf16(BM, p.BN, F16 ? 1 : 0);
    int s = 0, ph = 0, i = 0;
    for (int w = blockIdx.x; w < total_tiles; w += gridDim.x, ++i) {
      const int t = w / splits, sp = w - t * splits;
      const int buf = i & 1;
      mbar_wait(acc_empty_bar(buf), ((i >> 1) & 1) ^ 1);  // epilogue has drained this accumulator
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(buf * p.BN);
      const int kb_count = (t % tn_all) >= p.tiles_n ? p.num_kb2 : min(kb_per, p.num_kb - sp * kb_per);
      for (int kb = 0; kb < kb_count; ++kb) {
        mbar_wait(full_bar(s), ph);
        tc_fence_after();
        {
          const uint32_t a_s = smem_base + s * stage_bytes;
          const uint64_t adesc = umma_smem_desc_sw128(a_s);
          const uint64_t bdesc = umma_smem_desc_sw128(a_s + A_TILE_BYTES);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k)
            umma_bf16_elect(d_tmem, adesc + 2u * k, bdesc + 2u * k, idesc, (kb | k) != 0 ? 1u : 0u);
          umma_commit_elect(empty_bar(s));
          if (kb == kb_count - 1) umma_commit_elect(acc_full_bar(buf));
        }
        if (++s == p.stages) { s = 0; ph ^= 1; }
      }
    }
  } else {
    const int q = warp & 3;  // TMEM lane quarter this warp may read
    const int n_chunks = p.BN >> 5;
    int i = 0;
    for (int w = blockIdx.x; w < total_tiles; w += gridDim.x, ++i) {
      const int t = w / splits;
      const int buf = i & 1;
      const int tm = t / tn_all;
      const int nn = t - tm * tn_all;
      const bool second = nn >= p.tiles_n;
      const int m0 = tm * BM;
      const int n0 = (second ? nn - p.tiles_n : nn) * p.BN;
      const Epi epi = epi_of(p, second);
      mbar_wait(acc_full_bar(buf), (i >> 1) & 1);
      tc_fence_after();
      const int m = m0 + q * 32 + lane;
      const bool row_ok = m < p.M;
      const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(buf * p.BN);
      const int half = (warp - 2) >> 2;          // the two warps of a lane quarter alternate 32-column chunks
      const int last_c = ((n_chunks - 1 - half) & ~1) + half;
      if (last_c < 0) {                          // BN = 32: the second warp of the quarter has nothing to drain
        if (lane == 0) mbar_arrive(acc_empty_bar(buf));
        if (splits > 1) {                        // still part of the split-K hand-shake below
          asm volatile("bar.sync 1, 256;" ::: "memory");
          asm volatile("bar.sync 1, 256;" ::: "memory");
        }
        continue;
      }
      for (int c = half; c < n_chunks; c += 2) {
        uint32_t v[32];
        tmem_ld32(t_row + static_cast<uint32_t>(c * 32), v);
        tmem_ld_wait();
        if (c == last_c) {
          // every lane's TMEM reads are complete: hand the accumulator back before the stores
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(acc_empty_bar(buf));
        }
        if (splits == 1) {
          if (row_ok) epilogue_chunk<F16>(epi, m, n0 + c * 32, v);
        } else if (row_ok) {
          // this split's partial sums, in its own slice of the workspace (plain stores: the final sum below runs over
          // the slices in a fixed order, so the result does not depend on which CTA finishes when)
          float4* wsp = reinterpret_cast<float4*>(p.ws + (static_cast<size_t>(w - t * splits) * (p.tiles_m * BM) + m) * p.ldw + n0 + c * 32);
#pragma unroll
          for (int j = 0; j < 8; ++j)
            __stcg(wsp + j, make_float4(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]), __uint_as_float(v[4 * j + 2]),
                                        __uint_as_float(v[4 * j + 3])));
        }
      }
      if (splits > 1) {
        // the split that arrives last on the tile's counter owns the epilogue: it adds the slices up (read past L1,
        // split 0 first), writes the layer's output and puts the counter back to zero for the next launch
        __threadfence();
        asm volatile("bar.sync 1, 256;" ::: "memory");
        if (threadIdx.x == 64) {
          const int arrived = atomicAdd(p.ctr + t, 1);
          split_last = arrived == splits - 1;
          if (arrived == splits - 1) p.ctr[t] = 0;
        }
        asm volatile("bar.sync 1, 256;" ::: "memory");
        if (*reinterpret_cast<volatile int*>(&split_last)) {
          __threadfence();
          for (int c = half; c < n_chunks; c += 2) {
            if (!row_ok) continue;
            float acc[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) acc[j] = 0.f;
            for (int sp = 0; sp < splits; ++sp) {
              const float4* wsp = reinterpret_cast<const float4*>(p.ws + (static_cast<size_t>(sp) * (p.tiles_m * BM) + m) * p.ldw + n0 + c * 32);
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                const float4 f = __ldcg(wsp + j);
                acc[4 * j] += f.x; acc[4 * j + 1] += f.y; acc[4 * j + 2] += f.z; acc[4 * j + 3] += f.w;
              }
            }
            uint32_t v[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = __float_as_uint(acc[j]);
            epilogue_chunk<F16>(epi, m, n0 + c * 32, v);
          }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, p.tmem_cols);
  }
}

// ------------------------------------------------------------------------------------------------
// CTA-pair variant (cta_group::2): two CTAs on the two SMs of a TPC compute one 256 x BN tile.
// Each CTA loads its own 128 rows of A and HALF of the weight tile; one thread of the leader CTA
// issues M=256 MMAs that read both halves, so per SM the operand fetch per K step drops from
// 4 KB + 32*BN B to 4 KB + 16*BN B.  Barriers: the leader's `full` barrier collects the TMA bytes of
// both CTAs (peer bit cleared in the barrier address), tcgen05.commit is multicast to both CTAs'
// `empty` / `acc_full` barriers, and both CTAs' epilogue warps arrive on the leader's `acc_empty`.
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
constexpr uint32_t kPeerBitMask = 0xFEFFFFFFu;  // shared::cluster address of the same offset in the even CTA
__device__ __forceinline__ void tma_load_2d_2sm(uint32_t dst, const CUtensorMap* tm, uint32_t bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes "
      "[%0], [%1, {%3, %4}], [%2];" ::"r"(dst), "l"(tm), "r"(bar & kPeerBitMask), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_load_im2col_4d_2sm(uint32_t dst, const CUtensorMap* tm, uint32_t bar, int c,
                                                       int w, int h, int n, uint16_t off_w, uint16_t off_h) {
  asm volatile(
      "cp.async.bulk.tensor.4d.im2col.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes "
      "[%0], [%1, {%3, %4, %5, %6}], [%2], {%7, %8};" ::"r"(dst), "l"(tm), "r"(bar & kPeerBitMask), "r"(c), "r"(w),
      "r"(h), "r"(n), "h"(off_w), "h"(off_h)
      : "memory");
}
__device__ __forceinline__ void umma_bf16_2sm(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                              uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred pe, p;\n\t"
      "elect.sync _|pe, 0xffffffff;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "@pe tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit_2sm(uint32_t bar) {
  asm volatile(
      "{\n\t.reg .pred pe;\n\t"
      "elect.sync _|pe, 0xffffffff;\n\t"
      "@pe tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;\n\t}" ::"r"(bar),
      "h"(static_cast<uint16_t>(3))
      : "memory");
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t bar, uint32_t cta) {
  asm volatile(
      "{\n\t.reg .b32 ra;\n\t"
      "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
      "mbarrier.arrive.shared::cluster.b64 _, [ra];\n\t}" ::"r"(bar),
      "r"(cta)
      : "memory");
}

// TMA issue: an im2col-mode request takes 700-950 cycles from issue to completion whatever its size, and the
// requests of ONE issuing warp do not overlap each other, but requests of different warps do (tools/microbench/
// tma_rate.cu, two-issuer rows: the per-SM rate doubles with two warps or two CTAs).  So the k-blocks are dealt
// round-robin to `nprod` producer warps (warp 0 and warps 10 ..), each walking the same tile / stage sequence
// and issuing every nprod-th request: with three of them a request completes every ~300 cycles, around the MMA
// time of a k-block (256 cycles at tile-N 128, 512 at 256).
// -DB2I_PAIR_TRACE compiles clock stamps into the pair kernel (B2I_PAIR_TRACE=1 at run time prints them per launch)
#ifdef B2I_PAIR_TRACE
constexpr bool kPairTrace = true;
#else
constexpr bool kPairTrace = false;
#endif
constexpr int PAIR_MAX_PROD = 3;   // 384 threads: one more warp would push the epilogue warps under the registers they need (spills)
constexpr int PAIR_THREADS = PTHREADS + 32 * (PAIR_MAX_PROD - 1);
template <int AMODE, bool F16>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(PAIR_THREADS, 1)
conv_umma_pair_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                      const __grid_constant__ CUtensorMap tmB2, const KP p) {
  const long long t_start = kPairTrace ? clock64() : 0;
  griddep_launch();
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bars[2 * MAX_STAGES + 5];
  __shared__ uint32_t tmem_slot;
  __shared__ __align__(16) float bias_s[1024];   // bias of the layer (and of a riding downsample): constants, staged before the wait for the previous grid

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);   // warp-uniform, and the compiler knows it (MMA operands stay in uniform registers)
  const int lane = threadIdx.x & 31;
  const uint32_t rank = __shfl_sync(0xffffffffu, cluster_ctarank(), 0);
  const bool leader = rank == 0;
  const int cluster_id = blockIdx.x >> 1;
  const int num_clusters = gridDim.x >> 1;
  const int half_n = p.BN >> 1;
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t b_tile_bytes = static_cast<uint32_t>(half_n) * 128u;
  // Resident-weights mode (one N-tile per problem, e.g. ResNet layer2): this CTA's half of every weight
  // k-block is loaded once, in front of the A ring, and only the A operand streams.  The MMA then reads
  // 6 KB per K step and TMA writes 4 KB, instead of 8 KB + 8 KB for the one-CTA 128-wide tile.
  const bool resb = p.resb != 0;
  const uint32_t resb_bytes = resb ? static_cast<uint32_t>(p.num_kb + (p.tiles_n2 ? p.num_kb2 : 0)) * b_tile_bytes : 0u;
  // msub = 2: each CTA owns TWO 128-row sub-tiles (a 512-row tile per pair).  An im2col-mode TMA request costs
  // about 700 cycles whatever its size (tools/microbench/tma_rate.cu), more than the MMAs of one 128-row k-block
  // (512 / 256 cycles for N = 256 / 128): one 256-pixel request per k-block feeds twice the MMAs.
  const int msub = p.msub;
  const uint32_t a_bytes = static_cast<uint32_t>(msub) * A_TILE_BYTES;
  const uint32_t stage_bytes = resb ? a_bytes : a_bytes + b_tile_bytes;
  const int rows_cta = msub * BM;                       // output rows per CTA and tile
  const int acc_cols = msub * p.BN;                     // TMEM columns of one tile's accumulators
  const bool acc2 = 2 * acc_cols <= 512;                // two tiles' accumulators fit: epilogue overlaps the next tile
  const uint32_t ring_base = smem_base + resb_bytes;
  const uint32_t bar0 = smem_u32(bars);
  auto full_bar = [&](int s) { return bar0 + 8u * s; };
  auto empty_bar = [&](int s) { return bar0 + 8u * (MAX_STAGES + s); };
  auto acc_full_bar = [&](int b) { return bar0 + 8u * (2 * MAX_STAGES + b); };
  auto acc_empty_bar = [&](int b) { return bar0 + 8u * (2 * MAX_STAGES + 2 + b); };
  const uint32_t b_full_bar = bar0 + 8u * (2 * MAX_STAGES + 4);
  const int tn_all = p.tiles_n + p.tiles_n2;
  const int total_tiles = p.tiles_m * tn_all;       // tiles_m counts pair tiles (2 * msub * 128 rows) here

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(full_bar(s), 2);    // leader's expect_tx arrive + the peer's arrive
      mbar_init(empty_bar(s), 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(acc_full_bar(b), 1);
      mbar_init(acc_empty_bar(b), 16);  // 8 epilogue warps in each of the two CTAs
    }
    mbar_init(b_full_bar, 2);
    mbar_fence_init();
  }
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  const int nbias1 = p.tiles_n * p.BN, nbias2 = p.tiles_n2 * p.BN;
  const bool bias_staged = nbias1 + nbias2 <= 1024;
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)),
                 "r"(p.tmem_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();   // both CTAs' barriers are initialised before any remote arrive / multicast commit
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, tmem_slot, 0);
  if (resb && warp == 0 && lane == 0) {   // weights do not depend on the previous kernel: fetch them before waiting for it
    if (leader) mbar_expect_tx(b_full_bar, 2u * resb_bytes);
    else mbar_arrive_cluster(b_full_bar, 0);
    const int nrow = static_cast<int>(rank) * half_n;
    for (int kb = 0; kb < p.num_kb; ++kb) tma_load_2d_2sm(smem_base + kb * b_tile_bytes, &tmB, b_full_bar, kb * BK, nrow);
    if (p.tiles_n2)
      for (int j = 0; j < p.num_kb2; ++j)
        tma_load_2d_2sm(smem_base + (p.num_kb + j) * b_tile_bytes, &tmB2, b_full_bar, j * BK, nrow);
  }
  griddep_wait();   // everything above overlaps the previous kernel's tail; activations are read below
  if (kPairTrace && p.trace != nullptr && static_cast<int>(blockIdx.x) == p.trace_cta && threadIdx.x == 0) { p.trace[8000] = t_start; p.trace[8001] = clock64(); }

  const int prod_id = warp == 0 ? 0 : (warp >= PTHREADS / 32 ? warp - PTHREADS / 32 + 1 : -1);
  if (prod_id >= 0) {
    if (lane == 0 && prod_id < p.nprod) {
      const int cpb = (AMODE == A_IM2COL) ? (p.C >> 6) : 1;
      const int ohw = p.OH * p.OW;
      int s = 0, ph = 0;
      int turn = 0;   // whose k-block the next one is
      int g = 0;      // k-block sequence number (trace)
      const bool tr = kPairTrace && p.trace != nullptr && static_cast<int>(blockIdx.x) == p.trace_cta;
      for (int t0 = cluster_id; t0 < total_tiles; t0 += num_clusters) {
        const int t = p.rev ? total_tiles - 1 - t0 : t0;   // snake order: start with what the producer wrote last
        const int tm = t / tn_all;
        const int nn = t - tm * tn_all;
        const bool second = nn >= p.tiles_n;
        const int m0 = (tm * 2 + static_cast<int>(rank)) * rows_cta;
        const int n0 = (second ? nn - p.tiles_n : nn) * p.BN + static_cast<int>(rank) * half_n;
        const int kb_begin = second ? p.kb_first2 : 0;
        const int kb_count = second ? p.num_kb2 : p.num_kb;
        int iw_base = 0, ih_base = 0, n_base = 0;
        if (AMODE == A_IM2COL) {
          const int mc = m0 < p.M ? m0 : p.M;   // a CTA whose rows all lie past the batch still loads (zeros): start it just past the last image
          n_base = mc / ohw;
          const int rem = mc - n_base * ohw;
          const int oh = rem / p.OW;
          const int ow = rem - oh * p.OW;
          ih_base = oh * p.stride - p.pad;
          iw_base = ow * p.stride - p.pad;
        }
        const int tap0 = kb_begin / cpb;
        int cblk = kb_begin - tap0 * cpb, r = tap0 / p.S, sx = tap0 - (tap0 / p.S) * p.S;
        for (int j = 0; j < kb_count; ++j) {
          const int kb = kb_begin + j;
          const bool mine = turn == prod_id;
          if (++turn == p.nprod) turn = 0;
          if (mine) {
            if (tr) p.trace[4096 + g * 4] = clock64();
            mbar_wait(empty_bar(s), ph ^ 1);
            if (tr) p.trace[4096 + g * 4 + 1] = clock64();
            const uint32_t a_s = ring_base + s * stage_bytes;
            if (leader) mbar_expect_tx(full_bar(s), 2u * stage_bytes);
            else mbar_arrive_cluster(full_bar(s), 0);
            if (AMODE == A_IM2COL) {
              tma_load_im2col_4d_2sm(a_s, &tmA, full_bar(s), cblk << 6, iw_base, ih_base, n_base,
                                     static_cast<uint16_t>(sx), static_cast<uint16_t>(r));
            } else {
              tma_load_2d_2sm(a_s, &tmA, full_bar(s), kb * BK, m0);
            }
            if (!resb) {
              if (second) tma_load_2d_2sm(a_s + a_bytes, &tmB2, full_bar(s), j * BK, n0);
              else tma_load_2d_2sm(a_s + a_bytes, &tmB, full_bar(s), kb * BK, n0);
            }
            if (tr) { p.trace[4096 + g * 4 + 2] = clock64(); p.trace[4096 + g * 4 + 3] = prod_id; }
          }
          ++g;
          if (AMODE == A_IM2COL && ++cblk == cpb) {
            cblk = 0;
            if (++sx == p.S) { sx = 0; ++r; }
          }
          if (++s == p.stages) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (leader) {
      const uint32_t idesc = umma_idesc_bf16(256, p.BN, F16 ? 1 : 0);
      if (resb) mbar_wait(b_full_bar, 0);
      int s = 0, ph = 0, i = 0;
      int g = 0;
      const bool tr = kPairTrace && p.trace != nullptr && static_cast<int>(blockIdx.x) == p.trace_cta && lane == 0;
      for (int t0 = cluster_id; t0 < total_tiles; t0 += num_clusters, ++i) {
        const int t = p.rev ? total_tiles - 1 - t0 : t0;
        const int buf = acc2 ? (i & 1) : 0;
        const int use = acc2 ? (i >> 1) : i;          // how often this accumulator buffer has been used before
        if (tr) p.trace[g * 4 + 3] = clock64();
        mbar_wait(acc_empty_bar(buf), (use & 1) ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(buf * acc_cols);
        const bool second = (t % tn_all) >= p.tiles_n;
        const int kb_count = second ? p.num_kb2 : p.num_kb;
        for (int kb = 0; kb < kb_count; ++kb) {
          if (tr) p.trace[g * 4] = clock64();
          mbar_wait(full_bar(s), ph);
          if (tr) p.trace[g * 4 + 1] = clock64();
          tc_fence_after();
          {
            const uint32_t a_s = ring_base + s * stage_bytes;
            const uint64_t adesc = umma_smem_desc_sw128(a_s);
            const uint64_t bdesc = umma_smem_desc_sw128(resb ? smem_base + static_cast<uint32_t>((second ? p.num_kb : 0) + kb) * b_tile_bytes
                                                             : a_s + a_bytes);
            for (int sb = 0; sb < msub; ++sb) {        // the sub-tiles share the weight tile
              const uint64_t ad = adesc + static_cast<uint64_t>(sb * (A_TILE_BYTES >> 4));
              const uint32_t dt = d_tmem + static_cast<uint32_t>(sb * p.BN);
#pragma unroll
              for (int k = 0; k < BK / 16; ++k)
                umma_bf16_2sm(dt, ad + 2u * k, bdesc + 2u * k, idesc, (kb | k) != 0 ? 1u : 0u);
            }
            umma_commit_2sm(empty_bar(s));
            if (kb == kb_count - 1) umma_commit_2sm(acc_full_bar(buf));
          }
          if (tr) p.trace[g * 4 + 2] = clock64();
          ++g;
          if (++s == p.stages) { s = 0; ph ^= 1; }
        }
      }
    }
  } else {
    const int q = warp & 3;
    const int half = (warp - 2) >> 2;
    const int n_chunks = p.BN >> 5;
    const int last_c = ((n_chunks - 1 - half) & ~1) + half;
    if (bias_staged) {   // by the epilogue warps alone, while the first tile's operands are on their way
      for (int i = threadIdx.x - 64; i < nbias1; i += 256) bias_s[i] = __ldg(p.bias + i);
      for (int i = threadIdx.x - 64; i < nbias2; i += 256) bias_s[nbias1 + i] = __ldg(p.bias2 + i);
      asm volatile("bar.sync 1, 256;" ::: "memory");
    }
    int i = 0;
    for (int t0 = cluster_id; t0 < total_tiles; t0 += num_clusters, ++i) {
      const int t = p.rev ? total_tiles - 1 - t0 : t0;
      const int buf = acc2 ? (i & 1) : 0;
      const int use = acc2 ? (i >> 1) : i;
      const int tm = t / tn_all;
      const int nn = t - tm * tn_all;
      const bool second = nn >= p.tiles_n;
      const int m0 = (tm * 2 + static_cast<int>(rank)) * rows_cta;
      const int n0 = (second ? nn - p.tiles_n : nn) * p.BN;
      Epi epi = epi_of(p, second);
      if (bias_staged) epi.bias_smem = smem_u32(bias_s) + (second ? static_cast<uint32_t>(nbias1) * 4u : 0u);
      // the residual of the first sub-tile is fetched before the accumulator is waited for
      int m = m0 + q * 32 + lane;
      bool row_ok = m < p.M;
      bool pre = row_ok && epi.residual != nullptr && !epi.out_fp32;
      uint32_t rr[4][16];
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (pre && half + 2 * j < n_chunks && n0 + (half + 2 * j) * 32 < epi.Cout)
          prefetch_residual(epi, m, n0 + (half + 2 * j) * 32, rr[j]);
      mbar_wait(acc_full_bar(buf), use & 1);
      tc_fence_after();
      for (int sb = 0; sb < msub; ++sb) {
        if (sb > 0) {
          m = m0 + sb * BM + q * 32 + lane;
          row_ok = m < p.M;
          pre = false;
        }
        const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(buf * acc_cols + sb * p.BN);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int c = half + 2 * j;
          if (c >= n_chunks) break;
          uint32_t v[32];
          const bool tre = kPairTrace && p.trace != nullptr && static_cast<int>(blockIdx.x) == p.trace_cta && threadIdx.x == 64 && i < 4;
          if (tre) p.trace[6000 + (i * 4 + j) * 4] = clock64();
          tmem_ld32(t_row + static_cast<uint32_t>(c * 32), v);
          tmem_ld_wait();
          if (tre) p.trace[6000 + (i * 4 + j) * 4 + 1] = clock64() + (v[0] & 0);
          if (c == last_c && sb == msub - 1) {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
              if (leader) mbar_arrive(acc_empty_bar(buf));
              else mbar_arrive_cluster(acc_empty_bar(buf), 0);
            }
          }
          if (row_ok) epilogue_chunk<F16>(epi, m, n0 + c * 32, v, pre ? &rr[j] : nullptr);
          if (tre) p.trace[6000 + (i * 4 + j) * 4 + 2] = clock64();
        }
      }
    }
  }

  if (kPairTrace && p.trace != nullptr && static_cast<int>(blockIdx.x) == p.trace_cta && threadIdx.x == 64) p.trace[8002] = clock64();   // first epilogue thread done
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();   // the peer may still be arriving on / reading this CTA's barriers and smem
  if (kPairTrace && p.trace != nullptr && static_cast<int>(blockIdx.x) == p.trace_cta && threadIdx.x == 0) p.trace[8003] = clock64();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(p.tmem_cols) : "memory");
  }
}

// ------------------------------------------------------------------------------------------------
// 3x3 / stride 1 / pad 1 convolution over FLATTENED PADDED ROWS, CTA pair, resident weights
// (ResNet layer2: 128 -> 128 channels on 28x28 maps).
//
// The im2col kernels above fetch every input pixel nine times, through im2col-mode TMA requests that cost
// about 700 cycles each whatever their size: with 128-wide tiles that is 2.7x the MMA time of a k-block.
// Here the batch is seen as one long sequence of padded
// rows: row Y = n*(H+1) + y holds image n's row y as `pitch` pixels (pixel 0 = the left zero pad,
// pixels 1..W the row, the rest zero), and row n*(H+1) + H is all zero - the bottom pad of image n
// and the top pad of image n+1 at once.  Flatten those rows and a 3x3 tap (r, s) is a pure shift by
// (r-1)*pitch + (s-1) positions, so the nine A operands of a tile are nine 128-byte-aligned views
// of ONE halo patch (the absolute-address swizzle makes any such view valid, see the halo-patch
// kernel below).  A CTA owns 128 consecutive positions = 128/pitch rows; its patch is those rows
// plus one above and one below, fetched row by row with tiled TMA loads whose out-of-range rows and
// columns arrive as zeros.  Each pixel is read from L2 (rows+2)/rows times instead of nine.
// Positions in pad columns / pad rows are computed and dropped: W*H / (pitch*(H+1)) of the MMA
// work is useful (84 % for 28x28 with pitch 32).
// The two CTAs of a pair (cta_group::2, M = 256) each keep their half of ALL weight k-blocks in
// shared memory (9 * C/64 blocks of BN/2 x 64), so only the patches stream.
struct FlatP {
  int H, W, N, slices, BN, pitch, rpt, box_w;   // rpt = rows per 128-position tile
  int rows_total, total_pairs, nbuf;
  int ldo, relu, Cout, f16, rev;
  uint32_t tmem_cols;
  const float* bias;
  const __nv_bfloat16* residual;
  void* out;
};

__device__ __forceinline__ void tma_load_4d_2sm(uint32_t dst, const CUtensorMap* tm, uint32_t bar, int c0, int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes "
      "[%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(dst), "l"(tm), "r"(bar & kPeerBitMask), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}

constexpr int FLAT_MAXBUF = 4;

template <bool F16>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(PTHREADS, 1)
conv3x3_flat_pair_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmB, const FlatP p) {
  griddep_launch();
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bars[2 * FLAT_MAXBUF + 5];   // full[], empty[], acc_full[2], acc_empty[2], b_full
  __shared__ uint32_t tmem_slot;
  __shared__ __align__(16) float bias_s[256];   // the layer's bias, staged by the epilogue warps (one N tile of at most 256 channels)

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);   // warp-uniform, and the compiler knows it (MMA operands stay in uniform registers)
  const int lane = threadIdx.x & 31;
  const uint32_t rank = __shfl_sync(0xffffffffu, cluster_ctarank(), 0);
  const bool leader = rank == 0;
  const int cluster_id = blockIdx.x >> 1;
  const int num_clusters = gridDim.x >> 1;
  const int half_n = p.BN >> 1;
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t b_tile_bytes = static_cast<uint32_t>(half_n) * 128u;
  const int num_kb = 9 * p.slices;
  const uint32_t resb_bytes = static_cast<uint32_t>(num_kb) * b_tile_bytes;
  const uint32_t row_bytes = static_cast<uint32_t>(p.pitch) * 128u;
  // one 64-channel slice; the view of tap (2, 2) runs two pixels past the last patch row (they only feed
  // dropped pad-column outputs), hence the extra KB that keeps those reads inside the buffer
  const uint32_t patch_bytes = static_cast<uint32_t>(p.rpt + 2) * row_bytes + 1024u;
  const uint32_t tx_bytes = static_cast<uint32_t>(p.rpt + 2) * static_cast<uint32_t>(p.box_w) * 128u;
  const uint32_t ring_base = smem_base + resb_bytes;
  const uint32_t bar0 = smem_u32(bars);
  auto full_bar = [&](int s) { return bar0 + 8u * s; };
  auto empty_bar = [&](int s) { return bar0 + 8u * (FLAT_MAXBUF + s); };
  auto acc_full_bar = [&](int b) { return bar0 + 8u * (2 * FLAT_MAXBUF + b); };
  auto acc_empty_bar = [&](int b) { return bar0 + 8u * (2 * FLAT_MAXBUF + 2 + b); };
  const uint32_t b_full_bar = bar0 + 8u * (2 * FLAT_MAXBUF + 4);

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.nbuf; ++s) {
      mbar_init(full_bar(s), 2);    // leader's expect_tx arrive + the peer's arrive
      mbar_init(empty_bar(s), 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(acc_full_bar(b), 1);
      mbar_init(acc_empty_bar(b), 16);  // 8 epilogue warps in each of the two CTAs
    }
    mbar_init(b_full_bar, 2);
    mbar_fence_init();
  }
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmX);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)),
                 "r"(p.tmem_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, tmem_slot, 0);
  if (warp == 0 && lane == 0) {   // weights do not depend on the previous kernel: fetch them before waiting for it
    if (leader) mbar_expect_tx(b_full_bar, 2u * resb_bytes);
    else mbar_arrive_cluster(b_full_bar, 0);
    const int nrow = static_cast<int>(rank) * half_n;
    for (int kb = 0; kb < num_kb; ++kb) tma_load_2d_2sm(smem_base + kb * b_tile_bytes, &tmB, b_full_bar, kb * BK, nrow);
  }
  griddep_wait();

  const int hp = p.H + 1;   // padded rows per image
  if (warp == 0) {
    if (lane == 0) {
      int s = 0, ph = 0;
      for (int t0 = cluster_id; t0 < p.total_pairs; t0 += num_clusters) {
        const int t = p.rev ? p.total_pairs - 1 - t0 : t0;
        const int y0 = (2 * t + static_cast<int>(rank)) * p.rpt - 1;   // first patch row (one above the tile)
        for (int sl = 0; sl < p.slices; ++sl) {
          mbar_wait(empty_bar(s), ph ^ 1);
          const uint32_t a_s = ring_base + s * patch_bytes;
          if (leader) mbar_expect_tx(full_bar(s), 2u * tx_bytes);
          else mbar_arrive_cluster(full_bar(s), 0);
          for (int j = 0; j < p.rpt + 2; ++j) {
            const int Y = y0 + j;
            int n = -1, y = 0;                       // Y = -1: above the first image, all out of range -> zeros
            if (Y >= 0) { n = Y / hp; y = Y - n * hp; }   // y == H (or n >= N) is out of range too -> the shared zero row
            tma_load_4d_2sm(a_s + j * row_bytes, &tmX, full_bar(s), sl * 64, -1, y, n);
          }
          if (++s == p.nbuf) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (leader) {
      const uint32_t idesc = umma_idesc_bf16(256, p.BN, F16 ? 1 : 0);
      mbar_wait(b_full_bar, 0);
      int s = 0, ph = 0, i = 0;
      for (int t0 = cluster_id; t0 < p.total_pairs; t0 += num_clusters, ++i) {
        const int t = p.rev ? p.total_pairs - 1 - t0 : t0;
        const int buf = i & 1;
        mbar_wait(acc_empty_bar(buf), ((i >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(buf * p.BN);
        for (int sl = 0; sl < p.slices; ++sl) {
          mbar_wait(full_bar(s), ph);
          tc_fence_after();
          const uint32_t a_s = ring_base + s * patch_bytes;
#pragma unroll
          for (int tap = 0; tap < 9; ++tap) {
            const uint64_t adesc = umma_smem_desc_sw128(a_s + static_cast<uint32_t>((tap / 3) * p.pitch + (tap % 3)) * 128u);
            const uint64_t bdesc = umma_smem_desc_sw128(smem_base + static_cast<uint32_t>(tap * p.slices + sl) * b_tile_bytes);
#pragma unroll
            for (int k = 0; k < BK / 16; ++k)
              umma_bf16_2sm(d_tmem, adesc + 2u * k, bdesc + 2u * k, idesc, (sl | tap | k) != 0 ? 1u : 0u);
          }
          umma_commit_2sm(empty_bar(s));
          if (++s == p.nbuf) { s = 0; ph ^= 1; }
        }
        umma_commit_2sm(acc_full_bar(buf));
      }
    }
  } else {
    const int q = warp & 3;
    const int half = (warp - 2) >> 2;
    const int n_chunks = p.BN >> 5;
    const int last_c = ((n_chunks - 1 - half) & ~1) + half;
    for (int i = threadIdx.x - 64; i < p.BN; i += 256) bias_s[i] = __ldg(p.bias + i);
    asm volatile("bar.sync 1, 256;" ::: "memory");
    Epi epi;
    epi.bias = p.bias; epi.bias_smem = smem_u32(bias_s); epi.residual = p.residual; epi.out = p.out; epi.ldo = p.ldo; epi.relu = p.relu; epi.out_fp32 = 0; epi.Cout = p.Cout; epi.f16 = p.f16;
    int i = 0;
    for (int t0 = cluster_id; t0 < p.total_pairs; t0 += num_clusters, ++i) {
      const int t = p.rev ? p.total_pairs - 1 - t0 : t0;
      const int buf = i & 1;
      const int pos = q * 32 + lane;                                  // position inside this CTA's tile
      const int Y = (2 * t + static_cast<int>(rank)) * p.rpt + pos / p.pitch;
      const int xo = pos % p.pitch;
      const int n = Y / hp, y = Y - n * hp;
      const bool row_ok = xo < p.W && y < p.H && n < p.N;
      const int m = (n * p.H + y) * p.W + xo;
      const bool pre = row_ok && epi.residual != nullptr;
      uint32_t rr[4][16];
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (pre && half + 2 * j < n_chunks) prefetch_residual(epi, m, (half + 2 * j) * 32, rr[j]);
      mbar_wait(acc_full_bar(buf), (i >> 1) & 1);
      tc_fence_after();
      const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(buf * p.BN);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int c = half + 2 * j;
        if (c >= n_chunks) break;
        uint32_t v[32];
        tmem_ld32(t_row + static_cast<uint32_t>(c * 32), v);
        tmem_ld_wait();
        if (c == last_c) {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) {
            if (leader) mbar_arrive(acc_empty_bar(buf));
            else mbar_arrive_cluster(acc_empty_bar(buf), 0);
          }
        }
        if (row_ok) epilogue_chunk<F16>(epi, m, c * 32, v, pre ? &rr[j] : nullptr);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(p.tmem_cols) : "memory");
  }
}

// ------------------------------------------------------------------------------------------------
// The same flattened-row form with STREAMED weights and several N tiles (ResNet layer4: 512 -> 512 channels on 7x7
// maps, 4.7 MB of weights).  Why: an SM's TMA unit takes about 1.5 cycles per 128-byte row of a tiled box and more
// per row (pixel) of an im2col box (tools/microbench/l2_rate.cu: 85 B/clk per SM with three issuing warps of tiled
// loads; the im2col pair kernel above runs at 43 B/clk).  With 128-byte-swizzled operands every row is 128 bytes, so a
// k-block of the im2col kernel costs 128 pixel rows + 128 weight rows of TMA work = ~760 cycles against 512 cycles of
// MMAs.  Here a k-block costs 1/9 of a halo patch (16 rows for a 144-pixel patch) + 128 weight rows.
// Row pitch: W + 1 when that is a power of two (7x7 maps: pitch 8: pixel 0 of a row is the left zero pad and doubles
// as the right pad of the row above; 49 of every 64 positions are real outputs), else conv_flat_pitch(W).
// Warps: 0 = patch loads, 1 = MMA issue (leader CTA), 2..9 = epilogue, 10..11 = weight loads (k-blocks dealt in turn).
struct FlatSP {
  int H, W, N, slices, BN, pitch, rpt, box_w, tiles_n;
  int total_pairs, na, nb;
  int ldo, relu, Cout, f16, rev;
  uint32_t tmem_cols;
  const float* bias;
  const __nv_bfloat16* residual;
  void* out;
};

constexpr int FLATS_MAXA = 4, FLATS_MAXB = 8;
constexpr int FLATS_THREADS = PTHREADS + 64;

template <bool F16>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(FLATS_THREADS, 1)
conv3x3_flat_stream_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmB, const FlatSP p) {
  griddep_launch();
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bars[2 * FLATS_MAXA + 2 * FLATS_MAXB + 4];   // a_full[], a_empty[], b_full[], b_empty[], acc_full[2], acc_empty[2]
  __shared__ uint32_t tmem_slot;

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);   // warp-uniform, and the compiler knows it (MMA operands stay in uniform registers)
  const int lane = threadIdx.x & 31;
  const uint32_t rank = __shfl_sync(0xffffffffu, cluster_ctarank(), 0);
  const bool leader = rank == 0;
  const int cluster_id = blockIdx.x >> 1;
  const int num_clusters = gridDim.x >> 1;
  const int half_n = p.BN >> 1;
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t b_tile_bytes = static_cast<uint32_t>(half_n) * 128u;
  const uint32_t row_bytes = static_cast<uint32_t>(p.pitch) * 128u;
  // one 64-channel slice of the halo patch; the view of tap (2, 2) runs two pixels past the last patch row (they only
  // feed dropped pad outputs), hence the extra KB that keeps those reads inside the buffer
  const uint32_t patch_bytes = ((static_cast<uint32_t>(p.rpt + 2) * row_bytes + 1024u) + 1023u) & ~1023u;
  const uint32_t tx_bytes = static_cast<uint32_t>(p.rpt + 2) * static_cast<uint32_t>(p.box_w) * 128u;
  const uint32_t a_base = smem_base;
  const uint32_t b_base = smem_base + static_cast<uint32_t>(p.na) * patch_bytes;
  const uint32_t bar0 = smem_u32(bars);
  auto a_full = [&](int s) { return bar0 + 8u * s; };
  auto a_empty = [&](int s) { return bar0 + 8u * (FLATS_MAXA + s); };
  auto b_full = [&](int s) { return bar0 + 8u * (2 * FLATS_MAXA + s); };
  auto b_empty = [&](int s) { return bar0 + 8u * (2 * FLATS_MAXA + FLATS_MAXB + s); };
  auto acc_full_bar = [&](int b) { return bar0 + 8u * (2 * FLATS_MAXA + 2 * FLATS_MAXB + b); };
  auto acc_empty_bar = [&](int b) { return bar0 + 8u * (2 * FLATS_MAXA + 2 * FLATS_MAXB + 2 + b); };
  const int total_items = p.total_pairs * p.tiles_n;

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.na; ++s) {
      mbar_init(a_full(s), 2);    // leader's expect_tx arrive + the peer's arrive
      mbar_init(a_empty(s), 1);
    }
    for (int s = 0; s < p.nb; ++s) {
      mbar_init(b_full(s), 2);
      mbar_init(b_empty(s), 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(acc_full_bar(b), 1);
      mbar_init(acc_empty_bar(b), 16);  // 8 epilogue warps in each of the two CTAs
    }
    mbar_fence_init();
  }
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmX);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)),
                 "r"(p.tmem_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, tmem_slot, 0);
  const int hp = p.H + 1;   // padded rows per image

  if (warp >= PTHREADS / 32) {
    // weight k-blocks: they do not depend on the previous kernel, so the first ring-full goes out before the wait
    if (lane == 0) {
      const int me = warp - PTHREADS / 32;
      int s = 0, ph = 0, turn = 0;
      for (int w0 = cluster_id; w0 < total_items; w0 += num_clusters) {
        const int w = p.rev ? total_items - 1 - w0 : w0;
        const int nn = w % p.tiles_n;
        const int nrow = nn * p.BN + static_cast<int>(rank) * half_n;
        for (int sl = 0; sl < p.slices; ++sl) {
          for (int tap = 0; tap < 9; ++tap) {
            const bool mine = turn == me;
            turn ^= 1;
            if (mine) {
              mbar_wait(b_empty(s), ph ^ 1);
              if (leader) mbar_expect_tx(b_full(s), 2u * b_tile_bytes);
              else mbar_arrive_cluster(b_full(s), 0);
              tma_load_2d_2sm(b_base + s * b_tile_bytes, &tmB, b_full(s), (tap * p.slices + sl) * BK, nrow);
            }
            if (++s == p.nb) { s = 0; ph ^= 1; }
          }
        }
      }
    }
  } else if (warp == 0) {
    griddep_wait();   // activations of the previous kernel are read from here on
    if (lane == 0) {
      int s = 0, ph = 0;
      for (int w0 = cluster_id; w0 < total_items; w0 += num_clusters) {
        const int w = p.rev ? total_items - 1 - w0 : w0;
        const int t = w / p.tiles_n;
        const int y0 = (2 * t + static_cast<int>(rank)) * p.rpt - 1;   // first patch row (one above the tile)
        for (int sl = 0; sl < p.slices; ++sl) {
          mbar_wait(a_empty(s), ph ^ 1);
          const uint32_t a_s = a_base + s * patch_bytes;
          if (leader) mbar_expect_tx(a_full(s), 2u * tx_bytes);
          else mbar_arrive_cluster(a_full(s), 0);
          for (int j = 0; j < p.rpt + 2; ++j) {
            const int Y = y0 + j;
            int n = -1, y = 0;                       // Y = -1: above the first image, all out of range -> zeros
            if (Y >= 0) { n = Y / hp; y = Y - n * hp; }   // y == H (or n >= N) is out of range too -> the shared zero row
            tma_load_4d_2sm(a_s + j * row_bytes, &tmX, a_full(s), sl * 64, -1, y, n);
          }
          if (++s == p.na) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (leader) {
      const uint32_t idesc = umma_idesc_bf16(256, p.BN, F16 ? 1 : 0);
      int sa = 0, pha = 0, sb = 0, phb = 0, i = 0;
      for (int w0 = cluster_id; w0 < total_items; w0 += num_clusters, ++i) {
        const int buf = i & 1;
        mbar_wait(acc_empty_bar(buf), ((i >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(buf * p.BN);
        for (int sl = 0; sl < p.slices; ++sl) {
          mbar_wait(a_full(sa), pha);
          tc_fence_after();
          const uint32_t a_s = a_base + sa * patch_bytes;
#pragma unroll 1
          for (int tap = 0; tap < 9; ++tap) {
            mbar_wait(b_full(sb), phb);
            tc_fence_after();
            const uint64_t adesc = umma_smem_desc_sw128(a_s + static_cast<uint32_t>((tap / 3) * p.pitch + (tap % 3)) * 128u);
            const uint64_t bdesc = umma_smem_desc_sw128(b_base + sb * b_tile_bytes);
#pragma unroll
            for (int k = 0; k < BK / 16; ++k)
              umma_bf16_2sm(d_tmem, adesc + 2u * k, bdesc + 2u * k, idesc, (sl | tap | k) != 0 ? 1u : 0u);
            umma_commit_2sm(b_empty(sb));
            if (++sb == p.nb) { sb = 0; phb ^= 1; }
          }
          umma_commit_2sm(a_empty(sa));
          if (++sa == p.na) { sa = 0; pha ^= 1; }
        }
        umma_commit_2sm(acc_full_bar(buf));
      }
    }
  } else {
    griddep_wait();   // the residual is an earlier kernel's output
    const int q = warp & 3;
    const int half = (warp - 2) >> 2;
    const int n_chunks = p.BN >> 5;
    const int last_c = ((n_chunks - 1 - half) & ~1) + half;
    Epi epi;
    epi.bias = p.bias; epi.bias_smem = 0; epi.residual = p.residual; epi.out = p.out; epi.ldo = p.ldo; epi.relu = p.relu; epi.out_fp32 = 0; epi.Cout = p.Cout; epi.f16 = p.f16;
    int i = 0;
    for (int w0 = cluster_id; w0 < total_items; w0 += num_clusters, ++i) {
      const int w = p.rev ? total_items - 1 - w0 : w0;
      const int t = w / p.tiles_n;
      const int n0 = (w % p.tiles_n) * p.BN;
      const int buf = i & 1;
      const int pos = q * 32 + lane;                                  // position inside this CTA's tile
      const int Y = (2 * t + static_cast<int>(rank)) * p.rpt + pos / p.pitch;
      const int xo = pos % p.pitch;
      const int n = Y / hp, y = Y - n * hp;
      const bool row_ok = xo < p.W && y < p.H && n < p.N;
      const int m = (n * p.H + y) * p.W + xo;
      const bool pre = row_ok && epi.residual != nullptr;
      uint32_t rr[4][16];
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (pre && half + 2 * j < n_chunks) prefetch_residual(epi, m, n0 + (half + 2 * j) * 32, rr[j]);
      mbar_wait(acc_full_bar(buf), (i >> 1) & 1);
      tc_fence_after();
      const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(buf * p.BN);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int c = half + 2 * j;
        if (c >= n_chunks) break;
        uint32_t v[32];
        tmem_ld32(t_row + static_cast<uint32_t>(c * 32), v);
        tmem_ld_wait();
        if (c == last_c) {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) {
            if (leader) mbar_arrive(acc_empty_bar(buf));
            else mbar_arrive_cluster(acc_empty_bar(buf), 0);
          }
        }
        if (row_ok) epilogue_chunk<F16>(epi, m, n0 + c * 32, v, pre ? &rr[j] : nullptr);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(p.tmem_cols) : "memory");
  }
}

// ------------------------------------------------------------------------------------------------
// 3x3 / stride 1 / pad 1 convolution with 64 input and 64 output channels (ResNet layer1), reading
// the nine filter taps as SHIFTED VIEWS of one halo patch instead of nine im2col loads.
//
// A CTA owns 16 rows x 8 columns of output pixels.  One TMA tiled load brings the 18 x 10 pixel
// halo patch (128 bytes = 64 channels per pixel, 128-byte swizzle, zero fill outside the image =
// the conv's padding).  For tap (r, s) the A operand of the MMA is the patch starting at pixel
// (r, s): 8-row core-matrix groups are the 8 pixels of one output row (8 x 128 B contiguous), the
// next group is the next patch row (SBO = 10 x 128 B).  The nine 64x64 weight tiles stay resident in
// shared memory for the life of the CTA.  Per tile this moves 23 KB through L2 instead of 216 KB.
constexpr int PT_W = 10, PT_H = 18;
constexpr uint32_t PT_BYTES = ((PT_W * PT_H * 128 + 1023) / 1024) * 1024;   // 23552
constexpr uint32_t PT_WBYTES = 9 * 64 * 128;                                // resident weights, 73728
constexpr int PT_BUFS = 3;   // patches (+ residual tiles) in flight

struct PatchP {
  int N, H, W, tiles_w, tiles_h, total_tiles, relu, f16, rev;
  const float* bias;
  const __nv_bfloat16* residual;
  __nv_bfloat16* out;
  long long* trace;   // -DB2I_PAIR_TRACE + B2I_PATCH_TRACE=1: clock stamps of CTA 0
};

__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap* tm, uint32_t bar, int c0,
                                            int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes "
      "[%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(dst),
      "l"(tm), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}

constexpr int PATCH_THREADS = 64 + 256;   // TMA warp, MMA warp, 8 epilogue warps (two per TMEM lane quarter)
constexpr uint32_t PT_TILE_BYTES = 128 * 128;   // one 16x8-pixel x 64-channel bf16 tile (residual in / result out)

__device__ __forceinline__ void tma_store_4d(const CUtensorMap* tm, uint32_t src, int c0, int c1, int c2, int c3) {
  asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];" ::"l"(tm),
               "r"(src), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
               : "memory");
}

// Epilogue without per-thread global traffic: the residual tile arrives by TMA next to the patch,
// the result leaves by one TMA store from a 128B-swizzled staging tile (rows past the image edge
// are clipped by the TMA unit).
template <bool F16>
__global__ void __launch_bounds__(PATCH_THREADS, 1)
conv3x3_c64_patch_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmW,
                         const __grid_constant__ CUtensorMap tmR, const __grid_constant__ CUtensorMap tmO,
                         const PatchP p) {
  const long long t_entry = kPairTrace ? clock64() : 0;
  griddep_launch();
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bars[3 * PT_BUFS + 5];  // patch_full[], patch_empty[], res_empty[], acc_full[2], acc_empty[2], w_full
  __shared__ uint32_t tmem_slot;

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);   // warp-uniform, and the compiler knows it (MMA operands stay in uniform registers)
  const int lane = threadIdx.x & 31;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t w_smem = base;
  const uint32_t patch0 = base + PT_WBYTES;
  const uint32_t res0 = patch0 + PT_BUFS * PT_BYTES;
  const uint32_t out0 = res0 + PT_BUFS * PT_TILE_BYTES;
  const uint32_t bar0 = smem_u32(bars);
  auto patch_full = [&](int b) { return bar0 + 8u * b; };
  auto patch_empty = [&](int b) { return bar0 + 8u * (PT_BUFS + b); };
  auto res_empty = [&](int b) { return bar0 + 8u * (2 * PT_BUFS + b); };
  auto acc_full = [&](int b) { return bar0 + 8u * (3 * PT_BUFS + b); };
  auto acc_empty = [&](int b) { return bar0 + 8u * (3 * PT_BUFS + 2 + b); };
  const uint32_t w_full = bar0 + 8u * (3 * PT_BUFS + 4);
  const bool has_res = p.residual != nullptr;

  if (threadIdx.x == 0) {
    for (int b = 0; b < PT_BUFS; ++b) {
      mbar_init(patch_full(b), 1);
      mbar_init(patch_empty(b), 1);
      mbar_init(res_empty(b), 8);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(acc_full(b), 1);
      mbar_init(acc_empty(b), 8);
    }
    mbar_init(w_full, 1);
    mbar_fence_init();
  }
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmX);
    tma_prefetch_desc(&tmW);
    tma_prefetch_desc(&tmO);
    if (has_res) tma_prefetch_desc(&tmR);
  }
  if (warp == 1) tmem_alloc(smem_u32(&tmem_slot), 128);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, tmem_slot, 0);
  if (warp == 0 && lane == 0) {   // the weights do not depend on the previous kernel: fetch them before waiting for it
    mbar_expect_tx(w_full, PT_WBYTES);
    for (int tap = 0; tap < 9; ++tap) tma_load_2d(w_smem + tap * 8192u, &tmW, w_full, tap * 64, 0);
  }
  griddep_wait();   // everything above overlaps the previous kernel's tail; activations are read below
  if (kPairTrace && p.trace != nullptr && blockIdx.x == 0 && threadIdx.x == 0) { p.trace[1590] = t_entry; p.trace[1591] = clock64(); }
  const int tiles_per_img = p.tiles_w * p.tiles_h;

  if (warp == 0) {
    if (lane == 0) {
      int i = 0;
      for (int t = blockIdx.x; t < p.total_tiles; t += gridDim.x, ++i) {
        const int pb = i % PT_BUFS;
        const uint32_t par = ((i / PT_BUFS) & 1) ^ 1;
        const int tt = p.rev ? p.total_tiles - 1 - t : t;   // snake order: start with what the producer wrote last
        const int n = tt / tiles_per_img;
        const int rem = tt - n * tiles_per_img;
        const int ty = rem / p.tiles_w, tx = rem - ty * p.tiles_w;
        mbar_wait(patch_empty(pb), par);
        if (has_res) mbar_wait(res_empty(pb), par);
        mbar_expect_tx(patch_full(pb), PT_W * PT_H * 128 + (has_res ? PT_TILE_BYTES : 0u));
        tma_load_4d(patch0 + pb * PT_BYTES, &tmX, patch_full(pb), 0, tx * 8 - 1, ty * 16 - 1, n);
        if (has_res) tma_load_4d(res0 + pb * PT_TILE_BYTES, &tmR, patch_full(pb), 0, tx * 8, ty * 16, n);
      }
    }
  } else if (warp == 1) {
    const uint32_t idesc = umma_idesc_bf16(128, 64, F16 ? 1 : 0);
    mbar_wait(w_full, 0);
    int i = 0;
    for (int t = blockIdx.x; t < p.total_tiles; t += gridDim.x, ++i) {
      const int b = i & 1;
      const int pbuf = i % PT_BUFS;
      const bool tr = kPairTrace && p.trace != nullptr && blockIdx.x == 0 && lane == 0 && i < 200;
      if (tr) p.trace[i * 8] = clock64();
      mbar_wait(acc_empty(b), ((i >> 1) & 1) ^ 1);
      if (tr) p.trace[i * 8 + 1] = clock64();
      mbar_wait(patch_full(pbuf), (i / PT_BUFS) & 1);
      if (tr) p.trace[i * 8 + 2] = clock64();
      tc_fence_after();
      {
        const uint32_t pb = patch0 + pbuf * PT_BYTES;
        const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(b * 64);
#pragma unroll
        for (int tap = 0; tap < 9; ++tap) {
          const uint32_t a_addr = pb + static_cast<uint32_t>(((tap / 3) * PT_W + (tap % 3)) * 128);
          uint64_t adesc = umma_smem_desc_sw128(a_addr);
          adesc = (adesc & ~(static_cast<uint64_t>(0x3FFF) << 32)) | (static_cast<uint64_t>((PT_W * 128) >> 4) << 32);
          // base-offset field stays 0: measured on B200, the unit applies the 128B swizzle to absolute
          // shared-memory address bits, exactly as TMA wrote the patch, so any 128-byte-aligned start works
          const uint64_t bdesc = umma_smem_desc_sw128(w_smem + tap * 8192u);
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_bf16_elect(d_tmem, adesc + 2u * k, bdesc + 2u * k, idesc, (tap | k) != 0 ? 1u : 0u);
        }
        umma_commit_elect(patch_empty(pbuf));
        umma_commit_elect(acc_full(b));
      }
      if (tr) p.trace[i * 8 + 3] = clock64();
    }
  } else {
    const int q = warp & 3;
    const int c = (warp - 2) >> 2;               // which 32-channel half of the 64 outputs this warp drains
    const int mrow = q * 32 + lane;              // row of the 128-row tile: (output row, column) = (mrow / 8, mrow % 8)
    const uint32_t sw = static_cast<uint32_t>(mrow & 7);
    const bool storer = threadIdx.x == 64;       // first epilogue thread issues the TMA stores
    float bias[32];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float4 bv = __ldg(reinterpret_cast<const float4*>(p.bias + c * 32) + j);
      bias[4 * j] = bv.x; bias[4 * j + 1] = bv.y; bias[4 * j + 2] = bv.z; bias[4 * j + 3] = bv.w;
    }
    int i = 0;
    for (int t = blockIdx.x; t < p.total_tiles; t += gridDim.x, ++i) {
      const int b = i & 1;
      const int pbuf = i % PT_BUFS;
      const int tt = p.rev ? p.total_tiles - 1 - t : t;
      const int n = tt / tiles_per_img;
      const int rem = tt - n * tiles_per_img;
      const int ty = rem / p.tiles_w, tx = rem - ty * p.tiles_w;
      const bool tre = kPairTrace && p.trace != nullptr && blockIdx.x == 0 && threadIdx.x == 64 && i < 200;
      if (tre) p.trace[i * 8 + 4] = clock64();
      mbar_wait(acc_full(b), (i >> 1) & 1);
      if (tre) p.trace[i * 8 + 5] = clock64();
      tc_fence_after();
      const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(b * 64 + c * 32);
      uint32_t v[32];
      tmem_ld32(t_row, v);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(acc_empty(b));
      if (tre) p.trace[i * 8 + 6] = clock64() + (v[0] & 0);
      float f[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]) + bias[j];
      if (has_res) {
        const uint32_t r_row = res0 + pbuf * PT_TILE_BYTES + static_cast<uint32_t>(mrow) * 128u;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint32_t rr[4];
          asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(rr[0]), "=r"(rr[1]), "=r"(rr[2]), "=r"(rr[3])
                       : "r"(r_row + ((static_cast<uint32_t>(c * 4 + g) ^ sw) << 4)));
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            f[g * 8 + e * 2] += lo16_t<F16>(rr[e]);
            f[g * 8 + e * 2 + 1] += hi16_t<F16>(rr[e]);
          }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(res_empty(pbuf));
      }
      if (p.relu) {
#pragma unroll
        for (int j = 0; j < 32; ++j) f[j] = fmaxf(f[j], 0.f);
      }
      // staging tile b was handed to the TMA store two tiles ago: wait until that store has read it
      if (storer) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
      asm volatile("bar.sync 1, 256;" ::: "memory");
      const uint32_t o_row = out0 + b * PT_TILE_BYTES + static_cast<uint32_t>(mrow) * 128u;
#pragma unroll
      for (int g = 0; g < 4; ++g)
        asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(o_row + ((static_cast<uint32_t>(c * 4 + g) ^ sw) << 4)),
                     "r"(pack16x2_t<F16>(f[g * 8 + 0], f[g * 8 + 1])), "r"(pack16x2_t<F16>(f[g * 8 + 2], f[g * 8 + 3])),
                     "r"(pack16x2_t<F16>(f[g * 8 + 4], f[g * 8 + 5])), "r"(pack16x2_t<F16>(f[g * 8 + 6], f[g * 8 + 7]))
                     : "memory");
      fence_proxy_async_smem();
      asm volatile("bar.sync 1, 256;" ::: "memory");
      if (storer) {
        tma_store_4d(&tmO, out0 + b * PT_TILE_BYTES, 0, tx * 8, ty * 16, n);
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      }
      if (tre) p.trace[i * 8 + 7] = clock64();
    }
    if (kPairTrace && p.trace != nullptr && blockIdx.x == 0 && threadIdx.x == 64) p.trace[1592] = clock64();
    if (storer) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    if (kPairTrace && p.trace != nullptr && blockIdx.x == 0 && threadIdx.x == 64) p.trace[1593] = clock64();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 128);
  }
}

int persistent_stages(int BN) {
  int s = (210 * 1024) / (A_TILE_BYTES + BN * 128);
  if (s > MAX_STAGES) s = MAX_STAGES;
  return s;
}

void fill_kp(KP* p, const ConvArgs& a) {
  const ConvGeom& g = a.g;
  p->H = g.H; p->W = g.W; p->C = g.C; p->OH = g.OH; p->OW = g.OW; p->R = g.R; p->S = g.S;
  p->Sp = g.Sp; p->stride = g.stride; p->pad = g.pad; p->Cout = g.Cout; p->K = g.K; p->M = g.M;
  p->BN = a.BN; p->stages = a.stages; p->num_kb = g.K / BK; p->tiles_n = g.Cout_pad / a.BN;
  p->tiles_m = (g.M + BM - 1) / BM;
  p->ldo = a.ldo; p->relu = a.relu; p->out_fp32 = a.out_fp32;
  uint32_t cols = 32;
  while (cols < static_cast<uint32_t>(a.BN)) cols <<= 1;
  p->tmem_cols = cols;
  p->x = static_cast<const __nv_bfloat16*>(a.x);
  p->w = static_cast<const __nv_bfloat16*>(a.w);
  p->bias = a.bias;
  p->residual = static_cast<const __nv_bfloat16*>(a.residual);
  p->out = a.out;
  p->tiles_n2 = a.tmB2 ? g.Cout_pad / a.BN : 0;
  p->kb_first2 = a.kb_first2;
  p->num_kb2 = a.num_kb2;
  p->ldo2 = a.ldo2;
  p->relu2 = a.relu2;
  p->Cout2 = g.Cout;
  p->bias2 = a.bias2;
  p->out2 = a.out2;
  p->resb = 0;
  p->msub = 1;
  p->f16 = a.f16;
  p->rev = a.reverse;
  p->nprod = 1;
  p->a_rows = BM;
  p->trace = nullptr;
  p->trace_cta = 0;
  p->splits = 1;
  p->ldw = 0;
  p->ws = nullptr;
  p->ctr = nullptr;
}

template <int AMODE, bool F16>
cudaError_t launch_persistent(const ConvArgs& a, const CUtensorMap* tmA, const CUtensorMap* tmB,
                              cudaStream_t st) {
  static bool attr_set[16] = {};
  static int sm_count[16] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 16) return cudaErrorInvalidDevice;
  auto kern = conv_umma_persistent_kernel<AMODE, F16>;
  if (!attr_set[dev]) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024);
    if (e != cudaSuccess) return e;
    e = cudaDeviceGetAttribute(&sm_count[dev], cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess) return e;
    attr_set[dev] = true;
  }
  KP p;
  fill_kp(&p, a);
  p.stages = persistent_stages(a.BN);
  uint32_t cols = 32;
  while (cols < static_cast<uint32_t>(2 * a.BN)) cols <<= 1;  // two accumulators
  p.tmem_cols = cols;
  const int total = p.tiles_m * p.tiles_n;
  const int grid = total < sm_count[dev] ? total : sm_count[dev];
  const size_t smem = conv_umma_smem_bytes(a.BN, p.stages);
  if (a.splits > 1 && a.ws && a.ctr && p.tiles_n2 == 0 && AMODE == A_TILED) {
    // keep every split non-empty: kb_per = ceil(num_kb / splits) and (splits - 1) * kb_per < num_kb
    int sp = std::min(a.splits, p.num_kb);
    while (sp > 1 && (sp - 1) * ((p.num_kb + sp - 1) / sp) >= p.num_kb) --sp;
    p.splits = sp;
    p.ws = a.ws;
    p.ctr = a.ctr;
    p.ldw = a.ldw;
  }
  if (AMODE == A_TILED && a.a_rows >= 8 && a.a_rows < BM && (a.a_rows & 7) == 0 && p.tiles_m == 1) p.a_rows = a.a_rows;
  const int total2 = p.tiles_m * (p.tiles_n + p.tiles_n2) * p.splits;
  const int grid2 = total2 < sm_count[dev] ? total2 : sm_count[dev];
  (void)grid;
  static const int nprod_env = [] { const char* e = getenv("B2I_NPROD"); return e ? atoi(e) : 0; }();
  p.nprod = nprod_env >= 1 && nprod_env <= PERS_MAX_PROD ? nprod_env : PERS_MAX_PROD;
  p.nprod = std::min(p.nprod, p.stages);
  return launch_pdl(kern, dim3(static_cast<unsigned>(grid2)), dim3(PERS_THREADS), smem, st, *tmA, *tmB, a.tmB2 ? *a.tmB2 : *tmB, p);
}

template <int AMODE, bool F16>
cudaError_t launch_pair(const ConvArgs& a, const CUtensorMap* tmA, const CUtensorMap* tmB, cudaStream_t st) {
  static bool attr_set[16] = {};
  static int sm_count[16] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 16) return cudaErrorInvalidDevice;
  auto kern = conv_umma_pair_kernel<AMODE, F16>;
  if (!attr_set[dev]) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 221 * 1024);
    if (e != cudaSuccess) return e;
    e = cudaDeviceGetAttribute(&sm_count[dev], cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess) return e;
    attr_set[dev] = true;
  }
  KP p;
  fill_kp(&p, a);
  p.msub = (AMODE == A_IM2COL && (a.msub == 2 || a.msub == 4)) ? a.msub : 1;   // powers of two only: 384-pixel requests fault
  p.tiles_m = (a.g.M + 256 * p.msub - 1) / (256 * p.msub);
  const int a_bytes = p.msub * A_TILE_BYTES;
  const int stage = a_bytes + (a.BN / 2) * 128;
  p.stages = std::min(MAX_STAGES, (210 * 1024) / stage);
  uint32_t cols = 32;
  while (cols < static_cast<uint32_t>(std::min(512, 2 * p.msub * a.BN))) cols <<= 1;   // two tiles' accumulators when they fit
  p.tmem_cols = cols;
  const int total = p.tiles_m * (p.tiles_n + p.tiles_n2);
  int grid = std::min(2 * total, sm_count[dev] & ~1);
  size_t smem = static_cast<size_t>(p.stages) * stage + 1024;
  const size_t resb_bytes = static_cast<size_t>(p.num_kb + (p.tiles_n2 ? p.num_kb2 : 0)) * (a.BN / 2) * 128;
  if (a.resident_b && p.tiles_n == 1 && p.tiles_n2 <= 1 && resb_bytes + 3 * a_bytes + 1024 <= 221 * 1024) {
    p.resb = 1;
    p.stages = std::min<int>(MAX_STAGES, static_cast<int>((221 * 1024 - 1024 - resb_bytes) / a_bytes));
    smem = resb_bytes + static_cast<size_t>(p.stages) * a_bytes + 1024;
  }
  if (p.stages < 3) return cudaErrorInvalidValue;
  // B2I_NPROD=1..4 overrides the number of TMA-issuing warps (A/B timing); im2col layers default to 4, never more
  // than the ring has stages
  static const int nprod_env = [] { const char* e = getenv("B2I_NPROD"); return e ? atoi(e) : 0; }();
  p.nprod = nprod_env >= 1 && nprod_env <= PAIR_MAX_PROD ? nprod_env : (AMODE == A_IM2COL ? PAIR_MAX_PROD : 1);
  p.nprod = std::min(p.nprod, p.stages);
  static const bool trace_on = [] { const char* e = getenv("B2I_PAIR_TRACE"); return e && e[0] == '1'; }();
  if (kPairTrace && trace_on) {   // debugging: run this launch alone and print what CTA 0's producers and MMA warp saw, k-block by k-block
    static long long* d_trace = nullptr;
    if (!d_trace) cudaMalloc(&d_trace, 8192 * sizeof(long long));
    cudaMemsetAsync(d_trace, 0, 8192 * sizeof(long long), st);
    p.trace = d_trace;
    static const int trace_cta = [] { const char* e = getenv("B2I_PAIR_TRACE_CTA"); return e ? atoi(e) : 0; }();
    p.trace_cta = trace_cta < grid ? (trace_cta & ~1) : 0;
    cudaError_t e = launch_pdl(kern, dim3(static_cast<unsigned>(grid)), dim3(PAIR_THREADS), smem, st, *tmA, *tmB, a.tmB2 ? *a.tmB2 : *tmB, p);
    if (e != cudaSuccess) return e;
    e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) return e;
    static long long h[8192];
    cudaMemcpy(h, d_trace, sizeof h, cudaMemcpyDeviceToHost);
    const int my_tiles = (total - p.trace_cta / 2 + grid / 2 - 1) / (grid / 2);
    const int per_cta = my_tiles * p.num_kb;
    fprintf(stderr, "[pair trace] CTA %d: kernel entry -> past the wait for the previous grid %lld cycles; -> first MMA-warp wait %lld; last MMA issued at %lld; epilogue warps done at %lld; cluster released at %lld\n",
            p.trace_cta, h[8001] - h[8000], h[0] - h[8000], h[(std::min(per_cta, 1000) - 1) * 4 + 2] - h[8000], h[8002] - h[8000], h[8003] - h[8000]);
    const int n = std::min(per_cta, 1000);
    long long wait_full = 0, wait_empty = 0, issue = 0, lat = 0;
    int lat_n = 0;
    for (int g = 0; g < n; ++g) {
      wait_full += h[g * 4 + 1] - h[g * 4];
      if (h[4096 + g * 4 + 1]) {
        wait_empty += h[4096 + g * 4 + 1] - h[4096 + g * 4];
        issue += h[4096 + g * 4 + 2] - h[4096 + g * 4 + 1];
        if (h[g * 4 + 1] > h[4096 + g * 4 + 2]) { lat += h[g * 4 + 1] - h[4096 + g * 4 + 2]; ++lat_n; }
      }
    }
    const long long span = h[(n - 1) * 4 + 2] - h[0];
    fprintf(stderr, "[pair trace] M %d Cout %d K %d BN %d msub %d stages %d nprod %d grid %d: %d k-blocks in CTA 0, %lld cycles = %.0f per k-block; MMA warp waited for data %.0f per k-block; producers waited for a free stage %.0f, spent %.0f issuing; request issued -> MMA saw it full %.0f (n %d)\n",
            a.g.M, a.g.Cout, a.g.K, a.BN, p.msub, p.stages, p.nprod, grid, n, span, static_cast<double>(span) / n, static_cast<double>(wait_full) / n,
            static_cast<double>(wait_empty) / n, static_cast<double>(issue) / n, lat_n ? static_cast<double>(lat) / lat_n : 0.0, lat_n);
    for (int i = 0; i < my_tiles && i < 4; ++i)
      for (int j = 0; j < 4; ++j)
        if (h[6000 + (i * 4 + j) * 4])
          fprintf(stderr, "    epilogue warp 2, tile %d chunk %d: starts at %lld, accumulator read %lld cycles, bias / residual / store %lld cycles\n", i, j,
                  h[6000 + (i * 4 + j) * 4] - h[8000], h[6000 + (i * 4 + j) * 4 + 1] - h[6000 + (i * 4 + j) * 4], h[6000 + (i * 4 + j) * 4 + 2] - h[6000 + (i * 4 + j) * 4 + 1]);
    if (getenv("B2I_PAIR_TRACE_DUMP")) {
      for (int g = 0; g < std::min(n, 80); ++g)
        fprintf(stderr, "  kb %3d prod %lld: empty-wait %6lld..%6lld issued %6lld | mma full-wait %6lld..%6lld done-issue %6lld\n", g, h[4096 + g * 4 + 3],
                h[4096 + g * 4] - h[0], h[4096 + g * 4 + 1] - h[0], h[4096 + g * 4 + 2] - h[0], h[g * 4] - h[0], h[g * 4 + 1] - h[0], h[g * 4 + 2] - h[0]);
    }
    return cudaSuccess;
  }
  return launch_pdl(kern, dim3(static_cast<unsigned>(grid)), dim3(PAIR_THREADS), smem, st, *tmA, *tmB, a.tmB2 ? *a.tmB2 : *tmB, p);
}

template <bool F16>
cudaError_t launch_patch(const ConvArgs& a, const CUtensorMap* tmX, const CUtensorMap* tmW,
                         cudaStream_t st) {
  if (!a.tmO || (a.residual && !a.tmR)) return cudaErrorInvalidValue;
  static bool attr_set[16] = {};
  static int sm_count[16] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 16) return cudaErrorInvalidDevice;
  const size_t smem = PT_WBYTES + PT_BUFS * (PT_BYTES + PT_TILE_BYTES) + 2 * PT_TILE_BYTES + 1024;
  if (!attr_set[dev]) {
    cudaError_t e = cudaFuncSetAttribute(conv3x3_c64_patch_kernel<F16>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(smem));
    if (e != cudaSuccess) return e;
    e = cudaDeviceGetAttribute(&sm_count[dev], cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess) return e;
    attr_set[dev] = true;
  }
  const ConvGeom& g = a.g;
  PatchP p;
  p.N = g.N; p.H = g.H; p.W = g.W;
  p.tiles_w = (g.W + 7) / 8;
  p.tiles_h = (g.H + 15) / 16;
  p.total_tiles = g.N * p.tiles_w * p.tiles_h;
  p.relu = a.relu;
  p.f16 = a.f16;
  p.rev = a.reverse;
  p.bias = a.bias;
  p.residual = static_cast<const __nv_bfloat16*>(a.residual);
  p.out = static_cast<__nv_bfloat16*>(a.out);
  const int grid = p.total_tiles < sm_count[dev] ? p.total_tiles : sm_count[dev];
  p.trace = nullptr;
  static const bool trace_on = [] { const char* e = getenv("B2I_PATCH_TRACE"); return e && e[0] == '1'; }();
  if (kPairTrace && trace_on) {
    static long long* d_trace = nullptr;
    if (!d_trace) cudaMalloc(&d_trace, 1600 * sizeof(long long));
    cudaMemsetAsync(d_trace, 0, 1600 * sizeof(long long), st);
    p.trace = d_trace;
    cudaError_t e = launch_pdl(conv3x3_c64_patch_kernel<F16>, dim3(grid), dim3(PATCH_THREADS), smem, st, *tmX, *tmW, a.residual ? *a.tmR : *a.tmO, *a.tmO, p);
    if (e != cudaSuccess) return e;
    e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) return e;
    static long long h[1600];
    cudaMemcpy(h, d_trace, sizeof h, cudaMemcpyDeviceToHost);
    fprintf(stderr, "[patch trace] CTA 0 (residual %d): MMA warp: starts waiting / accumulator free / patch full / MMAs issued | epilogue warp 2: starts waiting / accumulator full / accumulator read / tile staged\n", a.residual ? 1 : 0);
    int last = 0;
    while (last < 170 && h[(last + 1) * 8]) ++last;
    fprintf(stderr, "  kernel entry -> past the wait for the previous grid %lld cycles; -> MMA warp's first wait %lld; last MMAs issued at %lld (%d tiles); last tile staged at %lld; its store complete at %lld\n",
            h[1591] - h[1590], h[0] - h[1590], h[last * 8 + 3] - h[1590], last + 1, h[1592] - h[1590], h[1593] - h[1590]);
    for (int i = 0; i < 170 && h[i * 8]; ++i)
      if (i < 4)
        fprintf(stderr, "  tile %3d: %7lld %7lld %7lld %7lld | %7lld %7lld %7lld %7lld\n", i, h[i * 8] - h[0], h[i * 8 + 1] - h[0], h[i * 8 + 2] - h[0], h[i * 8 + 3] - h[0],
                h[i * 8 + 4] - h[0], h[i * 8 + 5] - h[0], h[i * 8 + 6] - h[0], h[i * 8 + 7] - h[0]);
    return cudaSuccess;
  }
  return launch_pdl(conv3x3_c64_patch_kernel<F16>, dim3(grid), dim3(PATCH_THREADS), smem, st, *tmX, *tmW, a.residual ? *a.tmR : *a.tmO, *a.tmO, p);
}

template <bool F16>
cudaError_t launch_flat(const ConvArgs& a, const CUtensorMap* tmX, const CUtensorMap* tmB, cudaStream_t st) {
  static bool attr_set[16] = {};
  static int sm_count[16] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 16) return cudaErrorInvalidDevice;
  if (!attr_set[dev]) {
    cudaError_t e = cudaFuncSetAttribute(conv3x3_flat_pair_kernel<F16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 225 * 1024);
    if (e != cudaSuccess) return e;
    e = cudaDeviceGetAttribute(&sm_count[dev], cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess) return e;
    attr_set[dev] = true;
  }
  const ConvGeom& g = a.g;
  FlatP p;
  p.H = g.H; p.W = g.W; p.N = g.N; p.slices = g.C / 64; p.BN = a.BN;
  p.pitch = conv_flat_pitch(g.W);
  p.rpt = 128 / p.pitch;
  p.box_w = p.pitch;
  p.rows_total = g.N * (g.H + 1);
  p.total_pairs = (p.rows_total + 2 * p.rpt - 1) / (2 * p.rpt);
  const size_t resb = static_cast<size_t>(9) * p.slices * (a.BN / 2) * 128;
  const size_t patch = static_cast<size_t>(p.rpt + 2) * p.pitch * 128 + 1024;
  p.nbuf = static_cast<int>(std::min<size_t>(FLAT_MAXBUF, (225 * 1024 - 1024 - resb) / patch));
  p.ldo = a.ldo; p.relu = a.relu; p.Cout = g.Cout; p.f16 = a.f16; p.rev = a.reverse;
  uint32_t cols = 32;
  while (cols < static_cast<uint32_t>(2 * a.BN)) cols <<= 1;
  p.tmem_cols = cols;
  p.bias = a.bias;
  p.residual = static_cast<const __nv_bfloat16*>(a.residual);
  p.out = a.out;
  const int grid = std::min(2 * p.total_pairs, sm_count[dev] & ~1);
  const size_t smem = resb + p.nbuf * patch + 1024;
  return launch_pdl(conv3x3_flat_pair_kernel<F16>, dim3(static_cast<unsigned>(grid)), dim3(PTHREADS), smem, st, *tmX, *tmB, p);
}

template <bool F16>
cudaError_t launch_flat_stream(const ConvArgs& a, const CUtensorMap* tmX, const CUtensorMap* tmB, cudaStream_t st) {
  static bool attr_set[16] = {};
  static int sm_count[16] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 16) return cudaErrorInvalidDevice;
  if (!attr_set[dev]) {
    cudaError_t e = cudaFuncSetAttribute(conv3x3_flat_stream_kernel<F16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024);
    if (e != cudaSuccess) return e;
    e = cudaDeviceGetAttribute(&sm_count[dev], cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess) return e;
    attr_set[dev] = true;
  }
  const ConvGeom& g = a.g;
  FlatSP p;
  p.H = g.H; p.W = g.W; p.N = g.N; p.slices = g.C / 64; p.BN = a.BN;
  p.pitch = conv_flat_stream_pitch(g.W);
  p.rpt = 128 / p.pitch;
  p.box_w = p.pitch;
  p.tiles_n = g.Cout / a.BN;
  const int rows_total = g.N * (g.H + 1);
  p.total_pairs = (rows_total + 2 * p.rpt - 1) / (2 * p.rpt);
  const size_t patch = ((static_cast<size_t>(p.rpt + 2) * p.pitch * 128 + 1024) + 1023) & ~static_cast<size_t>(1023);
  const size_t btile = static_cast<size_t>(a.BN / 2) * 128;
  p.na = 3;
  p.nb = static_cast<int>(std::min<size_t>(FLATS_MAXB, (225 * 1024 - 1024 - p.na * patch) / btile));
  if (p.nb < 3) return cudaErrorInvalidValue;
  p.ldo = a.ldo; p.relu = a.relu; p.Cout = g.Cout; p.f16 = a.f16; p.rev = a.reverse;
  uint32_t cols = 32;
  while (cols < static_cast<uint32_t>(2 * a.BN)) cols <<= 1;
  p.tmem_cols = cols;
  p.bias = a.bias;
  p.residual = static_cast<const __nv_bfloat16*>(a.residual);
  p.out = a.out;
  const int grid = std::min(2 * p.total_pairs * p.tiles_n, sm_count[dev] & ~1);
  const size_t smem = p.na * patch + p.nb * btile + 1024;
  return launch_pdl(conv3x3_flat_stream_kernel<F16>, dim3(static_cast<unsigned>(grid)), dim3(FLATS_THREADS), smem, st, *tmX, *tmB, p);
}

template <int AMODE, bool B_TMA, bool F16>
cudaError_t launch_t(const ConvArgs& a, const CUtensorMap* tmA, const CUtensorMap* tmB,
                     cudaStream_t st) {
  static bool attr_set[16] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  auto kern = conv_umma_kernel<AMODE, B_TMA, F16>;
  if (dev < 16 && !attr_set[dev]) {
    cudaError_t e =
        cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024);  // 227 KB minus the static barriers
    if (e != cudaSuccess) return e;
    attr_set[dev] = true;
  }
  KP p;
  fill_kp(&p, a);
  const int tiles_m = p.tiles_m;
  const size_t smem = conv_umma_smem_bytes(a.BN, a.stages);
  static const CUtensorMap dummy = {};
  const dim3 grid(static_cast<unsigned>(tiles_m * p.tiles_n));
  if (AMODE == A_IM2COL || AMODE == A_TILED)   // activations arrive by TMA (L2): safe to start under the previous kernel's tail
    return launch_pdl(kern, grid, dim3(NTHREADS), smem, st, tmA ? *tmA : dummy, tmB ? *tmB : dummy, p);
  kern<<<grid, dim3(NTHREADS), smem, st>>>(tmA ? *tmA : dummy, tmB ? *tmB : dummy, p);   // gathered A goes through L1
  return cudaGetLastError();
}

}  // namespace

size_t conv_umma_smem_bytes(int BN, int stages) {
  return static_cast<size_t>(stages) * (A_TILE_BYTES + BN * 128) + 1024;
}

// Two CTAs per SM when three stages fit in half the shared memory (epilogue of one tile then
// overlaps the main loop of the other); otherwise one CTA with as many stages as fit.
int conv_umma_pick_stages(int BN) {
  const int stage = A_TILE_BYTES + BN * 128;
  const int half = 112 * 1024;
  int s;
  if (3 * stage <= half) s = half / stage;
  else s = (220 * 1024) / stage;
  if (s > 6) s = 6;
  if (s < 3) s = 3;
  return s;
}

template <bool F16>
static cudaError_t launch_conv_umma_t(const ConvArgs& a, const CUtensorMap* tmA, const CUtensorMap* tmB,
                                      cudaStream_t st) {
  const ConvGeom& g = a.g;
  if (a.BN < 32 || a.BN > 256 || (a.BN & 31) || g.Cout_pad % a.BN || g.K % BK || a.stages < 3 ||
      a.stages > MAX_STAGES || g.M <= 0)
    return cudaErrorInvalidValue;
  if (!a.out_fp32 && ((a.ldo & 15) || (g.Cout & 31))) return cudaErrorInvalidValue;
  if (a.tmB2 && !((a.amode == A_IM2COL) && (a.persistent || a.pair))) return cudaErrorInvalidValue;
  if (a.flat == 2) {
    if (!conv_flat_stream_applicable(g, a.BN) || a.out_fp32 || a.ldo != g.Cout || !tmA || !tmB) return cudaErrorInvalidValue;
    return launch_flat_stream<F16>(a, tmA, tmB, st);
  }
  if (a.flat) {
    if (!conv_flat_applicable(g, a.BN) || a.out_fp32 || a.ldo != g.Cout || !tmA || !tmB) return cudaErrorInvalidValue;
    return launch_flat<F16>(a, tmA, tmB, st);
  }
  if (a.patch) {
    if (!conv_patch_applicable(g) || a.out_fp32 || a.ldo != 64 || !tmA || !tmB) return cudaErrorInvalidValue;
    return launch_patch<F16>(a, tmA, tmB, st);
  }
  switch (a.amode) {
    case A_GATHER:
      if (g.C % 64 || g.K != g.R * g.S * g.C) return cudaErrorInvalidValue;
      return a.b_tma ? launch_t<A_GATHER, true, F16>(a, nullptr, tmB, st)
                     : launch_t<A_GATHER, false, F16>(a, nullptr, nullptr, st);
    case A_SMALLC:
      if (g.C != 4 || g.K < g.R * g.Sp * 4) return cudaErrorInvalidValue;
      return a.b_tma ? launch_t<A_SMALLC, true, F16>(a, nullptr, tmB, st)
                     : launch_t<A_SMALLC, false, F16>(a, nullptr, nullptr, st);
    case A_IM2COL:
      if (g.C % 64 || g.K != g.R * g.S * g.C || !tmA || !tmB) return cudaErrorInvalidValue;
      if (a.pair && a.BN >= 64) return launch_pair<A_IM2COL, F16>(a, tmA, tmB, st);
      return a.persistent ? launch_persistent<A_IM2COL, F16>(a, tmA, tmB, st)
                          : launch_t<A_IM2COL, true, F16>(a, tmA, tmB, st);
    case A_TILED:
      if (!tmA || !tmB) return cudaErrorInvalidValue;
      if (a.pair && a.BN >= 64) return launch_pair<A_TILED, F16>(a, tmA, tmB, st);
      return a.persistent ? launch_persistent<A_TILED, F16>(a, tmA, tmB, st)
                          : launch_t<A_TILED, true, F16>(a, tmA, tmB, st);
  }
  return cudaErrorInvalidValue;
}

cudaError_t launch_conv_umma(const ConvArgs& a, const CUtensorMap* tmA, const CUtensorMap* tmB,
                             cudaStream_t st) {
  return a.f16 ? launch_conv_umma_t<true>(a, tmA, tmB, st) : launch_conv_umma_t<false>(a, tmA, tmB, st);
}

// ---------------------------------------------------------------------------- tensor maps
namespace {
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
typedef CUresult (*EncodeIm2colFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                   const cuuint64_t*, const cuuint64_t*, const int*, const int*,
                                   cuuint32_t, cuuint32_t, const cuuint32_t*,
                                   CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

void* driver_fn(const char* name) {
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &fn, cudaEnableDefault, &q) != cudaSuccess ||
      q != cudaDriverEntryPointSuccess)
    return nullptr;
  return fn;
}
}  // namespace

int make_tmap_rows(CUtensorMap* tm, const void* a, int rows, int K, int box_rows,
                   const char** err) {
  static EncodeTiledFn fn = reinterpret_cast<EncodeTiledFn>(driver_fn("cuTensorMapEncodeTiled"));
  if (!fn) { *err = "cuTensorMapEncodeTiled entry point not found"; return -1; }
  cuuint64_t dims[2] = {static_cast<cuuint64_t>(K), static_cast<cuuint64_t>(rows)};
  cuuint64_t strides[1] = {static_cast<cuuint64_t>(K) * 2};
  cuuint32_t box[2] = {BK, static_cast<cuuint32_t>(box_rows)};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(a), dims, strides,
                  box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { *err = "cuTensorMapEncodeTiled failed"; return static_cast<int>(r); }
  return 0;
}

int make_tmap_weights(CUtensorMap* tm, const void* w, int Cout_pad, int K, int BN,
                      const char** err) {
  return make_tmap_rows(tm, w, Cout_pad, K, BN, err);
}

// Padded-row pitch of the flattened-row kernel: the smallest of 16 / 32 / 64 pixels holding W + 2.
int conv_flat_pitch(int W) { return W + 2 <= 16 ? 16 : (W + 2 <= 32 ? 32 : 64); }

// True when the flattened-row CTA-pair kernel handles this geometry: 3x3 / s1 / p1, channels a multiple of
// 64, one N tile, and this CTA's half of the weights plus two halo patches fitting in shared memory.
bool conv_flat_applicable(const ConvGeom& g, int BN) {
  if (g.R != 3 || g.S != 3 || g.stride != 1 || g.pad != 1 || g.OH != g.H || g.OW != g.W) return false;
  if (g.C % 64 || g.K != 9 * g.C || BN != g.Cout_pad || BN != g.Cout || BN < 64 || BN > 256 || (BN & 31) || g.W + 2 > 64) return false;
  const int pitch = conv_flat_pitch(g.W);
  const size_t resb = static_cast<size_t>(9) * (g.C / 64) * (BN / 2) * 128;
  const size_t patch = static_cast<size_t>(128 / pitch + 2) * pitch * 128 + 1024;
  return resb + 2 * patch + 1024 <= 226 * 1024;
}

// Row pitch of the streamed-weights flattened-row kernel: W + 1 when that is a power of two >= 8 (the left pad of a row
// is the right pad of the row above), else the pitch of the resident-weights kernel.
int conv_flat_stream_pitch(int W) {
  const int p1 = W + 1;
  return (p1 >= 8 && p1 <= 64 && (p1 & (p1 - 1)) == 0) ? p1 : conv_flat_pitch(W);
}

// True when conv3x3_flat_stream_kernel handles this geometry: 3x3 / s1 / p1, channels a multiple of 64, output
// channels a multiple of the tile width (128 or 256).
bool conv_flat_stream_applicable(const ConvGeom& g, int BN) {
  if (g.R != 3 || g.S != 3 || g.stride != 1 || g.pad != 1 || g.OH != g.H || g.OW != g.W) return false;
  if (g.C % 64 || g.K != 9 * g.C || (BN != 128 && BN != 256) || g.Cout % BN || g.Cout != g.Cout_pad) return false;
  const int pitch = conv_flat_stream_pitch(g.W);
  return pitch <= 64 && (g.W + 2 <= pitch || g.W + 1 == pitch);
}

int make_tmap_flat(CUtensorMap* tm, const void* x, int N, int H, int W, int C, const char** err) {
  return make_tmap_flat_pitch(tm, x, N, H, W, C, conv_flat_pitch(W), err);
}

// Tiled 4-D map over NHWC [N][H][W][C] whose box is one 64-channel slice of `pitch` pixels of one row.
int make_tmap_flat_pitch(CUtensorMap* tm, const void* x, int N, int H, int W, int C, int pitch, const char** err) {
  static EncodeTiledFn fn = reinterpret_cast<EncodeTiledFn>(driver_fn("cuTensorMapEncodeTiled"));
  if (!fn) { *err = "cuTensorMapEncodeTiled entry point not found"; return -1; }
  cuuint64_t dims[4] = {static_cast<cuuint64_t>(C), static_cast<cuuint64_t>(W), static_cast<cuuint64_t>(H), static_cast<cuuint64_t>(N)};
  cuuint64_t strides[3] = {static_cast<cuuint64_t>(C) * 2, static_cast<cuuint64_t>(W) * C * 2, static_cast<cuuint64_t>(H) * W * C * 2};
  cuuint32_t box[4] = {64, static_cast<cuuint32_t>(pitch), 1, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(x), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { *err = "cuTensorMapEncodeTiled (flattened rows) failed"; return static_cast<int>(r); }
  return 0;
}

// Tiled 4-D map over NHWC [N][H][W][C]: one 64-channel slice of `box_w` pixels of one row, taking every
// `stride`-th pixel (the input pixels of `box_w` consecutive output pixels of a stride-`stride` conv).
int make_tmap_pixrow(CUtensorMap* tm, const void* x, int N, int H, int W, int C, int box_w, int stride, const char** err) {
  static EncodeTiledFn fn = reinterpret_cast<EncodeTiledFn>(driver_fn("cuTensorMapEncodeTiled"));
  if (!fn) { *err = "cuTensorMapEncodeTiled entry point not found"; return -1; }
  cuuint64_t dims[4] = {static_cast<cuuint64_t>(C), static_cast<cuuint64_t>(W), static_cast<cuuint64_t>(H), static_cast<cuuint64_t>(N)};
  cuuint64_t strides[3] = {static_cast<cuuint64_t>(C) * 2, static_cast<cuuint64_t>(W) * C * 2, static_cast<cuuint64_t>(H) * W * C * 2};
  cuuint32_t box[4] = {64, static_cast<cuuint32_t>(box_w * stride), 1, 1};
  cuuint32_t estr[4] = {1, static_cast<cuuint32_t>(stride), 1, 1};
  CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(x), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { *err = "cuTensorMapEncodeTiled (pixel row) failed"; return static_cast<int>(r); }
  return 0;
}

bool conv_patch_applicable(const ConvGeom& g) {
  return g.C == 64 && g.Cout == 64 && g.Cout_pad == 64 && g.R == 3 && g.S == 3 && g.stride == 1 &&
         g.pad == 1 && g.K == 576 && g.OH == g.H && g.OW == g.W;
}

// Tiled 4-D map over NHWC [N][H][W][64] with an 18 x 10 pixel box: the halo patch of one 16 x 8
// output tile; out-of-image coordinates (the conv padding) are zero-filled by the TMA unit.
int make_tmap_patch(CUtensorMap* tm, const void* x, int N, int H, int W, const char** err) {
  return make_tmap_nhwc64(tm, x, N, H, W, PT_W, PT_H, err);
}

int make_tmap_tile(CUtensorMap* tm, const void* x, int N, int H, int W, const char** err) {
  return make_tmap_nhwc64(tm, x, N, H, W, 8, 16, err);
}

int make_tmap_nhwc64(CUtensorMap* tm, const void* x, int N, int H, int W, int box_w, int box_h, const char** err) {
  static EncodeTiledFn fn = reinterpret_cast<EncodeTiledFn>(driver_fn("cuTensorMapEncodeTiled"));
  if (!fn) { *err = "cuTensorMapEncodeTiled entry point not found"; return -1; }
  cuuint64_t dims[4] = {64, static_cast<cuuint64_t>(W), static_cast<cuuint64_t>(H), static_cast<cuuint64_t>(N)};
  cuuint64_t strides[3] = {128, static_cast<cuuint64_t>(W) * 128, static_cast<cuuint64_t>(H) * W * 128};
  cuuint32_t box[4] = {64, static_cast<cuuint32_t>(box_w), static_cast<cuuint32_t>(box_h), 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(x), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { *err = "cuTensorMapEncodeTiled (patch) failed"; return static_cast<int>(r); }
  return 0;
}

int make_tmap_im2col(CUtensorMap* tm, const void* x, const ConvGeom& g, int Nmax,
                     const char** err) {
  return make_tmap_im2col_px(tm, x, g, Nmax, BM, err);
}

int make_tmap_im2col_px(CUtensorMap* tm, const void* x, const ConvGeom& g, int Nmax, int pixels,
                        const char** err) {
  static EncodeIm2colFn fn =
      reinterpret_cast<EncodeIm2colFn>(driver_fn("cuTensorMapEncodeIm2col"));
  if (!fn) { *err = "cuTensorMapEncodeIm2col entry point not found"; return -1; }
  cuuint64_t dims[4] = {static_cast<cuuint64_t>(g.C), static_cast<cuuint64_t>(g.W),
                        static_cast<cuuint64_t>(g.H), static_cast<cuuint64_t>(Nmax)};
  cuuint64_t strides[3] = {static_cast<cuuint64_t>(g.C) * 2,
                           static_cast<cuuint64_t>(g.W) * g.C * 2,
                           static_cast<cuuint64_t>(g.H) * g.W * g.C * 2};
  // Bounding box of filter-window origins: [-pad, dim + pad - (window - 1)) per spatial axis.
  int lower[2] = {-g.pad, -g.pad};
  int upper[2] = {g.pad - (g.S - 1), g.pad - (g.R - 1)};
  cuuint32_t estr[4] = {1, static_cast<cuuint32_t>(g.stride), static_cast<cuuint32_t>(g.stride), 1};
  CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(x), dims, strides,
                  lower, upper, /*channelsPerPixel=*/BK, /*pixelsPerColumn=*/static_cast<cuuint32_t>(pixels), estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { *err = "cuTensorMapEncodeIm2col failed"; return static_cast<int>(r); }
  // Drivers up to CUDA 13.1 mis-encode im2col maps over tensors smaller than 128 KiB; the
  // publicly documented fix-up is to clear bit 21 of the second descriptor word.
  int drv = 0;
  cudaDriverGetVersion(&drv);
  const size_t bytes = static_cast<size_t>(Nmax) * g.H * g.W * g.C * 2;
  if (drv <= 13010 && bytes < 131072)
    reinterpret_cast<uint64_t*>(tm)[1] &= ~(1ull << 21);
  return 0;
}

}  // namespace b2i
