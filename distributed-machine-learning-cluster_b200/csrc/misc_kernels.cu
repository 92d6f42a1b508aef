// Bandwidth-bound kernels of the inference path (sm_100a): normalise, max-pool, adaptive
// average pool, fused softmax + top-k.  All are HBM-bound: coalesced 16-byte accesses along the
// NHWC channel axis, grids sized in multiples of the 148 SMs, warp shuffles for reductions.
#include "ops.h"
#include "ptx.cuh"

#include <cuda_bf16.h>
#include <float.h>

namespace b2i {
namespace {

constexpr int kSMs = 148;

__device__ __forceinline__ uint32_t pack2(float a, float b, int f16) { return pack16x2(a, b, f16); }

// ------------------------------------------------------------------ normalise
// tch: x.to_kind(Float) / 255.0, then (x - mean) / std.  True
// divisions, in that order, so the fp32 mode reproduces the reference's rounding.
__constant__ float kMean[3] = {0.485f, 0.456f, 0.406f};
__constant__ float kStd[3] = {0.229f, 0.224f, 0.225f};

template <bool F32>
__global__ void __launch_bounds__(256) normalize_u8_kernel(const uint8_t* __restrict__ in,
                                                            void* __restrict__ out,
                                                            long long quads, int f16) {
  // one thread = 4 pixels = 12 input bytes (three aligned 32-bit loads)
  for (long long q = blockIdx.x * 256ll + threadIdx.x; q < quads; q += 256ll * gridDim.x) {
    const uint32_t* src = reinterpret_cast<const uint32_t*>(in) + q * 3;
    const uint32_t w0 = __ldg(src), w1 = __ldg(src + 1), w2 = __ldg(src + 2);
    uint8_t b[12];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      b[i] = (w0 >> (8 * i)) & 0xFF;
      b[4 + i] = (w1 >> (8 * i)) & 0xFF;
      b[8 + i] = (w2 >> (8 * i)) & 0xFF;
    }
    float v[12];
#pragma unroll
    for (int i = 0; i < 12; ++i) {
      const int c = i % 3;
      v[i] = __fdiv_rn(__fsub_rn(__fdiv_rn(static_cast<float>(b[i]), 255.0f), kMean[c]), kStd[c]);
    }
    if (F32) {
      float4* o = reinterpret_cast<float4*>(out) + q * 4;
#pragma unroll
      for (int p = 0; p < 4; ++p) o[p] = make_float4(v[3 * p], v[3 * p + 1], v[3 * p + 2], 0.f);
    } else {
      uint4* o = reinterpret_cast<uint4*>(out) + q * 2;
      o[0] = make_uint4(pack2(v[0], v[1], f16), pack2(v[2], 0.f, f16), pack2(v[3], v[4], f16), pack2(v[5], 0.f, f16));
      o[1] = make_uint4(pack2(v[6], v[7], f16), pack2(v[8], 0.f, f16), pack2(v[9], v[10], f16), pack2(v[11], 0.f, f16));
    }
  }
}

// ------------------------------------------------------------------ max pool
// One thread = one output pixel x 8 bf16 channels (16 bytes) or 4 fp32 channels.
template <bool F32>
__global__ void __launch_bounds__(256) maxpool_kernel(const void* __restrict__ in,
                                                       void* __restrict__ out, int N, int H, int W,
                                                       int C, int k, int stride, int pad, int OH,
                                                       int OW, int f16) {
  const int vec = F32 ? 4 : 8;
  const int cv = C / vec;
  const long long total = static_cast<long long>(N) * OH * OW * cv;
  for (long long i = blockIdx.x * 256ll + threadIdx.x; i < total; i += 256ll * gridDim.x) {
    const int c = static_cast<int>(i % cv);
    long long pix = i / cv;
    const int ow = static_cast<int>(pix % OW);
    pix /= OW;
    const int oh = static_cast<int>(pix % OH);
    const int n = static_cast<int>(pix / OH);
    const int h0 = oh * stride - pad, w0 = ow * stride - pad;
    float m[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) m[j] = -FLT_MAX;
    if (k == 3) {
      // 3x3 window, branch-free: taps outside the image are clamped onto the nearest valid row / column (a
      // duplicate tap leaves the max unchanged; every window holds at least one valid pixel), so the nine loads
      // are all in flight together
      uint4 v[9];
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        const int ih = min(max(h0 + r, 0), H - 1);
#pragma unroll
        for (int s = 0; s < 3; ++s) {
          const int iw = min(max(w0 + s, 0), W - 1);
          const size_t off = ((static_cast<size_t>(n) * H + ih) * W + iw) * C + c * vec;
          if (F32) v[r * 3 + s] = __ldg(reinterpret_cast<const uint4*>(static_cast<const float*>(in) + off));
          else v[r * 3 + s] = __ldg(reinterpret_cast<const uint4*>(static_cast<const __nv_bfloat16*>(in) + off));
        }
      }
#pragma unroll
      for (int t = 0; t < 9; ++t) {
        const uint32_t u[4] = {v[t].x, v[t].y, v[t].z, v[t].w};
        if (F32) {
#pragma unroll
          for (int e = 0; e < 4; ++e) m[e] = fmaxf(m[e], __uint_as_float(u[e]));
        } else {
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            m[2 * e] = fmaxf(m[2 * e], lo16(u[e], f16));
            m[2 * e + 1] = fmaxf(m[2 * e + 1], hi16(u[e], f16));
          }
        }
      }
    } else
    for (int r = 0; r < k; ++r) {
      const int ih = h0 + r;
      if (ih < 0 || ih >= H) continue;
      for (int s = 0; s < k; ++s) {
        const int iw = w0 + s;
        if (iw < 0 || iw >= W) continue;
        const size_t off = ((static_cast<size_t>(n) * H + ih) * W + iw) * C + c * vec;
        if (F32) {
          const float4 v = __ldg(reinterpret_cast<const float4*>(static_cast<const float*>(in) + off));
          m[0] = fmaxf(m[0], v.x); m[1] = fmaxf(m[1], v.y);
          m[2] = fmaxf(m[2], v.z); m[3] = fmaxf(m[3], v.w);
        } else {
          const uint4 v =
              __ldg(reinterpret_cast<const uint4*>(static_cast<const __nv_bfloat16*>(in) + off));
          const uint32_t u[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            m[2 * e] = fmaxf(m[2 * e], lo16(u[e], f16));
            m[2 * e + 1] = fmaxf(m[2 * e + 1], hi16(u[e], f16));
          }
        }
      }
    }
    const size_t ooff = ((static_cast<size_t>(n) * OH + oh) * OW + ow) * C + c * vec;
    if (F32) {
      *reinterpret_cast<float4*>(static_cast<float*>(out) + ooff) =
          make_float4(m[0], m[1], m[2], m[3]);
    } else {
      *reinterpret_cast<uint4*>(static_cast<__nv_bfloat16*>(out) + ooff) =
          make_uint4(pack2(m[0], m[1], f16), pack2(m[2], m[3], f16), pack2(m[4], m[5], f16), pack2(m[6], m[7], f16));
    }
  }
}

// ------------------------------------------------------------------ adaptive average pool
template <bool F32>
__global__ void __launch_bounds__(256) adaptive_avgpool_kernel(const void* __restrict__ in,
                                                                void* __restrict__ out, int N,
                                                                int H, int W, int C, int OH,
                                                                int OW, int f16) {
  const int vec = F32 ? 4 : 8;
  const int cv = C / vec;
  const long long total = static_cast<long long>(N) * OH * OW * cv;
  for (long long i = blockIdx.x * 256ll + threadIdx.x; i < total; i += 256ll * gridDim.x) {
    const int c = static_cast<int>(i % cv);
    long long pix = i / cv;
    const int ow = static_cast<int>(pix % OW);
    pix /= OW;
    const int oh = static_cast<int>(pix % OH);
    const int n = static_cast<int>(pix / OH);
    const int h0 = (oh * H) / OH, h1 = ((oh + 1) * H + OH - 1) / OH;
    const int w0 = (ow * W) / OW, w1 = ((ow + 1) * W + OW - 1) / OW;
    float a[8] = {};
    for (int ih = h0; ih < h1; ++ih)
      for (int iw = w0; iw < w1; ++iw) {
        const size_t off = ((static_cast<size_t>(n) * H + ih) * W + iw) * C + c * vec;
        if (F32) {
          const float4 v = __ldg(reinterpret_cast<const float4*>(static_cast<const float*>(in) + off));
          a[0] += v.x; a[1] += v.y; a[2] += v.z; a[3] += v.w;
        } else {
          const uint4 v =
              __ldg(reinterpret_cast<const uint4*>(static_cast<const __nv_bfloat16*>(in) + off));
          const uint32_t u[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            a[2 * e] += lo16(u[e], f16);
            a[2 * e + 1] += hi16(u[e], f16);
          }
        }
      }
    const float cnt = static_cast<float>((h1 - h0) * (w1 - w0));
#pragma unroll
    for (int j = 0; j < 8; ++j) a[j] = __fdiv_rn(a[j], cnt);
    const size_t ooff = ((static_cast<size_t>(n) * OH + oh) * OW + ow) * C + c * vec;
    if (F32) {
      *reinterpret_cast<float4*>(static_cast<float*>(out) + ooff) =
          make_float4(a[0], a[1], a[2], a[3]);
    } else {
      *reinterpret_cast<uint4*>(static_cast<__nv_bfloat16*>(out) + ooff) =
          make_uint4(pack2(a[0], a[1], f16), pack2(a[2], a[3], f16), pack2(a[4], a[5], f16), pack2(a[6], a[7], f16));
    }
  }
}

// Global average (OH = OW = 1), one CTA per image: thread (pixel group g, channel chunk c) sums
// pixels g, g + G, ... with fully coalesced 16-byte loads across c, then the G partial sums meet in
// shared memory.  The generic kernel above gives one thread all 49 pixels of a chunk (64 CTAs total).
template <bool F32>
__global__ void __launch_bounds__(256) global_avgpool_kernel(const void* __restrict__ in,
                                                              void* __restrict__ out, int HW, int C, int f16) {
  constexpr int vec = F32 ? 4 : 8;
  __shared__ float part[256][8];
  const int cv = C / vec;
  const int groups = 256 / cv;
  const int c = threadIdx.x % cv, g = threadIdx.x / cv;
  const int n = blockIdx.x;
  float a[8] = {};
  griddep_launch();
  griddep_wait();   // launched under the tail of the last conv; its output is read below, past L1 (__ldcg)
  if (g < groups) {
    // four loads in flight per thread (the sum keeps its pixel order, so the result does not change)
    for (int px0 = g; px0 < HW; px0 += 4 * groups) {
      uint4 v[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int px = px0 + i * groups;
        v[i] = make_uint4(0u, 0u, 0u, 0u);
        if (px < HW) {
          const size_t off = (static_cast<size_t>(n) * HW + px) * C + c * vec;
          if (F32) v[i] = __ldcg(reinterpret_cast<const uint4*>(static_cast<const float*>(in) + off));
          else v[i] = __ldcg(reinterpret_cast<const uint4*>(static_cast<const __nv_bfloat16*>(in) + off));
        }
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        if (px0 + i * groups >= HW) break;
        const uint32_t u[4] = {v[i].x, v[i].y, v[i].z, v[i].w};
        if (F32) {
#pragma unroll
          for (int e = 0; e < 4; ++e) a[e] += __uint_as_float(u[e]);
        } else {
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            a[2 * e] += lo16(u[e], f16);
            a[2 * e + 1] += hi16(u[e], f16);
          }
        }
      }
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) part[threadIdx.x][j] = a[j];
  __syncthreads();
  if (g == 0) {
    for (int k = 1; k < groups; ++k)
#pragma unroll
      for (int j = 0; j < 8; ++j) a[j] += part[k * cv + c][j];
    const float cnt = static_cast<float>(HW);
#pragma unroll
    for (int j = 0; j < 8; ++j) a[j] = __fdiv_rn(a[j], cnt);
    const size_t ooff = static_cast<size_t>(n) * C + c * vec;
    if (F32) {
      *reinterpret_cast<float4*>(static_cast<float*>(out) + ooff) = make_float4(a[0], a[1], a[2], a[3]);
    } else {
      *reinterpret_cast<uint4*>(static_cast<__nv_bfloat16*>(out) + ooff) =
          make_uint4(pack2(a[0], a[1], f16), pack2(a[2], a[3], f16), pack2(a[4], a[5], f16), pack2(a[6], a[7], f16));
    }
  }
}

// ------------------------------------------------------------------ softmax + top-k
// One warp per row; each lane keeps classes lane, lane+32, ... in registers (<= 32 per lane).
constexpr int kMaxPerLane = 32;
constexpr int kMaxK = 8;

__global__ void __launch_bounds__(256) softmax_topk_kernel(const float* __restrict__ logits,
                                                            int rows, int classes, int ld, int k,
                                                            int32_t* __restrict__ idx,
                                                            float* __restrict__ prob) {
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * 8 + (threadIdx.x >> 5);
  griddep_launch();
  griddep_wait();
  if (row >= rows) return;
  const float* x = logits + static_cast<size_t>(row) * ld;
  float v[kMaxPerLane];
  float mx = -FLT_MAX;
#pragma unroll
  for (int j = 0; j < kMaxPerLane; ++j) {
    const int c = lane + 32 * j;
    v[j] = (c < classes) ? __ldcg(x + c) : -FLT_MAX;
    mx = fmaxf(mx, v[j]);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  float sum = 0.f;
#pragma unroll
  for (int j = 0; j < kMaxPerLane; ++j) {
    const int c = lane + 32 * j;
    if (c < classes) sum += expf(v[j] - mx);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);

  for (int t = 0; t < k; ++t) {
    // lane-local arg-max (first occurrence = lowest class index, classes ascend with j)
    float bv = -FLT_MAX;
    int bc = 0x7fffffff;
#pragma unroll
    for (int j = 0; j < kMaxPerLane; ++j) {
      const int c = lane + 32 * j;
      if (c < classes && v[j] > bv) { bv = v[j]; bc = c; }
    }
    // warp arg-max, ties -> lower class index
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
      const int oc = __shfl_xor_sync(0xffffffffu, bc, o);
      if (ov > bv || (ov == bv && oc < bc)) { bv = ov; bc = oc; }
    }
    if (lane == 0) {
      idx[static_cast<size_t>(row) * k + t] = bc;
      prob[static_cast<size_t>(row) * k + t] = __fdiv_rn(expf(bv - mx), sum);
    }
    if ((bc & 31) == lane) {
      const int jj = bc >> 5;
#pragma unroll
      for (int j = 0; j < kMaxPerLane; ++j)
        if (j == jj) v[j] = -FLT_MAX;
    }
  }
}

int grid_for(long long work_items) {
  long long blocks = (work_items + 255) / 256;
  const long long cap = static_cast<long long>(kSMs) * 16;  // 16 resident 256-thread CTAs / SM max
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return static_cast<int>(blocks);
}

}  // namespace

cudaError_t launch_normalize_u8(const uint8_t* in, void* out, int dtype, long long pixels,
                                cudaStream_t st) {
  if (pixels % 4) return cudaErrorInvalidValue;
  const long long quads = pixels / 4;
  if (dtype == DT_F32)
    normalize_u8_kernel<true><<<grid_for(quads), 256, 0, st>>>(in, out, quads, 0);
  else
    normalize_u8_kernel<false><<<grid_for(quads), 256, 0, st>>>(in, out, quads, dtype == DT_F16);
  return cudaGetLastError();
}

cudaError_t launch_maxpool(const void* in, void* out, int dtype, int N, int H, int W, int C,
                           int k, int stride, int pad, int OH, int OW, cudaStream_t st) {
  const int vec = dtype == DT_F32 ? 4 : 8;
  if (C % vec) return cudaErrorInvalidValue;
  const long long total = static_cast<long long>(N) * OH * OW * (C / vec);
  if (dtype == DT_F32)
    maxpool_kernel<true><<<grid_for(total), 256, 0, st>>>(in, out, N, H, W, C, k, stride, pad, OH, OW, 0);
  else
    maxpool_kernel<false><<<grid_for(total), 256, 0, st>>>(in, out, N, H, W, C, k, stride, pad, OH, OW, dtype == DT_F16);
  return cudaGetLastError();
}

cudaError_t launch_adaptive_avgpool(const void* in, void* out, int dtype, int N, int H, int W,
                                    int C, int OH, int OW, cudaStream_t st) {
  const int vec = dtype == DT_F32 ? 4 : 8;
  if (C % vec) return cudaErrorInvalidValue;
  const int cv = C / vec;
  if (OH == 1 && OW == 1 && cv <= 256 && 256 % cv == 0) {
    if (dtype == DT_F32) return launch_pdl(global_avgpool_kernel<true>, dim3(N), dim3(256), 0, st, in, out, H * W, C, 0);
    return launch_pdl(global_avgpool_kernel<false>, dim3(N), dim3(256), 0, st, in, out, H * W, C, static_cast<int>(dtype == DT_F16));
  }
  const long long total = static_cast<long long>(N) * OH * OW * (C / vec);
  if (dtype == DT_F32)
    adaptive_avgpool_kernel<true><<<grid_for(total), 256, 0, st>>>(in, out, N, H, W, C, OH, OW, 0);
  else
    adaptive_avgpool_kernel<false><<<grid_for(total), 256, 0, st>>>(in, out, N, H, W, C, OH, OW, dtype == DT_F16);
  return cudaGetLastError();
}

cudaError_t launch_softmax_topk(const float* logits, int rows, int classes, int ld, int k,
                                int32_t* idx, float* prob, cudaStream_t st) {
  if (classes > 32 * kMaxPerLane || k < 1 || k > kMaxK || k > classes) return cudaErrorInvalidValue;
  return launch_pdl(softmax_topk_kernel, dim3((rows + 7) / 8), dim3(256), 0, st, logits, rows, classes, ld, k, idx, prob);
}

}  // namespace b2i
