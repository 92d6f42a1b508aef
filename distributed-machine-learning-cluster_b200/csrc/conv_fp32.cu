// fp32 implicit-GEMM convolution / FC on the CUDA cores (FFMA), sm_100a.
//
// This is the "fp32 parity mode" of the path (BASELINE.json: logits within 1e-4 of the reference's
// fp32 CPU forward, reference src/services.rs:493).  TF32 tensor-core math keeps only 10 mantissa
// bits and cannot meet that bound on O(1..10) logits, so this mode accumulates in true fp32.
// Same geometry block, K ordering (r, s, c) and [Cout_pad][K] weight layout as the tcgen05 kernel,
// with fp32 activations (NHWC) and weights.
#include "conv.h"

namespace b2i {
namespace {

constexpr int TM = 64, TN = 64, TK = 16;

struct FP {
  int H, W, C, OH, OW, R, S, Sp, stride, pad, Cout, K, M, ldo, relu;
  const float* x;
  const float* w;
  const float* bias;
  const float* residual;
  float* out;
};

__global__ void __launch_bounds__(256) conv_fp32_kernel(const FP p) {
  __shared__ __align__(16) float As[2][TK][TM + 4];
  __shared__ __align__(16) float Bs[2][TK][TN + 4];

  const int t = threadIdx.x;
  const int m0 = blockIdx.x * TM;  // M tiles on x: conv1 at batch 1024 has > 65535 of them
  const int n0 = blockIdx.y * TN;
  const int kk = t & 15;   // k inside the K step this thread loads
  const int lr = t >> 4;   // rows lr + 16*i
  const int tx = t & 15, ty = t >> 4;

  // per-row pixel origin for the 4 A rows this thread gathers
  int ih0[4], iw0[4];
  const float* rowbase[4];
  const int ohw = p.OH * p.OW;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int m = m0 + lr + 16 * i;
    if (m < p.M) {
      const int n = m / ohw;
      const int rem = m - n * ohw;
      const int oh = rem / p.OW;
      const int ow = rem - oh * p.OW;
      ih0[i] = oh * p.stride - p.pad;
      iw0[i] = ow * p.stride - p.pad;
      rowbase[i] = p.x + static_cast<size_t>(n) * p.H * p.W * p.C;
    } else {
      ih0[i] = -(1 << 28);
      iw0[i] = 0;
      rowbase[i] = p.x;
    }
  }

  float acc[4][4] = {};
  float ra[4], rb[4];
  const int nk = p.K / TK;

  auto gload = [&](int kt) {
    const int k = kt * TK + kk;
    const int c = k % p.C;
    const int tap = k / p.C;
    const int s = tap % p.Sp;
    const int r = tap / p.Sp;
    const bool tap_ok = (r < p.R) && (s < p.S);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int ih = ih0[i] + r, iw = iw0[i] + s;
      const bool ok = tap_ok && ih >= 0 && ih < p.H && iw >= 0 && iw < p.W;
      ra[i] = ok ? __ldg(rowbase[i] + (static_cast<size_t>(ih) * p.W + iw) * p.C + c) : 0.f;
      rb[i] = __ldg(p.w + static_cast<size_t>(n0 + lr + 16 * i) * p.K + k);
    }
  };
  auto sstore = [&](int buf) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      As[buf][kk][lr + 16 * i] = ra[i];
      Bs[buf][kk][lr + 16 * i] = rb[i];
    }
  };

  gload(0);
  sstore(0);
  __syncthreads();
  for (int kt = 0; kt < nk; ++kt) {
    const int buf = kt & 1;
    if (kt + 1 < nk) gload(kt + 1);
#pragma unroll
    for (int k = 0; k < TK; ++k) {
      const float4 a4 = *reinterpret_cast<const float4*>(&As[buf][k][ty * 4]);
      const float4 b4 = *reinterpret_cast<const float4*>(&Bs[buf][k][tx * 4]);
      const float a[4] = {a4.x, a4.y, a4.z, a4.w};
      const float b[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    if (kt + 1 < nk) {
      sstore(buf ^ 1);
      __syncthreads();
    }
  }

#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int m = m0 + ty * 4 + i;
    if (m >= p.M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tx * 4 + j;
      if (n >= p.Cout) continue;
      const size_t off = static_cast<size_t>(m) * p.ldo + n;
      float v = acc[i][j] + __ldg(p.bias + n);
      if (p.residual) v += __ldg(p.residual + off);
      if (p.relu) v = fmaxf(v, 0.f);
      p.out[off] = v;
    }
  }
}

}  // namespace

cudaError_t launch_conv_fp32(const ConvArgs& a, cudaStream_t st) {
  const ConvGeom& g = a.g;
  if (g.K % TK || g.Cout_pad % TN || g.M <= 0) return cudaErrorInvalidValue;
  FP p;
  p.H = g.H; p.W = g.W; p.C = g.C; p.OH = g.OH; p.OW = g.OW; p.R = g.R; p.S = g.S; p.Sp = g.Sp;
  p.stride = g.stride; p.pad = g.pad; p.Cout = g.Cout; p.K = g.K; p.M = g.M; p.ldo = a.ldo;
  p.relu = a.relu;
  p.x = static_cast<const float*>(a.x);
  p.w = static_cast<const float*>(a.w);
  p.bias = a.bias;
  p.residual = static_cast<const float*>(a.residual);
  p.out = static_cast<float*>(a.out);
  dim3 grid((g.M + TM - 1) / TM, g.Cout_pad / TN);
  conv_fp32_kernel<<<grid, 256, 0, st>>>(p);
  return cudaGetLastError();
}

}  // namespace b2i
