// Convolution / FC as implicit GEMM: shared parameter block and launch entry points.
//
// GEMM view: M = N*OH*OW output pixels, N = Cout, K = R*S*C.
// Activations are NHWC; weights are [Cout_pad][K] (K-major) with K ordered (r, s, c) so a
// 64-wide K block of a C>=64 layer is one filter tap and 64 consecutive channels.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace b2i {

enum AMode : int {
  A_GATHER = 0,   // cp.async gather of 16-byte channel chunks, any geometry with C % 64 == 0
  A_IM2COL = 1,   // TMA im2col-mode tensor map, C % 64 == 0
  A_TILED = 2,    // plain 2-D TMA over an [M][K] matrix (FC layers)
  A_SMALLC = 3,   // cp.async gather of 8-byte pixels for C == 4 (first layer, K = R*Sp*4)
};

struct ConvGeom {
  int N, H, W, C;       // input, NHWC; C already padded (4 for the first layer, else multiple of 64)
  int OH, OW;           // output spatial
  int R, S, Sp;         // filter window; Sp = S padded (A_SMALLC only, else Sp == S)
  int stride, pad;
  int Cout;             // real output channels (stores are bounded by it)
  int Cout_pad;         // rows in the weight matrix (multiple of BN)
  int K;                // padded GEMM K, multiple of 64
  int M;                // N*OH*OW for the CURRENT batch
};

struct ConvArgs {
  ConvGeom g{};
  const void* x = nullptr;        // bf16 NHWC (or [M][K] for A_TILED); fp32 in the fp32 kernel
  const void* w = nullptr;        // [Cout_pad][K]
  const float* bias = nullptr;    // [Cout_pad] folded BN shift / conv bias
  const void* residual = nullptr; // optional, same layout/dtype as out
  void* out = nullptr;            // [M][ldo]
  int ldo = 0;                    // output row stride in elements
  int relu = 0;
  int out_fp32 = 0;               // bf16 kernel only: write fp32 (logits)
  int BN = 128;                   // tile N (multiple of 32, <= 256)
  int stages = 3;                 // one-tile-per-CTA kernel only; the persistent kernel sizes its own ring
  int amode = A_GATHER;
  int b_tma = 1;                  // 1: weights through TMA, 0: cp.async gather (A_GATHER / A_SMALLC only)
  int persistent = 1;             // TMA modes: persistent CTA-per-SM kernel with double-buffered TMEM
  // optional second problem on the same A operand (see conv_umma.cu): a 1x1 conv whose input pixel is
  // one tap of this conv; same Cout / BN; weights tmB2 [Cout_pad][num_kb2*64], k-blocks
  // [kb_first2, kb_first2 + num_kb2) of this conv's im2col stream
  const CUtensorMap* tmB2 = nullptr;
  const float* bias2 = nullptr;
  void* out2 = nullptr;
  int ldo2 = 0, relu2 = 0, kb_first2 = 0, num_kb2 = 0;
  int pair = 0;                   // TMA modes: CTA-pair kernel (cta_group::2); tmB must be built with BN/2 box rows
  int resident_b = 0;             // pair kernel: keep all weight k-blocks in shared memory when they fit (one N-tile)
  int reverse = 0;                // halo-patch / pair / flattened-row kernels: walk the tiles from the last to the first
  int f16 = 0;                    // tcgen05 path: 16-bit tensors (activations, weights) are IEEE half instead of bf16
  int msub = 1;                   // pair im2col kernel: 2 = two 128-row sub-tiles per CTA; tmA must then be a 256-pixel im2col map
  const CUtensorMap* tmR = nullptr;   // halo-patch kernel: 16 x 8 tile maps over the residual and the output
  const CUtensorMap* tmO = nullptr;
  int patch = 0;                  // 3x3/s1/p1, 64->64 only: halo-patch kernel (tmA = make_tmap_patch)
  int flat = 0;                   // 1: 3x3/s1/p1, one N tile: flattened-row CTA-pair kernel, resident weights (tmA = make_tmap_flat, tmB with BN/2 box rows); 2: streamed weights, several N tiles (tmA = make_tmap_flat_pitch with conv_flat_stream_pitch)
  // split-K (persistent kernel, A_TILED = FC layers): `splits` CTAs per output tile over disjoint k-block ranges;
  // ws: fp32 [splits][rows padded to 128][ldw] scratch; ctr: one int per output tile, all zero between launches
  int splits = 1, ldw = 0;
  float* ws = nullptr;
  int* ctr = nullptr;
  int a_rows = 128;               // A_TILED, persistent kernel: box rows of tmA (a multiple of 8 below 128 when the layer has one M tile)
};

// Flattened-row CTA-pair kernel (ResNet layer2 convolutions).
bool conv_flat_applicable(const ConvGeom& g, int BN);
int conv_flat_pitch(int W);
int make_tmap_flat(CUtensorMap* tm, const void* x, int N, int H, int W, int C, const char** err);
// streamed-weights form of the same kernel (a.flat = 2): several N tiles, weights through a ring, pitch W + 1 where that is a power of two
bool conv_flat_stream_applicable(const ConvGeom& g, int BN);
int conv_flat_stream_pitch(int W);
int make_tmap_flat_pitch(CUtensorMap* tm, const void* x, int N, int H, int W, int C, int pitch, const char** err);
// True when the halo-patch kernel handles this geometry (ResNet layer1 convolutions).
bool conv_patch_applicable(const ConvGeom& g);
int make_tmap_patch(CUtensorMap* tm, const void* x, int N, int H, int W, const char** err);   // 18 x 10 halo patch
int make_tmap_tile(CUtensorMap* tm, const void* x, int N, int H, int W, const char** err);    // 16 x 8 tile (residual in / result out)
int make_tmap_nhwc64(CUtensorMap* tm, const void* x, int N, int H, int W, int box_w, int box_h, const char** err);
int make_tmap_pixrow(CUtensorMap* tm, const void* x, int N, int H, int W, int C, int box_w, int stride, const char** err);

// bf16 tcgen05 path.  tmA / tmB may be null when the corresponding operand is gathered.
cudaError_t launch_conv_umma(const ConvArgs& a, const CUtensorMap* tmA, const CUtensorMap* tmB,
                             cudaStream_t st);
size_t conv_umma_smem_bytes(int BN, int stages);
int conv_umma_pick_stages(int BN);

// Fused ResNet stem (bf16 path): u8 [n][224][224][3] -> normalise -> conv1 7x7/2 + folded BN + ReLU ->
// max-pool 3x3/2 -> bf16 [n][56][56][64].  tmW: weight map over [64][256] (K order r, s<8, c<4).
// sched: two zero-initialised device words (the kernel's tile counter; it leaves them zero again), not
// shared by launches that may overlap.  in_u8 must be 4-byte aligned.  bias_host: 64 floats in HOST memory
// (they travel as kernel parameters).
cudaError_t launch_stem(const uint8_t* in_u8, void* out_bf16, const CUtensorMap* tmW,
                        const float* bias_host, int n_images, unsigned int* sched, int f16, cudaStream_t st);

// The same stem with the operands swapped (stem2.cu): weights as the MMA's A operand in tensor memory, rows = (conv row
// parity, channel), pixels along N, pooling in the epilogue's registers.  w2_dev: the [128][144] packed matrix
// stem2_pack_weights builds from the [64][256] 16-bit weights launch_stem uses.
void stem2_pack_weights(const uint16_t* w64x256, uint32_t* out128x144);
cudaError_t launch_stem2(const uint8_t* in_u8, void* out16, const uint32_t* w2_dev, const float* bias_host, int n_images,
                         unsigned int* sched, int f16, cudaStream_t st);

// Fused AlexNet stem (bf16 path): u8 [n][224][224][3] -> normalise -> features.0 11x11/4 pad 2 + bias + ReLU -> bf16
// [n][55][55][64].  tmW: weight map over [64][704] (K order: filter row, plane (R,G)/(B,0), 16 taps, 2 channels).
cudaError_t launch_alex_stem(const uint8_t* in_u8, void* out_bf16, const CUtensorMap* tmW,
                             const float* bias_host, int n_images, int f16, cudaStream_t st);

// fp32 FFMA path (the 1e-4 parity mode; also the on-device cross-check for the tcgen05 path).
cudaError_t launch_conv_fp32(const ConvArgs& a, cudaStream_t st);

// Tensor-map builders (driver entry points resolved at run time; no link against libcuda).
// Return 0 on success, else a CUresult / negative code; message in *err.
int make_tmap_weights(CUtensorMap* tm, const void* w, int Cout_pad, int K, int BN, const char** err);
int make_tmap_im2col(CUtensorMap* tm, const void* x, const ConvGeom& g, int Nmax, const char** err);
int make_tmap_im2col_px(CUtensorMap* tm, const void* x, const ConvGeom& g, int Nmax, int pixels, const char** err);   // `pixels` rows per request
int make_tmap_rows(CUtensorMap* tm, const void* a, int rows, int K, int box_rows, const char** err);

}  // namespace b2i
