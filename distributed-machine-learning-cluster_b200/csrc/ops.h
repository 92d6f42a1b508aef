// Bandwidth-bound kernels of the path: preprocessing, pooling, softmax + top-k.
// Each launcher cites the reference call it replaces; all layouts are NHWC.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace b2i {

enum DType : int { DT_BF16 = 0, DT_F32 = 1, DT_F16 = 2 };   // DT_F16: same kernels as DT_BF16, IEEE half storage

// u8 [N,224,224,3] (HWC, already resized + centre-cropped) -> NHWC4, (x/255 - mean)/std, pad
// channel = 0.  Replaces the normalise half of tch imagenet::load_image_and_resize
// (reference src/services.rs:492).
cudaError_t launch_normalize_u8(const uint8_t* in, void* out, int dtype, long long pixels,
                                cudaStream_t st);

// max_pool2d(k, stride, pad, dilation 1, ceil_mode false), NHWC, C % 8 == 0 (bf16) / C % 4 (fp32).
cudaError_t launch_maxpool(const void* in, void* out, int dtype, int N, int H, int W, int C,
                           int k, int stride, int pad, int OH, int OW, cudaStream_t st);

// adaptive_avg_pool2d to (OH, OW): window [floor(i*in/out), ceil((i+1)*in/out)), fp32 accumulate.
cudaError_t launch_adaptive_avgpool(const void* in, void* out, int dtype, int N, int H, int W,
                                    int C, int OH, int OW, cudaStream_t st);

// softmax(-1) over `classes` fp32 logits + top-k (largest, sorted, ties -> lower index), one warp
// per row.  Replaces `.softmax(-1, Float)` + imagenet::top (reference src/services.rs:493-494).
cudaError_t launch_softmax_topk(const float* logits, int rows, int classes, int ld, int k,
                                int32_t* idx, float* prob, cudaStream_t st);

}  // namespace b2i
