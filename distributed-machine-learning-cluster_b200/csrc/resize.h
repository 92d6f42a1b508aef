// Resize + centre crop of a decoded image, restating what tch's
// imagenet::load_image_and_resize(path, 224, 224) does after the JPEG decode (reference
// src/services.rs:492): target size rule and crop from SURVEY.md Appendix B.1 / 8c R4, resampling
// as stb_image_resize's stbir_resize_uint8 defaults (8c R3): Catmull-Rom when an axis grows,
// Mitchell-Netravali (B = C = 1/3, support widened by 1/scale, re-normalised) when it shrinks,
// edge clamp, linear u8 <-> float, horizontal pass first, fp32 arithmetic, taps summed in
// ascending input order.  [tch-recall]: the stb source is not on this machine, parity unpinned.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

namespace b2i {

// Gather form of one resampling axis restricted to the 224 output positions that survive the crop:
// output j reads taps [first[j], first[j] + count[j]) of idx/coef.
struct AxisPlan {
  int in_size = 0, out_size = 0, crop = 0;
  std::vector<int> first, count;
  std::vector<int> idx;      // clamped input positions
  std::vector<float> coef;
};

struct ResizeGeom { int rw, rh, crop_x, crop_y; };
ResizeGeom resize_target(int h, int w);
void plan_axis(int in_size, int out_size, int crop, AxisPlan* plan);

// Host image in, host 224x224x3 out; runs both passes on the current device. `st` may be null.
bool resize_crop_224_host(const uint8_t* image, int h, int w, uint8_t* out224, cudaStream_t st,
                          std::string* err);

// One image of a batch whose buffers the caller manages (predict_raw): every pointer is a device
// pointer; enqueues the two passes on `st`.
struct ResizeJob {
  const uint8_t* d_img = nullptr;
  int h = 0, w = 0;
  const int *xf = nullptr, *xc = nullptr, *xi = nullptr;
  const float* xw = nullptr;
  const int *yf = nullptr, *yc = nullptr, *yi = nullptr;
  const float* yw = nullptr;
  float* d_tmp = nullptr;     // float [h][224][3]
  uint8_t* d_out = nullptr;   // u8 [224][224][3]
};
cudaError_t launch_resize_job(const ResizeJob& j, cudaStream_t st);
void plan_image(int h, int w, AxisPlan* px, AxisPlan* py);

// Same two kernels, but no allocation of any kind on the GPU side: device in/out pointers from the
// caller, scratch in a static device array (limits: resize_static_max_rows() input rows, 16384
// resampling taps per axis).  Enqueues on `st` and returns without synchronising.
bool resize_crop_224_static(const uint8_t* d_image, int h, int w, uint8_t* d_out224, cudaStream_t st,
                            std::string* err);
int resize_static_max_rows();

}  // namespace b2i
