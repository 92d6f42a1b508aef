// See resize.h.
#include "resize.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <mutex>

namespace b2i {

namespace {

// stb_image_resize filter kernels, evaluated in float exactly in this operation order.
float catmullrom(float x) {
  x = std::fabs(x);
  if (x < 1.0f) return 1.0f - x * x * (2.5f - 1.5f * x);
  if (x < 2.0f) return 2.0f - x * (4.0f + x * (0.5f * x - 2.5f));
  return 0.0f;
}
float mitchell(float x) {
  x = std::fabs(x);
  if (x < 1.0f) return (16.0f + x * x * (21.0f * x - 36.0f)) / 18.0f;
  if (x < 2.0f) return (32.0f + x * (-60.0f + x * (36.0f - 7.0f * x))) / 18.0f;
  return 0.0f;
}
inline int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

}  // namespace

ResizeGeom resize_target(int h, int w) {
  ResizeGeom g;
  if (static_cast<long long>(w) * 224 == static_cast<long long>(h) * 224) {
    g.rw = 224; g.rh = 224; g.crop_x = 0; g.crop_y = 0;
    return g;
  }
  const double ratio = std::fmax(224.0 / static_cast<double>(w), 224.0 / static_cast<double>(h));
  long long rw = static_cast<long long>(ratio * static_cast<double>(w));
  long long rh = static_cast<long long>(ratio * static_cast<double>(h));
  if (rw < 224) rw = 224;
  if (rh < 224) rh = 224;
  g.rw = static_cast<int>(rw);
  g.rh = static_cast<int>(rh);
  g.crop_x = (g.rw - 224) / 2;
  g.crop_y = (g.rh - 224) / 2;
  return g;
}

void plan_axis(int in_size, int out_size, int crop, AxisPlan* p) {
  p->in_size = in_size; p->out_size = out_size; p->crop = crop;
  p->first.assign(224, 0);
  p->count.assign(224, 0);
  p->idx.clear();
  p->coef.clear();
  const float scale = static_cast<float>(out_size) / static_cast<float>(in_size);
  if (scale > 1.0f) {
    // growing axis: per output pixel, Catmull-Rom taps around its centre in input space
    const float radius = 2.0f * scale;
    for (int jj = 0; jj < 224; ++jj) {
      const int j = jj + crop;
      const float centre = static_cast<float>(j) + 0.5f;
      const float lo = (centre - radius) / scale;
      const float hi = (centre + radius) / scale;
      const float in_centre = centre / scale;
      int first = static_cast<int>(std::floor(lo + 0.5f));
      const int last = static_cast<int>(std::floor(hi - 0.5f));
      std::vector<float> c;
      for (int i = first; i <= last; ++i) {
        const float pc = static_cast<float>(i) + 0.5f;
        const float v = catmullrom(in_centre - pc);
        if (c.empty() && v == 0.0f) { first = i + 1; continue; }  // stb drops a leading zero tap
        c.push_back(v);
      }
      float total = 0.0f;
      for (float v : c) total += v;
      const float fs = 1.0f / total;
      while (!c.empty() && c.back() * fs == 0.0f) c.pop_back();
      p->first[jj] = static_cast<int>(p->idx.size());
      p->count[jj] = static_cast<int>(c.size());
      for (size_t i = 0; i < c.size(); ++i) {
        p->idx.push_back(clampi(first + static_cast<int>(i), 0, in_size - 1));
        p->coef.push_back(c[i] * fs);
      }
    }
    return;
  }
  // shrinking (or equal) axis: per INPUT pixel (plus clamped margin pixels), Mitchell weights onto
  // the output pixels inside its widened footprint; then each output's weights are normalised.
  const float radius = 2.0f / scale;
  const int margin = static_cast<int>(std::ceil(2.0f * 2.0f / scale)) / 2;
  const int ncontrib = in_size + 2 * margin;
  std::vector<int> n0(ncontrib), n1(ncontrib), off(ncontrib);
  std::vector<float> cf;
  for (int n = 0; n < ncontrib; ++n) {
    const int in_pixel = n - margin;
    const float centre = static_cast<float>(in_pixel) + 0.5f;
    const float lo = (centre - radius) * scale;
    const float hi = (centre + radius) * scale;
    const float out_centre = centre * scale;
    const int first = static_cast<int>(std::floor(lo + 0.5f));
    int last = static_cast<int>(std::floor(hi - 0.5f));
    off[n] = static_cast<int>(cf.size());
    for (int i = first; i <= last; ++i) {
      const float pc = static_cast<float>(i) + 0.5f;
      cf.push_back(mitchell(pc - out_centre) * scale);
    }
    while (last >= first && cf[off[n] + (last - first)] == 0.0f) { --last; }
    n0[n] = first;
    n1[n] = last;
  }
  // n0 and n1 never decrease with n, and j grows with jj: the contributors of output j are a window
  // [lo, ...) that only moves forward.  Sums run over ascending n exactly as a full scan would.
  int lo = 0;
  for (int jj = 0; jj < 224; ++jj) {
    const int j = jj + crop;
    while (lo < ncontrib && n1[lo] < j && n0[lo] <= j) ++lo;
    float total = 0.0f;
    for (int n = lo; n < ncontrib; ++n) {
      if (j >= n0[n] && j <= n1[n]) total += cf[off[n] + (j - n0[n])];
      else if (j < n0[n]) break;
    }
    const float fs = 1.0f / total;
    p->first[jj] = static_cast<int>(p->idx.size());
    int cnt = 0;
    for (int n = lo; n < ncontrib; ++n) {
      if (j >= n0[n] && j <= n1[n]) {
        const float v = cf[off[n] + (j - n0[n])] * fs;
        if (v != 0.0f) {
          p->idx.push_back(clampi(n - margin, 0, in_size - 1));
          p->coef.push_back(v);
          ++cnt;
        }
      } else if (j < n0[n]) break;
    }
    p->count[jj] = cnt;
  }
}

namespace {

// horizontal pass: u8 [h][w][3] -> float [h][224][3]
__global__ void __launch_bounds__(256) resize_h_kernel(const uint8_t* __restrict__ in, int h, int w,
                                                       const int* __restrict__ first,
                                                       const int* __restrict__ count,
                                                       const int* __restrict__ idx,
                                                       const float* __restrict__ coef,
                                                       float* __restrict__ tmp) {
  const long long total = static_cast<long long>(h) * 224;
  for (long long i = blockIdx.x * 256ll + threadIdx.x; i < total; i += 256ll * gridDim.x) {
    const int x = static_cast<int>(i % 224);
    const int y = static_cast<int>(i / 224);
    const uint8_t* row = in + static_cast<size_t>(y) * w * 3;
    const int f = first[x], n = count[x];
    float a0 = 0.f, a1 = 0.f, a2 = 0.f;
    for (int t = 0; t < n; ++t) {
      const uint8_t* px = row + idx[f + t] * 3;
      const float c = coef[f + t];
      a0 = __fadd_rn(a0, __fmul_rn(__fdiv_rn(static_cast<float>(px[0]), 255.0f), c));
      a1 = __fadd_rn(a1, __fmul_rn(__fdiv_rn(static_cast<float>(px[1]), 255.0f), c));
      a2 = __fadd_rn(a2, __fmul_rn(__fdiv_rn(static_cast<float>(px[2]), 255.0f), c));
    }
    float* o = tmp + i * 3;
    o[0] = a0; o[1] = a1; o[2] = a2;
  }
}

__device__ __forceinline__ uint8_t encode_u8(float v) {
  v = v < 0.f ? 0.f : (v > 1.f ? 1.f : v);
  return static_cast<uint8_t>(static_cast<int>(__fadd_rn(__fmul_rn(v, 255.0f), 0.5f)));
}

// vertical pass: float [h][224][3] -> u8 [224][224][3]
__global__ void __launch_bounds__(256) resize_v_kernel(const float* __restrict__ tmp,
                                                       const int* __restrict__ first,
                                                       const int* __restrict__ count,
                                                       const int* __restrict__ idx,
                                                       const float* __restrict__ coef,
                                                       uint8_t* __restrict__ out) {
  const int i = blockIdx.x * 256 + threadIdx.x;  // over 224*224*3 scalars, channel fastest
  if (i >= 224 * 224 * 3) return;
  const int y = i / (224 * 3);
  const int xc = i - y * (224 * 3);
  const int f = first[y], n = count[y];
  float a = 0.f;
  for (int t = 0; t < n; ++t)
    a = __fadd_rn(a, __fmul_rn(tmp[static_cast<size_t>(idx[f + t]) * (224 * 3) + xc], coef[f + t]));
  out[i] = encode_u8(a);
}

}  // namespace

bool resize_crop_224_host(const uint8_t* image, int h, int w, uint8_t* out224, cudaStream_t st,
                          std::string* err) {
  const ResizeGeom g = resize_target(h, w);
  AxisPlan px, py;
  plan_axis(w, g.rw, g.crop_x, &px);
  plan_axis(h, g.rh, g.crop_y, &py);
  struct Free { std::vector<void*> v; ~Free() { for (void* p : v) cudaFree(p); } } fr;
  auto dalloc = [&](size_t bytes, void** p) {
    cudaError_t e = cudaMalloc(p, bytes ? bytes : 16);
    if (e != cudaSuccess) { *err = std::string("cudaMalloc: ") + cudaGetErrorString(e); return false; }
    fr.v.push_back(*p);
    return true;
  };
  auto up = [&](const void* src, size_t bytes, void** p) {
    if (!dalloc(bytes, p)) return false;
    cudaError_t e = cudaMemcpyAsync(*p, src, bytes, cudaMemcpyHostToDevice, st);
    if (e != cudaSuccess) { *err = std::string("cudaMemcpy: ") + cudaGetErrorString(e); return false; }
    return true;
  };
  void *d_in, *d_tmp, *d_out, *xf, *xc, *xi, *xw, *yf, *yc, *yi, *yw;
  if (!up(image, static_cast<size_t>(h) * w * 3, &d_in)) return false;
  if (!dalloc(static_cast<size_t>(h) * 224 * 3 * 4, &d_tmp) || !dalloc(224 * 224 * 3, &d_out)) return false;
  if (!up(px.first.data(), 224 * 4, &xf) || !up(px.count.data(), 224 * 4, &xc) ||
      !up(px.idx.data(), px.idx.size() * 4, &xi) || !up(px.coef.data(), px.coef.size() * 4, &xw) ||
      !up(py.first.data(), 224 * 4, &yf) || !up(py.count.data(), 224 * 4, &yc) ||
      !up(py.idx.data(), py.idx.size() * 4, &yi) || !up(py.coef.data(), py.coef.size() * 4, &yw))
    return false;
  long long blocks = (static_cast<long long>(h) * 224 + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  resize_h_kernel<<<static_cast<int>(blocks), 256, 0, st>>>(
      static_cast<const uint8_t*>(d_in), h, w, static_cast<const int*>(xf), static_cast<const int*>(xc),
      static_cast<const int*>(xi), static_cast<const float*>(xw), static_cast<float*>(d_tmp));
  resize_v_kernel<<<(224 * 224 * 3 + 255) / 256, 256, 0, st>>>(
      static_cast<const float*>(d_tmp), static_cast<const int*>(yf), static_cast<const int*>(yc),
      static_cast<const int*>(yi), static_cast<const float*>(yw), static_cast<uint8_t*>(d_out));
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess) e = cudaMemcpyAsync(out224, d_out, 224 * 224 * 3, cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  if (e != cudaSuccess) { *err = std::string("resize kernels: ") + cudaGetErrorString(e); return false; }
  return true;
}


void plan_image(int h, int w, AxisPlan* px, AxisPlan* py) {
  const ResizeGeom g = resize_target(h, w);
  plan_axis(w, g.rw, g.crop_x, px);
  plan_axis(h, g.rh, g.crop_y, py);
}

cudaError_t launch_resize_job(const ResizeJob& j, cudaStream_t st) {
  long long blocks = (static_cast<long long>(j.h) * 224 + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  resize_h_kernel<<<static_cast<int>(blocks), 256, 0, st>>>(j.d_img, j.h, j.w, j.xf, j.xc, j.xi, j.xw, j.d_tmp);
  resize_v_kernel<<<(224 * 224 * 3 + 255) / 256, 256, 0, st>>>(j.d_tmp, j.yf, j.yc, j.yi, j.yw, j.d_out);
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------
// Allocation-free variant.  Some hosts of this library may not allocate GPU memory at run time: a
// serving process whose memory pool is carved up once at start-up, or a tenant under a fixed memory
// budget.  For them the scratch of the resize (the
// horizontally resampled rows and the two resampling plans) lives in one file-scope device array
// instead of cudaMalloc'ed buffers.  Every GPU that loads the library carries the array until the
// process exits.
constexpr int kStaticMaxH = 1024;        // tallest input image the scratch holds
constexpr int kStaticMaxTaps = 16384;    // resampling taps per axis (224 outputs x taps each)
constexpr size_t kStaticTmpBytes = static_cast<size_t>(kStaticMaxH) * 224 * 3 * sizeof(float);
constexpr size_t kStaticAxisBytes = 2 * 224 * sizeof(int) + kStaticMaxTaps * (sizeof(int) + sizeof(float));
constexpr size_t kStaticScratchBytes = kStaticTmpBytes + 2 * kStaticAxisBytes;   // 3,018,240 bytes

static __device__ __align__(256) unsigned char g_resize_scratch[kStaticScratchBytes];

bool resize_crop_224_static(const uint8_t* d_image, int h, int w, uint8_t* d_out224, cudaStream_t st,
                            std::string* err) {
  // In host code `g_resize_scratch` is only a handle for the symbol, not an address a kernel can
  // use; the device address is looked up once per GPU, with that GPU current.
  static unsigned char* scratch_of_device[64] = {};
  static std::mutex mu;
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= 64) { *err = "no current CUDA device"; return false; }
  unsigned char* scratch = nullptr;
  {
    std::lock_guard<std::mutex> lock(mu);
    if (!scratch_of_device[dev]) {
      unsigned char* ptr = nullptr;
      e = cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_resize_scratch);
      if (e != cudaSuccess || !ptr) {
        *err = std::string("cudaGetSymbolAddress(resize scratch): ") + cudaGetErrorString(e);
        return false;
      }
      scratch_of_device[dev] = ptr;
    }
    scratch = scratch_of_device[dev];
  }
  if (h < 1 || w < 1) { *err = "empty image"; return false; }
  if (h > kStaticMaxH) {
    *err = "image height " + std::to_string(h) + " exceeds the static scratch limit of " +
           std::to_string(kStaticMaxH) + " rows; use b2i_resize_crop_u8";
    return false;
  }
  const ResizeGeom g = resize_target(h, w);
  AxisPlan px, py;
  plan_axis(w, g.rw, g.crop_x, &px);
  plan_axis(h, g.rh, g.crop_y, &py);
  if (static_cast<int>(px.idx.size()) > kStaticMaxTaps || static_cast<int>(py.idx.size()) > kStaticMaxTaps) {
    *err = "resampling plan of " + std::to_string(std::max(px.idx.size(), py.idx.size())) +
           " taps exceeds the static scratch limit of " + std::to_string(kStaticMaxTaps) +
           " taps per axis (image " + std::to_string(h) + "x" + std::to_string(w) + "); use b2i_resize_crop_u8";
    return false;
  }
  // Refuse to run under stream capture: the plans below depend on this image's size and are uploaded from
  // host memory that dies with this call, so a captured copy would replay garbage.
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(st, &cap) == cudaSuccess && cap != cudaStreamCaptureStatusNone) {
    *err = "the stream is capturing; b2i_resize_crop_u8_static cannot be recorded into a CUDA graph";
    return false;
  }
  cudaGetLastError();
  // carve the array: [tmp rows][plans]; the plans are packed [xf xc yf yc | xi | yi | xw | yw] so that they
  // go up in ONE copy.  Every byte a kernel reads below is written earlier in this same call, so nothing
  // has to be cleared between calls.
  float* d_tmp = reinterpret_cast<float*>(scratch);
  const size_t nx = px.idx.size(), ny = py.idx.size();
  std::vector<int> pack(4 * 224 + 2 * (nx + ny));
  std::memcpy(&pack[0], px.first.data(), 224 * sizeof(int));
  std::memcpy(&pack[224], px.count.data(), 224 * sizeof(int));
  std::memcpy(&pack[448], py.first.data(), 224 * sizeof(int));
  std::memcpy(&pack[672], py.count.data(), 224 * sizeof(int));
  std::memcpy(&pack[896], px.idx.data(), nx * sizeof(int));
  std::memcpy(&pack[896 + nx], py.idx.data(), ny * sizeof(int));
  std::memcpy(&pack[896 + nx + ny], px.coef.data(), nx * sizeof(float));
  std::memcpy(&pack[896 + 2 * nx + ny], py.coef.data(), ny * sizeof(float));
  int* d_plan = reinterpret_cast<int*>(scratch + kStaticTmpBytes);
  const int *xf = d_plan, *xc = d_plan + 224, *yf = d_plan + 448, *yc = d_plan + 672;
  const int *xi = d_plan + 896, *yi = xi + nx;
  const float *xw = reinterpret_cast<const float*>(yi + ny), *yw = xw + nx;
  // `pack` is pageable and local: a host-to-device cudaMemcpyAsync from pageable memory returns only after
  // the runtime has copied the source into its own staging buffer, so the vector may die when this
  // function returns.  (That staging step is the one host-side wait in this call.)
  e = cudaMemcpyAsync(d_plan, pack.data(), pack.size() * sizeof(int), cudaMemcpyHostToDevice, st);
  if (e != cudaSuccess) { *err = std::string("plan upload: ") + cudaGetErrorString(e); return false; }
  long long blocks = (static_cast<long long>(h) * 224 + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  resize_h_kernel<<<static_cast<int>(blocks), 256, 0, st>>>(d_image, h, w, xf, xc, xi, xw, d_tmp);
  resize_v_kernel<<<(224 * 224 * 3 + 255) / 256, 256, 0, st>>>(d_tmp, yf, yc, yi, yw, d_out224);
  e = cudaGetLastError();
  if (e != cudaSuccess) { *err = std::string("resize kernels: ") + cudaGetErrorString(e); return false; }
  return true;
}

int resize_static_max_rows() { return kStaticMaxH; }

}  // namespace b2i
