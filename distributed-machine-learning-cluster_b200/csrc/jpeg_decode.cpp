// See jpeg_decode.h.
#include "jpeg_decode.h"

#include <dlfcn.h>
#include <nvjpeg.h>   // types and prototypes only: the library is reached through dlopen, never linked

#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <mutex>
#include <thread>
#include <vector>

#include "../../include/b200_infer.h"
#include "predict.h"

namespace b2i {

namespace {

// The entry points used, resolved once per process.
#define NVJ_SYMBOLS(X)               \
  X(nvjpegCreateEx)                  \
  X(nvjpegDestroy)                   \
  X(nvjpegJpegStateCreate)           \
  X(nvjpegJpegStateDestroy)          \
  X(nvjpegGetImageInfo)              \
  X(nvjpegDecode)                    \
  X(nvjpegDecodeBatchedInitialize)   \
  X(nvjpegDecodeBatched)             \
  X(nvjpegJpegStreamCreate)          \
  X(nvjpegJpegStreamDestroy)         \
  X(nvjpegJpegStreamParse)           \
  X(nvjpegDecodeBatchedSupported)

struct Api {
  void* dl = nullptr;
#define X(name) decltype(&::name) name = nullptr;
  NVJ_SYMBOLS(X)
#undef X
  std::string err;
};

Api* api() {
  static Api a;
  static std::once_flag once;
  std::call_once(once, [] {
    const char* env = std::getenv("B2I_NVJPEG_LIB");
    const char* names[] = {env, "libnvjpeg.so.12", "/usr/local/cuda/lib64/libnvjpeg.so.12",
                           "/usr/local/cuda/targets/x86_64-linux/lib/libnvjpeg.so.12", "libnvjpeg.so"};
    for (const char* n : names) {
      if (!n || !*n) continue;
      a.dl = dlopen(n, RTLD_NOW | RTLD_LOCAL);
      if (a.dl) break;
    }
    if (!a.dl) {
      a.err = "nvJPEG is not installed (libnvjpeg.so.12 not found; set B2I_NVJPEG_LIB to its path)";
      return;
    }
#define X(name)                                                           \
  a.name = reinterpret_cast<decltype(a.name)>(dlsym(a.dl, #name));        \
  if (!a.name && a.err.empty()) a.err = "libnvjpeg lacks " #name;
    NVJ_SYMBOLS(X)
#undef X
  });
  return &a;
}

struct Engine {
  nvjpegHandle_t handle = nullptr;
  nvjpegJpegState_t state = nullptr;
  int batch = -1;   // batch size of the last nvjpegDecodeBatchedInitialize
};

bool engine_open(Api* a, nvjpegBackend_t backend, Engine* e) {
  if (a->nvjpegCreateEx(backend, nullptr, nullptr, 0, &e->handle) != NVJPEG_STATUS_SUCCESS) {
    e->handle = nullptr;
    cudaGetLastError();
    return false;
  }
  if (a->nvjpegJpegStateCreate(e->handle, &e->state) != NVJPEG_STATUS_SUCCESS) {
    a->nvjpegDestroy(e->handle);
    e->handle = nullptr;
    e->state = nullptr;
    cudaGetLastError();
    return false;
  }
  return true;
}

void engine_close(Api* a, Engine* e) {
  if (e->state) a->nvjpegJpegStateDestroy(e->state);
  if (e->handle) a->nvjpegDestroy(e->handle);
  *e = Engine();
}

}  // namespace

struct JpegDecoder {
  Engine hw, gpu, cpu;          // NVJPG engines (batched) / GPU Huffman (batched) / hybrid, per image
  std::vector<Engine> pool;     // further hybrid engines, one per helper thread (entropy decode is host work)
  nvjpegJpegStream_t parsed = nullptr;
  int device = 0;
  std::string backend;
};

JpegDecoder* jpeg_decoder_create(std::string* err) {
  Api* a = api();
  if (!a->err.empty()) { *err = a->err; return nullptr; }
  auto d = new JpegDecoder();
  // B2I_JPEG_BACKEND = hardware | gpu | cpu restricts the choice (A/B measurements); default: all that open
  const char* want = std::getenv("B2I_JPEG_BACKEND");
  const std::string w = want ? want : "";
  if (!engine_open(a, NVJPEG_BACKEND_HYBRID, &d->cpu) && !engine_open(a, NVJPEG_BACKEND_DEFAULT, &d->cpu)) {
    *err = "nvjpegCreateEx failed for the hybrid backend";
    delete d;
    return nullptr;
  }
  if (w.empty() || w == "hardware") engine_open(a, NVJPEG_BACKEND_HARDWARE, &d->hw);
  if (w.empty() || w == "gpu") engine_open(a, NVJPEG_BACKEND_GPU_HYBRID, &d->gpu);
  if (a->nvjpegJpegStreamCreate(d->cpu.handle, &d->parsed) != NVJPEG_STATUS_SUCCESS) d->parsed = nullptr;
  cudaGetDevice(&d->device);
  // helper engines for the images no batched engine takes: B2I_JPEG_THREADS (default: half the host's
  // hardware threads, at most 16; all of that work when neither batched engine opened)
  const char* te = std::getenv("B2I_JPEG_THREADS");
  int threads = te ? std::atoi(te) : static_cast<int>(std::thread::hardware_concurrency() / 2);
  threads = threads < 1 ? 1 : (threads > 16 ? 16 : threads);
  for (int t = 1; t < threads; ++t) {
    Engine e;
    if (!engine_open(a, NVJPEG_BACKEND_HYBRID, &e)) break;
    d->pool.push_back(e);
  }
  d->backend = std::string(d->hw.handle ? "hardware+" : "") + (d->gpu.handle ? "gpu_hybrid+" : "") + "hybrid";
  return d;
}

void jpeg_decoder_destroy(JpegDecoder* d) {
  if (!d) return;
  Api* a = api();
  if (d->parsed) a->nvjpegJpegStreamDestroy(d->parsed);
  engine_close(a, &d->hw);
  engine_close(a, &d->gpu);
  engine_close(a, &d->cpu);
  for (Engine& e : d->pool) engine_close(a, &e);
  delete d;
}

const char* jpeg_decoder_backend(const JpegDecoder* d) { return d ? d->backend.c_str() : "none"; }

void jpeg_probe(JpegDecoder* d, const uint8_t* const* data, const size_t* len, int n, int* h, int* w,
                int* status) {
  Api* a = api();
  for (int i = 0; i < n; ++i) {
    h[i] = w[i] = 0;
    status[i] = B2I_EDECODE;
    if (!data[i] || len[i] < 4) continue;
    int ncomp = 0;
    nvjpegChromaSubsampling_t css;
    int ws[NVJPEG_MAX_COMPONENT] = {0}, hs[NVJPEG_MAX_COMPONENT] = {0};
    if (a->nvjpegGetImageInfo(d->cpu.handle, data[i], len[i], &ncomp, &css, ws, hs) != NVJPEG_STATUS_SUCCESS) continue;
    if ((ncomp != 1 && ncomp != 3) || ws[0] < 1 || hs[0] < 1) continue;
    if (ws[0] > kMaxRawSide || hs[0] > kMaxRawSide) { status[i] = B2I_EINVAL; continue; }
    w[i] = ws[0];
    h[i] = hs[0];
    status[i] = B2I_OK;
  }
}

bool jpeg_decode_batch(JpegDecoder* d, const uint8_t* const* data, const size_t* len, int n,
                       uint8_t* const* d_out, const int* h, const int* w, int* status, cudaStream_t st,
                       std::string* err) {
  Api* a = api();
  // route every image to the first engine that accepts its coding
  std::vector<int> to_hw, to_gpu, to_cpu;
  for (int i = 0; i < n; ++i) {
    if (status[i] != B2I_OK) continue;
    int where = 2;
    if ((d->hw.handle || d->gpu.handle) && d->parsed &&
        a->nvjpegJpegStreamParse(d->cpu.handle, data[i], len[i], 0, 0, d->parsed) == NVJPEG_STATUS_SUCCESS) {
      int ok = -1;
      if (d->hw.handle && a->nvjpegDecodeBatchedSupported(d->hw.handle, d->parsed, &ok) == NVJPEG_STATUS_SUCCESS && ok == 0) where = 0;
      else if (d->gpu.handle && a->nvjpegDecodeBatchedSupported(d->gpu.handle, d->parsed, &ok) == NVJPEG_STATUS_SUCCESS && ok == 0) where = 1;
    }
    (where == 0 ? to_hw : where == 1 ? to_gpu : to_cpu).push_back(i);
  }
  auto image_of = [&](int i) {
    nvjpegImage_t im;
    std::memset(&im, 0, sizeof im);
    im.channel[0] = d_out[i];
    im.pitch[0] = static_cast<size_t>(w[i]) * 3;
    return im;
  };
  auto batched = [&](Engine& e, std::vector<int>& set) {
    if (set.empty()) return;
    const int cnt = static_cast<int>(set.size());
    std::vector<const unsigned char*> ptr(cnt);
    std::vector<size_t> sz(cnt);
    std::vector<nvjpegImage_t> img(cnt);
    for (int j = 0; j < cnt; ++j) { ptr[j] = data[set[j]]; sz[j] = len[set[j]]; img[j] = image_of(set[j]); }
    nvjpegStatus_t rc = NVJPEG_STATUS_SUCCESS;
    if (e.batch != cnt) {
      rc = a->nvjpegDecodeBatchedInitialize(e.handle, e.state, cnt, 1, NVJPEG_OUTPUT_RGBI);
      e.batch = rc == NVJPEG_STATUS_SUCCESS ? cnt : -1;
    }
    if (rc == NVJPEG_STATUS_SUCCESS) rc = a->nvjpegDecodeBatched(e.handle, e.state, ptr.data(), sz.data(), img.data(), st);
    if (rc != NVJPEG_STATUS_SUCCESS) {
      // one damaged stream fails the whole batch: find it image by image with the hybrid decoder
      cudaGetLastError();
      e.batch = -1;
      to_cpu.insert(to_cpu.end(), set.begin(), set.end());
    }
  };
  batched(d->hw, to_hw);
  batched(d->gpu, to_gpu);
  auto one_by_one = [&](Engine& e, size_t first, size_t step) {
    for (size_t j = first; j < to_cpu.size(); j += step) {
      const int i = to_cpu[j];
      nvjpegImage_t im = image_of(i);
      if (a->nvjpegDecode(e.handle, e.state, data[i], len[i], NVJPEG_OUTPUT_RGBI, &im, st) != NVJPEG_STATUS_SUCCESS) {
        cudaGetLastError();
        status[i] = B2I_EDECODE;
      }
    }
  };
  const size_t helpers = to_cpu.size() >= 8 ? std::min(d->pool.size(), to_cpu.size() / 4) : 0;
  if (helpers == 0) {
    one_by_one(d->cpu, 0, 1);
  } else {
    // entropy decoding is host work: spread the images over the helper engines (each thread enqueues its
    // GPU part on the same stream; the images write disjoint buffers)
    std::vector<std::thread> th;
    const int dev = d->device;
    for (size_t t = 0; t < helpers; ++t)
      th.emplace_back([&, t, dev] { cudaSetDevice(dev); one_by_one(d->pool[t], t + 1, helpers + 1); });
    one_by_one(d->cpu, 0, helpers + 1);
    for (auto& x : th) x.join();
  }
  (void)err;
  (void)h;
  return true;
}

}  // namespace b2i
