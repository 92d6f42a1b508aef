// See jpeg_decode.h.
#include "jpeg_decode.h"

#include <dlfcn.h>
#include <nvjpeg.h>   // types and prototypes only: the library is reached through dlopen, never linked

#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <map>
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

// nvJPEG sizes its work buffers per batch and, left to itself, gets them from cudaMalloc / cudaFree -- both
// synchronise the device and serialise against every other allocation in the process.  These two allocators keep every block they were ever
// asked for on a free list (per process, never shrinking: a decoder's working set is a few hundred MB
// at most) and hand back the smallest block that fits.
struct BlockPool {
  std::mutex mu;
  std::multimap<size_t, void*> free_blocks;
  std::map<void*, size_t> live;
  bool pinned = false;
  int get(void** p, size_t bytes) {
    if (bytes == 0) bytes = 256;
    {
      std::lock_guard<std::mutex> lock(mu);
      auto it = free_blocks.lower_bound(bytes);
      if (it != free_blocks.end() && it->first <= 2 * bytes + (1u << 20)) {
        *p = it->second;
        live[*p] = it->first;
        free_blocks.erase(it);
        return 0;
      }
    }
    const size_t cap = (bytes + (bytes >> 2) + 65535) & ~static_cast<size_t>(65535);   // room to grow without a new block
    cudaError_t e = pinned ? cudaHostAlloc(p, cap, cudaHostAllocDefault) : cudaMalloc(p, cap);
    if (e != cudaSuccess) { cudaGetLastError(); *p = nullptr; return static_cast<int>(e); }
    std::lock_guard<std::mutex> lock(mu);
    live[*p] = cap;
    return 0;
  }
  int put(void* p) {
    if (!p) return 0;
    // cudaFree / cudaFreeHost wait for the device, and nvJPEG relies on that: it frees a staging buffer that an
    // asynchronous copy may still be reading.  A block goes back on the list only once the GPU is idle, or another
    // decoder thread could be handed it and write into it under that copy.  (Frees stop once the buffers have
    // grown to the working set, so this costs nothing in steady state.)
    cudaDeviceSynchronize();
    std::lock_guard<std::mutex> lock(mu);
    auto it = live.find(p);
    if (it == live.end()) return 0;
    free_blocks.emplace(it->second, p);
    live.erase(it);
    return 0;
  }
};
BlockPool* dev_pool(int device) {   // one free list per GPU
  static BlockPool pools[64];
  return &pools[device & 63];
}
BlockPool* pin_pool() {
  static BlockPool p;
  p.pinned = true;
  return &p;
}
int pool_dev_malloc(void** p, size_t n) {
  int d = 0;
  cudaGetDevice(&d);
  return dev_pool(d)->get(p, n);
}
int pool_dev_free(void* p) {
  if (!p) return 0;
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return 0; }
  return dev_pool(a.device)->put(p);   // back onto the list of the GPU that owns the block
}
int pool_pin_malloc(void** p, size_t n, unsigned int) { return pin_pool()->get(p, n); }
int pool_pin_free(void* p) { return pin_pool()->put(p); }

struct Engine {
  nvjpegHandle_t handle = nullptr;
  nvjpegJpegState_t state = nullptr;
  int batch = -1;   // batch size of the last nvjpegDecodeBatchedInitialize
  cudaEvent_t drained = nullptr;   // recorded after each per-image decode: the state's staging buffers are free again
};

bool engine_open(Api* a, nvjpegBackend_t backend, Engine* e) {
  static nvjpegDevAllocator_t dev_alloc = {pool_dev_malloc, pool_dev_free};
  static nvjpegPinnedAllocator_t pin_alloc = {pool_pin_malloc, pool_pin_free};
  static const bool pooled = [] { const char* v = std::getenv("B2I_JPEG_NO_POOL"); return !(v && v[0] == '1'); }();
  if (a->nvjpegCreateEx(backend, pooled ? &dev_alloc : nullptr, pooled ? &pin_alloc : nullptr, 0, &e->handle) != NVJPEG_STATUS_SUCCESS) {
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
  if (e->drained) cudaEventDestroy(e->drained);
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
  // B2I_JPEG_BACKEND = hardware | gpu | cpu picks the batched engine (A/B measurements).  Default: the NVJPG
  // hardware engines where the device exposes them, and the threaded hybrid decoder for everything else.  The
  // GPU-Huffman backend is opt-in: on B200 with ImageNet-sized files it decodes 2.7k images/s against 7-11k for
  // sixteen host threads of the hybrid decoder (profiles/r03_jpeg_decoder_arrangements.txt).
  const char* want = std::getenv("B2I_JPEG_BACKEND");
  const std::string w = want ? want : "";
  if (!engine_open(a, NVJPEG_BACKEND_HYBRID, &d->cpu) && !engine_open(a, NVJPEG_BACKEND_DEFAULT, &d->cpu)) {
    *err = "nvjpegCreateEx failed for the hybrid backend";
    delete d;
    return nullptr;
  }
  if (w.empty() || w == "hardware") engine_open(a, NVJPEG_BACKEND_HARDWARE, &d->hw);
  if (w == "gpu") engine_open(a, NVJPEG_BACKEND_GPU_HYBRID, &d->gpu);
  if (a->nvjpegJpegStreamCreate(d->cpu.handle, &d->parsed) != NVJPEG_STATUS_SUCCESS) d->parsed = nullptr;
  cudaGetDevice(&d->device);
  // helper engines for the images no batched engine takes: B2I_JPEG_THREADS (default: the host's hardware
  // threads, at most 32)
  const char* te = std::getenv("B2I_JPEG_THREADS");
  int threads = te ? std::atoi(te) : static_cast<int>(std::thread::hardware_concurrency());
  threads = threads < 1 ? 1 : (threads > 32 ? 32 : threads);
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
  static const bool trace = [] { const char* e = std::getenv("B2I_TRACE"); return e && e[0] == '1'; }();
  const auto t0 = std::chrono::steady_clock::now();
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
  const auto t1 = std::chrono::steady_clock::now();
  const size_t n_hw = to_hw.size(), n_gpu = to_gpu.size(), n_cpu0 = to_cpu.size();
  batched(d->hw, to_hw);
  batched(d->gpu, to_gpu);
  const auto t2 = std::chrono::steady_clock::now();
  auto one_by_one = [&](Engine& e, size_t first, size_t step) {
    for (size_t j = first; j < to_cpu.size(); j += step) {
      const int i = to_cpu[j];
      nvjpegImage_t im = image_of(i);
      // nvjpegDecode returns once the GPU part is enqueued; the state's pinned staging buffer is still being copied
      // from.  Wait for the previous image of THIS state to have left it before entropy-decoding the next one into
      // it (the GPU part of an image is tens of microseconds against a millisecond of host work: rarely a wait).
      if (!e.drained) cudaEventCreateWithFlags(&e.drained, cudaEventDisableTiming);
      else cudaEventSynchronize(e.drained);
      if (a->nvjpegDecode(e.handle, e.state, data[i], len[i], NVJPEG_OUTPUT_RGBI, &im, st) != NVJPEG_STATUS_SUCCESS) {
        cudaGetLastError();
        status[i] = B2I_EDECODE;
      }
      cudaEventRecord(e.drained, st);
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
  if (trace) {
    const auto t3 = std::chrono::steady_clock::now();
    auto ms = [](std::chrono::steady_clock::time_point x, std::chrono::steady_clock::time_point y) {
      return std::chrono::duration<double, std::milli>(y - x).count();
    };
    std::fprintf(stderr, "[b2i trace] jpeg decode of %d: routing %.2f ms; batched (%zu hardware, %zu gpu) %.2f ms; per image (%zu, of which %zu fell back) %.2f ms on %zu thread(s)\n",
                 n, ms(t0, t1), n_hw, n_gpu, ms(t1, t2), to_cpu.size(), to_cpu.size() - n_cpu0, ms(t2, t3), helpers + 1);
  }
  (void)err;
  (void)h;
  return true;
}

}  // namespace b2i
