// C ABI (include/b200_infer.h) over the engine: contexts, model load, batched predict with
// per-GPU pinned double-buffered staging and contiguous batch sharding (no collective), plus the
// single-kernel entry points the parity tests drive.
#include "../../include/b200_infer.h"

#include <cuda_runtime.h>

#include <algorithm>
#include <cstring>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "engine.h"
#include "half_host.h"
#include "jpeg_decode.h"
#include "predict.h"
#include "resize.h"

using namespace b2i;

namespace {

thread_local std::string g_err;

int fail(int code, const std::string& msg) {
  g_err = msg;
  return code;
}
int cuda_fail(cudaError_t e, const char* what) {
  return fail(e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver ? B2I_ENODEV : B2I_ECUDA,
              std::string(what) + ": " + cudaGetErrorString(e));
}
#define CAPI_CUDA(call)                                   \
  do {                                                    \
    cudaError_t e_ = (call);                              \
    if (e_ != cudaSuccess) return cuda_fail(e_, #call);   \
  } while (0)

inline uint16_t f2h16(float f, int prec) { return host_f32_to_16(f, prec == B2I_PREC_FP16); }
inline float h162f(uint16_t h, int prec) { return host_16_to_f32(h, prec == B2I_PREC_FP16); }
inline int dtype_of(int prec) { return prec == B2I_PREC_FP32 ? DT_F32 : (prec == B2I_PREC_FP16 ? DT_F16 : DT_BF16); }

struct DevBuf {
  void* p = nullptr;
  ~DevBuf() { if (p) cudaFree(p); }
  cudaError_t alloc(size_t bytes) { return cudaMalloc(&p, bytes ? bytes : 16); }
};

// fp32 host -> device in the kernel dtype
cudaError_t upload(DevBuf* d, const float* src, size_t n, int prec) {
  cudaError_t e = d->alloc(n * (prec == B2I_PREC_FP32 ? 4 : 2));
  if (e != cudaSuccess) return e;
  if (prec == B2I_PREC_FP32) return cudaMemcpy(d->p, src, n * 4, cudaMemcpyHostToDevice);
  std::vector<uint16_t> t(n);
  for (size_t i = 0; i < n; ++i) t[i] = f2h16(src[i], prec);
  return cudaMemcpy(d->p, t.data(), n * 2, cudaMemcpyHostToDevice);
}
cudaError_t download(const void* dsrc, float* dst, size_t n, int prec) {
  if (prec == B2I_PREC_FP32) return cudaMemcpy(dst, dsrc, n * 4, cudaMemcpyDeviceToHost);
  std::vector<uint16_t> t(n);
  cudaError_t e = cudaMemcpy(t.data(), dsrc, n * 2, cudaMemcpyDeviceToHost);
  if (e != cudaSuccess) return e;
  for (size_t i = 0; i < n; ++i) dst[i] = h162f(t[i], prec);
  return cudaSuccess;
}

// No C++ exception may cross the C boundary (std::bad_alloc from a vector, std::system_error from a
// thread that could not be started, ...): every entry point runs its body through this.
template <class F>
auto guarded(F&& body) -> decltype(body()) {
  try {
    return body();
  } catch (const std::bad_alloc&) {
    return fail(B2I_ENOMEM, "out of host memory");
  } catch (const std::exception& e) {
    return fail(B2I_EINTERNAL, std::string("internal error: ") + e.what());
  } catch (...) {
    return fail(B2I_EINTERNAL, "internal error: unknown exception");
  }
}

}  // namespace

extern "C" {

const char* b2i_last_error(void) { return g_err.c_str(); }

int b2i_device_count(void) {
  return guarded([&]() -> int {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
      cudaGetLastError();
      return 0;
    }
    return n;
  });
}

int b2i_init(const int* device_ids, int n_devices, b2i_ctx** out) {
  return guarded([&]() -> int {
    if (!out || n_devices < 1 || !device_ids) return fail(B2I_EINVAL, "b2i_init: need at least one device id");
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
      cudaGetLastError();
      return fail(B2I_ENODEV, std::string("no CUDA device: ") + (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0") +
                                  " (this library has no CPU path)");
    }
    auto ctx = std::make_unique<b2i_ctx>();
    for (int i = 0; i < n_devices; ++i) {
      const int d = device_ids[i];
      if (d < 0 || d >= count) return fail(B2I_EINVAL, "b2i_init: device ordinal " + std::to_string(d) + " out of range");
      cudaDeviceProp prop;
      CAPI_CUDA(cudaGetDeviceProperties(&prop, d));
      if (prop.major != 10)
        return fail(B2I_ENODEV, "device " + std::to_string(d) + " (" + prop.name + ") is not sm_100; kernels are sm_100a only");
      ctx->devices.push_back(d);
    }
    *out = ctx.release();
    return B2I_OK;
  });
}

void b2i_shutdown(b2i_ctx* ctx) { delete ctx; }

void b2i_default_config(b2i_model_config* cfg) {
  if (!cfg) return;
  cfg->precision = B2I_PREC_BF16;
  cfg->max_batch = 256;
  cfg->micro_batch = 0;
  cfg->alexnet_geometry = B2I_ALEXNET_TCH;
  cfg->conv_mode = -1;
  cfg->weights_tma = 1;
  cfg->tile_n_max = 0;
  cfg->keep_tensors = 0;
  cfg->fused_stem = 1;
  cfg->patch_conv = 1;
  cfg->pair_mma = 256;
  cfg->fuse_downsample = 1;
}

int b2i_ot_inspect(const char* path, char* names_out, size_t cap) {
  return guarded([&]() -> int {
    if (!path) return fail(B2I_EINVAL, "null path");
    std::map<std::string, HostTensor> t;
    std::string err;
    if (!read_ot_archive(path, &t, &err)) return fail(B2I_EIO, err);
    if (names_out && cap) {
      std::string s;
      for (auto& kv : t) {
        s += kv.first + ":";
        for (size_t i = 0; i < kv.second.shape.size(); ++i) s += (i ? "x" : "") + std::to_string(kv.second.shape[i]);
        s += "\n";
      }
      std::strncpy(names_out, s.c_str(), cap - 1);
      names_out[cap - 1] = 0;
    }
    return static_cast<int>(t.size());
  });
}

long long b2i_ot_read_tensor(const char* path, const char* name, float* out, long long cap) {
  return guarded([&]() -> long long {
    if (!path || !name) return fail(B2I_EINVAL, "null argument");
    std::map<std::string, HostTensor> t;
    std::string err;
    if (!read_ot_archive(path, &t, &err)) return fail(B2I_EIO, err);
    std::string key(name);
    for (char& c : key) if (c == '|') c = '.';
    auto it = t.find(key);
    if (it == t.end()) return fail(B2I_EINVAL, "no tensor named " + key);
    const long long n = it->second.numel();
    if (out) {
      if (cap < n) return fail(B2I_EINVAL, "output buffer too small");
      std::memcpy(out, it->second.data.data(), static_cast<size_t>(n) * 4);
    }
    return n;
  });
}

int b2i_model_load(b2i_ctx* ctx, const char* model_name, const char* ot_path,
                   const b2i_model_config* cfg_in, b2i_model** out) {
  return guarded([&]() -> int {
    if (!ctx || !model_name || !ot_path || !out) return fail(B2I_EINVAL, "b2i_model_load: null argument");
    b2i_model_config cfg;
    if (cfg_in) cfg = *cfg_in; else b2i_default_config(&cfg);
    const std::string name(model_name);
    if (name != "resnet18" && name != "alexnet")
      return fail(B2I_EINVAL, "unknown model '" + name + "' (expected resnet18 or alexnet)");
    if (cfg.precision != B2I_PREC_BF16 && cfg.precision != B2I_PREC_FP32 && cfg.precision != B2I_PREC_FP16) return fail(B2I_EINVAL, "bad precision");
    if (cfg.max_batch < 1) return fail(B2I_EINVAL, "max_batch must be >= 1");
    // the CTA-pair kernel splits a weight tile between two CTAs and needs tiles of at least 64 channels
    if (cfg.pair_mma != 0 && (cfg.pair_mma < 64 || cfg.pair_mma > 256))
      return fail(B2I_EINVAL, "pair_mma must be 0 (off) or a tile width in 64..256");
    std::map<std::string, HostTensor> tensors;
    std::string err;
    if (!read_ot_archive(ot_path, &tensors, &err)) return fail(B2I_EIO, err);
    NetSpec net;
    const bool fused = cfg.fused_stem != 0 && cfg.precision != B2I_PREC_FP32;
    const bool ok = name == "resnet18" ? build_resnet18(tensors, fused, &net, &err)
                                       : build_alexnet(tensors, cfg.alexnet_geometry, fused, &net, &err);
    if (!ok) return fail(B2I_EIO, err);
    tensors.clear();

    EngineConfig ec;
    ec.precision = cfg.precision;
    ec.max_batch = cfg.max_batch;
    ec.micro_batch = cfg.micro_batch;
    ec.alexnet_geometry = cfg.alexnet_geometry;
    ec.conv_amode = cfg.conv_mode == 2 ? 1 : cfg.conv_mode;
    ec.persistent = cfg.conv_mode == 2 ? 0 : 1;
    ec.patch_conv = cfg.patch_conv;
    ec.pair_mma = cfg.pair_mma;
    ec.fuse_downsample = cfg.fuse_downsample;
    ec.b_tma = cfg.weights_tma;
    if (cfg.tile_n_max > 0) ec.bn_max = cfg.tile_n_max;
    ec.keep_tensors = cfg.keep_tensors;
    ec.fused_stem = fused ? 1 : 0;

    auto m = std::make_unique<b2i_model>();
    m->engine_cfg = ec;
    m->ctx = ctx;
    m->name = name;
    m->flops = net.flops_per_image;
    m->max_batch = cfg.max_batch;
    m->slots.resize(ctx->devices.size());
    for (size_t i = 0; i < ctx->devices.size(); ++i) {
      DevSlot& s = m->slots[i];
      s.dm = std::make_unique<DeviceModel>();
      if (!s.dm->init(ctx->devices[i], net, ec, &err)) {
        for (auto& t : m->slots) slot_free(&t);
        return fail(B2I_ECUDA, err);
      }
      if (!slot_setup(&s, cfg.max_batch, &err)) {
        for (auto& t : m->slots) slot_free(&t);
        return fail(B2I_ECUDA, err);
      }
    }
    *out = m.release();
    return B2I_OK;
  });
}

int b2i_model_reload(b2i_model* m, const char* ot_path) {
  return guarded([&]() -> int {
    if (!m || !ot_path) return fail(B2I_EINVAL, "b2i_model_reload: null argument");
    std::map<std::string, HostTensor> tensors;
    std::string err;
    if (!read_ot_archive(ot_path, &tensors, &err)) return fail(B2I_EIO, err);
    NetSpec net;
    const bool ok = m->name == "resnet18" ? build_resnet18(tensors, m->engine_cfg.fused_stem != 0, &net, &err)
                                          : build_alexnet(tensors, m->engine_cfg.alexnet_geometry,
                                                          m->engine_cfg.fused_stem != 0 && m->engine_cfg.precision != PREC_FP32, &net, &err);
    if (!ok) return fail(B2I_EIO, err);
    tensors.clear();
    // build the new device copies first (the old ones keep serving), then swap them in under the lock
    std::vector<std::unique_ptr<DeviceModel>> fresh(m->slots.size());
    for (size_t i = 0; i < m->slots.size(); ++i) {
      fresh[i] = std::make_unique<DeviceModel>();
      if (!fresh[i]->init(m->slots[i].dm->device(), net, m->engine_cfg, &err)) return fail(B2I_ECUDA, err);
    }
    std::lock_guard<std::mutex> lock(m->mu);
    if (m->pending[0].active || m->pending[1].active) return fail(B2I_EINVAL, "requests are in flight; b2i_wait for them before reloading");
    for (size_t i = 0; i < m->slots.size(); ++i) {
      DevSlot& s = m->slots[i];
      cudaSetDevice(s.dm->device());
      cudaDeviceSynchronize();
      for (auto& kv : s.graphs) if (kv.second.exec) cudaGraphExecDestroy(kv.second.exec);   // captured pointers die with the old buffers
      s.graphs.clear();
      s.dm = std::move(fresh[i]);
    }
    m->flops = net.flops_per_image;
    return B2I_OK;
  });
}

void b2i_model_free(b2i_model* m) {
  if (!m) return;
  for (auto& s : m->slots) slot_free(&s);
  delete m;
}

double b2i_model_flops_per_image(const b2i_model* m) { return m ? m->flops : 0; }
int b2i_model_launches_per_microbatch(const b2i_model* m) {
  return guarded([&]() -> int {
    return m && !m->slots.empty() ? static_cast<int>(m->slots[0].dm->launches_per_microbatch()) : 0;
  });
}
int b2i_model_num_devices(const b2i_model* m) { return m ? static_cast<int>(m->slots.size()) : 0; }
long long b2i_model_weight_bytes(const b2i_model* m) {
  return guarded([&]() -> int {
    return m && !m->slots.empty() ? static_cast<long long>(m->slots[0].dm->weight_bytes()) : 0;
  });
}
long long b2i_model_activation_bytes(const b2i_model* m) {
  return guarded([&]() -> long long {
    return m && !m->slots.empty() ? static_cast<long long>(m->slots[0].dm->activation_bytes()) : 0;
  });
}

int b2i_predict_u8_on(b2i_model* m, int first_dev, int n_dev, const uint8_t* images, int n, int k,
                      int32_t* idx, float* prob, float* logits) {
  return guarded([&]() -> int {
    if (!m) return fail(B2I_EINVAL, "null model");
    if (n == 0) return B2I_OK;
    std::lock_guard<std::mutex> lock(m->mu);   // calls on one model are serialised (reference: Mutex<Box<dyn ModuleT>>)
    std::string err;
    // a blocking call drains whatever an asynchronous user left in flight on its lane first
    const int lane = static_cast<int>(m->next_ticket & 1);
    if (m->pending[lane].active) return fail(B2I_EINVAL, "an asynchronous request is still in flight; b2i_wait for it first");
    long long t = -1;
    int rc = predict_submit(m, first_dev, n_dev, images, n, k, idx, prob, logits, &t, &err);
    if (rc != B2I_OK) return fail(rc, err);
    rc = predict_wait(m, t, &err);
    if (rc != B2I_OK) return fail(rc, err);
    return B2I_OK;
  });
}

int b2i_predict_u8(b2i_model* m, const uint8_t* images, int n, int k, int32_t* idx, float* prob,
                   float* logits) {
  return guarded([&]() -> int {
    if (!m) return fail(B2I_EINVAL, "null model");
    return b2i_predict_u8_on(m, 0, static_cast<int>(m->slots.size()), images, n, k, idx, prob, logits);
  });
}

int b2i_predict_raw_u8(b2i_model* m, int first_dev, int n_dev, const uint8_t* const* images, const int* heights,
                       const int* widths, int n, int k, int32_t* idx, float* prob, float* logits) {
  return guarded([&]() -> int {
    if (!m) return fail(B2I_EINVAL, "null model");
    if (n == 0) return B2I_OK;
    if (n_dev <= 0) { first_dev = 0; n_dev = static_cast<int>(m->slots.size()); }
    std::string err;
    const int rc = predict_raw(m, first_dev, n_dev, images, heights, widths, n, k, idx, prob, logits, &err);
    if (rc != B2I_OK) return fail(rc, err);
    return B2I_OK;
  });
}

int b2i_predict_raw_u8_status(b2i_model* m, int first_dev, int n_dev, const uint8_t* const* images, const int* heights,
                              const int* widths, int n, int k, int32_t* idx, float* prob, float* logits, int* status) {
  return guarded([&]() -> int {
    if (!m) return fail(B2I_EINVAL, "null model");
    if (n == 0) return B2I_OK;
    if (n_dev <= 0) { first_dev = 0; n_dev = static_cast<int>(m->slots.size()); }
    std::string err;
    const int rc = predict_raw_status(m, first_dev, n_dev, images, heights, widths, n, k, idx, prob, logits, status, &err);
    if (rc != B2I_OK) return fail(rc, err);
    return B2I_OK;
  });
}

int b2i_predict_jpeg(b2i_model* m, int first_dev, int n_dev, const uint8_t* const* jpeg, const size_t* len, int n, int k,
                     int32_t* idx, float* prob, float* logits, int* status) {
  return guarded([&]() -> int {
    if (!m) return fail(B2I_EINVAL, "null model");
    if (n == 0) return B2I_OK;
    if (n_dev <= 0) { first_dev = 0; n_dev = static_cast<int>(m->slots.size()); }
    std::string err;
    const int rc = predict_jpeg(m, first_dev, n_dev, jpeg, len, n, k, idx, prob, logits, status, &err);
    if (rc != B2I_OK) return fail(rc, err);
    return B2I_OK;
  });
}

int b2i_jpeg_decode_u8(b2i_ctx* ctx, int dev, const uint8_t* jpeg, size_t len, uint8_t* out, size_t cap, int* h, int* w) {
  return guarded([&]() -> int {
    if (!ctx || dev < 0 || dev >= static_cast<int>(ctx->devices.size()) || !jpeg || !h || !w)
      return fail(B2I_EINVAL, "b2i_jpeg_decode_u8: bad argument");
    CAPI_CUDA(cudaSetDevice(ctx->devices[dev]));
    std::lock_guard<std::mutex> lock(ctx->jpeg_mu);
    if (ctx->jpeg.size() < ctx->devices.size()) ctx->jpeg.resize(ctx->devices.size(), nullptr);
    std::string err;
    if (!ctx->jpeg[dev]) {
      ctx->jpeg[dev] = jpeg_decoder_create(&err);
      if (!ctx->jpeg[dev]) return fail(B2I_ENODEV, err);
    }
    int st = B2I_EDECODE;
    jpeg_probe(ctx->jpeg[dev], &jpeg, &len, 1, h, w, &st);
    if (st != B2I_OK) return fail(st, "not a JPEG stream this decoder accepts");
    if (!out) return B2I_OK;
    const size_t bytes = static_cast<size_t>(*h) * *w * 3;
    if (cap < bytes) return fail(B2I_EINVAL, "output buffer too small for " + std::to_string(*h) + "x" + std::to_string(*w));
    DevBuf d;
    CAPI_CUDA(d.alloc(bytes));
    uint8_t* dst = static_cast<uint8_t*>(d.p);
    if (!jpeg_decode_batch(ctx->jpeg[dev], &jpeg, &len, 1, &dst, h, w, &st, nullptr, &err)) return fail(B2I_ECUDA, err);
    CAPI_CUDA(cudaDeviceSynchronize());
    if (st != B2I_OK) return fail(st, "the JPEG bitstream is damaged");
    CAPI_CUDA(cudaMemcpy(out, d.p, bytes, cudaMemcpyDeviceToHost));
    return B2I_OK;
  });
}

int b2i_jpeg_backend(b2i_ctx* ctx, int dev, char* out, size_t cap) {
  return guarded([&]() -> int {
    if (!ctx || dev < 0 || dev >= static_cast<int>(ctx->devices.size()) || !out || cap < 2)
      return fail(B2I_EINVAL, "b2i_jpeg_backend: bad argument");
    CAPI_CUDA(cudaSetDevice(ctx->devices[dev]));
    std::lock_guard<std::mutex> lock(ctx->jpeg_mu);
    if (ctx->jpeg.size() < ctx->devices.size()) ctx->jpeg.resize(ctx->devices.size(), nullptr);
    std::string err;
    if (!ctx->jpeg[dev]) {
      ctx->jpeg[dev] = jpeg_decoder_create(&err);
      if (!ctx->jpeg[dev]) return fail(B2I_ENODEV, err);
    }
    std::strncpy(out, jpeg_decoder_backend(ctx->jpeg[dev]), cap - 1);
    out[cap - 1] = 0;
    return B2I_OK;
  });
}

int b2i_submit_u8(b2i_model* m, const uint8_t* images, int n, int k, int32_t* idx, float* prob,
                  float* logits, long long* ticket) {
  return guarded([&]() -> int {
    if (!m || !ticket) return fail(B2I_EINVAL, "null model/ticket");
    std::lock_guard<std::mutex> lock(m->mu);
    std::string err;
    const int rc = predict_submit(m, 0, static_cast<int>(m->slots.size()), images, n, k, idx, prob, logits, ticket, &err);
    if (rc != B2I_OK) return fail(rc, err);
    return B2I_OK;
  });
}

int b2i_wait(b2i_model* m, long long ticket) {
  return guarded([&]() -> int {
    if (!m) return fail(B2I_EINVAL, "null model");
    std::lock_guard<std::mutex> lock(m->mu);
    std::string err;
    const int rc = predict_wait(m, ticket, &err);
    if (rc != B2I_OK) return fail(rc, err);
    return B2I_OK;
  });
}

int b2i_forward_device(b2i_model* m, int dev, const void* d_images, int n, int k, void* d_idx,
                       void* d_prob, void* d_logits, void* stream) {
  return guarded([&]() -> int {
    if (!m || dev < 0 || dev >= static_cast<int>(m->slots.size())) return fail(B2I_EINVAL, "bad model/device");
    if (!d_images || !d_idx || !d_prob) return fail(B2I_EINVAL, "null device buffer");
    // the lock orders this against b2i_model_reload (which destroys the device copy) and against the host
    // side of a predict on the same GPU; the kernels themselves run on the caller's stream
    std::lock_guard<std::mutex> lock(m->mu);
    std::string err;
    if (!m->slots[dev].dm->forward(static_cast<const uint8_t*>(d_images), n, k, static_cast<int32_t*>(d_idx),
                                   static_cast<float*>(d_prob), static_cast<float*>(d_logits),
                                   static_cast<cudaStream_t>(stream), &err))
      return fail(B2I_ECUDA, err);
    return B2I_OK;
  });
}

int b2i_model_profile(b2i_model* m, int dev, const void* d_images, int n, int k, void* d_idx,
                      void* d_prob, void* stream, float* ms, int cap, char* labels_out,
                      size_t labels_cap) {
  return guarded([&]() -> int {
    if (!m || dev < 0 || dev >= static_cast<int>(m->slots.size()) || !d_images || !d_idx || !d_prob || !ms)
      return fail(B2I_EINVAL, "b2i_model_profile: bad argument");
    std::lock_guard<std::mutex> lock(m->mu);
    std::vector<float> v;
    std::vector<std::string> names;
    std::string err;
    if (!m->slots[dev].dm->profile(static_cast<const uint8_t*>(d_images), n, k, static_cast<int32_t*>(d_idx),
                                   static_cast<float*>(d_prob), static_cast<cudaStream_t>(stream), &v, &names, &err))
      return fail(B2I_ECUDA, err);
    if (static_cast<int>(v.size()) > cap) return fail(B2I_EINVAL, "ms buffer too small");
    std::memcpy(ms, v.data(), v.size() * 4);
    if (labels_out && labels_cap) {
      std::string s;
      for (auto& nm : names) s += nm + "\n";
      std::strncpy(labels_out, s.c_str(), labels_cap - 1);
      labels_out[labels_cap - 1] = 0;
    }
    return static_cast<int>(v.size());
  });
}

void* b2i_host_alloc(size_t bytes) {
  void* p = nullptr;
  if (cudaMallocHost(&p, bytes) != cudaSuccess) {
    cudaGetLastError();
    try { g_err = "cudaMallocHost failed"; } catch (...) {}
    return nullptr;
  }
  return p;
}
void b2i_host_free(void* p) { if (p) cudaFreeHost(p); }

long long b2i_model_read_tensor(b2i_model* m, int dev, int tensor_id, int n, float* out, long long cap) {
  return guarded([&]() -> long long {
    if (!m || dev < 0 || dev >= static_cast<int>(m->slots.size()) || !out) return fail(B2I_EINVAL, "bad argument");
    std::lock_guard<std::mutex> lock(m->mu);
    std::vector<float> v;
    std::string err;
    if (!m->slots[dev].dm->read_tensor(tensor_id, n, &v, &err)) return fail(B2I_ECUDA, err);
    if (static_cast<long long>(v.size()) > cap) return fail(B2I_EINVAL, "output buffer too small");
    std::memcpy(out, v.data(), v.size() * 4);
    return static_cast<long long>(v.size());
  });
}
int b2i_model_num_tensors(const b2i_model* m) { return m && !m->slots.empty() ? m->slots[0].dm->num_tensors() : 0; }
int b2i_model_tensor_shape(const b2i_model* m, int id, int* h, int* w, int* c) {
  return guarded([&]() -> int {
    if (!m || m->slots.empty() || id < 0 || id >= m->slots[0].dm->num_tensors()) return fail(B2I_EINVAL, "bad tensor id");
    const TensorShape s = m->slots[0].dm->tensor_shape(id);
    if (h) *h = s.H;
    if (w) *w = s.W;
    if (c) *c = s.C;
    return B2I_OK;
  });
}

int b2i_resize_crop_u8(b2i_ctx* ctx, int dev, const uint8_t* image, int h, int w, uint8_t* out224) {
  return guarded([&]() -> int {
    if (!ctx || dev < 0 || dev >= static_cast<int>(ctx->devices.size()) || !image || !out224 || h < 1 || w < 1)
      return fail(B2I_EINVAL, "b2i_resize_crop_u8: bad argument");
    CAPI_CUDA(cudaSetDevice(ctx->devices[dev]));
    std::string err;
    if (!resize_crop_224_host(image, h, w, out224, nullptr, &err)) return fail(B2I_ECUDA, err);
    return B2I_OK;
  });
}

int b2i_resize_crop_u8_static(b2i_ctx* ctx, int dev, const void* d_image, int h, int w, void* d_out224,
                              void* cuda_stream) {
  return guarded([&]() -> int {
    if (!ctx || dev < 0 || dev >= static_cast<int>(ctx->devices.size()) || !d_image || !d_out224)
      return fail(B2I_EINVAL, "b2i_resize_crop_u8_static: bad argument");
    CAPI_CUDA(cudaSetDevice(ctx->devices[dev]));
    std::string err;
    if (!resize_crop_224_static(static_cast<const uint8_t*>(d_image), h, w, static_cast<uint8_t*>(d_out224),
                                static_cast<cudaStream_t>(cuda_stream), &err))
      return fail(err.find("exceeds the static scratch") != std::string::npos || err == "empty image" ||
                          err.find("capturing") != std::string::npos ? B2I_EINVAL : B2I_ECUDA, err);
    return B2I_OK;
  });
}

int b2i_resize_plan(int in_size, int out_size, int crop, int* first, int* count, int* idx,
                    float* coef, int cap) {
  return guarded([&]() -> int {
    if (in_size < 1 || out_size < 224 || crop < 0 || crop + 224 > out_size || !first || !count || !idx || !coef)
      return fail(B2I_EINVAL, "b2i_resize_plan: bad argument");
    AxisPlan p;
    plan_axis(in_size, out_size, crop, &p);
    if (static_cast<int>(p.idx.size()) > cap) return fail(B2I_EINVAL, "b2i_resize_plan: tap buffer too small");
    std::memcpy(first, p.first.data(), 224 * 4);
    std::memcpy(count, p.count.data(), 224 * 4);
    std::memcpy(idx, p.idx.data(), p.idx.size() * 4);
    std::memcpy(coef, p.coef.data(), p.coef.size() * 4);
    return static_cast<int>(p.idx.size());
  });
}

void b2i_resize_target(int h, int w, int* rw, int* rh, int* crop_x, int* crop_y) {
  const ResizeGeom g = resize_target(h, w);
  if (rw) *rw = g.rw;
  if (rh) *rh = g.rh;
  if (crop_x) *crop_x = g.crop_x;
  if (crop_y) *crop_y = g.crop_y;
}

// ------------------------------------------------------------------ single-kernel entry points
static int ctx_dev(b2i_ctx* ctx, int dev) {
  return guarded([&]() -> int {
    if (!ctx || dev < 0 || dev >= static_cast<int>(ctx->devices.size())) return fail(B2I_EINVAL, "bad context/device");
    CAPI_CUDA(cudaSetDevice(ctx->devices[dev]));
    return B2I_OK;
  });
}

int b2i_op_conv2d(b2i_ctx* ctx, int dev, int precision, int conv_mode, int weights_tma, int tile_n,
                  const float* x, int N, int H, int W, int C, const float* w, const float* bias,
                  const float* bn, int Cout, int R, int S, int stride, int pad,
                  const float* residual, int relu, float* out) {
  return guarded([&]() -> int {
    if (int rc = ctx_dev(ctx, dev)) return rc;
    if (!x || !w || !out || N < 1 || C < 1 || Cout < 1) return fail(B2I_EINVAL, "b2i_op_conv2d: bad argument");
    ConvSpec cs;
    build_conv_spec(&cs, "op", w, bias, bn, bn ? bn + Cout : nullptr, bn ? bn + 2 * Cout : nullptr,
                    bn ? bn + 3 * Cout : nullptr, Cout, C, R, S, stride, pad);
    const int OH = (H + 2 * pad - R) / stride + 1, OW = (W + 2 * pad - S) / stride + 1;
    if (OH < 1 || OW < 1) return fail(B2I_EINVAL, "empty output");
    // channel-pad the input
    const size_t npix = static_cast<size_t>(N) * H * W;
    std::vector<float> xp(npix * cs.Cin_pad, 0.f);
    for (size_t p = 0; p < npix; ++p) std::memcpy(&xp[p * cs.Cin_pad], &x[p * C], static_cast<size_t>(C) * 4);
    DevBuf dx, dw, db, dr, dout;
    CAPI_CUDA(upload(&dx, xp.data(), xp.size(), precision));
    CAPI_CUDA(upload(&dw, cs.w.data(), cs.w.size(), precision));
    CAPI_CUDA(upload(&db, cs.bias.data(), cs.bias.size(), B2I_PREC_FP32));
    const size_t nout = static_cast<size_t>(N) * OH * OW * Cout;
    const bool bf16 = precision != B2I_PREC_FP32;   // either 16-bit storage format: the tcgen05 path
    const bool out32 = bf16 && (Cout % 32 != 0);
    if (out32 && residual) return fail(B2I_EINVAL, "bf16 conv with a residual needs Cout % 32 == 0");
    if (residual) CAPI_CUDA(upload(&dr, residual, nout, precision));
    CAPI_CUDA(dout.alloc(nout * 4));
    CAPI_CUDA(cudaMemset(dout.p, 0xFF, nout * (bf16 && !out32 ? 2 : 4)));  // NaN canary
    ConvArgs a;
    ConvGeom& g = a.g;
    g.N = N; g.H = H; g.W = W; g.C = cs.Cin_pad; g.OH = OH; g.OW = OW; g.R = R; g.S = S; g.Sp = cs.Sp;
    g.stride = stride; g.pad = pad; g.Cout = Cout; g.Cout_pad = cs.Cout_pad; g.K = cs.K; g.M = N * OH * OW;
    a.x = dx.p; a.w = dw.p; a.bias = static_cast<const float*>(db.p); a.residual = residual ? dr.p : nullptr;
    a.out = dout.p; a.ldo = Cout; a.relu = relu; a.out_fp32 = out32 ? 1 : 0;
    a.f16 = precision == B2I_PREC_FP16 ? 1 : 0;
    int bnv = tile_n > 0 ? tile_n : 128;
    bnv = std::min(bnv, 256) / 32 * 32;
    while (bnv > 32 && cs.Cout_pad % bnv) bnv -= 32;
    a.BN = bnv;
    a.stages = conv_umma_pick_stages(bnv);
    a.amode = cs.smallc ? A_SMALLC : (conv_mode >= 1 ? A_IM2COL : A_GATHER);
    a.persistent = conv_mode == 2 ? 0 : 1;
    a.b_tma = (a.amode == A_IM2COL) ? 1 : weights_tma;
    if (conv_mode == 5 || conv_mode == 7 || conv_mode == 8) a.pair = 1;
    if (conv_mode == 8) a.msub = 2;   // CTA pair, two 128-row sub-tiles per CTA (256-pixel im2col requests)
    if (conv_mode == 7) a.resident_b = 1;   // the launcher falls back to streamed weights when they do not fit
    if (conv_mode == 6) {
      a.BN = cs.Cout_pad <= 256 ? cs.Cout_pad : a.BN;
      if (!bf16 || !conv_flat_applicable(g, a.BN)) return fail(B2I_EINVAL, "conv_mode 6 (flattened rows) needs bf16, 3x3 s1 p1, C % 64 == 0, one N tile of 64..256 channels, W <= 62");
      a.flat = 1;
      a.pair = 1;
    }
    if (conv_mode == 3) {
      if (!bf16 || !conv_patch_applicable(g)) return fail(B2I_EINVAL, "conv_mode 3 (halo patch) needs bf16, 3x3 s1 p1, 64 -> 64 channels");
      a.patch = 1;
      a.BN = 64;
    }
    if (!bf16) {
      CAPI_CUDA(launch_conv_fp32(a, nullptr));
    } else {
      CUtensorMap tmA, tmB, tmR, tmO;
      const char* terr = "";
      if (a.b_tma && make_tmap_weights(&tmB, dw.p, cs.Cout_pad, cs.K, (a.pair && a.BN >= 64) ? a.BN / 2 : a.BN, &terr)) return fail(B2I_ECUDA, terr);
      if (a.patch) {
        if (make_tmap_patch(&tmA, dx.p, N, H, W, &terr)) return fail(B2I_ECUDA, terr);
        if (make_tmap_tile(&tmO, dout.p, N, H, W, &terr)) return fail(B2I_ECUDA, terr);
        if (residual && make_tmap_tile(&tmR, dr.p, N, H, W, &terr)) return fail(B2I_ECUDA, terr);
        a.tmO = &tmO;
        a.tmR = residual ? &tmR : nullptr;
      } else if (a.flat) {
        if (make_tmap_flat(&tmA, dx.p, N, H, W, g.C, &terr)) return fail(B2I_ECUDA, terr);
      } else if (a.amode == A_IM2COL && make_tmap_im2col_px(&tmA, dx.p, g, N, 128 * a.msub, &terr)) {
        return fail(B2I_ECUDA, terr);
      }
      CAPI_CUDA(launch_conv_umma(a, (a.patch || a.flat || a.amode == A_IM2COL) ? &tmA : nullptr, a.b_tma ? &tmB : nullptr, nullptr));
    }
    CAPI_CUDA(cudaDeviceSynchronize());
    CAPI_CUDA(download(dout.p, out, nout, out32 ? B2I_PREC_FP32 : precision));
    return B2I_OK;
  });
}

static int op_linear(b2i_ctx* ctx, int dev, int precision, int tile_n, int splits, const float* x, int M, int K,
                     const float* w, const float* bias, int Cout, int relu, float* out) {
  return guarded([&]() -> int {
    if (int rc = ctx_dev(ctx, dev)) return rc;
    if (!x || !w || !out || M < 1 || K < 1 || Cout < 1) return fail(B2I_EINVAL, "b2i_op_linear: bad argument");
    ConvSpec cs;
    build_conv_spec(&cs, "op", w, bias, nullptr, nullptr, nullptr, nullptr, Cout, K, 1, 1, 1, 0);
    // K may have been padded to a multiple of 64 (and to 4 when K < 8): pad x to cs.K columns
    std::vector<float> xp(static_cast<size_t>(M) * cs.K, 0.f);
    for (int m = 0; m < M; ++m) std::memcpy(&xp[static_cast<size_t>(m) * cs.K], &x[static_cast<size_t>(m) * K], static_cast<size_t>(K) * 4);
    if (cs.smallc) return fail(B2I_EINVAL, "b2i_op_linear: K must be >= 8");
    DevBuf dx, dw, db, dout;
    CAPI_CUDA(upload(&dx, xp.data(), xp.size(), precision));
    CAPI_CUDA(upload(&dw, cs.w.data(), cs.w.size(), precision));
    CAPI_CUDA(upload(&db, cs.bias.data(), cs.bias.size(), B2I_PREC_FP32));
    const size_t nout = static_cast<size_t>(M) * Cout;
    CAPI_CUDA(dout.alloc(nout * 4));
    ConvArgs a;
    ConvGeom& g = a.g;
    g.N = M; g.H = g.W = 1; g.C = cs.K; g.OH = g.OW = 1; g.R = g.S = g.Sp = 1; g.stride = 1; g.pad = 0;
    g.Cout = Cout; g.Cout_pad = cs.Cout_pad; g.K = cs.K; g.M = M;
    a.x = dx.p; a.w = dw.p; a.bias = static_cast<const float*>(db.p); a.residual = nullptr; a.out = dout.p;
    a.ldo = Cout; a.relu = relu; a.out_fp32 = 1;
    a.f16 = precision == B2I_PREC_FP16 ? 1 : 0;
    int bnv = tile_n > 0 ? tile_n : 128;
    bnv = std::min(bnv, 256) / 32 * 32;
    while (bnv > 32 && cs.Cout_pad % bnv) bnv -= 32;
    a.BN = bnv; a.stages = conv_umma_pick_stages(bnv); a.amode = A_TILED; a.b_tma = 1;
    DevBuf dws, dctr;
    if (splits > 1 && precision != B2I_PREC_FP32) {
      const size_t rows = static_cast<size_t>((M + 127) / 128) * 128;
      CAPI_CUDA(dws.alloc(static_cast<size_t>(splits) * rows * cs.Cout_pad * 4));
      CAPI_CUDA(dctr.alloc(4096 * 4));
      CAPI_CUDA(cudaMemset(dctr.p, 0, 4096 * 4));
      a.splits = splits; a.ws = static_cast<float*>(dws.p); a.ctr = static_cast<int*>(dctr.p); a.ldw = cs.Cout_pad;
      if (M <= 120) a.a_rows = std::max(8, (M + 7) / 8 * 8);   // the A box no taller than the batch, as the engine builds it
    }
    if (precision == B2I_PREC_FP32) {
      CAPI_CUDA(launch_conv_fp32(a, nullptr));
    } else {
      CUtensorMap tmA, tmB;
      const char* terr = "";
      if (make_tmap_weights(&tmB, dw.p, cs.Cout_pad, cs.K, a.BN, &terr)) return fail(B2I_ECUDA, terr);
      if (make_tmap_rows(&tmA, dx.p, M, cs.K, a.a_rows, &terr)) return fail(B2I_ECUDA, terr);
      CAPI_CUDA(launch_conv_umma(a, &tmA, &tmB, nullptr));
    }
    CAPI_CUDA(cudaDeviceSynchronize());
    CAPI_CUDA(cudaMemcpy(out, dout.p, nout * 4, cudaMemcpyDeviceToHost));
    return B2I_OK;
  });
}

int b2i_op_linear(b2i_ctx* ctx, int dev, int precision, int tile_n, const float* x, int M, int K,
                  const float* w, const float* bias, int Cout, int relu, float* out) {
  return guarded([&]() -> int { return op_linear(ctx, dev, precision, tile_n, 1, x, M, K, w, bias, Cout, relu, out); });
}

int b2i_op_linear_splitk(b2i_ctx* ctx, int dev, int precision, int tile_n, int splits, const float* x, int M, int K,
                         const float* w, const float* bias, int Cout, int relu, float* out) {
  return guarded([&]() -> int {
    if (splits < 1 || splits > 64) return fail(B2I_EINVAL, "b2i_op_linear_splitk: splits must be in 1..64");
    return op_linear(ctx, dev, precision, tile_n, splits, x, M, K, w, bias, Cout, relu, out);
  });
}

int b2i_op_maxpool(b2i_ctx* ctx, int dev, int precision, const float* x, int N, int H, int W, int C,
                   int k, int stride, int pad, float* out) {
  return guarded([&]() -> int {
    if (int rc = ctx_dev(ctx, dev)) return rc;
    const int OH = (H + 2 * pad - k) / stride + 1, OW = (W + 2 * pad - k) / stride + 1;
    DevBuf dx, dout;
    const size_t nin = static_cast<size_t>(N) * H * W * C, nout = static_cast<size_t>(N) * OH * OW * C;
    CAPI_CUDA(upload(&dx, x, nin, precision));
    CAPI_CUDA(dout.alloc(nout * 4));
    CAPI_CUDA(launch_maxpool(dx.p, dout.p, dtype_of(precision), N, H, W, C, k, stride, pad, OH, OW, nullptr));
    CAPI_CUDA(cudaDeviceSynchronize());
    CAPI_CUDA(download(dout.p, out, nout, precision));
    return B2I_OK;
  });
}

int b2i_op_adaptive_avgpool(b2i_ctx* ctx, int dev, int precision, const float* x, int N, int H,
                            int W, int C, int OH, int OW, float* out) {
  return guarded([&]() -> int {
    if (int rc = ctx_dev(ctx, dev)) return rc;
    DevBuf dx, dout;
    const size_t nin = static_cast<size_t>(N) * H * W * C, nout = static_cast<size_t>(N) * OH * OW * C;
    CAPI_CUDA(upload(&dx, x, nin, precision));
    CAPI_CUDA(dout.alloc(nout * 4));
    CAPI_CUDA(launch_adaptive_avgpool(dx.p, dout.p, dtype_of(precision), N, H, W, C, OH, OW, nullptr));
    CAPI_CUDA(cudaDeviceSynchronize());
    CAPI_CUDA(download(dout.p, out, nout, precision));
    return B2I_OK;
  });
}

int b2i_op_normalize(b2i_ctx* ctx, int dev, int precision, const uint8_t* images, int n, float* out) {
  return guarded([&]() -> int {
    if (int rc = ctx_dev(ctx, dev)) return rc;
    if (!images || !out || n < 1) return fail(B2I_EINVAL, "b2i_op_normalize: bad argument");
    DevBuf din, dout;
    const size_t nin = static_cast<size_t>(n) * B2I_IMAGE_BYTES, nout = static_cast<size_t>(n) * 224 * 224 * 4;
    CAPI_CUDA(din.alloc(nin));
    CAPI_CUDA(cudaMemcpy(din.p, images, nin, cudaMemcpyHostToDevice));
    CAPI_CUDA(dout.alloc(nout * 4));
    CAPI_CUDA(launch_normalize_u8(static_cast<const uint8_t*>(din.p), dout.p, dtype_of(precision),
                                  static_cast<long long>(n) * 224 * 224, nullptr));
    CAPI_CUDA(cudaDeviceSynchronize());
    CAPI_CUDA(download(dout.p, out, nout, precision));
    return B2I_OK;
  });
}

int b2i_op_softmax_topk(b2i_ctx* ctx, int dev, const float* logits, int rows, int classes, int k,
                        int32_t* idx, float* prob) {
  return guarded([&]() -> int {
    if (int rc = ctx_dev(ctx, dev)) return rc;
    if (!logits || !idx || !prob || rows < 1) return fail(B2I_EINVAL, "b2i_op_softmax_topk: bad argument");
    DevBuf dl, di, dp;
    CAPI_CUDA(dl.alloc(static_cast<size_t>(rows) * classes * 4));
    CAPI_CUDA(cudaMemcpy(dl.p, logits, static_cast<size_t>(rows) * classes * 4, cudaMemcpyHostToDevice));
    CAPI_CUDA(di.alloc(static_cast<size_t>(rows) * k * 4));
    CAPI_CUDA(dp.alloc(static_cast<size_t>(rows) * k * 4));
    CAPI_CUDA(launch_softmax_topk(static_cast<const float*>(dl.p), rows, classes, classes, k,
                                  static_cast<int32_t*>(di.p), static_cast<float*>(dp.p), nullptr));
    CAPI_CUDA(cudaDeviceSynchronize());
    CAPI_CUDA(cudaMemcpy(idx, di.p, static_cast<size_t>(rows) * k * 4, cudaMemcpyDeviceToHost));
    CAPI_CUDA(cudaMemcpy(prob, dp.p, static_cast<size_t>(rows) * k * 4, cudaMemcpyDeviceToHost));
    return B2I_OK;
  });
}

}  // extern "C"
