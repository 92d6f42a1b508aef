// See predict.h.
#include "predict.h"

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>

#include "../../include/b200_infer.h"
#include "jpeg_decode.h"
#include "resize.h"

b2i_ctx::~b2i_ctx() {
  for (size_t i = 0; i < jpeg.size(); ++i)
    if (jpeg[i]) {
      cudaSetDevice(devices[i]);
      b2i::jpeg_decoder_destroy(jpeg[i]);
    }
}

namespace b2i {

namespace {

bool is_pinned(const void* p) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return a.type == cudaMemoryTypeHost;
}

struct Shard {
  int rc = B2I_OK;
  std::string err;
};

// H2D (chunked, own stream) + forward per chunk + D2H of the results, all asynchronous.
void enqueue_shard(DevSlot& s, int lane_id, const uint8_t* images, int n, int k, int32_t* idx,
                   float* prob, float* logits, Shard* out) {
  Lane& L = s.lane[lane_id];
  L.n = n; L.k = k; L.out_idx = idx; L.out_prob = prob; L.out_logits = logits;
  if (n == 0) return;
  auto cu = [&](cudaError_t e, const char* what) {
    if (e == cudaSuccess) return false;
    out->rc = B2I_ECUDA;
    out->err = std::string(what) + ": " + cudaGetErrorString(e);
    return true;
  };
  if (cu(cudaSetDevice(s.dm->device()), "cudaSetDevice")) return;
  const int classes = s.dm->classes();
  const size_t cap = static_cast<size_t>(s.dm->max_batch());
  const uint8_t* src = images;
  if (!is_pinned(images)) {
    // pageable source: stage through pinned memory so the copy is a real asynchronous DMA
    if (!L.h_stage && cu(cudaMallocHost(reinterpret_cast<void**>(&L.h_stage), cap * B2I_IMAGE_BYTES), "cudaMallocHost(stage)")) return;
    std::memcpy(L.h_stage, images, static_cast<size_t>(n) * B2I_IMAGE_BYTES);
    src = L.h_stage;
  }
  int c = 0;
  for (int off = 0; off < n; off += s.chunk, ++c) {
    const int nb = std::min(s.chunk, n - off);
    const size_t byte_off = static_cast<size_t>(off) * B2I_IMAGE_BYTES;
    if (cu(cudaMemcpyAsync(L.d_in + byte_off, src + byte_off, static_cast<size_t>(nb) * B2I_IMAGE_BYTES,
                           cudaMemcpyHostToDevice, s.copy_st), "H2D copy")) return;
    if (cu(cudaEventRecord(L.copied[c], s.copy_st), "cudaEventRecord")) return;
    if (cu(cudaStreamWaitEvent(s.comp_st, L.copied[c], 0), "cudaStreamWaitEvent")) return;
    auto run_forward = [&]() {
      return s.dm->forward(L.d_in + byte_off, nb, k, L.d_idx + static_cast<size_t>(off) * k,
                           L.d_prob + static_cast<size_t>(off) * k,
                           logits ? L.d_logits + static_cast<size_t>(off) * classes : nullptr, s.comp_st, &out->err);
    };
    bool done = false;
    if (nb <= s.graph_max_batch) {
      // key: every pointer the launches embed is a function of (lane, chunk index, nb, k, logits?)
      const uint64_t key = (static_cast<uint64_t>(lane_id) << 60) | (static_cast<uint64_t>(c) << 40) |
                           (static_cast<uint64_t>(nb) << 8) | (static_cast<uint64_t>(k) << 1) | (logits ? 1u : 0u);
      GraphEntry& ge = s.graphs[key];
      ++ge.uses;
      if (ge.uses == 2 && !ge.exec && s.graphs.size() <= 256) {
        // second use: capture (the first, eager use has run every one-time attribute / handle set-up)
        cudaGraph_t graph = nullptr;
        if (cudaStreamBeginCapture(s.comp_st, cudaStreamCaptureModeThreadLocal) == cudaSuccess) {
          const bool ok = run_forward();
          cudaError_t ce = cudaStreamEndCapture(s.comp_st, &graph);
          if (ok && ce == cudaSuccess && graph) {
            if (cudaGraphInstantiate(&ge.exec, graph, 0) != cudaSuccess) ge.exec = nullptr;
          }
          if (graph) cudaGraphDestroy(graph);
          if (!ok || ce != cudaSuccess) {
            cudaGetLastError();
            ge.exec = nullptr;
            ge.uses = 1 << 20;   // do not try again for this shape
          }
        }
      }
      if (ge.exec) {
        if (cu(cudaGraphLaunch(ge.exec, s.comp_st), "cudaGraphLaunch")) return;
        done = true;
      }
    }
    if (!done && !run_forward()) {
      out->rc = B2I_ECUDA;
      return;
    }
  }
  const size_t nk = static_cast<size_t>(n) * k;
  if (cu(cudaMemcpyAsync(L.h_idx, L.d_idx, nk * 4, cudaMemcpyDeviceToHost, s.comp_st), "D2H idx")) return;
  if (cu(cudaMemcpyAsync(L.h_prob, L.d_prob, nk * 4, cudaMemcpyDeviceToHost, s.comp_st), "D2H prob")) return;
  if (logits && cu(cudaMemcpyAsync(L.h_logits, L.d_logits, static_cast<size_t>(n) * classes * 4,
                                   cudaMemcpyDeviceToHost, s.comp_st), "D2H logits"))
    return;
  cu(cudaEventRecord(L.done, s.comp_st), "cudaEventRecord(done)");
}

void collect_shard(DevSlot& s, int lane_id, Shard* out) {
  Lane& L = s.lane[lane_id];
  if (L.n == 0) return;
  cudaError_t e = cudaSetDevice(s.dm->device());
  if (e == cudaSuccess) e = cudaEventSynchronize(L.done);
  if (e != cudaSuccess) {
    out->rc = B2I_ECUDA;
    out->err = std::string("waiting for the forward: ") + cudaGetErrorString(e);
    return;
  }
  const size_t nk = static_cast<size_t>(L.n) * L.k;
  std::memcpy(L.out_idx, L.h_idx, nk * 4);
  std::memcpy(L.out_prob, L.h_prob, nk * 4);
  if (L.out_logits) std::memcpy(L.out_logits, L.h_logits, static_cast<size_t>(L.n) * s.dm->classes() * 4);
}

}  // namespace

bool slot_setup(DevSlot* s, int max_batch, std::string* err) {
  cudaError_t e = cudaSetDevice(s->dm->device());
  auto chk = [&](cudaError_t r) { if (e == cudaSuccess) e = r; };
  s->chunk = s->dm->micro_batch();
  const int chunks = (max_batch + s->chunk - 1) / s->chunk;
  const size_t in_bytes = static_cast<size_t>(max_batch) * B2I_IMAGE_BYTES;
  const size_t nk = static_cast<size_t>(max_batch) * 8;
  chk(cudaStreamCreateWithFlags(&s->copy_st, cudaStreamNonBlocking));
  chk(cudaStreamCreateWithFlags(&s->comp_st, cudaStreamNonBlocking));
  for (Lane& L : s->lane) {
    chk(cudaMalloc(reinterpret_cast<void**>(&L.d_in), in_bytes));
    chk(cudaMalloc(reinterpret_cast<void**>(&L.d_idx), nk * 4));
    chk(cudaMalloc(reinterpret_cast<void**>(&L.d_prob), nk * 4));
    chk(cudaMallocHost(reinterpret_cast<void**>(&L.h_idx), nk * 4));
    chk(cudaMallocHost(reinterpret_cast<void**>(&L.h_prob), nk * 4));
    // logits staging up front as well (4 KB per image and lane): the first call that asks for logits must not
    // be the one that allocates
    chk(cudaMalloc(reinterpret_cast<void**>(&L.d_logits), static_cast<size_t>(max_batch) * B2I_CLASSES * 4));
    chk(cudaMallocHost(reinterpret_cast<void**>(&L.h_logits), static_cast<size_t>(max_batch) * B2I_CLASSES * 4));
    chk(cudaEventCreateWithFlags(&L.done, cudaEventDisableTiming));
    L.copied.assign(chunks, nullptr);
    for (auto& ev : L.copied) chk(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
  }
  if (e != cudaSuccess) {
    *err = std::string("staging allocation: ") + cudaGetErrorString(e);
    return false;
  }
  return true;
}

void slot_free(DevSlot* s) {
  if (!s->dm) return;
  cudaSetDevice(s->dm->device());
  cudaDeviceSynchronize();
  for (Lane& L : s->lane) {
    if (L.d_in) cudaFree(L.d_in);
    if (L.h_stage) cudaFreeHost(L.h_stage);
    if (L.d_idx) cudaFree(L.d_idx);
    if (L.d_prob) cudaFree(L.d_prob);
    if (L.d_logits) cudaFree(L.d_logits);
    if (L.h_idx) cudaFreeHost(L.h_idx);
    if (L.h_prob) cudaFreeHost(L.h_prob);
    if (L.h_logits) cudaFreeHost(L.h_logits);
    if (L.done) cudaEventDestroy(L.done);
    for (auto ev : L.copied) if (ev) cudaEventDestroy(ev);
    L = Lane();
  }
  for (RawLane& R : s->rawlane) {
    if (R.raw.d_img) cudaFree(R.raw.d_img);
    if (R.raw.h_img) cudaFreeHost(R.raw.h_img);
    if (R.raw.d_tab) cudaFree(R.raw.d_tab);
    if (R.raw.h_tab) cudaFreeHost(R.raw.h_tab);
    if (R.raw.d_tmp) cudaFree(R.raw.d_tmp);
    if (R.raw.d_jimg) cudaFree(R.raw.d_jimg);
    if (R.jpeg) jpeg_decoder_destroy(R.jpeg);
    if (R.d_in) cudaFree(R.d_in);
    if (R.d_idx) cudaFree(R.d_idx);
    if (R.d_prob) cudaFree(R.d_prob);
    if (R.d_logits) cudaFree(R.d_logits);
    if (R.h_idx) cudaFreeHost(R.h_idx);
    if (R.h_prob) cudaFreeHost(R.h_prob);
    if (R.h_logits) cudaFreeHost(R.h_logits);
    if (R.staged) cudaEventDestroy(R.staged);
    if (R.st) cudaStreamDestroy(R.st);
    R.raw = RawScratch(); R.jpeg = nullptr; R.d_in = nullptr; R.d_idx = nullptr; R.d_prob = nullptr; R.d_logits = nullptr;
    R.h_idx = nullptr; R.h_prob = nullptr; R.h_logits = nullptr; R.staged = nullptr; R.st = nullptr; R.busy = false;
  }
  for (auto& kv : s->graphs) if (kv.second.exec) cudaGraphExecDestroy(kv.second.exec);
  s->graphs.clear();
  if (s->copy_st) cudaStreamDestroy(s->copy_st);
  if (s->comp_st) cudaStreamDestroy(s->comp_st);
  s->dm.reset();
}

int predict_submit(b2i_model* m, int first_dev, int n_dev, const uint8_t* images, int n, int k,
                   int32_t* idx, float* prob, float* logits, long long* ticket, std::string* err) {
  if (n < 0 || (n > 0 && (!images || !idx || !prob))) { *err = "null buffer"; return B2I_EINVAL; }
  if (k < 1 || k > 8) { *err = "k must be in 1..8"; return B2I_EINVAL; }
  if (first_dev < 0 || n_dev < 1 || first_dev + n_dev > static_cast<int>(m->slots.size())) {
    *err = "device range outside the model's context";
    return B2I_EINVAL;
  }
  if (static_cast<long long>(n) > static_cast<long long>(n_dev) * m->max_batch) {
    *err = "batch " + std::to_string(n) + " exceeds " + std::to_string(n_dev) + " GPU(s) x max_batch " +
           std::to_string(m->max_batch);
    return B2I_EINVAL;
  }
  const long long t = m->next_ticket;
  const int lane = static_cast<int>(t & 1);
  if (m->pending[lane].active) {
    *err = "two requests are already in flight on this model; collect ticket " +
           std::to_string(m->pending[lane].ticket) + " with b2i_wait first";
    return B2I_EINVAL;
  }
  // contiguous shards: GPU g takes images [g*per, min(n, (g+1)*per))
  const int per = (n + n_dev - 1) / n_dev;
  std::vector<Shard> sh(n_dev);
  auto work = [&](int g) {
    const int lo = std::min(n, g * per), hi = std::min(n, (g + 1) * per);
    enqueue_shard(m->slots[first_dev + g], lane, images + static_cast<size_t>(lo) * B2I_IMAGE_BYTES, hi - lo, k,
                  idx + static_cast<size_t>(lo) * k, prob + static_cast<size_t>(lo) * k,
                  logits ? logits + static_cast<size_t>(lo) * B2I_CLASSES : nullptr, &sh[g]);
  };
  if (n_dev == 1) {
    work(0);
  } else {
    std::vector<std::thread> th;   // one enqueuing thread per GPU: launches are CPU-bound at small batch
    for (int g = 0; g < n_dev; ++g) th.emplace_back(work, g);
    for (auto& x : th) x.join();
  }
  m->pending[lane].active = true;
  m->pending[lane].ticket = t;
  m->pending[lane].first_dev = first_dev;
  m->pending[lane].n_dev = n_dev;
  m->next_ticket = t + 1;
  *ticket = t;
  for (int g = 0; g < n_dev; ++g)
    if (sh[g].rc != B2I_OK) {
      // leave the lane collectable so the caller's b2i_wait (or the blocking wrapper) clears it
      *err = "GPU " + std::to_string(first_dev + g) + ": " + sh[g].err;
      for (int q = 0; q < n_dev; ++q) m->slots[first_dev + q].lane[lane].n = 0;
      m->pending[lane].active = false;
      return sh[g].rc;
    }
  return B2I_OK;
}

namespace {

bool grow(void** dev, void** host, size_t* cap, size_t need, std::string* err) {
  if (need <= *cap) return true;
  const size_t ncap = need + need / 4 + 4096;
  if (*dev) cudaFree(*dev);
  if (host && *host) cudaFreeHost(*host);
  *dev = nullptr;
  if (host) *host = nullptr;
  *cap = 0;
  cudaError_t e = cudaMalloc(dev, ncap);
  if (e == cudaSuccess && host) e = cudaMallocHost(host, ncap);
  if (e != cudaSuccess) { *err = std::string("raw-image scratch: ") + cudaGetErrorString(e); return false; }
  *cap = ncap;
  return true;
}

// The raw-image / JPEG path runs in two phases so that the host-heavy part of one call overlaps the GPU part of
// another (a member serves several `predict` RPCs at once, reference src/main.rs:74-79):
//   stage   (no model lock; exclusive use of one of the slot's two RawLanes)  header probe, resampling plans,
//           entropy decoding / the copy of the pixels, all GPU work on the LANE's stream
//   finish  (under the model lock, like every other forward)  resize + crop into the lane's device batch, the
//           network, the copy of the answers, on the slot's compute stream
// With two callers, caller B decodes its batch while caller A's batch is on the GPU.
struct StagedShard {
  DevSlot* slot = nullptr;
  RawLane* lane = nullptr;
  int n = 0, k = 0;
  bool from_jpeg = false;
  int32_t* idx = nullptr;
  float* prob = nullptr;
  float* logits = nullptr;
  int* status = nullptr;               // per image, caller's array (JPEG path) or null
  std::vector<int> own_status;
  std::vector<ResizeJob> jobs;
  std::vector<char> usable;
  std::string trace;
  Shard result;
};

RawLane* lane_acquire(DevSlot& s) {
  std::unique_lock<std::mutex> lock(s.raw_mu);
  for (;;) {
    for (RawLane& L : s.rawlane)
      if (!L.busy) { L.busy = true; return &L; }
    s.raw_cv.wait(lock);
  }
}
void lane_release(DevSlot& s, RawLane* L) {
  {
    std::lock_guard<std::mutex> lock(s.raw_mu);
    L->busy = false;
  }
  s.raw_cv.notify_one();
}

bool lane_setup(DevSlot& s, RawLane& L, int max_batch, bool want_logits, std::string* err) {
  cudaError_t e = cudaSuccess;
  auto chk = [&](cudaError_t r) { if (e == cudaSuccess) e = r; };
  if (!L.st) {
    const size_t nk = static_cast<size_t>(max_batch) * 8;
    chk(cudaStreamCreateWithFlags(&L.st, cudaStreamNonBlocking));
    chk(cudaEventCreateWithFlags(&L.staged, cudaEventDisableTiming));
    chk(cudaMalloc(reinterpret_cast<void**>(&L.d_in), static_cast<size_t>(max_batch) * B2I_IMAGE_BYTES));
    chk(cudaMalloc(reinterpret_cast<void**>(&L.d_idx), nk * 4));
    chk(cudaMalloc(reinterpret_cast<void**>(&L.d_prob), nk * 4));
    chk(cudaMallocHost(reinterpret_cast<void**>(&L.h_idx), nk * 4));
    chk(cudaMallocHost(reinterpret_cast<void**>(&L.h_prob), nk * 4));
  }
  if (want_logits && !L.d_logits) {
    chk(cudaMalloc(reinterpret_cast<void**>(&L.d_logits), static_cast<size_t>(max_batch) * B2I_CLASSES * 4));
    chk(cudaMallocHost(reinterpret_cast<void**>(&L.h_logits), static_cast<size_t>(max_batch) * B2I_CLASSES * 4));
  }
  if (e != cudaSuccess) { *err = std::string("raw-path staging: ") + cudaGetErrorString(e); return false; }
  (void)s;
  return true;
}

// Phase 1 for one GPU's share.
//   images != null: decoded pixels in host memory, sizes in hs_in / ws_in
//   jpeg   != null: encoded streams; decoded on this GPU (jpeg_decode.h) straight into the lane's packed image
//                   buffer, sizes from the header probe, one status per image
void raw_stage(DevSlot& s, int device, int max_batch, const uint8_t* const* images, const uint8_t* const* jpeg,
               const size_t* jpeg_len, const int* hs_in, const int* ws_in, int n, int k, int32_t* idx, float* prob,
               float* logits, int* status, StagedShard* sh) {
  sh->slot = &s; sh->n = n; sh->k = k; sh->from_jpeg = jpeg != nullptr;
  sh->idx = idx; sh->prob = prob; sh->logits = logits; sh->status = status;
  if (n == 0) return;
  Shard* out = &sh->result;
  auto cu = [&](cudaError_t e, const char* what) {
    if (e == cudaSuccess) return false;
    out->rc = B2I_ECUDA;
    out->err = std::string(what) + ": " + cudaGetErrorString(e);
    return true;
  };
  if (cu(cudaSetDevice(device), "cudaSetDevice")) return;
  // B2I_TRACE=1: wall time of every host-side stage of this call on stderr (where does a batch from files go?)
  static const bool trace = [] { const char* e = getenv("B2I_TRACE"); return e && e[0] == '1'; }();
  using clk = std::chrono::steady_clock;
  auto t_last = clk::now();
  auto stamp = [&](const char* what) {
    if (!trace) return;
    const auto now = clk::now();
    char b[64];
    snprintf(b, sizeof b, " %s %.2f ms;", what, std::chrono::duration<double, std::milli>(now - t_last).count());
    sh->trace += b;
    t_last = now;
  };
  RawLane* Lp = lane_acquire(s);
  sh->lane = Lp;
  RawLane& L = *Lp;
  stamp("lane");
  if (!lane_setup(s, L, max_batch, logits != nullptr, &out->err)) { out->rc = B2I_ENOMEM; return; }
  RawScratch& R = L.raw;
  std::vector<int> hbuf, wbuf;
  const int* hs = hs_in;
  const int* ws = ws_in;
  if (jpeg) {
    if (!L.jpeg) {
      L.jpeg = jpeg_decoder_create(&out->err);
      if (!L.jpeg) { out->rc = B2I_ENODEV; return; }
    }
    hbuf.resize(n); wbuf.resize(n);
    if (!status) { sh->own_status.resize(n); status = sh->own_status.data(); sh->status = status; }
    jpeg_probe(L.jpeg, jpeg, jpeg_len, n, hbuf.data(), wbuf.data(), status);
    hs = hbuf.data();
    ws = wbuf.data();
    stamp("header probe");
  }
  sh->usable.assign(n, 1);
  if (jpeg)
    for (int i = 0; i < n; ++i) sh->usable[i] = status[i] == B2I_OK;
  auto usable = [&](int i) { return sh->usable[i] != 0; };
  std::vector<AxisPlan> px(n), py(n);
  std::vector<size_t> img_off(n), tab_off(n);
  size_t img_bytes = 0, tab_bytes = 0, tmp_elems = 0;
  for (int i = 0; i < n; ++i)
    if (usable(i) && !jpeg && (!images[i] || hs[i] < 1 || ws[i] < 1)) { out->rc = B2I_EINVAL; out->err = "image " + std::to_string(i) + " is empty"; return; }
  {
    // the resampling plans are independent per image: a few helper threads for a large batch
    const int nt = n >= 32 ? std::min<int>(8, std::max(1u, std::thread::hardware_concurrency())) : 1;
    auto plan_range = [&](int t) {
      for (int i = t; i < n; i += nt)
        if (usable(i)) plan_image(hs[i], ws[i], &px[i], &py[i]);
    };
    std::vector<std::thread> th;
    for (int t = 1; t < nt; ++t) th.emplace_back(plan_range, t);
    plan_range(0);
    for (auto& x : th) x.join();
  }
  stamp("resampling plans");
  for (int i = 0; i < n; ++i) {
    img_off[i] = img_bytes;
    tab_off[i] = tab_bytes;
    if (!usable(i)) continue;
    img_bytes += (static_cast<size_t>(hs[i]) * ws[i] * 3 + 255) & ~static_cast<size_t>(255);
    tab_bytes += ((4 * 224 * 4 + (px[i].idx.size() + py[i].idx.size()) * 8) + 255) & ~static_cast<size_t>(255);
    tmp_elems = std::max(tmp_elems, static_cast<size_t>(hs[i]) * 224 * 3);
  }
  if (!(jpeg ? grow(reinterpret_cast<void**>(&R.d_jimg), nullptr, &R.jimg_cap, img_bytes, &out->err)
             : grow(reinterpret_cast<void**>(&R.d_img), reinterpret_cast<void**>(&R.h_img), &R.img_cap, img_bytes, &out->err)) ||
      !grow(reinterpret_cast<void**>(&R.d_tab), reinterpret_cast<void**>(&R.h_tab), &R.tab_cap, tab_bytes, &out->err) ||
      !grow(reinterpret_cast<void**>(&R.d_tmp), nullptr, &R.tmp_cap, tmp_elems * 4, &out->err)) {
    out->rc = B2I_ENOMEM;
    return;
  }
  sh->jobs.assign(n, ResizeJob());
  {
    // per image: the pixels into pinned staging (decoded-pixel path: 0.5 MB per ImageNet-sized image, so a few
    // threads for a large batch) and the resampling tables behind one another in the table buffer
    const int nt = n >= 32 ? std::min<int>(8, std::max(1u, std::thread::hardware_concurrency())) : 1;
    auto fill_range = [&](int t) {
      for (int i = t; i < n; i += nt) {
        if (!usable(i)) continue;
        if (!jpeg) std::memcpy(R.h_img + img_off[i], images[i], static_cast<size_t>(hs[i]) * ws[i] * 3);
        uint8_t* ht = R.h_tab + tab_off[i];
        uint8_t* dt = R.d_tab + tab_off[i];
        size_t o = 0;
        auto put = [&](const void* src, size_t bytes) { std::memcpy(ht + o, src, bytes); const size_t at = o; o += bytes; return at; };
        ResizeJob& j = sh->jobs[i];
        j.d_img = (jpeg ? R.d_jimg : R.d_img) + img_off[i]; j.h = hs[i]; j.w = ws[i];
        j.xf = reinterpret_cast<const int*>(dt + put(px[i].first.data(), 224 * 4));
        j.xc = reinterpret_cast<const int*>(dt + put(px[i].count.data(), 224 * 4));
        j.yf = reinterpret_cast<const int*>(dt + put(py[i].first.data(), 224 * 4));
        j.yc = reinterpret_cast<const int*>(dt + put(py[i].count.data(), 224 * 4));
        j.xi = reinterpret_cast<const int*>(dt + put(px[i].idx.data(), px[i].idx.size() * 4));
        j.yi = reinterpret_cast<const int*>(dt + put(py[i].idx.data(), py[i].idx.size() * 4));
        j.xw = reinterpret_cast<const float*>(dt + put(px[i].coef.data(), px[i].coef.size() * 4));
        j.yw = reinterpret_cast<const float*>(dt + put(py[i].coef.data(), py[i].coef.size() * 4));
        j.d_tmp = R.d_tmp;
        j.d_out = L.d_in + static_cast<size_t>(i) * B2I_IMAGE_BYTES;
      }
    };
    std::vector<std::thread> th;
    for (int t = 1; t < nt; ++t) th.emplace_back(fill_range, t);
    fill_range(0);
    for (auto& x : th) x.join();
  }
  stamp("tables");
  if (cu(cudaMemcpyAsync(R.d_tab, R.h_tab, tab_bytes, cudaMemcpyHostToDevice, L.st), "H2D resize tables")) return;
  if (jpeg) {
    std::vector<uint8_t*> dst(n);
    for (int i = 0; i < n; ++i) dst[i] = R.d_jimg + img_off[i];
    if (!jpeg_decode_batch(L.jpeg, jpeg, jpeg_len, n, dst.data(), hs, ws, status, L.st, &out->err)) { out->rc = B2I_ECUDA; return; }
    for (int i = 0; i < n; ++i) sh->usable[i] = status[i] == B2I_OK;   // streams that turned out to be damaged
    stamp("decode (host side)");
  } else if (cu(cudaMemcpyAsync(R.d_img, R.h_img, img_bytes, cudaMemcpyHostToDevice, L.st), "H2D raw images")) {
    return;
  }
  cu(cudaEventRecord(L.staged, L.st), "cudaEventRecord(staged)");
}

// Phase 2 for one GPU's share; the caller holds the model lock.
void raw_finish(StagedShard* sh) {
  Shard* out = &sh->result;
  if (sh->n == 0 || out->rc != B2I_OK || !sh->lane) return;
  DevSlot& s = *sh->slot;
  RawLane& L = *sh->lane;
  const int n = sh->n, k = sh->k;
  auto cu = [&](cudaError_t e, const char* what) {
    if (e == cudaSuccess) return false;
    out->rc = B2I_ECUDA;
    out->err = std::string(what) + ": " + cudaGetErrorString(e);
    return true;
  };
  static const bool trace = [] { const char* e = getenv("B2I_TRACE"); return e && e[0] == '1'; }();
  const auto t0 = std::chrono::steady_clock::now();
  if (cu(cudaSetDevice(s.dm->device()), "cudaSetDevice")) return;
  if (n > s.dm->max_batch()) { out->rc = B2I_EINVAL; out->err = "shard exceeds max_batch"; return; }
  if (cu(cudaStreamWaitEvent(s.comp_st, L.staged, 0), "cudaStreamWaitEvent(staged)")) return;
  for (int i = 0; i < n; ++i)
    if (sh->usable[i] && cu(launch_resize_job(sh->jobs[i], s.comp_st), "resize kernels")) return;
  const int classes = s.dm->classes();
  for (int off = 0; off < n; off += s.chunk) {
    const int nb = std::min(s.chunk, n - off);
    if (!s.dm->forward(L.d_in + static_cast<size_t>(off) * B2I_IMAGE_BYTES, nb, k, L.d_idx + static_cast<size_t>(off) * k,
                       L.d_prob + static_cast<size_t>(off) * k,
                       sh->logits ? L.d_logits + static_cast<size_t>(off) * classes : nullptr, s.comp_st, &out->err)) {
      out->rc = B2I_ECUDA;
      return;
    }
  }
  const size_t nk = static_cast<size_t>(n) * k;
  if (cu(cudaMemcpyAsync(L.h_idx, L.d_idx, nk * 4, cudaMemcpyDeviceToHost, s.comp_st), "D2H idx")) return;
  if (cu(cudaMemcpyAsync(L.h_prob, L.d_prob, nk * 4, cudaMemcpyDeviceToHost, s.comp_st), "D2H prob")) return;
  if (sh->logits && cu(cudaMemcpyAsync(L.h_logits, L.d_logits, static_cast<size_t>(n) * classes * 4, cudaMemcpyDeviceToHost, s.comp_st), "D2H logits")) return;
  const auto t1 = std::chrono::steady_clock::now();
  if (cu(cudaStreamSynchronize(s.comp_st), "cudaStreamSynchronize")) return;
  if (trace) {
    const auto t2 = std::chrono::steady_clock::now();
    fprintf(stderr, "[b2i trace] %s shard of %d on GPU %d:%s launches %.2f ms; wait for the GPU %.2f ms\n", sh->from_jpeg ? "jpeg" : "raw", n,
            s.dm->device(), sh->trace.c_str(), std::chrono::duration<double, std::milli>(t1 - t0).count(),
            std::chrono::duration<double, std::milli>(t2 - t1).count());
  }
  std::memcpy(sh->idx, L.h_idx, nk * 4);
  std::memcpy(sh->prob, L.h_prob, nk * 4);
  if (sh->logits) std::memcpy(sh->logits, L.h_logits, static_cast<size_t>(n) * classes * 4);
  for (int i = 0; i < n; ++i)
    if (!sh->usable[i]) {   // its slot of the device batch held no image: blank the answer
      for (int j = 0; j < k; ++j) { sh->idx[static_cast<size_t>(i) * k + j] = -1; sh->prob[static_cast<size_t>(i) * k + j] = 0.f; }
      if (sh->logits) std::fill(sh->logits + static_cast<size_t>(i) * classes, sh->logits + static_cast<size_t>(i + 1) * classes, 0.f);
    }
}

// Both phases over the GPUs [first_dev, first_dev + n_dev): staging without the model lock, the forwards under it.
int raw_run(b2i_model* m, int first_dev, int n_dev, const uint8_t* const* images, const uint8_t* const* jpeg, const size_t* len,
            const int* heights, const int* widths, int n, int k, int32_t* idx, float* prob, float* logits, int* status,
            std::string* err) {
  if (k < 1 || k > 8) { *err = "k must be in 1..8"; return B2I_EINVAL; }
  if (first_dev < 0 || n_dev < 1 || first_dev + n_dev > static_cast<int>(m->slots.size())) { *err = "device range outside the model's context"; return B2I_EINVAL; }
  if (static_cast<long long>(n) > static_cast<long long>(n_dev) * m->max_batch) { *err = "batch exceeds GPUs x max_batch"; return B2I_EINVAL; }
  const int per = (n + n_dev - 1) / n_dev;
  std::vector<StagedShard> sh(n_dev);
  auto each = [&](auto&& fn) {
    if (n_dev == 1) { fn(0); return; }
    std::vector<std::thread> th;
    for (int g = 0; g < n_dev; ++g) th.emplace_back(fn, g);
    for (auto& x : th) x.join();
  };
  each([&](int g) {
    const int lo = std::min(n, g * per), hi = std::min(n, (g + 1) * per);
    raw_stage(m->slots[first_dev + g], m->ctx->devices[first_dev + g], m->max_batch, images ? images + lo : nullptr,
              jpeg ? jpeg + lo : nullptr, jpeg ? len + lo : nullptr, heights ? heights + lo : nullptr, widths ? widths + lo : nullptr,
              hi - lo, k, idx + static_cast<size_t>(lo) * k, prob + static_cast<size_t>(lo) * k,
              logits ? logits + static_cast<size_t>(lo) * B2I_CLASSES : nullptr, status ? status + lo : nullptr, &sh[g]);
  });
  int rc = B2I_OK;
  {
    std::lock_guard<std::mutex> lock(m->mu);   // forwards of one model are serialised, whatever path they come from
    if (m->pending[0].active || m->pending[1].active) {
      *err = "an asynchronous request is still in flight; b2i_wait for it first";
      rc = B2I_EINVAL;
    } else {
      each([&](int g) { raw_finish(&sh[g]); });
    }
  }
  for (int g = 0; g < n_dev; ++g) {
    if (sh[g].lane) {
      if (rc != B2I_OK || sh[g].result.rc != B2I_OK) {   // nothing may still be writing into the lane when it is handed on
        cudaSetDevice(m->ctx->devices[first_dev + g]);
        cudaStreamSynchronize(sh[g].lane->st);
      }
      lane_release(m->slots[first_dev + g], sh[g].lane);
    }
    if (rc == B2I_OK && sh[g].result.rc != B2I_OK) { *err = "GPU " + std::to_string(first_dev + g) + ": " + sh[g].result.err; rc = sh[g].result.rc; }
  }
  return rc;
}

}  // namespace

int predict_jpeg(b2i_model* m, int first_dev, int n_dev, const uint8_t* const* jpeg, const size_t* len, int n, int k,
                 int32_t* idx, float* prob, float* logits, int* status, std::string* err) {
  if (n < 0 || (n > 0 && (!jpeg || !len || !idx || !prob || !status))) { *err = "null buffer"; return B2I_EINVAL; }
  return raw_run(m, first_dev, n_dev, nullptr, jpeg, len, nullptr, nullptr, n, k, idx, prob, logits, status, err);
}

int predict_raw(b2i_model* m, int first_dev, int n_dev, const uint8_t* const* images, const int* heights,
                const int* widths, int n, int k, int32_t* idx, float* prob, float* logits, std::string* err) {
  if (n < 0 || (n > 0 && (!images || !heights || !widths || !idx || !prob))) { *err = "null buffer"; return B2I_EINVAL; }
  return raw_run(m, first_dev, n_dev, images, nullptr, nullptr, heights, widths, n, k, idx, prob, logits, nullptr, err);
}

int predict_raw_status(b2i_model* m, int first_dev, int n_dev, const uint8_t* const* images, const int* heights,
                       const int* widths, int n, int k, int32_t* idx, float* prob, float* logits, int* status,
                       std::string* err) {
  if (n < 0 || (n > 0 && (!images || !heights || !widths || !idx || !prob || !status))) { *err = "null buffer"; return B2I_EINVAL; }
  if (k < 1 || k > 8) { *err = "k must be in 1..8"; return B2I_EINVAL; }
  // images that cannot be used are set aside with a status of their own; the rest go through as one batch
  std::vector<const uint8_t*> img;
  std::vector<int> hs, ws, at;
  for (int i = 0; i < n; ++i) {
    if (!images[i] || heights[i] < 1 || widths[i] < 1) { status[i] = B2I_EDECODE; continue; }
    if (heights[i] > kMaxRawSide || widths[i] > kMaxRawSide) { status[i] = B2I_EINVAL; continue; }
    status[i] = B2I_OK;
    img.push_back(images[i]); hs.push_back(heights[i]); ws.push_back(widths[i]); at.push_back(i);
  }
  const int nv = static_cast<int>(at.size());
  for (int i = 0; i < n; ++i)
    if (status[i] != B2I_OK) {
      for (int j = 0; j < k; ++j) { idx[static_cast<size_t>(i) * k + j] = -1; prob[static_cast<size_t>(i) * k + j] = 0.f; }
      if (logits) std::fill(logits + static_cast<size_t>(i) * B2I_CLASSES, logits + static_cast<size_t>(i + 1) * B2I_CLASSES, 0.f);
    }
  if (nv == 0) return B2I_OK;
  if (nv == n) return predict_raw(m, first_dev, n_dev, images, heights, widths, n, k, idx, prob, logits, err);
  std::vector<int32_t> ci(static_cast<size_t>(nv) * k);
  std::vector<float> cp(static_cast<size_t>(nv) * k), cl(logits ? static_cast<size_t>(nv) * B2I_CLASSES : 0);
  const int rc = predict_raw(m, first_dev, n_dev, img.data(), hs.data(), ws.data(), nv, k, ci.data(), cp.data(),
                             logits ? cl.data() : nullptr, err);
  if (rc != B2I_OK) return rc;
  for (int v = 0; v < nv; ++v) {
    const size_t i = static_cast<size_t>(at[v]);
    std::memcpy(idx + i * k, &ci[static_cast<size_t>(v) * k], static_cast<size_t>(k) * 4);
    std::memcpy(prob + i * k, &cp[static_cast<size_t>(v) * k], static_cast<size_t>(k) * 4);
    if (logits) std::memcpy(logits + i * B2I_CLASSES, &cl[static_cast<size_t>(v) * B2I_CLASSES], B2I_CLASSES * 4);
  }
  return B2I_OK;
}

int predict_wait(b2i_model* m, long long ticket, std::string* err) {
  const int lane = static_cast<int>(ticket & 1);
  Pending& p = m->pending[lane];
  if (ticket < 0 || !p.active || p.ticket != ticket) {
    *err = "ticket " + std::to_string(ticket) + " is not in flight";
    return B2I_EINVAL;
  }
  std::vector<Shard> sh(p.n_dev);
  for (int g = 0; g < p.n_dev; ++g) collect_shard(m->slots[p.first_dev + g], lane, &sh[g]);
  p.active = false;
  for (int g = 0; g < p.n_dev; ++g)
    if (sh[g].rc != B2I_OK) {
      *err = "GPU " + std::to_string(p.first_dev + g) + ": " + sh[g].err;
      return sh[g].rc;
    }
  return B2I_OK;
}

}  // namespace b2i
