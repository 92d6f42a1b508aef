// Batched predict over the GPUs of a context: per-GPU lanes (two requests may be in flight so the
// host->device copy of request t+1 overlaps the forward of request t), contiguous batch sharding,
// pinned staging, results written back into the caller's arrays.  Replaces the reference's
// one-image-per-RPC scattering (reference src/services.rs:415-421) on one multi-GPU member.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <condition_variable>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "engine.h"

namespace b2i { struct JpegDecoder; }

struct b2i_ctx {
  std::vector<int> devices;
  std::mutex jpeg_mu;                          // guards the decoders below (b2i_jpeg_decode_u8)
  std::vector<b2i::JpegDecoder*> jpeg;         // one per device, created on first use
  ~b2i_ctx();
};

namespace b2i {

// One in-flight request's share on one GPU.
struct Lane {
  uint8_t* d_in = nullptr;       // device, max_batch images
  uint8_t* h_stage = nullptr;    // pinned, max_batch images; allocated the first time a pageable source shows up
  int32_t* d_idx = nullptr;
  float* d_prob = nullptr;
  float* d_logits = nullptr;     // device
  int32_t* h_idx = nullptr;      // pinned
  float* h_prob = nullptr;
  float* h_logits = nullptr;     // pinned
  std::vector<cudaEvent_t> copied;  // one per micro-batch chunk
  cudaEvent_t done = nullptr;
  // the request currently using this lane
  int n = 0, k = 0;
  int32_t* out_idx = nullptr;
  float* out_prob = nullptr;
  float* out_logits = nullptr;
};

// A captured forward: the kernel sequence of one (lane, chunk, batch, k, logits?) combination,
// replayed with a single cudaGraphLaunch.  Worth it where the sequence is launch-bound (small batch).
struct GraphEntry {
  int uses = 0;
  cudaGraphExec_t exec = nullptr;
};

// Grow-only buffers of the raw-image path (predict_raw): packed source images, resampling tables,
// horizontally resampled rows.
struct RawScratch {
  uint8_t* h_img = nullptr; uint8_t* d_img = nullptr; size_t img_cap = 0;   // pinned / device
  uint8_t* h_tab = nullptr; uint8_t* d_tab = nullptr; size_t tab_cap = 0;
  float* d_tmp = nullptr; size_t tmp_cap = 0;
  uint8_t* d_jimg = nullptr; size_t jimg_cap = 0;   // decoded images of the JPEG path (device only)
};

struct JpegDecoder;

// Everything one raw-image / JPEG request needs on one GPU before its forward: decoder, packed source images,
// resampling tables, the device batch the resize writes and the forward reads, result staging.  A slot has two,
// so that one request is being decoded while the other's forward runs (predict.cpp, raw_stage / raw_finish).
struct RawLane {
  RawScratch raw;
  JpegDecoder* jpeg = nullptr;                 // created by the lane's first JPEG request
  cudaStream_t st = nullptr;                   // decode kernels / pixel and table copies
  cudaEvent_t staged = nullptr;                // recorded on st when the lane's inputs are on the GPU
  uint8_t* d_in = nullptr;                     // max_batch cropped images
  int32_t* d_idx = nullptr; float* d_prob = nullptr; float* d_logits = nullptr;
  int32_t* h_idx = nullptr; float* h_prob = nullptr; float* h_logits = nullptr;   // pinned
  bool busy = false;                           // guarded by DevSlot::raw_mu
};

struct DevSlot {
  DevSlot() = default;
  DevSlot(DevSlot&& o) noexcept : dm(std::move(o.dm)), copy_st(o.copy_st), comp_st(o.comp_st), chunk(o.chunk),
                                  graph_max_batch(o.graph_max_batch), graphs(std::move(o.graphs)) {
    for (int i = 0; i < 2; ++i) { lane[i] = o.lane[i]; rawlane[i] = o.rawlane[i]; }
  }
  std::unique_ptr<DeviceModel> dm;
  RawLane rawlane[2];
  std::mutex raw_mu;
  std::condition_variable raw_cv;
  cudaStream_t copy_st = nullptr, comp_st = nullptr;
  Lane lane[2];
  int chunk = 0;
  int graph_max_batch = 64;                    // forwards of at most this many images are graph-replayed
  std::map<uint64_t, GraphEntry> graphs;
};

struct Pending {
  bool active = false;
  long long ticket = -1;
  int first_dev = 0, n_dev = 0;
};

}  // namespace b2i

struct b2i_model {
  b2i_ctx* ctx = nullptr;
  std::string name;
  double flops = 0;
  std::vector<b2i::DevSlot> slots;
  std::mutex mu;
  int max_batch = 0;
  long long next_ticket = 0;
  b2i::Pending pending[2];
  b2i::EngineConfig engine_cfg;   // kept for b2i_model_reload
};

namespace b2i {

bool slot_setup(DevSlot* s, int max_batch, std::string* err);   // streams, lanes, events (after dm is initialised)
void slot_free(DevSlot* s);

// Enqueue one request (H2D + forward + D2H of the results into pinned staging) on the GPUs
// [first_dev, first_dev + n_dev) and return its ticket without waiting.  Returns 0 or a B2I_E* code.
int predict_submit(b2i_model* m, int first_dev, int n_dev, const uint8_t* images, int n, int k,
                   int32_t* idx, float* prob, float* logits, long long* ticket, std::string* err);
// Blocking: decoded images of arbitrary size -> resize + centre crop on the GPU, straight into the
// device batch -> forward -> results.  images[i] is u8 [heights[i]][widths[i]][3] in host memory.
int predict_raw(b2i_model* m, int first_dev, int n_dev, const uint8_t* const* images, const int* heights,
                const int* widths, int n, int k, int32_t* idx, float* prob, float* logits, std::string* err);
// Same with one status per image (B2I_OK, B2I_EDECODE for a null / empty image, B2I_EINVAL for one larger
// than kMaxRawSide on a side): a bad image costs its own answer, not the batch's (the reference loses one
// query per bad file because it sends one id per RPC, reference src/services.rs:421,490-492).
constexpr int kMaxRawSide = 16384;
int predict_raw_status(b2i_model* m, int first_dev, int n_dev, const uint8_t* const* images, const int* heights,
                       const int* widths, int n, int k, int32_t* idx, float* prob, float* logits, int* status,
                       std::string* err);
// Blocking: encoded JPEG streams -> decode on the GPU (nvJPEG) -> resize + centre crop -> forward -> results,
// one status per image (B2I_EDECODE for a stream that cannot be decoded).
int predict_jpeg(b2i_model* m, int first_dev, int n_dev, const uint8_t* const* jpeg, const size_t* len, int n, int k,
                 int32_t* idx, float* prob, float* logits, int* status, std::string* err);
// Block until the request is complete and its results are in the arrays given at submit time.
int predict_wait(b2i_model* m, long long ticket, std::string* err);

}  // namespace b2i
