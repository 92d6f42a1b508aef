// Host-side engine: network description (ResNet-18 / AlexNet as the reference builds them with
// tch::vision::{resnet::resnet18, alexnet::alexnet}, reference src/services.rs:513-524; layer
// tables in SURVEY.md Appendix B/C), BatchNorm folding + weight re-layout, per-GPU weight upload,
// static activation-buffer plan and the per-batch kernel sequence.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <map>
#include <memory>
#include <string>
#include <vector>

#include "conv.h"
#include "ops.h"
#include "ot_reader.h"

namespace b2i {

enum Precision : int { PREC_BF16 = 0, PREC_FP32 = 1, PREC_FP16 = 2 };   // FP16: the tcgen05 path with IEEE-half tensors
enum AlexGeom : int { ALEX_TCH = 0, ALEX_TORCHVISION = 1 };

struct EngineConfig {
  int precision = PREC_BF16;
  int max_batch = 256;          // largest batch one forward() call accepts per GPU
  int micro_batch = 0;          // images per kernel sequence; 0 = max_batch
  int alexnet_geometry = ALEX_TCH;
  int conv_amode = -1;          // -1 auto (im2col TMA where C % 64 == 0), else AMode for those layers
  int b_tma = 1;                // weights through TMA
  int bn_max = 256;             // largest tile-N the planner may choose
  int fused_stem = 1;           // ResNet bf16: normalise + conv1 + max-pool as one kernel
  int fuse_downsample = 1;      // 1x1/s2 downsample convs ride on their sibling 3x3/s2 conv's kernel
  int pair_mma = 256;             // layers with tile-N >= 128 through the CTA-pair (cta_group::2) kernel
  int patch_conv = 1;           // 64->64 3x3/s1 layers through the halo-patch kernel
  int persistent = 1;           // TMA-fed layers: persistent kernel (0 = one tile per CTA)
  int keep_tensors = 0;         // 1: every activation gets its own buffer (per-layer parity tests)
};

// One convolution / FC with BatchNorm already folded, in the kernels' layout.
struct ConvSpec {
  std::string name;
  int Cin = 0, Cin_pad = 0, Cout = 0, Cout_pad = 0;
  int R = 1, S = 1, Sp = 1, stride = 1, pad = 0, K = 0;
  bool smallc = false;
  std::vector<float> w;     // [Cout_pad][K], k = (r*Sp + s)*Cin_pad + c
  std::vector<float> bias;  // [Cout_pad]
};

// OIHW fp32 weights (+ optional conv bias, + optional BN gamma/beta/mean/var, eps 1e-5) -> ConvSpec.
// `chw_to_hwc` > 0 re-orders an FC input axis laid out (C, H, W) to (H, W, C) with that C
// (AlexNet classifier.1 after an NHWC flatten).
void build_conv_spec(ConvSpec* cs, const std::string& name, const float* w_oihw, const float* bias,
                     const float* bn_gamma, const float* bn_beta, const float* bn_mean,
                     const float* bn_var, int Cout, int Cin, int R, int S, int stride, int pad,
                     int chw_to_hwc_channels = 0, int s_pad = 0);

struct TensorShape { int H = 0, W = 0, C = 0; size_t elems() const { return size_t(H) * W * C; } };

struct OpSpec {
  enum Kind { NORMALIZE, CONV, MAXPOOL, AVGPOOL, STEM, STEM_ALEX } kind = CONV;
  int in = -1, out = -1, res = -1;  // activation tensor ids
  int conv = -1;                    // index into NetSpec::convs
  bool relu = false, fc = false, logits = false;
  int k = 0, stride = 0, pad = 0;   // pooling
};

struct NetSpec {
  std::string model;
  std::vector<ConvSpec> convs;
  std::vector<TensorShape> tensors;  // tensor 0 is the normalised input
  std::vector<OpSpec> ops;
  int logits_tensor = -1;
  int classes = 1000;
  double flops_per_image = 0;        // 2 * MACs of conv + FC, no padding counted
};

bool build_resnet18(const std::map<std::string, HostTensor>& t, bool fused_stem, NetSpec* net,
                    std::string* err);
bool build_alexnet(const std::map<std::string, HostTensor>& t, int geometry, bool fused_stem, NetSpec* net,
                   std::string* err);

// Device-resident instance of a NetSpec on one GPU.
class DeviceModel {
 public:
  DeviceModel() = default;
  ~DeviceModel();
  DeviceModel(const DeviceModel&) = delete;
  DeviceModel& operator=(const DeviceModel&) = delete;

  bool init(int device, const NetSpec& net, const EngineConfig& cfg, std::string* err);
  // Inputs/outputs are DEVICE pointers on this model's GPU; enqueues on `st` and returns.
  //   u8: [n,224,224,3]; idx/prob: [n,k]; logits: [n,classes] or null.
  bool forward(const uint8_t* d_u8, int n, int k, int32_t* d_idx, float* d_prob, float* d_logits,
               cudaStream_t st, std::string* err);
  // One micro-batch with a CUDA-event pair around every launch: per-launch milliseconds, in
  // execution order (ops..., softmax+top-k), plus a label per launch. Synchronises.
  bool profile(const uint8_t* d_u8, int n, int k, int32_t* d_idx, float* d_prob, cudaStream_t st,
               std::vector<float>* ms, std::vector<std::string>* labels, std::string* err);
  // Copy of an intermediate activation (debug / per-layer parity): tensor id -> fp32 NHWC host.
  bool read_tensor(int tensor_id, int n, std::vector<float>* out, std::string* err);

  int device() const { return device_; }
  int max_batch() const { return cfg_.max_batch; }
  int micro_batch() const { return mb_; }
  int classes() const { return classes_; }
  int num_tensors() const { return static_cast<int>(tshape_.size()); }
  TensorShape tensor_shape(int id) const { return tshape_[id]; }
  long long launches_per_microbatch() const {   // kernels actually launched: ops folded into a sibling's kernel do not count
    long long n = 1;                              // softmax + top-k
    for (const DevOp& d : ops_) n += d.skip ? 0 : 1;
    return n;
  }
  size_t weight_bytes() const { return weight_bytes_; }
  size_t activation_bytes() const { return act_bytes_; }

 private:
  struct DevOp {
    OpSpec spec;
    ConvGeom g{};
    void* w = nullptr;
    uint32_t* w2 = nullptr;         // fused ResNet stem, second form: packed (parity, channel) x K matrix
    float* bias = nullptr;
    std::vector<float> bias_host;   // fused stem only: its bias travels as kernel parameters
    CUtensorMap tmA{}, tmB{}, tmR{}, tmO{};
    bool has_tmA = false, has_tmB = false;
    int amode = 0, BN = 0, stages = 0, b_tma = 0, patch = 0, pair = 0, resident_b = 0, flat = 0, msub = 1;
    int splits = 1, a_rows = 128;   // FC layers: split-K factor and rows of the A box (see conv_umma.cu, persistent kernel)
    int fused_second = -1;   // index of a 1x1 conv computed inside this op's kernel
    bool skip = false;       // computed by another op's kernel
  };
  bool run_microbatch(const uint8_t* d_u8, int nb, int k, int32_t* d_idx, float* d_prob,
                      float* d_logits, cudaStream_t st, std::string* err);
  std::vector<cudaEvent_t>* prof_events_ = nullptr;  // when set, run_microbatch records after every launch
  std::vector<std::string> op_names_;

  int device_ = -1;
  EngineConfig cfg_;
  int mb_ = 0, classes_ = 1000;
  size_t esize_ = 2;
  std::vector<DevOp> ops_;
  std::vector<TensorShape> tshape_;
  std::vector<int> tensor_buf_;      // tensor id -> buffer id
  std::vector<void*> bufs_;
  std::vector<size_t> buf_bytes_;
  std::vector<void*> owned_;         // weight allocations
  unsigned int* stem_sched_ = nullptr;
  float* splitk_ws_ = nullptr;       // FC split-K: per-split partial sums (fp32), see conv_umma.cu
  int* splitk_ctr_ = nullptr;        // one arrival counter per output tile, zero between launches
  size_t splitk_ws_bytes_ = 0;
  int logits_tensor_ = -1;
  size_t weight_bytes_ = 0, act_bytes_ = 0;
};

}  // namespace b2i
