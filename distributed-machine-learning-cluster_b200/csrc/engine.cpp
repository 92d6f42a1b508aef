// See engine.h.
#include "engine.h"
#include "half_host.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>

namespace b2i {

namespace {

inline int round_up(int v, int m) { return (v + m - 1) / m * m; }

const HostTensor* find(const std::map<std::string, HostTensor>& t, const std::string& name,
                       std::string* err) {
  auto it = t.find(name);
  if (it == t.end()) {
    *err = "weight archive is missing tensor '" + name + "'";
    return nullptr;
  }
  return &it->second;
}

bool shape_is(const HostTensor& h, std::initializer_list<int64_t> want) {
  return h.shape.size() == want.size() && std::equal(want.begin(), want.end(), h.shape.begin());
}

std::string shape_str(const HostTensor& h) {
  std::string s = "[";
  for (size_t i = 0; i < h.shape.size(); ++i) s += (i ? "," : "") + std::to_string(h.shape[i]);
  return s + "]";
}

// conv (no bias) + BN, the ResNet building block
bool add_conv_bn(const std::map<std::string, HostTensor>& t, const std::string& conv,
                 const std::string& bn, int Cout, int Cin, int k, int stride, int pad,
                 NetSpec* net, std::string* err, int s_pad = 0) {
  const HostTensor* w = find(t, conv + ".weight", err);
  const HostTensor* g = find(t, bn + ".weight", err);
  const HostTensor* b = find(t, bn + ".bias", err);
  const HostTensor* m = find(t, bn + ".running_mean", err);
  const HostTensor* v = find(t, bn + ".running_var", err);
  if (!w || !g || !b || !m || !v) return false;
  if (!shape_is(*w, {Cout, Cin, k, k})) {
    *err = conv + ".weight has shape " + shape_str(*w) + ", expected [" + std::to_string(Cout) +
           "," + std::to_string(Cin) + "," + std::to_string(k) + "," + std::to_string(k) + "]";
    return false;
  }
  for (const HostTensor* p : {g, b, m, v})
    if (!shape_is(*p, {Cout})) { *err = bn + " parameter has shape " + shape_str(*p); return false; }
  net->convs.emplace_back();
  build_conv_spec(&net->convs.back(), conv, w->data.data(), nullptr, g->data.data(),
                  b->data.data(), m->data.data(), v->data.data(), Cout, Cin, k, k, stride, pad, 0, s_pad);
  return true;
}

bool add_conv_bias(const std::map<std::string, HostTensor>& t, const std::string& name, int Cout,
                   int Cin, int k, int stride, int pad, int chw_c, NetSpec* net,
                   std::string* err) {
  const HostTensor* w = find(t, name + ".weight", err);
  const HostTensor* b = find(t, name + ".bias", err);
  if (!w || !b) return false;
  const bool ok = (k == 0) ? shape_is(*w, {Cout, Cin}) : shape_is(*w, {Cout, Cin, k, k});
  if (!ok || !shape_is(*b, {Cout})) {
    *err = name + " has unexpected shape " + shape_str(*w) + " / " + shape_str(*b);
    return false;
  }
  net->convs.emplace_back();
  const int kk = k == 0 ? 1 : k;
  build_conv_spec(&net->convs.back(), name, w->data.data(), b->data.data(), nullptr, nullptr,
                  nullptr, nullptr, Cout, Cin, kk, kk, stride, pad, chw_c);
  return true;
}

int new_tensor(NetSpec* net, int H, int W, int C) {
  TensorShape s;
  s.H = H; s.W = W; s.C = C;
  net->tensors.push_back(s);
  return static_cast<int>(net->tensors.size()) - 1;
}

int conv_out(int in, int k, int stride, int pad) { return (in + 2 * pad - k) / stride + 1; }

// Appends a conv op reading tensor `in`; returns the output tensor id.
int push_conv(NetSpec* net, int conv_idx, int in, int res, bool relu) {
  const ConvSpec& c = net->convs[conv_idx];
  const TensorShape ti = net->tensors[in];
  const int OH = conv_out(ti.H, c.R, c.stride, c.pad), OW = conv_out(ti.W, c.S, c.stride, c.pad);
  OpSpec op;
  op.kind = OpSpec::CONV;
  op.in = in; op.res = res; op.conv = conv_idx; op.relu = relu;
  op.out = new_tensor(net, OH, OW, c.Cout);
  net->ops.push_back(op);
  net->flops_per_image += 2.0 * OH * OW * c.Cout * c.Cin * c.R * c.S;
  return op.out;
}

int push_fc(NetSpec* net, int conv_idx, int in, bool relu, bool logits) {
  const ConvSpec& c = net->convs[conv_idx];
  OpSpec op;
  op.kind = OpSpec::CONV;
  op.in = in; op.conv = conv_idx; op.relu = relu; op.fc = true; op.logits = logits;
  op.out = new_tensor(net, 1, 1, c.Cout);
  if (logits) net->logits_tensor = op.out;
  net->ops.push_back(op);
  net->flops_per_image += 2.0 * c.Cout * c.Cin;
  return op.out;
}

int push_pool(NetSpec* net, OpSpec::Kind kind, int in, int k, int stride, int pad, int OH, int OW) {
  OpSpec op;
  op.kind = kind;
  op.in = in; op.k = k; op.stride = stride; op.pad = pad;
  op.out = new_tensor(net, OH, OW, net->tensors[in].C);
  net->ops.push_back(op);
  return op.out;
}

}  // namespace

void build_conv_spec(ConvSpec* cs, const std::string& name, const float* w, const float* bias,
                     const float* g, const float* b, const float* m, const float* v, int Cout,
                     int Cin, int R, int S, int stride, int pad, int chw_c, int s_pad) {
  cs->name = name;
  cs->Cin = Cin; cs->Cout = Cout; cs->R = R; cs->S = S; cs->Sp = s_pad > S ? s_pad : S; cs->stride = stride;
  cs->pad = pad;
  cs->smallc = Cin <= 4;  // first layers (RGB): one 8-byte pixel per filter tap
  cs->Cin_pad = cs->smallc ? 4 : round_up(Cin, 64);
  cs->Cout_pad = round_up(Cout, 64);
  cs->K = round_up(R * cs->Sp * cs->Cin_pad, 64);
  cs->w.assign(static_cast<size_t>(cs->Cout_pad) * cs->K, 0.f);
  cs->bias.assign(cs->Cout_pad, 0.f);
  const int hw = chw_c > 0 ? Cin / chw_c : 1;
  for (int co = 0; co < Cout; ++co) {
    float scale = 1.f, shift = bias ? bias[co] : 0.f;
    if (g) {
      // eval-mode batch_norm: y = (x - mean) / sqrt(var + eps) * gamma + beta, eps = 1e-5
      const double inv = 1.0 / std::sqrt(static_cast<double>(v[co]) + 1e-5);
      scale = static_cast<float>(g[co] * inv);
      shift = static_cast<float>(b[co] - m[co] * g[co] * inv + (bias ? bias[co] * g[co] * inv : 0.0));
    }
    cs->bias[co] = shift;
    float* dst = cs->w.data() + static_cast<size_t>(co) * cs->K;
    for (int c = 0; c < Cin; ++c)
      for (int r = 0; r < R; ++r)
        for (int s = 0; s < S; ++s) {
          const float val = w[((static_cast<size_t>(co) * Cin + c) * R + r) * S + s] * scale;
          int kc = c;
          if (chw_c > 0) kc = (c % hw) * chw_c + (c / hw);  // (c', hw') -> hw'*C + c'
          dst[(r * cs->Sp + s) * cs->Cin_pad + kc] = val;
        }
  }
}

bool build_resnet18(const std::map<std::string, HostTensor>& t, bool fused_stem, NetSpec* net,
                    std::string* err) {
  net->model = "resnet18";
  int cur;
  if (fused_stem) {
    // conv1 weights in the stem kernel's K order: (r, s padded to 8, c padded to 4) = 224 -> 256
    if (!add_conv_bn(t, "conv1", "bn1", 64, 3, 7, 2, 3, net, err, /*s_pad=*/8)) return false;
    OpSpec op;
    op.kind = OpSpec::STEM;
    op.conv = 0;
    op.relu = true;
    op.out = new_tensor(net, 56, 56, 64);
    net->ops.push_back(op);
    net->flops_per_image += 2.0 * 112 * 112 * 64 * 3 * 49;
    cur = op.out;
  } else {
    const int in0 = new_tensor(net, 224, 224, 4);
    OpSpec op;
    op.kind = OpSpec::NORMALIZE;
    op.out = in0;
    net->ops.push_back(op);
    if (!add_conv_bn(t, "conv1", "bn1", 64, 3, 7, 2, 3, net, err)) return false;
    cur = push_conv(net, 0, in0, -1, true);
    cur = push_pool(net, OpSpec::MAXPOOL, cur, 3, 2, 1, 56, 56);
  }
  int cin = 64;
  const int widths[4] = {64, 128, 256, 512};
  for (int L = 0; L < 4; ++L) {
    const int cout = widths[L];
    for (int B = 0; B < 2; ++B) {
      const int stride = (L > 0 && B == 0) ? 2 : 1;
      const std::string p = "layer" + std::to_string(L + 1) + "." + std::to_string(B);
      int identity = cur;
      if (stride != 1 || cin != cout) {
        if (!add_conv_bn(t, p + ".downsample.0", p + ".downsample.1", cout, cin, 1, stride, 0, net, err))
          return false;
        identity = push_conv(net, static_cast<int>(net->convs.size()) - 1, cur, -1, false);
      }
      if (!add_conv_bn(t, p + ".conv1", p + ".bn1", cout, cin, 3, stride, 1, net, err)) return false;
      const int mid = push_conv(net, static_cast<int>(net->convs.size()) - 1, cur, -1, true);
      if (!add_conv_bn(t, p + ".conv2", p + ".bn2", cout, cout, 3, 1, 1, net, err)) return false;
      cur = push_conv(net, static_cast<int>(net->convs.size()) - 1, mid, identity, true);
      cin = cout;
    }
  }
  cur = push_pool(net, OpSpec::AVGPOOL, cur, 0, 0, 0, 1, 1);
  if (!add_conv_bias(t, "fc", 1000, 512, 0, 1, 0, 0, net, err)) return false;
  push_fc(net, static_cast<int>(net->convs.size()) - 1, cur, false, true);
  return true;
}

bool build_alexnet(const std::map<std::string, HostTensor>& t, int geometry, bool fused_stem, NetSpec* net,
                   std::string* err) {
  net->model = "alexnet";
  int in0 = -1;
  if (!fused_stem) {
    in0 = new_tensor(net, 224, 224, 4);
    OpSpec op;
    op.kind = OpSpec::NORMALIZE;
    op.out = in0;
    net->ops.push_back(op);
  }
  auto last = [&]() { return static_cast<int>(net->convs.size()) - 1; };
  auto pool = [&](int in) {
    const TensorShape s = net->tensors[in];
    return push_pool(net, OpSpec::MAXPOOL, in, 3, 2, 0, conv_out(s.H, 3, 2, 0), conv_out(s.W, 3, 2, 0));
  };
  if (!add_conv_bias(t, "features.0", 64, 3, 11, 4, 2, 0, net, err)) return false;
  int cur;
  if (fused_stem) {   // normalise + features.0 + ReLU in one kernel (stem_alex.cu); the weights are re-packed at load
    OpSpec op;
    op.kind = OpSpec::STEM_ALEX;
    op.conv = last();
    op.relu = true;
    op.out = new_tensor(net, 55, 55, 64);
    net->ops.push_back(op);
    net->flops_per_image += 2.0 * 55 * 55 * 64 * 3 * 121;
    cur = pool(op.out);
  } else {
    cur = pool(push_conv(net, last(), in0, -1, true));
  }
  // features.3: tch 0.4.1 builds it as (k5, pad 1, stride 2); torchvision as (k5, pad 2, stride 1).
  // Same weight shape either way.
  const int p3 = geometry == ALEX_TORCHVISION ? 2 : 1, s3 = geometry == ALEX_TORCHVISION ? 1 : 2;
  if (!add_conv_bias(t, "features.3", 192, 64, 5, s3, p3, 0, net, err)) return false;
  cur = pool(push_conv(net, last(), cur, -1, true));
  if (!add_conv_bias(t, "features.6", 384, 192, 3, 1, 1, 0, net, err)) return false;
  cur = push_conv(net, last(), cur, -1, true);
  if (!add_conv_bias(t, "features.8", 256, 384, 3, 1, 1, 0, net, err)) return false;
  cur = push_conv(net, last(), cur, -1, true);
  if (!add_conv_bias(t, "features.10", 256, 256, 3, 1, 1, 0, net, err)) return false;
  cur = pool(push_conv(net, last(), cur, -1, true));
  const TensorShape s = net->tensors[cur];
  if (s.H != 6 || s.W != 6) cur = push_pool(net, OpSpec::AVGPOOL, cur, 0, 0, 0, 6, 6);
  // classifier.1 consumes a (C,H,W) flatten in the reference; activations here are (H,W,C).
  if (!add_conv_bias(t, "classifier.1", 4096, 9216, 0, 1, 0, 256, net, err)) return false;
  cur = push_fc(net, last(), cur, true, false);
  if (!add_conv_bias(t, "classifier.4", 4096, 4096, 0, 1, 0, 0, net, err)) return false;
  cur = push_fc(net, last(), cur, true, false);
  if (!add_conv_bias(t, "classifier.6", 1000, 4096, 0, 1, 0, 0, net, err)) return false;
  push_fc(net, last(), cur, false, true);
  return true;
}

// ------------------------------------------------------------------------------- DeviceModel
#define B2I_CUDA(call)                                                                    \
  do {                                                                                    \
    cudaError_t e_ = (call);                                                              \
    if (e_ != cudaSuccess) {                                                              \
      *err = std::string(#call) + ": " + cudaGetErrorString(e_);                          \
      return false;                                                                       \
    }                                                                                     \
  } while (0)

DeviceModel::~DeviceModel() {
  if (device_ < 0) return;
  cudaSetDevice(device_);
  for (void* p : bufs_) cudaFree(p);
  for (void* p : owned_) cudaFree(p);
}

bool DeviceModel::init(int device, const NetSpec& net, const EngineConfig& cfg, std::string* err) {
  device_ = device;
  cfg_ = cfg;
  classes_ = net.classes;
  logits_tensor_ = net.logits_tensor;
  if (cfg_.max_batch < 1) { *err = "max_batch must be >= 1"; return false; }
  mb_ = cfg_.micro_batch > 0 ? std::min(cfg_.micro_batch, cfg_.max_batch) : cfg_.max_batch;
  esize_ = cfg_.precision == PREC_FP32 ? 4 : 2;
  B2I_CUDA(cudaSetDevice(device));
  cudaDeviceProp prop;
  B2I_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10) {
    *err = "device " + std::to_string(device) + " is sm_" + std::to_string(prop.major) +
           std::to_string(prop.minor) + "; this library carries sm_100a kernels only";
    return false;
  }
  tshape_ = net.tensors;

  // ---- activation buffers: greedy reuse by liveness (last reader), sizes per micro-batch
  const int T = static_cast<int>(tshape_.size());
  std::vector<int> last_use(T, -1);
  for (size_t i = 0; i < net.ops.size(); ++i) {
    const OpSpec& o = net.ops[i];
    if (o.in >= 0) last_use[o.in] = static_cast<int>(i);
    if (o.res >= 0) last_use[o.res] = static_cast<int>(i);
  }
  last_use[logits_tensor_] = 1 << 30;
  tensor_buf_.assign(T, -1);
  std::vector<int> free_list;
  auto tensor_bytes = [&](int id) {
    const size_t es = (id == logits_tensor_) ? 4 : esize_;
    return tshape_[id].elems() * static_cast<size_t>(mb_) * es;
  };
  for (size_t i = 0; i < net.ops.size(); ++i) {
    const OpSpec& o = net.ops[i];
    const size_t need = tensor_bytes(o.out);
    int pick = -1;
    if (!cfg_.keep_tensors && o.out != logits_tensor_) {
      for (size_t f = 0; f < free_list.size(); ++f) {
        const int b = free_list[f];
        if (buf_bytes_[b] >= need && (pick < 0 || buf_bytes_[b] < buf_bytes_[pick])) pick = b;
      }
      if (pick >= 0) free_list.erase(std::find(free_list.begin(), free_list.end(), pick));
    }
    if (pick < 0) {
      buf_bytes_.push_back(need);
      pick = static_cast<int>(buf_bytes_.size()) - 1;
    }
    tensor_buf_[o.out] = pick;
    // inputs whose last reader is this op become reusable for LATER ops (never for this op's output)
    for (int id : {o.in, o.res})
      if (id >= 0 && last_use[id] == static_cast<int>(i) && id != logits_tensor_) {
        const int b = tensor_buf_[id];
        if (std::find(free_list.begin(), free_list.end(), b) == free_list.end()) free_list.push_back(b);
      }
  }
  bufs_.assign(buf_bytes_.size(), nullptr);
  for (size_t b = 0; b < buf_bytes_.size(); ++b) {
    B2I_CUDA(cudaMalloc(&bufs_[b], buf_bytes_[b]));
    B2I_CUDA(cudaMemset(bufs_[b], 0, buf_bytes_[b]));
    act_bytes_ += buf_bytes_[b];
  }

  // ---- weights + per-op launch plan
  const bool f16 = cfg_.precision == PREC_FP16;   // 16-bit tensors are IEEE half instead of bf16
  for (const OpSpec& o : net.ops) {
    DevOp d;
    d.spec = o;
    if (o.kind == OpSpec::STEM) {
      const ConvSpec& c = net.convs[o.conv];
      if (cfg_.precision == PREC_FP32 || c.K != 256 || c.Cout != 64) { *err = "fused stem needs the bf16 path and conv1 as [64][256]"; return false; }
      std::vector<uint16_t> wb(c.w.size());
      for (size_t i = 0; i < wb.size(); ++i) wb[i] = host_f32_to_16(c.w[i], f16);
      B2I_CUDA(cudaMalloc(&d.w, wb.size() * 2));
      B2I_CUDA(cudaMemcpy(d.w, wb.data(), wb.size() * 2, cudaMemcpyHostToDevice));
      owned_.push_back(d.w);
      weight_bytes_ += wb.size() * 2;
      {
        std::vector<uint32_t> w2(128 * 144);
        stem2_pack_weights(wb.data(), w2.data());
        void* p2 = nullptr;
        B2I_CUDA(cudaMalloc(&p2, w2.size() * 4));
        B2I_CUDA(cudaMemcpy(p2, w2.data(), w2.size() * 4, cudaMemcpyHostToDevice));
        owned_.push_back(p2);
        d.w2 = static_cast<uint32_t*>(p2);
        weight_bytes_ += w2.size() * 4;
      }
      void* bp = nullptr;
      B2I_CUDA(cudaMalloc(&bp, c.bias.size() * 4));
      B2I_CUDA(cudaMemcpy(bp, c.bias.data(), c.bias.size() * 4, cudaMemcpyHostToDevice));
      owned_.push_back(bp);
      d.bias = static_cast<float*>(bp);
      d.bias_host.assign(c.bias.begin(), c.bias.begin() + 64);
      const char* terr = "";
      if (make_tmap_weights(&d.tmB, d.w, 64, 256, 64, &terr)) { *err = std::string("stem: ") + terr; return false; }
      d.has_tmB = true;
      void* sp = nullptr;                      // the stem kernel's tile counter (two words, zero between launches)
      B2I_CUDA(cudaMalloc(&sp, 256));
      B2I_CUDA(cudaMemset(sp, 0, 256));
      owned_.push_back(sp);
      stem_sched_ = static_cast<unsigned int*>(sp);
    }
    if (o.kind == OpSpec::STEM_ALEX) {
      const ConvSpec& c = net.convs[o.conv];
      if (cfg_.precision == PREC_FP32 || c.R != 11 || c.S != 11 || c.Cin_pad != 4 || c.Cout != 64 || c.stride != 4 || c.pad != 2) {
        *err = "fused AlexNet stem needs the bf16 path and features.0 as 11x11, 3 -> 64";
        return false;
      }
      // [64][(r, s < 12, c < 4)] -> [64][704]: filter row r, plane (R,G) / (B,0), 16 taps (11 real), 2 channels
      std::vector<uint16_t> wb(static_cast<size_t>(64) * 704, 0);
      for (int oc = 0; oc < 64; ++oc)
        for (int r = 0; r < 11; ++r)
          for (int s = 0; s < 11; ++s)
            for (int ch = 0; ch < 3; ++ch) {
              const float v = c.w[static_cast<size_t>(oc) * c.K + (r * c.Sp + s) * c.Cin_pad + ch];
              wb[static_cast<size_t>(oc) * 704 + r * 64 + (ch >> 1) * 32 + s * 2 + (ch & 1)] = host_f32_to_16(v, f16);
            }
      B2I_CUDA(cudaMalloc(&d.w, wb.size() * 2));
      B2I_CUDA(cudaMemcpy(d.w, wb.data(), wb.size() * 2, cudaMemcpyHostToDevice));
      owned_.push_back(d.w);
      weight_bytes_ += wb.size() * 2;
      d.bias_host.assign(c.bias.begin(), c.bias.begin() + 64);
      const char* terr = "";
      if (make_tmap_weights(&d.tmB, d.w, 64, 704, 64, &terr)) { *err = std::string("alexnet stem: ") + terr; return false; }
      d.has_tmB = true;
    }
    if (o.kind == OpSpec::CONV) {
      const ConvSpec& c = net.convs[o.conv];
      const TensorShape ti = tshape_[o.in], to = tshape_[o.out];
      ConvGeom& g = d.g;
      if (o.fc) {
        g.H = g.W = 1; g.C = c.K; g.OH = g.OW = 1; g.R = g.S = g.Sp = 1; g.stride = 1; g.pad = 0;
        if (static_cast<int>(ti.elems()) != c.K) {
          *err = c.name + ": FC input has " + std::to_string(ti.elems()) + " features, weights expect " +
                 std::to_string(c.K);
          return false;
        }
      } else {
        g.H = ti.H; g.W = ti.W; g.C = c.Cin_pad; g.OH = to.H; g.OW = to.W;
        g.R = c.R; g.S = c.S; g.Sp = c.Sp; g.stride = c.stride; g.pad = c.pad;
        if (ti.C != c.Cin_pad) {
          *err = c.name + ": input tensor has " + std::to_string(ti.C) + " channels, layer expects " +
                 std::to_string(c.Cin_pad);
          return false;
        }
      }
      g.N = mb_; g.Cout = c.Cout; g.Cout_pad = c.Cout_pad; g.K = c.K; g.M = 0;
      // tile-N: largest multiple of 32 <= bn_max dividing Cout_pad
      int bn = std::min(cfg_.bn_max, 256) / 32 * 32;
      // FC layers have few rows (one micro-batch): 256-wide tiles would occupy a handful of SMs and run their
      // whole K loop in sequence; 64-wide tiles spread the same work over 4x the CTAs (ResNet fc 19 -> 12 us,
      // AlexNet classifier 132 -> 89 us at batch 256).  B2I_FC_BN overrides.
      static const int fc_bn = [] { const char* e = getenv("B2I_FC_BN"); return e ? atoi(e) : 0; }();
      if (o.fc) {
        const int tiles256 = ((mb_ + 127) / 128) * ((c.Cout_pad + 255) / 256);
        if (fc_bn >= 32) bn = std::min(bn, fc_bn / 32 * 32);
        else if (tiles256 < 148) bn = std::min(bn, 64);
      }
      while (bn > 32 && c.Cout_pad % bn) bn -= 32;
      // FC layers of the 16-bit path stream a weight matrix that is read once per batch.  Split-K gives every SM a
      // share when the layer has fewer tiles than SMs (64-wide tiles stay the best width at every batch; an A box
      // shorter than 128 rows and weights re-laid as contiguous tiles were tried and change nothing:
      // profiles/r03_fc_plan_matrix.txt).  The factor is fixed per layer at load time: an image's logits do not
      // depend on the size of the batch it arrives in.  B2I_NO_SPLITK=1 turns it off.
      static const bool fc_plan_on = [] { const char* e = getenv("B2I_NO_SPLITK"); return !(e && e[0] == '1'); }();
      bool fc_planned = false;
      // B2I_FC_SPLITS / B2I_FC_AROWS (with B2I_FC_BN) force a plan (A/B timing)
      static const int fc_splits_env = [] { const char* e = getenv("B2I_FC_SPLITS"); return e ? atoi(e) : 0; }();
      static const int fc_arows_env = [] { const char* e = getenv("B2I_FC_AROWS"); return e ? atoi(e) : 0; }();
      if (o.fc && cfg_.precision != PREC_FP32 && cfg_.persistent && fc_bn >= 32 && (fc_splits_env > 0 || fc_arows_env > 0)) {
        if (fc_splits_env > 0) d.splits = std::min(fc_splits_env, std::max(1, (c.K / 64) / 2));
        if (fc_arows_env >= 8 && (mb_ + 127) / 128 == 1) d.a_rows = std::min(128, std::max((mb_ + 7) / 8 * 8, fc_arows_env / 8 * 8));
        fc_planned = true;
      }
      if (o.fc && fc_bn < 32 && cfg_.precision != PREC_FP32 && fc_plan_on && cfg_.persistent) {
        // measured (profiles/r03_fc_plan_matrix.txt): 64-wide tiles win at every batch; splitting K pays while the
        // layer has fewer (tile, split) items than SMs and the partial sums stay small (one M tile)
        const int tiles_mn = ((mb_ + 127) / 128) * (c.Cout_pad / bn);
        int sp = std::max(1, std::min(8, 128 / std::max(1, tiles_mn)));
        sp = std::min(sp, std::max(1, (c.K / 64) / 16));   // at least 16 k-blocks per split (ResNet's 8-k-block FC is not split)
        while (sp > 1 && (sp - 1) * ((c.K / 64 + sp - 1) / sp) >= c.K / 64) --sp;
        d.splits = sp;
        fc_planned = true;
      }
      d.BN = bn;
      d.stages = conv_umma_pick_stages(bn);
      d.b_tma = cfg_.b_tma;
      if (o.fc) d.amode = A_TILED;
      else if (c.smallc) d.amode = A_SMALLC;
      else d.amode = (cfg_.conv_amode == A_GATHER) ? A_GATHER : A_IM2COL;
      if (d.amode == A_IM2COL || d.amode == A_TILED) d.b_tma = 1;
      d.patch = (cfg_.patch_conv && cfg_.precision != PREC_FP32 && !o.fc && d.amode == A_IM2COL && conv_patch_applicable(g))
                    ? 1 : 0;
      if (d.patch) d.BN = 64;
      d.pair = (cfg_.pair_mma && cfg_.precision != PREC_FP32 && !d.patch && d.BN >= cfg_.pair_mma && !fc_planned &&
                (d.amode == A_IM2COL || d.amode == A_TILED)) ? 1 : 0;
      // One N-tile whose weights fit in shared memory next to four A stages (ResNet layer2): CTA pairs with
      // resident weights, whatever pair_mma says about the tile width.  B2I_NO_RESB=1 turns this off.
      static const bool resb_on = [] { const char* e = getenv("B2I_NO_RESB"); return !(e && e[0] == '1'); }();
      if (resb_on && cfg_.pair_mma && cfg_.precision != PREC_FP32 && !d.patch && !o.fc && d.amode == A_IM2COL &&
          d.BN == c.Cout_pad && d.BN >= 64 &&
          static_cast<size_t>(c.K / 64 + c.Cin_pad / 64) * (d.BN / 2) * 128 + 4 * 16384 + 1024 <= 226 * 1024) {
        d.pair = 1;
        d.resident_b = 1;
      }
      // 3x3 / s1 convs of that kind go one step further: flattened padded rows, each pixel fetched from L2
      // 1.5 times instead of nine.  B2I_NO_FLAT=1 turns this off.
      static const bool flat_on = [] { const char* e = getenv("B2I_NO_FLAT"); return !(e && e[0] == '1'); }();
      if (flat_on && cfg_.pair_mma && cfg_.precision != PREC_FP32 && !d.patch && !o.fc && d.amode == A_IM2COL &&
          conv_flat_applicable(g, d.BN)) {
        d.flat = 1;
        d.pair = 1;
      }

      // 3x3 / s1 convs whose weights do not fit in shared memory (layer4, layer3): the same flattened rows with the weights
      // streamed through a ring (conv3x3_flat_stream_kernel).  B2I_FLATS_MAXW = widest map that takes this path
      // (default 0 = off: measured slower than the im2col form on layer4 and layer3, profiles/r03_layer4_tiling_matrix.txt; 7 = layer4, 14 = layer3 too).
      static const int flats_maxw = [] { const char* e = getenv("B2I_FLATS_MAXW"); return e ? atoi(e) : 0; }();
      if (flat_on && !d.flat && cfg_.pair_mma && cfg_.precision != PREC_FP32 && !d.patch && !o.fc && d.amode == A_IM2COL &&
          g.W <= flats_maxw && d.BN >= 128 && conv_flat_stream_applicable(g, d.BN)) {
        d.flat = 2;
        d.pair = 1;
        d.resident_b = 0;
      }

      // CTA-pair im2col layers: an im2col TMA request costs ~700 cycles whatever its size (tools/microbench/
      // tma_rate.cu), so a CTA that owns `msub` 128-row sub-tiles (one request of msub * 128 pixels per k-block)
      // feeds msub times the MMAs per request.  It also divides the number of tiles, and when msub * BN exceeds 256
      // TMEM columns only one accumulator set fits (the epilogue no longer overlaps the next tile).  Pick tile-N and
      // msub by an estimate of the launch time at the full micro-batch: waves x (k-blocks x max(request, MMAs) +
      // exposed epilogue + per-tile set-up).  At batch 256 it keeps (256, 1) for layer3 and layer4 and takes (128, 2)
      // for layer2.0.conv1, which is what an A/B over (msub, producers) measured (profiles/r03_msub_nprod_matrix.txt).  msub is 1, 2 or 4: a 384-pixel request faults on B200 (measured).
      // B2I_MSUB=1 / 2 / 4 forces msub (tile-N unchanged).
      d.msub = 1;
      if (d.pair && !d.flat && !o.fc && d.amode == A_IM2COL && g.R > 1) {
        static const int msub_env = [] { const char* e = getenv("B2I_MSUB"); return e ? atoi(e) : 0; }();
        const long long M = static_cast<long long>(mb_) * g.OH * g.OW;
        const int kb = c.K / 64;
        // a 3x3 / s2 conv will carry its block's 1x1 / s2 downsample as extra tiles of Cin/64 k-blocks each
        const bool carries = cfg_.fuse_downsample && g.R == 3 && g.stride == 2;
        auto fits = [&](int bn, int ms) {
          return ms * bn <= 512 && 3 * (ms * 16384 + (d.resident_b ? 0 : bn / 2 * 128)) <= 210 * 1024;
        };
        auto est = [&](int bn, int ms) {
          const long long units = ((M + 256 * ms - 1) / (256 * ms)) * (c.Cout_pad / bn) * (carries ? 2 : 1);
          const long long waves = (units + 73) / 74;
          // requests of different issuing warps overlap (tools/microbench/tma_rate.cu, two-issuer rows): with the pair
          // kernel's three producer warps one completes every req / 3 cycles
          const long long req = (ms <= 2 ? 700 : (ms == 3 ? 730 : 764)) / 3;
          // what a k-block costs: the request turn-around, its MMAs, or the time the SM needs to take in its
          // operands (this CTA's A rows and its half of the weight tile at ~40 B/clk: the im2col layers of layer3 /
          // layer4 run at that fill rate, profiles/r03_ncu_step_full.csv)
          const long long fill = (static_cast<long long>(ms) * 16384 + (d.resident_b ? 0 : bn / 2 * 128)) / 40;
          const long long kb_cycles = std::max<long long>(std::max<long long>(req, fill), static_cast<long long>(ms) * 4 * (bn / 2));
          // one accumulator set only: the MMAs of the next tile wait for the whole epilogue and the operand ring runs
          // dry meanwhile.  30000 reproduces layer4 at (256, 2): 118k cycles measured, against 111k at (256, 1)
          // (profiles/r03_layer4_tiling_matrix.txt)
          const long long epi = (2 * ms * bn > 512) ? 30000 : 0;
          const long long kb_avg = carries ? (kb + c.Cin_pad / 64) / 2 : kb;
          return waves * (kb_avg * kb_cycles + epi + 600);   // + per-tile set-up (barrier turn-around, bias / residual fetch)
        };
        if (msub_env == 1 || msub_env == 2 || msub_env == 4) {
          if (fits(d.BN, msub_env)) d.msub = msub_env;
        } else {
          long long best = est(d.BN, 1);
          const int bn0 = d.BN;
          for (int bn : {bn0, 128}) {
            // the downsample riding on a stride-2 conv shares its tile-N: leave that one alone
            if (bn > bn0 || c.Cout_pad % bn || bn < cfg_.pair_mma / 2 || ((d.resident_b || carries) && bn != bn0)) continue;
            for (int ms = 1; ms <= 4; ms *= 2) {   // 384-pixel requests fault: the pixel count must be a power of two
              if (!fits(bn, ms)) continue;
              const long long t = est(bn, ms);
              if (t < best - best / 50) { best = t; d.BN = bn; d.msub = ms; }   // needs a 2 % edge over the simpler choice
            }
          }
        }
      }

      if (getenv("B2I_PLAN_DEBUG")) fprintf(stderr, "[plan] %s BN %d pair %d msub %d resb %d flat %d patch %d splits %d a_rows %d\n", c.name.c_str(), d.BN, d.pair, d.msub, d.resident_b, d.flat, d.patch, d.splits, d.a_rows);
      const size_t wn = c.w.size();
      if (cfg_.precision == PREC_FP32) {
        B2I_CUDA(cudaMalloc(&d.w, wn * 4));
        B2I_CUDA(cudaMemcpy(d.w, c.w.data(), wn * 4, cudaMemcpyHostToDevice));
        weight_bytes_ += wn * 4;
      } else {
        std::vector<uint16_t> wb(wn);
        for (size_t i = 0; i < wn; ++i) wb[i] = host_f32_to_16(c.w[i], f16);
        B2I_CUDA(cudaMalloc(&d.w, wn * 2));
        B2I_CUDA(cudaMemcpy(d.w, wb.data(), wn * 2, cudaMemcpyHostToDevice));
        weight_bytes_ += wn * 2;
      }
      owned_.push_back(d.w);
      void* bp = nullptr;
      B2I_CUDA(cudaMalloc(&bp, c.bias.size() * 4));
      B2I_CUDA(cudaMemcpy(bp, c.bias.data(), c.bias.size() * 4, cudaMemcpyHostToDevice));
      owned_.push_back(bp);
      d.bias = static_cast<float*>(bp);

      if (cfg_.precision != PREC_FP32) {
        const char* terr = "";
        if (d.b_tma) {
          if (make_tmap_weights(&d.tmB, d.w, c.Cout_pad, c.K, d.pair ? d.BN / 2 : d.BN, &terr)) { *err = c.name + ": " + terr; return false; }
          d.has_tmB = true;
        }
        void* xin = bufs_[tensor_buf_[o.in]];
        if (d.patch) {
          if (make_tmap_patch(&d.tmA, xin, mb_, g.H, g.W, &terr)) { *err = c.name + ": " + terr; return false; }
          if (make_tmap_tile(&d.tmO, bufs_[tensor_buf_[o.out]], mb_, g.H, g.W, &terr)) { *err = c.name + ": " + terr; return false; }
          if (o.res >= 0 && make_tmap_tile(&d.tmR, bufs_[tensor_buf_[o.res]], mb_, g.H, g.W, &terr)) { *err = c.name + ": " + terr; return false; }
          d.has_tmA = true;
        } else if (d.flat) {
          if (make_tmap_flat_pitch(&d.tmA, xin, mb_, g.H, g.W, g.C, d.flat == 2 ? conv_flat_stream_pitch(g.W) : conv_flat_pitch(g.W), &terr)) { *err = c.name + ": " + terr; return false; }
          d.has_tmA = true;
        } else if (d.amode == A_IM2COL) {
          if (make_tmap_im2col_px(&d.tmA, xin, g, mb_, 128 * d.msub, &terr)) { *err = c.name + ": " + terr; return false; }
          d.has_tmA = true;
        } else if (d.amode == A_TILED) {
          if (make_tmap_rows(&d.tmA, xin, mb_, c.K, d.pair ? 128 : d.a_rows, &terr)) { *err = c.name + ": " + terr; return false; }
          d.has_tmA = true;
        }
      }
    }
    ops_.push_back(d);
    switch (o.kind) {
      case OpSpec::NORMALIZE: op_names_.push_back("normalize"); break;
      case OpSpec::STEM: op_names_.push_back("conv:stem(normalize+conv1+maxpool)"); break;
      case OpSpec::STEM_ALEX: op_names_.push_back("conv:stem(normalize+features.0)"); break;
      case OpSpec::MAXPOOL: op_names_.push_back("maxpool"); break;
      case OpSpec::AVGPOOL: op_names_.push_back("avgpool"); break;
      case OpSpec::CONV: op_names_.push_back((o.fc ? "fc:" : "conv:") + net.convs[o.conv].name); break;
    }
  }
  op_names_.push_back("softmax_topk");
  // Split-K scratch for the FC layers of the 16-bit path (few rows, many weights: the k-blocks of one output tile
  // are spread over several CTAs so that every SM streams a share of the weight matrix).  B2I_NO_SPLITK=1: off.
  // Split-K scratch for the FC layers (per-split partial sums + one arrival counter per output tile).
  {
    size_t need = 0;
    for (const DevOp& d : ops_)
      if (d.spec.kind == OpSpec::CONV && d.spec.fc && d.splits > 1)
        need = std::max(need, static_cast<size_t>(d.splits) * ((static_cast<size_t>(mb_) + 127) / 128) * 128 * d.g.Cout_pad * 4);
    if (need) {
      void* w = nullptr;
      void* c = nullptr;
      B2I_CUDA(cudaMalloc(&w, need));
      B2I_CUDA(cudaMalloc(&c, 4096 * 4));
      B2I_CUDA(cudaMemset(c, 0, 4096 * 4));
      owned_.push_back(w);
      owned_.push_back(c);
      splitk_ws_ = static_cast<float*>(w);
      splitk_ctr_ = static_cast<int*>(c);
      splitk_ws_bytes_ = need;
    }
  }
  // A 1x1 / stride-s / pad-0 conv followed by a 3x3 / stride-s / pad-1 conv on the same input (the
  // ResNet downsample and the block's first conv) read the same pixels: the 1x1 conv's input is the
  // centre tap of the 3x3 window.  Compute the former as extra N-tiles of the latter's kernel.
  if (cfg_.fuse_downsample && cfg_.precision != PREC_FP32) {
    for (size_t i = 0; i + 1 < ops_.size(); ++i) {
      DevOp& a = ops_[i];
      DevOp& b = ops_[i + 1];
      if (a.spec.kind != OpSpec::CONV || b.spec.kind != OpSpec::CONV || a.spec.fc || b.spec.fc) continue;
      if (a.spec.in != b.spec.in || a.spec.res >= 0 || a.spec.relu) continue;
      if (a.g.R != 1 || a.g.S != 1 || a.g.pad != 0 || b.g.R != 3 || b.g.S != 3 || b.g.pad != 1) continue;
      if (a.g.stride != b.g.stride || a.g.OH != b.g.OH || a.g.OW != b.g.OW || a.g.Cout != b.g.Cout) continue;
      if (a.amode != A_IM2COL || b.amode != A_IM2COL || a.patch || b.patch || a.BN != b.BN || a.pair != b.pair) continue;
      if (!a.has_tmB || (a.g.Cout & 31)) continue;
      b.fused_second = static_cast<int>(i);
      a.skip = true;
      op_names_[i + 1] += "+downsample";
    }
  }
  return true;
}

bool DeviceModel::run_microbatch(const uint8_t* d_u8, int nb, int k, int32_t* d_idx, float* d_prob,
                                 float* d_logits, cudaStream_t st, std::string* err) {
  const int dt = cfg_.precision == PREC_FP32 ? DT_F32 : (cfg_.precision == PREC_FP16 ? DT_F16 : DT_BF16);
  const int f16 = cfg_.precision == PREC_FP16 ? 1 : 0;
  auto buf = [&](int tensor) { return tensor >= 0 ? bufs_[tensor_buf_[tensor]] : nullptr; };
  size_t ev = 0;
  auto mark = [&]() {
    if (prof_events_ && ev < prof_events_->size()) cudaEventRecord((*prof_events_)[ev++], st);
  };
  mark();
  int patch_seq = 0;   // the stem writes ascending, so the first halo-patch conv walks descending, the next ascending, ...
  for (DevOp& d : ops_) {
    const OpSpec& o = d.spec;
    if (d.skip) { mark(); continue; }
    switch (o.kind) {
      case OpSpec::STEM: {
        // B2I_STEM=1: the first form of the kernel (pixels on the accumulator lanes, stem.cu); default: stem2.cu
        static const int stem_form = [] { const char* e = getenv("B2I_STEM"); return e ? atoi(e) : 2; }();
        if (stem_form != 1) B2I_CUDA(launch_stem2(d_u8, buf(o.out), d.w2, d.bias_host.data(), nb, stem_sched_, f16, st));
        else B2I_CUDA(launch_stem(d_u8, buf(o.out), &d.tmB, d.bias_host.data(), nb, stem_sched_, f16, st));
        break;
      }
      case OpSpec::STEM_ALEX:
        B2I_CUDA(launch_alex_stem(d_u8, buf(o.out), &d.tmB, d.bias_host.data(), nb, f16, st));
        break;
      case OpSpec::NORMALIZE:
        B2I_CUDA(launch_normalize_u8(d_u8, buf(o.out), dt, static_cast<long long>(nb) * 224 * 224, st));
        break;
      case OpSpec::MAXPOOL: {
        const TensorShape ti = tshape_[o.in], to = tshape_[o.out];
        B2I_CUDA(launch_maxpool(buf(o.in), buf(o.out), dt, nb, ti.H, ti.W, ti.C, o.k, o.stride,
                                o.pad, to.H, to.W, st));
        break;
      }
      case OpSpec::AVGPOOL: {
        const TensorShape ti = tshape_[o.in], to = tshape_[o.out];
        B2I_CUDA(launch_adaptive_avgpool(buf(o.in), buf(o.out), dt, nb, ti.H, ti.W, ti.C, to.H,
                                         to.W, st));
        break;
      }
      case OpSpec::CONV: {
        ConvArgs a;
        a.g = d.g;
        a.g.N = nb;
        a.g.M = nb * d.g.OH * d.g.OW;
        a.x = buf(o.in);
        a.w = d.w;
        a.bias = d.bias;
        a.residual = buf(o.res);
        a.out = buf(o.out);
        a.ldo = o.logits ? classes_ : d.g.Cout;
        a.relu = o.relu ? 1 : 0;
        a.out_fp32 = o.logits ? 1 : 0;
        a.BN = d.BN;
        a.stages = d.stages;
        a.amode = d.amode;
        a.b_tma = d.b_tma;
        a.persistent = cfg_.persistent;
        a.patch = d.patch;
        a.tmO = &d.tmO;
        a.tmR = o.res >= 0 ? &d.tmR : nullptr;
        a.pair = d.pair;
        a.resident_b = d.resident_b;
        a.flat = d.flat;
        a.msub = d.msub;
        // Snake order through the conv layers: a kernel starts with the tiles its producer wrote last, which are the
        // ones still in L2 (a layer1 tensor is 103 MB at batch 256, L2 holds 126 MB).  B2I_NO_SNAKE=1: off.
        static const bool snake = [] { const char* e = getenv("B2I_NO_SNAKE"); return !(e && e[0] == '1'); }();
        a.reverse = (snake && (d.patch || d.flat || d.pair) && !o.fc) ? (patch_seq++ & 1) ^ 1 : 0;
        a.f16 = f16;
        if (o.fc && !d.pair && d.amode == A_TILED) {
          a.a_rows = d.a_rows;
          if (d.splits > 1 && splitk_ws_) {
            a.splits = d.splits;
            a.ws = splitk_ws_;
            a.ctr = splitk_ctr_;
            a.ldw = d.g.Cout_pad;
          }
        }
        if (d.fused_second >= 0) {
          const DevOp& s2 = ops_[d.fused_second];
          const int cpb = d.g.C >> 6;
          a.tmB2 = &s2.tmB;
          a.bias2 = s2.bias;
          a.out2 = buf(s2.spec.out);
          a.ldo2 = s2.g.Cout;
          a.relu2 = 0;
          a.kb_first2 = (1 * d.g.S + 1) * cpb;   // centre tap (r, s) = (1, 1)
          a.num_kb2 = cpb;
        }
        if (cfg_.precision == PREC_FP32) {
          B2I_CUDA(launch_conv_fp32(a, st));
        } else {
          B2I_CUDA(launch_conv_umma(a, d.has_tmA ? &d.tmA : nullptr, d.has_tmB ? &d.tmB : nullptr, st));
        }
        break;
      }
    }
    mark();
  }
  const float* lg = static_cast<const float*>(buf(logits_tensor_));
  B2I_CUDA(launch_softmax_topk(lg, nb, classes_, classes_, k, d_idx, d_prob, st));
  mark();
  if (d_logits)
    B2I_CUDA(cudaMemcpyAsync(d_logits, lg, static_cast<size_t>(nb) * classes_ * 4,
                             cudaMemcpyDeviceToDevice, st));
  return true;
}

bool DeviceModel::forward(const uint8_t* d_u8, int n, int k, int32_t* d_idx, float* d_prob,
                          float* d_logits, cudaStream_t st, std::string* err) {
  if (n < 0 || n > cfg_.max_batch) {
    *err = "batch " + std::to_string(n) + " exceeds the model's max_batch " + std::to_string(cfg_.max_batch);
    return false;
  }
  if (k < 1 || k > 8) { *err = "k must be in 1..8"; return false; }
  B2I_CUDA(cudaSetDevice(device_));
  for (int off = 0; off < n; off += mb_) {
    const int nb = std::min(mb_, n - off);
    if (!run_microbatch(d_u8 + static_cast<size_t>(off) * 224 * 224 * 3, nb, k,
                        d_idx + static_cast<size_t>(off) * k, d_prob + static_cast<size_t>(off) * k,
                        d_logits ? d_logits + static_cast<size_t>(off) * classes_ : nullptr, st, err))
      return false;
  }
  return true;
}

bool DeviceModel::profile(const uint8_t* d_u8, int n, int k, int32_t* d_idx, float* d_prob,
                          cudaStream_t st, std::vector<float>* ms, std::vector<std::string>* labels,
                          std::string* err) {
  if (n < 1 || n > mb_) { *err = "profile: batch must be within one micro-batch"; return false; }
  B2I_CUDA(cudaSetDevice(device_));
  std::vector<cudaEvent_t> evs(ops_.size() + 2);
  for (auto& e : evs) B2I_CUDA(cudaEventCreate(&e));
  prof_events_ = &evs;
  const bool ok = run_microbatch(d_u8, n, k, d_idx, d_prob, nullptr, st, err);
  prof_events_ = nullptr;
  if (ok) {
    cudaError_t e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) { *err = std::string("profile sync: ") + cudaGetErrorString(e); }
    ms->clear();
    labels->clear();
    for (size_t i = 0; i + 1 < evs.size(); ++i) {
      if (i < ops_.size() && ops_[i].skip) continue;   // computed inside another op's kernel: no launch of its own
      float t = 0.f;
      cudaEventElapsedTime(&t, evs[i], evs[i + 1]);
      ms->push_back(t);
      labels->push_back(op_names_[i]);
    }
  }
  for (auto& e : evs) cudaEventDestroy(e);
  return ok;
}

bool DeviceModel::read_tensor(int id, int n, std::vector<float>* out, std::string* err) {
  if (id < 0 || id >= static_cast<int>(tshape_.size()) || n < 1 || n > mb_) {
    *err = "read_tensor: bad tensor id or batch";
    return false;
  }
  B2I_CUDA(cudaSetDevice(device_));
  B2I_CUDA(cudaDeviceSynchronize());
  const size_t ne = tshape_[id].elems() * n;
  out->resize(ne);
  const void* src = bufs_[tensor_buf_[id]];
  if (id == logits_tensor_ || cfg_.precision == PREC_FP32) {
    B2I_CUDA(cudaMemcpy(out->data(), src, ne * 4, cudaMemcpyDeviceToHost));
  } else {
    std::vector<uint16_t> tmp(ne);
    B2I_CUDA(cudaMemcpy(tmp.data(), src, ne * 2, cudaMemcpyDeviceToHost));
    for (size_t i = 0; i < ne; ++i) (*out)[i] = host_16_to_f32(tmp[i], cfg_.precision == PREC_FP16);
  }
  return true;
}

}  // namespace b2i
