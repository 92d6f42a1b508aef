/* b200_infer.h — C ABI of the B200-native inference path.
 *
 * Drop-in boundary for the five tch calls the reference makes under `Member::predict`
 * (reference src/services.rs:492-494) and for the model construction + weight load in
 * `MemberState::shared` (reference src/services.rs:513-524; dead twin src/model_utils.rs:24-49).
 * Plain pointers and sizes only; every buffer is caller-owned; no exceptions cross this line.
 * All entry points return 0 on success and a negative B2I_E* code on failure, with a
 * thread-local message available from b2i_last_error().
 *
 * There is NO CPU fallback: every compute entry point needs an sm_100 GPU and fails with
 * B2I_ENODEV / B2I_ECUDA otherwise.
 */
#ifndef B200_INFER_H_
#define B200_INFER_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2I_OK 0
#define B2I_EINVAL (-1)  /* bad argument (unknown model name, k out of range, null pointer ...) */
#define B2I_EIO (-2)     /* weight archive missing / malformed (the reference panics here: services.rs:516,522) */
#define B2I_ECUDA (-3)   /* CUDA runtime / driver error */
#define B2I_ENODEV (-4)  /* no usable sm_100 device */
#define B2I_ENOMEM (-5)
#define B2I_EINTERNAL (-6) /* a C++ exception was stopped at this boundary; the message says which */
#define B2I_EDECODE (-7)   /* per-image status only: this image could not be decoded / is empty (the reference's
                              `.ok()?` at services.rs:492); the other images of the batch were still answered */

/* Precision modes. The bounds are the ones tests/test_model_gpu.py enforces against the CPU restatement of
 * the reference path (oracle/, fp32); that restatement is itself not pinned by any reference-held vector
 * ("parity unpinned"). `unit` = max(1, standard deviation of the logits). */
#define B2I_PREC_BF16 0 /* tcgen05, bf16 storage, fp32 accumulate: logits within 2.5e-2 x unit of the same network
                          evaluated with bf16 storage on the CPU and within 4e-2 x unit of the fp32 restatement
                          (8 mantissa bits per stored activation: a FLAT 2e-2 holds only for logits of unit
                          scale); top-k identical wherever neighbouring logits differ by more than that */
#define B2I_PREC_FP32 1 /* FFMA fp32: logits within 1e-4 x unit of the fp32 restatement */
#define B2I_PREC_FP16 2 /* the tcgen05 path with IEEE-half storage (fp32 accumulate), same speed as bf16, three more
                          mantissa bits: logits within 2e-2 x unit of the fp32 restatement, and within a flat
                          2e-2 on logits of pretrained scale (DESIGN.md section 2 has the measured gaps) */

#define B2I_ALEXNET_TCH 0         /* features.3 = k5 pad1 stride2, as tch 0.4.1 builds it */
#define B2I_ALEXNET_TORCHVISION 1 /* features.3 = k5 pad2 stride1 */

#define B2I_IMAGE_BYTES (224 * 224 * 3)
#define B2I_CLASSES 1000

typedef struct b2i_ctx b2i_ctx;     /* a set of GPUs; replaces tch::Device::Cpu (services.rs:514,520) */
typedef struct b2i_model b2i_model; /* replaces Mutex<Box<dyn ModuleT>> (services.rs:455-456) */

typedef struct b2i_model_config {
  int precision;        /* B2I_PREC_* */
  int max_batch;        /* largest n one predict call may pass PER GPU (buffers are sized for it) */
  int micro_batch;      /* images per kernel sequence, 0 = max_batch */
  int alexnet_geometry; /* B2I_ALEXNET_* */
  int conv_mode;        /* bf16 path, C % 64 == 0 layers: -1 auto (= 1); 0 cp.async gather, one tile per CTA;
                           1 TMA im2col, persistent kernel; 2 TMA im2col, one tile per CTA
                           (b2i_op_conv2d only: 3 = halo-patch kernel, 3x3/s1/p1 64->64) */
  int weights_tma;      /* 1: weight tiles through TMA (default), 0: cp.async (gather modes only) */
  int tile_n_max;       /* largest output-channel tile, multiple of 32 in [32,256]; 0 = default */
  int keep_tensors;     /* 1: no activation-buffer reuse, so b2i_model_read_tensor sees every layer */
  int fused_stem;       /* ResNet-18 bf16: 1 (default) = normalise + conv1 + max-pool in one kernel, 0 = three */
  int patch_conv;       /* 3x3/s1 64->64 layers (ResNet layer1): 1 (default) = halo-patch kernel, 0 = im2col TMA */
  int pair_mma;         /* 0 = off; n = layers whose tile-N is >= n use the CTA-pair (cta_group::2) kernel; default 256 */
  int fuse_downsample;  /* 1 (default): a 1x1/s2 downsample conv is computed inside its sibling 3x3/s2 conv's kernel */
} b2i_model_config;

/* Thread-local description of the last failure on the calling thread. Never NULL. */
const char* b2i_last_error(void);

/* Number of CUDA devices visible (0 when there is no driver). */
int b2i_device_count(void);

/* Context over `n_devices` GPU ordinals. Replaces `VarStore::new(Device::Cpu)` (services.rs:514). */
int b2i_init(const int* device_ids, int n_devices, b2i_ctx** out);
void b2i_shutdown(b2i_ctx* ctx);

/* Fills cfg with the defaults (bf16, max_batch 256, tch AlexNet geometry, auto conv mode). */
void b2i_default_config(b2i_model_config* cfg);

/* Parse a `.ot` archive and report its tensors without touching a GPU (CPU-only; used by tests
 * and by callers that want to validate a file first). names_out receives '\n'-separated
 * "name:d0xd1x..." lines, truncated to cap bytes. Returns the tensor count or a negative code. */
int b2i_ot_inspect(const char* path, char* names_out, size_t cap);
/* Copy one tensor of the archive as fp32 (CPU-only). Returns numel or a negative code. */
long long b2i_ot_read_tensor(const char* path, const char* name, float* out, long long cap_elems);

/* model_name is "resnet18" or "alexnet" (anything else -> B2I_EINVAL, the reference returns None:
 * services.rs:477-481). Builds the network, folds BatchNorm, uploads weights to every GPU of ctx.
 * Replaces resnet::resnet18 / alexnet::alexnet + vs.load (services.rs:513-524). */
int b2i_model_load(b2i_ctx* ctx, const char* model_name, const char* ot_path,
                   const b2i_model_config* cfg, b2i_model** out);
void b2i_model_free(b2i_model* m);

/* Re-read the weights of a loaded model from `ot_path` (same architecture) and swap them in on every GPU:
 * the new device copies are built while the old ones keep serving, then exchanged under the model's lock,
 * so a concurrent predict sees either the old or the new weights, never a mixture. This is the operator
 * side of the reference's `train <sdfs_file> <model>` command, which lands a new `<model>.ot` on every
 * member (reference src/main.rs:231-256, src/services.rs:139-144) but never reloads it (load_model at
 * src/model_utils.rs:24-36 is dead code). On failure the old weights stay active. */
int b2i_model_reload(b2i_model* m, const char* ot_path);

/* Static facts: algorithmic FLOPs per image, kernel launches per micro-batch, bytes resident. */
double b2i_model_flops_per_image(const b2i_model* m);
int b2i_model_launches_per_microbatch(const b2i_model* m);
int b2i_model_num_devices(const b2i_model* m);
long long b2i_model_weight_bytes(const b2i_model* m);
long long b2i_model_activation_bytes(const b2i_model* m);

/* The batched form of the reference's per-image  forward_t(..) .softmax(-1) ; imagenet::top(..,k)
 * (services.rs:493-494), on HOST buffers:
 *   images : u8 [n][224][224][3] (HWC, RGB), already resized + centre-cropped
 *   idx    : int32 [n][k]   class indices, highest probability first (synset_words.txt order)
 *   prob   : float [n][k]   softmax probabilities
 *   logits : float [n][1000] or NULL
 * The batch is split contiguously over the model's GPUs (image i of GPU g's slice keeps its global
 * position), each GPU stages through pinned double buffers on its own streams, and only idx/prob
 * (and logits when asked) come back; there is no inter-GPU traffic. n may be anything from 0 to
 * num_devices * max_batch. Blocking; thread-safe (calls on one model are serialised, different
 * models run concurrently). */
int b2i_predict_u8(b2i_model* m, const uint8_t* images, int n, int k, int32_t* idx, float* prob,
                   float* logits);

/* Same, restricted to the GPUs [first_dev, first_dev + n_dev) of the model's context: the
 * dual-job fair share (reference src/services.rs:199-211) hands each job half of the devices. */
int b2i_predict_u8_on(b2i_model* m, int first_dev, int n_dev, const uint8_t* images, int n, int k,
                      int32_t* idx, float* prob, float* logits);

/* The body of `Member::predict` after the JPEG decode in one call (services.rs:492-494): images[i] is
 * a decoded u8 [heights[i]][widths[i]][3] image in host memory; each is resized + centre-cropped on
 * the GPU exactly as b2i_resize_crop_u8 does, written straight into the device batch, normalised,
 * run through the network and reduced to its top-k. Same sharding and outputs as b2i_predict_u8_on;
 * n_dev <= 0 means all GPUs of the model. Blocking. */
int b2i_predict_raw_u8(b2i_model* m, int first_dev, int n_dev, const uint8_t* const* images,
                       const int* heights, const int* widths, int n, int k, int32_t* idx, float* prob,
                       float* logits);

/* b2i_predict_raw_u8 with one status per image: status[i] is B2I_OK, B2I_EDECODE (images[i] is NULL or
 * has no pixels: the caller's decoder failed on it) or B2I_EINVAL (a side longer than 16384). An image with
 * a non-zero status gets idx -1 / prob 0 and does not stop the others, which are answered exactly as
 * b2i_predict_raw_u8 would answer them alone. The return value reports failures of the call itself. This is
 * how a batched handler keeps the reference's per-id behaviour: an empty directory skips that id
 * (services.rs:490) and a decode failure costs one query (one id per RPC, services.rs:421,492), not 32. */
int b2i_predict_raw_u8_status(b2i_model* m, int first_dev, int n_dev, const uint8_t* const* images,
                              const int* heights, const int* widths, int n, int k, int32_t* idx, float* prob,
                              float* logits, int* status);

/* `Member::predict` from the FILE CONTENTS on: the reference decodes each JPEG on the handler's thread
 * (stb, inside load_image_and_resize, services.rs:492); here jpeg[i] / len[i] is one encoded stream in
 * host memory and the whole batch is decoded on the GPU (nvJPEG, loaded with dlopen when this is first
 * called: NVJPG hardware engines where the device exposes them, else entropy decoding on a pool of host
 * threads with the IDCT on the GPU), resized + cropped by the b2i_resize_crop_u8 kernels straight into the device batch and run
 * through the network. status[i]: B2I_OK, B2I_EDECODE (not a decodable JPEG; that image gets idx -1 /
 * prob 0, the others are answered) or B2I_EINVAL (a side longer than 16384). Fails with B2I_ENODEV when
 * libnvjpeg.so.12 is not installed; callers then decode on the host (the thread pool in service.py) and
 * use b2i_predict_raw_u8_status. Decoders differ by a few LSB per pixel (stb, libjpeg,
 * nvJPEG), so results are compared with the CPU restatement on the pixels b2i_jpeg_decode_u8 returns.
 * Unlike the other predict calls this one (and b2i_predict_raw_u8[_status]) holds the model's lock only for
 * its GPU part: two threads calling it on one model overlap one batch's decode with the other's forward. */
int b2i_predict_jpeg(b2i_model* m, int first_dev, int n_dev, const uint8_t* const* jpeg, const size_t* len,
                     int n, int k, int32_t* idx, float* prob, float* logits, int* status);
/* The decode stage alone: *h, *w always; when out != NULL also the RGB pixels u8 [h][w][3] (cap >= h*w*3). */
int b2i_jpeg_decode_u8(b2i_ctx* ctx, int dev, const uint8_t* jpeg, size_t len, uint8_t* out, size_t cap,
                       int* h, int* w);
/* Which nvJPEG backends opened on GPU `dev`, e.g. "hardware+gpu_hybrid+hybrid". */
int b2i_jpeg_backend(b2i_ctx* ctx, int dev, char* out, size_t cap);

/* Asynchronous pair for throughput: b2i_submit_u8 enqueues the whole request (host->device copy,
 * forward, device->host copy of the results) on all GPUs of the model and returns a ticket at once;
 * b2i_wait blocks until that request is complete and idx/prob(/logits) given at submit time are
 * filled. Up to TWO requests may be in flight, so the copy of request t+1 overlaps the forward of
 * request t; submitting a third before waiting for the first fails with B2I_EINVAL. `images` should
 * be pinned (b2i_host_alloc) and must stay valid, with the output arrays, until b2i_wait returns.
 * This is what a member serving several concurrent `predict` RPCs (reference src/main.rs:74-79:
 * up to 10 channels) uses instead of serialising them. */
int b2i_submit_u8(b2i_model* m, const uint8_t* images, int n, int k, int32_t* idx, float* prob,
                  float* logits, long long* ticket);
int b2i_wait(b2i_model* m, long long ticket);

/* Device-resident form for one GPU of the model (dev = index inside the context): all pointers are
 * device pointers on that GPU, work is enqueued on `cuda_stream` (a cudaStream_t, may be NULL for
 * the default stream) and the call returns without synchronising. Used by the benchmark to time
 * the kernel sequence with inputs already in HBM. */
int b2i_forward_device(b2i_model* m, int dev, const void* d_images_u8, int n, int k, void* d_idx,
                       void* d_prob, void* d_logits, void* cuda_stream);

/* One forward over n <= micro_batch images on GPU `dev` with a CUDA-event pair around every kernel
 * launch (events on `cuda_stream`). ms[i] = duration of launch i in execution order, labels_out =
 * '\n'-separated launch names ("conv:layer1.0.conv1", "maxpool", ...). Returns the number of
 * launches or a negative code. Used by bench.py for the per-kernel share of a step. */
int b2i_model_profile(b2i_model* m, int dev, const void* d_images_u8, int n, int k, void* d_idx,
                      void* d_prob, void* cuda_stream, float* ms, int cap, char* labels_out,
                      size_t labels_cap);

/* Resize + centre-crop of one decoded image, the first half of
 * imagenet::load_image_and_resize(path, 224, 224) (services.rs:492):
 * u8 [h][w][3] -> u8 [224][224][3] on the GPU `dev` of ctx. */
int b2i_resize_crop_u8(b2i_ctx* ctx, int dev, const uint8_t* image, int h, int w, uint8_t* out224);

/* No counterpart in the reference: the same resize + centre crop as b2i_resize_crop_u8 for callers
 * that are not allowed to allocate GPU memory while they run -- a server that hands out one
 * pre-sized memory pool at start-up, or a process pinned to a fixed memory budget. Nothing is
 * allocated here (no cudaMalloc/cudaFree, no pinned memory): d_image (u8 [h][w][3]) and d_out224
 * (u8 [224][224][3]) are DEVICE buffers the caller owns, and the intermediate rows plus the two
 * resampling plans live in a 3,018,240-byte array compiled into the library, which every GPU that
 * loads the library holds until the process exits. Limits: h <= 1024 rows and <= 16384 resampling
 * taps per axis (with h <= 1024 the tap limit is not reached for any width); larger images are refused with B2I_EINVAL and a
 * message naming the limit. The first call on each GPU looks the array's device address up. The call
 * must not be made while `cuda_stream` is being captured into a CUDA graph (it is refused with
 * B2I_EINVAL): the plans are specific to this image's size and are uploaded from host memory that does
 * not outlive the call. Work is enqueued on `cuda_stream`; the call waits on the host only while the
 * runtime stages the plan upload (one copy of a few KB from pageable memory), never for the kernels.
 * All calls on one GPU share the array and it keeps whatever the previous call left in it; each call
 * overwrites every byte it later reads, so nothing needs clearing, but two calls that overlap on the
 * GPU (different streams, or different threads) would corrupt each other's rows: keep calls on one
 * GPU in one stream, or order them with events. */
int b2i_resize_crop_u8_static(b2i_ctx* ctx, int dev, const void* d_image, int h, int w, void* d_out224,
                              void* cuda_stream);

/* CPU-only: the resampling plan b2i_resize_crop_u8 uses for one axis (in_size -> out_size, keeping
 * the 224 outputs from `crop` on). first/count: int[224]; idx/coef: up to cap taps. Returns the
 * number of taps or a negative code. Lets the host logic be checked without a GPU. */
int b2i_resize_plan(int in_size, int out_size, int crop, int* first, int* count, int* idx,
                    float* coef, int cap);
/* CPU-only: target size and crop offsets for an h x w image. */
void b2i_resize_target(int h, int w, int* rw, int* rh, int* crop_x, int* crop_y);

/* Pinned host memory for the staging-free fast path of b2i_predict_u8. */
void* b2i_host_alloc(size_t bytes);
void b2i_host_free(void* p);

/* Copy activation tensor `tensor_id` (0 = normalised input, then one per layer in execution order)
 * of the last predict on GPU `dev` as fp32 NHWC. Needs keep_tensors = 1. Returns numel or < 0. */
long long b2i_model_read_tensor(b2i_model* m, int dev, int tensor_id, int n, float* out,
                                long long cap_elems);
int b2i_model_num_tensors(const b2i_model* m);
int b2i_model_tensor_shape(const b2i_model* m, int tensor_id, int* h, int* w, int* c);

/* ---- single-kernel entry points on HOST buffers (parity tests drive each kernel through these;
 * fp32 NHWC in and out, converted to the kernel's dtype inside). ---------------------------- */

/* conv2d (+ optional eval-BatchNorm fold, bias, residual, ReLU). x: [N][H][W][C]; w: OIHW
 * [Cout][C][R][S]; bn: NULL or [4][Cout] = gamma, beta, running_mean, running_var; residual/out:
 * [N][OH][OW][Cout]. conv_mode as in b2i_model_config (0, 1 or 2), or one of the kernels the model
 * chooses by geometry: 3 halo patch (3x3 s1 p1, 64 -> 64), 5 CTA pair, 6 flattened padded rows (3x3 s1 p1,
 * C % 64 == 0, one N tile, W <= 62), 7 CTA pair with resident weights, 8 CTA pair with two sub-tiles per CTA; C <= 4 always uses the
 * small-channel gather. tile_n 0 = default. */
int b2i_op_conv2d(b2i_ctx* ctx, int dev, int precision, int conv_mode, int weights_tma, int tile_n,
                  const float* x, int N, int H, int W, int C, const float* w, const float* bias,
                  const float* bn, int Cout, int R, int S, int stride, int pad,
                  const float* residual, int relu, float* out);
/* linear (+bias, ReLU): x [M][K], w [Cout][K], out [M][Cout]. */
int b2i_op_linear(b2i_ctx* ctx, int dev, int precision, int tile_n, const float* x, int M, int K,
                  const float* w, const float* bias, int Cout, int relu, float* out);
/* The same with the K dimension of every output tile split over `splits` CTAs (the FC layers of a small batch:
 * few rows, a large weight matrix that all SMs should stream); partial sums are added in split order, so the
 * result is reproducible run to run. 16-bit precisions only; splits = 1 is b2i_op_linear. */
int b2i_op_linear_splitk(b2i_ctx* ctx, int dev, int precision, int tile_n, int splits, const float* x, int M,
                         int K, const float* w, const float* bias, int Cout, int relu, float* out);
int b2i_op_maxpool(b2i_ctx* ctx, int dev, int precision, const float* x, int N, int H, int W, int C,
                   int k, int stride, int pad, float* out);
int b2i_op_adaptive_avgpool(b2i_ctx* ctx, int dev, int precision, const float* x, int N, int H,
                            int W, int C, int OH, int OW, float* out);
int b2i_op_normalize(b2i_ctx* ctx, int dev, int precision, const uint8_t* images, int n,
                     float* out_nhwc4);
int b2i_op_softmax_topk(b2i_ctx* ctx, int dev, const float* logits, int rows, int classes, int k,
                        int32_t* idx, float* prob);

#ifdef __cplusplus
}
#endif
#endif /* B200_INFER_H_ */
