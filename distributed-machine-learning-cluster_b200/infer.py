"""Python face of the C ABI: contexts, models and the single-kernel entry points, on numpy arrays.
Everything here forwards to libb200infer.so; nothing is computed in Python."""
import ctypes

import numpy as np

from . import _lib
from ._lib import (ALEXNET_TCH, ALEXNET_TORCHVISION, CLASSES, IMAGE_BYTES, PREC_BF16, PREC_FP16, PREC_FP32, B2IError,
                   ModelConfig, check)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _ptr(a):
    return a.ctypes.data_as(ctypes.c_void_p) if a is not None else None


def device_count():
    return _lib.lib().b2i_device_count()


class Context:
    """A set of GPUs (replaces `tch::Device::Cpu`, reference src/services.rs:514,520)."""

    def __init__(self, devices=(0,)):
        devices = list(devices)
        arr = (ctypes.c_int * len(devices))(*devices)
        h = ctypes.c_void_p()
        check(_lib.lib().b2i_init(arr, len(devices), ctypes.byref(h)))
        self._h = h
        self.devices = devices

    def close(self):
        if self._h:
            _lib.lib().b2i_shutdown(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # ---- single-kernel entry points (parity tests) -------------------------------------------
    def conv2d(self, x, w, bias=None, bn=None, stride=1, pad=0, residual=None, relu=False, precision=PREC_BF16,
               conv_mode=0, weights_tma=1, tile_n=0, dev=0):
        """x NHWC fp32, w OIHW fp32, bn = (gamma, beta, mean, var) or None -> NHWC fp32."""
        x, w = _f32(x), _f32(w)
        N, H, W, C = x.shape
        Cout, Cin, R, S = w.shape
        assert Cin == C
        OH, OW = (H + 2 * pad - R) // stride + 1, (W + 2 * pad - S) // stride + 1
        out = np.empty((N, OH, OW, Cout), np.float32)
        b = _f32(bias) if bias is not None else None
        bnp = _f32(np.stack(bn)) if bn is not None else None
        r = _f32(residual) if residual is not None else None
        check(_lib.lib().b2i_op_conv2d(self._h, dev, precision, conv_mode, weights_tma, tile_n, _ptr(x), N, H, W, C,
                                       _ptr(w), _ptr(b), _ptr(bnp), Cout, R, S, stride, pad, _ptr(r), int(relu),
                                       _ptr(out)))
        return out

    def linear(self, x, w, bias=None, relu=False, precision=PREC_BF16, tile_n=0, dev=0, splits=1):
        """x [M,K] fp32, w [Cout,K] fp32 -> [M,Cout] fp32; splits > 1: split-K over that many CTAs per output tile."""
        x, w = _f32(x), _f32(w)
        M, K = x.shape
        Cout = w.shape[0]
        out = np.empty((M, Cout), np.float32)
        b = _f32(bias) if bias is not None else None
        if splits > 1:
            check(_lib.lib().b2i_op_linear_splitk(self._h, dev, precision, tile_n, splits, _ptr(x), M, K, _ptr(w), _ptr(b),
                                                  Cout, int(relu), _ptr(out)))
        else:
            check(_lib.lib().b2i_op_linear(self._h, dev, precision, tile_n, _ptr(x), M, K, _ptr(w), _ptr(b), Cout,
                                           int(relu), _ptr(out)))
        return out

    def maxpool(self, x, k, stride, pad, precision=PREC_BF16, dev=0):
        x = _f32(x)
        N, H, W, C = x.shape
        OH, OW = (H + 2 * pad - k) // stride + 1, (W + 2 * pad - k) // stride + 1
        out = np.empty((N, OH, OW, C), np.float32)
        check(_lib.lib().b2i_op_maxpool(self._h, dev, precision, _ptr(x), N, H, W, C, k, stride, pad, _ptr(out)))
        return out

    def adaptive_avgpool(self, x, oh, ow, precision=PREC_BF16, dev=0):
        x = _f32(x)
        N, H, W, C = x.shape
        out = np.empty((N, oh, ow, C), np.float32)
        check(_lib.lib().b2i_op_adaptive_avgpool(self._h, dev, precision, _ptr(x), N, H, W, C, oh, ow, _ptr(out)))
        return out

    def normalize(self, images_u8, precision=PREC_BF16, dev=0):
        img = np.ascontiguousarray(images_u8, dtype=np.uint8)
        n = img.shape[0]
        assert img.shape[1:] == (224, 224, 3)
        out = np.empty((n, 224, 224, 4), np.float32)
        check(_lib.lib().b2i_op_normalize(self._h, dev, precision, _ptr(img), n, _ptr(out)))
        return out

    def softmax_topk(self, logits, k, dev=0):
        lg = _f32(logits)
        rows, classes = lg.shape
        idx = np.empty((rows, k), np.int32)
        prob = np.empty((rows, k), np.float32)
        check(_lib.lib().b2i_op_softmax_topk(self._h, dev, _ptr(lg), rows, classes, k, _ptr(idx), _ptr(prob)))
        return idx, prob

    def resize_crop(self, image_u8, dev=0):
        """u8 [H,W,3] -> u8 [224,224,3] (tch load_image_and_resize minus decode and normalise)."""
        img = np.ascontiguousarray(image_u8, dtype=np.uint8)
        h, w, c = img.shape
        assert c == 3
        out = np.empty((224, 224, 3), np.uint8)
        check(_lib.lib().b2i_resize_crop_u8(self._h, dev, _ptr(img), h, w, _ptr(out)))
        return out

    def jpeg_decode(self, data, dev=0):
        """One encoded JPEG (bytes) -> u8 [H,W,3] RGB as the GPU decoder produces it (the decode stage of
        Model.predict_jpeg alone; raises B2IError with code EDECODE for a stream it cannot decode)."""
        buf = np.frombuffer(data, dtype=np.uint8)
        h, w = ctypes.c_int(), ctypes.c_int()
        check(_lib.lib().b2i_jpeg_decode_u8(self._h, dev, _ptr(buf), buf.size, None, 0, ctypes.byref(h), ctypes.byref(w)))
        out = np.empty((h.value, w.value, 3), np.uint8)
        check(_lib.lib().b2i_jpeg_decode_u8(self._h, dev, _ptr(buf), buf.size, _ptr(out), out.size, ctypes.byref(h),
                                            ctypes.byref(w)))
        return out

    def jpeg_backend(self, dev=0):
        """Which nvJPEG backends opened on this GPU ("hardware+gpu_hybrid+hybrid", ...)."""
        buf = ctypes.create_string_buffer(128)
        check(_lib.lib().b2i_jpeg_backend(self._h, dev, buf, len(buf)))
        return buf.value.decode()

    def resize_crop_static(self, d_image, h, w, d_out224, stream=None, dev=0):
        """Allocation-free resize + crop for callers that may not allocate GPU memory: `d_image`
        (u8 [h,w,3]) and `d_out224` (u8 [224,224,3]) are device addresses (ints) owned by the
        caller, the scratch is a fixed array inside the library (h <= 1024). Enqueues on `stream`
        and returns; calls on one GPU must not overlap."""
        check(_lib.lib().b2i_resize_crop_u8_static(self._h, dev, d_image, h, w, d_out224, stream))


def resize_target(h, w):
    rw, rh, cx, cy = ctypes.c_int(), ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    _lib.lib().b2i_resize_target(h, w, ctypes.byref(rw), ctypes.byref(rh), ctypes.byref(cx), ctypes.byref(cy))
    return rw.value, rh.value, cx.value, cy.value


def resize_plan(in_size, out_size, crop):
    cap = 224 * (8 + 4 * (in_size // max(out_size, 1) + 2) + 8)
    first = np.empty(224, np.int32)
    count = np.empty(224, np.int32)
    idx = np.empty(cap, np.int32)
    coef = np.empty(cap, np.float32)
    n = check(_lib.lib().b2i_resize_plan(in_size, out_size, crop, _ptr(first), _ptr(count), _ptr(idx), _ptr(coef), cap))
    return [(idx[f:f + c].copy(), coef[f:f + c].copy()) for f, c in zip(first, count)], n


def ot_inspect(path):
    buf = ctypes.create_string_buffer(1 << 16)
    n = check(_lib.lib().b2i_ot_inspect(str(path).encode(), buf, len(buf)))
    out = {}
    for line in buf.value.decode().splitlines():
        name, _, shape = line.partition(":")
        out[name] = tuple(int(d) for d in shape.split("x")) if shape else ()
    return out, n


def ot_read_tensor(path, name):
    n = check(_lib.lib().b2i_ot_read_tensor(str(path).encode(), name.encode(), None, 0))
    out = np.empty(n, np.float32)
    check(_lib.lib().b2i_ot_read_tensor(str(path).encode(), name.encode(), _ptr(out), n))
    return out


class Model:
    """A loaded network on every GPU of a context (replaces `Mutex<Box<dyn ModuleT>>`,
    reference src/services.rs:455-456, built at 513-524)."""

    def __init__(self, ctx, model_name, ot_path, precision=PREC_BF16, max_batch=256, micro_batch=0,
                 alexnet_geometry=ALEXNET_TCH, conv_mode=-1, weights_tma=1, tile_n_max=0, keep_tensors=False,
                 fused_stem=True, patch_conv=None, pair_mma=None,
                 fuse_downsample=None):
        cfg = ModelConfig()
        _lib.lib().b2i_default_config(ctypes.byref(cfg))
        cfg.precision, cfg.max_batch, cfg.micro_batch = precision, max_batch, micro_batch
        cfg.alexnet_geometry, cfg.conv_mode, cfg.weights_tma = alexnet_geometry, conv_mode, weights_tma
        cfg.tile_n_max, cfg.keep_tensors, cfg.fused_stem = tile_n_max, int(keep_tensors), int(fused_stem)
        if patch_conv is not None:
            cfg.patch_conv = int(patch_conv)
        if pair_mma is not None:
            cfg.pair_mma = int(pair_mma)
        if fuse_downsample is not None:
            cfg.fuse_downsample = int(fuse_downsample)
        h = ctypes.c_void_p()
        check(_lib.lib().b2i_model_load(ctx._h, model_name.encode(), str(ot_path).encode(), ctypes.byref(cfg),
                                        ctypes.byref(h)))
        self._h = h
        self.ctx = ctx
        self.name = model_name
        self.max_batch = max_batch
        self.precision = precision

    def reload(self, ot_path):
        """Swap in the weights of another `.ot` archive of the same architecture (the `train` flow)."""
        check(_lib.lib().b2i_model_reload(self._h, str(ot_path).encode()))

    def close(self):
        if self._h:
            _lib.lib().b2i_model_free(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    @property
    def flops_per_image(self):
        return _lib.lib().b2i_model_flops_per_image(self._h)

    @property
    def launches_per_microbatch(self):
        return _lib.lib().b2i_model_launches_per_microbatch(self._h)

    @property
    def num_devices(self):
        return _lib.lib().b2i_model_num_devices(self._h)

    @property
    def weight_bytes(self):
        return _lib.lib().b2i_model_weight_bytes(self._h)

    @property
    def activation_bytes(self):
        return _lib.lib().b2i_model_activation_bytes(self._h)

    def predict_u8(self, images_u8, k=5, want_logits=False, devices=None):
        """images u8 [n,224,224,3] (host) -> (idx int32 [n,k], prob f32 [n,k][, logits f32 [n,1000]])."""
        img = images_u8 if (isinstance(images_u8, np.ndarray) and images_u8.dtype == np.uint8 and
                            images_u8.flags.c_contiguous) else np.ascontiguousarray(images_u8, dtype=np.uint8)
        n = img.shape[0]
        assert img.shape[1:] == (224, 224, 3), img.shape
        idx = np.empty((n, k), np.int32)
        prob = np.empty((n, k), np.float32)
        logits = np.empty((n, CLASSES), np.float32) if want_logits else None
        if devices is None:
            check(_lib.lib().b2i_predict_u8(self._h, _ptr(img), n, k, _ptr(idx), _ptr(prob), _ptr(logits)))
        else:
            first, count = devices
            check(_lib.lib().b2i_predict_u8_on(self._h, first, count, _ptr(img), n, k, _ptr(idx), _ptr(prob),
                                               _ptr(logits)))
        return (idx, prob, logits) if want_logits else (idx, prob)

    def predict_raw(self, images, k=5, want_logits=False, devices=None):
        """Decoded images of any size (list of u8 [H,W,3] arrays) -> GPU resize + crop -> forward ->
        (idx, prob[, logits]); the whole Member::predict body after the JPEG decode in one call."""
        imgs = [np.ascontiguousarray(im, dtype=np.uint8) for im in images]
        n = len(imgs)
        ptrs = (ctypes.c_void_p * max(n, 1))(*[im.ctypes.data for im in imgs])
        hs = (ctypes.c_int * max(n, 1))(*[im.shape[0] for im in imgs])
        ws = (ctypes.c_int * max(n, 1))(*[im.shape[1] for im in imgs])
        idx = np.empty((n, k), np.int32)
        prob = np.empty((n, k), np.float32)
        logits = np.empty((n, CLASSES), np.float32) if want_logits else None
        first, count = devices if devices is not None else (0, 0)
        check(_lib.lib().b2i_predict_raw_u8(self._h, first, count, ptrs, hs, ws, n, k, _ptr(idx), _ptr(prob), _ptr(logits)))
        return (idx, prob, logits) if want_logits else (idx, prob)

    def predict_raw_status(self, images, k=5, want_logits=False, devices=None):
        """predict_raw with one status per image: entries of `images` may be None (the caller's decoder
        failed); those come back with status EDECODE and idx -1 while the rest are answered.
        -> (idx, prob[, logits], status int32 [n])."""
        imgs = [None if im is None else np.ascontiguousarray(im, dtype=np.uint8) for im in images]
        n = len(imgs)
        ptrs = (ctypes.c_void_p * max(n, 1))(*[None if im is None else im.ctypes.data for im in imgs])
        hs = (ctypes.c_int * max(n, 1))(*[0 if im is None else im.shape[0] for im in imgs])
        ws = (ctypes.c_int * max(n, 1))(*[0 if im is None else im.shape[1] for im in imgs])
        idx = np.empty((n, k), np.int32)
        prob = np.empty((n, k), np.float32)
        status = np.zeros(n, np.int32)
        logits = np.empty((n, CLASSES), np.float32) if want_logits else None
        first, count = devices if devices is not None else (0, 0)
        check(_lib.lib().b2i_predict_raw_u8_status(self._h, first, count, ptrs, hs, ws, n, k, _ptr(idx), _ptr(prob),
                                                   _ptr(logits), _ptr(status)))
        return (idx, prob, logits, status) if want_logits else (idx, prob, status)

    def predict_jpeg(self, blobs, k=5, want_logits=False, devices=None):
        """Encoded JPEG streams (list of bytes; None / b"" = nothing to decode) -> GPU decode -> resize +
        crop -> forward -> (idx, prob[, logits], status int32 [n]).  The whole body of Member::predict
        from the file contents on."""
        bufs = [np.frombuffer(b, dtype=np.uint8) if b else None for b in blobs]
        n = len(bufs)
        ptrs = (ctypes.c_void_p * max(n, 1))(*[None if b is None else b.ctypes.data for b in bufs])
        lens = (ctypes.c_size_t * max(n, 1))(*[0 if b is None else b.size for b in bufs])
        idx = np.empty((n, k), np.int32)
        prob = np.empty((n, k), np.float32)
        status = np.zeros(n, np.int32)
        logits = np.empty((n, CLASSES), np.float32) if want_logits else None
        first, count = devices if devices is not None else (0, 0)
        check(_lib.lib().b2i_predict_jpeg(self._h, first, count, ptrs, lens, n, k, _ptr(idx), _ptr(prob), _ptr(logits),
                                          _ptr(status)))
        return (idx, prob, logits, status) if want_logits else (idx, prob, status)

    def submit_u8(self, images_u8, idx, prob, logits=None):
        """Asynchronous predict: enqueue and return a ticket. `images_u8` (ideally a PinnedBuffer
        view), `idx` int32 [n,k] and `prob` f32 [n,k] must stay alive until wait(ticket) returns."""
        n, k = idx.shape
        t = ctypes.c_longlong(-1)
        check(_lib.lib().b2i_submit_u8(self._h, _ptr(images_u8), n, k, _ptr(idx), _ptr(prob), _ptr(logits), ctypes.byref(t)))
        return t.value

    def wait(self, ticket):
        check(_lib.lib().b2i_wait(self._h, ticket))

    def forward_device(self, d_images, n, k, d_idx, d_prob, d_logits=None, stream=None, dev=0):
        """Raw device pointers (ints) on GPU `dev`; enqueues on `stream` (cudaStream_t as int)."""
        check(_lib.lib().b2i_forward_device(self._h, dev, d_images, n, k, d_idx, d_prob, d_logits, stream))

    def profile(self, d_images, n, k, d_idx, d_prob, stream=None, dev=0):
        """[(launch name, milliseconds)] for one micro-batch, CUDA events around every launch."""
        ms = np.zeros(256, np.float32)
        names = ctypes.create_string_buffer(1 << 14)
        cnt = check(_lib.lib().b2i_model_profile(self._h, dev, d_images, n, k, d_idx, d_prob, stream, _ptr(ms), 256,
                                                 names, len(names)))
        return list(zip(names.value.decode().splitlines(), ms[:cnt].tolist()))

    def num_tensors(self):
        return _lib.lib().b2i_model_num_tensors(self._h)

    def tensor_shape(self, tid):
        h, w, c = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
        check(_lib.lib().b2i_model_tensor_shape(self._h, tid, ctypes.byref(h), ctypes.byref(w), ctypes.byref(c)))
        return h.value, w.value, c.value

    def read_tensor(self, tid, n, dev=0):
        h, w, c = self.tensor_shape(tid)
        out = np.empty((n, h, w, c), np.float32)
        check(_lib.lib().b2i_model_read_tensor(self._h, dev, tid, n, _ptr(out), out.size))
        return out


class PinnedBuffer:
    """Page-locked host memory viewed as a numpy u8 array (staging-free path of predict_u8)."""

    def __init__(self, nbytes):
        p = _lib.lib().b2i_host_alloc(nbytes)
        if not p:
            raise B2IError(_lib.ENOMEM, "pinned allocation of %d bytes failed" % nbytes)
        self._p = p
        self.array = np.ctypeslib.as_array((ctypes.c_uint8 * nbytes).from_address(p))

    def close(self):
        if self._p:
            self.array = None
            _lib.lib().b2i_host_free(self._p)
            self._p = None
