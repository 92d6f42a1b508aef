"""ctypes binding of include/b200_infer.h.  The shared library is the product: if it is missing
this module raises (there is no Python or CPU fallback)."""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# B2I_LIB: another build of the same library (A/B timing of two builds inside one GPU session)
LIB_PATH = os.environ.get("B2I_LIB") or os.path.join(_HERE, "libb200infer.so")

OK, EINVAL, EIO, ECUDA, ENODEV, ENOMEM, EINTERNAL, EDECODE = 0, -1, -2, -3, -4, -5, -6, -7
PREC_BF16, PREC_FP32, PREC_FP16 = 0, 1, 2
ALEXNET_TCH, ALEXNET_TORCHVISION = 0, 1
IMAGE_BYTES = 224 * 224 * 3
CLASSES = 1000


class ModelConfig(ctypes.Structure):
    _fields_ = [
        ("precision", ctypes.c_int),
        ("max_batch", ctypes.c_int),
        ("micro_batch", ctypes.c_int),
        ("alexnet_geometry", ctypes.c_int),
        ("conv_mode", ctypes.c_int),
        ("weights_tma", ctypes.c_int),
        ("tile_n_max", ctypes.c_int),
        ("keep_tensors", ctypes.c_int),
        ("fused_stem", ctypes.c_int),
        ("patch_conv", ctypes.c_int),
        ("pair_mma", ctypes.c_int),
        ("fuse_downsample", ctypes.c_int),
    ]


class B2IError(RuntimeError):
    def __init__(self, code, message):
        super().__init__("b200_infer error %d: %s" % (code, message))
        self.code = code


_c = ctypes
_vp, _i, _f, _ll, _sz = _c.c_void_p, _c.c_int, _c.c_float, _c.c_longlong, _c.c_size_t
_ip = _c.POINTER(_c.c_int)

# name -> (restype, argtypes); mirrors include/b200_infer.h one to one
SIGNATURES = {
    "b2i_last_error": (_c.c_char_p, []),
    "b2i_device_count": (_i, []),
    "b2i_init": (_i, [_ip, _i, _c.POINTER(_vp)]),
    "b2i_shutdown": (None, [_vp]),
    "b2i_default_config": (None, [_c.POINTER(ModelConfig)]),
    "b2i_ot_inspect": (_i, [_c.c_char_p, _c.c_char_p, _sz]),
    "b2i_ot_read_tensor": (_ll, [_c.c_char_p, _c.c_char_p, _vp, _ll]),
    "b2i_model_load": (_i, [_vp, _c.c_char_p, _c.c_char_p, _c.POINTER(ModelConfig), _c.POINTER(_vp)]),
    "b2i_model_free": (None, [_vp]),
    "b2i_model_reload": (_i, [_vp, _c.c_char_p]),
    "b2i_model_flops_per_image": (_c.c_double, [_vp]),
    "b2i_model_launches_per_microbatch": (_i, [_vp]),
    "b2i_model_num_devices": (_i, [_vp]),
    "b2i_model_weight_bytes": (_ll, [_vp]),
    "b2i_model_activation_bytes": (_ll, [_vp]),
    "b2i_predict_u8": (_i, [_vp, _vp, _i, _i, _vp, _vp, _vp]),
    "b2i_predict_u8_on": (_i, [_vp, _i, _i, _vp, _i, _i, _vp, _vp, _vp]),
    "b2i_predict_raw_u8": (_i, [_vp, _i, _i, _vp, _vp, _vp, _i, _i, _vp, _vp, _vp]),
    "b2i_predict_raw_u8_status": (_i, [_vp, _i, _i, _vp, _vp, _vp, _i, _i, _vp, _vp, _vp, _vp]),
    "b2i_predict_jpeg": (_i, [_vp, _i, _i, _vp, _vp, _i, _i, _vp, _vp, _vp, _vp]),
    "b2i_jpeg_decode_u8": (_i, [_vp, _i, _vp, _sz, _vp, _sz, _ip, _ip]),
    "b2i_jpeg_backend": (_i, [_vp, _i, _c.c_char_p, _sz]),
    "b2i_submit_u8": (_i, [_vp, _vp, _i, _i, _vp, _vp, _vp, _c.POINTER(_ll)]),
    "b2i_wait": (_i, [_vp, _ll]),
    "b2i_forward_device": (_i, [_vp, _i, _vp, _i, _i, _vp, _vp, _vp, _vp]),
    "b2i_model_profile": (_i, [_vp, _i, _vp, _i, _i, _vp, _vp, _vp, _vp, _i, _c.c_char_p, _sz]),
    "b2i_resize_crop_u8": (_i, [_vp, _i, _vp, _i, _i, _vp]),
    "b2i_resize_crop_u8_static": (_i, [_vp, _i, _vp, _i, _i, _vp, _vp]),
    "b2i_resize_plan": (_i, [_i, _i, _i, _vp, _vp, _vp, _vp, _i]),
    "b2i_resize_target": (None, [_i, _i, _ip, _ip, _ip, _ip]),
    "b2i_host_alloc": (_vp, [_sz]),
    "b2i_host_free": (None, [_vp]),
    "b2i_model_read_tensor": (_ll, [_vp, _i, _i, _i, _vp, _ll]),
    "b2i_model_num_tensors": (_i, [_vp]),
    "b2i_model_tensor_shape": (_i, [_vp, _i, _ip, _ip, _ip]),
    "b2i_op_conv2d": (_i, [_vp, _i, _i, _i, _i, _i, _vp, _i, _i, _i, _i, _vp, _vp, _vp, _i, _i, _i, _i, _i, _vp, _i, _vp]),
    "b2i_op_linear": (_i, [_vp, _i, _i, _i, _vp, _i, _i, _vp, _vp, _i, _i, _vp]),
    "b2i_op_linear_splitk": (_i, [_vp, _i, _i, _i, _i, _vp, _i, _i, _vp, _vp, _i, _i, _vp]),
    "b2i_op_maxpool": (_i, [_vp, _i, _i, _vp, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "b2i_op_adaptive_avgpool": (_i, [_vp, _i, _i, _vp, _i, _i, _i, _i, _i, _i, _vp]),
    "b2i_op_normalize": (_i, [_vp, _i, _i, _vp, _i, _vp]),
    "b2i_op_softmax_topk": (_i, [_vp, _i, _vp, _i, _i, _i, _vp, _vp]),
}

_lib = None


def lib():
    """The loaded library; raises if the extension has not been built (no fallback exists)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                "%s is missing: build it with distributed-machine-learning-cluster_b200/build.sh "
                "(or __graft_entry__.build()); this package has no CPU or Python fallback" % LIB_PATH)
        handle = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)  # AttributeError here = header and library disagree
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def check(rc):
    if rc < 0:
        raise B2IError(rc, lib().b2i_last_error().decode("utf-8", "replace"))
    return rc
