"""B200-native inference path for the cluster's `predict` / `jobs` commands.

The product is `libb200infer.so` (hand-written sm_100a CUDA + C++ host, C ABI in
include/b200_infer.h).  This package is the thin host-side mirror of the reference's operator
interface over that ABI; importing it without the built library raises.
"""
from . import _lib
from ._lib import B2IError, LIB_PATH
from .infer import (ALEXNET_TCH, ALEXNET_TORCHVISION, CLASSES, IMAGE_BYTES, PREC_BF16, PREC_FP16, PREC_FP32, Context, Model,
                    PinnedBuffer, device_count, ot_inspect, ot_read_tensor, resize_plan, resize_target)
from .service import (MISSING, SKIPPED, UNDECODABLE, Job, Leader, MemberState, assign_members, decode_bytes, decode_image, jobs_report, load_labels, load_model,
                      predict_image, sample_predict)

_lib.lib()  # fail at import time if the extension is missing
