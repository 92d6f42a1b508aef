"""torch-CPU fp32 restatement of the five tch calls under `Member::predict`
(reference src/services.rs:492-494) and of the two model graphs built at src/services.rs:513-524.

Model definitions follow tch 0.4.1's `vision::resnet::resnet18` / `vision::alexnet::alexnet`
([tch-recall]); each op is the ATen operator tch binds (conv2d,
batch_norm eval, relu, max_pool2d, adaptive_avg_pool2d, linear, softmax, topk).  Weight dicts use
'.'-separated names as produced by oracle.ot_format.read_ot.
"""
import numpy as np
import torch
import torch.nn.functional as F

MEAN = (0.485, 0.456, 0.406)
STD = (0.229, 0.224, 0.225)
BN_EPS = 1e-5


def normalize_u8(images_u8):
    """tch imagenet::normalize (second half of load_image_and_resize, services.rs:492):
    u8 HWC [N,224,224,3] -> f32 NCHW, x/255 then (x-mean)/std."""
    x = torch.as_tensor(np.asarray(images_u8)).to(torch.float32).permute(0, 3, 1, 2) / 255.0
    mean = torch.tensor(MEAN, dtype=torch.float32).view(1, 3, 1, 1)
    std = torch.tensor(STD, dtype=torch.float32).view(1, 3, 1, 1)
    return (x - mean) / std


def _bn(x, w, p):
    return F.batch_norm(x, w[p + ".running_mean"], w[p + ".running_var"], w[p + ".weight"], w[p + ".bias"],
                        training=False, eps=BN_EPS)


def _basic_block(x, w, p, stride, has_down):
    y = F.relu(_bn(F.conv2d(x, w[p + ".conv1.weight"], None, stride, 1), w, p + ".bn1"))
    y = _bn(F.conv2d(y, w[p + ".conv2.weight"], None, 1, 1), w, p + ".bn2")
    if has_down:
        x = _bn(F.conv2d(x, w[p + ".downsample.0.weight"], None, stride, 0), w, p + ".downsample.1")
    return F.relu(x + y)


def resnet18_forward(w, x, taps=None):
    """tch resnet18(p, 1000).forward_t(x, train=false). x: f32 [N,3,224,224] -> logits [N,1000].
    `taps`, if a list, receives (name, NCHW tensor) after every conv-level stage."""
    def tap(name, t):
        if taps is not None:
            taps.append((name, t))
        return t
    x = tap("conv1", F.relu(_bn(F.conv2d(x, w["conv1.weight"], None, 2, 3), w, "bn1")))
    x = tap("maxpool", F.max_pool2d(x, 3, 2, 1))
    cin = 64
    for L, cout in enumerate((64, 128, 256, 512), start=1):
        for B in range(2):
            stride = 2 if (L > 1 and B == 0) else 1
            x = tap("layer%d.%d" % (L, B), _basic_block(x, w, "layer%d.%d" % (L, B), stride, stride != 1 or cin != cout))
            cin = cout
    x = tap("avgpool", F.adaptive_avg_pool2d(x, (1, 1))).flatten(1)
    return F.linear(x, w["fc.weight"], w["fc.bias"])


def alexnet_forward(w, x, geometry="tch", taps=None):
    """tch alexnet(p, 1000).forward_t(x, train=false); dropout is the identity in eval.
    geometry 'tch': features.3 = conv(k5, pad 1, stride 2) as tch 0.4.1 builds it [tch-recall];
    'torchvision': conv(k5, pad 2, stride 1)."""
    def tap(name, t):
        if taps is not None:
            taps.append((name, t))
        return t
    p3, s3 = (1, 2) if geometry == "tch" else (2, 1)
    x = tap("features.0", F.relu(F.conv2d(x, w["features.0.weight"], w["features.0.bias"], 4, 2)))
    x = tap("pool0", F.max_pool2d(x, 3, 2))
    x = tap("features.3", F.relu(F.conv2d(x, w["features.3.weight"], w["features.3.bias"], s3, p3)))
    x = tap("pool1", F.max_pool2d(x, 3, 2))
    x = tap("features.6", F.relu(F.conv2d(x, w["features.6.weight"], w["features.6.bias"], 1, 1)))
    x = tap("features.8", F.relu(F.conv2d(x, w["features.8.weight"], w["features.8.bias"], 1, 1)))
    x = tap("features.10", F.relu(F.conv2d(x, w["features.10.weight"], w["features.10.bias"], 1, 1)))
    x = tap("pool2", F.max_pool2d(x, 3, 2))
    x = tap("avgpool", F.adaptive_avg_pool2d(x, (6, 6))).flatten(1)
    x = F.relu(F.linear(x, w["classifier.1.weight"], w["classifier.1.bias"]))
    x = F.relu(F.linear(x, w["classifier.4.weight"], w["classifier.4.bias"]))
    return F.linear(x, w["classifier.6.weight"], w["classifier.6.bias"])


def forward(model_name, w, x, alexnet_geometry="tch"):
    if model_name == "resnet18":
        return resnet18_forward(w, x)
    if model_name == "alexnet":
        return alexnet_forward(w, x, alexnet_geometry)
    raise ValueError("unknown model %r" % model_name)  # the reference returns None (services.rs:480)


def softmax_topk(logits, k):
    """`.softmax(-1, Float)` + imagenet::top(&out, k) (services.rs:493-494): probabilities of the k
    largest classes, sorted descending, with their class indices."""
    p = torch.softmax(logits.to(torch.float32), dim=-1)
    v, i = torch.topk(p, k, dim=-1, largest=True, sorted=True)
    return i.to(torch.int32), v


def predict_u8(model_name, w, images_u8, k=5, alexnet_geometry="tch", batch=64):
    """Whole operator on already-cropped u8 images: (idx [N,k], prob [N,k], logits [N,1000])."""
    outs = []
    with torch.no_grad():
        for i in range(0, len(images_u8), batch):
            outs.append(forward(model_name, w, normalize_u8(images_u8[i:i + batch]), alexnet_geometry))
    logits = torch.cat(outs) if outs else torch.zeros(0, 1000)
    idx, prob = softmax_topk(logits, k)
    return idx.numpy(), prob.numpy(), logits.numpy()


# ---------------------------------------------------------------------------------------------
# bf16 evaluation of the same graphs: what the reference's arithmetic becomes when weights (after
# the BatchNorm fold) and every stored activation are rounded to bf16 while sums stay fp32.  Used by
# the GPU tests to separate "the kernel computes what it should" (tight bound against this) from
# "bf16 storage costs this much against the fp32 reference" (the gap of this function itself).
def _q(t):
    return t.to(torch.bfloat16).to(torch.float32)


def _fold(w, conv, bn):
    inv = 1.0 / torch.sqrt(w[bn + ".running_var"].double() + BN_EPS)
    scale = (w[bn + ".weight"].double() * inv)
    wf = w[conv + ".weight"] * scale.float().view(-1, 1, 1, 1)
    shift = (w[bn + ".bias"].double() - w[bn + ".running_mean"].double() * scale).float()
    return _q(wf), shift


def resnet18_forward_bf16(w, images_u8):
    x = _q(normalize_u8(images_u8))
    wf, b = _fold(w, "conv1", "bn1")
    x = _q(F.relu(F.conv2d(x, wf, b, 2, 3)))
    x = F.max_pool2d(x, 3, 2, 1)
    cin = 64
    for L, cout in enumerate((64, 128, 256, 512), start=1):
        for B in range(2):
            p = "layer%d.%d" % (L, B)
            stride = 2 if (L > 1 and B == 0) else 1
            idn = x
            if stride != 1 or cin != cout:
                wd, bd = _fold(w, p + ".downsample.0", p + ".downsample.1")
                idn = _q(F.conv2d(x, wd, bd, stride, 0))
            wa, ba = _fold(w, p + ".conv1", p + ".bn1")
            t = _q(F.relu(F.conv2d(x, wa, ba, stride, 1)))
            wb, bb = _fold(w, p + ".conv2", p + ".bn2")
            x = _q(F.relu(F.conv2d(t, wb, bb, 1, 1) + idn))
            cin = cout
    x = _q(F.adaptive_avg_pool2d(x, (1, 1))).flatten(1)
    return F.linear(x, _q(w["fc.weight"]), w["fc.bias"])


def alexnet_forward_bf16(w, images_u8, geometry="tch"):
    p3, s3 = (1, 2) if geometry == "tch" else (2, 1)
    x = _q(normalize_u8(images_u8))
    x = F.max_pool2d(_q(F.relu(F.conv2d(x, _q(w["features.0.weight"]), w["features.0.bias"], 4, 2))), 3, 2)
    x = F.max_pool2d(_q(F.relu(F.conv2d(x, _q(w["features.3.weight"]), w["features.3.bias"], s3, p3))), 3, 2)
    x = _q(F.relu(F.conv2d(x, _q(w["features.6.weight"]), w["features.6.bias"], 1, 1)))
    x = _q(F.relu(F.conv2d(x, _q(w["features.8.weight"]), w["features.8.bias"], 1, 1)))
    x = F.max_pool2d(_q(F.relu(F.conv2d(x, _q(w["features.10.weight"]), w["features.10.bias"], 1, 1))), 3, 2)
    if x.shape[-1] != 6:
        x = _q(F.adaptive_avg_pool2d(x, (6, 6)))
    x = x.flatten(1)
    x = _q(F.relu(F.linear(x, _q(w["classifier.1.weight"]), w["classifier.1.bias"])))
    x = _q(F.relu(F.linear(x, _q(w["classifier.4.weight"]), w["classifier.4.bias"])))
    return F.linear(x, _q(w["classifier.6.weight"]), w["classifier.6.bias"])
