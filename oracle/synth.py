"""Seeded synthetic weights with the reference's tensor names and shapes (SURVEY.md Appendix A;
distributions from SURVEY.md 8d config 3).  The real pretrained_models/*.ot are Git-LFS pointers
in the reference checkout, so these stand in for them; the same `.ot` file feeds oracle and GPU."""
import collections

import numpy as np
import torch

# Authentication of the real files, should they ever be present (pretrained_models/*.ot:2-3).
LFS = {
    "resnet18": ("f3491340", 46831783),
    "alexnet": ("8fb43520", 244408272),
}


def _conv(g, cout, cin, k):
    std = (2.0 / (cin * k * k)) ** 0.5  # He normal
    return torch.randn(cout, cin, k, k, generator=g) * std


def _bn(g, out, prefix, c):
    out[prefix + ".weight"] = torch.rand(c, generator=g) + 0.5          # gamma ~ U(0.5, 1.5)
    out[prefix + ".bias"] = torch.randn(c, generator=g) * 0.1           # beta  ~ N(0, 0.1)
    out[prefix + ".running_mean"] = torch.randn(c, generator=g) * 0.1   # mu    ~ N(0, 0.1)
    out[prefix + ".running_var"] = torch.rand(c, generator=g) + 0.5     # var   ~ U(0.5, 1.5)


def resnet18_weights(seed=0):
    g = torch.Generator().manual_seed(seed)
    w = collections.OrderedDict()
    w["conv1.weight"] = _conv(g, 64, 3, 7)
    _bn(g, w, "bn1", 64)
    cin = 64
    for L, cout in enumerate((64, 128, 256, 512), start=1):
        for B in range(2):
            p = "layer%d.%d" % (L, B)
            stride = 2 if (L > 1 and B == 0) else 1
            w[p + ".conv1.weight"] = _conv(g, cout, cin, 3)
            _bn(g, w, p + ".bn1", cout)
            w[p + ".conv2.weight"] = _conv(g, cout, cout, 3)
            _bn(g, w, p + ".bn2", cout)
            if stride != 1 or cin != cout:
                w[p + ".downsample.0.weight"] = _conv(g, cout, cin, 1)
                _bn(g, w, p + ".downsample.1", cout)
            cin = cout
    w["fc.weight"] = torch.randn(1000, 512, generator=g) * 0.01
    w["fc.bias"] = torch.randn(1000, generator=g) * 0.01
    return w


def alexnet_weights(seed=1):
    g = torch.Generator().manual_seed(seed)
    w = collections.OrderedDict()
    for name, cout, cin, k in (("features.0", 64, 3, 11), ("features.3", 192, 64, 5), ("features.6", 384, 192, 3),
                               ("features.8", 256, 384, 3), ("features.10", 256, 256, 3)):
        w[name + ".weight"] = _conv(g, cout, cin, k)
        w[name + ".bias"] = torch.randn(cout, generator=g) * 0.01
    for name, cout, cin in (("classifier.1", 4096, 9216), ("classifier.4", 4096, 4096)):
        w[name + ".weight"] = torch.randn(cout, cin, generator=g) * (2.0 / cin) ** 0.5
        w[name + ".bias"] = torch.randn(cout, generator=g) * 0.01
    w["classifier.6.weight"] = torch.randn(1000, 4096, generator=g) * 0.01
    w["classifier.6.bias"] = torch.randn(1000, generator=g) * 0.01
    return w


def weights(model_name, seed=None):
    if model_name == "resnet18":
        return resnet18_weights(0 if seed is None else seed)
    if model_name == "alexnet":
        return alexnet_weights(1 if seed is None else seed)
    raise ValueError(model_name)


def synthetic_images(n, seed=0):
    """u8 [n,224,224,3], i.i.d. uniform{0..255} from a seeded torch CPU generator (8d config 3)."""
    g = torch.Generator().manual_seed(seed)
    return torch.randint(0, 256, (n, 224, 224, 3), generator=g, dtype=torch.uint8).numpy()


def ensure_ot(model_name, path, seed=None):
    """Write the synthetic archive for `model_name` at `path` unless it is already there."""
    import os
    from .ot_format import write_ot
    if not os.path.exists(path):
        os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
        tmp = "%s.tmp%d" % (path, os.getpid())
        write_ot(tmp, weights(model_name, seed), stem=model_name)
        os.replace(tmp, path)
    return path
