"""Seeded synthetic weights with the reference's tensor names and shapes (SURVEY.md Appendix A;
distributions from SURVEY.md 8d config 3).  The real pretrained_models/*.ot are Git-LFS pointers
in the reference checkout, so these stand in for them; the same `.ot` file feeds oracle and GPU."""
import collections

import numpy as np
import torch

# The real files, should they ever be present: Git-LFS object id (sha256 of the content) and size from the
# pointer files the reference commits in their place (reference pretrained_models/resnet18.ot:2-3,
# pretrained_models/alexnet.ot:2-3).
LFS = {
    "resnet18": ("f349134076ce5f2dc50f1003b798fe068c7ab0695c757450723d259441b20ebb", 46831783),
    "alexnet": ("8fb43520a45c164aba438753cdc5fd4b6f5c5dd5436db6d69efdf1e5532ff0e8", 244408272),
}


def real_weights(model_name, search=None):
    """Path of the reference's pretrained `<model>.ot` if a file with the LFS pointer's size AND sha256 is
    present, else None.  Looked for in $B2I_WEIGHT_DIR, ./pretrained_models and <repo>/pretrained_models
    (where the reference keeps them, src/services.rs:516,522).  Tests and bench.py switch to the real
    weights by themselves the moment such a file appears; a file that does not authenticate is ignored
    (a Git-LFS pointer, a truncated download, other weights)."""
    import hashlib
    import os
    oid, size = LFS[model_name]
    here = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    dirs = list(search or []) + [os.environ.get("B2I_WEIGHT_DIR", ""), "pretrained_models",
                                 os.path.join(here, "pretrained_models")]
    for d in dirs:
        path = os.path.join(d, model_name + ".ot") if d else ""
        if not path or not os.path.isfile(path) or os.path.getsize(path) != size:
            continue
        h = hashlib.sha256()
        with open(path, "rb") as f:
            for block in iter(lambda: f.read(1 << 22), b""):
                h.update(block)
        if h.hexdigest() == oid:
            return os.path.abspath(path)
    return None


def _conv(g, cout, cin, k):
    std = (2.0 / (cin * k * k)) ** 0.5  # He normal
    return torch.randn(cout, cin, k, k, generator=g) * std


def _bn(g, out, prefix, c):
    out[prefix + ".weight"] = torch.rand(c, generator=g) + 0.5          # gamma ~ U(0.5, 1.5)
    out[prefix + ".bias"] = torch.randn(c, generator=g) * 0.1           # beta  ~ N(0, 0.1)
    out[prefix + ".running_mean"] = torch.randn(c, generator=g) * 0.1   # mu    ~ N(0, 0.1)
    out[prefix + ".running_var"] = torch.rand(c, generator=g) + 0.5     # var   ~ U(0.5, 1.5)


# The default synthetic ResNet-18 produces logits of standard deviation 9 (|max| 41): tolerances quoted for
# "logits" (BASELINE.json: 2e-2 / 1e-4 max-abs) were written with a pretrained network in mind, whose
# logits have a standard deviation of 2-3.  The "ps" (pretrained-scale) variant is the same network with
# its classifier scaled so that the logits land there; flat absolute bounds are checked on it.
PS_FC_GAIN = 0.28


def resnet18_weights(seed=0, fc_gain=1.0):
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
    w["fc.weight"] = torch.randn(1000, 512, generator=g) * 0.01 * fc_gain
    w["fc.bias"] = torch.randn(1000, generator=g) * 0.01 * fc_gain
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
    if model_name == "resnet18_ps":
        return resnet18_weights(0 if seed is None else seed, fc_gain=PS_FC_GAIN)
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
        write_ot(tmp, weights(model_name, seed), stem=model_name.split("_")[0])
        os.replace(tmp, path)
    return path
