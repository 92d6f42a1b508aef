"""The product's C++ `.ot` reader (zip central directory + minimal pickle VM) against a genuinely
libtorch-written archive, against python-written archives that libtorch itself accepts, and
against malformed input.  CPU only: goes through the C ABI's b2i_ot_* entry points."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from oracle import ot_format, synth


def test_reads_libtorch_written_archive(pkg):
    path = os.path.join(GOLDEN, "tiny_libtorch.ot")
    shapes, n = pkg.ot_inspect(path)
    assert n == 5
    assert shapes == {"conv1.weight": (4, 3, 2, 2), "bn1.running_mean": (4,), "layer1.0.conv1.weight": (300,),
                      "fc.bias": (66000,), "strided.view": (6, 4)}
    ref = ot_format.read_ot(path)  # libtorch's own reader
    for name in shapes:
        got = pkg.ot_read_tensor(path, name).reshape(shapes[name])
        assert np.array_equal(got, ref[name].numpy()), name
    # the strided tensor is the transpose of arange(24).reshape(4, 6): strides were honoured
    assert np.array_equal(pkg.ot_read_tensor(path, "strided.view").reshape(6, 4),
                          np.arange(24, dtype=np.float32).reshape(4, 6).T)
    # '|' and '.' are both accepted
    assert np.array_equal(pkg.ot_read_tensor(path, "conv1|weight"), pkg.ot_read_tensor(path, "conv1.weight"))


def test_python_writer_roundtrips_through_libtorch_and_the_product(pkg, tmp_path):
    rng = np.random.default_rng(3)
    tensors = {"a.b.weight": rng.standard_normal((5, 3, 3, 3)).astype(np.float32),
               "a.b.bias": rng.standard_normal(5).astype(np.float32),
               "big": rng.standard_normal(70001).astype(np.float32),
               "scalarish": np.float32([1.5])}
    path = str(tmp_path / "w.ot")
    ot_format.write_ot(path, tensors)
    back = ot_format.read_ot(path)
    for k, v in tensors.items():
        assert np.array_equal(back[k].numpy(), v)
        assert np.array_equal(pkg.ot_read_tensor(path, k).reshape(v.shape), v)


def test_synthetic_resnet18_archive_matches_reference_inventory(pkg, resnet18_ot):
    """102 fp32 tensors, 11,699,112 floats."""
    shapes, n = pkg.ot_inspect(resnet18_ot)
    assert n == 102
    assert sum(int(np.prod(s)) for s in shapes.values()) == 11_699_112
    assert shapes["layer4.0.downsample.0.weight"] == (512, 256, 1, 1)
    w = synth.resnet18_weights()
    assert np.array_equal(pkg.ot_read_tensor(resnet18_ot, "layer2.0.conv1.weight"),
                          w["layer2.0.conv1.weight"].numpy().ravel())


def test_synthetic_alexnet_archive_matches_reference_inventory(pkg, alexnet_ot):
    """16 fp32 tensors, 61,100,840 floats."""
    shapes, n = pkg.ot_inspect(alexnet_ot)
    assert n == 16
    assert sum(int(np.prod(s)) for s in shapes.values()) == 61_100_840


def test_malformed_archives_fail_with_a_message(pkg, tmp_path):
    lfs = tmp_path / "ptr.ot"
    lfs.write_text("version https://git-lfs.github.com/spec/v1\noid sha256:f349\nsize 46831783\n")
    with pytest.raises(pkg.B2IError) as e:
        pkg.ot_inspect(str(lfs))
    assert e.value.code == -2 and "Git-LFS pointer" in str(e.value)
    with pytest.raises(pkg.B2IError) as e:
        pkg.ot_inspect(str(tmp_path / "missing.ot"))
    assert e.value.code == -2
    good = open(os.path.join(GOLDEN, "tiny_libtorch.ot"), "rb").read()
    trunc = tmp_path / "trunc.ot"
    trunc.write_bytes(good[: len(good) // 2])
    with pytest.raises(pkg.B2IError):
        pkg.ot_inspect(str(trunc))
    junk = tmp_path / "junk.ot"
    junk.write_bytes(os.urandom(4096))
    with pytest.raises(pkg.B2IError):
        pkg.ot_inspect(str(junk))
    with pytest.raises(pkg.B2IError):
        pkg.ot_read_tensor(os.path.join(GOLDEN, "tiny_libtorch.ot"), "no.such.tensor")
