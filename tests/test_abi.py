"""The C-ABI library loads and exports every symbol include/b200_infer.h declares; without a GPU
the compute entry points fail loudly instead of falling back."""
import ctypes
import os
import re

import pytest

from conftest import ROOT


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "b200_infer.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b2i_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_the_boundary():
    syms = declared_symbols()
    for must in ("b2i_init", "b2i_model_load", "b2i_predict_u8", "b2i_predict_u8_on", "b2i_model_free",
                 "b2i_last_error", "b2i_resize_crop_u8", "b2i_forward_device"):
        assert must in syms


def test_library_exports_every_declared_symbol(pkg):
    handle = ctypes.CDLL(pkg.LIB_PATH)
    missing = [s for s in declared_symbols() if not hasattr(handle, s)]
    assert not missing, missing


def test_python_binding_covers_every_declared_symbol(pkg):
    from importlib import import_module
    sigs = import_module("distributed-machine-learning-cluster_b200._lib").SIGNATURES
    assert sorted(sigs) == declared_symbols()


def test_no_torch_in_the_boundary():
    text = open(os.path.join(ROOT, "include", "b200_infer.h")).read()
    assert "torch" not in text.lower().replace("torchvision", "")
    assert "at::" not in text and "std::" not in text


def test_library_does_not_link_torch_or_cpu_blas(pkg):
    import subprocess
    out = subprocess.run(["ldd", pkg.LIB_PATH], capture_output=True, text=True).stdout
    for banned in ("libtorch", "libc10", "libcudnn", "libcublas", "libmkl", "libopenblas"):
        assert banned not in out, out


@pytest.mark.skipif(os.path.exists("/dev/nvidia0"), reason="GPU present")
def test_compute_fails_loudly_without_gpu(pkg):
    assert pkg.device_count() == 0
    with pytest.raises(pkg.B2IError) as e:
        pkg.Context((0,))
    assert e.value.code == -4 and "no CPU path" in str(e.value)


def test_product_never_imports_the_oracle():
    pkgdir = os.path.join(ROOT, "distributed-machine-learning-cluster_b200")
    for dirpath, _, files in os.walk(pkgdir):
        for fn in files:
            if fn.endswith((".py", ".cpp", ".cu", ".h", ".cuh")):
                src = open(os.path.join(dirpath, fn)).read()
                assert "oracle" not in src.lower(), (fn, "mentions the oracle")


def _struct_fields(text, opener, field_re):
    body = text[text.index(opener):]
    body = body[:body.index("}")]
    return re.findall(field_re, body)


def test_model_config_layout_agrees_everywhere():
    """b2i_model_config is passed by pointer and filled by b2i_default_config: a binding with fewer fields
    gets written past its end.  The header, the ctypes table, the Rust -sys crate and the block a maintainer
    would copy out of INTEGRATION.md must list the same fields in the same order."""
    from importlib import import_module
    header = open(os.path.join(ROOT, "include", "b200_infer.h")).read()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    h_fields = _struct_fields(header, "typedef struct b2i_model_config", r"\bint\s+(\w+)\s*;")
    assert len(h_fields) >= 12 and h_fields[0] == "precision"
    py_fields = [n for n, _ in import_module("distributed-machine-learning-cluster_b200._lib").ModelConfig._fields_]
    rust = open(os.path.join(ROOT, "integration", "rust", "b200-infer-sys", "src", "lib.rs")).read()
    rs_fields = _struct_fields(rust, "pub struct b2i_model_config", r"pub\s+(\w+)\s*:\s*c_int")
    doc = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    md_fields = _struct_fields(doc, "pub struct b2i_model_config", r"pub\s+(\w+)\s*:\s*c_int")
    assert py_fields == h_fields
    assert rs_fields == h_fields
    assert md_fields == h_fields


def test_error_codes_agree(pkg):
    from importlib import import_module
    header = open(os.path.join(ROOT, "include", "b200_infer.h")).read()
    codes = {m.group(1): int(m.group(2)) for m in re.finditer(r"#define B2I_(E[A-Z]+) \((-\d+)\)", header)}
    lib = import_module("distributed-machine-learning-cluster_b200._lib")
    assert codes and all(getattr(lib, name) == val for name, val in codes.items()), codes
    rust = open(os.path.join(ROOT, "integration", "rust", "b200-infer-sys", "src", "lib.rs")).read()
    for name, val in codes.items():
        assert re.search(r"B2I_%s: c_int = %d;" % (name, val), rust), name


def test_every_entry_point_stops_exceptions():
    """No C++ exception may cross the C boundary: every extern "C" function with a body of more than one
    statement runs it inside guarded(...)."""
    src = open(os.path.join(ROOT, "distributed-machine-learning-cluster_b200", "csrc", "capi.cpp")).read()
    ext = src[src.index('extern "C" {'):]
    heads = list(re.finditer(r"^(?:static )?(?:int|long long) (b2i_\w+|ctx_dev)\([^;{]*\{\s*$", ext, flags=re.M))
    assert len(heads) >= 25
    for h in heads:
        nxt = ext[h.end():h.end() + 60]
        assert "return guarded(" in nxt, h.group(1)
