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


def _central_directory(blob):
    """[(entry offset, name, name length, extra length)] of a (non-ZIP64) archive."""
    import struct
    eocd = blob.rfind(b"PK\x05\x06")
    count, = struct.unpack_from("<H", blob, eocd + 10)
    p, = struct.unpack_from("<I", blob, eocd + 16)
    out = []
    for _ in range(count):
        nlen, elen, clen = struct.unpack_from("<HHH", blob, p + 28)
        out.append((p, blob[p + 46:p + 46 + nlen].decode(), nlen, elen))
        p += 46 + nlen + elen + clen
    return eocd, out


def test_crafted_sizes_and_offsets_are_refused_not_followed(pkg, tmp_path):
    """Archives arrive through the `train` file copy, so the reader treats every size and offset in them
    as untrusted: values that would wrap a 64-bit sum, ZIP64 records that claim to reach past their extra
    field and local headers at the very end of the file all fail with B2I_EIO instead of being read."""
    import struct
    good = bytearray(open(os.path.join(GOLDEN, "tiny_libtorch.ot"), "rb").read())
    eocd, entries = _central_directory(bytes(good))
    payload = [e for e in entries if "/data/" in e[1]][0]
    pkl = [e for e in entries if e[1].endswith("data.pkl")][0]

    def write(name, blob):
        p = tmp_path / name
        p.write_bytes(bytes(blob))
        return str(p)

    def rebuild_with_zip64_extra(entry, usize64, record_size=None):
        """Central-directory entry `entry` gets usize = 0xFFFFFFFF and a ZIP64 extra record carrying usize64."""
        p, _, nlen, elen = entry
        b = bytearray(good)
        struct.pack_into("<I", b, p + 24, 0xFFFFFFFF)
        extra = struct.pack("<HHQ", 1, 8 if record_size is None else record_size, usize64)
        old_end = p + 46 + nlen + elen
        b[p + 46 + nlen:old_end] = extra
        struct.pack_into("<H", b, p + 30, len(extra))
        grow = len(extra) - elen
        ne = eocd + grow
        cd_size, = struct.unpack_from("<I", b, ne + 12)
        struct.pack_into("<I", b, ne + 12, cd_size + grow)
        if bytes(b[ne - 20:ne - 16]) == b"PK\x06\x07":      # ZIP64 locator + end record (libtorch always writes them)
            e64, = struct.unpack_from("<Q", b, ne - 12)
            struct.pack_into("<Q", b, ne - 12, e64 + grow)
            size64, = struct.unpack_from("<Q", b, e64 + grow + 40)
            struct.pack_into("<Q", b, e64 + grow + 40, size64 + grow)
        return b

    cases = {
        # (a) sizes near 2^64: offset + size wraps to a small number
        "wrap_payload.ot": rebuild_with_zip64_extra(payload, 2 ** 64 - 16),
        "wrap_pickle.ot": rebuild_with_zip64_extra(pkl, 2 ** 64 - 8),
        "huge_payload.ot": rebuild_with_zip64_extra(payload, 2 ** 40),
        # (b) a ZIP64 record whose length field reaches past the entry's extra field
        "long_record.ot": rebuild_with_zip64_extra(payload, 1234, record_size=0xFFF0),
    }
    # (c) local-header offset pointing at the last bytes of the file
    b = bytearray(good)
    struct.pack_into("<I", b, payload[0] + 42, len(b) - 8)
    cases["late_local_header.ot"] = b
    # central-directory offset / size that wrap
    b = bytearray(good)
    struct.pack_into("<I", b, eocd + 16, 0xFFFFFFF0)
    if bytes(b[eocd - 20:eocd - 16]) == b"PK\x06\x07":
        e64, = struct.unpack_from("<Q", b, eocd - 12)
        struct.pack_into("<Q", b, e64 + 48, 2 ** 64 - 16)
    cases["cd_offset.ot"] = b
    for name, blob in cases.items():
        path = write(name, blob)
        if name == "long_record.ot":      # the record is clamped to the extra field; 1234 bytes is then simply too short or enough
            try:
                pkg.ot_inspect(path)
            except pkg.B2IError as e:
                assert e.code == -2
            continue
        with pytest.raises(pkg.B2IError) as e:
            pkg.ot_inspect(path)
            pytest.fail("%s was accepted" % name)
        assert e.value.code == -2, name
        assert "zip64 end record" not in str(e.value), "%s: the crafted field was never reached (%s)" % (name, e.value)


def test_mutated_archives_never_crash_the_reader(pkg, tmp_path):
    """300 seeded single-field corruptions of the central directory and of data.pkl: each either still
    parses or fails with an error code."""
    good = open(os.path.join(GOLDEN, "tiny_libtorch.ot"), "rb").read()
    eocd, entries = _central_directory(good)
    pkl = [e for e in entries if e[1].endswith("data.pkl")][0]
    pkl_at = good.find(b"\x80\x02")
    rng = np.random.default_rng(11)
    path = str(tmp_path / "m.ot")
    regions = [(entries[0][0], eocd + 22), (pkl_at, pkl_at + 600)]
    for i in range(300):
        b = bytearray(good)
        lo, hi = regions[i % 2]
        for _ in range(int(rng.integers(1, 4))):
            at = int(rng.integers(lo, hi))
            b[at] = int(rng.integers(0, 256))
        with open(path, "wb") as f:
            f.write(b)
        try:
            pkg.ot_inspect(path)
        except pkg.B2IError as e:
            assert e.code in (-1, -2)
    assert pkl[1].endswith("data.pkl")
