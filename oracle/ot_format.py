"""`.ot` weight archives (reference pretrained_models/*.ot, loaded at src/services.rs:516,522).

tch writes these with torch::serialize::OutputArchive (a TorchScript module whose parameters are
the named tensors).  `write_ot` emits the same container byte layout in pure
Python (ZIP, STORED entries, 64-byte aligned payloads, protocol-2 pickle with persistent storage
ids) so fixtures can be produced anywhere; `read_ot` reads an archive back through libtorch's own
reader (`torch.jit.load`), the same reader family tch uses, which is what pins the writer.
"""
import io
import struct
import zipfile

import numpy as np
import torch


def _binunicode(s: str) -> bytes:
    b = s.encode("utf-8")
    return b"X" + struct.pack("<I", len(b)) + b


def _int(v: int) -> bytes:
    if 0 <= v < 256:
        return b"K" + struct.pack("<B", v)
    if 0 <= v < 65536:
        return b"M" + struct.pack("<H", v)
    return b"J" + struct.pack("<i", v)


def _tuple_of_ints(vals) -> bytes:
    return b"(" + b"".join(_int(int(v)) for v in vals) + b"t"


def _data_pkl(named) -> bytes:
    """Protocol-2 pickle shaped like libtorch's: Module() state dict of _rebuild_tensor_v2 calls."""
    out = io.BytesIO()
    memo = 0

    def put():
        nonlocal memo
        b = b"q" + struct.pack("<B", memo) if memo < 256 else b"r" + struct.pack("<I", memo)
        memo += 1
        return b

    def get(i):
        return b"h" + struct.pack("<B", i) if i < 256 else b"j" + struct.pack("<I", i)

    out.write(b"\x80\x02")
    out.write(b"c__torch__\nModule\n" + put())  # memo 0
    out.write(b")\x81}(")
    ids = {}
    for k, (name, arr) in enumerate(named):
        out.write(_binunicode(name) + put())
        if "rebuild" not in ids:
            out.write(b"ctorch._utils\n_rebuild_tensor_v2\n")
            ids["rebuild"] = memo
            out.write(put())
        else:
            out.write(get(ids["rebuild"]))
        out.write(b"((")
        for key, blob in (("storage", _binunicode("storage")), ("ftype", b"ctorch\nFloatStorage\n")):
            if key not in ids:
                out.write(blob)
                ids[key] = memo
                out.write(put())
            else:
                out.write(get(ids[key]))
        out.write(_binunicode(str(k)) + put())
        if "cpu" not in ids:
            out.write(_binunicode("cpu"))
            ids["cpu"] = memo
            out.write(put())
        else:
            out.write(get(ids["cpu"]))
        out.write(_int(arr.size) + b"tQ" + put())
        out.write(_int(0))
        out.write(_tuple_of_ints(arr.shape))
        strides = []
        acc = 1
        for d in reversed(arr.shape):
            strides.append(acc)
            acc *= d
        out.write(_tuple_of_ints(reversed(strides)))
        out.write(b"\x89")
        if "odict" not in ids:
            out.write(b"ccollections\nOrderedDict\n")
            ids["odict"] = memo
            out.write(put())
        else:
            out.write(get(ids["odict"]))
        out.write(b")RtR")
    out.write(b"ub" + put() + b".")
    return out.getvalue()


def write_ot(path, named_tensors, stem=None):
    """named_tensors: ordered mapping / list of (name, float32 array-like). Names may use '.' or '|';
    they are stored with '|' like tch does."""
    items = list(named_tensors.items()) if hasattr(named_tensors, "items") else list(named_tensors)
    named = []
    for name, t in items:
        a = t.detach().cpu().numpy() if isinstance(t, torch.Tensor) else np.asarray(t)
        named.append((name.replace(".", "|"), np.ascontiguousarray(a, dtype="<f4")))
    if stem is None:
        stem = str(path).rsplit("/", 1)[-1].rsplit(".", 1)[0]
    code = "class Module(Module):\n  __parameters__ = [" + "".join('"%s", ' % n for n, _ in named) + "]\n"
    code += "  __buffers__ = []\n  __annotations__ = []\n"
    code += "".join('  __annotations__["%s"] = Tensor\n' % n for n, _ in named)
    with open(path, "wb") as f:
        central = []

        def add(name, payload, align=False):
            name_b = (stem + "/" + name).encode()
            off = f.tell()
            extra = b""
            if align:
                # pad the local extra field so the payload starts on a 64-byte boundary
                base = off + 30 + len(name_b)
                pad = (-(base + 4)) % 64
                extra = struct.pack("<HH", 0x4246, pad) + b"Z" * pad
            crc = zipfile.crc32(payload) & 0xFFFFFFFF
            hdr = struct.pack("<IHHHHHIIIHH", 0x04034B50, 20, 0x0800, 0, 0, 0x21, crc, len(payload),
                              len(payload), len(name_b), len(extra))
            f.write(hdr + name_b + extra + payload)
            central.append((name_b, crc, len(payload), off))

        for k, (_, arr) in enumerate(named):
            add("data/%d" % k, arr.tobytes(), align=True)
        add("data.pkl", _data_pkl(named))
        add("code/__torch__.py", code.encode())
        add("constants.pkl", b"\x80\x02).")
        add("version", b"3\n")
        add("byteorder", b"little")
        cd_off = f.tell()
        for name_b, crc, size, off in central:
            if size >= 0xFFFFFFFF or off >= 0xFFFFFFFF:
                raise ValueError("archive too large for the non-ZIP64 writer")
            f.write(struct.pack("<IHHHHHHIIIHHHHHII", 0x02014B50, 20, 20, 0x0800, 0, 0, 0x21, crc, size, size,
                                len(name_b), 0, 0, 0, 0, 0, off) + name_b)
        cd_size = f.tell() - cd_off
        f.write(struct.pack("<IHHHHIIH", 0x06054B50, 0, 0, len(central), len(central), cd_size, cd_off, 0))


def read_ot(path):
    """name ('.'-separated) -> float32 torch tensor, through libtorch's archive reader."""
    m = torch.jit.load(str(path), map_location="cpu")
    out = {}
    for n, p in list(m.named_parameters()) + list(m.named_buffers()):
        out[n.replace("|", ".")] = p.detach().to(torch.float32).contiguous()
    if not out:
        raise ValueError("no tensors in %s" % path)
    return out
