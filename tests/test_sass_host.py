"""What the compiled kernels look like (cuobjdump on the objects build() leaves in-tree; no GPU needed).

The round's largest gain came from the MMA-issuing warps: unless the compiler knows that the role indices are the same in
every lane it moves each tcgen05.mma operand into uniform registers through an ELECT / R2UR.BROADCAST chain, ~20
instructions per MMA at ~4 cycles each, which capped every tensor-core kernel of the library.  This test keeps the issue
loops lean: few instructions between consecutive UTCHMMA in the kernels that matter, and it checks that the kernels are
tcgen05 / TMA code at all (UTCHMMA, UTMALDG present, no mma.sync-class HMMA in the conv kernels)."""
import os
import re
import shutil
import statistics
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BUILD = os.path.join(ROOT, "distributed-machine-learning-cluster_b200", "build")


def _sass(obj):
    exe = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    path = os.path.join(BUILD, obj)
    if not os.path.exists(exe) or not os.path.exists(path):
        pytest.skip("cuobjdump or %s not available (run __graft_entry__.build() first)" % obj)
    return subprocess.run([exe, "-sass", path], capture_output=True, text=True, timeout=300).stdout


def _functions(sass):
    """name -> list of (address, instruction text)"""
    out, name = {}, None
    for ln in sass.splitlines():
        m = re.search(r"Function : (\S+)", ln)
        if m:
            name = m.group(1)
            out[name] = []
            continue
        m = re.search(r"/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
        if m and name:
            out[name].append((int(m.group(1), 16), m.group(2)))
    return out


def _mma_gaps(instrs):
    addrs = [a for a, t in instrs if "UTCHMMA" in t]
    return [(b - a) // 16 for a, b in zip(addrs, addrs[1:])]


@pytest.fixture(scope="module")
def conv_functions():
    return _functions(_sass("conv_umma.o"))


@pytest.mark.parametrize("needle,limit", [("conv_umma_pair_kernel", 8), ("conv3x3_c64_patch_kernel", 10),
                                          ("conv3x3_flat_pair_kernel", 8), ("conv_umma_persistent_kernel", 6)])
def test_mma_issue_loops_are_lean(conv_functions, needle, limit):
    hits = {n: f for n, f in conv_functions.items() if needle in n}
    assert hits, needle
    for name, instrs in hits.items():
        gaps = _mma_gaps(instrs)
        assert gaps, name
        assert statistics.median(gaps) <= limit, (name, statistics.median(gaps), "instructions between consecutive MMAs")


def test_conv_kernels_are_tcgen05_and_tma(conv_functions):
    text = "\n".join(t for f in conv_functions.values() for _, t in f)
    assert "UTCHMMA" in text and "UTMALDG" in text and "LDTM" in text      # tcgen05.mma, TMA loads, tcgen05.ld
    assert "UTMALDG.4D.IM2COL" in text                                      # im2col-mode TMA
    assert not re.search(r"\bHMMA\b", text)                                 # no mma.sync / wmma path in the conv kernels


@pytest.mark.parametrize("obj", ["stem.o", "stem2.o", "stem_alex.o"])
def test_stem_kernels_issue_leanly(obj):
    fns = _functions(_sass(obj))
    checked = 0
    for name, instrs in fns.items():
        gaps = _mma_gaps(instrs)
        if gaps:
            assert statistics.median(gaps) <= 10, (name, statistics.median(gaps))
            checked += 1
    assert checked
