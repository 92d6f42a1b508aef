"""bench.py on a box without a GPU: the reference arm (CPU restatement of the path) prints the contract's JSON line with
the same config as the product arm, ranks other than 0 stay silent, and the product arm refuses to run without the
CUDA library's device instead of falling back to anything."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(args, env=None, timeout=600):
    e = dict(os.environ)
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        e.pop(k, None)
    e.update(env or {})
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py")] + args, env=e, capture_output=True, text=True,
                          timeout=timeout, cwd=ROOT)


def test_reference_arm_line_and_config():
    p = _run(["--impl", "reference", "--steps", "1", "--warmup", "1", "--batch", "8"])
    assert p.returncode == 0, p.stderr[-400:]
    lines = [ln for ln in p.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "images/s" and d["higher_is_better"] is True
    assert d["value"] > 0 and d["steps"] == 1 and d["n_gpus"] == 1 and d["gpu_launches"] == 0
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1 and d["cpu_baseline"]["value"] == d["value"]
    assert d["e2e"] == {"value": d["value"], "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    sys.path.insert(0, ROOT)
    import bench
    # the product arm builds its config with the same function and arguments
    kind = "pretrained" if "pretrained weights" in d["config"]["workload"] else "seeded synthetic"
    assert d["config"] == bench.make_config("resnet18", 8, None, kind)
    assert d["metric"] == bench.metric_name("resnet18", 8)


def test_reference_arm_other_ranks_print_nothing():
    p = _run(["--impl", "reference", "--steps", "1", "--warmup", "1", "--batch", "8", "--gpus", "2"],
             env={"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"})
    assert p.returncode == 0, p.stderr[-400:]
    assert not [ln for ln in p.stdout.splitlines() if ln.startswith("{")]


def test_product_arm_fails_loudly_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("this box has a GPU")
    p = _run(["--steps", "1", "--warmup", "1", "--no-extras"], timeout=300)
    assert p.returncode != 0
    assert not [ln for ln in p.stdout.splitlines() if ln.startswith("{")]
