import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")
CACHE = os.path.join(ROOT, "tests", "_cache")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def pkg():
    import dmlc_b200
    return dmlc_b200


@pytest.fixture(scope="session")
def ctx(pkg):
    """One-GPU context on cuda:0; fails (does not skip) when there is no GPU: no fallback exists."""
    c = pkg.Context((0,))
    yield c
    c.close()


def ot_path(model_name):
    """Seeded synthetic weights in genuine `.ot` format, written once per checkout."""
    from oracle import synth
    return synth.ensure_ot(model_name, os.path.join(CACHE, model_name + ".ot"))


@pytest.fixture(scope="session")
def resnet18_ot():
    return ot_path("resnet18")


@pytest.fixture(scope="session")
def alexnet_ot():
    return ot_path("alexnet")


@pytest.fixture(scope="session")
def resnet18_ps_ot():
    """The synthetic ResNet-18 with its classifier scaled to pretrained-like logits (std ~2.5)."""
    return ot_path("resnet18_ps")


def fixture_paths():
    """[(wnid, path)] of the committed fixture JPEGs, in the order of golden_vectors.npz's rows."""
    import numpy as np
    gv = np.load(os.path.join(GOLDEN, "golden_vectors.npz"))
    root = os.path.join(GOLDEN, "imagenet_1k", "train")
    return [(str(w), os.path.join(root, str(w), sorted(os.listdir(os.path.join(root, str(w))))[0])) for w in gv["wnids"]]


_CROPS = {}


def fixture_crops():
    """u8 [n,224,224,3]: the oracle's decode + resize + centre crop of every committed fixture (computed once
    per session by a process pool: the numpy restatement takes ~80 ms per image)."""
    import numpy as np
    from concurrent.futures import ProcessPoolExecutor
    from oracle import stb_resize
    if "crops" not in _CROPS:
        paths = [p for _, p in fixture_paths()]
        with ProcessPoolExecutor(max_workers=min(16, os.cpu_count() or 1)) as ex:
            _CROPS["crops"] = np.stack(list(ex.map(stb_resize.load_image_and_resize, paths, chunksize=4)))
    return _CROPS["crops"]
