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
