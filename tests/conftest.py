import json
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

DATA = os.path.join(ROOT, "data", "lego_rust")
CAM_JSON = os.path.join(DATA, "tf_reference_samples.json")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (sm_100a) GPU; run with -m gpu on the GPU box")


def _have_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _have_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def golden():
    with open(CAM_JSON) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def onets():
    """The CPU oracle's two networks (test infrastructure)."""
    from oracle import pyoracle as O
    return O.Network(os.path.join(DATA, "coarse")), O.Network(os.path.join(DATA, "fine"))


@pytest.fixture(scope="session")
def ocam():
    from oracle import pyoracle as O
    return O.camera_from_json(CAM_JSON)


@pytest.fixture(scope="session")
def renderer():
    """The product: libnerf_b200.so on cuda:0 (GPU tests only)."""
    import nerf_rs_b200 as nb
    r = nb.Renderer(DATA, device=0)
    yield r
    r.close()


@pytest.fixture(scope="session")
def cam():
    import nerf_rs_b200 as nb
    return nb.camera_from_json(CAM_JSON)
