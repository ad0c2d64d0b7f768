import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
if os.path.join(ROOT, "tests") not in sys.path:
    sys.path.insert(1, os.path.join(ROOT, "tests"))          # test modules import helpers from each other by bare name


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def golden():
    import json
    import numpy as np
    g = os.path.join(ROOT, "tests", "golden")
    with open(os.path.join(g, "manifest.json")) as f:
        manifest = json.load(f)
    return manifest["cases"], np.load(os.path.join(g, "small_cases.npz"))


@pytest.fixture(scope="session")
def enhancer():
    """One context for the whole GPU session (vex_create is not free)."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from video_enhancer_proto_b200 import B200VideoEnhancer
    e = B200VideoEnhancer(device="cuda:0", seed=0)
    e._ensure_ctx()
    yield e
    e.close()
