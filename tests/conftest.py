import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden_dir():
    return os.path.join(ROOT, "tests", "golden")


@pytest.fixture(autouse=True)
def _fresh_operator_capacity_hints(request):
    """rasterize_gaussians_sum sizes its record buffer from the intersection counts it has seen for an
    image size; a test must not inherit (or leave behind) the hints and unread mail of another one"""
    yield
    if request.node.get_closest_marker("gpu") is None:
        return
    import sys
    rs = sys.modules.get("gsvc_b200.rasterize_sum")
    if rs is None:
        return
    import torch
    torch.cuda.synchronize()
    for dctx in rs._ctx_by_device.values():
        dctx.pending.clear()
        dctx.cap.clear()
        dctx.cap_n.clear()
