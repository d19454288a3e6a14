import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN


@pytest.fixture(autouse=True)
def _device_buffer_guards():
    """With MMD_B200_GUARD=1 (tools/gpu_guard.sh) every test ends with a check of the guard bands around the
    library's device buffers; a no-op otherwise."""
    yield
    if os.environ.get("MMD_B200_GUARD", "0") not in ("", "0"):
        from multimodal_decoding_b200.engine import Engine
        for eng in list(Engine.live):
            eng.check_guards()
