import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def built_extension():
    """Build (incrementally) and load the in-tree sm_100a extension; tests fail -- not skip -- if it cannot load."""
    import nm_vllm_b200
    if not nm_vllm_b200.is_built():
        from nm_vllm_b200 import build
        build.build_all()
    nm_vllm_b200.load_extension()
    return nm_vllm_b200
