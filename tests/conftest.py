import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

import mcrt_b200  # noqa: E402,F401  registers the hyphenated package directory


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def ref_oracle():
    """The compiled reference (oracle/_ref/libmcrt_ref.so); skipped where it was never built."""
    from oracle import ref
    if not ref.available():
        pytest.skip("oracle/_ref/libmcrt_ref.so not built (needs /root/reference at build time)")
    return ref.RefOracle()


@pytest.fixture(scope="session")
def cornell():
    from mcrt_b200 import scene_library
    return scene_library.cornell_box()
