import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    config.addinivalue_line("markers", "slow: larger inputs")


def _has_gpu():
    try:
        import mhm2_kmer_analysis_v2_b200._lib as L

        return L.lib().kc_device_count() > 0
    except Exception:
        return False


HAS_GPU = _has_gpu()


def pytest_collection_modifyitems(config, items):
    # a GPU test on a box without a GPU is an error of the invocation, not a skip -- but the CPU run selects with
    # -m "not gpu", so only guard against silently passing
    if HAS_GPU:
        return
    skip = pytest.mark.skip(reason="no CUDA device here; run with -m gpu on the B200 box")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def oracle_mod():
    import oracle as O

    O.lib()
    return O
