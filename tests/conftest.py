import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def built():
    """Build libksareplay.so / the oracle if stale (nvcc cross-compiles without a GPU)."""
    import __graft_entry__ as g

    g.build()
    return g


@pytest.fixture(scope="session")
def engine(built):
    import torch

    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import kube_sqs_autoscaler_b200 as ksa

    return ksa.ReplayEngine(device=0)
