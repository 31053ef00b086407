import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def port():
    from oracle import bindings as ob
    ob.build(port=True, ref=False)
    return ob.Port()


@pytest.fixture(scope="session")
def cv2_golden():
    return np.load(os.path.join(GOLDEN, "cv2_steps.npz"))


@pytest.fixture(scope="session")
def gradmag_golden():
    return np.load(os.path.join(GOLDEN, "gradmag.npz"))


@pytest.fixture(scope="session")
def ref_golden():
    return np.load(os.path.join(GOLDEN, "ref_flows.npz"))


@pytest.fixture(scope="session")
def dis():
    """The product bindings; the library must already be built (python -m of_dis_b200.build)."""
    import of_dis_b200 as d
    d.lib()
    return d


@pytest.fixture(scope="session")
def gpu(dis):
    if dis.device_count() < 1:
        pytest.fail("gpu-marked test on a machine without a CUDA device (the product has no CPU path)")
    return dis
