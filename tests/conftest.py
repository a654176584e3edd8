import os, sys
import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def rel_err(a, b):
    a = np.asarray(a, dtype=np.float64); b = np.asarray(b, dtype=np.float64)
    if a.size == 0:
        return 0.0
    return float(np.abs(a - b).max() / max(1.0, np.abs(a).max()))


def load_golden(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))


def golden_cases():
    sys.path.insert(0, GOLDEN)
    import make_golden
    return make_golden.CASES, make_golden.checksum


@pytest.fixture(scope="session")
def oracle_mod():
    import oracle
    oracle.lib()          # builds the C restatement on first use
    return oracle


@pytest.fixture(scope="session")
def cuda_lib():
    import qpdo_b200
    return qpdo_b200.lib()
