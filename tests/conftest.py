import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA GPU (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden_traj():
    import numpy as np
    return np.load(os.path.join(GOLDEN, "ref_traj.npz"))


@pytest.fixture(scope="session")
def golden_field():
    import numpy as np
    return np.load(os.path.join(GOLDEN, "ref_field.npz"))


@pytest.fixture(scope="session")
def golden_sweep():
    import numpy as np
    return np.load(os.path.join(GOLDEN, "ref_sweep.npz"))


@pytest.fixture(scope="session")
def sensitivity():
    import json
    return json.load(open(os.path.join(GOLDEN, "sensitivity.json")))


@pytest.fixture(scope="session")
def engine():
    import raytracing_b200 as rt
    return rt.get_engine(0)
