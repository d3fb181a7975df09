import os
import sys

import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "slow: CPU test that takes more than a few seconds")


@pytest.fixture(scope="session")
def oracle():
    from cpu_checkers import Oracle
    return Oracle()


@pytest.fixture(scope="session")
def golden():
    import json
    import numpy as np
    here = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    with open(os.path.join(here, "golden_ref.json")) as f:
        meta = json.load(f)
    return meta, np.load(os.path.join(here, "golden_ref.npz"))
