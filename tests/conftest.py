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


@pytest.fixture(autouse=True)
def _single_kernel_path_only_where_asked(request, monkeypatch):
    """Frames of at most 2,048 lights on an unsharded context run as ONE kernel by default (csrc/hvcf_small.cu).  The parity tests
    of the hierarchical kernels use many such small frames on purpose (0, 1, 31, 64, 1,025 lights ...), so the switch is OFF
    unless a test is parametrized with mode "small" (tests/test_gpu_parity.py: MODES) or sets HVCF_SMALL itself."""
    mode = getattr(getattr(request.node, "callspec", None), "params", {}).get("mode")
    monkeypatch.setenv("HVCF_SMALL", "1" if mode == "small" else "0")
