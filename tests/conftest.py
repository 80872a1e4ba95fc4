import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture
def cpu_extend(monkeypatch):
    """Lets host-logic tests build ExtendedSystems without a GPU: the oracle's
    extend_matrices stands in for the K1 kernel (CPU tests only)."""
    from oracle import qp_oracle
    import mpc_interface_b200.tools as tools

    monkeypatch.setattr(tools, "extend_matrices", qp_oracle.extend_matrices)
    return qp_oracle.extend_matrices
