"""Shared test configuration.

`-m "not gpu"` : oracle vs golden vectors / compiled reference, host logic, C-ABI symbols.
`-m gpu`       : parity of the CUDA path (through the C-ABI) against the oracle.
"""
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
PKG_DIR = ROOT / "stereo-matching_b200"
for p in (str(ROOT), str(PKG_DIR)):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = Path(__file__).resolve().parent / "golden"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def golden():
    import numpy as np

    def load(name):
        return dict(np.load(GOLDEN / f"{name}.npz"))
    return load
