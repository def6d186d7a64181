"""pytest configuration: registers the `gpu` marker and shared fixtures/helpers."""
import sys
from pathlib import Path

import numpy as np
import pytest

REPO = Path(__file__).resolve().parents[1]
if str(REPO) not in sys.path:
    sys.path.insert(0, str(REPO))
GOLDEN = REPO / "tests" / "golden"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def split_blob(blob: np.ndarray):
    return bytes(blob.tobytes()).decode().split("\x00")


@pytest.fixture(scope="session")
def binary_golden():
    return np.load(GOLDEN / "binary_golden.npz")


@pytest.fixture(scope="session")
def formulation_golden():
    return np.load(GOLDEN / "formulation_golden.npz")


@pytest.fixture(scope="session")
def ibmpg1_lines():
    z = np.load(GOLDEN / "ibmpg1_lines.npz")
    return list(z["names"]), split_blob(z["csv_blob"])


@pytest.fixture(scope="session")
def oracle():
    from oracle import restatement
    restatement.load()
    return restatement
