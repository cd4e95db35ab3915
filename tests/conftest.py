import json
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

GOLDEN_DIR = Path(__file__).resolve().parent / "golden"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with `-m gpu`)")


@pytest.fixture(scope="session")
def golden():
    return json.loads((GOLDEN_DIR / "golden.json").read_text())


@pytest.fixture(scope="session")
def wo():
    """The CPU oracle (test infrastructure): oracle/wharf_oracle.py over oracle/wharf_oracle.c."""
    from oracle import wharf_oracle
    wharf_oracle.build()
    return wharf_oracle


@pytest.fixture(scope="session")
def cora():
    z = np.load(GOLDEN_DIR / "cora_graph.npz")
    return int(z["offsets"].shape[0]), z["offsets"], z["edges"]


@pytest.fixture(scope="session")
def wharf():
    """The product package; the CUDA library must be present and a GPU usable."""
    import dynamicgraphrepresentationlearning_b200 as pkg
    return pkg


SIX_OFFSETS = np.array([0, 2, 5, 10, 13, 15], dtype=np.uint64)
SIX_EDGES = np.array([1, 2, 0, 2, 3, 0, 1, 3, 4, 5, 1, 2, 5, 2, 5, 2, 3, 4], dtype=np.uint32)
