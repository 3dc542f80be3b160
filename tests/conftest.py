import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

GOLDEN = ROOT / "tests" / "golden"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    """GPU tests fail loudly (not skip) when selected on a machine without CUDA, except that
    the default CPU run deselects them via -m "not gpu"."""
    return


@pytest.fixture(scope="session")
def golden():
    import numpy as np

    return {p.stem: np.load(p) for p in GOLDEN.glob("*.npz")}
