"""pytest configuration: the `gpu` marker (tests that need a real B200) and import paths."""
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

GOLDEN = Path(__file__).resolve().parent / "golden"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN


@pytest.fixture(scope="session")
def built_library():
    """The in-tree C-ABI library, (re)built if a source is newer (nvcc cross-compiles on CPU)."""
    from cid_b200.build import build_library

    return build_library()
