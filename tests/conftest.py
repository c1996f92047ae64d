import os
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Build the checkers and the product library if they are not there yet."""
    from oracle import build as obuild

    obuild.build_c_oracle()
    from flashgmm_b200 import build as pbuild

    if not pbuild.LIB.exists():
        pbuild.build(verbose=False)
    yield


def have_gpu():
    try:
        import torch

        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.fixture(scope="session")
def ref_available():
    from oracle import ref

    return ref.available()
