import os
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    # `-m gpu` on a box without a device must fail loudly, not silently skip: the product
    # path has no CPU fallback.  Without -m gpu (CPU suite) gpu tests are deselected by -m.
    pass


@pytest.fixture(scope="session")
def oracle():
    from oracle import orb_oracle
    orb_oracle.build()
    return orb_oracle
