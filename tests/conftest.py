import os
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def oracle():
    from oracle import loader

    return loader.load_oracle()


@pytest.fixture(scope="session")
def reference():
    from oracle import loader

    lib = loader.load_reference()
    if lib is None:
        pytest.skip("oracle/_ref/libpinyon_ref.so not built (needs /root/reference)")
    return lib


@pytest.fixture(scope="session")
def ctx():
    from pinyon_b200 import Context

    c = Context(0)  # raises without an sm_100 GPU: there is no CPU fallback
    yield c
    c.close()
