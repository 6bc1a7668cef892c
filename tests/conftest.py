import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def port():
    import bb_oracle
    return bb_oracle.Port()


@pytest.fixture(scope="session")
def ref():
    import bb_oracle
    if not bb_oracle.ref_available() and not os.path.isdir("/root/reference/Code/C"):
        pytest.skip("oracle/_ref/libbb_ref.so not built (reference sources absent)")
    return bb_oracle.Ref()
