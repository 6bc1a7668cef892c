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


def pytest_sessionfinish(session, exitstatus):
    """achieved parity errors of this session -> gpurun_out/parity_errors.json (merged with what is there)"""
    import json
    try:
        from helpers import PARITY
    except ImportError:
        return
    if not PARITY:
        return
    out = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out, exist_ok=True)
    path = os.path.join(out, "parity_errors.json")
    rows = {}
    if os.path.exists(path):
        try:
            rows = {(r["row"], r["config"], r["against"]): r for r in json.load(open(path))}
        except (ValueError, KeyError):
            rows = {}
    for (row, config, against), v in PARITY.items():
        rows[(row, config, against)] = {"row": row, "config": config, "against": against, **v}   # the latest session wins
    json.dump(sorted(rows.values(), key=lambda r: (r["row"], r["config"], r["against"])), open(path, "w"), indent=1)
