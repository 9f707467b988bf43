import json
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for p in (ROOT, HERE):
    if p not in sys.path:
        sys.path.insert(0, p)

import __graft_entry__ as entry  # noqa: E402


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def _cuda_devices():
    try:
        gs = entry.load_package()
        n = gs.load().gs_device_count()
        return max(n, 0)
    except Exception:
        return 0


def pytest_collection_modifyitems(config, items):
    if _cuda_devices() > 0:
        return
    skip = pytest.mark.skip(reason="no CUDA device visible")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def gs():
    """the product package with its C-ABI library built and bound (no compute without a GPU)"""
    mod = entry.load_package()
    if not os.path.exists(mod.LIB):
        mod.build()
    mod.load()
    return mod


@pytest.fixture(scope="session")
def oracle_mod():
    orc = entry.load_oracle()
    if not orc.available("port"):
        orc.build(ref=False)
    return orc


@pytest.fixture(scope="session")
def golden():
    with open(os.path.join(HERE, "golden", "golden.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def golden_fields():
    return np.load(os.path.join(HERE, "golden", "fields.npz"))
