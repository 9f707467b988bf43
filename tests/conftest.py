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


def _gpu_count():
    """NVIDIA GPUs on this machine (nvidia-smi -L).  Deliberately NOT derived from loading the product library: on a GPU
    box a broken build or an unloadable .so must FAIL the GPU suite, not skip it."""
    import shutil
    import subprocess
    exe = shutil.which("nvidia-smi")
    if not exe:
        return 0
    try:
        r = subprocess.run([exe, "-L"], capture_output=True, text=True, timeout=60)
    except Exception:
        return 0
    return sum(1 for line in r.stdout.splitlines() if line.startswith("GPU ")) if r.returncode == 0 else 0


def pytest_collection_modifyitems(config, items):
    n = _gpu_count()
    if n == 0:
        skip = pytest.mark.skip(reason="no NVIDIA GPU on this machine (nvidia-smi)")
        for item in items:
            if "gpu" in item.keywords:
                item.add_marker(skip)
        return
    # multi-GPU cases that this box cannot hold are not part of its suite: deselect them (a skip would read as a test
    # that did not run for want of something fixable)
    keep, drop = [], []
    for item in items:
        need = getattr(item, "callspec", None) and item.callspec.params.get("world")
        if item.name.startswith("test_two_devices_in_one_process"):
            need = 2
        (drop if (isinstance(need, int) and "test_multigpu" in item.nodeid and need > n) else keep).append(item)
    if drop:
        config.hook.pytest_deselected(items=drop)
        items[:] = keep


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
