import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    # Build-on-demand so a fresh checkout works; normally __graft_entry__.build() has done this.
    lib = os.path.join(ROOT, "constraint_solver_b200", "lib", "libxpbd_b200.so")
    ora = os.path.join(ROOT, "oracle", "_build", "liboracle.so")
    if not os.path.exists(ora):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle")])
    if not os.path.exists(lib):
        subprocess.check_call(["make", "-C", ROOT, "lib"])


def _has_gpu() -> bool:
    from constraint_solver_b200 import _abi
    import ctypes
    n = ctypes.c_int32(0)
    return _abi.load().xpbd_device_count(ctypes.byref(n)) == 0 and n.value > 0


def pytest_collection_modifyitems(config, items):
    # `-m gpu` on a box without a device should fail loudly rather than skip silently; only skip when
    # the gpu tests were not asked for explicitly.
    if "gpu" in (config.getoption("-m") or ""):
        return
    if _has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)
