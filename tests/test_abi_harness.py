"""tests/abi_harness.c: the header as plain C11 (-Wall -Werror), record sizes / offsets as rust/src/ffi.rs assumes, and the
create -> add_hull -> upload -> step -> download round trip driven from C the way a maintainer's FFI would."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIBDIR = os.path.join(ROOT, "constraint_solver_b200", "lib")


def _build(tmp_path):
    exe = str(tmp_path / "abi_harness")
    subprocess.check_call(["gcc", "-std=c11", "-Wall", "-Werror", "-pedantic", "-I" + os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tests", "abi_harness.c"), "-o", exe, "-L" + LIBDIR, "-lxpbd_b200",
                           "-Wl,-rpath," + LIBDIR])
    return exe


def test_header_is_c11_and_layouts_match(tmp_path):
    exe = _build(tmp_path)
    out = subprocess.run([exe, "layout"], capture_output=True, text=True)
    assert out.returncode == 0 and "layout ok" in out.stdout, out.stderr


def test_c_caller_without_device_is_refused_loudly(tmp_path):
    import ctypes as C
    from constraint_solver_b200 import _abi
    n = C.c_int32(0)
    if _abi.load().xpbd_device_count(C.byref(n)) == 0 and n.value > 0:
        pytest.skip("a CUDA device is present")
    out = subprocess.run([_build(tmp_path), "run"], capture_output=True, text=True)
    assert out.returncode == 0 and "no device" in out.stdout, out.stderr


@pytest.mark.gpu
def test_c_caller_round_trip_on_the_device(tmp_path):
    out = subprocess.run([_build(tmp_path), "run"], capture_output=True, text=True)
    assert out.returncode == 0 and "run ok" in out.stdout, out.stdout + out.stderr
