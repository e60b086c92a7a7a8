"""xpbd_solve_batch_static — the entry point that allocates nothing (its device working copies live in a static device
array inside the library; no counterpart in the reference) — against
xpbd_solve_batch, through the C ABI, bit for bit (the tolerance of every parity test here: fp32 arithmetic, identical
kernel), on calls in a row that share the array: first at both limits (XPBD_STATIC_MAX_BODIES bodies x 8 slots =
XPBD_STATIC_MAX_CONSTRAINTS), then a smaller and differently shaped batch over the stale contents, then the first again.

Named test_zz_* so that pytest collects it last; it needs no state from other tests and
passes alone in a fresh process as well as after them."""
import ctypes as C

import numpy as np
import pytest

import constraint_solver_b200 as cs
from constraint_solver_b200 import _abi, scenes

pytestmark = pytest.mark.gpu
L = _abi.load
DT = 1 / 60 / 25


def _case(n, poly, seed):
    """Tumbling bodies near the ground and their ground constraints (xpbd_ground_batch: n x n_vertices slots)."""
    rng = np.random.default_rng(seed)
    rec = scenes.c2_perturbed(n).reshape(n)
    rec["position"][:, 2] = rng.uniform(-0.3, 1.0, n)
    rec["internal_force"] = rng.uniform(-1, 1, (n, 3))
    rec["internal_torque"] = rng.uniform(-0.1, 0.1, (n, 3))
    past = np.zeros(n, dtype=_abi.FRAME_DTYPE)
    past["position"] = rec["position"] + 0.01
    past["rotation"] = scenes.random_unit_quaternions(rng, n)
    nv = len(poly.vertices)
    cons = np.zeros((n, nv), dtype=_abi.CONSTRAINT_DTYPE)
    counts = np.zeros(n, dtype=np.int32)
    c = poly.as_c()
    _abi.check(L().xpbd_ground_batch(_abi.ptr(rec), _abi.ptr(past), n, C.byref(c), _abi.ptr(cons), _abi.ptr(counts)))
    assert counts.sum() > n // 4  # the scene does produce contacts
    return rec, cons, counts, nv


def _both(rec, cons, counts, nv):
    want, got = rec.copy(), rec.copy()
    _abi.check(L().xpbd_solve_batch(_abi.ptr(want), len(rec), _abi.ptr(cons), _abi.ptr(counts), nv, DT, 1e-6))
    _abi.check(L().xpbd_solve_batch_static(_abi.ptr(got), len(rec), _abi.ptr(cons), _abi.ptr(counts), nv, DT, 1e-6))
    assert not np.array_equal(want.view(np.uint8), rec.view(np.uint8))  # the sweep moved bodies
    assert np.array_equal(got.view(np.uint8), want.view(np.uint8))      # every byte of every record
    return got


def test_static_scratch_solve_matches_the_allocating_entry_point_on_calls_in_a_row():
    big = _case(_abi.XPBD_STATIC_MAX_BODIES, cs.Polytope.new_cube(), 5)            # 4096 x 8 slots: both limits exactly
    assert big[0].shape[0] * big[3] == _abi.XPBD_STATIC_MAX_CONSTRAINTS
    small = _case(300, cs.Polytope.new_icosahedron(), 6)                            # 300 x 12 slots over stale contents
    first = _both(*big)
    _both(*small)
    again = _both(*big)
    assert np.array_equal(first.view(np.uint8), again.view(np.uint8))


def test_static_scratch_edge_sizes_and_limits():
    lib = L()
    rec, cons, counts, nv = _case(64, cs.Polytope.new_cube(), 9)
    # n = 0: nothing to do, nothing touched
    assert lib.xpbd_solve_batch_static(_abi.ptr(rec), 0, _abi.ptr(cons), _abi.ptr(counts), nv, DT, 1e-6) == _abi.XPBD_OK
    # n = 1
    _both(rec[:1].copy(), cons[:1].copy(), counts[:1].copy(), nv)
    # one body too many, and too many constraint slots: a clear error, records untouched
    n = _abi.XPBD_STATIC_MAX_BODIES + 1
    many = np.zeros(n, dtype=rec.dtype); many[:] = rec[0]
    before = many.copy()
    mc = np.zeros((n, 1), dtype=_abi.CONSTRAINT_DTYPE); mk = np.zeros(n, dtype=np.int32)
    with pytest.raises(_abi.XpbdError) as e:
        _abi.check(lib.xpbd_solve_batch_static(_abi.ptr(many), n, _abi.ptr(mc), _abi.ptr(mk), 1, DT, 1e-6))
    assert e.value.status == _abi.XPBD_ERR_CAPACITY and "at most 4096" in str(e.value)
    n, stride = 3000, 12  # 36 000 slots > 32 768
    mc = np.zeros((n, stride), dtype=_abi.CONSTRAINT_DTYPE); mk = np.zeros(n, dtype=np.int32)
    with pytest.raises(_abi.XpbdError) as e:
        _abi.check(lib.xpbd_solve_batch_static(_abi.ptr(many), n, _abi.ptr(mc), _abi.ptr(mk), stride, DT, 1e-6))
    assert e.value.status == _abi.XPBD_ERR_CAPACITY and "at most 32768" in str(e.value)
    assert np.array_equal(many.view(np.uint8), before.view(np.uint8))
    # and the array still works after the rejected calls
    _both(rec, cons, counts, nv)
