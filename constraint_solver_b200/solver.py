"""Host-side mirror of the reference's `solver` module (src/solver.rs)."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _abi
from .constraint import Constraint
from .geometry import Polytope
from .rigid import Rigid

COMPLIANCE = 1e-6  # src/solver.rs:20


def step(rigid: Rigid, polytope: Polytope, dt: float, substep_count: int) -> None:
    """solver::step (src/solver.rs:3-17): all substeps fused in one CUDA launch."""
    c = polytope.as_c()
    _abi.check(_abi.load().xpbd_solver_step(C.c_void_p(rigid.rec.ctypes.data), C.byref(c), float(dt), int(substep_count)))


def solve(rigid: Rigid, constraints: list[Constraint], dt: float) -> None:
    """solver::solve (src/solver.rs:19-27): one Gauss-Seidel sweep in list order."""
    n = len(constraints)
    recs = np.zeros(max(n, 1), dtype=_abi.CONSTRAINT_DTYPE)
    for i, c in enumerate(constraints):
        recs[i] = c.as_record()
    counts = np.array([n], dtype=np.int32)
    _abi.check(_abi.load().xpbd_solve_batch(C.c_void_p(rigid.rec.ctypes.data), 1, _abi.ptr(recs), _abi.ptr(counts),
                                            max(n, 1), float(dt), COMPLIANCE))
