"""constraint_solver_b200 — B200-native batched XPBD substep path behind the API of
jim-ec/constraint_solver (rigid.rs, constraint.rs, solver.rs, world.rs, collision.rs, geometry.rs).

All arithmetic on the step path runs in hand-written CUDA kernels (sm_100a) inside
lib/libxpbd_b200.so, reached through the C ABI in include/xpbd_b200.h.  This Python package is
the thin host-side mirror used by tests and bench.py; it has no CPU or PyTorch fallback.
"""
from . import _abi, collision, solver
from ._abi import XpbdError
from .constraint import Constraint
from .frame import Frame
from .geometry import Plane, Polytope, RigidMetrics
from .rigid import Rigid
from .world import World, Worlds, shard_range
from .driver import HeadlessDriver

__all__ = ["Constraint", "Frame", "Plane", "Polytope", "Rigid", "RigidMetrics", "World", "Worlds", "XpbdError", "HeadlessDriver", "shard_range",
           "collision", "solver", "_abi"]
