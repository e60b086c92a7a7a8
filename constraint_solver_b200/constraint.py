"""Host-side mirror of `Constraint` (src/constraint.rs:6-37), extended to two bodies (rigid_b)."""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

from . import _abi


@dataclass
class Constraint:
    rigid: int = 0
    contacts: tuple = field(default_factory=lambda: (np.zeros(3), np.zeros(3)))
    distance: float = 0.0
    rigid_b: int = -1  # < 0: the world (the reference's single-body form)

    def current_distance(self) -> float:  # src/constraint.rs:21
        d = np.asarray(self.contacts[1], dtype=np.float64) - np.asarray(self.contacts[0], dtype=np.float64)
        return float(np.sqrt((d[0] * d[0] + d[1] * d[1]) + d[2] * d[2]))

    def as_record(self) -> np.ndarray:
        r = np.zeros((), dtype=_abi.CONSTRAINT_DTYPE)
        r["rigid"], r["rigid_b"] = self.rigid, self.rigid_b
        r["contact0"], r["contact1"] = self.contacts[0], self.contacts[1]
        r["distance"] = self.distance
        return r

    @staticmethod
    def from_record(r) -> "Constraint":
        return Constraint(int(r["rigid"]), (np.array(r["contact0"]), np.array(r["contact1"])), float(r["distance"]),
                          int(r["rigid_b"]))
