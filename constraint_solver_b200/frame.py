"""Host-side mirror of `Frame` (src/frame.rs:7-10): position + rotation (cgmath order v.x v.y v.z s)."""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

from . import _abi


@dataclass
class Frame:
    position: np.ndarray = field(default_factory=lambda: np.zeros(3))
    rotation: np.ndarray = field(default_factory=lambda: np.array([0.0, 0.0, 0.0, 1.0]))  # v.x v.y v.z s

    def as_record(self) -> np.ndarray:
        r = np.zeros((), dtype=_abi.FRAME_DTYPE)
        r["position"] = self.position
        r["rotation"] = self.rotation
        return r

    @staticmethod
    def from_record(r) -> "Frame":
        return Frame(np.array(r["position"], dtype=np.float64), np.array(r["rotation"], dtype=np.float64))
