"""Headless driver + state history: what is left of the reference's `app::run` loop
(src/app.rs:201-215) once the window, input and renderer are gone.

The app keeps `states: Vec<(World, DebugLines)>` — every stepped world is pushed and can be rewound with the
Down-arrow (src/app.rs:48-49,146-148,204-215); `World: Copy` is what makes that cheap.  Here a frame's state
is the packed fp32 dynamic snapshot (13 floats per body: xpbd_batch_download_dynamic_f32), and rewinding is an
upload of an earlier snapshot (xpbd_batch_upload_dynamic_f32).
"""
from __future__ import annotations

import numpy as np

from .world import SUBSTEPS, Worlds

FRAME_TIME = 1.0 / 60.0  # src/app.rs:15
DYN_FIELDS = ("px", "py", "pz", "qx", "qy", "qz", "qs", "vx", "vy", "vz", "wx", "wy", "wz")


class HeadlessDriver:
    def __init__(self, worlds: Worlds, dt: float = FRAME_TIME, substeps: int = SUBSTEPS, history: bool = True,
                 first_world: int = 0, n_worlds: int | None = None):
        """`first_world`/`n_worlds` restrict the recorded history to a world range (the whole batch steps anyway)."""
        self.worlds, self.dt, self.substeps = worlds, float(dt), int(substeps)
        self.first_world = int(first_world)
        self.n_recorded = worlds.n_worlds - self.first_world if n_worlds is None else int(n_worlds)
        self.keep_history = bool(history)
        self.states: list[np.ndarray] = []  # states[k] = snapshot after k frames (states[0] = initial)
        self.current_state = 0
        if self.keep_history:
            self.states.append(self._snapshot())

    def _snapshot(self) -> np.ndarray:
        return self.worlds.download_dynamic(self.first_world, self.n_recorded)

    def step(self, n_frames: int = 1) -> None:
        """src/app.rs:204-215: `time_speed()` frames, each pushed onto the history."""
        for _ in range(int(n_frames)):
            if self.keep_history and self.current_state + 1 < len(self.states):
                del self.states[self.current_state + 1:]  # stepping after a rewind discards the redo tail
            self.worlds.integrate(self.dt, self.substeps, 1)
            self.current_state += 1
            if self.keep_history:
                self.states.append(self._snapshot())

    def rewind(self, n_frames: int = 1) -> None:
        """src/app.rs:146-148: go back in the history.  Only valid when the history covers the whole batch."""
        if not self.keep_history:
            raise RuntimeError("history is disabled")
        if self.first_world != 0 or self.n_recorded != self.worlds.n_worlds:
            raise RuntimeError("rewind needs a history of the whole batch")
        self.current_state = max(0, self.current_state - int(n_frames))
        self.worlds.upload_dynamic(self.states[self.current_state])
        self.worlds.sync()

    def trajectory(self, world: int = 0, body: int = 0) -> np.ndarray:
        """(frames + 1) x 13 array of one body's recorded states."""
        return np.stack([s[world - self.first_world, body] for s in self.states[: self.current_state + 1]])

    def dump_csv(self, path: str, world: int = 0) -> None:
        """One row per (frame, body) of `world`: frame, body, px..wz — the parity tooling's trajectory format."""
        with open(path, "w") as f:
            f.write("frame,body," + ",".join(DYN_FIELDS) + "\n")
            for k, s in enumerate(self.states[: self.current_state + 1]):
                for b in range(s.shape[1]):
                    f.write(f"{k},{b}," + ",".join(repr(float(x)) for x in s[world - self.first_world, b]) + "\n")
