"""Action-space objects.  The reference uses ``gymnasium.spaces.Box`` and
``gymnasium.vector.utils.batch_space`` (controllers/base_controller.py:9-10);
gymnasium is used when importable, otherwise this minimal Box keeps the same
attributes the controllers and their callers read (low, high, shape, dtype, sample)."""
from __future__ import annotations

import numpy as np

try:  # pragma: no cover - gymnasium is absent in the build image
    from gymnasium import spaces as _gs
    from gymnasium.vector.utils import batch_space as _batch_space

    Box = _gs.Box
    batch_space = _batch_space
except Exception:

    class Box:
        def __init__(self, low, high, shape=None, dtype=np.float32):
            if shape is None:
                shape = np.broadcast(np.asarray(low), np.asarray(high)).shape
            self.dtype = np.dtype(dtype)
            self.low = np.broadcast_to(np.asarray(low, dtype=self.dtype), shape).copy()
            self.high = np.broadcast_to(np.asarray(high, dtype=self.dtype), shape).copy()
            self.shape = tuple(shape)
            self._rng = np.random.default_rng()

        def seed(self, seed=None):
            self._rng = np.random.default_rng(seed)

        def sample(self):
            lo = np.where(np.isfinite(self.low), self.low, -1.0)
            hi = np.where(np.isfinite(self.high), self.high, 1.0)
            return self._rng.uniform(lo, hi).astype(self.dtype)

        def contains(self, x) -> bool:
            x = np.asarray(x)
            return x.shape == self.shape and bool(np.all(x >= self.low) and np.all(x <= self.high))

        def __repr__(self):
            return f"Box({self.low.min()}, {self.high.max()}, {self.shape}, {self.dtype})"

    def batch_space(space: "Box", n: int = 1) -> "Box":
        return Box(np.tile(space.low, (n,) + (1,) * space.low.ndim), np.tile(space.high, (n,) + (1,) * space.high.ndim),
                   dtype=space.dtype)


def normalize_action_space(space: "Box") -> "Box":
    """gym_utils.normalize_action_space (utils/gym_utils.py:108-110)."""
    return Box(-1, 1, shape=space.shape, dtype=space.dtype)
