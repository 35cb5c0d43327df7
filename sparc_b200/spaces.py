"""Action/observation spaces.  Uses gymnasium when it is installed; otherwise a small built-in
`Box` / `Dict` with the same constructor and `.sample()` (gymnasium is an optional dependency —
the reference needs only these two classes, wire_edm.py:84-101)."""
from __future__ import annotations

import numpy as np

try:  # pragma: no cover - depends on the environment
    import gymnasium as _gym
    from gymnasium.spaces import Box, Dict  # noqa: F401

    EnvBase = _gym.Env
    HAVE_GYMNASIUM = True
except Exception:  # gymnasium absent: minimal stand-ins
    HAVE_GYMNASIUM = False

    class Box:
        def __init__(self, low, high, shape=None, dtype=np.float32, seed=None):
            self.dtype = np.dtype(dtype)
            self.shape = tuple(shape) if shape is not None else np.shape(low)
            self.low = np.full(self.shape, low, dtype=self.dtype)
            self.high = np.full(self.shape, high, dtype=self.dtype)
            self._rng = np.random.default_rng(seed)

        def seed(self, seed=None):
            self._rng = np.random.default_rng(seed)

        def sample(self):
            if np.issubdtype(self.dtype, np.integer):
                return self._rng.integers(self.low, self.high + 1, size=self.shape).astype(self.dtype)
            return self._rng.uniform(self.low, self.high, size=self.shape).astype(self.dtype)

        def contains(self, x):
            x = np.asarray(x)
            return x.shape == self.shape and bool(np.all(x >= self.low) and np.all(x <= self.high))

    class Dict(dict):
        def __init__(self, spaces=None, seed=None):
            super().__init__(spaces or {})
            self.spaces = self

        def seed(self, seed=None):
            for k, sp in self.items():
                sp.seed(None if seed is None else seed + hash(k) % 1000)

        def sample(self):
            return {k: sp.sample() for k, sp in self.items()}

        def contains(self, x):
            return isinstance(x, dict) and all(k in x and sp.contains(x[k]) for k, sp in self.items())

    class EnvBase:
        """What gymnasium.Env provides that the reference relies on: metadata, np_random seeding."""

        metadata: dict = {}
        np_random = None

        def reset(self, *, seed=None, options=None):
            if seed is not None or self.np_random is None:
                self.np_random = np.random.default_rng(seed)
