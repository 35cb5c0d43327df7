"""`spaces.Box` / `spaces.Dict` with `.sample()` — test infrastructure only."""
import numpy as np

_rng = np.random.default_rng(0)


class Box:
    def __init__(self, low, high, shape=None, dtype=np.float32):
        self.low, self.high = low, high
        self.shape = tuple(shape) if shape is not None else (1,)
        self.dtype = np.dtype(dtype)

    def sample(self):
        if np.issubdtype(self.dtype, np.integer):
            return _rng.integers(int(self.low), int(self.high) + 1, size=self.shape).astype(self.dtype)
        return _rng.uniform(self.low, self.high, size=self.shape).astype(self.dtype)

    def seed(self, seed=None):
        global _rng
        _rng = np.random.default_rng(seed)


class Dict(dict):
    def __init__(self, spaces=None):
        super().__init__(spaces or {})
        self.spaces = self

    def sample(self):
        return {k: v.sample() for k, v in self.items()}

    def seed(self, seed=None):
        Box.seed(None, seed)
