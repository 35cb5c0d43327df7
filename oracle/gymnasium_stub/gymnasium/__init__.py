"""Minimal stand-in for the `gymnasium` package (TEST INFRASTRUCTURE ONLY).

gymnasium is not installed in this image and cannot be installed (no network).
The reference (`/root/reference/src/wedm/envs/wire_edm.py:4-6,17,84-101,107`)
touches only `gym.Env`, `spaces.Box` and `spaces.Dict`; this stub supplies just
those so the unmodified reference can be imported to generate golden vectors.
It is never imported by the product package `sparc_b200`.
"""
from . import spaces  # noqa: F401
import numpy as _np


class Env:
    metadata: dict = {}
    np_random = None

    def reset(self, *, seed=None, options=None):
        # gymnasium.Env.reset reseeds np_random only when a seed is given
        if seed is not None or self.np_random is None:
            self.np_random = _np.random.default_rng(seed)
        return None


__all__ = ["Env", "spaces"]
