"""Minimal ``gym.spaces`` stand-ins.

The reference builds ``gym.spaces.Box`` / ``gym.spaces.MultiDiscrete`` objects
(gym_microrts/envs/vec_env.py:242-252).  ``gym`` is not a dependency of this package: when it is
importable the real classes are used, otherwise these shims provide the attributes callers read
(``shape``, ``dtype``, ``low``, ``high``, ``nvec``, ``sample``, ``seed``, ``contains``).
"""
from __future__ import annotations

import numpy as np

try:  # pragma: no cover - gym is absent in the build image
    from gym.spaces import Box, Discrete, MultiDiscrete  # type: ignore
except Exception:  # noqa: BLE001

    class _Space:
        def __init__(self, shape, dtype):
            self.shape = tuple(shape)
            self.dtype = np.dtype(dtype)
            self._rng = np.random.default_rng()

        def seed(self, seed=None):
            self._rng = np.random.default_rng(seed)
            return [seed]

    class Box(_Space):
        def __init__(self, low, high, shape, dtype=np.float32):
            super().__init__(shape, dtype)
            self.low = np.full(self.shape, low, self.dtype)
            self.high = np.full(self.shape, high, self.dtype)

        def sample(self):
            if np.issubdtype(self.dtype, np.integer):
                return self._rng.integers(self.low, self.high + 1, dtype=self.dtype)
            return self._rng.uniform(self.low, self.high).astype(self.dtype)

        def contains(self, x):
            x = np.asarray(x)
            return x.shape == self.shape and bool(np.all(x >= self.low) and np.all(x <= self.high))

        def __repr__(self):
            return f"Box({self.low.min()}, {self.high.max()}, {self.shape}, {self.dtype})"

    class MultiDiscrete(_Space):
        def __init__(self, nvec, dtype=np.int64):
            self.nvec = np.asarray(nvec, dtype=dtype)
            super().__init__(self.nvec.shape, dtype)

        def sample(self):
            return (self._rng.random(self.nvec.shape) * self.nvec).astype(self.dtype)

        def contains(self, x):
            x = np.asarray(x)
            return x.shape == self.shape and bool(np.all(x >= 0) and np.all(x < self.nvec))

        def __repr__(self):
            return f"MultiDiscrete({self.nvec})"

    class Discrete(_Space):
        def __init__(self, n):
            self.n = int(n)
            super().__init__((), np.int64)

        def sample(self):
            return int(self._rng.integers(0, self.n))

        def contains(self, x):
            return 0 <= int(x) < self.n

        def __repr__(self):
            return f"Discrete({self.n})"
