"""Pure-Python stand-in models used to pin util_b200.wrappers against the reference's wrappers
(make_golden.py builds them on the reference's base classes, test_wrappers.py on ours)."""
import numpy as np


def nearest_two(weighted_base):
    class NearestTwo(weighted_base):
        def _fit(self, x):
            self.x = x

        def _predict(self, z):
            ids, wts = [], []
            for p in z:
                ids.append(list(np.argsort(((self.x - p) ** 2).sum(1), kind="stable")[:2]))
                wts.append([0.75, 0.25])
            return ids, wts
    return NearestTwo


def linear(base):
    class Linear(base):
        def _fit(self, x, y):
            self.c = np.linalg.lstsq(x, y, rcond=None)[0]

        def _predict(self, z):
            return z @ self.c
    return Linear


def wrapper_inputs():
    rng = np.random.default_rng(3)
    xu = rng.integers(0, 3, size=(40, 2)).astype(float)  # many duplicated rows
    yu = rng.random(40)
    zu = rng.random((7, 2)) * 2
    xs, ys, zs = rng.random((50, 3)), rng.random(50), rng.random((9, 3))
    return xu, yu, zu, xs, ys, zs
