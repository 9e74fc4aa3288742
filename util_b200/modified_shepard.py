"""ShepMod, interface of util/approximate/modified_shepard/__init__.py:15-65.

_fit = SHEPMOD (radius of influence per node) and _predict = SHEPMODVAL + the
wrapper's `wts > 0` filter, both as batched CUDA calls (csrc/shepmod.cu).
"""
import numpy as np

from . import _lib
from .base import WeightedApproximator


class TooFewDataPoints(Exception):
    def __init__(self, d, n):
        super().__init__(f"Only {n} points were provided but {d+2} are needed to define a modified shepard interpolant.")


class ShepMod(WeightedApproximator):
    def __init__(self, device=None):
        self.ierrors = {}
        self.x = self.rw = None
        self.device = device
        self.model = None
        self.capacity = 64  # hit-list width per query; grown automatically on overflow

    def _fit(self, control_points, **kwargs):
        self.x = np.asfortranarray(control_points.T)  # (M, N) like the reference (:30)
        if self.model is not None: self.model.close()
        self.model = _lib.Model(control_points, device=self.device)
        self.rw, ier = self.model.shepmod_fit()
        if ier != 0:
            raise TooFewDataPoints(*self.x.shape)

    def _run(self, x, y=None):
        cap = max(self.capacity, self.x.shape[0] + 1)
        ids, wts, cnt, ier, out = self.model.shepmod_predict(x, cap=cap, y=y, want_weights=y is None)
        self.capacity = max(self.capacity, self.model.last_cap)
        for code, num in zip(*np.unique(ier, return_counts=True)):
            self.ierrors[int(code)] = self.ierrors.get(int(code), 0) + int(num)  # :50
        return ids, wts, cnt, out

    def _predict(self, x):
        ids, wts, cnt, _ = self._run(x)
        return ([ids[i, :cnt[i]] for i in range(len(ids))],
                [wts[i, :cnt[i]] for i in range(len(ids))])

    def _predict_values(self, x):
        return self._run(x, y=True)[3][:, 0]

    def errors(self):
        print("%i normal returns." % self.ierrors.get(0, 0))
        print(("%i points outside the radius of influence " +
               "of all other points.") % self.ierrors.get(1, 0))
        return self.ierrors.get(0, 0), self.ierrors.get(1, 0)
