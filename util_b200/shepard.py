"""Shepard inverse-distance weighting, interface of util/approximate/shepard/__init__.py:20-47.

As `NearestNeighbor(method=Shepard)` the weights over the k neighbours are
computed by avg_shepard_kernel (WEIGHT, shepard/weight.f90:1-34).  Stand-alone
`Shepard()` weights every fitted point; it is served by the same kernel with
k = n, which the kernel supports up to n = 128 (the reference's demo uses 20
points, shepard/__init__.py:53); larger stand-alone fits are not built yet.
"""
import numpy as np

from . import _lib
from .base import WeightedApproximator


class Shepard(WeightedApproximator):
    _b200_method = _lib.SHEPARD
    _MAX_N = 128
    points = None
    model = None

    def __init__(self, *args, device=None, **kwargs):
        self.device = device

    def _fit(self, points):
        # the reference keeps points as (d, n) Fortran order (shepard/__init__.py:32)
        self.points = np.asarray(points.copy().T, order="F")
        if points.shape[0] > self._MAX_N:
            raise NotImplementedError(
                f"stand-alone Shepard over n={points.shape[0]} > {self._MAX_N} points is not built; "
                "use NearestNeighbor(k, method=Shepard)")
        if self.model is not None: self.model.close()
        self.model = _lib.Model(points, device=self.device)

    def _predict(self, points, display_wait_sec=3.):
        n = self.points.shape[1]
        ids, wts, _, _, _ = self.model.nn_predict(points, n, _lib.SHEPARD)
        # the reference returns indices 0..n-1 with weights in that order (:36-46)
        indices = np.zeros((points.shape[0], n), dtype=int) + np.arange(n)
        weights = np.zeros(indices.shape)
        np.put_along_axis(weights, ids, wts, axis=1)
        return indices, weights
