"""Voronoi-cell weights, interface of util/approximate/voronoi/__init__.py:22-62.

As `NearestNeighbor(k, method=Voronoi)` (BASELINE.json config 4, "candidate-pruned
by kNN") the weights on the k neighbours come from voronoi_kernel (PREDICT,
voronoi/voronoi.f90:57-155).  Stand-alone `Voronoi()` over all n fitted points is
served by the same kernel with k = n while its Gram matrix fits in shared
memory (n <= 96 or so); the un-pruned large-n mesh is not built yet.
"""
import numpy as np

from . import _lib
from .base import WeightedApproximator


class DuplicateInterpolationPoints(Exception): pass


class Voronoi(WeightedApproximator):
    _b200_method = _lib.VORONOI
    _MAX_N = 96
    points = None
    dots = None  # the reference's lazily filled n x n cache; results never depended on it
    model = None

    def __init__(self, device=None):
        self.device = device

    def _fit(self, points):
        self.points = points.copy().T
        if points.shape[0] > self._MAX_N:
            raise NotImplementedError(
                f"stand-alone Voronoi over n={points.shape[0]} > {self._MAX_N} points is not built; "
                "use NearestNeighbor(k, method=Voronoi)")
        if self.model is not None: self.model.close()
        self.model = _lib.Model(points, device=self.device)

    def _predict(self, points, display_wait_sec=3.):
        n = self.points.shape[1]
        ids, wts, cnt, err, _ = self.model.nn_predict(points, n, _lib.VORONOI)
        if (err == 1).any():
            raise OverflowError("The scale of the data caused a squared float to overflow, consider normalizing data.")
        if (err == 2).any():
            raise DuplicateInterpolationPoints("Some fit points were duplicated.")
        indices, weights = [], []
        for i in range(len(ids)):
            o = np.argsort(ids[i, :cnt[i]])  # idx[support > 0]: ascending point index (:58-61)
            indices.append(ids[i, :cnt[i]][o])
            weights.append(wts[i, :cnt[i]][o])
        return indices, weights
