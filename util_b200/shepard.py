"""Shepard inverse-distance weighting, interface of util/approximate/shepard/__init__.py:20-47.

As `NearestNeighbor(method=Shepard)` the weights over the k neighbours are computed by
avg_shepard_kernel; stand-alone `Shepard()` weights EVERY fitted point per query
(WEIGHT, shepard/weight.f90:1-34) through shepard_all_kernel: the weighted sums are fused
into the scan, and the dense (q, n) weight matrix the reference returns is only
materialised when a caller asks for it (`_predict`).
"""
import numpy as np

from . import _lib
from .base import WeightedApproximator


class Shepard(WeightedApproximator):
    _b200_method = _lib.SHEPARD
    points = None
    model = None

    def __init__(self, *args, device=None, **kwargs):
        self.device = device

    def _fit(self, points):
        # the reference keeps points as (d, n) Fortran order (shepard/__init__.py:32)
        self.points = np.asarray(points.copy().T, order="F")
        if self.model is not None: self.model.close()
        self.model = _lib.Model(points, device=self.device)

    def _predict(self, points, display_wait_sec=3.):
        n = self.points.shape[1]
        weights, _ = self.model.shepard_predict(points, want_weights=True)
        indices = np.zeros((points.shape[0], n), dtype=int) + np.arange(n)  # :36-37
        return indices, weights

    def _predict_values(self, points, display_wait_sec=3.):
        return self.model.shepard_predict(points, y=True)[1][:, 0]
