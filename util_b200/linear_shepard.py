"""LSHEP, interface of util/approximate/linear_shepard/__init__.py:19-89.

_fit = LSHEP (non-robust: the reference wrapper never forwards `rlshep`, :36,53)
per response column; _predict = LSHEPVAL per query and column; both batched CUDA
calls (csrc/lshep.cu).
"""
import numpy as np

from . import _lib
from .base import Approximator
from .modified_shepard import TooFewDataPoints


class LSHEP(Approximator):
    def __init__(self, radius=2, device=None, robust=False, ripple=False):
        """`robust=True` (not in the reference's wrapper, which never passes RLSHEP to the Fortran,
        linear_shepard/__init__.py:36,53) selects the module's robust fit: IRLS with Huber and bisquare
        weights, linear_shepard.f95:136-175.  `ripple=True` selects the module's RIPPLE fit (:474-665)."""
        self.ierrors = {}
        self.robust = robust
        self.ripple = ripple
        self.radius = radius
        self.x = self.f = self.a = self.rw = None
        self.device = device
        self.model = None

    def _fit(self, control_points, values, rlshep=True, **kwargs):
        self.x = np.asfortranarray(control_points.T)
        if self.x.shape[1] <= self.x.shape[0] + 1:
            raise TooFewDataPoints(*self.x.shape)
        if self.model is not None: self.model.close()
        self.model = _lib.Model(control_points, device=self.device)
        if self.ripple:
            if values.shape[1] != 1:
                raise ValueError("RIPPLE takes one response column")
            a, rw, ier = self.model.ripple_fit(values[:, 0])
        else:
            a, rw, ier = self.model.lshep_fit(values, robust=self.robust)
        self.fit_ier = ier
        r = values.shape[1]
        # per-column lists like the reference (:43-55); a[i] is (M, N), rw[i] already * radius
        self.f = [np.asfortranarray(values[:, i]) for i in range(r)]
        self.a = [np.asfortranarray(a[i].T) for i in range(r)]
        self.rw = [rw * self.radius for _ in range(r)]

    def _predict(self, x, display_wait_sec=.5):
        out, ier = self.model.lshep_predict(x, self.radius, len(self.f))
        for code, num in zip(*np.unique(ier, return_counts=True)):
            self.ierrors[int(code)] = self.ierrors.get(int(code), 0) + int(num)  # :74
        return out

    def errors(self):
        print("%i normal returns." % self.ierrors.get(0, 0))
        print(("%i points outside the radius of influence " +
               "of all other points.") % self.ierrors.get(1, 0))
        print(("%i points at which hyperplane normals are significantly" +
               " different.\n  This is only hazardous for piecewise " +
               "linear underlying functions.") % self.ierrors.get(2, 0))
        return self.ierrors.get(0, 0), self.ierrors.get(1, 0), self.ierrors.get(2, 0)
