"""Model combinators with the interface of util/approximate/wrappers.py.

`Split` (:13-36) and `unique` (:42-116) are host-side glue around any model of this
package; the distance work underneath stays on the GPU.  `condition` (:122-184) with an
explicit linear map: the reference derives its map from PLRM / util.stats.pca / mpca,
which are outside this package's scope, so here the caller supplies the
d x d' matrix (the "PCA"/"MPCA" branch's `self.conditioner`, :170) and the wrapper applies
it before fit and predict exactly as :172,183 do.
"""
import numpy as np

from .base import Approximator, UnexpectedType


class Split(Approximator):
    """wrappers.py:13-36: k models, each fitted on every k-th row of one seeded shuffle of
    the data, predictions averaged."""

    def __init__(self, model_type, k=5, seed=0, *args, **kwargs):
        self._seed = seed
        self._count = k
        self._models = [model_type(*args, **kwargs) for _ in range(k)]

    def _fit(self, x, y):
        # np.random.seed + np.random.shuffle of arange(n), as :22-24 (same permutation)
        np.random.seed(self._seed)
        order = np.arange(len(x))
        np.random.shuffle(order)
        for i, model in enumerate(self._models):
            rows = order[i:len(x):self._count]
            model.fit(x[rows], y[rows])

    def _predict(self, z):
        return np.mean(np.array([model(z) for model in self._models]), axis=0)


def unique(weighted_approximator):
    """wrappers.py:42-116: fit the wrapped weighted model on the distinct rows only (first
    occurrence, ascending), responses of identical rows averaged; without responses,
    predict maps the weights back onto every original row (duplicates share the weight of
    their representative, then the row is renormalised)."""

    class UniquePointsApproximator(weighted_approximator):
        original_points = None
        unique_points = None
        unique_indices = None
        original_values = None

        def fit(self, points, values=None, *args, **kwargs):
            if (not issubclass(type(points), np.ndarray)) or (len(points.shape) != 2):
                raise UnexpectedType("Expected 2D numpy array as first argument.")
            self.original_points = points
            groups = {}
            for i, pt in enumerate(points):
                groups.setdefault(tuple(pt), []).append(i)
            self.unique_points = groups
            self.unique_indices = np.array(sorted(rows[0] for rows in groups.values()))
            if values is not None:
                self.original_values = values
                # one averaged response per distinct row, in order of first occurrence (:68-76)
                averaged = []
                for i in self.unique_indices:
                    rows = groups[tuple(points[i])]
                    wt = 1. / len(rows)
                    averaged.append(sum(values[j] * wt for j in rows))
                args = args + (averaged,)
            return super().fit(points[self.unique_indices, :], *args, **kwargs)

        def predict(self, points, *args, **kwargs):
            if (not issubclass(type(points), np.ndarray)) or (0 >= len(points.shape) < 2):
                raise UnexpectedType("Provided 'points' should be a 1D or 2D numpy array.")
            if self.original_values is not None:
                return super().predict(points)
            single = len(points.shape) == 1
            if single:
                points = np.reshape(points, (1, len(points)))
            indices, weights = self._predict(points, *args, **kwargs)
            response = []
            for ids, wts in zip(indices, weights):
                orig_ids, orig_wts = [], []
                for i, w in zip(ids, wts):
                    rows = self.unique_points[tuple(self.original_points[self.unique_indices[i]])]
                    orig_ids += rows
                    orig_wts += [w] * len(rows)
                total = sum(orig_wts)
                response.append((orig_ids, [w / total for w in orig_wts]))
            return response[0] if single else response

    return UniquePointsApproximator


def condition(approximator, conditioner):
    """wrappers.py:122-184 for a given linear map: the model is fitted on, and evaluated
    at, `x @ conditioner` (:172,183).  `conditioner` is the (d, d') matrix
    `(diag(values) @ components).T` of :170."""
    conditioner = np.asarray(conditioner, dtype=float)
    if conditioner.ndim != 2:
        raise UnexpectedType("conditioner should be a 2D (d, d') array.")

    class ConditionedApproximator(approximator):
        def _fit(self, x, *args, **kwargs):
            self.conditioner = conditioner
            return super()._fit(np.matmul(x, conditioner), *args, **kwargs)

        def _predict(self, x, *args, **kwargs):
            return super()._predict(np.matmul(x, self.conditioner), *args, **kwargs)

        def _predict_values(self, x, *args, **kwargs):
            return super()._predict_values(np.matmul(x, self.conditioner), *args, **kwargs)

    return ConditionedApproximator
