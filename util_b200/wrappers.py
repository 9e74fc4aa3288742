"""Model combinators with the interface of util/approximate/wrappers.py.

`Split` (:13-36) and `unique` (:42-116) are host-side glue around any model of this
package; the distance work underneath stays on the GPU.  `condition` (:117-184) keeps the
reference's signature: method="PCA" builds the linear map on the host (sklearn PCA + the
per-axis error scaling of util.stats.normalize_error), an explicit `conditioner=` matrix is
accepted as well, and the map is applied before fit and predict exactly as :172,183 do.
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


def _axis_slopes_to_scale(points, values, metric):
    """util.stats.normalize_error (stats/metric_pca.py:23-49): per axis the total metric variation between
    neighbours along that axis, divided by their spacing; the axes are then scaled by slope^-2,
    normalised by the median and to unit geometric mean."""
    n, d = points.shape
    slope = np.zeros(d)
    for axis in range(d):
        order = np.argsort(points[:, axis])
        for lo, hi in zip(order[:-1], order[1:]):
            gap = points[hi, axis] - points[lo, axis]
            if gap > 0:
                slope[axis] += metric(values[lo], values[hi]) / (gap * d)
    if slope.min() <= 0.0:  # axes without any slope are the only ones needed
        slope = np.where(slope == 0, 1., float("inf"))
    fix = slope ** (-2)
    fix /= np.median(fix)
    return fix / np.prod(fix) ** (1 / d)


def condition(approximator, metric=None, method="PLR", dim=None, samples=None, scale=True,
              display=False, seed=None, conditioner=None, **cond_kwargs):
    """wrappers.py:117-184: the model is fitted on, and evaluated at, `x @ conditioner` (:172,183).

    method="PCA" (:145-149): the principal axes of x (sklearn PCA, as the reference's util.stats.pca
    calls it) scaled by `normalize_error` so that the response has unit metric slope along each axis.
    `conditioner=` (an extension) supplies the (d, d') matrix `(diag(values) @ components).T` of :170
    directly.  The reference's default "PLR" trains its PLRM network and "MPCA" samples random point
    pairs through util.random; both lie outside this package and raise."""
    from .base import abs_diff
    if metric is None:
        metric = abs_diff
    if conditioner is not None:
        conditioner = np.asarray(conditioner, dtype=float)
        if conditioner.ndim != 2:
            raise UnexpectedType("conditioner should be a 2D (d, d') array.")
    elif method != "PCA":
        raise NotImplementedError(
            f"condition(method={method!r}): only method='PCA' and an explicit conditioner= matrix are "
            "available; the reference's PLR / MPCA conditioners need its PLRM network / random pair sampler")

    class ConditionedApproximator(approximator):
        def _fit(self, x, *args, num_comps=None, **kwargs):
            if conditioner is not None:
                self.conditioner = conditioner
            else:
                y = args[0] if args else getattr(self, "y", None)
                if y is None:
                    raise UnexpectedType("condition(method='PCA') needs response values to scale the axes.")
                if num_comps is None: num_comps = min(x.shape + (16,))   # :128
                if dim is not None: num_comps = min(dim, num_comps)       # :129
                from sklearn.decomposition import PCA
                num_comps = min(num_comps, *x.shape)
                components = PCA(n_components=num_comps).fit(x).components_    # :146, stats/metric_pca.py:82-94
                values = _axis_slopes_to_scale(np.matmul(x, components.T), y, metric)   # :148
                if not scale: values[:] = 1.                                           # :157
                self.conditioner = np.matmul(np.diag(values), components).T           # :170
            return super()._fit(np.matmul(x, self.conditioner), *args, **kwargs)

        def _predict(self, x, *args, **kwargs):
            return super()._predict(np.matmul(x, self.conditioner), *args, **kwargs)

        def _predict_values(self, x, *args, **kwargs):
            return super()._predict_values(np.matmul(x, self.conditioner), *args, **kwargs)

    return ConditionedApproximator
