"""NearestNeighbor / Average with the interface of
util/approximate/nearest_neighbor/__init__.py:7-132.

The reference builds an sklearn BallTree in _fit and then, per query point in a
Python loop, queries the tree, constructs `self.method()`, fits it on the k
neighbours and predicts (:69-83).  Here _fit uploads the points to the GPU and
_predict is one batched call: fused tensor-core distance + candidate selection,
fp64 re-rank (indices bit-exact, ties to the lowest index), and the sub-model
weights (Average / Shepard / Voronoi) evaluated by CUDA kernels.
"""
import numpy as np

from . import _lib
from .base import WeightedApproximator, abs_diff


class Average(WeightedApproximator):
    """nearest_neighbor/__init__.py:7-14: equal weights over the fitted points."""
    _b200_method = _lib.AVERAGE

    def _fit(self, points):
        self.points = points

    def _predict(self, points):
        n = len(self.points)
        indices = np.tile(np.arange(n), (len(points), 1))
        weights = np.full((len(points), n), 1 / n)
        return indices, weights


def _method_code(method):
    code = getattr(method, "_b200_method", None)
    if code is None:
        raise NotImplementedError(
            f"NearestNeighbor(method={getattr(method, '__name__', method)}): only Average, Shepard "
            "and Voronoi sub-models have a CUDA kernel")
    return code


class NearestNeighbor(WeightedApproximator):
    def __init__(self, k=1, method=Average, device=None, **auto_kwargs):
        self.points = None
        self.method = method
        self.num_neighbors = k
        self.auto_kwargs = auto_kwargs
        self.device = device
        self.model = None  # _lib.Model: the device-resident points (stands in for self.tree)

    def _fit(self, control_points, k=None, display=True, **kwargs):
        if k is not None: self.num_neighbors = k
        self.points = control_points.copy()
        if self.model is not None: self.model.close()
        self.model = _lib.Model(self.points, device=self.device)
        self.tree = self.model  # attribute name users of the reference read
        # nearest_neighbor/__init__.py:42-55: automatic k
        if self.num_neighbors is None:
            if self.y is not None:
                self.auto_kwargs.update(kwargs)
                if "mean" not in self.auto_kwargs:
                    self.auto_kwargs["mean"] = not self.classifier
                end = ("\n" if display else "\r")
                print("Nearest neighbor, estimating best value for 'k'..", end=end, flush=True)
                self.num_neighbors = auto(self.points, self.y, **self.auto_kwargs)
                print(f"  chose k = {self.num_neighbors}", end=end, flush=True)
                print("                                                 ", end=end, flush=True)
            else:
                self.num_neighbors = 1

    def _check(self):
        if self.model is None:
            raise RuntimeError("NearestNeighbor: call fit() first")
        k = int(self.num_neighbors)
        if k > len(self.points):
            raise ValueError(f"k={k} exceeds the number of fitted points ({len(self.points)})")
        return k

    def _predict(self, points, max_size=2**20, display=True):
        """-> (pts, wts): neighbour indices closest[guess_pts] and weights per query.

        Rectangular (q,k) arrays for Average and Shepard as in the reference; for
        Voronoi (only positive-weight neighbours are kept, a varying number per
        query) lists of 1-D arrays -- np.array() of those rows is what raises
        "inhomogeneous shape" in the reference on numpy >= 1.24.
        """
        k = self._check()
        code = _method_code(self.method)
        ids, wts, cnt, err, _ = self.model.nn_predict(points, k, code)
        self._raise_submodel_errors(code, err)
        if code == _lib.VORONOI:
            return ([ids[i, :cnt[i]] for i in range(len(ids))],
                    [wts[i, :cnt[i]] for i in range(len(ids))])
        return ids, wts

    def _predict_values(self, points, max_size=2**20, display=True):
        k = self._check()
        code = _method_code(self.method)
        _, _, _, err, out = self.model.nn_predict(points, k, code, y=self._y_numeric, want_weights=False)
        self._raise_submodel_errors(code, err)
        return out[:, 0]

    @staticmethod
    def _raise_submodel_errors(code, err):
        if code == _lib.VORONOI and err.any():
            from .voronoi import DuplicateInterpolationPoints
            # voronoi/__init__.py:54-57
            if (err == 1).any():
                raise OverflowError("The scale of the data caused a squared float to overflow, "
                                    "consider normalizing data.")
            raise DuplicateInterpolationPoints("Some fit points were duplicated.")

    def neighbors(self, points, k=None, return_distance=False):
        """Batched BallTree.query: indices (q,k) ascending by distance (ties: lowest index)."""
        k = int(k or self.num_neighbors)
        idx, d2 = self.model.knn(points, k)
        return (np.sqrt(d2), idx) if return_distance else idx


def auto(points, values, metric=abs_diff, max_k=None, samples=100, mean=True, k_step=1,
         model=NearestNeighbor, display=False):
    """nearest_neighbor/__init__.py:93-132: pick the power-of-two k minimising the mean
    (or mean-min) response difference to the k nearest neighbours of sampled points."""
    from math import floor, log2
    from random import sample
    if max_k is None:
        max_k = 2 ** floor(log2(len(points) // 2))
        # the CUDA kNN keeps at most 1024 neighbours per query; the reference's default (half the data
        # set) is capped accordingly -- the error curve is flat long before that in practice
        max_k = min(max_k, 512)
    model = NearestNeighbor(k=max_k + 1)
    model.fit(points)
    # util.random.random_range(len(points), count=samples) (:104)
    indices = sample(range(len(points)), min(samples, len(points)))
    neighbors = model.neighbors(points[indices])[:, 1:]
    differences = np.array([[metric(values[i1], values[i2]) for i2 in neighbors[i]]
                            for i, i1 in enumerate(indices)])
    k_values = {}
    for k_pow in range(0, int(log2(max_k)) + 1, k_step):
        k = 2 ** k_pow
        if mean: k_values[k] = np.mean(differences[:, :k])
        else: k_values[k] = np.mean(np.min(differences[:, :k], axis=1))
    if 2 ** k_pow != max_k:
        k = 2 ** int(log2(max_k))
        if mean: k_values[k] = np.mean(differences[:, :k])
        else: k_values[k] = np.mean(np.min(differences[:, :k], axis=1))
    best_k = min(k_values.items(), key=lambda i: i[1])[0]
    if display:
        name = "mean" if mean else "minimum"
        from math import log10, ceil
        print('-' * 52)
        print(" Estimated " + name + " error for various choices of 'k':")
        for k in sorted(k_values):
            extra = "  <-- chosen 'k'" if k == best_k else ""
            print(f"  k = {k:{ceil(log10(max_k))}d} ~ {k_values[k]:.4e}" + extra)
        print('-' * 52)
    return best_k
