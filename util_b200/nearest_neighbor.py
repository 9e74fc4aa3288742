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
    """Average / Shepard / Voronoi sub-models have a fused CUDA kernel; None for any other class."""
    return getattr(method, "_b200_method", None)


class NearestNeighbor(WeightedApproximator):
    def __init__(self, k=1, method=Average, device=None, devices=None, shard="query", **auto_kwargs):
        """`device` / `devices` / `shard` are additions to the reference's constructor: one GPU ordinal, or
        a list of ordinals over which the queries (shard="query", the default: every GPU holds all fitted
        points) or the fitted points (shard="data": local top-k, NCCL exchange, on-device merge) are split."""
        self.points = None
        self.method = method
        self.num_neighbors = k
        self.auto_kwargs = auto_kwargs
        self.device = device
        self.devices = list(devices) if devices is not None else None
        self.shard = shard
        self.model = None  # _lib.Model: the device-resident points (stands in for self.tree)

    def _fit(self, control_points, k=None, display=True, **kwargs):
        if k is not None: self.num_neighbors = k
        self.points = control_points.copy()
        if self.model is not None: self.model.close()
        if self.devices is not None and len(self.devices) > 1:
            if _method_code(self.method) not in (_lib.AVERAGE, _lib.SHEPARD):
                raise NotImplementedError("several devices: Average and Shepard sub-models")
            self.model = _lib.MultiModel(self.points, self.devices, shard=self.shard)
        else:
            dev = self.devices[0] if self.devices else self.device
            self.model = _lib.Model(self.points, device=dev)
        self.tree = self.model  # attribute name users of the reference read
        if self.num_neighbors is None:
            self.num_neighbors = self._choose_k(display, kwargs)

    def _choose_k(self, display, overrides):
        """k=None (nearest_neighbor/__init__.py:42-55): without responses k = 1, otherwise auto() with the
        constructor's keyword arguments, those of this fit() call on top, and `mean` defaulting to
        "not a classifier".  Progress lines as the reference prints them."""
        if self.y is None:
            return 1
        self.auto_kwargs.update(overrides)
        self.auto_kwargs.setdefault("mean", not self.classifier)
        end = "\n" if display else "\r"
        print("Nearest neighbor, estimating best value for 'k'..", end=end, flush=True)
        k = auto(self.points, self.y, **self.auto_kwargs)
        print(f"  chose k = {k}", end=end, flush=True)
        print(" " * 49, end=end, flush=True)
        return k

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
        if code is None:
            return self._predict_any_submodel(points, k)
        ids, wts, cnt, err, _ = self.model.nn_predict(points, k, code)
        self._raise_submodel_errors(code, err)
        if code == _lib.VORONOI:
            return ([ids[i, :cnt[i]] for i in range(len(ids))],
                    [wts[i, :cnt[i]] for i in range(len(ids))])
        return ids, wts

    def _predict_any_submodel(self, points, k):
        """Any other WeightedApproximator class as `method` (nearest_neighbor/__init__.py:78-83): the k
        neighbours of all queries come from one batched GPU call; the sub-model is then constructed,
        fitted on each query's neighbours and evaluated there, as the reference does per point."""
        closest = self.model.knn(points, k, want_d2=False)
        pts, wts = [], []
        for nbrs, pt in zip(closest, points):
            sub = self.method()
            sub.fit(self.points[nbrs])
            guess_pts, guess_wts = sub.predict(pt)
            pts.append(nbrs[np.asarray(guess_pts, dtype=np.int64)])
            wts.append(np.asarray(guess_wts))
        return pts, wts

    def _predict_values(self, points, max_size=2**20, display=True):
        k = self._check()
        code = _method_code(self.method)
        if code is None:
            return None  # no fused path: base.predict forms the sums from (ids, wts)
        _, _, _, err, out = self.model.nn_predict(points, k, code, y=True, want_weights=False)
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


def _default_sample(n, count):
    """Row sample of the reference's auto(): util.random.random_range(n, count=count) draws
    random.sample(range(n), count) whenever n <= 2**15 (util/random/random.py:55-58), so with the
    same `random` seed this picks the same rows there; for larger n the reference walks a linear
    congruential generator, which a caller who needs the identical rows can pass in as `indices`."""
    import random
    return random.sample(range(n), min(count, n))


def auto(points, values, metric=abs_diff, max_k=None, samples=100, mean=True, k_step=1,
         model=None, display=False, indices=None):
    """Choice of k as in nearest_neighbor/__init__.py:93-132: for a sample of the fitted rows, the
    mean (or mean of the minimum) response difference to their k nearest other rows, over
    k = 1, 2, 4, ... <= max_k; the k with the lowest value wins (the smallest on ties).

    One batched kNN call (k = max_k + 1, first column dropped as in :107) replaces the reference's
    per-point loop.  `indices` (optional) fixes the sampled rows; `model` is accepted for signature
    compatibility (the reference ignores it as well, :104)."""
    n = len(points)
    if max_k is None:
        max_k = 1 << ((n // 2).bit_length() - 1)  # largest power of two <= n/2  (:98-100)
    rows = np.asarray(list(indices) if indices is not None else _default_sample(n, samples), dtype=np.int64)
    knn = NearestNeighbor(k=max_k + 1)
    knn.fit(points)
    nbrs = knn.neighbors(points[rows], k=max_k + 1)[:, 1:]
    knn.model.close()
    if metric is abs_diff and isinstance(values, np.ndarray) and values.ndim == 1 and values.dtype.kind in "fiu":
        diffs = np.abs(values[rows, None] - values[nbrs])
    else:  # any pairwise metric over arbitrary response objects
        diffs = np.array([[metric(values[i], values[j]) for j in nbrs[r]] for r, i in enumerate(rows)])
    top = max_k.bit_length() - 1  # int(log2(max_k))
    ks = [1 << e for e in range(0, top + 1, k_step)]
    if ks[-1] != max_k and (1 << top) not in ks:
        ks.append(1 << top)  # (:118-121) the largest power of two is always a candidate
    score = (lambda k: np.mean(diffs[:, :k])) if mean else (lambda k: np.mean(np.min(diffs[:, :k], axis=1)))
    table = {k: score(k) for k in ks}
    best_k = min(table.items(), key=lambda kv: kv[1])[0]
    if display:
        width = len(str(max_k - 1)) if max_k > 1 else 0
        print("-" * 52)
        print(f" Estimated {'mean' if mean else 'minimum'} error for various choices of 'k':")
        for k in sorted(table):
            print(f"  k = {k:{width}d} ~ {table[k]:.4e}" + ("  <-- chosen 'k'" if k == best_k else ""))
        print("-" * 52)
    return best_k
