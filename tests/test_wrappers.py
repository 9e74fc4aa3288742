"""Host logic of util_b200.wrappers (Split / unique / condition, util/approximate/wrappers.py) with
pure-Python stand-in models: no GPU needed."""
import numpy as np
import pytest

from util_b200.base import Approximator, WeightedApproximator, UnexpectedType
from util_b200.wrappers import Split, unique, condition


class NearestOne(WeightedApproximator):
    """weight 1 on the nearest fitted row"""
    def _fit(self, x):
        self.x = x
    def _predict(self, z):
        ids = [[int(np.argmin(((self.x - p) ** 2).sum(1)))] for p in z]
        return ids, [[1.0] for _ in z]


class MeanModel(Approximator):
    fitted_rows = None
    def _fit(self, x, y):
        self.fitted_rows = len(x)
        self.m = y.mean(0)
    def _predict(self, z):
        return np.tile(self.m, (len(z), 1))


def test_split_partitions_rows_and_averages():
    rng = np.random.default_rng(0)
    x, y = rng.random((23, 3)), rng.random(23)
    s = Split(MeanModel, k=4, seed=7)
    s.fit(x, y)
    sizes = [m.fitted_rows for m in s._models]
    assert sum(sizes) == 23 and max(sizes) - min(sizes) <= 1
    np.random.seed(7)
    order = np.arange(23)
    np.random.shuffle(order)
    want = np.mean([y[order[i::4]].mean() for i in range(4)])
    got = s(rng.random((5, 3)))
    assert got.shape == (5,) and np.allclose(got, want, rtol=0, atol=1e-15)


def test_unique_averages_duplicates_and_maps_weights_back():
    x = np.array([[0., 0.], [1., 0.], [0., 0.], [0., 1.], [1., 0.], [0., 0.]])
    y = np.array([1., 10., 2., 5., 20., 3.])
    U = unique(NearestOne)
    m = U()
    m.fit(x, y)
    assert list(m.unique_indices) == [0, 1, 3]
    assert np.array_equal(m.x, x[[0, 1, 3]])
    assert np.allclose(m(np.array([[0.1, 0.], [0.9, 0.1], [0., 0.8]])), [2., 15., 5.])
    m2 = U()
    m2.fit(x)
    ids, wts = m2(np.array([0.1, 0.1]))
    assert ids == [0, 2, 5] and np.allclose(wts, [1 / 3] * 3)
    with pytest.raises(UnexpectedType):
        m2.fit([1, 2, 3])


def test_condition_applies_the_map_on_both_sides():
    rng = np.random.default_rng(1)
    x, z = rng.random((30, 4)), rng.random((6, 4))
    A = np.diag([1., 0., 3., 0.])[:, [0, 2]]  # keep and scale two coordinates
    C = condition(NearestOne, conditioner=A)
    m = C()
    m.fit(x)
    assert m.x.shape == (30, 2)
    got = [ids[0] for ids, _ in m(z)]
    want = [int(np.argmin((((x - p) @ A) ** 2).sum(1))) for p in z]
    assert got == want


def test_wrappers_match_the_reference_golden():
    """tests/golden/wrappers.npz: the reference's own Split / unique (wrappers.py:13-116) run by
    tests/golden/make_golden.py around the stand-in models of tests/golden/standins.py."""
    import os, sys
    here = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    sys.path.insert(0, here)
    import standins
    g = np.load(os.path.join(here, "wrappers.npz"))
    xu, yu, zu, xs, ys, zs = standins.wrapper_inputs()
    a = unique(standins.nearest_two(WeightedApproximator))()
    a.fit(xu, yu)
    assert np.array_equal(np.array(a(zu)), g["unique_values"])
    b = unique(standins.nearest_two(WeightedApproximator))()
    b.fit(xu)
    assert np.array_equal(b.unique_indices, g["unique_indices"])
    for row, ids, wts in zip(b(zu), g["unique_ids"], g["unique_wts"]):
        m = len(row[0])
        assert list(row[0]) == list(ids[:m]) and np.all(ids[m:] == -1)
        assert np.array_equal(np.array(row[1]), wts[:m])
    s = Split(standins.linear(Approximator), k=4, seed=5)
    s.fit(xs, ys)
    assert np.array_equal(s(zs), g["split"])


def test_condition_pca_matches_the_reference_golden():
    """tests/golden/wrappers.npz `cond_*`: the reference's condition(method="PCA") (wrappers.py:117-184 with
    util.stats.pca / normalize_error) around the stand-in linear model."""
    import os, sys
    here = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    sys.path.insert(0, here)
    import standins
    g = np.load(os.path.join(here, "wrappers.npz"))
    xu, yu, zu, xs, ys, zs = standins.wrapper_inputs()
    for tag, kw in (("a", {}), ("b", dict(dim=2)), ("c", dict(scale=False))):
        m = condition(standins.linear(Approximator), method="PCA", **kw)()
        m.fit(xs, ys)
        assert np.allclose(m.conditioner, g[f"cond_{tag}_matrix"], rtol=1e-12, atol=1e-14)
        assert np.allclose(m(zs), g[f"cond_{tag}_pred"], rtol=1e-10, atol=1e-12)
    with pytest.raises(NotImplementedError):
        condition(standins.linear(Approximator))
