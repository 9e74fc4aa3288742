"""BASELINE.json configs C2, C3, C4 at their stated sizes through the drop-in classes,
checked against the CPU oracle on a seeded sample of >= 256 queries (and >= 256 fitted nodes for the
fitted state).  C1 is tests/test_gpu_parity.py::test_nn_golden (the reference's own output), C5 is
tests/test_gpu_tensor.py::test_tensor_knn_full_size_against_exact_kernel and bench.py's parity leg.
"""
import numpy as np
import pytest

from conftest import fun, synth

pytestmark = pytest.mark.gpu

RTOL = 1e-12  # north_star: predicted values within 1e-12 relative of the reference
SAMPLE = 256


def relerr(a, b):
    a, b = np.asarray(a, dtype=float), np.asarray(b, dtype=float)
    return np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-3 * max(1.0, np.abs(b).max())))


@pytest.fixture(scope="module")
def ub():
    import util_b200
    from util_b200 import _lib
    assert _lib.lib().b200_device_count() >= 1
    return util_b200


def test_c2_shepmod_100k(ub, oracle):
    """configs[1]: ShepMod, 100k fitted points x 100k queries, 16-D.  rw of sampled nodes = distance to
    the (M+1)-th nearest other node (modified_shepard.f95:45-57), bit for bit; predictions, supports,
    weights and IER of sampled queries against SHEPMODVAL (f95:83-176)."""
    n, q, d = 100000, 100000, 16
    x, z = synth(n, q, d)
    y = fun(x)
    m = ub.ShepMod()
    m.fit(x, y)
    pred = np.array(m.predict(z))
    rng = np.random.default_rng(5)
    nodes = rng.choice(n, SAMPLE, replace=False)
    _, d2 = oracle.knn(x, x[nodes], d + 2)  # self + the M+1 nearest others
    assert np.array_equal(np.sqrt(d2[:, -1]), m.rw[nodes])
    sel = rng.choice(q, SAMPLE, replace=False)
    oi, ow, oc, oe = oracle.shepmod_predict(x, m.rw, z[sel], cap=256)
    assert oc.max() <= 256
    ref = oracle.weighted_sum(oi, ow, oc, y)
    assert relerr(pred[sel], ref) < RTOL
    ids, wts = m._predict(z[sel])
    for r in range(SAMPLE):
        assert np.array_equal(ids[r], oi[r, :oc[r]])
        assert np.max(np.abs(wts[r] - ow[r, :oc[r]])) < 1e-13


def test_c3_lshep_200k(ub, oracle):
    """configs[2]: LSHEP(radius=2) fit + predict, 200k points, 10-D.  Gradients of sampled nodes against
    the real LAPACK DGELSS on the same weighted systems (linear_shepard.f95:116-134, RCOND = NP*eps),
    rw against the NP-th neighbour distance; predictions of sampled queries against LSHEPVAL
    (f95:365-472) evaluated by the oracle on the fitted arrays (whose gradients the DGELSS sample
    pins).  All at 1e-12."""
    from scipy.linalg import lapack
    n, q, d = 200000, 200000, 10
    x, z = synth(n, q, d)
    y = fun(x)
    m = ub.LSHEP(radius=2)
    m.fit(x, y)
    assert m.fit_ier == 0
    pred = m.predict(z)
    npn = (3 * d + 1) // 2
    rng = np.random.default_rng(6)
    nodes = rng.choice(n, SAMPLE, replace=False)
    idx, d2 = oracle.knn(x, x[nodes], npn + 1)
    a_ref = np.empty((SAMPLE, d))
    for r, k in enumerate(nodes):
        nb, dist = idx[r, 1:], np.sqrt(d2[r, 1:])
        assert m.rw[0][k] == 2 * dist[-1]  # the clamp sqrt(DIAM)/2 is inactive here (:177)
        rp = 1.1 * dist[-1]
        om = (rp - dist) / (rp * dist)
        A = om[:, None] * (x[nb] - x[k])
        b = om * (y[nb] - y[k])
        _, sol, _, rank, _, info = lapack.dgelss(A, b, cond=npn * np.finfo(float).eps)
        assert info == 0 and rank == d
        a_ref[r] = sol[:d]
    got = m.a[0][:, nodes].T
    assert np.max(np.abs(got - a_ref) / np.abs(a_ref).max(axis=1, keepdims=True)) < RTOL
    sel = rng.choice(q, SAMPLE, replace=False)
    ref, ier = oracle.lshep_predict(x, y, m.a[0].T, m.rw[0], z[sel])
    assert relerr(pred[sel], ref) < RTOL
    out2, ier2 = m.model.lshep_predict(z[sel], m.radius, 1)  # the per-query IER codes (f95:381-391)
    assert np.array_equal(ier2[:, 0], ier)
    assert np.array_equal(out2[:, 0], pred[sel])


def test_c4_voronoi_pruned_50k(ub, oracle):
    """configs[3]: NearestNeighbor(k=64, method=Voronoi), 50k x 50k, 32-D -- the reference's own
    composition of kNN pruning and PREDICT (voronoi.f90:57-155) on the neighbours."""
    n, q, d, k = 50000, 50000, 32, 64
    x, z = synth(n, q, d)
    y = fun(x)
    m = ub.NearestNeighbor(k=k, method=ub.Voronoi)
    m.fit(x, y)
    pred = np.array(m.predict(z))
    rng = np.random.default_rng(7)
    sel = rng.choice(q, SAMPLE, replace=False)
    ids, w, cnt, err = oracle.nn_predict(x, z[sel], k, "voronoi")
    assert not err.any()
    ref = oracle.weighted_sum(ids, w, cnt, y)
    assert relerr(pred[sel], ref) < RTOL
    gi, gw = m._predict(z[sel])
    for r in range(SAMPLE):
        assert np.array_equal(gi[r], ids[r, :cnt[r]])
        assert np.max(np.abs(gw[r] - w[r, :cnt[r]])) < 1e-12
    # a second point on the curve: k = d + 1 neighbours
    m2 = ub.NearestNeighbor(k=d + 1, method=ub.Voronoi)
    m2.fit(x, y)
    p2 = np.array(m2.predict(z[sel]))
    ids, w, cnt, err = oracle.nn_predict(x, z[sel], d + 1, "voronoi")
    assert relerr(p2, oracle.weighted_sum(ids, w, cnt, y)) < RTOL
