"""GPU parity tests: the CUDA path, called through the C-ABI / the drop-in classes,
against the CPU oracle on the same seeded inputs and against the golden vectors
the real reference produced.  Bit-exact for neighbour indices and squared
distances; values within 1e-12 relative (BASELINE.json north_star)."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, NN_CASES, fun, load_golden, synth

pytestmark = pytest.mark.gpu

RTOL = 1e-12  # north_star: predicted values within 1e-12 relative of the reference


def relerr(a, b):
    a, b = np.asarray(a, dtype=float), np.asarray(b, dtype=float)
    return np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-3 * max(1.0, np.abs(b).max())))


@pytest.fixture(scope="module")
def ub():
    import util_b200
    from util_b200 import _lib
    assert _lib.lib().b200_device_count() >= 1
    return util_b200


@pytest.mark.parametrize("path", ["exact", "auto"])
@pytest.mark.parametrize("name", NN_CASES)
def test_nn_golden(ub, oracle, name, path):
    """Drop-in NearestNeighbor vs the outputs of util.approximate.NearestNeighbor."""
    g, x, z = load_golden(name)
    k = int(g["k"])
    y = fun(x)
    m = ub.NearestNeighbor(k=k)
    m.fit(x, y)
    m.model.set_option("knn_path", 1 if path == "exact" else 0)
    pred = m.predict(z)
    assert isinstance(pred, list) and len(pred) == len(z)
    assert np.array_equal(np.array(pred), g["pred"])
    m2 = ub.NearestNeighbor(k=k)
    m2.fit(x)
    m2.model.set_option("knn_path", 1 if path == "exact" else 0)
    pw = m2.predict(z)
    ids = np.array([p[0] for p in pw])
    wts = np.array([p[1] for p in pw])
    assert np.array_equal(ids, g["ids"])          # bit-exact neighbour indices
    assert np.array_equal(wts, g["wts"])
    dist, idx = m2.neighbors(z, return_distance=True)
    assert np.array_equal(dist, g["dist"])        # 0 ulp vs BallTree distances
    single = m.predict(z[0])
    assert single == g["pred"][0]


def test_nn_demo_latin(ub):
    g = np.load(os.path.join(GOLDEN, "nn_demo_latin.npz"))
    m = ub.NearestNeighbor()
    m.fit(g["x"], g["y"])
    assert np.array_equal(np.array(m(g["z"])), g["pred"])


@pytest.mark.parametrize("n,q,d,k", [(3000, 257, 3, 4), (1000, 1, 5, 1), (517, 300, 7, 17),
                                     (200, 50, 2, 200), (5000, 64, 70, 9), (900, 33, 130, 5),
                                     (200000, 20, 16, 16), (150000, 50, 64, 33)])  # few queries: ~1000-way split + warp merge
def test_knn_exact_vs_oracle(ub, oracle, n, q, d, k):
    from util_b200 import _lib
    x, z = synth(n, q, d)
    mdl = _lib.Model(x)
    mdl.set_option("knn_path", 1)
    idx, d2 = mdl.knn(z, k)
    oi, od = oracle.knn(x, z, k)
    assert np.array_equal(idx, oi) and np.array_equal(d2, od)
    # self-excluding scan used by the fits
    if k < n:
        idx, d2 = mdl.knn(x[:100], k, self_offset=0)
        assert not (idx == np.arange(100)[:, None]).any()


def test_knn_ties_lowest_index(ub, oracle):
    from util_b200 import _lib
    # integer lattice: many exactly equal distances
    g = np.array([[a, b, c] for a in range(6) for b in range(6) for c in range(6)], dtype=float)
    x = np.vstack([g, g[:50]])  # duplicated rows too
    z = g[::7] + 0.5
    mdl = _lib.Model(x)
    for path in (1, 0):
        mdl.set_option("knn_path", path)
        idx, d2 = mdl.knn(z, 12)
        oi, od = oracle.knn(x, z, 12)
        assert np.array_equal(idx, oi) and np.array_equal(d2, od)


@pytest.mark.parametrize("method", ["average", "shepard", "voronoi"])
@pytest.mark.parametrize("n,q,d,k", [(2000, 100, 4, 10), (4000, 200, 16, 33), (3000, 64, 32, 64)])
def test_nn_submodels_vs_oracle(ub, oracle, method, n, q, d, k):
    x, z = synth(n, q, d)
    y = fun(x)
    cls = {"average": ub.Average, "shepard": ub.Shepard, "voronoi": ub.Voronoi}[method]
    oi, ow, oc, oe = oracle.nn_predict(x, z, k, method)
    op = oracle.weighted_sum(oi, ow, oc, y)
    m = ub.NearestNeighbor(k=k, method=cls)
    m.fit(x, y)
    pred = np.array(m.predict(z))
    assert relerr(pred, op) < RTOL
    m2 = ub.NearestNeighbor(k=k, method=cls)
    m2.fit(x)
    pw = m2.predict(z)
    for p in range(q):
        ids, wts = pw[p]
        assert np.array_equal(np.asarray(ids), oi[p, :oc[p]])
        assert np.max(np.abs(np.asarray(wts) - ow[p, :oc[p]])) < 1e-12
        assert abs(np.sum(wts) - 1) < 1e-13
    if method != "voronoi":
        assert np.array_equal(pred, op)  # same summation order -> bit-exact


def test_nn_exact_hit_and_edge_cases(ub, oracle):
    x, z = synth(500, 40, 3)
    y = fun(x)
    for cls in (ub.Average, ub.Shepard, ub.Voronoi):
        m = ub.NearestNeighbor(k=6, method=cls)
        m.fit(x, y)
        # interpolation at fitted nodes (weight.f90:25-28; Voronoi weight 1 at its own centre)
        if cls is not ub.Average:
            assert np.allclose(np.array(m(x[:30])), y[:30], rtol=0, atol=1e-12)
    m = ub.NearestNeighbor(k=3)
    m.fit(x, y)
    assert m.predict(np.empty((0, 3))) == []
    with pytest.raises(ub.UnexpectedType):
        m.fit(list(x), y)
    mk = ub.NearestNeighbor(k=501)
    mk.fit(x, y)
    with pytest.raises(ValueError):
        mk.predict(z)
    # NaN / infinite queries are refused on both search paths, as BallTree.query refuses them
    for bad in (np.nan, np.inf):
        zb = z.copy()
        zb[7, 1] = bad
        with pytest.raises(ValueError):
            m.predict(zb)
    xb, zb = synth(20000, 20000, 8)
    mb = ub.NearestNeighbor(k=4)
    mb.fit(xb, fun(xb))
    zb[12345, 3] = np.nan
    with pytest.raises(ValueError):
        mb.predict(zb)
    # duplicated fit points make the Voronoi sub-model raise, as voronoi/__init__.py:56-57 does
    xd = np.vstack([x, x[:5]])
    mv = ub.NearestNeighbor(k=8, method=ub.Voronoi)
    mv.fit(xd, fun(xd))
    with pytest.raises(ub.DuplicateInterpolationPoints):
        mv.predict(x[:5] + 1e-9)
    # non-numeric y goes through the reference's own Python sum on GPU (ids, wts)
    from fractions import Fraction
    yf = [Fraction(int(1000 * v), 1000) for v in y]
    mf = ub.NearestNeighbor(k=2)
    mf.fit(x, yf)
    out = mf(z[:5])
    ref_ids = oracle.knn(x, z[:5], 2)[0]
    assert all(abs(float(out[i]) - sum(float(yf[j]) * .5 for j in ref_ids[i])) < 1e-12 for i in range(5))


@pytest.mark.parametrize("n,q,d", [(1500, 300, 3), (4000, 500, 8), (6000, 300, 16)])
def test_shepmod_vs_oracle(ub, oracle, n, q, d):
    x, z = synth(n, q, d)
    z[:20] = x[:20]                       # exact hits (modified_shepard.f95:136-140)
    z[20:25] = 3.0 + np.arange(5)[:, None]  # outside every radius -> IER 1 fallback (:163-170)
    y = fun(x)
    rw_o, ier = oracle.shepmod_fit(x)
    m = ub.ShepMod()
    m.fit(x, y)
    assert np.array_equal(m.rw, rw_o)     # bit-exact radii
    oi, ow, oc, oe = oracle.shepmod_predict(x, rw_o, z, cap=n)
    op = oracle.weighted_sum(oi, ow, oc, y)
    pred = np.array(m(z))
    assert relerr(pred, op) < RTOL
    assert m.ierrors.get(1, 0) == int((oe == 1).sum()) and m.ierrors.get(0, 0) == int((oe == 0).sum())
    m2 = ub.ShepMod()
    m2.fit(x)
    pw = m2(z)
    for p in range(q):
        ids, wts = pw[p]
        assert np.array_equal(ids, oi[p, :oc[p]])
        assert np.max(np.abs(wts - ow[p, :oc[p]])) < 1e-13
    with pytest.raises(ub.TooFewDataPoints):
        ub.ShepMod().fit(x[:d + 1])


@pytest.mark.parametrize("n,q,d", [(6000, 42000, 64), (5000, 52000, 29)])
def test_shepmod_tensor_filter_high_dimension(ub, oracle, n, q, d):
    """The radius filter on the tensor path (n*q >= 2.5e8) where a 128-row tile is large: its hit lists stay in
    shared memory whatever the kNN kernel of that shape does with its lists.  Sample against the oracle, everything
    against the library's dense scan."""
    from util_b200 import _lib
    x, z = synth(n, q, d)
    z[:50] = x[:50]
    y = fun(x)
    res = []
    for path in (_lib.PATH_AUTO, _lib.PATH_EXACT):
        m = _lib.Model(x)
        m.set_option("knn_path", path)
        rw, ier = m.shepmod_fit()
        res.append(m.shepmod_predict(z, cap=128, y=y))
        m.close()
    (i1, w1, c1, e1, o1), (i0, w0, c0, e0, o0) = res
    assert np.array_equal(c0, c1) and np.array_equal(e0, e1) and np.array_equal(o0, o1)
    sel = np.random.default_rng(3).choice(q, 200, replace=False)
    oi, ow, oc, oe = oracle.shepmod_predict(x, rw, z[sel], cap=n)
    ref = oracle.weighted_sum(oi, ow, oc, y)
    assert relerr(o1[sel, 0], ref) < RTOL
    for r, p in enumerate(sel):
        assert np.array_equal(i1[p, :c1[p]], oi[r, :oc[r]])
        assert np.array_equal(i0[p, :c0[p]], oi[r, :oc[r]])


def test_shepmod_capacity_growth(ub, oracle):
    x, z = synth(3000, 100, 2)
    m = ub.ShepMod()
    m.capacity = 3
    m.fit(x)
    rw_o, _ = oracle.shepmod_fit(x)
    oi, ow, oc, oe = oracle.shepmod_predict(x, rw_o, z, cap=3000)
    pw = m(z)
    assert all(np.array_equal(pw[p][0], oi[p, :oc[p]]) for p in range(100))


@pytest.mark.parametrize("n,q,d,r", [(1200, 200, 3, 1), (3000, 300, 10, 2), (2000, 100, 16, 1)])
def test_lshep_vs_oracle(ub, oracle, n, q, d, r):
    x, z = synth(n, q, d)
    z[:10] = x[:10]
    z[10:13] = 4.0
    Y = np.stack([fun(x) * (c + 1) + c * x[:, 1] for c in range(r)], axis=1)
    m = ub.LSHEP()
    m.fit(x, Y if r > 1 else Y[:, 0])
    out = m(z)
    assert out.shape == ((q, r) if r > 1 else (q,))
    out = out.reshape(q, r)
    ier_hist = {}
    for c in range(r):
        a_o, rw_o, ier, rank = oracle.lshep_fit(x, Y[:, c])
        assert np.array_equal(m.rw[c], rw_o * 2)
        # QR vs DGELSS on full-rank systems: relative to the gradient norm
        assert np.max(np.abs(m.a[c].T - a_o)) / np.max(np.abs(a_o)) < 1e-12
        o_out, o_ier = oracle.lshep_predict(x, Y[:, c], a_o, rw_o * 2, z)
        # evaluation parity on identical fitted state: feed the oracle the GPU's gradients
        g_out, g_ier = oracle.lshep_predict(x, Y[:, c], m.a[c].T, rw_o * 2, z)
        assert relerr(out[:, c], g_out) < RTOL
        # the whole chain -- GPU fit + GPU evaluation against DGELSS fit + oracle evaluation
        assert relerr(out[:, c], o_out) < RTOL
        for e in g_ier:
            ier_hist[int(e)] = ier_hist.get(int(e), 0) + 1
    assert m.ierrors == ier_hist
    with pytest.raises(ub.TooFewDataPoints):
        ub.LSHEP().fit(x[:d + 1], Y[:d + 1, 0])


def test_lshep_rank_deficient_nodes_minimum_norm(ub, oracle):
    """Nodes whose NP nearest neighbours do not span the space (a lattice that lives in a 2-D plane of
    5-D, a collinear cluster): DGELSS returns the minimum-norm gradient and IER = 2
    (linear_shepard.f95:132-135); the wrapper uses that gradient.  The GPU fit must agree with the
    oracle's real DGELSS on every node, flagged ones included."""
    rng = np.random.default_rng(3)
    g = np.array([[a, b] for a in range(30) for b in range(30)], dtype=float) / 30.0
    basis = rng.standard_normal((2, 5))
    plane = g @ basis                                  # every neighbourhood has rank 2 in 5-D
    line = np.outer(np.linspace(0, 1, 60), rng.standard_normal(5)) + 3.0   # rank 1
    blob = rng.random((400, 5)) + 10.0                 # full rank
    x = np.vstack([plane, line, blob])
    f = np.sin(x @ rng.standard_normal(5)) + x[:, 0]
    m = ub.LSHEP()
    m.fit(x, f)
    a_o, rw_o, ier, rank = oracle.lshep_fit(x, f)
    assert ier == 2 and (rank < 5).sum() >= len(plane)           # the oracle's DGELSS sees the deficiency
    assert np.array_equal(m.rw[0], rw_o * 2)
    # Exactly singular systems leave rounding-noise singular values around eps*sigma_1, i.e. right at the
    # cut NP*eps*sigma_1: whether such a value is kept is decided by rounding in the reference too, so
    # nodes are compared where every singular value is either well below the cut or well-conditioned.
    npn = min(len(x) - 1, (3 * 5 + 1) // 2)
    idx, d2 = oracle.knn(x, x, npn + 1)
    clear = np.zeros(len(x), dtype=bool)
    for k in range(len(x)):
        nb, dist = idx[k, 1:], np.sqrt(d2[k, 1:])
        rp = 1.1 * dist[-1]
        om = (rp - dist) / (rp * dist)
        sv = np.linalg.svd(om[:, None] * (x[nb] - x[k]), compute_uv=False)
        ratio = sv / sv[0]
        cut = npn * np.finfo(float).eps
        clear[k] = not np.any((ratio > cut / 3) & (ratio < 1e-8))  # also no kept direction that is pure rounding noise
    assert (clear & (rank < 5)).sum() > 200 and (clear & (rank == 5)).sum() > 300
    scale = np.maximum(np.abs(a_o).max(axis=1, keepdims=True), 1e-3 * np.abs(a_o[clear]).max())
    err = np.abs(m.a[0].T - a_o) / scale
    assert err[clear].max() < 1e-11, err[clear].max()
    zc = x[clear][::5] + 1e-3
    out = m(zc)
    o_out, _ = oracle.lshep_predict(x, f, m.a[0].T, rw_o * 2, zc)
    assert relerr(out, o_out) < RTOL
    assert m.fit_ier == 2


@pytest.mark.parametrize("n,d", [(3000, 3), (4000, 10)])
def test_lshep_robust_fit_vs_oracle(ub, oracle, n, d):
    """LSHEP(robust=True): the RLSHEP branch (linear_shepard.f95:136-175; IRLS :181-282, MAD :284-308,
    QUANTILE :310-335) against the oracle's restatement (real DGELSS inside).  IRLS is a chain of eleven
    least-squares solves with discrete decisions (median picks, the 0.8 weight cut, the bisquare
    convergence test), so a node whose decision sits on a rounding tie may differ; all but a handful must
    agree to 1e-10 in the gradients and exactly in RW."""
    x, z = synth(n, 300, d)
    rng = np.random.default_rng(9)
    f = fun(x) + 0.05 * rng.standard_normal(n)
    f[::23] += 3.0  # gross outliers
    m = ub.LSHEP(robust=True)
    m.fit(x, f)
    a_o, rw_o, ier_o, om = oracle.lshep_fit_robust(x, f)
    got = m.a[0].T
    scale = np.maximum(np.abs(a_o).max(axis=1, keepdims=True), 1e-3 * np.abs(a_o).max())
    err = (np.abs(got - a_o) / scale).max(axis=1)
    same_rw = m.rw[0] == rw_o * 2
    bad = (err > 1e-10) | ~same_rw
    assert bad.sum() <= max(2, n // 500), (bad.sum(), err.max())
    assert m.fit_ier == ier_o or bad.any()
    out = m(z)
    o_out, _ = oracle.lshep_predict(x, f, got, m.rw[0], z)
    assert relerr(out, o_out) < RTOL


@pytest.mark.parametrize("n,d", [(1500, 2), (2500, 5)])
def test_ripple_fit_vs_oracle(ub, oracle, n, d):
    """LSHEP(ripple=True): RIPPLE (linear_shepard.f95:474-665) against the oracle's restatement.  The start
    is the argmin over ~NP (M+1)(M+2)/2 small least-squares residuals (QR here, DGELSS there) followed by ten
    IRLS solves with discrete decisions; nodes whose choice sits on a rounding tie may pick another
    subset, so all but a small fraction must agree to 1e-9 in the gradients and exactly in RW."""
    x, z = synth(n, 200, d)
    rng = np.random.default_rng(4)
    f = fun(x) + 0.02 * rng.standard_normal(n)
    f[::19] += 2.0
    m = ub.LSHEP(ripple=True)
    m.fit(x, f)
    a_o, rw_o, ier_o = oracle.ripple(x, f)
    got = m.a[0].T
    scale = np.maximum(np.abs(a_o).max(axis=1, keepdims=True), 1e-3 * np.abs(a_o).max())
    err = (np.abs(got - a_o) / scale).max(axis=1)
    bad = (err > 1e-9) | (m.rw[0] != rw_o * 2)
    assert bad.mean() < 0.01, (bad.sum(), err.max())
    out = m(z)
    o_out, _ = oracle.lshep_predict(x, f, got, m.rw[0], z)
    assert relerr(out, o_out) < RTOL


def test_standalone_shepard_voronoi_demo_sizes(ub, oracle):
    # the reference's own demos: test_plot(model, N=20) (shepard/__init__.py:53, voronoi/__init__.py:68)
    x, z = synth(20, 60, 2)
    y = fun(x)
    ms = ub.Shepard()
    ms.fit(x, y)
    ps = np.array(ms(z))
    ref = np.array([sum(y[i] * w for i, w in enumerate(oracle.shepard_weight(x, p))) for p in z])
    assert relerr(ps, ref) < RTOL
    mv = ub.Voronoi()
    mv.fit(x, y)
    pv = np.array(mv(z))
    refv = []
    for p in z:
        w, e = oracle.voronoi_predict(x, p)
        refv.append(sum(y[i] * w[i] for i in range(20) if w[i] > 0))
    assert relerr(pv, np.array(refv)) < RTOL


def test_standalone_voronoi_all_points(ub, oracle):
    """Voronoi() over all fitted points beyond what one block's shared memory holds (voronoi_all_kernel against
    the device-resident n x n inner-product cache): supporting points, weights and predictions against the
    oracle's PREDICT; interpolation at nodes; duplicated fit points raise as voronoi/__init__.py:56-57."""
    for (n, q, d) in [(300, 40, 5), (700, 16, 2)]:
        x, z = synth(n, q, d)
        y = fun(x)
        m = ub.Voronoi()
        m.fit(x)
        got = m(z)
        for p, (ids, wts) in zip(z, got):
            w, e = oracle.voronoi_predict(x, p)
            assert e == 0
            assert np.array_equal(ids, np.nonzero(w > 0)[0])
            assert relerr(wts, w[w > 0]) < RTOL
        m2 = ub.Voronoi()
        m2.fit(x, y)
        pv = np.array(m2(z))
        ref = np.array([sum(y[i] * w for i, w in zip(ids, wts)) for ids, wts in got])
        assert relerr(pv, ref) < RTOL
        at_nodes = np.array(m2(x[:10]))
        assert relerr(at_nodes, y[:10]) < RTOL
    xd = np.vstack([x, x[:1]])
    md = ub.Voronoi()
    md.fit(xd)
    with pytest.raises(ub.DuplicateInterpolationPoints):
        md(z[:2])


@pytest.mark.parametrize("n,d,qs", [(5000, 6, 24), (20000, 4, 12)])
def test_standalone_voronoi_beyond_4096_points(ub, oracle, n, d, qs):
    """Voronoi() over all n fitted points for n beyond one block's shared memory (persistent blocks, Gram
    cache in HBM): supports, weights and predictions against PREDICT (voronoi.f90:57-155) as the oracle
    restates it; SURVEY.md 8d asks for the N = 5 000 point."""
    x, z = synth(n, qs, d)
    y = fun(x)
    m = ub.Voronoi()
    m.fit(x, y)
    pred = np.array(m(z))
    gi, gw = m._predict(z)
    ref = []
    for r, p in enumerate(z):
        w, e = oracle.voronoi_predict(x, p)
        assert e == 0
        sup = np.nonzero(w > 0)[0]
        assert np.array_equal(gi[r], sup)
        assert np.max(np.abs(gw[r] - w[sup])) < 1e-12
        ref.append(sum(y[i] * w[i] for i in sup))
    assert relerr(pred, np.array(ref)) < RTOL


def test_nn_with_any_weighted_submodel(ub, oracle):
    """NearestNeighbor(method=cls) instantiates whatever WeightedApproximator class it is given
    (nearest_neighbor/__init__.py:78-83; SURVEY.md 3.1 lists ShepMod, which needs k >= d + 2)."""
    x, z = synth(600, 40, 3)
    y = fun(x)
    k = 8
    m = ub.NearestNeighbor(k=k, method=ub.ShepMod)
    m.fit(x, y)
    pred = np.array(m(z))
    ids, w, cnt, _ = oracle.nn_predict(x, z, k, "shepmod")
    ref = oracle.weighted_sum(ids, w, cnt, y)
    assert relerr(pred, ref) < RTOL
    gi, gw = m._predict(z)
    for r in range(len(z)):
        assert np.array_equal(gi[r], ids[r, :cnt[r]])


def test_auto_k_selection(ub):
    """nearest_neighbor.auto (nearest_neighbor/__init__.py:93-132): k=None picks a power of two from the
    neighbours of 100 sampled points; a response that only depends on the nearest point wants k = 1."""
    import random
    x, z = synth(4000, 10, 4)
    y = fun(x)
    random.seed(3)
    m = ub.NearestNeighbor(k=None)
    m.fit(x, y, display=False)
    assert m.num_neighbors in (1, 2, 4, 8, 16, 32, 64, 128, 256, 512, 1024, 2048)
    assert m.num_neighbors <= 8          # smooth low-noise response: small k wins
    assert len(m.predict(z)) == 10
    k = ub.auto(x, y, max_k=16, samples=50)
    assert k in (1, 2, 4, 8, 16)
    assert ub.auto(x, y, max_k=10, samples=50) in (1, 2, 4, 8)  # the reference raises here (log2 unbound, :112)


def test_auto_matches_reference_fixture(ub):
    """auto() against tests/golden/auto_k.npz: the k chosen by the reference's own function
    (nearest_neighbor/__init__.py:93-132, run by tests/golden/make_golden.py) on the same data and the
    same sampled rows -- including n > 2048 (more than 1024 neighbours per sampled row) and the default
    max_k = largest power of two <= n/2."""
    import os
    import random
    from conftest import GOLDEN
    g = np.load(os.path.join(GOLDEN, "auto_k.npz"))
    cases = [dict(n=300, d=2, kw={}), dict(n=2000, d=5, kw=dict(mean=False)), dict(n=2500, d=3, kw=dict(k_step=2)),
             dict(n=900, d=4, kw=dict(samples=40)), dict(n=6000, d=3, kw=dict(samples=30)),
             dict(n=40000, d=6, kw=dict(samples=50, mean=False))]
    for i, c in enumerate(cases):
        x = np.random.default_rng(100 + i).random((c["n"], c["d"]))
        assert x.sum() == float(g[f"x_sum{i}"]), "numpy RNG stream changed"
        y = fun(x) + (0.3 * np.random.default_rng(200 + i).standard_normal(c["n"]) if i % 2 else 0.0)
        kw = dict(c["kw"])
        assert ub.auto(x, y, indices=g[f"rows{i}"], **kw) == int(g[f"k{i}"]), (i, c)
        if c["n"] <= 2 ** 15:
            # the default sampler draws the same rows as the reference under the same `random` seed
            random.seed(i)
            assert ub.auto(x, y, **kw) == int(g[f"k{i}"]), (i, c)


def test_standalone_shepard_all_points(ub, oracle):
    """Shepard() over every fitted point (shepard/__init__.py:35-47 + weight.f90) at a size where the
    dense loop matters, incl. exact hits (weight.f90:25-28) and the dense weight matrix."""
    x, z = synth(3000, 120, 5)
    z[:7] = x[10:17]
    y = fun(x)
    m = ub.Shepard()
    m.fit(x, y)
    pred = np.array(m(z))
    W = np.array([oracle.shepard_weight(x, p) for p in z])
    ref = np.array([sum(y[i] * w for i, w in enumerate(row) if w != 0) for row in W])
    assert relerr(pred, ref) < RTOL
    assert np.array_equal(pred[:7], y[10:17])
    m2 = ub.Shepard()
    m2.fit(x)
    pw = m2(z[:20])
    ids, wts = pw[3]
    assert np.array_equal(ids, np.arange(3000)) and np.max(np.abs(wts - W[3])) < 1e-15
    assert all(abs(w.sum() - 1) < 1e-12 for _, w in pw)


def test_nn_predict_chunked_upload_matches_single_shot():
    """Large batches from page-locked memory are uploaded in chunks under the previous chunk's kernels
    (api.cu b200_nn_predict); the result must not depend on it."""
    import torch
    from util_b200 import _lib
    k = 3
    # several chunks / one short chunk and the rest; d = 4: exact kernel, d = 5: tensor path, whose
    # uncertified queries are resolved together after the last chunk
    for n, d, q in ((5000, 4, 148 * 256 * 6 + 777), (5000, 4, 148 * 256 * 2 + 500), (6000, 5, 148 * 256 * 3 + 99)):
        x, z = synth(n, q, d)
        y = fun(x)
        m = _lib.Model(x)
        ref = m.nn_predict(z, k, _lib.SHEPARD, y=y, want_weights=True)          # pageable: one copy
        zp = torch.from_numpy(z).pin_memory().numpy()
        unc0 = m.stat("uncertified")
        got = m.nn_predict(zp, k, _lib.SHEPARD, y=y, want_weights=True)         # pinned: chunked
        for a, b in zip(ref, got):
            assert np.array_equal(a, b)
        if d == 5:
            assert m.stat("uncertified") > unc0  # the deferred path was exercised


def test_wrappers_around_gpu_models(ub, oracle):
    """Split / unique / condition (util/approximate/wrappers.py) around the CUDA-backed models."""
    x, z = synth(3000, 50, 4)
    y = fun(x)
    # Split: mean of k models, each on every k-th row of the seeded shuffle (:21-36)
    s = ub.Split(ub.NearestNeighbor, k=3, seed=1)
    s.fit(x, y)
    np.random.seed(1)
    order = np.arange(len(x))
    np.random.shuffle(order)
    parts = []
    for i in range(3):
        rows = order[i::3]
        ids, w, cnt, _ = oracle.nn_predict(x[rows], z, 1, "average")
        parts.append(oracle.weighted_sum(ids, w, cnt, y[rows]))
    assert relerr(s(z), np.mean(parts, axis=0)) < RTOL
    # unique: duplicated rows collapse to their first occurrence, responses averaged (:52-80)
    xd = np.vstack([x[:500], x[:100]])
    yd = np.concatenate([y[:500], y[:100] + 2.0])
    U = ub.unique(ub.NearestNeighbor)
    m = U(k=1)
    m.fit(xd, yd)
    assert len(m.unique_indices) == 500
    assert relerr(np.array(m(x[:100])), y[:100] + 1.0) < RTOL
    # condition with an explicit map: neighbours are found in the mapped space (:172,183)
    A = np.diag([2.0, 1.0, 0.5, 0.0])
    C = ub.condition(ub.NearestNeighbor, conditioner=A)
    c = C(k=5)
    c.fit(x, y)
    ids, w, cnt, _ = oracle.nn_predict(x @ A, z @ A, 5, "average")
    assert relerr(np.array(c(z)), oracle.weighted_sum(ids, w, cnt, y)) < RTOL


def test_eval_mesh_bit_exact(oracle):
    """b200_eval_mesh against the oracle's EVAL_MESH: same operand order, so the same bits."""
    from util_b200.voronoi import eval_mesh
    for (n, q, d) in [(40, 12, 3), (333, 70, 6)]:
        x, z = synth(n, q, d)
        ip, ap = x @ x.T, z @ x.T
        got = eval_mesh(ip, ap)
        assert np.array_equal(got, oracle.eval_mesh(ip, ap))


@pytest.mark.parametrize("shift,scale", [(1e6, 1.0), (0.0, 1e-6), (-3e3, 50.0)])
def test_shepmod_lshep_prefilter_on_shifted_scaled_data(ub, oracle, shift, scale):
    """The fp32 prefilter of the ShepMod / LSHEP scans must never change a result: far-offset data
    (coordinates huge against the distances: its rounding bound swallows the radii and everything goes to
    the exact path), tiny scales, exact hits, queries outside every radius, clustered nodes."""
    rng = np.random.default_rng(7)
    n, q, d = 2500, 160, 5
    x = rng.random((n, d))
    x[:400] = x[0] + 1e-3 * rng.random((400, d))   # a tight cluster: very small radii next to large ones
    z = rng.random((q, d))
    z[:15] = x[500:515]                              # exact hits
    z[15:20] = 5.0                                   # outside every radius
    z[20:60] = x[0] + 1e-3 * rng.random((40, d))    # inside the cluster
    x = x * scale + shift
    z = z * scale + shift
    y = fun((x - shift) / scale)
    rw_o, _ = oracle.shepmod_fit(x)
    m = ub.ShepMod()
    m.fit(x)
    assert np.array_equal(m.rw, rw_o)
    oi, ow, oc, oe = oracle.shepmod_predict(x, rw_o, z, cap=n)
    pw = m(z)
    for p in range(q):
        assert np.array_equal(pw[p][0], oi[p, :oc[p]])
        assert np.max(np.abs(pw[p][1] - ow[p, :oc[p]])) < 1e-13
    assert m.ierrors.get(1, 0) == int((oe == 1).sum())
    ml = ub.LSHEP()
    ml.fit(x, y)
    a_o, rwl, _, _ = oracle.lshep_fit(x, y)
    g_out, g_ier = oracle.lshep_predict(x, y, ml.a[0].T, rwl * 2, z)
    assert relerr(np.asarray(ml(z)), g_out) < RTOL
    hist = {}
    for e in g_ier:
        hist[int(e)] = hist.get(int(e), 0) + 1
    assert ml.ierrors == hist
