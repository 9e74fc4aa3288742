"""CPU oracle vs the reference's golden vectors and the invariants its source implies
(SURVEY.md 8c pins i-vi).  No GPU needed."""
import numpy as np
import pytest

from conftest import NN_CASES, fun, load_golden, synth, GOLDEN
import os


@pytest.mark.parametrize("name", NN_CASES)
def test_knn_matches_reference_nearest_neighbor(oracle, name):
    # pin (vi): real util.approximate.NearestNeighbor (sklearn BallTree) index for index
    g, x, z = load_golden(name)
    k = int(g["k"])
    idx, d2 = oracle.knn(x, z, k)
    assert np.array_equal(idx, g["ids"])
    assert np.array_equal(np.sqrt(d2), g["dist"])  # 0 ulp vs BallTree distances
    ids, w, cnt, err = oracle.nn_predict(x, z, k, "average")
    assert np.array_equal(ids, g["ids"]) and np.array_equal(w, g["wts"])
    pred = oracle.weighted_sum(ids, w, cnt, fun(x))
    assert np.array_equal(pred, g["pred"])  # bit-exact left-to-right sum (base.py:169)


def test_demo_latin(oracle):
    g = np.load(os.path.join(GOLDEN, "nn_demo_latin.npz"))
    ids, w, cnt, _ = oracle.nn_predict(g["x"], g["z"], 1, "average")
    assert np.array_equal(oracle.weighted_sum(ids, w, cnt, g["y"]), g["pred"])


def test_pairwise_gemm_form_pin(oracle):
    # pin (i): util/math/pairs.py:64-83 -- GEMM-form distance within 2^-32 of direct differences
    g = np.load(os.path.join(GOLDEN, "pairwise.npz"))
    a, b = g["a"], g["b"]
    idx, d2 = oracle.knn(b, a, b.shape[0])
    for i in range(a.shape[0]):
        assert np.all(np.abs(np.sqrt(d2[i]) - g["d"][i, idx[i]]) < 2.0 ** -32)


def test_shepard_weight_branches(oracle):
    pts = np.array([[0., 0.], [1., 0.], [0., 1.], [1., 1.]])
    w = oracle.shepard_weight(pts, np.array([.25, .25]))
    d = np.sqrt(((pts - [.25, .25]) ** 2).sum(1))
    assert np.allclose(w, (1 / d ** 2) / (1 / d ** 2).sum(), rtol=1e-15) and abs(w.sum() - 1) < 1e-15
    w = oracle.shepard_weight(pts, np.array([1., 0.]))          # weight.f90:25-28 exact hit
    assert np.array_equal(w, [0, 1, 0, 0])
    w = oracle.shepard_weight(np.zeros((3, 2)), np.zeros(2))    # weight.f90:22-24 all coincident
    assert np.allclose(w, 1 / 3)


def test_shepmod_invariants(oracle):
    x, z = synth(1500, 200, 3)
    rw, ier = oracle.shepmod_fit(x)
    assert ier == 0
    # pin (iv): rw[k] is the (M+1)-th smallest non-self distance
    D = np.sqrt(((x[:, None, :] - x[None, :, :]) ** 2).sum(-1))
    np.fill_diagonal(D, np.inf)
    assert np.allclose(np.sort(D, axis=1)[:, 3], rw, rtol=1e-15)
    ids, w, cnt, ie = oracle.shepmod_predict(x, rw, z, cap=128)
    assert cnt.max() <= 128
    for p in range(len(z)):
        ww = w[p, :cnt[p]]
        assert np.all(ww > 0) and abs(ww.sum() - 1) < 1e-14          # pin (iii)
        assert np.all(np.diff(ids[p, :cnt[p]]) > 0)
    # pin (ii): predicting at a node is one-hot on that node
    ids, w, cnt, ie = oracle.shepmod_predict(x, rw, x[:50], cap=8)
    assert np.array_equal(cnt, np.ones(50)) and np.array_equal(ids[:, 0], np.arange(50)) and np.all(w[:, 0] == 1)
    # too few points (modified_shepard.f95:37-40)
    assert oracle.shepmod_fit(x[:4])[1] == 1
    # fallback: far-away query -> IER 1, M+1 positive weights (f95:163-170)
    ids, w, cnt, ie = oracle.shepmod_predict(x, rw, np.full((1, 3), 5.0), cap=8)
    assert ie[0] == 1 and cnt[0] == 4 and abs(w[0, :4].sum() - 1) < 1e-15


def test_lshep_invariants(oracle):
    x, z = synth(1200, 150, 3)
    # a linear function is reproduced exactly by local linear fits
    f = 2 * x[:, 0] - x[:, 1] + .5 * x[:, 2] + 1
    a, rw, ier, rank = oracle.lshep_fit(x, f)
    assert ier == 0 and rank.min() == 3
    assert np.allclose(a, [2, -1, .5], atol=1e-9)
    out, ie = oracle.lshep_predict(x, f, a, rw * 2, z)
    assert np.allclose(out, 2 * z[:, 0] - z[:, 1] + .5 * z[:, 2] + 1, atol=1e-9)
    # QR / lstsq agreement on the weighted systems (SURVEY probe: ~1e-14)
    y = fun(x)
    a, rw, ier, rank = oracle.lshep_fit(x, y)
    k = 7
    npn = (3 * 3 + 1) // 2
    idx, d2 = oracle.knn(x, x[k:k + 1], npn + 1)
    nb = idx[0, 1:]
    dist = np.sqrt(d2[0, 1:])
    om = (1.1 * dist[-1] - dist) / (1.1 * dist[-1] * dist)
    sol = np.linalg.lstsq(om[:, None] * (x[nb] - x[k]), om * (y[nb] - y[k]), rcond=None)[0]
    assert np.allclose(a[k], sol, rtol=1e-10)
    D = ((x[:, None, :] - x[None, :, :]) ** 2).sum(-1)
    assert rw[k] == min(np.sqrt(D.max()) / 2, dist[-1])
    # interpolation at nodes (f95:428-431)
    out, ie = oracle.lshep_predict(x, y, a, rw * 2, x[:40])
    assert np.array_equal(out, y[:40])


def test_voronoi_closed_form(oracle):
    # SURVEY 3.4 probe: control flow is pruning only; closed form agrees to ~1e-13
    for d in (2, 8, 32):
        x, z = synth(120, 20, d)
        for p in z:
            w, err = oracle.voronoi_predict(x, p)
            assert err == 0 and abs(w.sum() - 1) < 1e-14 and np.all(w >= 0)
            num = ((p - x)[:, None, :] * (x[None, :, :] - x[:, None, :])).sum(-1)
            den = ((x[None, :, :] - x[:, None, :]) ** 2).sum(-1)
            np.fill_diagonal(den, 1)
            r = np.maximum(num / den, 0)
            np.fill_diagonal(r, 0)
            wc = np.maximum(0, 1 - r.max(1))
            wc /= wc.sum()
            assert np.array_equal(w > 0, wc > 0) and np.abs(w - wc).max() < 1e-12
    # duplicates -> ERROR 2 (voronoi.f90:145-148)
    x2 = np.vstack([x[:10], x[:1]])
    assert oracle.voronoi_predict(x2, z[0])[1] == 2


def test_nn_submodels(oracle):
    x, z = synth(2000, 100, 4)
    y = fun(x)
    for m in ("shepard", "voronoi", "shepmod"):
        ids, w, cnt, err = oracle.nn_predict(x, z, 10, m)
        assert np.all(err == 0) or m == "shepmod"
        for p in range(len(z)):
            assert abs(w[p, :cnt[p]].sum() - 1) < 1e-13 and np.all(w[p, :cnt[p]] > 0)
        if m == "shepard":
            assert np.all(cnt == 10)


def test_oracle_eval_mesh_agrees_with_predict(oracle):
    """EVAL_MESH (voronoi.f90:5-44) from inner products and PREDICT (:57-155) from coordinates are the same
    mesh up to rounding (different operand grouping); rows are convex combinations."""
    x, z = synth(40, 12, 3)
    mesh = oracle.eval_mesh(x @ x.T, z @ x.T)
    assert mesh.shape == (12, 40) and np.all(mesh >= 0)
    assert np.allclose(mesh.sum(1), 1.0, rtol=0, atol=1e-14)
    for p, row in zip(z, mesh):
        w, e = oracle.voronoi_predict(x, p)
        assert e == 0 and np.allclose(row, w, rtol=1e-10, atol=1e-13)


def test_oracle_robust_lshep_invariants(oracle):
    """Robust LSHEP pieces (linear_shepard.f95:136-175, 181-335) -- parity unpinned, so the invariants the
    source implies: QUANTILE is the k-th order statistic (with ties), MAD = median |r - median r| / 0.6745,
    IRLS weights lie in [0, 1], gross outliers get small weights and shrink RW, clean linear data keeps the
    plane's gradient."""
    rng = np.random.default_rng(0)
    for a in (rng.standard_normal(15), np.round(rng.standard_normal(20), 1), np.zeros(7)):
        for k in range(1, len(a) + 1):
            assert oracle.quantile(a, k) == np.sort(a)[k - 1]
    for n in (15, 16, 5):
        r = rng.standard_normal(n)
        assert oracle.mad(r) == pytest.approx(np.median(np.abs(r - np.median(r))) / 0.6745, rel=1e-15)
    x = rng.random((600, 3))
    g = np.array([1.0, 2.0, -3.0])
    f = x @ g + 0.5
    f[::17] += 5.0
    a, rw, ier, om = oracle.lshep_fit_robust(x, f)
    a0, rw0, _, _ = oracle.lshep_fit(x, f)
    assert om.min() >= 0.0 and om.max() <= 1.0
    clean = np.ones(600, dtype=bool)
    clean[::17] = False
    off = lambda grads: (np.abs(grads[clean] - g).max(axis=1) > 1e-6).mean()
    assert off(a0) > 0.15                                               # the plain fit is pulled by outlier neighbours
    assert off(a) < 0.75 * off(a0)                                      # the robust fit less often (NP = 5 leaves little redundancy)
    assert np.all(rw <= rw0 + 1e-15) and (rw < rw0).mean() > 0.1        # RW is cut at down-weighted neighbours


def test_oracle_ripple_invariants(oracle):
    """RIPPLE (linear_shepard.f95:474-665), parity unpinned: a noise-free plane is reproduced, gross outliers
    pull far fewer nodes than the plain LSHEP fit, radii only shrink."""
    rng = np.random.default_rng(0)
    x = rng.random((800, 3))
    g = np.array([1.0, 2.0, -3.0])
    f = x @ g + 0.5
    a, rw, ier = oracle.ripple(x, f)
    assert np.abs(a - g).max() < 1e-10
    f2 = f.copy()
    f2[::13] += 4.0
    a2, rw2, _ = oracle.ripple(x, f2)
    a0, rw0, _, _ = oracle.lshep_fit(x, f2)
    clean = np.ones(800, dtype=bool)
    clean[::13] = False
    off = lambda grads: (np.abs(grads[clean] - g).max(axis=1) > 1e-6).mean()
    assert off(a0) > 0.2 and off(a2) < 0.05
    assert np.all(rw2 <= rw0 * (1 + 1e-15))
