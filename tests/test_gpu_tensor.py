"""GPU tests of the tensor-core kNN path (tcgen05 fused distance + selection,
fp64 re-rank, certificate, exact fallback): the result must be the exact
(d2, index)-ordered neighbours, bit for bit, whatever the filter does."""
import numpy as np
import pytest

from conftest import fun, synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def L():
    from util_b200 import _lib
    assert _lib.lib().b200_device_count() >= 1
    return _lib


def test_stage1_dots_match_fp64_within_bound(L):
    """The raw stage-1 values must be -2 s^2 (z-mu).(y-mu) up to the fp16 operand rounding;
    checks the UMMA descriptors and the K-major tile images."""
    for (n, q, d) in [(1000, 300, 8), (700, 256, 64), (513, 100, 29), (900, 130, 100)]:
        x, z = synth(n, q, d)
        m = L.Model(x)
        s = m.tc_debug_scores(z, k=1)
        sc = m.stat("tc_scale")
        K = int(m.stat("tc_K"))
        assert K % 16 == 0 and K >= d and K < d + 16
        mu = x.mean(0)
        exact = -2.0 * sc ** 2 * (z - mu) @ (x - mu).T
        err = np.abs(s[:, :n].astype(np.float64) - exact)
        ymax = m.stat("tc_ymax")
        xn = 2.0 * sc * np.sqrt(((z - mu) ** 2).sum(1))[:, None]
        bound = (2 * 2.0 ** -11 + 2.0 ** -16) * 1.01 * xn * ymax + 2.0 ** -25 * (2 * sc * np.abs(z - mu).sum(1)[:, None] + np.sqrt(d) * ymax) + 1e-7
        assert np.all(err <= bound), (n, q, d, (err / bound).max())
        assert np.all(s[:, n:] == 0)  # padding columns are zero rows


def test_tensor_accumulation_error_bound(L):
    """GAMMA_ACC (knn_tc.cu): the certificate assumes |tensor-core sum - exact sum of the fp16 products|
    <= 2^-16 * sum |products|.  Measure it on operands that are exact in fp16 (so operand rounding is
    zero and the whole difference is accumulation error) with adversarial structure: alternating signs,
    magnitudes spanning fp16's normal range, cancellation, for K = 16 ... 192."""
    rng = np.random.default_rng(7)
    worst = 0.0
    for d in (16, 64, 80, 128, 192):
        n, q = 640, 256
        # fp16-exact values: 11-bit integers times powers of two, all normal in fp16 after scaling
        def fp16_exact(shape, emin):
            mant = rng.integers(1024, 2048, size=shape).astype(np.float64)
            ex = rng.integers(emin, 0, size=shape)
            sign = rng.choice([-1.0, 1.0], size=shape)
            return sign * mant * 2.0 ** (ex - 11)
        x = fp16_exact((n, d), -8)
        x[: n // 4] = np.abs(x[: n // 4]) * np.where(np.arange(d) % 2 == 0, 1.0, -1.0)  # alternating signs
        x[n // 4: n // 2, : d // 2] *= 2.0 ** -2                                         # mixed magnitudes
        z = fp16_exact((q, d), -10) * 0.25
        z[: q // 2] = np.abs(z[: q // 2])                                                 # cancellation against x
        # make the centring exact: symmetrise the data so the mean is exactly zero
        x = np.vstack([x, -x])
        m = L.Model(x)
        s = m.tc_debug_scores(z, k=1)[:, : 2 * n].astype(np.float64)
        sc = m.stat("tc_scale")  # set by the first tensor-path call
        xt = -2.0 * z * sc
        yt = x * sc
        assert np.array_equal(xt.astype(np.float16).astype(np.float64), xt)
        assert np.array_equal(yt.astype(np.float16).astype(np.float64), yt)
        exact = xt @ yt.T
        mag = np.abs(xt) @ np.abs(yt).T
        rel = np.abs(s - exact) / mag
        worst = max(worst, rel.max())
        assert rel.max() <= 2.0 ** -16, (d, rel.max())
    print(f"worst accumulation error / sum|terms| = {worst:.3e} (bound 2^-16 = {2.0 ** -16:.3e})")


@pytest.mark.parametrize("n,q,d,k", [(20000, 1000, 64, 16), (30000, 700, 16, 17), (9000, 513, 8, 1),
                                     (12000, 300, 32, 33), (5000, 260, 10, 15), (40000, 300, 3, 8),
                                     (8000, 300, 100, 5), (6000, 130, 125, 3),  # K = 112 / 128: generic K loop, one half per CTA
                                     (5000, 260, 10, 40), (7000, 200, 20, 100),  # 128-entry append-and-trim lists in global memory
                                     (15000, 400, 24, 24), (50000, 600, 32, 64),  # 64-entry append lists; C4's shape
                                     (9000, 200, 100, 33),   # high d: the 64-entry append lists move to global memory
                                     (6000, 300, 12, 110)])  # keep > 116: sorted 128-entry lists in shared memory, one half per CTA
def test_tensor_knn_bit_exact(L, oracle, n, q, d, k):
    x, z = synth(n, q, d)
    m = L.Model(x)
    m.set_option("knn_path", L.PATH_TENSOR)
    idx, d2 = m.knn(z, k)
    oi, od = oracle.knn(x, z, k)
    assert np.array_equal(idx, oi)
    assert np.array_equal(d2, od)
    assert m.stat("tc_queries") == q
    # the filter should certify nearly everything on uniform data
    # (in very low dimension the worst-case error bound is loose against the tiny
    # neighbour gaps: those queries fall back to the exact kernel, still bit-exact)
    if d >= 8:
        assert m.stat("uncertified") <= 0.2 * q
    print(f"uncertified {int(m.stat('uncertified'))}/{q} (n={n}, d={d}, k={k}, cap={int(m.stat('tc_cap'))})")


def test_tensor_knn_uncertified_fallbacks_agree(L, oracle):
    """A few uncertified queries are resolved by the rescan (every point within the candidates' k-th
    distance, ranked exactly); many, or option "fallback" = 1, by the split brute-force scan.  Same answer."""
    x, z = synth(40000, 300, 3)  # 3-D: the error bound is loose against the tiny gaps
    oi, od = oracle.knn(x, z, 8)
    for nq, flags in [(300, 0), (300, 1), (40, 0), (40, 1)]:
        m = L.Model(x)
        m.set_option("knn_path", L.PATH_TENSOR)
        m.set_option("fallback", flags)
        idx, d2 = m.knn(z[:nq], 8)
        assert m.stat("uncertified") > 0
        assert np.array_equal(idx, oi[:nq]) and np.array_equal(d2, od[:nq])


def test_tensor_knn_self_excluding(L, oracle):
    x, _ = synth(15000, 1, 16)
    m = L.Model(x)
    m.set_option("knn_path", L.PATH_TENSOR)
    idx, d2 = m.knn(x[:2000], 17, self_offset=0)
    oi, od = oracle.knn(x, x[:2000], 18)
    assert np.array_equal(idx, oi[:, 1:]) and np.array_equal(d2, od[:, 1:])


def test_tensor_knn_hard_inputs(L, oracle):
    """Clustered / duplicated / far-offset data: the certificate must send what the
    fp16 filter cannot separate to the exact path, never return a wrong index."""
    rng = np.random.default_rng(5)
    # tight clusters far from the origin
    centres = rng.random((20, 12)) * 100 + 1000
    x = np.repeat(centres, 500, axis=0) + rng.normal(scale=1e-3, size=(10000, 12))
    z = np.repeat(centres, 10, axis=0) + rng.normal(scale=1e-3, size=(200, 12))
    m = L.Model(x)
    m.set_option("knn_path", L.PATH_TENSOR)
    idx, d2 = m.knn(z, 8)
    oi, od = oracle.knn(x, z, 8)
    assert np.array_equal(idx, oi) and np.array_equal(d2, od)
    # exact duplicates and lattice ties -> lowest index wins
    g = np.array([[a, b, c] for a in range(16) for b in range(16) for c in range(16)], dtype=float)
    x = np.vstack([g, g])
    z = g[::9] + 0.5
    m = L.Model(x)
    m.set_option("knn_path", L.PATH_TENSOR)
    idx, d2 = m.knn(z, 10)
    oi, od = oracle.knn(x, z, 10)
    assert np.array_equal(idx, oi) and np.array_equal(d2, od)
    # the same with the unsorted append-and-trim lists (k = 24: 64 entries, k = 48: 128): tied scores at
    # the trim threshold are dropped together and the certificate must notice when too few are left
    for k in (24, 48):
        m = L.Model(x)
        m.set_option("knn_path", L.PATH_TENSOR)
        idx, d2 = m.knn(z, k)
        oi, od = oracle.knn(x, z, k)
        assert np.array_equal(idx, oi) and np.array_equal(d2, od)
    # queries far outside the data (fp16 overflow guard) and n not a multiple of 128
    x, z = synth(4999, 130, 5)
    z[:5] *= 1e6
    m = L.Model(x)
    m.set_option("knn_path", L.PATH_TENSOR)
    idx, d2 = m.knn(z, 4)
    oi, od = oracle.knn(x, z, 4)
    assert np.array_equal(idx, oi) and np.array_equal(d2, od)


def test_tensor_knn_round_trip_at_size(L):
    """Size-independent properties at a larger size (no oracle): every fitted row is its own
    nearest neighbour at distance 0, distances ascend, indices are in range and distinct."""
    n, d, k = 200000, 64, 16
    x = np.random.default_rng(0).random((n, d))
    m = L.Model(x)
    m.set_option("knn_path", L.PATH_TENSOR)
    sel = np.random.default_rng(2).choice(n, 4096, replace=False)
    idx, d2 = m.knn(x[sel], k)
    assert np.array_equal(idx[:, 0], sel) and np.all(d2[:, 0] == 0)
    assert np.all(np.diff(d2, axis=1) >= 0)
    assert idx.min() >= 0 and idx.max() < n
    assert all(len(set(r)) == k for r in idx[:200])
    # and d2 is exactly the reference arithmetic on the returned rows
    j = idx[:50]
    ref = np.zeros((50, k))
    for c in range(d):
        t = x[sel[:50], None, c] - x[j, c]
        ref += t * t
    assert np.array_equal(ref, d2[:50])


def test_tensor_knn_full_size_against_exact_kernel(L, oracle):
    """BASELINE C5 data size (1M x 64): the tensor path (norm-ordered images, selection queues, the tail round
    split over idle SMs, rescan fallback) must give, bit for bit, what the CPU oracle gives on a sample of
    the queries and what the fp64 brute-force kernel gives on a larger one, and every fitted row must find
    itself first."""
    n, d, k = 1000000, 64, 16
    x = np.random.default_rng(0).random((n, d))
    q = 148 * 256 + 3000  # one full round of the persistent grid + a short tail round
    z = np.random.default_rng(1).random((q, d))
    sel = np.random.default_rng(2).choice(n, 512, replace=False)
    z[:512] = x[sel]
    m = L.Model(x)
    m.set_option("knn_path", L.PATH_TENSOR)
    idx, d2 = m.knn(z, k)
    assert m.stat("tc_queries") == q
    assert np.array_equal(idx[:512, 0], sel) and np.all(d2[:512, 0] == 0)
    assert np.all(np.diff(d2, axis=1) >= 0) and idx.min() >= 0 and idx.max() < n
    pick = np.concatenate([np.arange(0, 700), np.arange(q - 700, q)])  # both launches
    m.set_option("knn_path", L.PATH_EXACT)
    ei, ed = m.knn(z[pick], k)
    assert np.array_equal(idx[pick], ei) and np.array_equal(d2[pick], ed)
    osel = np.concatenate([np.arange(0, 128), np.arange(q - 128, q)])
    oi, od = oracle.knn(x, z[osel], k)
    assert np.array_equal(idx[osel], oi) and np.array_equal(d2[osel], od)
