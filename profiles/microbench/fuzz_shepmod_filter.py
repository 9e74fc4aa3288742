"""Randomised cross-check of ShepMod predict on the tensor path (radius filter + hit evaluation + handed-back
queries) against the dense fp64 scan of the same library (itself pinned to the oracle by the test suite):
ids, weights, counts, IER codes and fused sums must be identical bit for bit.
    python profiles/microbench/fuzz_shepmod_filter.py [cases] [seed]
"""
import sys
import numpy as np
sys.path.insert(0, ".")
from util_b200 import _lib

cases = int(sys.argv[1]) if len(sys.argv) > 1 else 24
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 0)
bad = 0
for c in range(cases):
    d = int(rng.choice([5, 8, 13, 16, 29, 32, 61]))
    n = int(rng.integers(5000, 60000))
    q = int(max(2.6e8 // n + 1, rng.choice([5000, 20000, 60000])))
    kind = rng.choice(["uniform", "normal", "clusters", "dups", "lattice", "onnodes"])
    if kind == "uniform":
        x = rng.random((n, d)); z = rng.random((q, d))
    elif kind == "normal":
        x = rng.normal(size=(n, d)); z = rng.normal(size=(q, d))
    elif kind == "clusters":
        cen = rng.random((30, d)) * 10
        x = cen[rng.integers(0, 30, n)] + 0.01 * rng.normal(size=(n, d))
        z = cen[rng.integers(0, 30, q)] + 0.01 * rng.normal(size=(q, d))
    elif kind == "dups":
        base = rng.random((n // 3 + 1, d))
        x = base[rng.integers(0, len(base), n)]
        z = rng.random((q, d)); z[: q // 4] = x[rng.integers(0, n, q // 4)]
    elif kind == "onnodes":  # queries on and next to fitted nodes: exact-hit rule, tiny distances
        x = rng.random((n, d)); z = x[rng.integers(0, n, q)] + rng.choice([0.0, 1e-12, 1e-6], size=(q, 1))
    else:
        x = rng.integers(0, 6, size=(n, d)).astype(float); z = rng.integers(0, 6, size=(q, d)) + 0.5
    scale = float(rng.choice([1.0, 1e-4, 1e3])); shift = float(rng.choice([0.0, 0.0, 50.0, -1e4]))
    x = x * scale + shift; z = z * scale + shift
    y = np.sin(x[:, 0]) + x[:, -1]
    res = []
    try:
        for path in (_lib.PATH_EXACT, _lib.PATH_AUTO):
            m = _lib.Model(x)
            m.set_option("knn_path", path)
            m.shepmod_fit()
            ids, wts, cnt, ier, out = m.shepmod_predict(z, y=y)
            res.append((ids, wts, cnt, ier, out, m.stat("tc_queries")))
            m.close()
    except Exception as e:
        print(f"case {c}: {kind} n={n} q={q} d={d}: {e}")
        continue
    (i0, w0, c0, e0, o0, _), (i1, w1, c1, e1, o1, tq) = res
    ok = np.array_equal(c0, c1) and np.array_equal(e0, e1) and np.array_equal(o0, o1, equal_nan=True)
    if ok:
        for r in rng.choice(q, 400, replace=False):
            ok &= np.array_equal(i0[r, :c0[r]], i1[r, :c1[r]]) and np.array_equal(w0[r, :c0[r]], w1[r, :c1[r]])
    bad += not ok
    print(f"case {c}: {kind:8s} n={n:6d} q={q:6d} d={d:3d} scale={scale:g} shift={shift:g} mean hits {c0.mean():7.1f} "
          f"max {c0.max():5d} ier!=0 {int((e0 != 0).sum()):6d}  {'ok' if ok else 'MISMATCH'}", flush=True)
print("mismatches:", bad)
