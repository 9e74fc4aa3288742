"""Randomised cross-check of the tensor kNN path against the fp64 brute-force kernel (itself pinned to the
oracle by the test suite): shapes, k (sorted, append-and-trim and 128-entry lists), scales, offsets, clustered
and duplicated data.
    python profiles/microbench/fuzz_knn.py [cases] [seed]
"""
import sys
import numpy as np
sys.path.insert(0, ".")
from util_b200 import _lib

cases = int(sys.argv[1]) if len(sys.argv) > 1 else 40
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 0)
bad = 0
for c in range(cases):
    d = int(rng.choice([2, 3, 5, 8, 13, 16, 29, 32, 61, 64, 77, 100, 125]))
    n = int(rng.integers(4200, 120000))
    q = int(rng.choice([130, 700, 5000, 40000]))
    k = int(rng.choice([1, 2, 5, 16, 17, 24, 32, 33, 48, 64, 100]))
    kind = rng.choice(["uniform", "normal", "clusters", "dups", "lattice"])
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
    else:
        x = rng.integers(0, 6, size=(n, d)).astype(float); z = rng.integers(0, 6, size=(q, d)) + 0.5
    scale = float(rng.choice([1.0, 1e-4, 1e3])); shift = float(rng.choice([0.0, 0.0, 50.0, -1e4]))
    x = x * scale + shift; z = z * scale + shift
    flags = 0
    m = _lib.Model(x)
    m.set_option("knn_path", _lib.PATH_TENSOR)
    try:
        it, dt = m.knn(z, k)
    except Exception as e:  # unsupported shape for the tensor path is fine, anything else is not
        print(f"case {c}: {kind} n={n} q={q} d={d} k={k} flags={flags}: {e}")
        continue
    unc = int(m.stat("uncertified"))
    m.set_option("knn_path", _lib.PATH_EXACT)
    sel = rng.choice(q, min(q, 600), replace=False)
    ie, de = m.knn(z[sel], k)
    ok = np.array_equal(it[sel], ie) and np.array_equal(dt[sel], de)
    bad += not ok
    print(f"case {c}: {kind:8s} n={n:6d} q={q:5d} d={d:3d} k={k:2d} scale={scale:g} shift={shift:g} flags={flags:2d} "
          f"uncertified {unc:5d}  {'ok' if ok else 'MISMATCH'}")
    m.close()
print("mismatches:", bad)
sys.exit(1 if bad else 0)
