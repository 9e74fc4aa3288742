"""Time the fp64 brute-force kNN path (knn_path = exact) at a few shapes."""
import sys, time, numpy as np
sys.path.insert(0, ".")
from util_b200 import _lib
for (n, q, d, k) in [(100000, 100000, 8, 16), (200000, 50000, 3, 8), (50000, 50000, 32, 16), (100000, 20000, 64, 16), (10000, 1000, 8, 1)]:
    x = np.random.default_rng(0).random((n, d)); z = np.random.default_rng(1).random((q, d))
    m = _lib.Model(x); m.set_option("knn_path", _lib.PATH_EXACT)
    m.knn(z[:2000], k)
    t0 = time.perf_counter(); idx, d2 = m.knn(z, k); t1 = time.perf_counter()
    print(f"n={n} q={q} d={d} k={k}: {1e3*(t1-t0):8.1f} ms wall   checksum {int(idx.sum())} {d2.sum():.12e}")
    m.close()
