"""Tensor path (with fallback) vs exact path at low dimension: where should knn_dispatch switch?"""
import sys, time, numpy as np
sys.path.insert(0, ".")
from util_b200 import _lib
for d in (2, 3, 4, 5, 6, 8, 12, 16):
    n, q, k = 200000, 100000, 8
    x = np.random.default_rng(0).random((n, d)); z = np.random.default_rng(1).random((q, d))
    m = _lib.Model(x)
    res = {}
    for name, path in (("tensor", _lib.PATH_TENSOR), ("exact", _lib.PATH_EXACT)):
        m.set_option("knn_path", path)
        m.knn(z[:40000], k)
        m.set_option("reset_stats", 1)
        t0 = time.perf_counter(); m.knn(z, k); t1 = time.perf_counter()
        res[name] = (1e3 * (t1 - t0), int(m.stat("uncertified")))
    print(f"d={d:2d}: tensor {res['tensor'][0]:7.1f} ms (uncertified {res['tensor'][1]:6d} of {q}), exact {res['exact'][0]:7.1f} ms")
    m.close()
