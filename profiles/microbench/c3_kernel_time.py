import sys, time, numpy as np
sys.path.insert(0, ".")
from util_b200 import _lib
n = q = 200000; d = 10
x = np.random.default_rng(0).random((n, d)); z = np.random.default_rng(1).random((q, d))
f = 3 * x[:, 0] + .5 * np.cos(8 * x[:, 0]) + np.sin(5 * x[:, -1])
m = _lib.Model(x)
m.lshep_fit(f)
for rep in range(2):
    t0 = time.perf_counter()
    out, ier = m.lshep_predict(z, 2.0, 1)
    t1 = time.perf_counter()
    print("lshep predict wall %.1f ms" % ((t1 - t0) * 1e3))
