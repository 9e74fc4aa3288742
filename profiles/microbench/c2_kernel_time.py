import sys, time, numpy as np
sys.path.insert(0, ".")
from util_b200 import _lib
n = q = 100000; d = 16
x = np.random.default_rng(0).random((n, d)); z = np.random.default_rng(1).random((q, d))
y = 3 * x[:, 0]
m = _lib.Model(x)
m.shepmod_fit()
for rep in range(2):
    m.set_option("reset_stats", 1)
    t0 = time.perf_counter()
    r = m.shepmod_predict(z, cap=64, y=y)
    t1 = time.perf_counter()
    print("shepmod predict wall %.1f ms, kernel %.1f ms" % ((t1 - t0) * 1e3, m.stat("ms_weights")))
