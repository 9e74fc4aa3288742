import sys, time, numpy as np
sys.path.insert(0, ".")
from util_b200 import _lib
n = q = 50000; d = 32; k = 64
x = np.random.default_rng(0).random((n, d)); z = np.random.default_rng(1).random((q, d))
y = 3 * x[:, 0]
m = _lib.Model(x)
for rep in range(2):
    m.set_option("reset_stats", 1)
    t0 = time.perf_counter()
    r = m.nn_predict(z, k, _lib.VORONOI, y=y, want_weights=False)
    t1 = time.perf_counter()
    print("C4 wall %.1f ms: prep %.2f stage1 %.2f rerank %.2f exact %.2f weights %.2f uncert %d cap %d halves %d" % (
        (t1 - t0) * 1e3, m.stat("ms_prep"), m.stat("ms_stage1"), m.stat("ms_rerank"), m.stat("ms_exact"),
        m.stat("ms_weights"), m.stat("uncertified"), m.stat("tc_cap"), m.stat("tc_halves")))
