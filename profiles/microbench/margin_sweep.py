import sys, numpy as np
sys.path.insert(0, ".")
from util_b200 import _lib
for (n, d, k) in ((50000, 32, 64), (50000, 32, 33), (100000, 16, 40)):
    x = np.random.default_rng(0).random((n, d)); z = np.random.default_rng(1).random((n, d))
    for mg in (-1, 4, 8, 12, 24):
        m = _lib.Model(x); m.set_option("knn_path", 2); m.set_option("margin", mg)
        m.knn(z, k); m.set_option("reset_stats", 1); m.knn(z, k)
        print(f"n={n} d={d} k={k} margin={mg}: keep {int(m.stat('tc_keep'))} cap {int(m.stat('tc_cap'))} stage1 {m.stat('ms_stage1'):.2f} rerank {m.stat('ms_rerank'):.2f} exact {m.stat('ms_exact'):.2f} unc {int(m.stat('uncertified'))}", flush=True)
        m.close()
