import sys, numpy as np
sys.path.insert(0, ".")
from util_b200 import _lib
for (n, q, d, k) in ((1000000, 37888, 64, 16), (200000, 200000, 64, 16), (100000, 100000, 16, 18)):
    x = np.random.default_rng(0).random((n, d)); z = np.random.default_rng(1).random((q, d))
    for mg in (-1, 32 - k, 33 - k, 40 - k):
        m = _lib.Model(x); m.set_option("knn_path", 2); m.set_option("margin", mg)
        m.knn(z, k); m.set_option("reset_stats", 1); m.knn(z, k)
        print(f"n={n} q={q} d={d} k={k} margin={mg}: keep {int(m.stat('tc_keep'))} cap {int(m.stat('tc_cap'))} halves {int(m.stat('tc_halves'))} slots {int(m.stat('tc_slots'))} stage1 {m.stat('ms_stage1'):.2f} rerank {m.stat('ms_rerank'):.2f} unc {int(m.stat('uncertified'))}", flush=True)
        m.close()
