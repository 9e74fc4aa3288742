import sys, time, numpy as np
sys.path.insert(0, ".")
from util_b200 import _lib
n, d = 100000, 16
x = np.random.default_rng(0).random((n, d))
m = _lib.Model(x)
for margin in (-1, 12, 16):
    m.set_option("margin", margin); m.set_option("reset_stats", 1)
    m.shepmod_fit()
    t0 = time.perf_counter(); m.set_option("reset_stats", 1); m.shepmod_fit(); dt = time.perf_counter() - t0
    print("margin", margin, "fit s %.4f" % dt, {k: round(m.stat(k), 2) for k in ("uncertified", "ms_stage1", "ms_rerank", "ms_exact", "tc_cap", "tc_keep")})
