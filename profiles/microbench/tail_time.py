"""Stage-1 time at query counts with a short last round (8-GPU shard of C5, C5 itself), tail handled by the
data-split launches (default) or left as one partly filled round (option tail_split = 0); results must be identical."""
import sys, numpy as np
sys.path.insert(0, ".")
from util_b200 import _lib
n, d, k = 1000000, 64, 16
x = np.random.default_rng(0).random((n, d))
m = _lib.Model(x); m.set_option("knn_path", 2)
for q in (125000, 250000, 500000):
    z = np.random.default_rng(1).random((q, d))
    ref = None
    for fl in (1, 0, 1, 0):
        m.set_option("tail_split", fl)
        m.knn(z[:1000], k)
        m.set_option("reset_stats", 1)
        idx, d2 = m.knn(z, k)
        if ref is None: ref = (idx, d2)
        same = np.array_equal(idx, ref[0]) and np.array_equal(d2, ref[1])
        print("q", q, "tail_split", fl, "stage1 %.2f rerank %.2f exact %.2f uncert %d same %s" % (m.stat("ms_stage1"), m.stat("ms_rerank"), m.stat("ms_exact"), m.stat("uncertified"), same))
