"""Per-tile period of CTA 0 in windows along the data stream (how the selection cost decays as tau tightens)."""
import sys, numpy as np
sys.path.insert(0, ".")
from util_b200 import _lib
n, q, d, k = 1000000, 37888, 64, 16
x = np.random.default_rng(0).random((n, d)); z = np.random.default_rng(1).random((q, d))
m = _lib.Model(x)
m.set_option("knn_path", 2)
for mode in (0, 1):
    m.set_option("tc_mode", mode)
    for start in (0, 96, 300, 800, 2000, 4000, 7000):
        m.set_option("tc_trace_start", start)
        m.tc_debug_trace(z, k)
        t = m.tc_debug_trace(z, k)
        a = t[2, :, 0]
        ok = a > 0
        per = (a[ok][-1] - a[ok][0]) / max(ok.sum() - 1, 1)
        cm = t[2, :, 3] > 0
        print(f"mode {mode} tiles {start:5d}..{start+95:5d}: epilogue period {per:7.0f} clk/tile, tiles with compaction in warp 0: {cm.sum():2d}/96")
