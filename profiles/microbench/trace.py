"""Print the producer / MMA / epilogue timeline of CTA 0 (clock64 stamps) for a few tiles."""
import sys, numpy as np
sys.path.insert(0, ".")
from util_b200 import _lib
n, q, d, k = 1000000, 37888, 64, 16
x = np.random.default_rng(0).random((n, d)); z = np.random.default_rng(1).random((q, d))
m = _lib.Model(x)
m.set_option("knn_path", 2)
m.set_option("tc_trace_start", 4000)
for mode, flags, halves in [(0, 0, 2)]:
    m.set_option("tc_mode", mode); m.set_option("tc_flags", flags); m.set_option("tc_halves", halves)
    m.tc_debug_trace(z, k)
    t = m.tc_debug_trace(z, k)
    t0 = t[t > 0].min()
    t = np.where(t > 0, t - t0, -1)
    print(f"--- mode {mode} flags {flags} halves {halves}: per-tile period (MMA commit) "
          f"{(t[1, 80, 3] - t[1, 40, 3]) / 40:.0f} clk")
    print("tile | prod: empty_ok | mma: first_issue all_issued - end | epi: accf_ok released")
    e = t[2, 8:88, 1] - t[2, 8:88, 0]
    print("epilogue warp 0: mean busy per tile %.0f clk (max %d), mean wait for accf %.0f clk" % (e.mean(), e.max(), (t[2, 9:88, 0] - t[2, 8:87, 1]).mean()))
    cm = t[2, 8:88, 3] > 0
    cd = (t[2, 8:88, 3] - t[2, 8:88, 2])[cm]
    print("tiles with a compaction in warp 0: %d of 80; last compaction in such a tile took %s clk" % (cm.sum(), cd))
    print("busy per tile:", e)
    for i in []:
        print(i, "|", t[0, i, 0], "|", *t[1, i], "|", t[2, i, 0], t[2, i, 1])
