"""Per-tile epilogue period of CTA 0 over the first tiles of the stream (start-up cost of the selection)."""
import sys, numpy as np
sys.path.insert(0, ".")
from util_b200 import _lib
n, q, d, k = 1000000, 37888, 64, 16
x = np.random.default_rng(0).random((n, d)); z = np.random.default_rng(1).random((q, d))
m = _lib.Model(x)
m.set_option("knn_path", 2)
m.set_option("tc_mode", 0)
m.set_option("tc_trace_start", 0)
m.tc_debug_trace(z, k)
t = m.tc_debug_trace(z, k)
a = t[2, :, 0]
e = t[2, :, 1]
c0 = t[2, :, 2]
c1 = t[2, :, 3]
for i in range(0, 95):
    print(f"tile {i:3d}: start->next start {a[i+1]-a[i]:7d}  start->release {e[i]-a[i]:7d}  last compaction {c1[i]-c0[i] if c1[i] > 0 else 0:6d}")
print("total first 95 tiles", a[95] - a[0])
