"""Per-tile timeline (CTA 0) in the steady state: producer, MMA issuers, epilogue warp 0."""
import sys, numpy as np
sys.path.insert(0, ".")
from util_b200 import _lib
n, q, d, k = 1000000, 37888, 64, 16
x = np.random.default_rng(0).random((n, d)); z = np.random.default_rng(1).random((q, d))
m = _lib.Model(x)
m.set_option("knn_path", 2)
start = int(sys.argv[1]) if len(sys.argv) > 1 else 4000
mode = int(sys.argv[2]) if len(sys.argv) > 2 else 0
flags = int(sys.argv[3]) if len(sys.argv) > 3 else 0
m.set_option("tc_flags", flags)
m.set_option("tc_trace_start", start)
m.set_option("tc_mode", mode)
m.tc_debug_trace(z, k)
t = m.tc_debug_trace(z, k)
t0 = t[t > 0].min()
t = np.where(t > 0, t - t0, -1)
print(f"mode {mode} start {start}: tile | prod empty_ok | mma first_issue, all_issued, committed | epi0 accf_ok, released | d(first_issue) d(accf)")
for i in range(1, 90):
    print(f"{i:3d} | {t[0,i,0]:7d} | acce {t[0,i,1]:7d} full {t[0,i,2]:7d} fence {t[0,i,3]:7d} | {t[1,i,0]:7d} {t[1,i,1]:7d} {t[1,i,3]:7d} | {t[2,i,0]:7d} {t[2,i,1]:7d} | {t[1,i,0]-t[1,i-1,0]:5d} {t[2,i,0]-t[2,i-1,0]:5d}  issue {t[1,i,1]-t[1,i,0]:5d}  accf-issued {t[2,i,0]-t[1,i,1]:5d}  busy {t[2,i,1]-t[2,i,0]:5d}")
