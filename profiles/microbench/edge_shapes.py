import sys, numpy as np
sys.path.insert(0, ".")
from util_b200 import _lib
bad = 0
for (n, q, d, k) in ((4097, 1, 5, 1), (4097, 257, 5, 128), (5000, 513, 16, 113), (5000, 255, 16, 117), (20000, 1000, 64, 52), (20000, 1000, 64, 53),
                     (20000, 1000, 64, 36), (20000, 1000, 64, 37), (6000, 300, 200, 16), (6000, 300, 125, 64), (130, 50, 8, 128), (129, 300, 8, 128)):
    x = np.random.default_rng(n).random((n, d)); z = np.random.default_rng(q).random((q, d))
    m = _lib.Model(x); m.set_option("knn_path", 2)
    try:
        it, dt = m.knn(z, k)
    except Exception as e:
        print(f"n={n} q={q} d={d} k={k}: {e}"); continue
    info = f"cap {int(m.stat('tc_cap'))} keep {int(m.stat('tc_keep'))} halves {int(m.stat('tc_halves'))} slots {int(m.stat('tc_slots'))} unc {int(m.stat('uncertified'))}"
    m.set_option("knn_path", 1)
    ie, de = m.knn(z, k)
    ok = np.array_equal(it, ie) and np.array_equal(dt, de)
    bad += not ok
    print(f"n={n} q={q} d={d} k={k}: {info} {'ok' if ok else 'MISMATCH'}", flush=True)
    m.close()
print("mismatches:", bad)
