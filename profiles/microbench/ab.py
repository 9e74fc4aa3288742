"""A/B of two builds of the library on the tensor kNN stage (same box, alternating):
    python profiles/microbench/ab.py libA.so libB.so n q d k [reps]
"""
import os, subprocess, sys
a, b, n, q, d, k = sys.argv[1:7]
reps = int(sys.argv[7]) if len(sys.argv) > 7 else 3
code = f"""
import numpy as np
from util_b200 import _lib
x = np.random.default_rng(0).random(({n}, {d})); z = np.random.default_rng(1).random(({q}, {d}))
m = _lib.Model(x); m.set_option("knn_path", 2)
m.knn(z, {k})
ts = []
for _ in range(5):
    m.set_option("reset_stats", 1); m.knn(z, {k}); ts.append(m.stat("ms_stage1"))
print(" ".join(f"{{t:.3f}}" for t in ts), "unc", m.stat("uncertified"))
"""
root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for r in range(reps):
    for name, lib in (("A", a), ("B", b)):
        env = dict(os.environ, B200_LIB=os.path.abspath(lib), PYTHONPATH=root)
        out = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True)
        print(name, os.path.basename(lib), out.stdout.strip() or out.stderr[-300:], flush=True)
