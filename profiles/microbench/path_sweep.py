"""Exact (fp64 brute force + fp32 prefilter) against tensor kNN path over problem sizes: kernel ms per call.
    python profiles/microbench/path_sweep.py
"""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from util_b200 import _lib

for d in (8, 16, 32, 64):
    for n in (20000, 50000, 100000, 200000):
        x = np.random.default_rng(0).random((n, d)); z = np.random.default_rng(1).random((n, d))
        for k in (1, 16, 64):
            row = []
            for path in (1, 2):
                m = _lib.Model(x)
                m.set_option("knn_path", path)
                try:
                    m.knn(z, k); m.set_option("reset_stats", 1)
                    m.knn(z, k)
                    ms = sum(m.stat(s) for s in ("ms_prep", "ms_stage1", "ms_rerank", "ms_exact"))
                    row.append((ms, m.stat("uncertified")))
                except Exception as e:
                    row.append((float("nan"), str(e)[:30]))
                m.close()
            print(f"d={d:3d} n=q={n:7d} k={k:3d}  exact {row[0][0]:9.2f} ms   tensor {row[1][0]:9.2f} ms (uncertified {row[1][1]})", flush=True)
