"""The one-process multi-GPU handle (b200_multi_*) at C5: queries/s for query and data sharding over the
visible GPUs, with the exchange / merge split.   python profiles/microbench/multi_handle.py [n_gpus]"""
import json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from util_b200 import _lib

G = int(sys.argv[1]) if len(sys.argv) > 1 else _lib.lib().b200_device_count()
n, q, d, k = 1_000_000, 1_000_000, 64, 16
x = np.random.default_rng(0).random((n, d)); z = np.random.default_rng(1).random((q, d))
y = 3 * x[:, 0] + .5 * np.cos(8 * x[:, 0]) + np.sin(5 * x[:, -1])
for shard in ("query", "data"):
    m = _lib.MultiModel(x, list(range(G)), shard=shard)
    m.set_values(y)
    for _ in range(2):
        m.nn_predict(z, k, _lib.SHEPARD, y=True, want_weights=False)
    m.set_option("reset_stats", 1)
    steps = 3
    t0 = time.perf_counter()
    for _ in range(steps):
        out = m.nn_predict(z, k, _lib.SHEPARD, y=True, want_weights=False)[4]
    dt = (time.perf_counter() - t0) / steps
    print(json.dumps({"handle": "b200_multi", "n_gpus": G, "shard": shard, "queries_per_s_e2e": q / dt, "ms_per_step": dt * 1e3,
                      "ms_stage1_sum_over_devices": m.stat("ms_stage1") / steps, "ms_rerank_sum": m.stat("ms_rerank") / steps,
                      "ms_local_max": m.stat("ms_local") / steps, "ms_collective_max": m.stat("ms_collective") / steps,
                      "ms_merge_max": m.stat("ms_merge") / steps, "uncertified": m.stat("uncertified") / steps,
                      "checksum": float(out.sum())}), flush=True)
    m.close()
