"""Selection statistics of the tensor kNN kernel (instrumented build) for one shape.
    python profiles/microbench/tc_stats.py n q d k
"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from util_b200 import _lib

n, q, d, k = (int(a) for a in sys.argv[1:5])
x = np.random.default_rng(0).random((n, d)); z = np.random.default_rng(1).random((q, d))
names = ["events", "events_hit", "survivors", "rank_merges", "sel_busy_clk", "trims", "paired_events", "clk_read_ballot", "clk_insert_few", "clk_store_sync", "clk_rank_merge", "pushes", "clk_to_tile64", "clk_to_tile512", "clk_to_tile2048", "clk_block"]
for stats in (0, 1):
    m = _lib.Model(x)
    m.set_option("knn_path", 2)
    m.knn(z, k)
    m.set_option("reset_stats", 1)
    m.set_option("tc_stats", stats)
    m.knn(z, k)
    ms = m.stat("ms_stage1")
    line = f"n={n} q={q} d={d} k={k} tc_stats={stats}: stage1 {ms:.3f} ms rerank {m.stat('ms_rerank'):.3f} ms keep {int(m.stat('tc_keep'))} cap {int(m.stat('tc_cap'))} halves {int(m.stat('tc_halves'))}"
    if stats:
        vals = {nm: m.stat(f"tc_stat{i}") for i, nm in enumerate(names) if nm != "-"}
        rows = q
        blocks = -(-q // (128 * int(m.stat("tc_halves"))))
        line += "\n   per block: " + ", ".join(f"{nm} {vals[nm] / blocks:.0f}" for nm in names[12:])
        line += "\n   per row: " + ", ".join(f"{nm} {v / rows:.1f}" for nm, v in vals.items())
        if vals["events"]:
            line += f"\n   selection busy clk per event {vals['sel_busy_clk'] / vals['events']:.0f}"
    print(line, flush=True)
    m.close()
