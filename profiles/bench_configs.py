"""Time BASELINE.json configs[0..3] (C1-C4) through the drop-in classes on one B200,
next to the CPU oracle on a bounded sample of the same workload, and check parity on that sample.
(configs[4] is bench.py's headline.)  Prints one JSON line per config.

    python profiles/bench_configs.py [c1 c2 c3 c4]
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import util_b200 as ub  # noqa: E402
from oracle import harness as H  # noqa: E402


def fun(x):
    return 3 * x[:, 0] + .5 * np.cos(8 * x[:, 0]) + np.sin(5 * x[:, -1])


def synth(n, q, d):
    return np.random.default_rng(0).random((n, d)), np.random.default_rng(1).random((q, d))


def relerr(a, b):
    a, b = np.asarray(a, float), np.asarray(b, float)
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-3 * max(1.0, np.abs(b).max()))))


def timed(f, reps=1):
    f()  # warm-up (operand images, allocations)
    t0 = time.perf_counter()
    for _ in range(reps):
        r = f()
    return (time.perf_counter() - t0) / reps, r


def c1():
    n, q, d, k = 10000, 1000, 8, 1
    x, z = synth(n, q, d)
    y = fun(x)
    m = ub.NearestNeighbor(k=k)
    t_fit, _ = timed(lambda: m.fit(x, y))
    t_pred, pred = timed(lambda: m.predict(z), reps=5)
    t0 = time.perf_counter()
    ids, w, cnt, _ = H.nn_predict(x, z, k, "average")
    ref = H.weighted_sum(ids, w, cnt, y)
    t_cpu = time.perf_counter() - t0
    return dict(config="C1 NearestNeighbor(k=1) 10k x 1k, 8-D", fit_s=t_fit, predict_s=t_pred, qps=q / t_pred,
                cpu_qps=q / t_cpu, cpu_sample=q, bit_exact=bool(np.array_equal(np.array(pred), ref)))


def c2():
    n, q, d = 100000, 100000, 16
    x, z = synth(n, q, d)
    y = fun(x)
    m = ub.ShepMod()
    t_fit, _ = timed(lambda: m.fit(x, y))
    t_pred, pred = timed(lambda: m.predict(z))
    s = 400  # oracle sample (the CPU scan is O(n) per query; the fit O(n^2) is sampled through rw of s nodes)
    t0 = time.perf_counter()
    oi, ow, oc, oe = H.shepmod_predict(x, m.rw, z[:s], cap=256)
    ref = H.weighted_sum(oi, ow, oc, y)
    t_cpu = time.perf_counter() - t0
    idx, d2 = H.knn(x, x[:s], d + 2)
    rw_ok = bool(np.array_equal(np.sqrt(d2[:, -1]), m.rw[:s]))
    return dict(config="C2 ShepMod 100k x 100k, 16-D", fit_s=t_fit, predict_s=t_pred, qps=q / t_pred,
                cpu_qps=s / t_cpu, cpu_sample=s, rw_bit_exact_on_sample=rw_ok,
                max_rel_err=relerr(np.array(pred[:s]), ref), ierrors=m.ierrors,
                contributors_mean=float(oc.mean()), contributors_max=int(oc.max()))


def c3():
    n, q, d = 200000, 200000, 10
    x, z = synth(n, q, d)
    y = fun(x)
    m = ub.LSHEP(radius=2)
    t_fit, _ = timed(lambda: m.fit(x, y))
    t_pred, pred = timed(lambda: m.predict(z))
    s = 200
    t0 = time.perf_counter()
    ref, ier = H.lshep_predict(x, y, m.a[0].T, m.rw[0], z[:s])  # evaluation parity on the GPU's fitted state
    t_cpu = time.perf_counter() - t0
    # fit parity on a sample of nodes: DGELSS on the same neighbourhoods
    npn = (3 * d + 1) // 2
    nodes = np.arange(0, n, n // 100)[:100]
    idx, d2 = H.knn(x, x[nodes], npn + 1)
    worst = 0.0
    for r, kk in enumerate(nodes):
        nb, dist = idx[r, 1:], np.sqrt(d2[r, 1:])
        om = (1.1 * dist[-1] - dist) / (1.1 * dist[-1] * dist)
        sol = np.linalg.lstsq(om[:, None] * (x[nb] - x[kk]), om * (y[nb] - y[kk]), rcond=None)[0]
        worst = max(worst, float(np.max(np.abs(m.a[0][:, kk] - sol)) / np.max(np.abs(sol))))
    return dict(config="C3 LSHEP(radius=2) fit + predict 200k x 200k, 10-D", fit_s=t_fit, predict_s=t_pred,
                qps=q / t_pred, cpu_qps=s / t_cpu, cpu_sample=s, max_rel_err=relerr(pred[:s], ref),
                fit_grad_rel_err_vs_lstsq=worst, fit_ier=m.fit_ier, ierrors=m.ierrors)


def c4():
    n, q, d, k = 50000, 50000, 32, 64
    x, z = synth(n, q, d)
    y = fun(x)
    m = ub.NearestNeighbor(k=k, method=ub.Voronoi)
    t_fit, _ = timed(lambda: m.fit(x, y))
    t_pred, pred = timed(lambda: m.predict(z))
    s = 300
    t0 = time.perf_counter()
    ids, w, cnt, err = H.nn_predict(x, z[:s], k, "voronoi")
    ref = H.weighted_sum(ids, w, cnt, y)
    t_cpu = time.perf_counter() - t0
    return dict(config="C4 NearestNeighbor(k=64, method=Voronoi) 50k x 50k, 32-D", fit_s=t_fit, predict_s=t_pred,
                qps=q / t_pred, cpu_qps=s / t_cpu, cpu_sample=s, max_rel_err=relerr(np.array(pred[:s]), ref),
                support_mean=float(cnt.mean()))


if __name__ == "__main__":
    H.set_threads(os.cpu_count() or 1)
    which = sys.argv[1:] or ["c1", "c2", "c3", "c4"]
    for name in which:
        r = {"c1": c1, "c2": c2, "c3": c3, "c4": c4}[name]()
        r["cpu_cores"] = os.cpu_count()
        print(json.dumps(r), flush=True)
