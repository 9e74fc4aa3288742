#!/usr/bin/env python
"""bench.py -- query points/sec of util.approximate's distance-based predict hot path on B200.

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path, headline config C5
    python bench.py --config c1|c2|c3|c4                     # the other BASELINE.json configs (one GPU)
    python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU algorithm on the host cores

Headline workload (BASELINE.json configs[4], SURVEY.md 8d "C5"): X 1,000,000 x 64 uniform [0,1) (seed 0),
Z 1,000,000 x 64 (seed 1), k = 16, NearestNeighbor(method=Shepard), y = 3x0 + .5cos(8x0) + sin(5x_{d-1}).
One step = one pass of the hot path over the whole query batch.  `value` is timed on the device with the
inputs resident in HBM; `e2e` goes through the host-buffer C-ABI call (pinned host memory in, host results
out, copies inside the timed region).  For N > 1 the default (`--shard auto`) gives every rank all fitted
points and a slice of the queries -- the data set fits one GPU, which is the case north_star assigns to
query sharding; `--shard data` splits the fitted points instead (local top-k, NCCL all-to-all of the
partial lists, on-device merge).  Total work is fixed either way, so scaling is "strong".

C1-C4 go through the drop-in classes: fit and predict are timed separately, `value` is the predict rate
from the library's own kernel timers (inputs resident), `e2e` the class call with host arrays, and every
stage gets a roofline entry: tensor peak where the stage is a contraction, HBM bytes for the weighting
stages, and the measured fp64 / fp32 FMA rate of the CUDA cores (b200_debug_fma_peak) for the dense scans.

cpu_baseline: the oracle port (OpenMP, all host cores) on a bounded sample; when the unmodified reference
is installed under baseline/_ref (pip --target) the real util.approximate.NearestNeighbor
(pure Python + sklearn BallTree, one thread) is timed too and reported as `reference_class`.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

CONFIGS = {
    # name: (n, q, d, k)
    "c5": (1_000_000, 1_000_000, 64, 16),
    "c5small": (200_000, 100_000, 64, 16),
    "c5q": (1_000_000, 37_888, 64, 16),   # one 256-query block per SM: quick kernel experiments
}
# the other BASELINE.json configs, run through the drop-in classes: (n, q, d, k, model)
CLASS_CONFIGS = {
    "c1": (10_000, 1_000, 8, 1, "NearestNeighbor(k=1)"),
    "c2": (100_000, 100_000, 16, None, "ShepMod()"),
    "c3": (200_000, 200_000, 10, None, "LSHEP(radius=2)"),
    "c4": (50_000, 50_000, 32, 64, "NearestNeighbor(k=64, method=Voronoi)"),
}


def fun(x):
    return 3 * x[:, 0] + .5 * np.cos(8 * x[:, 0]) + np.sin(5 * x[:, -1])


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        j = json.load(open(p))
        return dict(tensor=float(j["bf16_tflops_sustained"]), tensor_burst=float(j["bf16_tflops"]),
                    hbm=float(j["hbm_gbs"]), src="measured")
    return dict(tensor=1400.0, tensor_burst=1590.0, hbm=6650.0, src="fallback")


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False

    def run(self):
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                self.rows.append([c.strip() for c in out.strip().split(",")])
            except Exception:
                pass
            time.sleep(0.2)

    def summary(self):
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) >= 6 and r[2 + i].lower() == "active" for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def cpu_port(x, y, z, k, seconds=15.0, threads=None):
    """The oracle (CPU restatement of the reference algorithm) on a bounded query sample sized to
    take about `seconds` on all host cores."""
    from oracle import harness as H
    cores = threads or os.cpu_count() or 1
    H.set_threads(cores)
    s = max(2, cores)
    per = None
    for _ in range(3):  # grow the probe until the rate estimate is trustworthy
        t0 = time.perf_counter()
        H.nn_predict(x, z[:s], k, "shepard")
        dt = time.perf_counter() - t0
        per = dt / s
        if dt > 1.0 or s >= len(z):
            break
        s = min(len(z), max(s * 4, int(1.5 / max(per, 1e-9))))
    s = int(max(cores, min(len(z), seconds / max(per, 1e-9))))
    t0 = time.perf_counter()
    ids, w, cnt, _ = H.nn_predict(x, z[:s], k, "shepard")
    out = H.weighted_sum(ids, w, cnt, y)
    dt = time.perf_counter() - t0
    return s / dt, s, dt, cores, (ids, out)


def reference_class_rate(x, y, z, k, seconds=25.0, queries=None):
    """The unmodified reference, when it is installed under baseline/_ref: the real
    util.approximate.NearestNeighbor (pure Python + sklearn BallTree; its Fortran sub-models cannot be
    built here, so method=Average) on a bounded query sample, one thread.  None if not installed."""
    ref = os.path.join(ROOT, "baseline", "_ref")
    if not os.path.isdir(os.path.join(ref, "util")):
        return None
    import types
    import tempfile
    stub = types.ModuleType("fmodpy")  # the reference imports fmodpy at package import; nothing is compiled
    stub.config = types.SimpleNamespace(load_config=lambda *a, **kw: None)

    def fimport(*a, **kw):
        raise ImportError("fmodpy stub: no Fortran toolchain in this image")
    stub.fimport = fimport
    sys.modules.setdefault("fmodpy", stub)
    sys.path.insert(0, ref)
    cwd = os.getcwd()
    os.chdir(tempfile.mkdtemp())
    try:
        import util.approximate as ua
    except Exception as e:  # e.g. sklearn missing
        os.chdir(cwd)
        return {"unavailable": repr(e)[:200]}
    os.chdir(cwd)
    m = ua.NearestNeighbor(k=k)
    t0 = time.perf_counter()
    m.fit(x, y)
    t_fit = time.perf_counter() - t0
    s = 4
    t0 = time.perf_counter()
    m.predict(z[:s], display=False)
    per = (time.perf_counter() - t0) / s
    s = int(queries or max(8, min(len(z), seconds / max(per, 1e-9))))
    t0 = time.perf_counter()
    pred = m.predict(z[:s], display=False)
    dt = time.perf_counter() - t0
    return {"value": s / dt, "unit": "queries/s", "cores": 1, "kind": "reference",
            "sample": f"first {s} queries, util.approximate.NearestNeighbor(k={k}) from baseline/_ref "
                      f"(method=Average: the Fortran sub-models cannot be built), BallTree build {t_fit:.1f} s not included",
            "fit_s": t_fit, "pred": np.asarray(pred, dtype=float)}


def run_class_config(args, name):
    """C1-C4 through the drop-in classes on one GPU: fit and predict timed separately."""
    import util_b200 as ub
    from util_b200 import _lib
    from oracle import harness as H
    n, q, d, k, label = CLASS_CONFIGS[name]
    x = np.random.default_rng(0).random((n, d))
    z = np.random.default_rng(1).random((q, d))
    y = fun(x)
    pk = peaks()
    f64 = C.c_double()
    f32 = C.c_double()
    _lib.check(_lib.lib().b200_debug_fma_peak(0, C.byref(f64), C.byref(f32)))
    make = {"c1": lambda: ub.NearestNeighbor(k=1), "c2": lambda: ub.ShepMod(), "c3": lambda: ub.LSHEP(radius=2),
            "c4": lambda: ub.NearestNeighbor(k=64, method=ub.Voronoi)}[name]
    m = make()
    m.fit(x, y)  # warm-up fit (operand images, allocations)
    t0 = time.perf_counter()
    m.fit(x, y)
    fit_s = time.perf_counter() - t0
    fit_stats = {s: m.model.stat(s) for s in ("ms_prep", "ms_stage1", "ms_rerank", "ms_exact", "ms_weights")}
    sampler = ClockSampler(0)
    sampler.start()
    for _ in range(max(args.warmup, 3)):
        m.predict(z)
    m.model.set_option("reset_stats", 1)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        pred = m.predict(z)
    dt = (time.perf_counter() - t0) / args.steps
    sampler.stop_flag = True
    st = {s: m.model.stat(s) / args.steps for s in
          ("ms_prep", "ms_stage1", "ms_rerank", "ms_exact", "ms_weights", "launches", "uncertified")}
    ms_dev = st["ms_prep"] + st["ms_stage1"] + st["ms_rerank"] + st["ms_exact"] + st["ms_weights"]
    if name == "c2":
        ms_dev = st["ms_weights"]  # ShepMod's stage timer spans the whole predict, radius filter (ms_stage1) included
    pred = np.asarray(pred, dtype=float).reshape(q)
    # ---- roofline of the dominant predict kernel
    if name in ("c2", "c3"):
        # dense scan of every node by every query: fp32 prefilter (2 q n d flops) in front of the fp64 arithmetic;
        # SURVEY 8d counts C2 as a contraction against the tensor peak and C3 as (5M+12) q n fp64 flops
        if name == "c2":
            flops, peak, unit_peak = 2.0 * q * n * d, pk["tensor"], "dense bf16 tensor peak (north_star assigns this stage to the tensor pipe)"
            kern = "knn_tc_kernel<THR> radius filter + shepmod_cand_kernel" if st["ms_stage1"] > 0 else "shepmod_eval_kernel"
        else:
            flops, peak, unit_peak = (5.0 * d + 12.0) * q * n, f64.value, "measured fp64 FMA rate (b200_debug_fma_peak)"
            kern = "lshep_eval_kernel"
        ach = flops / (st["ms_weights"] * 1e-3) / 1e12
        roof = {"bound": "tensor" if name == "c2" else "fp64", "kernel": kern, "achieved": ach, "peak": peak,
                "unit": "TFLOP/s", "frac": ach / peak, "traffic": None, "peak_source": unit_peak,
                "ms_per_launch": st["ms_weights"], "algorithmic_flops_per_launch": flops,
                "fp32_fma_peak_tflops": f32.value, "fp64_fma_peak_tflops": f64.value}
    else:
        flops = 2.0 * q * n * d
        ms_knn = st["ms_stage1"] if st["ms_stage1"] > 0 else st["ms_exact"]
        ach = flops / (max(ms_knn, 1e-9) * 1e-3) / 1e12
        roof = {"bound": "tensor", "kernel": "knn_tc_kernel" if st["ms_stage1"] > 0 else "knn_exact_kernel (fp64 CUDA cores)",
                "achieved": ach, "peak": pk["tensor"], "unit": "TFLOP/s", "frac": ach / pk["tensor"], "traffic": None,
                "peak_source": pk["src"] + " bf16_tflops_sustained", "ms_per_launch": ms_knn,
                "algorithmic_flops_per_launch": flops}
    stages = {"weights": {"bound": "hbm", "ms": st["ms_weights"]}}
    if name == "c4":
        byts = q * (k * d * 8 + d * 8 + k * 4 + k * 12)  # k neighbour rows + query + ids in, ids/weights out
        stages["weights"].update(bytes=byts, achieved=byts / 1e9 / (st["ms_weights"] * 1e-3), peak=pk["hbm"], unit="GB/s")
    # ---- CPU baseline: the oracle port on a bounded sample, parity on that sample
    H.set_threads(os.cpu_count() or 1)
    s = {"c1": q, "c2": 400, "c3": 200, "c4": 300}[name]
    t0 = time.perf_counter()
    if name == "c1":
        ids, w, cnt, _ = H.nn_predict(x, z[:s], 1, "average")
        ref = H.weighted_sum(ids, w, cnt, y)
    elif name == "c2":
        oi, ow, oc, _ = H.shepmod_predict(x, m.rw, z[:s], cap=256)
        ref = H.weighted_sum(oi, ow, oc, y)
    elif name == "c3":
        ref, _ = H.lshep_predict(x, y, m.a[0].T, m.rw[0], z[:s])
    else:
        ids, w, cnt, _ = H.nn_predict(x, z[:s], k, "voronoi")
        ref = H.weighted_sum(ids, w, cnt, y)
    t_cpu = time.perf_counter() - t0
    rel = float(np.max(np.abs(pred[:s] - ref) / np.maximum(np.abs(ref), 1e-3 * max(1.0, np.abs(ref).max()))))
    line = {"metric": "query points/sec (%s predict)" % label, "value": q / (ms_dev * 1e-3), "unit": "queries/s",
            "n_gpus": 1, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_dev,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64 (f16 tensor filter / f32 prefilter in front)", "data": "synthetic",
            "config": {"workload": f"{label} fit + predict, {n}x{d} data, {q} queries", "n": n, "q": q, "d": d, "k": k,
                       "parallelism": "single", "l2": "results of the previous step are overwritten; inputs re-uploaded every step"},
            "fit_s": fit_s, "fit_stages_ms": fit_stats,
            "roofline": roof, "stages_ms": {kk: vv for kk, vv in st.items() if kk.startswith("ms_")},
            "roofline_stages": stages, "gpu_launches": int(round(st["launches"] * args.steps)),
            "e2e": {"value": q / dt, "unit": "queries/s", "h2d_bytes_per_step": int(z.nbytes),
                    "d2h_bytes_per_step": int(q * 8), "ms_per_step": dt * 1e3},
            "clocks": sampler.summary(),
            "cpu_baseline": {"value": s / t_cpu, "unit": "queries/s", "cores": os.cpu_count(), "kind": "port",
                             "sample": f"first {s} of {q} queries against all {n} points in {t_cpu:.1f} s, OpenMP over queries"},
            "parity_on_sample": {"max_rel_err": rel}}
    if name == "c1" and not args.no_ref_class:
        rc = reference_class_rate(x, y, z, 1, seconds=args.ref_seconds)
        if rc and "pred" in rc:
            p_ref = rc.pop("pred")
            rc["predictions_equal"] = bool(np.array_equal(p_ref, pred[:len(p_ref)]))
        line["reference_class"] = rc
    print(json.dumps(line), flush=True)


def run_reference(args, n, q, d, k):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    x = np.random.default_rng(0).random((n, d))
    z = np.random.default_rng(1).random((q, d))
    y = fun(x)
    from oracle import harness as H
    cores = os.cpu_count() or 1
    H.set_threads(cores)
    # size one step at ~ (120 s budget) / (steps + warmup)
    s0 = max(2, cores)
    per = None
    for _ in range(3):
        t0 = time.perf_counter()
        H.nn_predict(x, z[:s0], k, "shepard")
        dt0 = time.perf_counter() - t0
        per = dt0 / s0
        if dt0 > 1.0 or s0 >= q:
            break
        s0 = min(q, max(s0 * 4, int(1.5 / max(per, 1e-9))))
    budget = 120.0 / (args.steps + args.warmup)
    s = int(max(cores, min(q, budget / max(per, 1e-9))))
    for _ in range(args.warmup):
        H.nn_predict(x, z[:s], k, "shepard")
    t0 = time.perf_counter()
    for i in range(args.steps):
        ids, w, cnt, _ = H.nn_predict(x, z[:s], k, "shepard")
        H.weighted_sum(ids, w, cnt, y)
    dt = (time.perf_counter() - t0) / args.steps
    val = s / dt
    line = {"impl": "reference", "metric": "query points/sec (kNN+Shepard predict)", "value": val,
            "unit": "queries/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"NearestNeighbor(k={k}, method=Shepard) predict, {n}x{d} data, {q} queries",
                       "n": n, "q": q, "d": d, "k": k},
            "cpu_baseline": {"value": val, "unit": "queries/s", "cores": cores, "kind": "port",
                             "sample": f"{s} of {q} queries per step against all {n} points, OpenMP over queries"},
            "e2e": {"value": val, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    if not args.no_ref_class:
        rc = reference_class_rate(x, y, z, k, seconds=args.ref_seconds, queries=args.ref_queries or None)
        if rc:
            rc.pop("pred", None)
        line["reference_class"] = rc
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--config", default="c5")
    ap.add_argument("--shard", default="auto", choices=["auto", "data", "query"],
                    help="N>1: shard the fitted points (NCCL all-to-all + merge) or the queries; auto = "
                         "queries when the data set fits one GPU's HBM (north_star), else data")
    ap.add_argument("--cpu-seconds", type=float, default=15.0)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-ref-class", action="store_true", help="skip timing the real reference class from baseline/_ref")
    ap.add_argument("--ref-seconds", type=float, default=25.0)
    ap.add_argument("--ref-queries", type=int, default=0, help="fixed query sample for the reference class (0: by time)")
    ap.add_argument("--no-e2e", action="store_true", help="kernel experiments only: skip the host-buffer leg")
    ap.add_argument("--path", type=int, default=0, help="0 auto, 1 exact, 2 tensor")
    ap.add_argument("--margin", type=int, default=-1)
    ap.add_argument("--tc-halves", type=int, default=0)
    ap.add_argument("--tc-slots", type=int, default=0)
    ap.add_argument("--tc-parts", type=int, default=1)
    ap.add_argument("--tc-stats", type=int, default=0, help="instrumented stage-1 kernel: event / stall counters")
    ap.add_argument("--tc-mode", type=int, default=0, help="ceilings: 1 no selection, 2 no epilogue")
    args = ap.parse_args()
    if args.config in CLASS_CONFIGS:
        if args.impl == "reference":
            n, q, d, k, _ = CLASS_CONFIGS[args.config]
            return run_reference(args, n, q, d, k or 16)
        return run_class_config(args, args.config)
    n, q, d, k = CONFIGS[args.config]
    if args.shard == "auto":
        args.shard = "query" if n * d * 8 < 60e9 else "data"
    if args.warmup < 3:
        args.warmup = 3

    if args.impl == "reference":
        return run_reference(args, n, q, d, k)

    import torch
    import torch.distributed as dist
    from util_b200 import _lib
    lib = _lib.lib()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    G = world

    x = np.random.default_rng(0).random((n, d))
    z = np.random.default_rng(1).random((q, d))
    y = fun(x)

    # ---- data layout: under data sharding rank r owns a contiguous block of the fitted points
    # (util_b200.sharding); otherwise every rank holds all of them (and, with --shard query, a
    # slice of the queries).
    from util_b200 import sharding
    data_sharded = G > 1 and args.shard == "data"
    if data_sharded:
        sh = sharding.DataShardedKNN(x, local)
        model, lo, hi = sh.model, sh.lo, sh.hi
    else:
        model, lo, hi = _lib.Model(x, device=local), 0, n
    model.set_option("knn_path", args.path)
    model.set_option("margin", args.margin)
    model.set_option("tc_mode", args.tc_mode)
    model.set_option("tc_halves", args.tc_halves)
    model.set_option("tc_slots", args.tc_slots)
    model.set_option("tc_parts", args.tc_parts)
    model.set_option("tc_stats", args.tc_stats)
    h = model.handle
    if G > 1 and args.shard == "query":
        qs = (q + G - 1) // G
        zl = z[rank * qs:min(q, (rank + 1) * qs)]
    else:
        zl = z
    ql = zl.shape[0]
    z_pin = torch.from_numpy(zl).pin_memory()
    z_dev = z_pin.to(dev, non_blocking=False)
    y_dev = torch.from_numpy(y).to(dev)
    idx = torch.empty((ql, k), dtype=torch.int32, device=dev)
    d2 = torch.empty((ql, k), dtype=torch.float64, device=dev)
    qm = q // G if data_sharded else ql
    out = torch.empty((qm,), dtype=torch.float64, device=dev)
    last = {"out": out, "idx": idx}
    stream = torch.cuda.current_stream().cuda_stream

    wev = []  # CUDA events around the weighting kernel of each step

    def step_device():
        if data_sharded:
            # local top-k -> NCCL all-to-all of the partial lists -> on-device merge -> weights
            last["out"], last["idx"], _ = sh.predict(z_dev, k, _lib.SHEPARD, y_dev, stream)
        else:
            _lib.check(lib.b200_knn_dev(h, z_dev.data_ptr(), ql, k, -1, idx.data_ptr(), d2.data_ptr(), stream))
            w0, w1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            w0.record()
            _lib.check(lib.b200_nn_weights_dev(h, None, ql, k, _lib.SHEPARD, idx.data_ptr(), d2.data_ptr(),
                                               y_dev.data_ptr(), 1, None, None, None, out.data_ptr(), stream))
            w1.record()
            wev.append((w0, w1))

    def sync_all():
        torch.cuda.synchronize()
        if G > 1:
            dist.barrier()
            torch.cuda.synchronize()

    for _ in range(args.warmup):
        step_device()
    sync_all()
    model.set_option("reset_stats", 1)
    wev.clear()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step_device()
    e1.record()
    sync_all()
    ms = e0.elapsed_time(e1) / args.steps
    stats = {s: model.stat(s) / args.steps for s in
             ("ms_prep", "ms_stage1", "ms_rerank", "ms_exact", "launches", "uncertified")}
    stats["ms_weights"] = sum(a.elapsed_time(b) for a, b in wev) / max(len(wev), 1)
    if data_sharded:
        sh.ms_collective = sh.ms_merge = 0.0
        sh._timed = sh._timed[-args.steps:]
        sh.collect_times()
        stats["ms_collective"] = sh.ms_collective / args.steps
        stats["ms_merge"] = sh.ms_merge / args.steps
    if G > 1:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    value = q / (ms * 1e-3) if not (G > 1 and args.shard == "query") else q / (ms * 1e-3)

    # ---- end to end through the host-buffer C-ABI (pinned inputs, host outputs, copies timed)
    e2e = None
    h2d = d2h = 0
    if args.no_e2e:
        dt = float("nan")
    elif not data_sharded:
        zp = z_pin.numpy()
        model.set_values(y)  # as WeightedApproximator.fit does: the responses go to the device once
        for _ in range(2):
            model.nn_predict(zp, k, _lib.SHEPARD, y=True, want_weights=False)
        sync_all()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            res = model.nn_predict(zp, k, _lib.SHEPARD, y=True, want_weights=False)
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / args.steps
        h2d = zp.nbytes
        d2h = ql * 8 + ql * 4 * 2
    else:
        # host queries -> every rank (each needs all of them), host <- this rank's slice of predictions
        out_host = torch.empty((qm,), dtype=torch.float64).pin_memory()

        def step_e2e():
            z_dev.copy_(z_pin, non_blocking=True)
            step_device()
            out_host.copy_(last["out"], non_blocking=True)
            torch.cuda.synchronize()
        for _ in range(2):
            step_e2e()
        sync_all()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            step_e2e()
        dt = (time.perf_counter() - t0) / args.steps
        h2d = z_pin.numel() * 8
        d2h = qm * 8
    if G > 1:
        t = torch.tensor([dt], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t.item())
    e2e = {"value": q / dt, "unit": "queries/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
           "ms_per_step": dt * 1e3}
    sampler.stop_flag = True

    if rank == 0:
        pk = peaks()
        nloc = hi - lo
        flops = 2.0 * ql * nloc * d  # algorithmic: the -2 Z X^T term only
        s1 = stats["ms_stage1"]
        tensor_path = s1 > 0
        # DRAM bytes per launch of the current kernel: not measurable outside a profiler, so it is the figure of
        # the newest committed `ncu --set full` capture of this config (profiles/rNN/knn_tc_traffic.json names it)
        traffic, traffic_src = None, None
        for rd in sorted((p for p in os.listdir(os.path.join(ROOT, "profiles")) if p.startswith("r")), reverse=True):
            tp = os.path.join(ROOT, "profiles", rd, "knn_tc_traffic.json")
            if os.path.exists(tp) and G == 1:
                tj = json.load(open(tp))
                if tj.get("config") == args.config:
                    traffic = tj["dram_bytes_read"] + tj["dram_bytes_write"]
                    traffic_src = f"profiles/{rd}/knn_tc_traffic.json"
                break
        if tensor_path:
            ach = flops / (s1 * 1e-3) / 1e12
            roof = {"bound": "tensor", "kernel": "knn_tc_kernel", "achieved": ach, "peak": pk["tensor"],
                    "unit": "TFLOP/s", "frac": ach / pk["tensor"], "traffic": traffic,
                    "traffic_source": traffic_src,
                    "peak_source": pk["src"] + " bf16_tflops_sustained", "ms_per_launch": s1,
                    "algorithmic_flops_per_launch": flops}
        else:
            ach = flops / (stats["ms_exact"] * 1e-3) / 1e12 if stats["ms_exact"] else 0.0
            roof = {"bound": "tensor", "kernel": "knn_exact_kernel (fp64 CUDA cores)", "achieved": ach,
                    "peak": pk["tensor"], "unit": "TFLOP/s", "frac": ach / pk["tensor"], "traffic": None,
                    "peak_source": pk["src"]}
        line = {"metric": "query points/sec (kNN+Shepard predict)", "value": value, "unit": "queries/s",
                "n_gpus": G, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
                "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f16 filter + f64 re-rank",
                "data": "synthetic",
                "config": {"workload": f"NearestNeighbor(k={k}, method=Shepard) predict, {n}x{d} data, {q} queries",
                           "n": n, "q": q, "d": d, "k": k, "parallelism": (f"{args.shard}-shard x{G}" if G > 1 else "single"),
                           "l2": "inputs larger than L2 (queries %d MB, operand images %d MB)" % (zl.nbytes >> 20, (nloc * int(model.stat('tc_K') or d) * 2) >> 20)},
                "roofline": roof,
                "stages_ms": {kk: vv for kk, vv in stats.items() if kk.startswith("ms_")},
                # HBM-bound stages: re-rank gathers c candidate rows per query
                # (c*d*8 + d*8 + k*12 bytes); weighting reads k*(8+4+8) and writes 8 bytes per query
                # (a row's buffer ends with keep..CAP candidates: keep is the lower bound used here)
                "roofline_stages": {
                    "rerank": {"bound": "hbm", "bytes_per_query": int(model.stat("tc_keep")) * d * 8 + d * 8 + k * 12,
                               "ms": stats["ms_rerank"],
                               "achieved": (ql * (model.stat("tc_keep") * d * 8 + d * 8 + k * 12) / 1e9) / (stats["ms_rerank"] * 1e-3)
                               if stats["ms_rerank"] > 0 else None, "peak": pk["hbm"], "unit": "GB/s"},
                    "weights": {"bound": "hbm", "bytes_per_query": k * 20 + 8, "ms": stats["ms_weights"],
                                "achieved": (ql * (k * 20 + 8) / 1e9) / (stats["ms_weights"] * 1e-3)
                                if stats["ms_weights"] > 0 else None, "peak": pk["hbm"], "unit": "GB/s"}},
                "uncertified_per_step": stats["uncertified"],
                "tc": {"cap": model.stat("tc_cap"), "keep": model.stat("tc_keep"), "slots": model.stat("tc_slots"), "halves": model.stat("tc_halves"),
                       "K": model.stat("tc_K"), "mode": args.tc_mode},
                "gpu_launches": int(round(stats["launches"] * args.steps)),
                **({"tc_stats": {nm: model.stat(f"tc_stat{i}") / args.steps for i, nm in enumerate(
                    ["events", "events_hit", "survivors", "rank_merges", "sel_busy_clk", "trims",
                     "paired_events", "acc_wait_clk", "push_clk", "release_clk", "tile_clk", "pushes",
                     "push_clk_after_tile64", "pushes_after_tile64", "clk_to_tile512", "clk_to_tile64"])}} if args.tc_stats else {}),
                "e2e": e2e, "clocks": sampler.summary()}
        if G == 1 and not args.no_cpu:
            v, s, dtc, cores, (cids, cout) = cpu_port(x, y, z, k, seconds=args.cpu_seconds)
            line["cpu_baseline"] = {"value": v, "unit": "queries/s", "cores": cores, "kind": "port",
                                    "sample": f"first {s} of {q} queries against all {n} points in {dtc:.1f} s, OpenMP over queries"}
            # the timed GPU outputs must agree with the CPU sample: indices bit-exact, values 1e-12
            gi = last["idx"][:s].cpu().numpy().astype(np.int64)
            go = last["out"][:s].cpu().numpy()
            line["parity_on_sample"] = {"indices_equal": bool(np.array_equal(gi, cids)),
                                        "max_rel_err": float(np.max(np.abs(go - cout) / np.maximum(np.abs(cout), 1e-3)))}
            if not args.no_ref_class:
                # the reference's own class (one thread, method=Average); its neighbours must be ours
                rc = reference_class_rate(x, y, z, k, seconds=args.ref_seconds, queries=args.ref_queries or None)
                if rc and "pred" in rc:
                    p_ref = rc.pop("pred")
                    mine = y[last["idx"][:len(p_ref)].cpu().numpy().astype(np.int64)].mean(axis=1)
                    rc["max_rel_err_vs_ours"] = float(np.max(np.abs(mine - p_ref) / np.maximum(np.abs(p_ref), 1e-3)))
                line["reference_class"] = rc
        print(json.dumps(line), flush=True)
    if G > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
