"""Generate golden vectors from the REAL reference (run in the build container only).

    python tests/golden/make_golden.py

Imports util.approximate from /root/reference with a stub `fmodpy` module (the
Fortran-backed models cannot be built here: no gfortran), runs the reference's
own NearestNeighbor (pure Python + sklearn BallTree) and util.math.pairs
.pairwise_distance, and stores inputs seeds + outputs as small .npz fixtures.
The GPU box has no /root/reference; tests read only the committed fixtures.
"""
import os
import sys
import tempfile
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"


def _import_reference():
    stub = types.ModuleType("fmodpy")
    stub.config = types.SimpleNamespace(load_config=lambda *a, **k: None)

    def fimport(*a, **k):
        raise ImportError("fmodpy stub: no Fortran toolchain in this image")
    stub.fimport = fimport
    sys.modules["fmodpy"] = stub
    sys.path.insert(0, REF)
    cwd = os.getcwd()
    os.chdir(tempfile.mkdtemp())  # the reference writes nothing, but stay out of the repo
    import util.approximate as ua
    from util.math.pairs import pairwise_distance
    os.chdir(cwd)
    return ua, pairwise_distance


def fun(x):
    # util/approximate/testing.py:25
    return 3 * x[:, 0] + .5 * np.cos(8 * x[:, 0]) + np.sin(5 * x[:, -1])


def synth(n, q, d, sx=0, sz=1):
    # uniform [0,1)^d, data seed 0, query seed 1
    x = np.random.default_rng(sx).random((n, d))
    z = np.random.default_rng(sz).random((q, d))
    return x, z


def main():
    ua, pairwise_distance = _import_reference()
    cases = {
        "nn_c1": dict(n=10000, q=1000, d=8, k=1),        # BASELINE.json configs[0]
        "nn_k5_d16": dict(n=5000, q=300, d=16, k=5),
        "nn_k16_d64": dict(n=20000, q=200, d=64, k=16),  # configs[4] shape, reduced n
        "nn_k32_d16": dict(n=8000, q=200, d=16, k=32),
        "nn_k3_d2": dict(n=700, q=500, d=2, k=3),
    }
    for name, c in cases.items():
        x, z = synth(c["n"], c["q"], c["d"])
        y = fun(x)
        m = ua.NearestNeighbor(k=c["k"])
        m.fit(x, y)
        pred = np.array(m.predict(z, display=False), dtype=np.float64)
        m2 = ua.NearestNeighbor(k=c["k"])
        m2.fit(x)
        pw = m2.predict(z, display=False)
        ids = np.array([p[0] for p in pw], dtype=np.int64)
        wts = np.array([p[1] for p in pw], dtype=np.float64)
        dist = m.tree.query(z, k=c["k"])[0]
        np.savez_compressed(os.path.join(HERE, name + ".npz"), n=c["n"], q=c["q"], d=c["d"], k=c["k"],
                            x_sum=x.sum(), z_sum=z.sum(), ids=ids, wts=wts, pred=pred, dist=dist)
        print(name, ids.shape, pred[:2])

    # the reference's own demo input (nearest_neighbor/__init__.py:166: test_plot(m, N=300))
    from util.random import latin
    np.random.seed(0)
    x = latin(300, 2)
    y = fun(x)
    g = np.linspace(-.1, 1.1, 25)
    z = np.array([[a, b] for a in g for b in g])
    m = ua.NearestNeighbor()
    m.fit(x, y)
    pred = np.array(m.predict(z, display=False))
    np.savez_compressed(os.path.join(HERE, "nn_demo_latin.npz"), x=x, y=y, z=z, pred=pred)
    print("nn_demo_latin", pred[:2])

    # util/math/pairs.py:39-57 GEMM-form distances (+ its own assert test :64-83)
    rng = np.random.default_rng(7)
    a, b = rng.random((100, 30)), rng.random((10, 30))
    d = pairwise_distance(a, b)
    for i in range(len(a)):
        for j in range(len(b)):
            assert abs(d[i, j] - np.linalg.norm(a[i] - b[j])) < 2 ** (-32)
    np.savez_compressed(os.path.join(HERE, "pairwise.npz"), a=a, b=b, d=d)
    print("pairwise", d.shape)


def gen_wrappers():
    """Split / unique of the reference (wrappers.py:13-116) around pure-Python stand-in models."""
    from util.approximate.wrappers import Split, unique
    from util.approximate.base import Approximator, WeightedApproximator
    sys.path.insert(0, HERE)
    import standins
    xu, yu, zu, xs, ys, zs = standins.wrapper_inputs()
    a = unique(standins.nearest_two(WeightedApproximator))()
    a.fit(xu, yu)
    with_values = np.array(a(zu))
    b = unique(standins.nearest_two(WeightedApproximator))()
    b.fit(xu)
    rows = b(zu)
    s = Split(standins.linear(Approximator), k=4, seed=5)
    s.fit(xs, ys)
    from util.approximate.wrappers import condition
    cond = {}
    for tag, kw in (("a", {}), ("b", dict(dim=2)), ("c", dict(scale=False))):
        c = condition(standins.linear(Approximator), method="PCA", **kw)()
        c.fit(xs, ys)
        cond[f"cond_{tag}_matrix"] = c.conditioner
        cond[f"cond_{tag}_pred"] = c(zs)
    np.savez(os.path.join(HERE, "wrappers.npz"), **cond, unique_values=with_values,
             unique_ids=np.array([list(r[0]) + [-1] * (40 - len(r[0])) for r in rows]),
             unique_wts=np.array([list(r[1]) + [0.0] * (40 - len(r[1])) for r in rows]),
             unique_indices=b.unique_indices, split=s(zs))


def gen_auto():
    """nearest_neighbor.auto (nearest_neighbor/__init__.py:93-132) of the reference on seeded inputs; the
    rows it sampled are recorded (util.random.random_range wrapped) so that the test can hand them in."""
    import random
    import util.random as ur
    import util.approximate.nearest_neighbor as rnn
    real = ur.random_range
    drawn = []

    def recording(*a, **k):
        for v in real(*a, **k):
            drawn.append(v)
            yield v
    ur.random_range = recording
    cases = [dict(n=300, d=2, kw={}), dict(n=2000, d=5, kw=dict(mean=False)), dict(n=2500, d=3, kw=dict(k_step=2)),
             dict(n=900, d=4, kw=dict(samples=40)), dict(n=6000, d=3, kw=dict(samples=30)),
             dict(n=40000, d=6, kw=dict(samples=50, mean=False))]  # the last one draws rows with the LCG walk
    # (a caller-supplied max_k cannot be pinned: the reference only imports log2 when max_k is None and
    # raises UnboundLocalError at :112 otherwise)
    out = {}
    for i, c in enumerate(cases):
        x = np.random.default_rng(100 + i).random((c["n"], c["d"]))
        y = fun(x) + (0.3 * np.random.default_rng(200 + i).standard_normal(c["n"]) if i % 2 else 0.0)
        random.seed(i)
        drawn.clear()
        k = rnn.auto(x, y, **c["kw"])
        out[f"k{i}"] = k
        out[f"rows{i}"] = np.array(drawn, dtype=np.int64)
        out[f"x_sum{i}"] = x.sum()
        print("auto", i, c, "->", k, len(drawn))
    ur.random_range = real
    np.savez_compressed(os.path.join(HERE, "auto_k.npz"), **out)


if __name__ == "__main__":
    main()
    gen_wrappers()
    gen_auto()
