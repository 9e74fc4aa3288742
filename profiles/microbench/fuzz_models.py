"""Randomised parity of the drop-in classes against the CPU oracle at small sizes: NearestNeighbor with each
sub-model, ShepMod, LSHEP, stand-alone Shepard / Voronoi, over random n, d, k, scales and shifts.
    python profiles/microbench/fuzz_models.py [cases] [seed]
"""
import sys
import numpy as np
sys.path.insert(0, ".")
import util_b200 as ub
from oracle import harness as H

cases = int(sys.argv[1]) if len(sys.argv) > 1 else 30
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 0)


def relerr(a, b):
    a, b = np.asarray(a, float), np.asarray(b, float)
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-3 * max(1.0, np.abs(b).max()))))


bad = 0
for c in range(cases):
    d = int(rng.integers(2, 12)); n = int(rng.integers(200, 3000)); q = int(rng.integers(5, 120))
    scale = float(rng.choice([1.0, 1e-3, 1e2])); shift = float(rng.choice([0.0, 5.0, -300.0]))
    x = rng.random((n, d)) * scale + shift; z = rng.random((q, d)) * scale + shift
    y = np.sin(3 * (x[:, 0] - shift) / scale) + (x[:, -1] - shift) / scale
    kind = rng.choice(["nn_average", "nn_shepard", "nn_voronoi", "shepmod", "lshep", "shepard", "voronoi"])
    try:
        if kind.startswith("nn_"):
            meth = kind[3:]; k = int(rng.integers(1, min(20, n)))
            if meth == "voronoi": k = max(k, 2)
            m = ub.NearestNeighbor(k=k, method={"average": ub.Average, "shepard": ub.Shepard, "voronoi": ub.Voronoi}[meth])
            m.fit(x, y); got = np.array(m(z))
            ids, w, cnt, err = H.nn_predict(x, z, k, meth); ref = H.weighted_sum(ids, w, cnt, y)
        elif kind == "shepmod":
            m = ub.ShepMod(); m.fit(x, y); got = np.array(m(z))
            rw, _ = H.shepmod_fit(x); oi, ow, oc, oe = H.shepmod_predict(x, rw, z, cap=n); ref = H.weighted_sum(oi, ow, oc, y)
        elif kind == "lshep":
            m = ub.LSHEP(); m.fit(x, y); got = np.asarray(m(z))
            a_o, rwl, _, _ = H.lshep_fit(x, y); ref, _ = H.lshep_predict(x, y, m.a[0].T, rwl * 2, z)
        elif kind == "shepard":
            m = ub.Shepard(); m.fit(x, y); got = np.array(m(z))
            ref = np.array([sum(y[i] * w for i, w in enumerate(H.shepard_weight(x, p))) for p in z])
        else:
            xs = x[:200]; ys = y[:200]
            m = ub.Voronoi(); m.fit(xs, ys); got = np.array(m(z))
            ref = []
            for p in z:
                w, e = H.voronoi_predict(xs, p); ref.append(sum(ys[i] * w[i] for i in range(len(xs)) if w[i] > 0))
            ref = np.array(ref)
        e = relerr(got, ref)
        ok = e < 1e-11
    except ub.DuplicateInterpolationPoints as ex:
        # tiny scales put squared distances under the reference's sqrt(eps) duplicate test (voronoi.f90:137):
        # fine when the oracle flags ERROR 2 on the same input
        e = str(ex)
        if kind == "nn_voronoi":
            ok = bool((H.nn_predict(x, z, k, "voronoi")[3] == 2).any())
        else:
            ok = kind == "voronoi" and any(H.voronoi_predict(x[:200], p)[1] == 2 for p in z)
    except Exception as ex:
        e, ok = str(ex), False
    bad += not ok
    print(f"case {c}: {kind:10s} n={n} q={q} d={d} scale={scale:g} shift={shift:g}  err {e}  {'ok' if ok else 'MISMATCH'}")
print("mismatches:", bad)
