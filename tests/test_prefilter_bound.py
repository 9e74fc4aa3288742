"""The fp32 prefilter of the exact kNN / ShepMod / LSHEP scans (util_b200/csrc/common.cuh: pre_gap, pre_thr,
pre_eps1, pre_up) restated in numpy float32: whenever it rules a point out (af >= thr(L)), the point's true
distance must be >= L.  Checked on coordinates of very different magnitudes and on distances right at L."""
import numpy as np

f32 = np.float32


def pre_gap(d, cm):
    return f32(2.0001 * 5.9604644775390625e-08 * np.sqrt(float(d)) * cm) * f32(1.000001)


def pre_eps1(d):
    return f32(1.0) + f32(d + 16) * f32(2.384185791015625e-07)


def pre_up(L):
    return f32(L) * f32(1.0000005)


def pre_thr(L_up, g0, eps1):
    lim = (L_up + g0) * f32(1.0000005)
    return lim * lim * eps1


def af32(z, x):
    """the kernels' fp32 squared distance: float-rounded coordinates, fma chain in dimension order"""
    zf, xf = z.astype(f32), x.astype(f32)
    acc = np.zeros(len(x), dtype=f32)
    for j in range(x.shape[1]):
        t = (zf[j] - xf[:, j]).astype(f32)
        acc = (t.astype(np.float64) * t.astype(np.float64) + acc.astype(np.float64)).astype(f32)  # fmaf: one rounding
    return acc


def test_prefilter_never_rules_out_a_point_inside():
    rng = np.random.default_rng(0)
    ruled_out = 0
    for trial in range(300):
        d = int(rng.choice([1, 2, 3, 5, 8, 16, 33, 64]))
        scale = float(10.0 ** rng.integers(-6, 7))
        shift = float(rng.choice([0.0, 0.0, 1.0, -37.5, 1e4])) * scale
        z = rng.random(d) * scale + shift
        x = rng.random((400, d)) * scale + shift
        # half the points at distances very close to the limit L
        L = float(np.linalg.norm(x[0] - z))
        dirs = rng.normal(size=(200, d))
        dirs /= np.linalg.norm(dirs, axis=1)[:, None]
        x[200:] = z + dirs * L * (1.0 + rng.uniform(-3e-6, 3e-6, size=(200, 1)))
        cm = max(np.abs(x).max(), np.abs(z).max())
        thr = pre_thr(pre_up(L), pre_gap(d, cm), pre_eps1(d))
        af = af32(z, x)
        # exact distances in extended precision
        dist = np.sqrt(((x.astype(np.longdouble) - z.astype(np.longdouble)) ** 2).sum(1))
        out = af >= thr
        ruled_out += int(out.sum())
        assert np.all(dist[out] >= np.longdouble(L)), (trial, d, scale, shift)
    assert ruled_out > 10000  # the filter does rule points out on these inputs
