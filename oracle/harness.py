"""ctypes front end of the CPU oracle (oracle/liboracle.so).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py, never by util_b200.

Each function names the reference code it restates; see oracle.cpp and
oracle_lshep_voronoi.cpp for line-level citations.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

METHODS = {"average": 0, "shepard": 1, "voronoi": 2, "shepmod": 3}


def build(force=False):
    """Compile liboracle.so with the recipe in oracle/Makefile."""
    so = os.path.join(_HERE, "liboracle.so")
    srcs = [os.path.join(_HERE, f) for f in
            ("oracle.cpp", "oracle_lshep_voronoi.cpp", "oracle_robust.cpp", "oracle_common.h", "Makefile")]
    if force or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B", "liboracle.so"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(build())
        _LIB.orc_lshep_val.restype = C.c_double
    return _LIB


def _p(a, t=C.c_double):
    return a.ctypes.data_as(C.POINTER(t))


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def set_threads(t):
    lib().orc_set_threads(int(t))


def max_threads():
    return lib().orc_max_threads()


def knn(x, z, k):
    """Brute-force k nearest rows of x for each row of z: (idx int64 (q,k), d2 (q,k))."""
    x, z = _f64(x), _f64(z)
    q = z.shape[0]
    idx = np.empty((q, k), dtype=np.int64)
    d2 = np.empty((q, k), dtype=np.float64)
    rc = lib().orc_knn(_p(x), C.c_int64(x.shape[0]), x.shape[1], _p(z), C.c_int64(q), k,
                       _p(idx, C.c_int64), _p(d2))
    if rc:
        raise ValueError("orc_knn: bad k/d")
    return idx, d2


def shepard_weight(pts, pt):
    pts, pt = _f64(pts), _f64(pt)
    w = np.empty(pts.shape[0])
    lib().orc_shepard_weight(_p(pts), C.c_int64(pts.shape[0]), pts.shape[1], _p(pt), _p(w))
    return w


def shepmod_fit(x):
    """SHEPMOD: returns (rw, ier)."""
    x = _f64(x)
    rw = np.ones(x.shape[0])
    ier = lib().orc_shepmod_fit(_p(x), C.c_int64(x.shape[0]), x.shape[1], _p(rw))
    return rw, ier


def shepmod_val(x, rw, xp):
    """SHEPMODVAL on one point: returns (dense wts (n), ier)."""
    x, rw, xp = _f64(x), _f64(rw), _f64(xp)
    w = np.empty(x.shape[0])
    ier = lib().orc_shepmod_val(_p(xp), x.shape[1], C.c_int64(x.shape[0]), _p(x), _p(rw), _p(w))
    return w, ier


def shepmod_predict(x, rw, z, cap=None):
    """ShepMod._predict over a batch: (ids (q,cap) int64, wts (q,cap), cnt (q), ier (q))."""
    x, rw, z = _f64(x), _f64(rw), _f64(z)
    q = z.shape[0]
    cap = int(cap or x.shape[0])
    ids = np.full((q, cap), -1, dtype=np.int64)
    w = np.zeros((q, cap))
    cnt = np.zeros(q, dtype=np.int32)
    ier = np.zeros(q, dtype=np.int32)
    lib().orc_shepmod_predict(_p(x), C.c_int64(x.shape[0]), x.shape[1], _p(rw), _p(z), C.c_int64(q),
                              cap, _p(ids, C.c_int64), _p(w), _p(cnt, C.c_int32), _p(ier, C.c_int32))
    return ids, w, cnt, ier


def _dgelss_ptr():
    """Address of LAPACK DGELSS as shipped in scipy (linear_shepard.f95:132 links
    the system LAPACK; scipy's OpenBLAS build is the one available here)."""
    from scipy.linalg import cython_lapack
    cap = cython_lapack.__pyx_capi__["dgelss"]
    C.pythonapi.PyCapsule_GetName.restype = C.c_char_p
    C.pythonapi.PyCapsule_GetName.argtypes = [C.py_object]
    C.pythonapi.PyCapsule_GetPointer.restype = C.c_void_p
    C.pythonapi.PyCapsule_GetPointer.argtypes = [C.py_object, C.c_char_p]
    return C.pythonapi.PyCapsule_GetPointer(cap, C.pythonapi.PyCapsule_GetName(cap))


def lshep_fit(x, f):
    """LSHEP (non-robust): returns (a (n,d), rw (n) before the radius multiplier, ier, rank (n))."""
    x, f = _f64(x), _f64(f)
    n, d = x.shape
    a = np.ones((n, d))
    rw = np.ones(n)
    rank = np.zeros(n, dtype=np.int32)
    ier = lib().orc_lshep_fit(_p(x), C.c_int64(n), d, _p(f), _p(a), _p(rw), _p(rank, C.c_int32),
                              C.c_void_p(_dgelss_ptr()))
    return a, rw, ier, rank


def quantile(a, k):
    """QUANTILE (linear_shepard.f95:310-335): the k-th smallest (1-based) element."""
    a = _f64(a)
    lib().orc_quantile.restype = C.c_double
    return lib().orc_quantile(_p(a), len(a), int(k))


def mad(res):
    """MAD (linear_shepard.f95:284-308)."""
    res = _f64(res)
    lib().orc_mad.restype = C.c_double
    return lib().orc_mad(_p(res), len(res))


def lshep_fit_robust(x, f):
    """LSHEP with RLSHEP present (linear_shepard.f95:136-175): (a (n,d), rw (n), ier, omega (n,np))."""
    x, f = _f64(x), _f64(f)
    n, d = x.shape
    npn = min(n - 1, (3 * d + 1) // 2)
    a = np.ones((n, d))
    rw = np.ones(n)
    om = np.ones((n, npn))
    ier = lib().orc_lshep_fit_robust(_p(x), C.c_int64(n), d, _p(f), _p(a), _p(rw), _p(om),
                                     C.c_void_p(_dgelss_ptr()))
    return a, rw, ier, om


def ripple(x, f):
    """RIPPLE (linear_shepard.f95:474-665): (a (n,d), rw (n), ier)."""
    x, f = _f64(x), _f64(f)
    n, d = x.shape
    a = np.ones((n, d))
    rw = np.ones(n)
    ier = lib().orc_ripple(_p(x), C.c_int64(n), d, _p(f), _p(a), _p(rw), C.c_void_p(_dgelss_ptr()))
    return a, rw, ier


def lshep_predict(x, f, a, rw, z):
    """LSHEPVAL over a batch: (values (q), ier (q))."""
    x, f, a, rw, z = _f64(x), _f64(f), _f64(a), _f64(rw), _f64(z)
    q = z.shape[0]
    out = np.empty(q)
    ier = np.zeros(q, dtype=np.int32)
    lib().orc_lshep_predict(_p(x), C.c_int64(x.shape[0]), x.shape[1], _p(f), _p(a), _p(rw), _p(z),
                            C.c_int64(q), _p(out), _p(ier, C.c_int32))
    return out, ier


def voronoi_predict(pts, pt):
    """Voronoi PREDICT on one point: (dense weights (n), error)."""
    pts, pt = _f64(pts), _f64(pt)
    w = np.empty(pts.shape[0])
    err = lib().orc_voronoi_predict(_p(pts), C.c_int64(pts.shape[0]), pts.shape[1], _p(pt), _p(w))
    return w, err


def eval_mesh(int_prods, app_int_prods):
    """EVAL_MESH (voronoi.f90:5-44): (n,n) and (q,n) inner products -> (q,n) mesh."""
    ip = np.asfortranarray(int_prods, dtype=np.float64)
    ap = np.asfortranarray(app_int_prods, dtype=np.float64)
    mesh = np.empty(ap.shape, order="F")
    lib().orc_eval_mesh(ip.ctypes.data_as(C.POINTER(C.c_double)), C.c_int64(ip.shape[0]),
                        ap.ctypes.data_as(C.POINTER(C.c_double)), C.c_int64(ap.shape[0]),
                        mesh.ctypes.data_as(C.POINTER(C.c_double)))
    return mesh


def nn_predict(x, z, k, method="average"):
    """NearestNeighbor._predict with a sub-model: (ids (q,k), wts (q,k), cnt (q), err (q))."""
    x, z = _f64(x), _f64(z)
    q = z.shape[0]
    ids = np.empty((q, k), dtype=np.int64)
    w = np.empty((q, k))
    cnt = np.zeros(q, dtype=np.int32)
    err = np.zeros(q, dtype=np.int32)
    rc = lib().orc_nn_predict(_p(x), C.c_int64(x.shape[0]), x.shape[1], _p(z), C.c_int64(q), k,
                              METHODS[method], _p(ids, C.c_int64), _p(w), _p(cnt, C.c_int32),
                              _p(err, C.c_int32))
    if rc:
        raise ValueError("orc_nn_predict: bad k")
    return ids, w, cnt, err


def weighted_sum(ids, wts, cnt, y):
    """WeightedApproximator.predict's left-to-right sum(y[i]*w) (base.py:169)."""
    ids = np.ascontiguousarray(ids, dtype=np.int64)
    wts, y = _f64(wts), _f64(y)
    cnt = np.ascontiguousarray(cnt, dtype=np.int32)
    out = np.empty(ids.shape[0])
    lib().orc_weighted_sum(_p(ids, C.c_int64), _p(wts), _p(cnt, C.c_int32), C.c_int64(ids.shape[0]),
                           ids.shape[1], _p(y), _p(out))
    return out
