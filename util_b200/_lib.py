"""ctypes binding of libb200approx.so (include/b200approx.h).

This is the stub a maintainer of the reference would add in place of the
fmodpy-generated `<name>_from_py.py` wrappers
(util/approximate/shepard/weight/weight_from_py.py:16-21,54-85).  The library is
the only compute path: if it is missing, or no CUDA device is present, calls
raise -- there is no CPU fallback.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# B200_LIB: developer hook for A/B runs of two builds of the same library (profiles/); never a different backend
LIB_PATH = os.environ.get("B200_LIB") or os.path.join(_HERE, "csrc", "libb200approx.so")

AVERAGE, SHEPARD, VORONOI = 0, 1, 2
PATH_AUTO, PATH_EXACT, PATH_TENSOR = 0, 1, 2

c_dp = C.POINTER(C.c_double)
c_i64p = C.POINTER(C.c_int64)
c_i32p = C.POINTER(C.c_int32)

_SIGNATURES = {
    "b200_version": (C.c_int, []),
    "b200_last_error": (C.c_char_p, []),
    "b200_device_count": (C.c_int, []),
    "b200_create": (C.c_int, [c_dp, C.c_int64, C.c_int32, C.c_int32, C.POINTER(C.c_void_p)]),
    "b200_destroy": (C.c_int, [C.c_void_p]),
    "b200_set_option": (C.c_int, [C.c_void_p, C.c_char_p, C.c_double]),
    "b200_get_stat": (C.c_int, [C.c_void_p, C.c_char_p, c_dp]),
    "b200_set_values": (C.c_int, [C.c_void_p, c_dp, C.c_int32]),
    "b200_knn": (C.c_int, [C.c_void_p, c_dp, C.c_int64, C.c_int32, C.c_int64, c_i64p, c_dp]),
    "b200_nn_predict": (C.c_int, [C.c_void_p, c_dp, C.c_int64, C.c_int32, C.c_int32, c_dp, C.c_int32,
                                  c_i64p, c_dp, c_i32p, c_i32p, c_dp]),
    "b200_shepard_predict": (C.c_int, [C.c_void_p, c_dp, C.c_int64, c_dp, C.c_int32, c_dp, c_dp]),
    "b200_shepmod_fit": (C.c_int, [C.c_void_p, c_dp, c_i32p]),
    "b200_eval_mesh": (C.c_int, [C.c_int32, c_dp, C.c_int32, c_dp, C.c_int64, c_dp]),
    "b200_voronoi_predict": (C.c_int, [C.c_void_p, c_dp, C.c_int64, C.c_int32, c_i64p, c_dp, c_i32p,
                                       c_i32p, c_dp, C.c_int32, c_dp]),
    "b200_shepmod_predict": (C.c_int, [C.c_void_p, c_dp, C.c_int64, C.c_int32, c_i64p, c_dp, c_i32p,
                                       c_i32p, c_dp, C.c_int32, c_dp]),
    "b200_lshep_fit": (C.c_int, [C.c_void_p, c_dp, C.c_int32, c_dp, c_dp, c_i32p]),
    "b200_lshep_fit_robust": (C.c_int, [C.c_void_p, c_dp, c_dp, c_dp, c_i32p]),
    "b200_ripple_fit": (C.c_int, [C.c_void_p, c_dp, c_dp, c_dp, c_i32p]),
    "b200_lshep_predict": (C.c_int, [C.c_void_p, C.c_double, c_dp, C.c_int64, c_dp, c_i32p]),
    "b200_knn_dev": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int64, C.c_void_p,
                               C.c_void_p, C.c_void_p]),
    "b200_nn_weights_dev": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int32,
                                      C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p,
                                      C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "b200_knn_exact_noalloc": (C.c_int, [C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_int64, C.c_int32,
                                         C.c_void_p, C.c_void_p, C.c_void_p]),
    "b200_tc_debug_scores": (C.c_int, [C.c_void_p, c_dp, C.c_int64, C.c_int32, C.POINTER(C.c_float),
                                       C.c_int64]),
    "b200_multi_create": (C.c_int, [c_dp, C.c_int64, C.c_int32, C.c_int32, c_i32p, C.c_int32, C.POINTER(C.c_void_p)]),
    "b200_multi_destroy": (C.c_int, [C.c_void_p]),
    "b200_multi_set_option": (C.c_int, [C.c_void_p, C.c_char_p, C.c_double]),
    "b200_multi_get_stat": (C.c_int, [C.c_void_p, C.c_char_p, c_dp]),
    "b200_multi_set_values": (C.c_int, [C.c_void_p, c_dp, C.c_int32]),
    "b200_multi_nn_predict": (C.c_int, [C.c_void_p, c_dp, C.c_int64, C.c_int32, C.c_int32, c_i64p, c_dp, c_dp]),
    "b200_debug_fma_peak": (C.c_int, [C.c_int32, c_dp, c_dp]),
    "b200_merge_topk_dev": (C.c_int, [C.c_int32, C.c_void_p, C.c_void_p, C.c_int32, C.c_int64,
                                      C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]),
}

_lib = None


class B200Error(RuntimeError):
    pass


class NonFiniteInput(B200Error, ValueError):
    """NaN or infinity in the query points; sklearn's BallTree.query raises ValueError for the same input."""


def lib():
    """Load the shared library; fail loudly when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise B200Error(
                f"{LIB_PATH} is missing: build it with `python -m util_b200.build` "
                "(nvcc, sm_100a). util_b200 has no CPU fallback.")
        l = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(l, name)
            fn.restype = res
            fn.argtypes = args
        _lib = l
    return _lib


def check(rc):
    if rc != 0:
        text = lib().b200_last_error().decode("utf-8", "replace") or f"libb200approx error {rc}"
        raise (NonFiniteInput if "NaN or infinity" in text else B200Error)(text)


def f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def ptr(a, t=c_dp):
    return None if a is None else a.ctypes.data_as(t)


class MultiModel:
    """Owner of one `b200_multi` handle: the fitted points on several GPUs of this box, driven from this
    process (queries sharded, or the fitted points sharded with an NCCL exchange of the partial lists)."""

    def __init__(self, x, devices, shard="query"):
        x = f64(x)
        if shard not in ("query", "data"):
            raise ValueError("shard must be 'query' or 'data'")
        self.n, self.d = x.shape
        self.devices = [int(v) for v in devices]
        self.shard = shard
        devs = np.asarray(self.devices, dtype=np.int32)
        self._h = C.c_void_p()
        check(lib().b200_multi_create(ptr(x), self.n, self.d, len(devs), ptr(devs, c_i32p), int(shard == "data"),
                                      C.byref(self._h)))
        self.values_r = 0

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            lib().b200_multi_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_option(self, key, value):
        check(lib().b200_multi_set_option(self._h, key.encode(), float(value)))

    def stat(self, key):
        v = C.c_double()
        check(lib().b200_multi_get_stat(self._h, key.encode(), C.byref(v)))
        return v.value

    def set_values(self, y):
        y = f64(y).reshape(self.n, -1)
        check(lib().b200_multi_set_values(self._h, ptr(y), y.shape[1]))
        self.values_r = y.shape[1]

    def knn(self, z, k, self_offset=-1, want_d2=False):
        """Neighbour indices (q,k) ascending by (distance, index); distances are not exchanged."""
        if self_offset != -1 or want_d2:
            raise NotImplementedError("MultiModel.knn returns indices only")
        return self.nn_predict(z, k, AVERAGE)[0]

    def nn_predict(self, z, k, method, y=None, want_weights=True):
        """Same return convention as Model.nn_predict: (ids, wts, cnt, err, out); y=True uses set_values()."""
        z = f64(z)
        q = z.shape[0]
        ids = wts = out = None
        if want_weights:
            ids = np.empty((q, k), dtype=np.int64)
            wts = np.empty((q, k), dtype=np.float64)
        if y is not None:
            if y is not True:
                self.set_values(y)
            out = np.empty((q, self.values_r))
        check(lib().b200_multi_nn_predict(self._h, ptr(z), q, k, method, ptr(ids, c_i64p), ptr(wts), ptr(out)))
        return ids, wts, np.full(q, k, dtype=np.int32), np.zeros(q, dtype=np.int32), out


class Model:
    """Owner of one `b200_model` handle (the fitted points resident on one GPU)."""

    def __init__(self, x, device=None):
        x = f64(x)
        if x.ndim != 2:
            raise ValueError("x must be 2-D")
        if device is None:
            device = int(os.environ.get("LOCAL_RANK", "0")) if os.environ.get("B200_USE_LOCAL_RANK") else 0
        self.n, self.d = x.shape
        self.device = device
        self._h = C.c_void_p()
        check(lib().b200_create(ptr(x), self.n, self.d, device, C.byref(self._h)))

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            lib().b200_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def handle(self):
        return self._h

    def set_option(self, key, value):
        check(lib().b200_set_option(self._h, key.encode(), float(value)))

    def stat(self, key):
        v = C.c_double()
        check(lib().b200_get_stat(self._h, key.encode(), C.byref(v)))
        return v.value

    def set_values(self, y):
        """Keep the response values (n,) or (n,r) on the device; predict calls made with y=True use them."""
        y = f64(y).reshape(self.n, -1)
        check(lib().b200_set_values(self._h, ptr(y), y.shape[1]))
        self.values_r = y.shape[1]

    def _y_args(self, y, q):
        """(host y or None, r, out): y=True selects the values left on the device by set_values()."""
        if y is None:
            return None, 0, None
        if y is True:
            return None, self.values_r, np.empty((q, self.values_r))
        y = f64(y).reshape(self.n, -1)
        return y, y.shape[1], np.empty((q, y.shape[1]))

    def knn(self, z, k, self_offset=-1, want_d2=True):
        z = f64(z)
        q = z.shape[0]
        idx = np.empty((q, k), dtype=np.int64)
        d2 = np.empty((q, k), dtype=np.float64) if want_d2 else None
        check(lib().b200_knn(self._h, ptr(z), q, k, self_offset, ptr(idx, c_i64p), ptr(d2)))
        return (idx, d2) if want_d2 else idx

    def nn_predict(self, z, k, method, y=None, want_weights=True):
        """-> ids (q,k), wts (q,k), cnt (q), err (q), out (q,r) or None."""
        z = f64(z)
        q = z.shape[0]
        ids = wts = None
        if want_weights:
            ids = np.empty((q, k), dtype=np.int64)
            wts = np.empty((q, k), dtype=np.float64)
        cnt = np.empty(q, dtype=np.int32)
        err = np.empty(q, dtype=np.int32)
        y, r, out = self._y_args(y, q)
        check(lib().b200_nn_predict(self._h, ptr(z), q, k, method, ptr(y), r, ptr(ids, c_i64p), ptr(wts),
                                    ptr(cnt, c_i32p), ptr(err, c_i32p), ptr(out)))
        return ids, wts, cnt, err, out

    def tc_debug_scores(self, z, k=1):
        """Raw stage-1 scores of the tensor path (tests only)."""
        z = f64(z)
        cols = ((self.n + 127) // 128) * 128
        out = np.empty((z.shape[0], cols), dtype=np.float32)
        check(lib().b200_tc_debug_scores(self._h, ptr(z), z.shape[0], k, ptr(out, C.POINTER(C.c_float)), cols))
        return out

    def shepard_predict(self, z, y=None, want_weights=False):
        """Stand-alone Shepard over all fitted points: (wts (q,n) or None, out (q,r) or None)."""
        z = f64(z)
        q = z.shape[0]
        wts = np.empty((q, self.n)) if want_weights else None
        y, r, out = self._y_args(y, q)
        check(lib().b200_shepard_predict(self._h, ptr(z), q, ptr(y), r, ptr(wts), ptr(out)))
        return wts, out

    def shepmod_fit(self):
        rw = np.empty(self.n)
        ier = C.c_int32(0)
        check(lib().b200_shepmod_fit(self._h, ptr(rw), C.byref(ier)))
        return rw, ier.value

    def shepmod_predict(self, z, cap=64, y=None, want_weights=True):
        """want_weights=False (with y): only the fused sums, counts and codes come back; the hit lists stay
        on the device (returned ids/wts are then None, `cap` still grows until the longest list fits)."""
        z = f64(z)
        q = z.shape[0]
        want_sums = y is not None
        y, r, out = self._y_args(y, q)
        while True:
            keep = want_weights or not want_sums
            ids = np.empty((q, cap), dtype=np.int64) if keep else None
            wts = np.empty((q, cap)) if keep else None
            cnt = np.empty(q, dtype=np.int32)
            ier = np.empty(q, dtype=np.int32)
            check(lib().b200_shepmod_predict(self._h, ptr(z), q, cap, ptr(ids, c_i64p), ptr(wts),
                                             ptr(cnt, c_i32p), ptr(ier, c_i32p), ptr(y), r, ptr(out)))
            mx = int(cnt.max()) if q else 0
            if mx <= cap:
                self.last_cap = cap
                return ids, wts, cnt, ier, out
            cap = max(2 * cap, mx)  # a hit list overflowed: repeat with room for the longest

    def voronoi_predict(self, z, cap=32, y=None):
        """Stand-alone Voronoi over all fitted points: per query the supporting points (ascending
        index) and weights; the list capacity grows until the longest fits."""
        z = f64(z)
        q = z.shape[0]
        y, r, out = self._y_args(y, q)
        cap = max(1, min(cap, self.n))
        while True:
            ids = np.empty((q, cap), dtype=np.int64)
            wts = np.empty((q, cap))
            cnt = np.empty(q, dtype=np.int32)
            err = np.empty(q, dtype=np.int32)
            check(lib().b200_voronoi_predict(self._h, ptr(z), q, cap, ptr(ids, c_i64p), ptr(wts),
                                             ptr(cnt, c_i32p), ptr(err, c_i32p), ptr(y), r, ptr(out)))
            mx = int(cnt.max()) if q else 0
            if mx <= cap:
                return ids, wts, cnt, err, out
            cap = min(self.n, max(2 * cap, mx))

    def ripple_fit(self, f):
        """RIPPLE (linear_shepard.f95:474-665) for one response column: (a (1,n,d), rw (n), ier)."""
        f = f64(f).reshape(self.n)
        a = np.empty((1, self.n, self.d))
        rw = np.empty(self.n)
        ier = C.c_int32(0)
        check(lib().b200_ripple_fit(self._h, ptr(f), ptr(a), ptr(rw), C.byref(ier)))
        return a, rw, ier.value

    def lshep_fit(self, f, robust=False):
        f = f64(f).reshape(self.n, -1)
        r = f.shape[1]
        a = np.empty((r, self.n, self.d))
        rw = np.empty(self.n)
        ier = C.c_int32(0)
        if robust:
            if r != 1:
                raise ValueError("the robust LSHEP fit takes one response column (its radii depend on the response)")
            check(lib().b200_lshep_fit_robust(self._h, ptr(f), ptr(a), ptr(rw), C.byref(ier)))
        else:
            check(lib().b200_lshep_fit(self._h, ptr(f), r, ptr(a), ptr(rw), C.byref(ier)))
        return a, rw, ier.value

    def lshep_predict(self, z, radius, r):
        z = f64(z)
        q = z.shape[0]
        out = np.empty((q, r))
        ier = np.empty((q, r), dtype=np.int32)
        check(lib().b200_lshep_predict(self._h, float(radius), ptr(z), q, ptr(out), ptr(ier, c_i32p)))
        return out, ier
