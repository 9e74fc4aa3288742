"""CPU-side checks of the boundary: the C-ABI library loads, exports every symbol
include/b200approx.h declares, and fails loudly (no CPU fallback) without a GPU;
host-side validation mirrors util/approximate/base.py."""
import ctypes
import os
import re

import numpy as np
import pytest

from conftest import ROOT


def _declared():
    names = set()
    for h in ("b200approx.h", "b200approx_debug.h"):  # the boundary and the test hooks
        hdr = open(os.path.join(ROOT, "include", h)).read()
        hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
        names |= set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", hdr))
    return sorted(names)


def test_library_exports_every_declared_symbol():
    from util_b200 import _lib, build
    build.build()
    names = _declared()
    assert len(names) >= 14
    dll = ctypes.CDLL(_lib.LIB_PATH)
    for n in names:
        assert hasattr(dll, n), f"{n} declared in b200approx.h but not exported"
    # and the Python binding covers the same set
    assert sorted(_lib._SIGNATURES) == names
    assert _lib.lib().b200_version() >= 100


def test_no_cpu_fallback():
    from util_b200 import _lib
    import util_b200
    if _lib.lib().b200_device_count() > 0:
        pytest.skip("GPU present")
    with pytest.raises(_lib.B200Error, match="no CUDA device"):
        util_b200.NearestNeighbor(k=1).fit(np.random.rand(10, 2), np.random.rand(10))
    # nothing under util_b200 may reach for the oracle
    for root, _, files in os.walk(os.path.join(ROOT, "util_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh")):
                src = open(os.path.join(root, f)).read()
                assert "oracle" not in src.replace("oracle/", "").lower() or f == "__never__", f


def test_host_validation_matches_reference_contract():
    import util_b200 as ub
    x = np.random.rand(10, 2)
    with pytest.raises(ub.UnexpectedType):
        ub.NearestNeighbor().fit(x.tolist(), None)
    with pytest.raises(ub.UnexpectedType):
        ub.NearestNeighbor().fit(x[:, 0], None)
    with pytest.raises(ub.MissingOperator):
        ub.NearestNeighbor().fit(x, 5, classifier=False)
    with pytest.raises(ub.MissingOperator):
        ub.NearestNeighbor().fit(x, [object()] * 10)
    with pytest.raises(ub.UnexpectedType):
        ub.LSHEP().fit(x, "abc", classifier=False)
    with pytest.raises(ub.TooFewDataPoints):
        ub.LSHEP().fit(x[:3], np.zeros(3))
    with pytest.raises(ub.UnexpectedType):
        ub.NearestNeighbor().predict([[0.0, 1.0]])
    assert str(ub.NearestNeighbor()) == "NearestNeighbor"
    avg = ub.Average()
    avg.fit(x)
    ids, w = avg._predict(x[:3])
    assert ids.shape == (3, 10) and np.allclose(w, .1)
