"""b200_knn_exact_noalloc: the exact kNN entry point that performs no device allocation.

Who needs it: callers that cannot let a library call cudaMalloc while they run -- a
serving loop that cannot afford the device-wide synchronisation an allocation implies,
or a process whose device memory is already fully handed out, where a late allocation
would fail.  Its split-scan scratch is a fixed 12 MiB array inside the library, which
bounds a call to q*k <= 262144 and k <= 128.

The test drives it through the C-ABI with caller-owned device buffers and compares
with the existing allocating entry point (b200_knn_dev, exact path) -- indices and
squared distances must be identical -- on two calls in a row with different shapes,
so that whatever the first call left in the shared scratch cannot leak into the
second.  Collected last (file name) so a failure here cannot hide other tests.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_noalloc_knn_matches_allocating_entry_point_twice(oracle):
    import torch
    from util_b200 import _lib
    lib = _lib.lib()
    assert lib.b200_device_count() >= 1
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    stream = torch.cuda.current_stream().cuda_stream
    rng = np.random.default_rng(11)
    # (n, q, d, k): the first call fills much of the scratch, the second is smaller and differently shaped
    # d = 120 and d = 300 take the launcher's 64- and 32-query block shapes
    for (n, q, d, k) in [(60000, 4000, 16, 16), (9000, 700, 5, 3), (5000, 300, 120, 4), (3000, 200, 300, 2)]:
        x = rng.random((n, d))
        z = rng.random((q, d))
        x_dev = torch.from_numpy(x).to(dev)
        z_dev = torch.from_numpy(z).to(dev)
        idx_ref = torch.empty((q, k), dtype=torch.int32, device=dev)
        d2_ref = torch.empty((q, k), dtype=torch.float64, device=dev)
        model = _lib.Model(x, device=0)
        model.set_option("knn_path", _lib.PATH_EXACT)
        _lib.check(lib.b200_knn_dev(model.handle, z_dev.data_ptr(), q, k, -1, idx_ref.data_ptr(),
                                    d2_ref.data_ptr(), stream))
        idx = torch.full((q, k), -7, dtype=torch.int32, device=dev)
        d2 = torch.full((q, k), -1.0, dtype=torch.float64, device=dev)
        _lib.check(lib.b200_knn_exact_noalloc(x_dev.data_ptr(), n, d, z_dev.data_ptr(), q, k,
                                              idx.data_ptr(), d2.data_ptr(), stream))
        torch.cuda.synchronize()
        # integer indices and fp64 distances from identical arithmetic: exact equality, as in the other tests
        assert torch.equal(idx, idx_ref)
        assert torch.equal(d2, d2_ref)
        # and against the CPU oracle on a sample of the queries (not only CUDA against CUDA)
        sel = rng.choice(q, size=min(q, 256), replace=False)
        oi, od = oracle.knn(x, z[sel], k)
        assert np.array_equal(idx.cpu().numpy()[sel], oi)
        assert np.array_equal(d2.cpu().numpy()[sel], od)
        model.close()
    # requests beyond the static scratch are refused with a clear message, not truncated
    rc = lib.b200_knn_exact_noalloc(x_dev.data_ptr(), n, d, z_dev.data_ptr(), 300000, 3, idx.data_ptr(),
                                    d2.data_ptr(), stream)
    assert rc != 0 and b"static scratch" in lib.b200_last_error()
    rc = lib.b200_knn_exact_noalloc(x_dev.data_ptr(), n, d, z_dev.data_ptr(), 10, 200, idx.data_ptr(),
                                    d2.data_ptr(), stream)
    assert rc != 0 and b"k=200" in lib.b200_last_error()
    # a dimension whose query block cannot fit shared memory is refused cleanly as well
    # (nothing is launched, so the buffers need not match the claimed shape)
    rc = lib.b200_knn_exact_noalloc(x_dev.data_ptr(), 4, 700, z_dev.data_ptr(), 2, 1, idx.data_ptr(),
                                    d2.data_ptr(), stream)
    assert rc != 0 and b"too large" in lib.b200_last_error()
