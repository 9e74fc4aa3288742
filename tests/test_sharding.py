"""Host-side logic of the data-sharded path on CPU: two processes, gloo backend.
Each rank plays a GPU: local exact top-k on its shard comes from the CPU oracle (the
checker, standing in for the kernel), the partial lists go through the SAME
shard arithmetic and all-to-all plumbing the GPU path uses (util_b200.sharding),
and the merged result must equal the global oracle answer index for index."""
import os
import socket

import numpy as np
import pytest

from conftest import synth


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n, q, d, k, ret):
    import torch
    import torch.distributed as dist
    from oracle import harness as H
    from util_b200 import sharding
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    H.set_threads(2)
    x, z = synth(n, q, d)
    lo, hi = sharding.shard_bounds(n, world, rank)
    idx, d2 = H.knn(x[lo:hi], z, k)                       # local top-k of ALL queries on my shard
    idx = torch.from_numpy((idx + lo).astype(np.int32))   # global row ids
    d2 = torch.from_numpy(d2)
    rd2, ridx = sharding.exchange_partials(d2, idx)
    qlo, qhi = sharding.query_slice(q, world, rank)
    # checker-side merge of the `world` sorted lists by (d2, index)
    pd2 = rd2.numpy().transpose(1, 0, 2).reshape(qhi - qlo, world * k)
    pidx = ridx.numpy().transpose(1, 0, 2).reshape(qhi - qlo, world * k)
    order = np.lexsort((pidx, pd2), axis=1)[:, :k]
    mi = np.take_along_axis(pidx, order, axis=1)
    md = np.take_along_axis(pd2, order, axis=1)
    gi, gd = H.knn(x, z[qlo:qhi], k)
    ret[rank] = bool(np.array_equal(mi, gi) and np.array_equal(md, gd))
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_data_sharded_merge_matches_global(world):
    import torch.multiprocessing as mp
    n, q, d, k = 3001, 6 * world * 5, 6, 7
    port = _free_port()
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, q, d, k, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert all(ret[r] for r in range(world))


def test_shard_arithmetic():
    from util_b200 import sharding
    for n in (1, 7, 1000, 1_000_000):
        for w in (1, 2, 4, 8):
            b = [sharding.shard_bounds(n, w, r) for r in range(w)]
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[i][1] == b[i + 1][0] for i in range(w - 1))
    assert sharding.query_slice(16, 4, 3) == (12, 16)
    with pytest.raises(ValueError):
        sharding.query_slice(10, 4, 0)
