"""Multi-GPU tests (run only when >= 2 CUDA devices are visible): the data-sharded kNN +
Shepard prediction over NCCL must be bit-identical to the single-GPU path."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu

SCRIPT = r'''
import os, sys, numpy as np, torch, torch.distributed as dist
sys.path.insert(0, os.environ["B200_ROOT"])
from util_b200 import _lib, sharding
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dev = torch.device("cuda", rank)
dist.init_process_group("nccl", device_id=dev)
n, q, d, k = 50000, 4096, 16, 8
x = np.random.default_rng(0).random((n, d)); z = np.random.default_rng(1).random((q, d))
y = 3 * x[:, 0] + .5 * np.cos(8 * x[:, 0]) + np.sin(5 * x[:, -1])
sh = sharding.DataShardedKNN(x, rank)
z_dev = torch.from_numpy(z).to(dev); y_dev = torch.from_numpy(y).to(dev)
out, midx, md2 = sh.predict(z_dev, k, _lib.SHEPARD, y_dev)
torch.cuda.synchronize()
full = _lib.Model(x, device=rank)
ids, wts, cnt, err, ref = full.nn_predict(z, k, _lib.SHEPARD, y=y)
lo, hi = sharding.query_slice(q, world, rank)
ok = np.array_equal(midx.cpu().numpy(), ids[lo:hi].astype(np.int32)) and np.array_equal(out.cpu().numpy(), ref[lo:hi, 0])
t = torch.tensor([1 if ok else 0], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MIN)
if rank == 0: print("SHARDED_OK" if int(t.item()) == 1 else "SHARDED_MISMATCH")
dist.destroy_process_group()
'''


def test_data_sharded_two_gpus(tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    f = tmp_path / "w.py"
    f.write_text(SCRIPT)
    env = dict(os.environ, B200_ROOT=ROOT)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", str(port), str(f)],
                       capture_output=True, text=True, timeout=600, env=env)
    assert "SHARDED_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


def test_one_process_handle_over_two_gpus(oracle):
    """b200_multi_*: NearestNeighbor(devices=[0, 1]) from one process, no torch -- queries sharded, and the
    fitted points sharded with the NCCL exchange + on-device merge.  Both must reproduce the oracle's
    neighbours bit for bit and the single-GPU predictions exactly."""
    from util_b200 import _lib
    import util_b200 as ub
    if _lib.lib().b200_device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    n, q, d, k = 60000, 5001, 16, 8   # q not a multiple of the device count: padded slices
    x = np.random.default_rng(0).random((n, d)); z = np.random.default_rng(1).random((q, d))
    y = 3 * x[:, 0] + .5 * np.cos(8 * x[:, 0]) + np.sin(5 * x[:, -1])
    oi, od = oracle.knn(x, z[:300], k)
    one = ub.NearestNeighbor(k=k, method=ub.Shepard)
    one.fit(x, y)
    ref = np.array(one(z))
    for shard in ("query", "data"):
        m = ub.NearestNeighbor(k=k, method=ub.Shepard, devices=[0, 1], shard=shard)
        m.fit(x, y)
        pred = np.array(m(z))
        assert np.array_equal(pred, ref), shard
        ids, wts = m._predict(z[:300])
        assert np.array_equal(ids, oi), shard
        if shard == "data":
            assert m.model.stat("ms_collective") > 0 and m.model.stat("ms_merge") > 0
