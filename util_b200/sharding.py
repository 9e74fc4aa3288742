"""Multi-GPU partitioning of the kNN + weighting path (one process per GPU).

The reference has no distributed code; this is the B200 design
for BASELINE.json configs[4]: the fitted points are sharded over the ranks,
every rank finds the exact local top-k of all queries against its shard, the
partial lists are exchanged with ONE all-to-all over NCCL (rank r receives the
lists of its slice of the queries from everybody) and merged on the device by
(d2, global index) -- bit-identical to the single-GPU answer -- and each rank
weights its slice.  When the data set fits one GPU the queries can be sharded
instead (no collective at all).

Only the plumbing lives here (shard arithmetic + the torch.distributed calls);
every computation is a kernel of libb200approx.  `torch` is imported lazily so
the single-GPU package has no torch dependency.
"""
import numpy as np


def shard_bounds(n, world, rank):
    """Contiguous block [lo, hi) of `n` rows owned by `rank`."""
    per = (n + world - 1) // world
    lo = min(n, rank * per)
    return lo, min(n, lo + per)


def query_slice(q, world, rank):
    """Queries a rank merges and weights under data sharding; q must divide evenly."""
    if q % world:
        raise ValueError(f"query count {q} must be a multiple of the number of ranks {world}")
    per = q // world
    return rank * per, (rank + 1) * per


def exchange_partials(d2, idx, group=None):
    """All-to-all of per-rank partial top-k lists.

    d2 (q,k) float64 and idx (q,k) int32 are this rank's lists for ALL queries with
    GLOBAL row indices.  Returns (parts_d2, parts_idx) of shape (world, q/world, k):
    entry [g] is rank g's list for this rank's slice of the queries.  Works on CUDA
    tensors over NCCL and on CPU tensors over gloo (tests).
    """
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    q, k = d2.shape
    qm = q // world
    rd2 = torch.empty((world, qm, k), dtype=d2.dtype, device=d2.device)
    ridx = torch.empty((world, qm, k), dtype=idx.dtype, device=idx.device)
    dist.all_to_all_single(rd2.view(-1), d2.contiguous().view(-1), group=group)
    dist.all_to_all_single(ridx.view(-1), idx.contiguous().view(-1), group=group)
    return rd2, ridx


class DataShardedKNN:
    """kNN (+ Average/Shepard weighting) with the fitted points sharded over the ranks of a
    torch.distributed NCCL group.  All tensors are CUDA tensors on this rank's device."""

    def __init__(self, x_full, device_index, group=None):
        import torch
        import torch.distributed as dist
        from . import _lib
        self._lib = _lib
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.n = x_full.shape[0]
        self.lo, self.hi = shard_bounds(self.n, self.world, self.rank)
        self.device_index = device_index
        self.model = _lib.Model(np.ascontiguousarray(x_full[self.lo:self.hi]), device=device_index)
        self.torch = torch
        self.ms_collective = self.ms_merge = 0.0  # accumulated CUDA-event times of the exchange and the merge
        self._timed = []

    def knn(self, z_dev, k, stream=None):
        """-> merged (d2, idx) for this rank's query slice: tensors (q/world, k)."""
        torch, lib = self.torch, self._lib.lib()
        q = z_dev.shape[0]
        st = stream if stream is not None else torch.cuda.current_stream().cuda_stream
        idx = torch.empty((q, k), dtype=torch.int32, device=z_dev.device)
        d2 = torch.empty((q, k), dtype=torch.float64, device=z_dev.device)
        self._lib.check(lib.b200_knn_dev(self.model.handle, z_dev.data_ptr(), q, k, -1, idx.data_ptr(),
                                         d2.data_ptr(), st))
        idx.add_(self.lo)  # local -> global row ids
        e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        e[0].record()
        rd2, ridx = exchange_partials(d2, idx, self.group)
        e[1].record()
        qm = q // self.world
        md2 = torch.empty((qm, k), dtype=torch.float64, device=z_dev.device)
        midx = torch.empty((qm, k), dtype=torch.int32, device=z_dev.device)
        self._lib.check(lib.b200_merge_topk_dev(self.device_index, rd2.data_ptr(), ridx.data_ptr(), self.world,
                                                qm, k, md2.data_ptr(), midx.data_ptr(), st))
        e[2].record()
        self._timed.append(e)
        return md2, midx

    def collect_times(self):
        """Fold the recorded CUDA events into ms_collective / ms_merge (call after a synchronize)."""
        for e in self._timed:
            self.ms_collective += e[0].elapsed_time(e[1])
            self.ms_merge += e[1].elapsed_time(e[2])
        n = len(self._timed)
        self._timed = []
        return n

    def predict(self, z_dev, k, method, y_dev, stream=None):
        """Weighted prediction for this rank's query slice (y_dev: all n responses, replicated)."""
        torch, lib = self.torch, self._lib.lib()
        st = stream if stream is not None else torch.cuda.current_stream().cuda_stream
        md2, midx = self.knn(z_dev, k, st)
        qm = md2.shape[0]
        out = torch.empty((qm,), dtype=torch.float64, device=z_dev.device)
        self._lib.check(lib.b200_nn_weights_dev(self.model.handle, None, qm, k, method, midx.data_ptr(),
                                                md2.data_ptr(), y_dev.data_ptr(), 1, None, None, None,
                                                out.data_ptr(), st))
        return out, midx, md2
