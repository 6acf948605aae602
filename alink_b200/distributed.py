"""Row-sharded execution over torch.distributed: one process per GPU, rows split by Alink's
DefaultDistributedInfo formula, centroids replicated, one SUM all-reduce of the k x (d+1) accumulator per
superstep (the reference's only parallelism strategy: KMeansTrainBatchOp.java:69-76, AllReduce.java:85-120).

torch.distributed is plumbing only: it carries the NCCL unique id to every rank (the all-reduce itself runs
inside libakm on the handle's stream) and provides the host-buffer all-reduce used by the Java-like flow
and by the CPU (gloo) tests of the host logic.
"""
import numpy as np

from .kmeans import shard_bounds


class TorchDistributed:
    def __init__(self, device=None):
        import torch
        import torch.distributed as dist
        if not dist.is_initialized():
            raise RuntimeError("torch.distributed is not initialised")
        self.torch, self.dist = torch, dist
        self.rank, self.world = dist.get_rank(), dist.get_world_size()
        self.backend = dist.get_backend()
        self.device = device if device is not None else ("cuda" if self.backend == "nccl" else "cpu")

    def _reduce(self, arr, op):
        t = self.torch.as_tensor(np.ascontiguousarray(arr, dtype=np.float64)).to(self.device)
        self.dist.all_reduce(t, op=op)
        return t.cpu().numpy()

    def all_sum(self, arr):
        return self._reduce(arr, self.dist.ReduceOp.SUM)

    def all_max(self, arr):
        return self._reduce(arr, self.dist.ReduceOp.MAX)

    # AllReduce(CENTROID_ALL_REDUCE) on a host buffer (deterministic ring/tree order, unlike the
    # arrival-order sum of AllReduce.java:307-316)
    allreduce = all_sum

    def gather(self, obj):
        """[obj of rank 0, obj of rank 1, ...] on every rank."""
        out = [None] * self.world
        self.dist.all_gather_object(out, obj)
        return out

    def shard(self, total):
        return shard_bounds(self.rank, self.world, total)

    def attach(self, handle):
        """Give `handle` an NCCL communicator: rank 0 creates the id, everybody receives it."""
        from . import native
        obj = [native.unique_id() if self.rank == 0 else None]
        self.dist.broadcast_object_list(obj, src=0)
        handle.comm_init(obj[0], self.rank, self.world)

    def sample_rows(self, X, k, rng):
        """k distinct global rows (same choice on every rank), gathered from their owners."""
        counts = [None] * self.world
        self.dist.all_gather_object(counts, int(X.shape[0]))
        total = int(sum(counts))
        starts = np.concatenate([[0], np.cumsum(counts)])
        pick = np.sort(rng.choice(total, size=min(k, total), replace=False))
        mine = [(int(g), X[g - starts[self.rank]].copy()) for g in pick
                if starts[self.rank] <= g < starts[self.rank + 1]]
        allp = [None] * self.world
        self.dist.all_gather_object(allp, mine)
        rows = dict(p for part in allp for p in part)
        return np.vstack([rows[int(g)] for g in pick])
