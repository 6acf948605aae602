"""Row-sharded training over 2 GPUs (NCCL all-reduce inside akm_step) against one GPU and the oracle.
Skipped on a single-GPU box."""
import os

import numpy as np
import pytest

import oracle as orc
from alink_b200 import native

pytestmark = pytest.mark.gpu


def _blobs(n, d, k, seed):
    rng = np.random.default_rng(seed)
    centers = rng.uniform(-10, 10, size=(k, d))
    return (centers[rng.integers(0, k, size=n)] + rng.normal(size=(n, d))).astype(np.float32).astype(np.float64)


def _worker(rank, world, port, X, C0, metric, iters, q):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    from alink_b200.distributed import TorchDistributed
    td = TorchDistributed()
    start, cnt = td.shard(X.shape[0])
    with native.Handle(X.shape[1], metric, device=rank) as h:
        h.upload_rows(X[start:start + cnt])
        td.attach(h)
        m = h.train(C0, max_iter=iters, tol=0.0)
        labels = h.get_labels()
    q.put((rank, start, labels, m["C"], m["k"], m["weights"], m["objective"]))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("metric", [native.EUCLIDEAN, native.COSINE])
def test_two_gpus_equal_one_gpu_and_the_oracle(metric):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp
    n, d, k, iters = 40001, 32, 24, 6                         # odd n: the shards differ in size
    X = _blobs(n, d, 12, seed=31)
    C0 = X[np.random.default_rng(1).choice(n, k, replace=False)].copy()
    ref = orc.train(metric, X, C0, max_iter=iters, tol=0.0, threads=2)      # two "subtasks", summed in rank order
    with native.Handle(d, metric) as h:
        h.upload_rows(X)
        one = h.train(C0, max_iter=iters, tol=0.0)
        one_labels = h.get_labels()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, X, C0, metric, iters, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = sorted([q.get(timeout=300) for _ in range(2)], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    labels = np.concatenate([g[2] for g in got])
    assert (labels == ref["labels"]).all() and (labels == one_labels).all()
    for g in got:                                             # every rank holds the same model
        assert g[4] == ref["k"] == one["k"]
        np.testing.assert_allclose(g[3][: g[4]], ref["C"][: ref["k"]], rtol=1e-10, atol=1e-12)
        np.testing.assert_allclose(g[3][: g[4]], one["C"][: one["k"]], rtol=1e-10, atol=1e-12)
        assert (g[5] == ref["weights"]).all()
        np.testing.assert_allclose(g[6], ref["objective"], rtol=1e-9)
    assert (got[0][3] == got[1][3]).all()
