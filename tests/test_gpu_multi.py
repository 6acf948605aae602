"""Row-sharded training over 2 (and 4) GPUs (NCCL all-reduce inside akm_step and akm_superstep) against one GPU and
the oracle.  Skipped on a box with fewer GPUs."""
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


def _worker(rank, world, port, X, C0, metric, iters, q, mode="step"):
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
        if mode == "step":                                   # device-resident loop
            m = h.train(C0, max_iter=iters, tol=0.0)
        elif mode == "superstep":                            # host-buffer supersteps, all-reduce on the device buffer
            C, _ = orc.prepare_centroids(metric, C0)
            k, objective, s = C0.shape[0], [], None
            for step in range(1, iters + 1):
                r = h.superstep(C, k=k, step_no=step, tol=0.0, want_sum=True)
                C, k, s = r["C"], r["k"], r["sum"]
                objective.append(r["objective"])
            m = dict(C=C, k=k, weights=s.reshape(C0.shape[0], -1)[:, -1].copy(), objective=np.array(objective))
        else:                                                # "nofinite": one rank holds a NaN row -> every rank must fail
            h.set_centroids(C0)
            try:
                h.step(0.0)
                q.put((rank, "no error"))
            except native.AkmError as e:
                q.put((rank, e.code))
            dist.barrier()
            dist.destroy_process_group()
            return
        labels = h.get_labels()
    q.put((rank, start, labels, m["C"], m["k"], m["weights"], m["objective"]))
    dist.barrier()
    dist.destroy_process_group()




@pytest.mark.parametrize("mode", ["step", "superstep"])
@pytest.mark.parametrize("world,n", [(2, 40001), (4, 40003)])        # odd n: DefaultDistributedInfo remainders
@pytest.mark.parametrize("metric", [native.EUCLIDEAN, native.COSINE])
def test_row_shards_equal_one_gpu_and_the_oracle(metric, world, n, mode):
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip("needs %d GPUs" % world)
    import torch.multiprocessing as mp
    d, k, iters = 32, 24, 6
    X = _blobs(n, d, 12, seed=31)
    C0 = X[np.random.default_rng(1).choice(n, k, replace=False)].copy()
    ref = orc.train(metric, X, C0, max_iter=iters, tol=0.0, threads=world)   # `world` subtasks, summed in rank order
    with native.Handle(d, metric) as h:
        h.upload_rows(X)
        one = h.train(C0, max_iter=iters, tol=0.0)
        one_labels = h.get_labels()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, world, port, X, C0, metric, iters, q, mode)) for r in range(world)]
    for p in procs:
        p.start()
    got = sorted([q.get(timeout=300) for _ in range(world)], key=lambda t: t[0])
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
    for g in got[1:]:
        assert (got[0][3] == g[3]).all()


def test_a_row_without_a_finite_distance_fails_every_rank():
    # the reference fails the whole job (ArrayIndexOutOfBounds in one subtask); a rank that returned alone would
    # leave the others waiting in the next all-reduce
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp
    n, d, k = 20001, 16, 8
    X = _blobs(n, d, 8, seed=7)
    X[n - 3, 2] = np.nan                                      # lives in rank 1's shard only
    C0 = X[:k].copy()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + ((os.getpid() + 1) % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, X, C0, native.EUCLIDEAN, 1, q, "nofinite")) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=300) for _ in range(2))
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert got == {0: -5, 1: -5}
