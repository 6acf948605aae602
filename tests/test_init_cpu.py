"""Host logic of the initial-centroid path (alink_b200/init.py) without a GPU."""
import os
import sys

import numpy as np
import pytest

import oracle as orc
from alink_b200 import init as init_mod

sys.path.insert(0, os.path.dirname(__file__))
from init_double import InitDouble  # noqa: E402


def blobs(n, d, k, seed, spread=10.0):
    rng = np.random.default_rng(seed)
    centers = rng.uniform(-spread, spread, size=(k, d))
    return (centers[rng.integers(0, k, size=n)] + rng.normal(size=(n, d))).astype(np.float32).astype(np.float64)


def test_java_random_stream_of_seed_zero():
    # java.util.Random(0): the values every JDK prints (LocalKmeansFunc.java:39 uses exactly this generator)
    r = init_mod.JavaRandom(0)
    assert [r.next_double() for _ in range(3)] == [0.730967787376657, 0.24053641567148587, 0.6374174253501083]
    r = init_mod.JavaRandom(0)
    assert [r._next(32) for _ in range(3)] == [-1155484576, -723955400, 1033096058]
    r = init_mod.JavaRandom(0)
    assert [r.next_int(10) for _ in range(8)] == [0, 8, 9, 7, 5, 3, 1, 1]
    r = init_mod.JavaRandom(0)
    assert r.next_int(16) == (16 * ((-1155484576 & 0xFFFFFFFF) >> 1)) >> 31       # power-of-two branch


def test_pick_weight_matches_the_reference_loop():
    # LocalKmeansFunc.pickWeight (:107-121) restated as the literal loop
    def literal(w, c, r01):
        s = [0.0]
        for a, b in zip(w, c):
            s.append(s[-1] + a * b)
        r = r01 * s[-1]
        res = len(c)
        for i, v in enumerate(s):
            if v >= r:
                res = i
                break
        return 0 if res == 0 else res - 1

    rng = np.random.default_rng(1)
    for _ in range(50):
        m = int(rng.integers(1, 40))
        w = rng.integers(1, 9, size=m).astype(np.float64)
        c = rng.random(m)
        a, b = init_mod.JavaRandom(7), init_mod.JavaRandom(7)
        assert init_mod.pick_weight(w, c, a) == literal(w, c, b.next_double())


def test_local_kmeans_reproduces_the_reference_junit_values():
    # TEST/operator/common/clustering/kmeans/LocalKmeansFuncTest.java:19-34 (dense) and :36-54 (sparse): samples
    # i * ones, weights i, k = 2 -> centroid norms.  Pins the weighted k-means++ picks (new Random(0)), pickWeight's
    # off-by-one and the never-cleared sum matrix all at once.
    w = np.arange(10, dtype=np.float64)
    dense = np.array([[i, i] for i in range(10)], dtype=np.float64)
    C = init_mod.local_kmeans(2, w, dense, orc.EUCLIDEAN)
    assert abs(np.linalg.norm(C[0]) - 10.842) < 0.001 and abs(np.linalg.norm(C[1]) - 5.185) < 0.001
    sparse = np.zeros((10, 20))
    for i in range(10):
        sparse[i, i] = sparse[i, i + 1] = i
    C = init_mod.local_kmeans(2, w, sparse, orc.EUCLIDEAN)
    assert abs(np.linalg.norm(C[0]) - 8.615) < 0.001 and abs(np.linalg.norm(C[1]) - 4.128) < 0.001


@pytest.mark.parametrize("metric", [orc.EUCLIDEAN, orc.COSINE])
def test_local_kmeans_finds_the_blobs(metric):
    k, d = 6, 8
    X = blobs(300, d, k, seed=3, spread=20.0)
    Xp, _ = orc.prepare_rows(metric, X)
    w = np.ones(300)
    C = init_mod.local_kmeans(k, w, Xp, metric)
    assert C.shape == (k, d) and np.isfinite(C).all()
    Cp, cl = orc.prepare_centroids(metric, C)
    _, labels, _ = orc.assign_accumulate(metric, Xp, np.einsum("ij,ij->i", Xp, Xp), Cp, cl)
    assert len(set(labels.tolist())) == k                       # every centroid owns rows


@pytest.mark.parametrize("metric", [orc.EUCLIDEAN, orc.COSINE])
def test_kmeans_parallel_on_the_cpu_double(metric):
    k, d, n = 12, 10, 4000
    X = blobs(n, d, k, seed=5, spread=15.0)
    info = {}
    C0 = init_mod.kmeans_parallel_init(InitDouble(X, metric), k, 2, seed=11, metric=metric, info=info)
    assert C0.shape == (k, d)
    assert k < info["candidates"] < 8 * k + 2                  # about 2k picks per round
    # seeding quality: Lloyd from these centroids reaches (nearly) the blob optimum
    ref = orc.train(metric, X, C0, max_iter=10, tol=1e-4)
    rnd = orc.train(metric, X, X[:k].copy(), max_iter=10, tol=1e-4)
    assert ref["objective"][-1] <= rnd["objective"][-1] * 1.05 or ref["k"] == k


def test_initsteps_one_means_a_single_candidate():
    X = blobs(500, 4, 3, seed=9)
    info = {}
    C0 = init_mod.kmeans_parallel_init(InitDouble(X, orc.EUCLIDEAN), 3, 1, seed=0, metric=orc.EUCLIDEAN, info=info)
    assert info["candidates"] == 1 and C0.shape == (1, 4)       # samples.size() <= k: the candidates are the model


def _worker(rank, world, port, X, k, seed, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from alink_b200.distributed import TorchDistributed
    td = TorchDistributed()
    start, cnt = td.shard(X.shape[0])
    C0 = init_mod.kmeans_parallel_init(InitDouble(X[start:start + cnt], orc.EUCLIDEAN), k, 3, seed, orc.EUCLIDEAN, td)
    R0 = init_mod.random_init(InitDouble(X[start:start + cnt], orc.EUCLIDEAN), k, seed, td)
    if rank == 0:
        q.put((C0, R0))
    dist.destroy_process_group()


def test_sharding_does_not_change_the_initial_centroids():
    # picks are a function of the global row index, the candidate order is rank order = global row order
    import torch.multiprocessing as mp
    k, seed = 5, 21
    X = blobs(1500, 6, k, seed=8)
    solo = init_mod.kmeans_parallel_init(InitDouble(X, orc.EUCLIDEAN), k, 3, seed, orc.EUCLIDEAN)
    solo_r = init_mod.random_init(InitDouble(X, orc.EUCLIDEAN), k, seed)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, X, k, seed, q)) for r in range(2)]
    for p in procs:
        p.start()
    C0, R0 = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    np.testing.assert_allclose(C0, solo, rtol=1e-9, atol=1e-12)
    assert (R0 == solo_r).all()
