"""CPU tests of the host-side logic: vector parsing, Java number formatting, model wire format,
shard formula, the iterateICQ driver (against the oracle), and the world_size-2 gloo path."""
import os
import sys

import numpy as np
import pytest

import oracle as orc
from alink_b200 import model as model_io
from alink_b200.javafmt import java_double_to_string
from alink_b200.kmeans import iterate_icq, shard_bounds
from alink_b200.vector import parse_vector, rows_to_matrix


class OraclePartition:
    """Test double for native.Handle: the same host-buffer interface backed by the CPU oracle."""

    def __init__(self, X, metric):
        self.metric = metric
        self.Xp, self.xl = orc.prepare_rows(metric, X)
        self.d = X.shape[1]

    def assign_accumulate(self, C, k=None, want_labels=False):
        C = np.ascontiguousarray(C, dtype=np.float64)
        cl = (C * 0).sum(axis=1)
        if self.metric == orc.EUCLIDEAN:
            cl = np.array([orc.dot(c, c) for c in C])
        if self.Xp.shape[0] == 0:
            return np.zeros(C.shape[0] * (self.d + 1)), 0.0, None
        s, labels, mind = orc.assign_accumulate(self.metric, self.Xp, self.xl, C, cl, k=k)
        obj = float((mind ** 2).sum()) if self.metric == orc.EUCLIDEAN else float(mind.sum())
        return s, obj, labels

    def update_centroids(self, sum_, k0, k, step_no=1, prev=None):
        C, cl, nk = orc.update_centroids(self.metric, sum_, k, k0, self.d)
        shift = None
        if prev is not None:
            pl = np.array([orc.dot(c, c) for c in prev]) if self.metric == orc.EUCLIDEAN else np.zeros(k0)
            if step_no % 2 == 1:
                _, shift = orc.converged(self.metric, C, cl, prev, pl, nk, 0.0)
            else:
                _, shift = orc.converged(self.metric, prev, pl, C, cl, nk, 0.0)
        return C, nk, shift


def test_parse_vector_like_vectorutil():
    assert parse_vector("1 2 3").tolist() == [1.0, 2.0, 3.0]
    assert parse_vector("0.1,0.2, 0.3").tolist() == [0.1, 0.2, 0.3]
    assert parse_vector("$5$1:0.1 3:0.4").tolist() == [0, 0.1, 0, 0.4, 0]
    assert parse_vector("1:2.0 0:1.0", size=3).tolist() == [1.0, 2.0, 0.0]
    assert parse_vector(None) is None
    X, mask = rows_to_matrix(["1 2", None, "3 4"])
    assert mask.tolist() == [False, True, False] and X[2].tolist() == [3.0, 4.0]


def test_sparse_vector_column_to_csr():
    from alink_b200.vector import is_sparse_text, parse_vector, rows_to_csr
    col = ["$5$1:0.1 3:0.4", "0:2,4:-1", "$5$", "2:7 2:8"]      # header optional, comma or space, empty row, repeated index
    assert all(is_sparse_text(v) for v in col) and not is_sparse_text("1 2 3")
    indptr, indices, data, d = rows_to_csr(col)
    assert d == 5 and indptr.tolist() == [0, 2, 4, 4, 5]
    assert indices.tolist() == [1, 3, 0, 4, 2] and data.tolist() == [0.1, 0.4, 2.0, -1.0, 8.0]
    dense = np.zeros((4, 5))
    for r in range(4):
        dense[r, indices[indptr[r]:indptr[r + 1]]] = data[indptr[r]:indptr[r + 1]]
    for r, v in enumerate(col):
        assert (dense[r] == parse_vector(v, 5)).all()           # same values as the dense parse
    with pytest.raises(ValueError):
        rows_to_csr(col, vector_size=3)


def test_java_double_to_string():
    cases = {1.0: "1.0", 0.001: "0.001", 1e-4: "1.0E-4", 1e7: "1.0E7", 9999999.0: "9999999.0", -0.0: "-0.0",
             float("nan"): "NaN", float("inf"): "Infinity", 123456.789: "123456.789",
             1.7976931348623157e308: "1.7976931348623157E308", 0.010869565217391353: "0.010869565217391353"}
    for v, s in cases.items():
        assert java_double_to_string(v) == s, (v, java_double_to_string(v), s)


def test_model_wire_format_roundtrip_and_reference_fixture(golden):
    rows = golden["kmeans_util"]["model_rows"]
    md = model_io.load_model(rows)
    assert md["k"] == 2 and md["vectorSize"] == 3 and md["distanceType"] == "EUCLIDEAN" and md["vectorCol"] == "Y"
    assert md["centroids"].tolist() == [[9.1, 9.1, 9.1], [0.1, 0.1, 0.1]] and md["weights"].tolist() == [3.0, 3.0]
    out = model_io.save_model(md["centroids"], md["weights"], 2, 3, "EUCLIDEAN", "Y")
    assert [list(r) for r in out] == rows                   # byte-identical to KMeansUtilTest.java:28-33
    # legacy centroid row (KMeansUtil.java:282-310)
    legacy = [rows[0], [1048576, '{"clusterId":0,"weight":3.0,"center":"{\\"data\\":[9.1,9.1,9.1]}"}'],
              [2097152, '{"clusterId":1,"weight":3.0,"center":"[0.1,0.1,0.1]"}']]
    assert model_io.load_model(legacy)["centroids"].tolist() == [[9.1, 9.1, 9.1], [0.1, 0.1, 0.1]]
    # the literal fixtures of KMeansOldModelMapper1Test.java:19-26 and KMeansOldModelMapper2Test.java:19-25
    old1 = [[0, '{"vectorCol":"\\"Y\\"","latitudeCol":null,"longitudeCol":null,"distanceType":"\\"EUCLIDEAN\\"","k":"2",'
                '"modelSchema":"\\"model_id bigint,model_info string\\"","isNewFormat":"true","vectorSize":"3"}'],
            [1048576, '{"center":"{\\"data\\":[9.1,9.1,9.1]}","clusterId":0,"weight":3.0}'],
            [2097152, '{"center":"{\\"data\\":[0.1,0.1,0.1]}","clusterId":1,"weight":3.0}']]
    old2 = [[0, '{"vectorCol":"\\"Y\\"","latitudeCol":null,"longitudeCol":null,"distanceType":"\\"EUCLIDEAN\\"","k":"2","vectorSize":"3"}'],
            [1048576, '{"clusterId":0,"weight":3.0,"center":"[9.1, 9.1, 9.1]","vec":null}'],
            [2097152, '{"clusterId":1,"weight":3.0,"center":"[0.1, 0.1, 0.1]","vec":null}']]
    for fixture in (old1, old2):
        m = model_io.load_model(fixture)
        assert m["k"] == 2 and m["vectorSize"] == 3 and m["distanceType"] == "EUCLIDEAN" and m["vectorCol"] == "Y"
        assert m["centroids"].tolist() == [[9.1, 9.1, 9.1], [0.1, 0.1, 0.1]] and m["clusterIds"].tolist() == [0, 1]
    # long vectors are sliced at 32 KiB and reassembled by model_id
    big = np.arange(6000, dtype=np.float64)[None, :] + 0.5
    sl = model_io.save_model(big, [1.0], 1, 6000, "COSINE", "v")
    assert len(sl) > 2 and all(len(s) <= 32 * 1024 for _, s in sl)
    assert (model_io.load_model(sl[::-1])["centroids"] == big).all()


def test_shard_bounds(golden):
    for (task, par, total), (start, cnt) in golden["shard"]["cases"]:
        assert shard_bounds(task, par, total) == (start, cnt)
    for total, par in [(10, 3), (7, 8), (100, 8)]:
        spans = [shard_bounds(t, par, total) for t in range(par)]
        assert sum(c for _, c in spans) == total
        assert all(spans[i][0] + spans[i][1] == spans[i + 1][0] for i in range(par - 1))


@pytest.mark.parametrize("metric", [orc.EUCLIDEAN, orc.COSINE])
def test_iterate_icq_driver_equals_oracle_train(metric):
    rng = np.random.default_rng(3)
    X = rng.normal(size=(400, 7)) + rng.integers(0, 4, size=(400, 1)) * 3.0
    C0 = X[:6].copy()
    ref = orc.train(metric, X, C0, max_iter=15, tol=1e-4)
    got = iterate_icq(OraclePartition(X, metric), C0, 15, 1e-4)
    assert got["iters"] == ref["iters"] and got["k"] == ref["k"]
    assert (got["C"] == ref["C"]).all() and (got["weights"] == ref["weights"]).all()
    np.testing.assert_allclose(got["objective"], ref["objective"], rtol=1e-12)


def _gloo_worker(rank, world, port, tmp):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from alink_b200.distributed import TorchDistributed
    td = TorchDistributed()
    rng = np.random.default_rng(5)
    X = rng.normal(size=(301, 5)) + rng.integers(0, 3, size=(301, 1)) * 4.0
    C0 = X[[0, 100, 200, 300]].copy()
    start, cnt = td.shard(X.shape[0])
    part = OraclePartition(X[start:start + cnt], orc.EUCLIDEAN)
    got = iterate_icq(part, C0, 12, 1e-4, allreduce=td.allreduce)
    picked = td.sample_rows(X[start:start + cnt], 4, np.random.default_rng(0))
    np.save(os.path.join(tmp, "C%d.npy" % rank), got["C"])
    np.save(os.path.join(tmp, "w%d.npy" % rank), got["weights"])
    np.save(os.path.join(tmp, "p%d.npy" % rank), picked)
    np.save(os.path.join(tmp, "it%d.npy" % rank), np.array([got["iters"], got["k"], start, cnt]))
    dist.destroy_process_group()


def test_world_size_2_gloo_sharded_training_matches_single_process(tmp_path):
    import torch.multiprocessing as mp
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_gloo_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    rng = np.random.default_rng(5)
    X = rng.normal(size=(301, 5)) + rng.integers(0, 3, size=(301, 1)) * 4.0
    C0 = X[[0, 100, 200, 300]].copy()
    ref = orc.train(orc.EUCLIDEAN, X, C0, max_iter=12, tol=1e-4, threads=2)   # same shards, rank-order sum
    C = [np.load(tmp_path / ("C%d.npy" % r)) for r in range(2)]
    w = [np.load(tmp_path / ("w%d.npy" % r)) for r in range(2)]
    it = [np.load(tmp_path / ("it%d.npy" % r)) for r in range(2)]
    assert (C[0] == C[1]).all() and (w[0] == w[1]).all()            # replicas agree bit for bit
    assert it[0][0] == ref["iters"] and it[0][1] == ref["k"]
    assert it[0][2:].tolist() == [0, 151] and it[1][2:].tolist() == [151, 150]   # DefaultDistributedInfo
    np.testing.assert_allclose(C[0], ref["C"], rtol=1e-12, atol=1e-14)
    assert (w[0] == ref["weights"]).all()
    p = [np.load(tmp_path / ("p%d.npy" % r)) for r in range(2)]
    assert (p[0] == p[1]).all() and p[0].shape == (4, 5)


def test_recording_window_encoding_properties():
    """The search kernel's per-row window slot carries W' = W * 2^e with e in the five low mantissa bits
    (kernels_tensor.cu: plain_window / encode_window / decode_window).  A numpy model of the same bit operations:
    the decoded certifying window is never below W (and at most 63 ulps above), the recording window never below it
    and within a factor of two of its target."""
    rng = np.random.default_rng(3)
    W = np.exp(rng.uniform(np.log(1e-6), np.log(1e6), size=20000)).astype(np.float32)
    target = (W * np.exp(rng.uniform(0.0, np.log(3e5), size=W.size))).astype(np.float32)

    def bits(x):
        return x.view(np.uint32)

    def plain(w):
        return ((bits(w) + np.uint32(31)) & np.uint32(0xFFFFFFE0)).view(np.float32)

    def encode(w, t):
        _, e = np.frexp((t / w).astype(np.float32))
        e = np.clip(e - 1, 0, 24).astype(np.int32)
        scaled = np.ldexp(w, e).astype(np.float32)
        enc = (((bits(scaled) + np.uint32(31)) & np.uint32(0xFFFFFFE0)) | e.astype(np.uint32)).view(np.float32)
        return np.where(t > w, enc, plain(w)), e

    def decode(wenc):
        e = (bits(wenc) & np.uint32(31)).astype(np.int32)
        return (wenc * np.ldexp(np.float32(1.0), -e)).astype(np.float32), e

    p = plain(W)
    dp, e0 = decode(p)
    assert (e0 == 0).all() and (dp == p).all() and (p >= W).all() and (p <= W * np.float32(1 + 2.0 ** -17)).all()
    enc, e = encode(W, target)
    dec, e_back = decode(enc)
    assert (e_back == e).all()
    assert (dec >= W).all() and (dec <= W * np.float32(1 + 2.0 ** -17)).all()        # certifies with at least W
    assert (enc >= dec).all()                                                          # records at least what it certifies
    assert (enc <= target * np.float32(1 + 2.0 ** -17)).all()
    assert ((enc > target * np.float32(0.5)) | (e == 24)).all()                        # within a factor two of the target
