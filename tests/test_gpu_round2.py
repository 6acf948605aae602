"""Round-2 additions through the C ABI: the fused host-buffer superstep, incremental accumulation, uploads of rows
the caller has already prepared, the asynchronous upload pipeline, and the regressions the code review asked for
(pinned-arena lifetime, compaction maps of the two code paths, worklist sizing)."""
import numpy as np
import pytest

import oracle as orc
from alink_b200 import native

pytestmark = pytest.mark.gpu


def blobs(n, d, k, seed, fp32=True, spread=10.0):
    rng = np.random.default_rng(seed)
    centers = rng.uniform(-spread, spread, size=(k, d))
    X = centers[rng.integers(0, k, size=n)] + rng.normal(size=(n, d))
    if fp32:
        X = X.astype(np.float32).astype(np.float64)
    return X


# ------------------------------------------------------------------ akm_superstep
@pytest.mark.parametrize("engine", [native.ENGINE_EXACT, native.ENGINE_AUTO])
@pytest.mark.parametrize("metric", [native.EUCLIDEAN, native.COSINE])
def test_superstep_equals_the_three_fine_grained_calls_and_the_oracle(engine, metric):
    n, d, k0 = 6000, 40, 33
    X = blobs(n, d, 9, seed=21)                               # fewer blobs than centroids: clusters get dropped
    C0 = X[np.random.default_rng(2).choice(n, k0, replace=False)].copy()
    ref = orc.train(metric, X, C0, max_iter=5, tol=0.0)
    Cp, _ = orc.prepare_centroids(metric, C0)
    with native.Handle(d, metric) as ha, native.Handle(d, metric) as hb:
        for h in (ha, hb):
            h.set_engine(engine)
            h.upload_rows(X)
        Ca, ka = Cp.copy(), k0
        Cb, kb = Cp.copy(), k0
        for step in range(1, 6):
            r = ha.superstep(Ca, k=ka, step_no=step, tol=1e-4, want_sum=True, want_shift=True)
            s, obj, _ = hb.assign_accumulate(Cb, k=kb)
            Cn, kn, shift = hb.update_centroids(s, k0, kb, step_no=step, prev=Cb)
            assert r["k"] == kn
            # two handles: the order of the fp64 atomics differs, so sums of non-fp32 values agree to the last bits only
            np.testing.assert_allclose(r["C"], Cn, rtol=1e-13, atol=1e-15)
            np.testing.assert_allclose(r["shift"], shift, rtol=1e-9, atol=1e-12)
            assert (r["sum"].reshape(k0, d + 1)[:, d] == s.reshape(k0, d + 1)[:, d]).all()
            np.testing.assert_allclose(r["sum"], s, rtol=1e-12, atol=1e-9)
            assert abs(r["objective"] - obj) <= 1e-9 * abs(obj)
            assert abs(r["objective"] - ref["objective"][step - 1]) <= 1e-9 * abs(obj)
            assert r["converged"] == bool((r["shift"] < 1e-4).all())
            Ca, ka, Cb, kb = r["C"], r["k"], Cn, kn
        labels = ha.get_labels()
    assert ka == ref["k"]
    assert (labels == ref["labels"]).all()
    np.testing.assert_allclose(Ca[:ka], ref["C"][:ka], rtol=1e-10, atol=1e-12)


# ------------------------------------------------------------------ incremental accumulation
@pytest.mark.parametrize("engine", [native.ENGINE_EXACT, native.ENGINE_AUTO])
@pytest.mark.parametrize("metric", [native.EUCLIDEAN, native.COSINE])
@pytest.mark.parametrize("fp32", [True, False])
def test_delta_accumulation_equals_full_accumulation_and_the_oracle(engine, metric, fp32):
    n, d, k0, iters = 30000, 24, 40, 12
    X = blobs(n, d, 25, seed=5, fp32=fp32, spread=3.0)        # overlapping blobs: labels keep moving for many passes
    C0 = X[np.random.default_rng(3).choice(n, k0, replace=False)].copy()
    ref = orc.train(metric, X, C0, max_iter=iters, tol=0.0)
    out = {}
    for delta in (1, 0):
        with native.Handle(d, metric) as h:
            h.set_engine(engine)
            h.set_option("delta_accumulate", delta)
            h.set_option("reorder_min_rows", 1024)            # regrouping on: the label record has to move with the rows
            h.upload_rows(X)
            m = h.train(C0, max_iter=iters, tol=0.0)
            out[delta] = (m, h.get_labels())
    for delta in (1, 0):
        m, labels = out[delta]
        assert m["k"] == ref["k"]
        assert (labels == ref["labels"]).all()
        assert (m["weights"] == ref["weights"]).all()         # counts stay exact
        np.testing.assert_allclose(m["C"][: m["k"]], ref["C"][: ref["k"]], rtol=1e-11, atol=1e-12)
        np.testing.assert_allclose(m["objective"], ref["objective"], rtol=1e-9)
    np.testing.assert_allclose(out[1][0]["C"], out[0][0]["C"], rtol=1e-12, atol=1e-13)


def test_delta_accumulation_survives_calls_that_rewrite_the_labels():
    # prediction and cluster evaluation borrow the label buffer: the next superstep has to notice
    n, d, k0 = 20000, 16, 12
    X = blobs(n, d, 12, seed=8, spread=4.0)
    C0 = X[:k0].copy()
    ref = orc.train(native.EUCLIDEAN, X, C0, max_iter=8, tol=0.0)
    with native.Handle(d, native.EUCLIDEAN) as h:
        h.upload_rows(X)
        h.set_centroids(C0)
        for s in range(8):
            h.step(0.0)
            if s == 4:
                h.predict_cached(C0, want_dist=False)
            if s == 5:
                h.eval_clusters(np.zeros(n, dtype=np.int32), 3)
            if s == 6:
                h.predict(X[:5000], C0)
        m = h.get_model()
        labels = h.get_labels()
    assert (labels == ref["labels"]).all()
    assert (m["weights"] == ref["weights"]).all()
    np.testing.assert_allclose(m["C"][: m["k"]], ref["C"][: ref["k"]], rtol=1e-11, atol=1e-12)


# ------------------------------------------------------------------ rows the caller prepared (COSINE)
@pytest.mark.parametrize("engine", [native.ENGINE_EXACT, native.ENGINE_AUTO])
def test_prepared_cosine_rows_are_taken_as_given(engine):
    # what the ComQueue hands GpuKMeansAssignCluster: vectors already scaled by CosineDistance.updateLabel
    n, d, k0 = 5000, 32, 11
    X = blobs(n, d, 11, seed=13)
    X[7] = 0.0                                                # a zero vector stays untouched (norm not > 0)
    C0 = X[100:100 + k0].copy()
    ref = orc.train(orc.COSINE, X, C0, max_iter=4, tol=0.0)
    Xp, _ = orc.prepare_rows(orc.COSINE, X)                   # Java's in-place normalised doubles
    Cp, cl = orc.prepare_centroids(orc.COSINE, C0)
    s_ref, l_ref, _ = orc.assign_accumulate(orc.COSINE, Xp, np.zeros(n), Cp, cl)
    with native.Handle(d, native.COSINE) as h:
        h.set_engine(engine)
        h.upload_rows_prepared(Xp[:3000])
        h.upload_rows_prepared(Xp[3000:])
        assert h.num_rows() == (n, 8)
        pick = np.array([0, 7, 2999, 3000, n - 1], dtype=np.int64)
        assert (h.get_rows(pick) == Xp[pick]).all()           # stored bit for bit, not normalised again
        s, _, labels = h.assign_accumulate(Cp, want_labels=True)
        assert (labels == l_ref).all()
        np.testing.assert_allclose(s, s_ref, rtol=1e-12, atol=1e-12)
        m = h.train(C0, max_iter=4, tol=0.0)
        assert (h.get_labels() == ref["labels"]).all()
    np.testing.assert_allclose(m["C"][: m["k"]], ref["C"][: ref["k"]], rtol=1e-11, atol=1e-13)


def test_prepared_euclidean_rows_equal_plain_upload():
    X = blobs(3000, 12, 5, seed=17, fp32=False)
    C = X[:5].copy()
    with native.Handle(12, native.EUCLIDEAN) as h1, native.Handle(12, native.EUCLIDEAN) as h2:
        h1.upload_rows(X)
        h2.upload_rows_prepared(X)
        s1, o1, l1 = h1.assign_accumulate(C, want_labels=True)
        s2, o2, l2 = h2.assign_accumulate(C, want_labels=True)
    assert (l1 == l2).all() and (s1.reshape(5, 13)[:, 12] == s2.reshape(5, 13)[:, 12]).all()
    np.testing.assert_allclose(s1, s2, rtol=1e-13)


# ------------------------------------------------------------------ upload pipeline
@pytest.mark.parametrize("metric", [native.EUCLIDEAN, native.COSINE])
def test_upload_returns_once_staged_and_mixes_fp32_and_fp64_chunks(metric):
    d = 96
    rng = np.random.default_rng(31)
    exact = rng.normal(size=(250_000, d)).astype(np.float32).astype(np.float64)     # 192 MB: three staging chunks
    inexact = rng.normal(size=(30_000, d))
    Xall = np.vstack([exact, inexact, exact[:1000]])
    Xp, _ = orc.prepare_rows(metric, Xall)
    buf = np.empty_like(exact)
    with native.Handle(d, metric) as h:
        np.copyto(buf, exact)
        h.upload_rows(buf)
        buf[:] = -1.0                                         # the caller's buffer is free again after the call returns
        assert h.num_rows() == (250_000, 4)
        h.upload_rows(inexact)                                # a chunk that is not fp32-representable: fp64 from here on
        assert h.num_rows() == (280_000, 8)
        np.copyto(buf[:1000], exact[:1000])
        h.upload_rows(buf[:1000])
        buf[:1000] = 7.0
        h.upload_wait()
        pick = np.array([0, 1, 87_380, 87_381, 174_762, 249_999, 250_000, 279_999, 280_000, 280_999], dtype=np.int64)
        assert (h.get_rows(pick) == Xp[pick]).all()
        C = Xall[:7].copy()
        Cp, cl = orc.prepare_centroids(metric, C)
        sub = np.arange(0, Xall.shape[0], 97)
        _, l_ref, _ = orc.assign_accumulate(metric, Xp[sub], orc.prepare_rows(metric, Xall[sub])[1], Cp, cl)
        _, _, labels = h.assign_accumulate(Cp, want_labels=True)
    assert (labels[sub] == l_ref).all()


@pytest.mark.parametrize("metric", [native.EUCLIDEAN, native.COSINE])
def test_upload_from_page_locked_memory_skips_staging_and_repairs_misfits(metric):
    d = 64
    rng = np.random.default_rng(41)
    n_a, n_b = 300_000, 200_000                               # 154 MB + 102 MB of doubles: several 64 MB chunks each
    Xa = rng.normal(size=(n_a, d)).astype(np.float32).astype(np.float64)
    Xb = rng.normal(size=(n_b, d)).astype(np.float32).astype(np.float64)
    Xb[150_000, 3] = 0.1                                      # one value that is not an fp32 value, in the second chunk
    Xall = np.vstack([Xa, Xb, Xa[:500]])
    Xp, _ = orc.prepare_rows(metric, Xall)
    with native.PinnedArray((n_a, d)) as pa, native.PinnedArray((n_b, d)) as pb, native.Handle(d, metric) as h:
        np.copyto(pa.array, Xa)
        np.copyto(pb.array, Xb)
        h.upload_rows(pa.array)
        assert h.num_rows() == (n_a, 4)                       # all fp32 values: fp32 storage (checked on the device)
        h.upload_rows(pb.array)                               # returns at once; the misfit is found after the fact
        h.upload_rows(pa.array[:500])
        h.upload_wait()
        assert h.num_rows() == (n_a + n_b + 500, 8)
        pick = np.array([0, n_a - 1, n_a, n_a + 131_071, n_a + 131_072, n_a + 150_000, n_a + n_b - 1, n_a + n_b + 499],
                        dtype=np.int64)
        assert (h.get_rows(pick) == Xp[pick]).all()


@pytest.mark.parametrize("metric", [native.EUCLIDEAN, native.COSINE])
@pytest.mark.parametrize("fp32", [True, False])
def test_csr_upload_equals_dense_upload(metric, fp32):
    # SparseVector samples (BLAS.java:217-240, KMeansUtil.java:79-82): expanded on the device, bit-identical to dense
    rng = np.random.default_rng(77)
    n, d, k0 = 70_000, 300, 9                                 # 168 MB dense: three staging chunks
    dense = np.zeros((n, d))
    nnz_per_row = rng.integers(0, 12, size=n)                 # some rows are empty
    indptr = np.zeros(n + 1, dtype=np.int64)
    indices, values = [], []
    for r in range(n):
        cols = np.sort(rng.choice(d, size=nnz_per_row[r], replace=False))
        vals = rng.normal(size=cols.size)
        if fp32:
            vals = vals.astype(np.float32).astype(np.float64)
        dense[r, cols] = vals
        indices.append(cols); values.append(vals)
        indptr[r + 1] = indptr[r] + cols.size
    indices = np.concatenate(indices).astype(np.int32)
    values = np.concatenate(values)
    C = dense[rng.choice(n, k0, replace=False)] + 0.01
    with native.Handle(d, metric) as ha, native.Handle(d, metric) as hb:
        ha.upload_rows_csr(indptr, indices, values)
        hb.upload_rows(dense)
        assert ha.num_rows() == hb.num_rows() == (n, 4 if fp32 else 8)
        pick = np.array([0, 1, 27_961, 27_962, 55_923, n - 1], dtype=np.int64)
        assert (ha.get_rows(pick) == hb.get_rows(pick)).all()
        Cp, _ = orc.prepare_centroids(metric, C)
        sa, oa, la = ha.assign_accumulate(Cp, want_labels=True)
        sb, ob, lb = hb.assign_accumulate(Cp, want_labels=True)
    assert (la == lb).all()
    np.testing.assert_allclose(sa, sb, rtol=1e-13, atol=1e-13)
    with native.Handle(d, metric) as h:
        with pytest.raises(native.AkmError) as e:
            h.upload_rows_csr(np.array([0, 1]), np.array([d], dtype=np.int32), np.array([1.0]))
        assert e.value.code == -1


def test_train_op_takes_a_sparse_vector_column_through_csr():
    from alink_b200.kmeans import KMeansTrainBatchOp, Table
    rows = ["$6$0:%g 3:%g" % (a, b) for a, b in [(0.0, 0.1), (0.1, 0.0), (0.2, 0.2), (9.0, 9.1), (9.1, 9.0), (9.2, 9.2)]]
    dense = ["%g 0 0 %g 0 0" % (a, b) for a, b in [(0.0, 0.1), (0.1, 0.0), (0.2, 0.2), (9.0, 9.1), (9.1, 9.0), (9.2, 9.2)]]
    models = []
    for col in (rows, dense):
        t = Table([(v,) for v in col], ["vec"])
        op = KMeansTrainBatchOp().setVectorCol("vec").setK(2).setInitCentroids(np.array([[0, 0, 0, 0, 0, 0], [9, 0, 0, 9, 0, 0.0]]))
        op.linkFrom(t)
        models.append(op.model_rows)
    assert models[0] == models[1]


def test_nan_rows_go_through_the_fp64_path():
    X = blobs(2000, 8, 3, seed=1)
    X[5, 3] = np.nan
    with native.Handle(8, native.EUCLIDEAN) as h:
        h.upload_rows(X)
        assert h.num_rows() == (2000, 8)
        with pytest.raises(native.AkmError) as e:
            h.assign_accumulate(X[:3].copy())
        assert e.value.code == -5


# ------------------------------------------------------------------ code-review regressions
def test_pinned_arena_survives_growth_of_the_staging_buffers():
    d, k0 = 16, 8
    X = blobs(4000, d, k0, seed=3)
    C = X[:k0].copy()
    with native.Handle(d, native.EUCLIDEAN) as h:
        h.upload_rows(X[:100])                                # tiny staging buffers
        s, _, _ = h.assign_accumulate(C)
        C1, k1, _ = h.update_centroids(s, k0, k0, step_no=1, prev=C)       # allocates the pinned arena
        big = blobs(1_200_000, d, k0, seed=4)                 # 154 MB: the staging buffers have to grow
        idx, dist, _ = h.predict(big, C1[:k1])
        s2, _, _ = h.assign_accumulate(C1, k=k1)              # arena again
        C2, k2, _ = h.update_centroids(s2, k0, k1, step_no=2, prev=C1)
    Xp, xl = orc.prepare_rows(orc.EUCLIDEAN, X[:100])
    Cp, cl = orc.prepare_centroids(orc.EUCLIDEAN, C1)
    s_ref, _, _ = orc.assign_accumulate(orc.EUCLIDEAN, Xp, xl, Cp, cl, k=k1)
    np.testing.assert_allclose(s2, s_ref, rtol=1e-12, atol=1e-9)
    C_ref, _, k_ref = orc.update_centroids(orc.EUCLIDEAN, s_ref, k1, k0, d)
    assert k2 == k_ref and (C2 == C_ref).all()
    ridx, rdist, _ = orc.predict(orc.EUCLIDEAN, big[:2000], C1[:k1])
    assert (idx[:2000] == ridx).all() and (dist[:2000] == rdist).all()


def test_labels_out_of_a_large_partition_generated_on_the_device():
    # labels_out >= 8 MB goes through the pipelined download while the accumulator sits in the pinned arena
    d, k0, n = 8, 6, 2_300_000
    with native.Handle(d, native.EUCLIDEAN) as h:
        h.blobs_append(77, k0, 0, n)
        rows = np.arange(0, n, 1151, dtype=np.int64)
        Xs = h.blobs_rows(77, k0, rows).astype(np.float64)
        C = h.blobs_rows(77, k0, np.arange(k0, dtype=np.int64) * (n // k0)).astype(np.float64)
        s, obj, labels = h.assign_accumulate(C, want_labels=True)
    Xp, xl = orc.prepare_rows(orc.EUCLIDEAN, Xs)
    Cp, cl = orc.prepare_centroids(orc.EUCLIDEAN, C)
    _, l_ref, _ = orc.assign_accumulate(orc.EUCLIDEAN, Xp, xl, Cp, cl)
    assert (labels[rows] == l_ref).all()
    cnt = s.reshape(k0, d + 1)[:, d]
    assert cnt.sum() == n and (cnt == np.bincount(labels, minlength=k0)).all()


def test_host_buffer_update_between_device_steps_does_not_touch_the_loop_state():
    # akm_update_centroids used to overwrite the compaction map the next akm_step derives its seeds from
    n, d, k0 = 40000, 32, 300
    X = blobs(n, d, 40, seed=9)                               # 300 centroids on 40 blobs: clusters die, the map matters
    C0 = X[np.random.default_rng(5).choice(n, k0, replace=False)].copy()
    ref = orc.train(orc.EUCLIDEAN, X, C0, max_iter=6, tol=0.0)
    rng = np.random.default_rng(6)
    with native.Handle(d, native.EUCLIDEAN) as h:
        h.set_option("reorder_min_rows", 1024)
        h.upload_rows(X)
        h.set_centroids(C0)
        for s in range(6):
            h.step(0.0)
            buf = rng.normal(size=(k0, d + 1))
            buf[:, d] = rng.integers(0, 2, size=k0)           # a very different set of empty clusters
            h.update_centroids(buf.ravel(), k0, k0, step_no=1, prev=C0)
        m = h.get_model()
        labels = h.get_labels()
    assert m["k"] == ref["k"] and (labels == ref["labels"]).all()
    np.testing.assert_allclose(m["C"][: m["k"]], ref["C"][: ref["k"]], rtol=1e-11, atol=1e-12)


def test_every_row_on_the_worklist():
    # identical centroids: every row is a true tie, so every row of the partition takes the exact path
    n, d, k0 = 70000, 16, 4
    X = blobs(n, d, 4, seed=12)
    C = np.tile(X[:1], (k0, 1))
    with native.Handle(d, native.EUCLIDEAN) as h:
        h.set_engine(native.ENGINE_TENSOR)
        h.upload_rows(X)
        s, _, labels = h.assign_accumulate(C, want_labels=True)
        st = h.stats()
    assert (labels == 0).all()                                # first minimum
    assert st["rows_refined"] == n and s.reshape(k0, d + 1)[0, d] == n


@pytest.mark.parametrize("metric", [native.EUCLIDEAN, native.COSINE])
def test_bounds_hold_for_whatever_centroids_the_host_passes_next(metric):
    # Host-buffer supersteps keep seeds and second-nearest bounds between calls; they are bounds on how far the
    # centroids the caller passes NEXT lie from the previous ones, so centroids that are not the ones the library
    # returned (perturbed, permuted, partly replaced) must give the exhaustive answer as well.
    n, d, k = 60000, 20, 400
    rng = np.random.default_rng(21)
    centers = rng.uniform(-8, 8, size=(k, d))
    X = (centers[rng.integers(0, k, size=n)] + rng.normal(size=(n, d))).astype(np.float32).astype(np.float64)
    C = X[rng.choice(n, k, replace=False)].copy()
    with native.Handle(d, metric) as ht, native.Handle(d, metric) as he:
        ht.set_engine(native.ENGINE_TENSOR); he.set_engine(native.ENGINE_EXACT)
        ht.set_option("reorder_min_rows", 1)
        ht.upload_rows(X); he.upload_rows(X)
        k_cur = k
        for it in range(12):
            rt = ht.superstep(C, k=k_cur, step_no=it + 1, tol=0.0, want_sum=True)
            re_ = he.superstep(C, k=k_cur, step_no=it + 1, tol=0.0, want_sum=True)
            assert rt["k"] == re_["k"], "pass %d" % it
            k_cur = re_["k"]                                             # (the big move at the end empties a few clusters)
            cnt_t = rt["sum"].reshape(k, d + 1)[:, d]
            cnt_e = re_["sum"].reshape(k, d + 1)[:, d]
            assert (cnt_t == cnt_e).all(), "pass %d: counts differ in %d clusters" % (it, int((cnt_t != cnt_e).sum()))
            np.testing.assert_allclose(rt["sum"], re_["sum"], rtol=1e-11, atol=1e-9)
            np.testing.assert_allclose(rt["objective"], re_["objective"], rtol=1e-9)
            C = re_["C"].copy()
            if it == 4:
                C += rng.normal(scale=0.05, size=C.shape)              # a small nudge: most bounds survive it
            elif it == 6:
                C = C[rng.permutation(k)]                                # every label changes, nothing else does
            elif it == 8:
                C[::7] = X[rng.choice(n, C[::7].shape[0], replace=False)]  # some centroids jump anywhere
            elif it == 9:
                C += rng.normal(scale=1.5, size=C.shape)               # a big move of all of them
