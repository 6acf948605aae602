"""Parity of the CUDA path (through the C ABI) with the CPU oracle and the reference's golden vectors.

Bars (BASELINE.json north_star): assignments bit-exact (documented ties aside -- none are needed
here: both engines reproduce Java's first-minimum rule exactly), centroids and objective within
1e-6 relative (the tests use 1e-9 or tighter; the update step is bit-exact)."""
import numpy as np
import pytest

import oracle as orc
from alink_b200 import native

pytestmark = pytest.mark.gpu

ENGINES = [native.ENGINE_EXACT, native.ENGINE_AUTO]


def blobs(n, d, k, seed, fp32=True, spread=10.0):
    rng = np.random.default_rng(seed)
    centers = rng.uniform(-spread, spread, size=(k, d))
    X = centers[rng.integers(0, k, size=n)] + rng.normal(size=(n, d))
    if fp32:
        X = X.astype(np.float32).astype(np.float64)
    return X


def oracle_pass(metric, X, C, k=None):
    Xp, xl = orc.prepare_rows(metric, X)
    Cp, cl = orc.prepare_centroids(metric, C)
    s, labels, mind = orc.assign_accumulate(metric, Xp, xl, Cp, cl, k=k)
    obj = float((mind ** 2).sum()) if metric == orc.EUCLIDEAN else float(mind.sum())
    return Cp, s, labels, mind, obj


@pytest.mark.parametrize("engine", ENGINES)
@pytest.mark.parametrize("metric", [native.EUCLIDEAN, native.COSINE])
@pytest.mark.parametrize("n,d,k", [(1, 3, 1), (63, 5, 2), (64, 64, 37), (65, 130, 100), (5000, 128, 257),
                                   (3000, 784, 10), (4097, 16, 1000)])
def test_assign_accumulate_matches_oracle(engine, metric, n, d, k):
    X = blobs(n, d, max(k, 2), seed=n + d + k)
    C = X[np.linspace(0, n - 1, k).astype(int)].copy() if n >= k else blobs(k, d, k, seed=1)
    Cp, s_ref, l_ref, _, obj_ref = oracle_pass(metric, X, C)
    with native.Handle(d, metric) as h:
        h.set_engine(engine)
        h.upload_rows(X)
        s, obj, labels = h.assign_accumulate(Cp, want_labels=True)
    assert (labels == l_ref).all()
    L = s.reshape(k, d + 1)
    R = s_ref.reshape(k, d + 1)
    assert (L[:, d] == R[:, d]).all()                       # counts are exact
    np.testing.assert_allclose(L, R, rtol=1e-12, atol=1e-9)
    assert abs(obj - obj_ref) <= 1e-9 * max(1.0, abs(obj_ref))


@pytest.mark.parametrize("engine", ENGINES)
@pytest.mark.parametrize("metric", [native.EUCLIDEAN, native.COSINE])
def test_storage_fp32_exact_vs_fp64_and_lossless_migration(engine, metric):
    # COSINE: an fp32 partition keeps the RAW rows and |x|, and recomputes x * (1.0 / |x|) where it reads them;
    # the migration to fp64 stores the normalised rows.  Either way the values are the ones Java holds.
    d, k = 24, 9
    Xa = blobs(700, d, k, seed=3, fp32=True)
    Xb = blobs(900, d, k, seed=4, fp32=False)                # not fp32-representable
    C = Xa[:k].copy()
    X = np.vstack([Xa, Xb])
    Cp, s_ref, l_ref, _, obj_ref = oracle_pass(metric, X, C)
    with native.Handle(d, metric) as h:
        h.set_engine(engine)
        h.upload_rows(Xa)
        assert h.num_rows() == (700, 4)
        Xp, _ = orc.prepare_rows(metric, Xa)
        pick = np.array([0, 13, 699], dtype=np.int64)
        assert (h.get_rows(pick) == Xp[pick]).all()          # prepared vectors bit for bit
        h.upload_rows(Xb)                                    # switches the partition to fp64, losslessly
        assert h.num_rows() == (1600, 8)
        assert (h.get_rows(pick) == Xp[pick]).all()
        s, obj, labels = h.assign_accumulate(Cp, want_labels=True)
    assert (labels == l_ref).all()
    np.testing.assert_allclose(s, s_ref, rtol=1e-12, atol=1e-9)
    assert abs(obj - obj_ref) <= 1e-9 * abs(obj_ref)


@pytest.mark.parametrize("metric", [native.EUCLIDEAN, native.COSINE])
def test_float32_upload_equals_double_upload(metric):
    d, k = 20, 6
    X = blobs(1000, d, k, seed=5)
    C, _ = orc.prepare_centroids(metric, X[:k].copy())
    with native.Handle(d, metric) as h1, native.Handle(d, metric) as h2:
        h1.upload_rows(X)
        h2.upload_rows(X.astype(np.float32))
        s1, o1, l1 = h1.assign_accumulate(C, want_labels=True)
        s2, o2, l2 = h2.assign_accumulate(C, want_labels=True)
    assert (l1 == l2).all()
    if metric == native.EUCLIDEAN:
        assert (s1 == s2).all()                              # sums of fp32-exact values are exact in fp64: order-free
    else:
        np.testing.assert_allclose(s1, s2, rtol=1e-13, atol=1e-13)   # normalised values: the atomics' order shows in the last bits


def test_update_centroids_golden_and_bit_exact(golden):
    g = golden["update_centroids"]
    with native.Handle(g["d"], native.EUCLIDEAN) as h:
        C, nk, _ = h.update_centroids(g["buffer"], g["k"], g["k"])
    assert nk == g["new_k"]
    np.testing.assert_allclose(C.ravel(), g["centroids"], atol=g["tol"])
    # random buffers with empty clusters: bit-exact against the oracle, both metrics, incl. shift
    rng = np.random.default_rng(11)
    for metric in (native.EUCLIDEAN, native.COSINE):
        k0, d = 57, 19
        buf = rng.normal(size=(k0, d + 1))
        buf[:, d] = rng.integers(0, 5, size=k0)
        prev = rng.normal(size=(k0, d))
        prev_p, prev_l = orc.prepare_centroids(metric, prev)
        C_ref, cl_ref, nk_ref = orc.update_centroids(metric, buf.ravel(), k0, k0, d)
        for step_no in (1, 2):
            with native.Handle(d, metric) as h:
                C, nk, shift = h.update_centroids(buf.ravel(), k0, k0, step_no=step_no, prev=prev_p)
            assert nk == nk_ref
            assert (C == C_ref).all()
            if step_no % 2 == 1:   # new buffer is CENTROID1
                _, diag = orc.converged(metric, C_ref, cl_ref, prev_p, prev_l, nk_ref, 1e-4)
            else:
                _, diag = orc.converged(metric, prev_p, prev_l, C_ref, cl_ref, nk_ref, 1e-4)
            assert (shift == diag).all()


@pytest.mark.parametrize("engine", ENGINES)
@pytest.mark.parametrize("metric", [native.EUCLIDEAN, native.COSINE])
def test_training_loop_matches_oracle(engine, metric):
    n, d, k = 6000, 48, 40
    X = blobs(n, d, 25, seed=21)                            # fewer blobs than centroids: splits + near ties
    C0 = X[np.random.default_rng(2).choice(n, k, replace=False)].copy()
    ref = orc.train(metric, X, C0, max_iter=12, tol=1e-4)
    with native.Handle(d, metric) as h:
        h.set_engine(engine)
        h.upload_rows(X)
        m = h.train(C0, max_iter=12, tol=1e-4)
        labels = h.get_labels()
    assert m["iters"] == ref["iters"] and m["k"] == ref["k"]
    assert (labels == ref["labels"]).all()
    np.testing.assert_allclose(m["C"][: m["k"]], ref["C"][: ref["k"]], rtol=1e-9, atol=1e-12)
    assert (m["weights"] == ref["weights"]).all()
    np.testing.assert_allclose(m["objective"], ref["objective"], rtol=1e-9)


def test_empty_clusters_are_dropped_like_the_reference():
    X = np.array([[0.0], [0.1], [10.0], [10.1]])
    C0 = np.array([[0.0], [100.0], [10.0]])
    ref = orc.train(orc.EUCLIDEAN, X, C0, max_iter=5, tol=0.0)
    with native.Handle(1, native.EUCLIDEAN) as h:
        h.upload_rows(X)
        m = h.train(C0, max_iter=5, tol=0.0)
        labels = h.get_labels()
    assert m["k"] == ref["k"] == 2
    assert (m["C"] == ref["C"]).all()
    assert labels.tolist() == [0, 0, 1, 1]
    assert (m["weights"] == ref["weights"]).all()


@pytest.mark.parametrize("engine", ENGINES)
def test_exact_ties_first_index_wins(engine):
    d = 32
    X = blobs(2000, d, 4, seed=8)
    base = X[:5].copy()
    C = np.vstack([base, base, base[::-1]])                 # every centroid appears three times
    _, s_ref, l_ref, _, _ = oracle_pass(orc.EUCLIDEAN, X, C)
    assert l_ref.max() < 5                                  # duplicates never win
    with native.Handle(d, native.EUCLIDEAN) as h:
        h.set_engine(engine)
        h.upload_rows(X)
        s, _, labels = h.assign_accumulate(C, want_labels=True)
    assert (labels == l_ref).all()


@pytest.mark.parametrize("engine", ENGINES)
@pytest.mark.parametrize("metric", [native.EUCLIDEAN, native.COSINE])
def test_near_ties_are_resolved_exactly(engine, metric):
    # centroids that differ by a few ulps / 1e-7 relative: the tensor-core search must hand these rows
    # to the exact evaluation, and the exact evaluation must reproduce Java's rounding.
    rng = np.random.default_rng(5)
    d, kb = 128, 16
    base = rng.uniform(-10, 10, size=(kb, d))
    C = np.vstack([base, base * (1 + 1e-7), base + 1e-9, base * (1 - 3e-8), np.nextafter(base, np.inf)])
    X = (base[rng.integers(0, kb, size=4000)] + 0.5 * rng.normal(size=(4000, d))).astype(np.float32).astype(np.float64)
    Cp, s_ref, l_ref, _, _ = oracle_pass(metric, X, C)
    with native.Handle(d, metric) as h:
        h.set_engine(engine)
        h.upload_rows(X)
        s, _, labels = h.assign_accumulate(Cp, want_labels=True)
        st = h.stats()
    assert (labels == l_ref).all(), "mismatches: %d" % int((labels != l_ref).sum())
    if st["engine_used"] == native.ENGINE_TENSOR:
        assert st["rows_refined"] > 0


@pytest.mark.parametrize("metric", [native.EUCLIDEAN, native.COSINE])
@pytest.mark.parametrize("n,d,k", [(3000, 256, 500), (2500, 784, 10), (1500, 200, 300), (700, 1000, 70), (3000, 136, 1100)])
def test_wide_rows_use_the_k_streaming_tensor_search(metric, n, d, k):
    # d > 128 (BASELINE configs 2 and 5): both operands stream through shared memory in 64-half blocks
    X = blobs(n, d, max(k // 3, 2), seed=n + d)             # several centroids per blob: real near-ties
    C = X[np.random.default_rng(d).choice(n, k, replace=False)].copy()
    Cp, s_ref, l_ref, _, obj_ref = oracle_pass(metric, X, C)
    with native.Handle(d, metric) as h:
        h.set_engine(native.ENGINE_TENSOR)
        h.upload_rows(X)
        s, obj, labels = h.assign_accumulate(Cp, want_labels=True)
        st = h.stats()
    assert st["engine_used"] == native.ENGINE_TENSOR and st["rows_certified"] > 0
    assert (labels == l_ref).all(), "mismatches: %d" % int((labels != l_ref).sum())
    np.testing.assert_allclose(s, s_ref, rtol=1e-12, atol=1e-9)
    assert abs(obj - obj_ref) <= 1e-9 * max(1.0, abs(obj_ref))


@pytest.mark.parametrize("metric", [native.EUCLIDEAN, native.COSINE])
def test_wide_rows_training_loop_matches_oracle(metric):
    n, d, k = 5000, 256, 60
    X = blobs(n, d, 30, seed=77)
    C0 = X[np.random.default_rng(3).choice(n, k, replace=False)].copy()
    ref = orc.train(metric, X, C0, max_iter=8, tol=1e-4)
    with native.Handle(d, metric) as h:
        h.set_engine(native.ENGINE_TENSOR)
        h.upload_rows(X)
        m = h.train(C0, max_iter=8, tol=1e-4)
        labels = h.get_labels()
        st = h.stats()
    assert st["engine_used"] == native.ENGINE_TENSOR
    assert m["iters"] == ref["iters"] and m["k"] == ref["k"]
    assert (labels == ref["labels"]).all()
    np.testing.assert_allclose(m["C"][: m["k"]], ref["C"][: ref["k"]], rtol=1e-9, atol=1e-12)
    assert (m["weights"] == ref["weights"]).all()


@pytest.mark.parametrize("metric", [native.EUCLIDEAN, native.COSINE])
def test_tile_pruning_changes_nothing_but_the_work(metric):
    # k > 256 so that there are several centroid tiles; a cluster-ordered partition so that groups agree
    n, d, k = 60000, 32, 700
    X = blobs(n, d, k, seed=41, spread=20.0)
    C0 = X[np.random.default_rng(4).choice(n, k, replace=False)].copy()
    ref = orc.train(metric, X, C0, max_iter=7, tol=0.0, threads=4)
    out = {}
    for prune in (1, 0):
        with native.Handle(d, metric) as h:
            h.set_engine(native.ENGINE_TENSOR)
            h.set_option("reorder_min_rows", 1)
            h.set_option("tile_pruning", prune)
            h.upload_rows(X)
            m = h.train(C0, max_iter=7, tol=0.0)
            out[prune] = (h.get_labels(), m, h.stats())
    for prune in (1, 0):
        labels, m, st = out[prune]
        assert (labels == ref["labels"]).all() and m["k"] == ref["k"]
        np.testing.assert_allclose(m["C"][: m["k"]], ref["C"][: ref["k"]], rtol=1e-9, atol=1e-12)
    assert out[0][2]["tiles_computed"] == out[0][2]["tiles_total"] > 0
    assert 0 <= out[1][2]["tiles_computed"] < out[1][2]["tiles_total"]     # something was skipped
    assert (out[1][1]["C"] == out[0][1]["C"]).all() or metric == native.COSINE


@pytest.mark.parametrize("metric,k,nb,centred", [(native.EUCLIDEAN, 700, 700, True), (native.EUCLIDEAN, 700, 700, False),
                                                 (native.EUCLIDEAN, 200, 120, False), (native.COSINE, 300, 300, True),
                                                 (native.EUCLIDEAN, 520, 260, False),
                                                 # k <= 128: the two-team kernel
                                                 (native.EUCLIDEAN, 100, 100, True), (native.COSINE, 90, 60, False),
                                                 (native.EUCLIDEAN, 128, 300, False)])
def test_whole_group_skipping_pass_by_pass(metric, k, nb, centred):
    # Row groups whose rows provably keep their centroid are not searched at all (no tile of theirs is evaluated).
    # Every pass of a long run must still give the labels and sums of the exact engine; k < nb leaves clusters that
    # span several blobs (far rows: no skipping there), k > nb clusters that split a blob (close centroids).
    n, d, iters = 50000, 24, 14
    X = blobs(n, d, nb, seed=5, spread=15.0)
    C0 = X[np.random.default_rng(6).choice(n, k, replace=False)].copy()
    if centred:                                  # one start per blob (the first draw of blobs() is the centres)
        C0 = np.random.default_rng(5).uniform(-15.0, 15.0, size=(nb, d)) + 0.3
    hs = []
    for engine in (native.ENGINE_TENSOR, native.ENGINE_EXACT):
        h = native.Handle(d, metric)
        h.set_engine(engine)
        h.set_option("reorder_min_rows", 1)
        h.upload_rows(X)
        h.set_centroids(C0)
        hs.append(h)
    skipped_whole = 0
    try:
        for it in range(iters):
            ra, rb = hs[0].step(0.0), hs[1].step(0.0)
            la, lb = hs[0].get_labels(), hs[1].get_labels()
            assert (la == lb).all(), "pass %d: %d labels differ" % (it, int((la != lb).sum()))
            assert ra[1] == rb[1]                                          # k after the update
            np.testing.assert_allclose(ra[2], rb[2], rtol=1e-9)            # objective
            st = hs[0].stats()
            ntiles = (ra[1] + 255) // 256
            if st["tiles_total"] and st["tiles_computed"] * ntiles < st["tiles_total"]:
                skipped_whole += 1                                         # fewer tiles than one per row group
        ma, mb = hs[0].get_model(), hs[1].get_model()
        assert (ma["weights"] == mb["weights"]).all()
        np.testing.assert_allclose(ma["C"][: ma["k"]], mb["C"][: mb["k"]], rtol=1e-10, atol=1e-12)
    finally:
        for h in hs:
            h.close()
    if centred and metric == native.EUCLIDEAN:
        assert skipped_whole > 0                                           # well-separated blobs: most groups drop out


def test_only_the_first_k_centroids_are_candidates():
    d = 8
    X = blobs(500, d, 3, seed=9)
    C = np.vstack([X[:4], np.zeros((3, d))])                # stale zero rows beyond k, as after a drop
    Cp, s_ref, l_ref, _, _ = oracle_pass(orc.EUCLIDEAN, X, C, k=4)
    with native.Handle(d, native.EUCLIDEAN) as h:
        h.upload_rows(X)
        s, _, labels = h.assign_accumulate(Cp, k=4, want_labels=True)
    assert (labels == l_ref).all() and labels.max() < 4
    np.testing.assert_allclose(s, s_ref, rtol=1e-12, atol=1e-9)


def test_nan_rows_fail_like_java():
    d = 4
    X = blobs(100, d, 2, seed=1)
    X[17, 2] = np.nan
    with native.Handle(d, native.EUCLIDEAN) as h:
        h.upload_rows(X)
        with pytest.raises(native.AkmError) as e:
            h.assign_accumulate(X[:2].copy())
        assert e.value.code == -5                            # Java: ArrayIndexOutOfBounds on index -1


def test_empty_partition_and_argument_errors():
    with native.Handle(3, native.EUCLIDEAN) as h:
        s, obj, _ = h.assign_accumulate(np.ones((2, 3)))
        assert (s == 0).all() and obj == 0.0
        with pytest.raises(native.AkmError):
            h.step()                                         # no centroids yet -> STATE
        with pytest.raises(ValueError):
            h.upload_rows(np.ones((2, 4)))


def test_predict_golden_detail_string_bit_exact(golden):
    from alink_b200.javafmt import java_vector_to_string
    g = golden["model_mapper"]
    C = np.array(golden["kmeans_util"]["centroids"])
    with native.Handle(3, native.EUCLIDEAN) as h:
        idx, dist, probs = h.predict(np.array([g["row"]], dtype=np.float64), C, want_dist=True, want_probs=True)
    assert idx[0] == g["pred"]
    assert abs(dist[0] - g["distance"]) <= g["distance_tol"]
    assert java_vector_to_string(probs[0]) == g["detail_exact"]


@pytest.mark.parametrize("engine", ENGINES)
@pytest.mark.parametrize("metric", [native.EUCLIDEAN, native.COSINE])
def test_predict_matches_oracle_bitwise(engine, metric):
    n, d, k = 3000, 40, 23
    X = blobs(n, d, k, seed=31, fp32=False)
    C = blobs(k, d, k, seed=32, fp32=False)
    idx_r, dist_r, probs_r = orc.predict(metric, X, C, want_dist=True, want_probs=True)
    with native.Handle(d, metric) as h:
        h.set_engine(engine)
        idx, dist, probs = h.predict(X, C, want_dist=True, want_probs=True)
        idx2, dist2, _ = h.predict(X, C, want_dist=True, want_probs=False)
    assert (idx == idx_r).all() and (idx2 == idx_r).all()
    assert (dist == dist_r).all() and (dist2 == dist_r).all()   # Math.sqrt / fp64 order reproduced bit for bit
    assert (probs == probs_r).all()


def test_six_point_pipeline_golden(golden):
    g = golden["pipeline_6_points"]
    X = np.array(g["rows"], dtype=np.float64)
    with native.Handle(3, native.EUCLIDEAN) as h:
        h.upload_rows(X)
        m = h.train(X[[0, 3]], max_iter=20, tol=1e-4)
        idx, dist = h.predict_cached(m["C"][: m["k"]])
    np.testing.assert_allclose(dist, g["pred_distances"], atol=g["tol"])
    want = np.array(g["pred_up_to_permutation"])
    assert (idx == want).all() or (idx == 1 - want).all()
    assert abs(float((dist ** 2).sum()) - g["ssw"]) <= 0.01


def test_iris_config1():
    # BASELINE config 1: iris k=3 (sklearn's bundled copy; the reference fetches it over HTTP)
    from sklearn.datasets import load_iris
    X = load_iris().data.astype(np.float64)
    C0 = X[[0, 50, 100]].copy()
    ref = orc.train(orc.EUCLIDEAN, X, C0, max_iter=20, tol=1e-4)
    with native.Handle(4, native.EUCLIDEAN) as h:
        h.upload_rows(X)
        assert h.num_rows()[1] == 8                          # 5.1 etc. are not fp32-representable
        m = h.train(C0, max_iter=20, tol=1e-4)
        labels = h.get_labels()
    assert m["iters"] == ref["iters"] and (labels == ref["labels"]).all()
    np.testing.assert_allclose(m["C"][:3], ref["C"][:3], rtol=1e-12)


@pytest.mark.parametrize("metric", [native.EUCLIDEAN, native.COSINE])
def test_cluster_ordered_layout_is_invisible_to_the_caller(metric):
    # the library regroups rows by cluster internally; labels, predictions and appended rows must still come
    # back in the caller's order, and the training result must not change
    n, d, k = 9000, 32, 24
    X = blobs(n, d, 12, seed=77)
    C0 = X[np.random.default_rng(1).choice(n, k, replace=False)].copy()
    ref = orc.train(metric, X, C0, max_iter=6, tol=0.0)
    with native.Handle(d, metric) as h:
        h.set_engine(native.ENGINE_TENSOR)
        h.set_option("reorder_min_rows", 1)
        h.upload_rows(X[:7000])
        h.set_centroids(C0)
        h.step(0.0)                                     # rows get regrouped after this pass
        h.upload_rows(X[7000:])                         # appended after a regrouping
        h.set_centroids(C0)
        for _ in range(6):
            conv, kk, obj = h.step(0.0)
        labels = h.get_labels()
        m = h.get_model()
        idx, dist = h.predict_cached(m["C"][: m["k"]])
        s, _, l2 = h.assign_accumulate(ref["C"], k=ref["k"], want_labels=True)
    assert (labels == ref["labels"]).all()
    np.testing.assert_allclose(m["C"][: m["k"]], ref["C"][: ref["k"]], rtol=1e-9, atol=1e-12)
    idx_r, dist_r, _ = orc.predict(metric, X, m["C"][: m["k"]])
    assert (idx == idx_r).all() and (dist == dist_r).all()
    Cp, s_ref, l_ref, _, _ = oracle_pass(metric, X, ref["C"], k=ref["k"]) if metric == native.EUCLIDEAN else (None, None, None, None, None)
    if l_ref is not None:
        assert (l2 == l_ref).all()


@pytest.mark.parametrize("cfg,n,d,k,metric,need_gb", [
    ("config2", 1_000_000, 784, 10, native.EUCLIDEAN, 8),
    ("config3", 10_000_000, 64, 100, native.EUCLIDEAN, 8),
    ("config4", 100_000_000, 128, 1000, native.EUCLIDEAN, 120),
    ("config5", 20_000_000, 256, 500, native.COSINE, 70),
])
def test_full_size_baseline_configs_tensor_engine_equals_exact_engine(cfg, n, d, k, metric, need_gb):
    """Every BASELINE.json shape at full size: the certified tensor-core path and the Java-order fp64 engine
    (itself pinned on the oracle above) must agree on every label, and the usual size-independent invariants
    must hold.  Rows come from the library's counter-based generator, so nothing is uploaded."""
    seed = 20260101
    free, _ = __import__("torch").cuda.mem_get_info()
    if free < need_gb * (1 << 30):
        pytest.skip("needs ~%d GB of free HBM" % need_gb)
    init_rows = (np.arange(k, dtype=np.int64) * n) // k
    out = {}
    for name, engine, steps in (("tensor", native.ENGINE_TENSOR, 3), ("exact", native.ENGINE_EXACT, 3)):
        with native.Handle(d, metric) as h:
            h.set_engine(engine)
            h.reserve_rows(n)
            h.blobs_append(seed, k, 0, n)
            C0 = h.blobs_rows(seed, k, init_rows).astype(np.float64)
            h.set_centroids(C0)
            objs, ks = [], []
            for _ in range(steps):
                _, kk, obj = h.step(0.0)
                objs.append(obj); ks.append(kk)
                if len(objs) == 2:
                    labels2 = h.get_labels()
                    model2 = h.get_model()
            out[name] = dict(objs=objs, ks=ks, labels=labels2, model=model2, labels3=h.get_labels(), stats=h.stats())
            if name == "tensor":
                # the third pass runs on the cluster-ordered layout with seeded thresholds: invariants only
                lab3 = h.get_labels()
                m3 = h.get_model()
                assert objs[2] <= objs[1] <= objs[0]
                assert int(m3["weights"].sum()) == n                      # every row counted exactly once
                assert lab3.min() >= 0 and lab3.max() < ks[1]
                assert (np.bincount(lab3, minlength=k) == m3["weights"]).all()   # model weights = counts of the last pass
    t, e = out["tensor"], out["exact"]
    assert t["ks"] == e["ks"]
    assert (t["labels"] == e["labels"]).all()
    assert (t["labels3"] == e["labels3"]).all()          # third pass: cluster-ordered rows, seeded windows, pruned tiles
    if k > 256:
        assert t["stats"]["tiles_computed"] <= t["stats"]["tiles_total"]
    for a, b in zip(t["objs"], e["objs"]):
        assert abs(a - b) <= 1e-9 * abs(b)
    kk = e["ks"][1]
    np.testing.assert_allclose(t["model"]["C"][:kk], e["model"]["C"][:kk], rtol=1e-9, atol=1e-12)
    assert (t["model"]["weights"] == e["model"]["weights"]).all()


@pytest.mark.parametrize("metric,d,k", [(native.EUCLIDEAN, 2, 300), (native.EUCLIDEAN, 3, 700), (native.COSINE, 3, 260),
                                        (native.EUCLIDEAN, 5, 100)])
def test_bounds_on_unstructured_low_dimensional_data(metric, d, k):
    # Uniform points: no cluster structure, every row has many centroids within 1.5 x the distance of its own (the
    # recording window of the second-nearest bound fills up, rings overflow), margins are tiny and centroids keep
    # moving for many passes.  Labels, k and objective must equal the exact engine's after every pass.
    n, iters = 70000, 16
    rng = np.random.default_rng(1234 + d + k)
    X = rng.uniform(-1.0, 1.0, size=(n, d))
    if metric == native.COSINE:
        X += 2.0                                                   # keep away from the origin
    X = X.astype(np.float32).astype(np.float64)
    C0 = X[rng.choice(n, k, replace=False)].copy()
    hs = []
    for engine in (native.ENGINE_TENSOR, native.ENGINE_EXACT):
        h = native.Handle(d, metric)
        h.set_engine(engine)
        h.set_option("reorder_min_rows", 1)
        h.upload_rows(X)
        h.set_centroids(C0)
        hs.append(h)
    try:
        for it in range(iters):
            ra, rb = hs[0].step(0.0), hs[1].step(0.0)
            la, lb = hs[0].get_labels(), hs[1].get_labels()
            assert (la == lb).all(), "pass %d: %d labels differ" % (it, int((la != lb).sum()))
            assert ra[1] == rb[1]
            np.testing.assert_allclose(ra[2], rb[2], rtol=1e-9)
        ma, mb = hs[0].get_model(), hs[1].get_model()
        assert (ma["weights"] == mb["weights"]).all()
        np.testing.assert_allclose(ma["C"][: ma["k"]], mb["C"][: mb["k"]], rtol=1e-10, atol=1e-12)
    finally:
        for h in hs:
            h.close()


@pytest.mark.parametrize("metric,fp32", [(native.EUCLIDEAN, True), (native.COSINE, False), (native.EUCLIDEAN, False)])
def test_bounds_stay_valid_over_a_long_run(metric, fp32):
    # 70 passes: second-nearest bounds decay and are refreshed many times, the incremental accumulation goes through
    # its periodic full pass (every 64), clusters that split blobs keep exchanging rows.  Every pass is compared.
    n, d, k, nb, iters = 300000, 48, 600, 420, 70
    X = blobs(n, d, nb, seed=11, spread=6.0, fp32=fp32)      # fp32=False: the partition is held in fp64
    C0 = X[np.random.default_rng(12).choice(n, k, replace=False)].copy()
    hs = []
    for engine in (native.ENGINE_TENSOR, native.ENGINE_EXACT):
        h = native.Handle(d, metric)
        h.set_engine(engine)
        h.upload_rows(X)                                      # above reorder_min_rows: regrouped like a big partition
        h.set_centroids(C0)
        hs.append(h)
    skipped = 0
    try:
        for it in range(iters):
            ra, rb = hs[0].step(0.0), hs[1].step(0.0)
            assert ra[1] == rb[1], "pass %d: k %d vs %d" % (it, ra[1], rb[1])
            np.testing.assert_allclose(ra[2], rb[2], rtol=1e-9, err_msg="pass %d" % it)
            st = hs[0].stats()
            skipped += st["tiles_computed"] * 3 < st["tiles_total"]
            if it % 10 == 9 or it == iters - 1:
                la, lb = hs[0].get_labels(), hs[1].get_labels()
                assert (la == lb).all(), "pass %d: %d labels differ" % (it, int((la != lb).sum()))
        ma, mb = hs[0].get_model(), hs[1].get_model()
        assert (ma["weights"] == mb["weights"]).all()
        np.testing.assert_allclose(ma["C"][: ma["k"]], mb["C"][: mb["k"]], rtol=1e-10, atol=1e-12)
    finally:
        for h in hs:
            h.close()
    if metric == native.EUCLIDEAN:
        assert skipped > 20                                   # the bounds did carry most of the run
