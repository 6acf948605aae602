"""Pins the CPU oracle against every golden vector the reference's own JUnit
tests hold for the KMeans hot path.  CPU only."""
import numpy as np
import pytest

import oracle as orc


def test_blas_golden(golden):
    g = golden["blas"]
    tol = g["tol"]
    mat, dv1, dv2 = g["mat_2x3_colmajor"], g["dv1"], g["dv2"]
    ones3, ones2 = np.ones(3), np.ones(2)
    np.testing.assert_allclose(orc.gemv(1.0, mat, 2, 3, True, dv1, 0.0, ones3), g["gemv_t"]["beta0"], atol=tol)
    np.testing.assert_allclose(orc.gemv(1.0, mat, 2, 3, True, dv1, 1.0, ones3), g["gemv_t"]["beta1_from_ones"], atol=tol)
    np.testing.assert_allclose(orc.gemv(2.0, mat, 2, 3, False, dv2, 0.0, ones2), g["gemv_n"]["beta0"], atol=tol)
    np.testing.assert_allclose(orc.gemv(2.0, mat, 2, 3, False, dv2, 1.0, ones2), g["gemv_n"]["beta1_from_ones"], atol=tol)
    assert abs(orc.dot(dv1, ones2) - g["dot_dv1_ones"]["value"]) <= tol
    np.testing.assert_allclose(orc.axpy(1.0, dv1, ones2), g["axpy_1_dv1_ones"]["value"], atol=tol)
    np.testing.assert_allclose(orc.scal(0.5, dv1), g["scal_half_dv1"]["value"], atol=tol)


def test_gemm_tn_equals_naive_triple_loop():
    # BLASTest.testGemm :72-92 compares gemm with a naive loop at 1e-8; the oracle's
    # diagonal-of-gemm in orc_converged is the same loop, checked here against numpy.
    rng = np.random.default_rng(0)
    C1, C2 = rng.random((5, 7)), rng.random((5, 7))
    _, l1 = orc.prepare_centroids(orc.EUCLIDEAN, C1)
    _, l2 = orc.prepare_centroids(orc.EUCLIDEAN, C2)
    _, diag = orc.converged(orc.EUCLIDEAN, C1, l1, C2, l2, 5, 1e-4)
    np.testing.assert_allclose(diag, np.linalg.norm(C1 - C2, axis=1), atol=1e-8)


@pytest.mark.parametrize("name,metric", [("euclidean", orc.EUCLIDEAN), ("cosine", orc.COSINE)])
def test_distance_golden(golden, name, metric):
    g = golden[name]
    tol = g["tol"]
    M = np.array([g["v1"], g["v2"]], dtype=np.float64)
    Mp, ml = orc.prepare_centroids(metric, M)
    if metric == orc.COSINE:
        np.testing.assert_allclose(np.linalg.norm(Mp, axis=1), g["unit_norm_after_update_label"]["value"], atol=tol)
    else:
        np.testing.assert_allclose(ml, (M * M).sum(axis=1), atol=1e-3)   # EuclideanDistanceTest :43-54
    # vec x vec via vec x matrix
    Xp, xl = orc.prepare_rows(metric, M[:1])
    d = orc.distances(metric, Xp[0], xl[0], Mp, ml)
    assert abs(d[1] - g["vec_vec"]["value"]) <= tol
    # matrix x matrix (column-major m x n result == [d00, d10, d01, d11])
    Xp, xl = orc.prepare_rows(metric, M)
    mm = [orc.distances(metric, Xp[i], xl[i], Mp, ml)[j] for i in range(2) for j in range(2)]
    np.testing.assert_allclose(mm, g["matrix_matrix"]["value"], atol=tol)
    # sparse vector (densified: adding x_i = 0 terms leaves the sequential sum unchanged)
    sp = np.array([g["sparse1_dense"]], dtype=np.float64)
    Sp, sl = orc.prepare_rows(metric, sp)
    np.testing.assert_allclose(orc.distances(metric, Sp[0], sl[0], Mp, ml), g["sparse1_vs_matrix"]["value"], atol=tol)


def test_update_sum_matrix_golden(golden):
    g = golden["kmeans_util"]
    C = np.array(g["centroids"], dtype=np.float64)
    Cp, cl = orc.prepare_centroids(orc.EUCLIDEAN, C)
    X = np.array([[i, i, i] for i in range(10)], dtype=np.float64)
    Xp, xl = orc.prepare_rows(orc.EUCLIDEAN, X)
    s, labels, _ = orc.assign_accumulate(orc.EUCLIDEAN, Xp, xl, Cp, cl)
    assert s.tolist() == g["update_sum_matrix"]["sum"]            # small integers: exact
    # sparse variant, densified
    Xs = np.zeros((10, 3))
    for i in range(10):
        Xs[i, i % 3] = i * i
    Xp, xl = orc.prepare_rows(orc.EUCLIDEAN, Xs)
    s, _, _ = orc.assign_accumulate(orc.EUCLIDEAN, Xp, xl, Cp, cl)
    assert s.tolist() == g["update_sum_matrix_sparse"]["sum"]
    # threads (Flink subtasks) only change the summation grouping
    s2, l2, _ = orc.assign_accumulate(orc.EUCLIDEAN, Xp, xl, Cp, cl, threads=3)
    assert s2.tolist() == g["update_sum_matrix_sparse"]["sum"]


def test_closest_cluster_golden(golden):
    g = golden["kmeans_util"]
    c = g["closest"]
    Cp, cl = orc.prepare_centroids(orc.EUCLIDEAN, np.array(g["centroids"], dtype=np.float64))
    Xp, xl = orc.prepare_rows(orc.EUCLIDEAN, np.array([c["sample"]], dtype=np.float64))
    d = orc.distances(orc.EUCLIDEAN, Xp[0], xl[0], Cp, cl)
    assert orc.min_index(d) == c["index"]
    assert abs(d[0] - c["d0"]) <= c["tol"] and abs(d[1] - c["d1"]) <= c["tol"]


def test_update_centroids_golden(golden):
    g = golden["update_centroids"]
    C, cl, nk = orc.update_centroids(orc.EUCLIDEAN, g["buffer"], g["k"], g["k"], g["d"])
    assert nk == g["new_k"]
    np.testing.assert_allclose(C.ravel(), g["centroids"], atol=g["tol"])
    np.testing.assert_allclose(cl, (C * C).sum(axis=1), atol=1e-12)


def test_model_mapper_golden_bit_exact(golden):
    from alink_b200.javafmt import java_vector_to_string
    g = golden["model_mapper"]
    C = np.array(golden["kmeans_util"]["centroids"], dtype=np.float64)
    idx, dist, probs = orc.predict(orc.EUCLIDEAN, np.array([g["row"]], dtype=np.float64), C,
                                   want_dist=True, want_probs=True)
    assert idx[0] == g["pred"]
    assert abs(dist[0] - g["distance"]) <= g["distance_tol"]
    # the only bit-exact pin on this path (KMeansModelMapperTest.java:58)
    assert java_vector_to_string(probs[0]) == g["detail_exact"]


def test_pipeline_six_points(golden):
    g = golden["pipeline_6_points"]
    X = np.array(g["rows"], dtype=np.float64)
    # init parity is unpinned (Flink sampler); any init with one centroid per group converges the same
    res = orc.train(orc.EUCLIDEAN, X, X[[0, 3]], max_iter=20, tol=1e-4)
    assert res["k"] == 2
    idx, dist, _ = orc.predict(orc.EUCLIDEAN, X, res["C"][:res["k"]])
    np.testing.assert_allclose(dist, g["pred_distances"], atol=g["tol"])
    want = np.array(g["pred_up_to_permutation"])
    assert (idx == want).all() or (idx == 1 - want).all()
    # SSW = sum of squared distances to the cluster mean (ClusterEvaluationUtil.java:208-218)
    assert abs(float((dist ** 2).sum()) - g["ssw"]) <= 0.01
    assert abs(res["objective"][-1] - g["ssw"]) <= 0.01
    assert res["weights"][:2].tolist() == [3.0, 3.0]


def test_shard_formula(golden):
    for (task, par, total), (start, cnt) in golden["shard"]["cases"]:
        assert orc.shard(task, par, total) == (start, cnt)


def test_empty_cluster_is_dropped_and_ids_shift():
    # KMeansUpdateCentroids.java:58-66: weight==0 clusters are compacted away.
    X = np.array([[0.0], [0.1], [10.0], [10.1]])
    C0 = np.array([[0.0], [100.0], [10.0]])          # middle centroid attracts nothing
    res = orc.train(orc.EUCLIDEAN, X, C0, max_iter=5, tol=0.0)
    assert res["k"] == 2
    np.testing.assert_allclose(res["C"][:2, 0], [0.05, 10.05], rtol=1e-12)
    assert (res["C"][2] == 0).all()
    assert res["labels"].tolist() == [0, 0, 1, 1]


def test_first_minimum_wins_on_exact_ties_and_nan_never_selected():
    assert orc.min_index([3.0, 1.0, 1.0, 2.0]) == 1
    assert orc.min_index([float("nan"), 2.0, float("nan")]) == 1
    assert orc.min_index([float("nan"), float("nan")]) == -1
    assert orc.min_index([1.0, 0.5, 0.25], end=2) == 1


def test_cosine_training_normalises_samples_in_place():
    rng = np.random.default_rng(1)
    X = rng.normal(size=(50, 6)) + 3.0
    res = orc.train(orc.COSINE, X, X[:3], max_iter=4, tol=0.0)
    np.testing.assert_allclose(np.linalg.norm(res["Xp"], axis=1), 1.0, atol=1e-12)
    np.testing.assert_allclose(np.linalg.norm(res["C"][:res["k"]], axis=1), 1.0, atol=1e-12)
