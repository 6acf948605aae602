"""EvalClusterBatchOp: the oracle against the reference's JUnit values, and the host-only metrics."""
import numpy as np

import oracle as orc
from alink_b200.evaluation import label_metrics, summary_to_metrics

# TEST/operator/batch/evaluation/EvalClusterBatchOpTest.java:16-51
POINTS = np.array([[0, 0, 0], [0.1, 0.1, 0.1], [0.2, 0.2, 0.2], [9, 9, 9], [9.1, 9.1, 9.1], [9.2, 9.2, 9.2]], dtype=np.float64)
PRED = np.array([0, 0, 0, 1, 1, 1])


def test_oracle_reproduces_the_junit_metrics():
    r = orc.eval_clusters(orc.EUCLIDEAN, POINTS, PRED, 2)
    assert abs(r["calinski_harabaz"] - 12150.00) < 0.01
    assert abs(r["compactness_mean"] - 0.115) < 0.01
    assert r["total"] == 6 and r["k"] == 2
    assert abs(r["davies_bouldin"] - 0.014) < 0.01
    assert abs(r["separation"] - 15.58) < 0.01
    assert abs(r["ssb"] - 364.5) < 0.01
    assert abs(r["ssw"] - 0.119) < 0.01
    assert abs(r["silhouette"] - 0.99) < 0.01


def test_label_metrics_of_a_perfect_clustering_are_one():
    m = label_metrics(["0", "0", "0", "1", "1", "1"], ["1", "1", "1", "0", "0", "0"])
    assert m["purity"] == 1.0 and abs(m["NMI"] - 1.0) < 1e-12 and m["RI"] == 1.0 and abs(m["ARI"] - 1.0) < 1e-12


def test_label_metrics_against_hand_computed_values():
    # contingency (pred x actual) = [[2, 1], [0, 3]]
    act = ["a", "a", "b", "b", "b", "b"]
    pred = ["x", "x", "x", "y", "y", "y"]
    m = label_metrics(act, pred)
    assert abs(m["purity"] - 5.0 / 6.0) < 1e-15
    tp = 1 + 0 + 0 + 3                          # C(2,2) + C(1,2) + C(0,2) + C(3,2)
    tp_fp, tp_fn, tot = 1 + 6, 3 + 3, 15        # C(2,2)+C(4,2), C(3,2)+C(3,2), C(6,2)
    ri = (tp + (tot - tp - (tp_fp - tp) - (tp_fn - tp))) / tot
    exp = tp_fp * tp_fn / tot
    ari = (tp - exp) / ((tp_fp + tp_fn) / 2 - exp)
    assert abs(m["RI"] - ri) < 1e-15 and abs(m["ARI"] - ari) < 1e-15
    assert 0.0 < m["NMI"] < 1.0


def test_host_remainder_matches_the_oracle():
    rng = np.random.default_rng(0)
    for metric in (orc.EUCLIDEAN, orc.COSINE):
        X = rng.normal(size=(400, 7)) + rng.integers(0, 5, size=(400, 1)) * 3.0
        lab = rng.integers(0, 6, size=400)
        lab[lab == 4] = 5                        # an id that never occurs
        r = orc.eval_clusters(metric, X, lab, 6)
        m = summary_to_metrics(metric, list(range(6)), r["count"], r["mean"], r["compactness"] * np.maximum(r["count"], 1),
                               r["dist_sq_sum"], r["silhouette"] * r["total"])
        assert m["k"] == r["k"] == 5 and m["count"] == 400
        for a, b in (("SSB", "ssb"), ("SSW", "ssw"), ("compactness", "compactness_mean"), ("seperation", "separation"),
                     ("daviesBouldin", "davies_bouldin"), ("calinskiHarabaz", "calinski_harabaz"),
                     ("silhouetteCoefficient", "silhouette")):
            assert abs(m[a] - r[b]) <= 1e-10 * max(1.0, abs(r[b])), (a, m[a], r[b])


def test_cluster_statistics_junit_values():
    # TEST/operator/common/evaluation/ClusterEvaluationUtilTest.java:36-76
    r = orc.eval_clusters(orc.EUCLIDEAN, POINTS[:3], np.zeros(3, dtype=int), 1)
    assert r["count"][0] == 3 and abs(r["compactness"][0] - 0.115) < 0.001
    assert abs(r["dist_sq_sum"][0] - 0.06) < 0.01 and abs(r["norm_sq_sum"][0] - 0.15) < 0.01
    np.testing.assert_allclose(r["mean"][0], [0.1, 0.1, 0.1], rtol=0, atol=1e-15)
    r = orc.eval_clusters(orc.COSINE, POINTS[3:], np.zeros(3, dtype=int), 1)
    assert r["count"][0] == 3 and abs(r["compactness"][0]) < 0.001 and abs(r["dist_sq_sum"][0]) < 0.01
    assert abs(r["norm_sq_sum"][0] - 3.0) < 0.01 and abs((r["mean"][0] ** 2).sum() - 1.0) < 0.01


def test_confusion_matrix_junit_values():
    # ClusterEvaluationUtilTest.java:104-114: matrix {{5,1,2},{1,4,0},{0,1,3}}, rows = predictions, columns = labels
    mat = [[5, 1, 2], [1, 4, 0], [0, 1, 3]]
    act, pred = [], []
    for i, row in enumerate(mat):
        for j, c in enumerate(row):
            act += ["a%d" % j] * c
            pred += ["p%d" % i] * c
    m = label_metrics(act, pred)
    assert abs(m["NMI"] - 0.364) < 0.001 and abs(m["purity"] - 0.705) < 0.001
    assert abs(m["RI"] - 0.68) < 0.01 and abs(m["ARI"] - 0.24) < 0.01
