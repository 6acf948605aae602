"""EvalClusterBatchOp on the GPU (akm_eval_clusters) against the oracle and the reference's JUnit values."""
import numpy as np
import pytest

import oracle as orc
from alink_b200 import native
from alink_b200.evaluation import EvalClusterBatchOp
from alink_b200.kmeans import KMeans, MemSourceBatchOp

pytestmark = pytest.mark.gpu


def test_junit_example_end_to_end():
    # TEST/operator/batch/evaluation/EvalClusterBatchOpTest.java:16-51
    rows = [(0, "0,0,0"), (0, "0.1,0.1,0.1"), (0, "0.2,0.2,0.2"), (1, "9,9,9"), (1, "9.1,9.1,9.1"), (1, "9.2,9.2,9.2")]
    src = MemSourceBatchOp(rows, ["label", "Y"])
    model = KMeans().setVectorCol("Y").setPredictionCol("pred").setK(2).fit(src)
    op = EvalClusterBatchOp().setPredictionCol("pred").setVectorCol("Y").setLabelCol("label")
    m = op.linkFrom(model.transform(src)).collectMetrics()
    assert abs(m.getCalinskiHarabaz() - 12150.00) < 0.01
    assert abs(m.getCompactness() - 0.115) < 0.01
    assert m.getCount() == 6 and m.getK() == 2
    assert abs(m.getDaviesBouldin() - 0.014) < 0.01
    assert abs(m.getSeperation() - 15.58) < 0.01
    assert abs(m.getSsb() - 364.5) < 0.01
    assert abs(m.getSsw() - 0.119) < 0.01
    assert abs(m.getPurity() - 1.0) < 0.01 and abs(m.getNmi() - 1.0) < 0.01
    assert abs(m.getAri() - 1.0) < 0.01 and abs(m.getRi() - 1.0) < 0.01
    assert abs(m.getSilhouetteCoefficient() - 0.99) < 0.01


@pytest.mark.parametrize("metric", [native.EUCLIDEAN, native.COSINE])
@pytest.mark.parametrize("n,d,k", [(5000, 24, 9), (700, 130, 70), (3000, 5, 3)])
def test_device_passes_match_the_oracle(metric, n, d, k):
    rng = np.random.default_rng(n + d)
    centers = rng.uniform(-10, 10, size=(k, d))
    blob = rng.integers(0, k, size=n)
    X = (centers[blob] + rng.normal(size=(n, d))).astype(np.float32).astype(np.float64)
    lab = blob.copy()
    lab[rng.random(n) < 0.05] = rng.integers(0, k)          # some rows in the wrong cluster
    lab[::97] = -1                                           # rows without a prediction
    ref = orc.eval_clusters(metric, X, lab, k + 1)           # id k never occurs
    with native.Handle(d, metric) as h:
        h.upload_rows(X)
        r = h.eval_clusters(lab, k + 1)
    assert (r["count"] == ref["count"]).all()
    np.testing.assert_allclose(r["mean"], ref["mean"], rtol=1e-11, atol=1e-13)
    np.testing.assert_allclose(r["dist_sum"], ref["compactness"] * ref["count"], rtol=1e-10, atol=1e-12)
    np.testing.assert_allclose(r["dist_sq_sum"], ref["dist_sq_sum"], rtol=1e-10, atol=1e-12)
    np.testing.assert_allclose(r["norm_sq_sum"], ref["norm_sq_sum"], rtol=1e-11)
    assert abs(r["silhouette_sum"] / ref["total"] - ref["silhouette"]) <= 1e-9


def test_cluster_ordered_partition_gives_the_same_metrics():
    n, d, k = 6000, 16, 7
    rng = np.random.default_rng(4)
    X = (rng.uniform(-10, 10, size=(k, d))[rng.integers(0, k, size=n)] + rng.normal(size=(n, d))).astype(np.float32).astype(np.float64)
    with native.Handle(d, native.EUCLIDEAN) as h:
        h.set_option("reorder_min_rows", 1)
        h.upload_rows(X)
        h.train(X[:k].copy(), max_iter=5, tol=0.0)
        lab = h.get_labels()
        r = h.eval_clusters(lab, k)
    ref = orc.eval_clusters(orc.EUCLIDEAN, X, lab, k)
    assert (r["count"] == ref["count"]).all()
    np.testing.assert_allclose(r["dist_sq_sum"], ref["dist_sq_sum"], rtol=1e-10)
    assert abs(r["silhouette_sum"] / n - ref["silhouette"]) <= 1e-9
