"""The reference's own operator-level tests, re-run against the GPU path through the mirrored API
(KMeansTest.java, KMeansModelMapperTest.java, EvalClusterBatchOpTest.java, docs/en/kmeans.md)."""
import numpy as np
import pytest

import oracle as orc
from alink_b200 import model as model_io
from alink_b200.kmeans import (KMeans, KMeansPredictBatchOp, KMeansTrainBatchOp, MemSourceBatchOp, Pipeline, Table,
                               VectorAssembler)

pytestmark = pytest.mark.gpu

ROWS = [("0 0 0",), ("0.1 0.1 0.1",), ("0.2 0.2 0.2",), ("9 9 9",), ("9.1 9.1 9.1",), ("9.2 9.2 9.2",)]


def test_kmeans_pipeline_like_KMeansTest(golden):
    g = golden["pipeline_6_points"]
    data = MemSourceBatchOp(ROWS, ["vector"])
    kmeans = KMeans().setVectorCol("vector").setPredictionCol("pred").setPredictionDistanceCol("distance").setK(2)
    model = Pipeline().add(kmeans).fit(data)
    res = model.transform(data)
    assert res.getColNames() == ["vector", "pred", "distance"]
    dist = [r[2] for r in res.collect()]
    np.testing.assert_allclose(dist, g["pred_distances"], atol=g["tol"])
    pred = np.array([r[1] for r in res.collect()])
    want = np.array(g["pred_up_to_permutation"])
    assert (pred == want).all() or (pred == 1 - want).all()          # docs/en/kmeans.md:44-52
    assert abs(sum(x * x for x in dist) - g["ssw"]) <= 0.01          # EvalClusterBatchOpTest: SSW 0.119


def test_model_mapper_like_KMeansModelMapperTest(golden):
    g = golden["model_mapper"]
    model = Table(golden["kmeans_util"]["model_rows"], ["model_id", "model_info"])
    data = Table([("0 0 0",)], ["Y"])
    out = KMeansPredictBatchOp().setPredictionCol("pred").linkFrom(model, data)
    assert out.getColNames() == ["Y", "pred"] and out.collect()[0][1] == g["pred"]
    out = KMeansPredictBatchOp().setPredictionCol("pred").setPredictionDetailCol("detail").linkFrom(model, data)
    assert out.getColNames() == ["Y", "pred", "detail"] and out.collect()[0][2] == g["detail_exact"]
    out = KMeansPredictBatchOp().setPredictionCol("pred").setPredictionDistanceCol("distance").linkFrom(model, data)
    assert out.getColNames() == ["Y", "pred", "distance"]
    assert abs(out.collect()[0][2] - g["distance"]) <= g["distance_tol"]
    out = (KMeansPredictBatchOp().setPredictionCol("pred").setPredictionDetailCol("detail")
           .setPredictionDistanceCol("distance").linkFrom(model, data))
    row = out.collect()[0]
    assert out.getColNames() == ["Y", "pred", "detail", "distance"]
    assert row[2] == g["detail_exact"] and abs(row[3] - g["distance"]) <= g["distance_tol"]


def test_train_op_writes_reference_wire_format_and_matches_oracle():
    rng = np.random.default_rng(4)
    X = (rng.normal(size=(500, 6)) + rng.integers(0, 3, size=(500, 1)) * 5.0).round(3)
    rows = [(i, " ".join(repr(float(v)) for v in x)) for i, x in enumerate(X)]
    data = Table(rows, ["id", "vec"])
    C0 = X[[0, 200, 400]].copy()
    op = KMeansTrainBatchOp().setVectorCol("vec").setK(3).setMaxIter(30).setInitCentroids(C0).linkFrom(data)
    md = model_io.load_model(op.collect())
    ref = orc.train(orc.EUCLIDEAN, X, C0, max_iter=30, tol=1e-4)
    assert md["k"] == ref["k"] and md["vectorSize"] == 6 and md["vectorCol"] == "vec"
    np.testing.assert_allclose(md["centroids"], ref["C"][: ref["k"]], rtol=1e-12)
    assert md["weights"].tolist() == ref["weights"][: ref["k"]].tolist()
    assert op.collect()[0][0] == 0 and op.collect()[1][0] == 1048576
    # predict with reserved columns + a null vector row
    pdata = Table(rows[:5] + [(99, None)], ["id", "vec"])
    out = (KMeansPredictBatchOp().setPredictionCol("cluster").setPredictionDistanceCol("dist").setReservedCols(["id"])
           .linkFrom(op, pdata))
    assert out.getColNames() == ["id", "cluster", "dist"]
    got = out.collect()
    assert got[5] == (99, None, None)
    idx_ref, dist_ref, _ = orc.predict(orc.EUCLIDEAN, X[:5], md["centroids"])     # same model bytes on both sides
    assert [r[1] for r in got[:5]] == idx_ref.tolist() and [r[2] for r in got[:5]] == dist_ref.tolist()


def test_cosine_operator_and_errors():
    rng = np.random.default_rng(8)
    X = np.abs(rng.normal(size=(300, 5))) + np.eye(5)[rng.integers(0, 5, size=300)] * 6
    data = Table([(" ".join(map(str, x)),) for x in X], ["v"])
    C0 = X[:5].copy()
    op = KMeansTrainBatchOp().setVectorCol("v").setK(5).setDistanceType("cosine").setInitCentroids(C0).linkFrom(data)
    md = model_io.load_model(op.collect())
    ref = orc.train(orc.COSINE, X, C0, max_iter=20, tol=1e-4)
    assert md["distanceType"] == "COSINE" and md["k"] == ref["k"]
    np.testing.assert_allclose(md["centroids"], ref["C"][: ref["k"]], rtol=1e-12)
    with pytest.raises(RuntimeError, match="The train dataset is empty!"):
        KMeansTrainBatchOp().setVectorCol("v").linkFrom(Table([], ["v"]))
    with pytest.raises(ValueError):
        KMeansTrainBatchOp().setDistanceType("MANHATTAN")          # HasKMeansDistanceType admits two metrics
    with pytest.raises(ValueError, match="Can not find column"):
        KMeansTrainBatchOp().setVectorCol("nope").linkFrom(data)


def test_iris_pipeline_config1():
    # BASELINE config 1: iris k=3 via Pipeline(VectorAssembler + KMeans), parallelism 1
    from sklearn.datasets import load_iris
    iris = load_iris()
    cols = ["sepal_length", "sepal_width", "petal_length", "petal_width"]
    data = Table([tuple(map(float, r)) for r in iris.data], cols)
    va = VectorAssembler().setSelectedCols(cols).setOutputCol("features")
    km = KMeans().setVectorCol("features").setK(3).setPredictionCol("cluster_id").setInitCentroids(iris.data[[0, 50, 100]])
    model = Pipeline().add(va).add(km).fit(data)
    out = model.transform(data)
    pred = np.array([r[-1] for r in out.collect()])
    ref = orc.train(orc.EUCLIDEAN, iris.data, iris.data[[0, 50, 100]], max_iter=20, tol=1e-4)
    idx_ref, _, _ = orc.predict(orc.EUCLIDEAN, iris.data, ref["C"][: ref["k"]])
    assert (pred == idx_ref).all()
    assert len(set(pred.tolist())) == 3
