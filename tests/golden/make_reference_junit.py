#!/usr/bin/env python
"""Writes tests/golden/reference_junit.json.

The reference (ZhangYuef/Alink) is Java and this image has no JVM, so golden
vectors cannot be produced by running it.  Every value below is a literal
transcribed from the reference's own JUnit tests / docs for this path; the
`src` of each case is the reference file:line that holds the literal
(TEST = core/src/test/java/com/alibaba/alink).  `tol` is the tolerance the
reference test itself uses; `exact` marks the bit-exact pins.
"""
import json
import os

cases = {
    "blas": {
        "src": "TEST/common/linalg/BLASTest.java:12-17",
        "tol": 1.0e-8,
        "mat_2x3_colmajor": [1, 4, 2, 5, 3, 6],
        "dv1": [1, 2],
        "dv2": [1, 2, 3],
        "gemv_t": {"src": "BLASTest.java:124-132", "alpha": 1.0,
                   "beta0": [9, 12, 15], "beta1_from_ones": [10, 13, 16]},
        "gemv_n": {"src": "BLASTest.java:113-121", "alpha": 2.0,
                   "beta0": [28, 64], "beta1_from_ones": [29, 65]},
        "dot_dv1_ones": {"src": "BLASTest.java:41-44", "value": 3.0},
        "axpy_1_dv1_ones": {"src": "BLASTest.java:47-50", "value": [2, 3]},
        "scal_half_dv1": {"src": "BLASTest.java:29-32", "value": [0.5, 1.0]},
    },
    "euclidean": {
        "src": "TEST/operator/common/distance/EuclideanDistanceTest.java:20-23",
        "tol": 0.01,
        "v1": [1, 2, 4, 1, 3],
        "v2": [4, 6, 1, 2, 4],
        "sparse1_dense": [0, 0.1, 0, 0.4, 0],
        "vec_vec": {"src": ":27-28,63", "value": 6.0},
        "matrix_matrix": {"src": ":76", "value": [0.0, 6.0, 6.0, 0.0]},
        "sparse1_vs_matrix": {"src": ":89", "value": [5.47, 8.38]},
    },
    "cosine": {
        "src": "TEST/operator/common/distance/CosineDistanceTest.java:18-21",
        "tol": 0.01,
        "v1": [1, 2, 4, 1, 3],
        "v2": [4, 6, 1, 2, 4],
        "sparse1_dense": [0, 0.1, 0, 0.4, 0],
        "vec_vec": {"src": ":24-25,56", "value": 0.2852},
        "matrix_matrix": {"src": ":67", "value": [0.0, 0.2852, 0.2852, 0.0]},
        "sparse1_vs_matrix": {"src": ":80", "value": [0.73, 0.60]},
        "unit_norm_after_update_label": {"src": ":40-49", "value": 1.0},
    },
    "kmeans_util": {
        "src": "TEST/operator/common/clustering/kmeans/KMeansUtilTest.java:28-33",
        "model_rows": [
            [0, "{\"vectorCol\":\"\\\"Y\\\"\",\"latitudeCol\":null,\"longitudeCol\":null,"
                "\"distanceType\":\"\\\"EUCLIDEAN\\\"\",\"k\":\"2\",\"vectorSize\":\"3\"}"],
            [1048576, "{\"clusterId\":0,\"weight\":3.0,\"vec\":{\"data\":[9.1,9.1,9.1]}}"],
            [2097152, "{\"clusterId\":1,\"weight\":3.0,\"vec\":{\"data\":[0.1,0.1,0.1]}}"],
        ],
        "centroids": [[9.1, 9.1, 9.1], [0.1, 0.1, 0.1]],
        "update_sum_matrix": {
            "src": ":53-79 (literal at :79; upstream forgets to assert it)",
            "samples": "i*ones(3), i=0..9",
            "sum": [35.0, 35.0, 35.0, 5.0, 10.0, 10.0, 10.0, 5.0]},
        "update_sum_matrix_sparse": {
            "src": ":82-109 (literal at :109; upstream forgets to assert it)",
            "samples": "SparseVector(3, [i%3], [i*i]), i=0..9",
            "sum": [117.0, 65.0, 89.0, 6.0, 9.0, 1.0, 4.0, 4.0]},
        "closest": {"src": ":128-144", "sample": [9, 9, 9], "index": 0,
                    "d0": 0.173, "d1": 15.415, "tol": 0.01},
    },
    "update_centroids": {
        "src": "TEST/operator/common/clustering/kmeans/KMeansUpdateCentroidsTest.java:19-39",
        "tol": 0.01, "d": 2, "k": 5,
        "buffer": [0.5, 0.5, 2.0, 2.5, 2.5, 2.0, 2.0, 2.0, 1.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0],
        "new_k": 3,
        "centroids": [0.25, 0.25, 1.25, 1.25, 2.0, 2.0, 0.0, 0.0, 0.0, 0.0],
    },
    "model_mapper": {
        "src": "TEST/operator/common/clustering/kmeans/KMeansModelMapperTest.java:41,58,75,93-95",
        "row": [0, 0, 0],
        "pred": 1,
        "detail_exact": "0.010869565217391353 0.9891304347826086",
        "distance": 0.173, "distance_tol": 0.001,
    },
    "pipeline_6_points": {
        "src": "TEST/pipeline/clustering/KMeansTest.java:21-49; docs/en/kmeans.md:44-52; "
               "TEST/operator/batch/evaluation/EvalClusterBatchOpTest.java:16-51",
        "rows": [[0, 0, 0], [0.1, 0.1, 0.1], [0.2, 0.2, 0.2], [9, 9, 9], [9.1, 9.1, 9.1], [9.2, 9.2, 9.2]],
        "k": 2,
        "pred_distances": [0.173, 0, 0.173, 0.173, 0, 0.173], "tol": 0.01,
        "pred_up_to_permutation": [1, 1, 1, 0, 0, 0],
        "ssw": 0.119, "ssb": 364.5, "calinski_harabaz": 12150.0, "compactness": 0.115,
    },
    "shard": {
        "src": "core/src/main/java/com/alibaba/alink/common/io/directreader/DefaultDistributedInfo.java:10-35",
        "cases": [  # (task, parallelism, total) -> (start, count), from the formula itself
            [[0, 8, 100000000], [0, 12500000]], [[7, 8, 100000000], [87500000, 12500000]],
            [[0, 3, 10], [0, 4]], [[1, 3, 10], [4, 3]], [[2, 3, 10], [7, 3]],
            [[3, 4, 2], [2, 0]], [[1, 4, 2], [1, 1]],
        ],
    },
}

if __name__ == "__main__":
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "reference_junit.json")
    with open(out, "w") as f:
        json.dump(cases, f, indent=1, sort_keys=True)
    print("wrote", out)
