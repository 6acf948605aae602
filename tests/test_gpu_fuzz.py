"""Randomised shapes through the whole training loop: tensor engine (regrouping, incremental accumulation, tile
pruning, the two-team epilogue for small k, fp32 and fp64 storage, rows whose byte length is not a multiple of 16)
against the Java-order exact engine and, for the smaller cases, the oracle."""
import numpy as np
import pytest

import oracle as orc
from alink_b200 import native

pytestmark = pytest.mark.gpu

CASES = [  # (seed, n, d, k0, blobs, metric, fp32, spread)
    (1, 3001, 1, 2, 2, native.EUCLIDEAN, True, 10.0),
    (2, 9000, 3, 7, 5, native.COSINE, True, 5.0),
    (3, 40007, 17, 33, 20, native.EUCLIDEAN, False, 4.0),
    (4, 70001, 64, 100, 60, native.EUCLIDEAN, True, 3.0),
    (5, 50000, 130, 129, 100, native.COSINE, False, 6.0),
    (6, 120013, 31, 300, 150, native.EUCLIDEAN, True, 3.0),
    (7, 33333, 129, 513, 400, native.EUCLIDEAN, True, 8.0),
    (8, 25000, 8, 128, 64, native.COSINE, True, 2.0),
]


@pytest.mark.parametrize("seed,n,d,k0,nb,metric,fp32,spread", CASES)
def test_training_loop_tensor_engine_equals_exact_engine(seed, n, d, k0, nb, metric, fp32, spread):
    rng = np.random.default_rng(seed)
    centers = rng.uniform(-spread, spread, size=(nb, d))
    X = centers[rng.integers(0, nb, size=n)] + rng.normal(size=(n, d))
    if fp32:
        X = X.astype(np.float32).astype(np.float64)
    if metric == native.COSINE:
        X[np.abs(X).sum(axis=1) == 0] = 1.0
    C0 = X[rng.choice(n, k0, replace=False)].copy()
    iters = 7
    out = {}
    for engine in (native.ENGINE_AUTO, native.ENGINE_EXACT):
        with native.Handle(d, metric) as h:
            h.set_engine(engine)
            h.set_option("reorder_min_rows", 2048)
            cut = n // 3
            h.upload_rows(X[:cut])                          # two uploads: the store grows (and keeps its headroom)
            h.upload_rows(X[cut:])
            m = h.train(C0, max_iter=iters, tol=0.0)
            out[engine] = (m, h.get_labels(), h.stats())
    ma, la, sa = out[native.ENGINE_AUTO]
    me, le, _ = out[native.ENGINE_EXACT]
    assert ma["k"] == me["k"]
    assert (la == le).all()
    assert (ma["weights"] == me["weights"]).all()
    np.testing.assert_allclose(ma["C"][: ma["k"]], me["C"][: me["k"]], rtol=1e-11, atol=1e-12)
    np.testing.assert_allclose(ma["objective"], me["objective"], rtol=1e-9)
    if n * k0 <= 3_000_000:                                   # the oracle runs these in a second or two
        ref = orc.train(metric, X, C0, max_iter=iters, tol=0.0)
        assert ref["k"] == ma["k"] and (ref["labels"] == la).all()
        np.testing.assert_allclose(ma["C"][: ma["k"]], ref["C"][: ref["k"]], rtol=1e-10, atol=1e-12)
