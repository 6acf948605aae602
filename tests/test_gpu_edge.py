"""Edge shapes and value ranges through both engines: the certificate must hand everything it cannot decide
to the Java-order evaluation, whatever the magnitudes."""
import numpy as np
import pytest

import oracle as orc
from alink_b200 import native

pytestmark = pytest.mark.gpu

ENGINES = [native.ENGINE_EXACT, native.ENGINE_AUTO]


def check(metric, X, C, engine, k=None):
    Xp, xl = orc.prepare_rows(metric, X)
    Cp, cl = orc.prepare_centroids(metric, C)
    s_ref, l_ref, mind = orc.assign_accumulate(metric, Xp, xl, Cp, cl, k=k)
    with native.Handle(X.shape[1], metric) as h:
        h.set_engine(engine)
        h.upload_rows(X)
        s, obj, labels = h.assign_accumulate(Cp, k=k, want_labels=True)
        st = h.stats()
    assert (labels == l_ref).all(), "mismatches: %d" % int((labels != l_ref).sum())
    cnt = s.reshape(C.shape[0], -1)[:, -1]
    assert (cnt == s_ref.reshape(C.shape[0], -1)[:, -1]).all()
    return st


@pytest.mark.parametrize("engine", ENGINES)
@pytest.mark.parametrize("metric", [native.EUCLIDEAN, native.COSINE])
@pytest.mark.parametrize("scale", [1e-30, 1e-12, 1.0, 1e6, 1e18, 1e30])
def test_value_ranges(engine, metric, scale):
    rng = np.random.default_rng(int(abs(np.log10(scale))) + 3)
    n, d, k = 1500, 40, 37
    X = (rng.normal(size=(n, d)) + rng.integers(0, 6, size=(n, 1)) * 4.0) * scale
    X = X.astype(np.float32).astype(np.float64)
    C = X[rng.choice(n, k, replace=False)].copy()
    check(metric, X, C, engine)


@pytest.mark.parametrize("engine", ENGINES)
@pytest.mark.parametrize("metric", [native.EUCLIDEAN, native.COSINE])
def test_mixed_magnitudes_in_one_partition(engine, metric):
    rng = np.random.default_rng(12)
    n, d, k = 2000, 24, 16
    X = rng.normal(size=(n, d)) * 10.0 ** rng.integers(-8, 9, size=(n, 1))    # rows 16 orders of magnitude apart
    C = X[rng.choice(n, k, replace=False)].copy()
    check(metric, X, C, engine)


@pytest.mark.parametrize("engine", ENGINES)
@pytest.mark.parametrize("metric", [native.EUCLIDEAN, native.COSINE])
def test_degenerate_shapes(engine, metric):
    rng = np.random.default_rng(5)
    for n, d, k in [(1, 1, 1), (2, 1, 5), (300, 1, 3), (257, 2, 1), (1000, 3, 300), (513, 129, 2)]:
        X = rng.normal(size=(n, d)) + 3.0
        C = rng.normal(size=(k, d)) + 3.0
        check(metric, X, C, engine)


@pytest.mark.parametrize("engine", ENGINES)
def test_identical_rows_and_centroids(engine):
    X = np.tile(np.array([[1.5, -2.0, 0.25, 8.0]]), (1000, 1))
    C = np.tile(np.array([[1.5, -2.0, 0.25, 8.0]]), (7, 1))                    # every distance ties: index 0 wins
    st = check(native.EUCLIDEAN, X, C, engine)
    assert st["rows"] == 1000


@pytest.mark.parametrize("engine", ENGINES)
def test_zero_vectors_under_cosine(engine):
    rng = np.random.default_rng(7)
    X = rng.normal(size=(600, 8))
    X[::5] = 0.0                                                               # norm 0: left untouched by updateLabel
    C = rng.normal(size=(9, 8))
    C[3] = 0.0
    check(native.COSINE, X, C, engine)


@pytest.mark.parametrize("engine", ENGINES)
def test_only_first_k_of_k0_centroids_compete(engine):
    rng = np.random.default_rng(9)
    X = rng.normal(size=(800, 16)) * 3
    C = rng.normal(size=(40, 16)) * 3
    check(native.EUCLIDEAN, X, C, engine, k=11)


@pytest.mark.parametrize("metric", [native.EUCLIDEAN, native.COSINE])
def test_infinite_values_take_the_exact_path(metric):
    rng = np.random.default_rng(2)
    X = rng.normal(size=(700, 12))
    X[10, 3] = np.inf
    X[11, 0] = -np.inf
    C = rng.normal(size=(5, 12))
    Xp, xl = orc.prepare_rows(metric, X)
    Cp, cl = orc.prepare_centroids(metric, C)
    try:
        _, l_ref, _ = orc.assign_accumulate(metric, Xp, xl, Cp, cl)
    except ArithmeticError:                                                     # Java: getMinPointIndex returns -1
        l_ref = None
    with native.Handle(12, metric) as h:
        h.upload_rows(X)
        if l_ref is None:
            with pytest.raises(native.AkmError):
                h.assign_accumulate(Cp, want_labels=True)
        else:
            _, _, labels = h.assign_accumulate(Cp, want_labels=True)
            assert (labels == l_ref).all()
