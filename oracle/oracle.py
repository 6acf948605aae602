"""ctypes loader for oracle/libkmeans_oracle.so (TEST INFRASTRUCTURE ONLY).

Every wrapper names the reference function it restates; see kmeans_oracle.c for
file:line citations.  numpy arrays are float64, C-contiguous, row-major
(n x d rows == Alink's d x n column-major DenseMatrix).
"""
import ctypes
import os
import subprocess

import numpy as np

EUCLIDEAN = 0
COSINE = 1

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libkmeans_oracle.so")
_lib = None

_dp = ctypes.POINTER(ctypes.c_double)
_ip = ctypes.POINTER(ctypes.c_int32)
_lp = ctypes.POINTER(ctypes.c_int64)


def build(force=False):
    """Compile the oracle with gcc (-O2 -ffp-contract=off). Building the checker is not using it."""
    src = os.path.join(_HERE, "kmeans_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s", "libkmeans_oracle.so"] + (["-B"] if force else []))
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        L = ctypes.CDLL(_SO)
        L.orc_dot.restype = ctypes.c_double
        L.orc_prepare_vector.restype = ctypes.c_double
        L.orc_shard_start.restype = ctypes.c_int64
        L.orc_shard_count.restype = ctypes.c_int64
        L.orc_shard_start.argtypes = [ctypes.c_int64] * 3
        L.orc_shard_count.argtypes = [ctypes.c_int64] * 3
        _lib = L
    return _lib


def _d(a):
    return a.ctypes.data_as(_dp) if a is not None else None


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def dot(x, y):
    x, y = _f64(x), _f64(y)
    return lib().orc_dot(_d(x), _d(y), ctypes.c_int(x.size))


def axpy(a, x, y):
    x = _f64(x)
    y = _f64(y).copy()
    lib().orc_daxpy(ctypes.c_int(x.size), ctypes.c_double(a), _d(x), _d(y))
    return y


def scal(a, x):
    x = _f64(x).copy()
    lib().orc_dscal(ctypes.c_int(x.size), ctypes.c_double(a), _d(x))
    return x


def gemv(alpha, A_colmajor, m, n, trans, x, beta, y):
    """BLAS.gemv on a column-major m x n matrix given as its flat data array."""
    A, x, y = _f64(A_colmajor), _f64(x), _f64(y).copy()
    f = lib().orc_dgemv_t if trans else lib().orc_dgemv_n
    f(ctypes.c_int(m), ctypes.c_int(n), ctypes.c_double(alpha), _d(A), _d(x), ctypes.c_double(beta), _d(y))
    return y


def prepare_rows(metric, X):
    """FastDistance.prepareVectorData per row. Returns (X', xlabel); X' normalised for COSINE."""
    X = _f64(X).copy()
    n, d = X.shape
    xl = np.zeros(n, dtype=np.float64)
    lib().orc_prepare_rows(ctypes.c_int(metric), _d(X), ctypes.c_int64(n), ctypes.c_int(d), _d(xl))
    return X, xl


def prepare_centroids(metric, C):
    """updateLabel on a FastDistanceMatrixData. Returns (C', clabel)."""
    C = _f64(C).copy()
    k0, d = C.shape
    cl = np.zeros(k0, dtype=np.float64)
    lib().orc_prepare_centroids(ctypes.c_int(metric), _d(C), ctypes.c_int(k0), ctypes.c_int(d), _d(cl))
    return C, cl


def distances(metric, x, xlabel, C, clabel):
    x, C, clabel = _f64(x), _f64(C), _f64(clabel)
    k0, d = C.shape
    res = np.zeros(k0, dtype=np.float64)
    lib().orc_distances(ctypes.c_int(metric), _d(x), ctypes.c_double(xlabel), _d(C), _d(clabel),
                        ctypes.c_int(k0), ctypes.c_int(d), _d(res))
    return res


def min_index(data, end=None):
    data = _f64(data)
    return lib().orc_min_index(_d(data), ctypes.c_int(data.size if end is None else end))


def assign_accumulate(metric, Xp, xlabel, C, clabel, k=None, dup=False, threads=1):
    """KMeansAssignCluster.calc on prepared rows/centroids.
    Returns (sum[k0*(d+1)], labels[n], mind[n])."""
    Xp, C, clabel = _f64(Xp), _f64(C), _f64(clabel)
    xlabel = _f64(xlabel)
    n, d = Xp.shape
    k0 = C.shape[0]
    k = k0 if k is None else k
    s = np.zeros(k0 * (d + 1), dtype=np.float64)
    labels = np.zeros(n, dtype=np.int32)
    mind = np.zeros(n, dtype=np.float64)
    rc = lib().orc_assign_accumulate_mt(
        ctypes.c_int(metric), _d(Xp), _d(xlabel), ctypes.c_int64(n), ctypes.c_int(d), _d(C), _d(clabel),
        ctypes.c_int(k0), ctypes.c_int(k), _d(s), labels.ctypes.data_as(_ip), _d(mind),
        ctypes.c_int(1 if dup else 0), ctypes.c_int(threads))
    if rc:
        raise ArithmeticError("no finite distance for some row (Java: index -1)")
    return s, labels, mind


def update_centroids(metric, sum_, k, k0, d):
    """KMeansUpdateCentroids.updateCentroids. Returns (C_new[k0,d], clabel_new[k0], new_k)."""
    sum_ = _f64(sum_)
    C = np.zeros((k0, d), dtype=np.float64)
    cl = np.zeros(k0, dtype=np.float64)
    nk = lib().orc_update_centroids(ctypes.c_int(metric), _d(sum_), ctypes.c_int(k), ctypes.c_int(k0),
                                    ctypes.c_int(d), _d(C), _d(cl))
    return C, cl, nk


def converged(metric, C1, l1, C2, l2, k, tol):
    """KMeansIterTermination.calc. Returns (bool, diag[k])."""
    C1, l1, C2, l2 = _f64(C1), _f64(l1), _f64(C2), _f64(l2)
    d = C1.shape[1]
    diag = np.zeros(k, dtype=np.float64)
    c = lib().orc_converged(ctypes.c_int(metric), _d(C1), _d(l1), _d(C2), _d(l2), ctypes.c_int(k),
                            ctypes.c_int(d), ctypes.c_double(tol), _d(diag))
    return bool(c), diag


def train(metric, X, C0, max_iter=20, tol=1e-4, threads=1):
    """Whole KMeansTrainBatchOp.iterateICQ loop at P = threads.
    Returns dict(C, k, weights, objective, labels, iters, Xp)."""
    Xp = _f64(X).copy()
    C0 = _f64(C0)
    n, d = Xp.shape
    k0 = C0.shape[0]
    outC = np.zeros((k0, d), dtype=np.float64)
    outk = ctypes.c_int(0)
    w = np.zeros(k0, dtype=np.float64)
    obj = np.zeros(max(max_iter, 1), dtype=np.float64)
    labels = np.zeros(n, dtype=np.int32)
    it = lib().orc_train(ctypes.c_int(metric), _d(Xp), ctypes.c_int64(n), ctypes.c_int(d), _d(C0),
                         ctypes.c_int(k0), ctypes.c_int(max_iter), ctypes.c_double(tol),
                         ctypes.c_int(threads), _d(outC), ctypes.byref(outk), _d(w), _d(obj),
                         labels.ctypes.data_as(_ip))
    if it < 0:
        raise ArithmeticError("no finite distance for some row (Java: index -1)")
    return dict(C=outC, k=outk.value, weights=w, objective=obj[:it], labels=labels, iters=it, Xp=Xp)


def probs(dist):
    dist = _f64(dist)
    p = np.zeros_like(dist)
    lib().orc_probs(_d(dist), ctypes.c_int(dist.size), _d(p))
    return p


def predict(metric, X, C, want_dist=True, want_probs=False):
    """KMeansModelMapper.map over a batch. Returns (idx int64[n], dist|None, probs|None)."""
    X, C = _f64(X), _f64(C)
    n, d = X.shape
    k = C.shape[0]
    idx = np.zeros(n, dtype=np.int64)
    dist = np.zeros(n, dtype=np.float64) if want_dist else None
    pr = np.zeros((n, k), dtype=np.float64) if want_probs else None
    lib().orc_predict(ctypes.c_int(metric), _d(X), ctypes.c_int64(n), ctypes.c_int(d), _d(C),
                      ctypes.c_int(k), idx.ctypes.data_as(_lp), _d(dist), _d(pr))
    return idx, dist, pr


def shard(task, par, total):
    """DefaultDistributedInfo.startPos / localRowCnt."""
    L = lib()
    return L.orc_shard_start(task, par, total), L.orc_shard_count(task, par, total)


def eval_clusters(metric, X, labels, k):
    """EvalClusterBatchOp's vector metrics (getClusterStatistics + toMetrics + silhouette) on raw rows.
    Returns dict(count[k], mean[k,d], compactness[k], dist_sq_sum[k], norm_sq_sum[k], ssb, ssw, compactness_mean,
    separation, davies_bouldin, calinski_harabaz, silhouette, k, total)."""
    X = _f64(X)
    n, d = X.shape
    lab = np.ascontiguousarray(labels, dtype=np.int32)
    cnt = np.zeros(k); mean = np.zeros((k, d)); comp = np.zeros(k); dss = np.zeros(k); ns = np.zeros(k)
    out = np.zeros(9)
    L = lib()
    L.orc_eval_clusters.restype = ctypes.c_int
    rc = L.orc_eval_clusters(ctypes.c_int(metric), _d(X), ctypes.c_int64(n), ctypes.c_int(d),
                             lab.ctypes.data_as(ctypes.POINTER(ctypes.c_int)), ctypes.c_int(k),
                             _d(cnt), _d(mean), _d(comp), _d(dss), _d(ns), _d(out))
    if rc != 0:
        raise MemoryError("orc_eval_clusters")
    return dict(count=cnt, mean=mean, compactness=comp, dist_sq_sum=dss, norm_sq_sum=ns, ssb=out[0], ssw=out[1],
                compactness_mean=out[2], separation=out[3], davies_bouldin=out[4], calinski_harabaz=out[5],
                silhouette=out[6], k=int(out[7]), total=int(out[8]))
