"""ctypes binding of the C ABI in include/akm.h (alink_b200/csrc/libakm.so).

There is no CPU fallback: importing works without a GPU (so the CPU test suite can check
that the library loads and exports every declared symbol), but every compute entry point
needs a CUDA device and raises AkmError otherwise.  A missing library raises at load time.
"""
import ctypes
import os
import re

import numpy as np

EUCLIDEAN = 0
COSINE = 1
ENGINE_AUTO, ENGINE_EXACT, ENGINE_TENSOR = 0, 1, 2
UNIQUE_ID_BYTES = 128

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("AKM_LIB") or os.path.join(_HERE, "csrc", "libakm.so")   # AKM_LIB: A/B builds
HEADER_PATH = os.path.join(os.path.dirname(_HERE), "include", "akm.h")

ERR_NAMES = {-1: "BADARG", -2: "CUDA", -3: "NCCL", -4: "OOM", -5: "NOFINITE", -6: "STATE", -7: "UNSUPPORTED"}


class AkmError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("akm error %s (%d): %s" % (ERR_NAMES.get(code, "?"), code, msg))
        self.code = code


class Stats(ctypes.Structure):
    _fields_ = [("rows", ctypes.c_int64), ("rows_certified", ctypes.c_int64), ("rows_refined", ctypes.c_int64),
                ("kernel_launches", ctypes.c_int64), ("engine_used", ctypes.c_int32), ("storage_bytes", ctypes.c_int32),
                ("rows_exact_all", ctypes.c_int64), ("tiles_total", ctypes.c_int64), ("tiles_computed", ctypes.c_int64),
                ("rows_moved", ctypes.c_int64), ("accumulate_mode", ctypes.c_int32), ("reserved0", ctypes.c_int32)]


_lib = None
_c_dp = ctypes.POINTER(ctypes.c_double)
_c_fp = ctypes.POINTER(ctypes.c_float)
_c_ip = ctypes.POINTER(ctypes.c_int32)
_c_lp = ctypes.POINTER(ctypes.c_int64)
_H = ctypes.c_void_p


def declared_symbols():
    """Names of every function include/akm.h declares (used by the CPU export test)."""
    with open(HEADER_PATH) as f:
        text = f.read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(akm_[a-z0-9_]+)\s*\(", text)))


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError("libakm.so is not built (%s); run `python -c 'import __graft_entry__ as g; g.build()'` "
                              "or `make -C alink_b200/csrc` -- there is no CPU fallback" % LIB_PATH)
        L = ctypes.CDLL(LIB_PATH, mode=ctypes.RTLD_GLOBAL)
        L.akm_last_error.restype = ctypes.c_char_p
        L.akm_last_error.argtypes = [_H]
        L.akm_version.restype = ctypes.c_char_p
        L.akm_create.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.POINTER(_H)]
        L.akm_destroy.argtypes = [_H]
        L.akm_set_stream.argtypes = [_H, ctypes.c_void_p]
        L.akm_set_engine.argtypes = [_H, ctypes.c_int]
        L.akm_reserve_rows.argtypes = [_H, ctypes.c_int64]
        L.akm_upload_rows.argtypes = [_H, _c_dp, ctypes.c_int64]
        L.akm_upload_rows_f32.argtypes = [_H, _c_fp, ctypes.c_int64]
        L.akm_upload_rows_prepared.argtypes = [_H, _c_dp, ctypes.c_int64]
        L.akm_upload_rows_csr.argtypes = [_H, _c_lp, _c_ip, _c_dp, ctypes.c_int64]
        L.akm_upload_wait.argtypes = [_H]
        L.akm_host_alloc.argtypes = [ctypes.c_size_t, ctypes.POINTER(ctypes.c_void_p)]
        L.akm_host_free.argtypes = [ctypes.c_void_p]
        L.akm_superstep.argtypes = [_H, _c_dp, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_double, _c_dp,
                                    ctypes.POINTER(ctypes.c_int), ctypes.POINTER(ctypes.c_int), _c_dp, _c_dp, _c_dp]
        L.akm_upload_rows_device_f32.argtypes = [_H, ctypes.c_void_p, ctypes.c_int64]
        L.akm_clear_rows.argtypes = [_H]
        L.akm_num_rows.argtypes = [_H, _c_lp, ctypes.POINTER(ctypes.c_int)]
        L.akm_assign_accumulate.argtypes = [_H, _c_dp, ctypes.c_int, ctypes.c_int, _c_dp, _c_dp, _c_ip]
        L.akm_update_centroids.argtypes = [_H, _c_dp, ctypes.c_int, ctypes.c_int, ctypes.c_int, _c_dp, _c_dp,
                                           ctypes.POINTER(ctypes.c_int), _c_dp]
        L.akm_set_centroids.argtypes = [_H, _c_dp, ctypes.c_int]
        L.akm_step.argtypes = [_H, ctypes.c_double, ctypes.POINTER(ctypes.c_int), ctypes.POINTER(ctypes.c_int), _c_dp]
        L.akm_get_model.argtypes = [_H, _c_dp, ctypes.POINTER(ctypes.c_int), _c_dp, ctypes.POINTER(ctypes.c_int)]
        L.akm_get_labels.argtypes = [_H, _c_ip]
        L.akm_comm_unique_id.argtypes = [ctypes.c_void_p]
        L.akm_comm_init.argtypes = [_H, ctypes.c_void_p, ctypes.c_int, ctypes.c_int]
        L.akm_allreduce_sum.argtypes = [_H, _c_dp, ctypes.c_int64]
        L.akm_predict.argtypes = [_H, _c_dp, ctypes.c_int64, _c_dp, ctypes.c_int, _c_lp, _c_dp, _c_dp]
        L.akm_predict_cached.argtypes = [_H, _c_dp, ctypes.c_int, _c_lp, _c_dp]
        L.akm_get_stats.argtypes = [_H, ctypes.POINTER(Stats)]
        L.akm_set_option.argtypes = [_H, ctypes.c_char_p, ctypes.c_double]
        L.akm_init_reset.argtypes = [_H]
        L.akm_init_update.argtypes = [_H, _c_dp, ctypes.c_int, ctypes.c_int, _c_dp]
        L.akm_init_pick.argtypes = [_H, ctypes.c_double, ctypes.c_uint64, ctypes.c_int, ctypes.c_int64, ctypes.c_int, _c_lp,
                                    ctypes.POINTER(ctypes.c_int)]
        L.akm_init_counts.argtypes = [_H, ctypes.c_int, _c_dp]
        L.akm_get_rows.argtypes = [_H, _c_lp, ctypes.c_int, _c_dp]
        L.akm_eval_clusters.argtypes = [_H, _c_ip, ctypes.c_int, _c_dp, _c_dp, _c_dp, _c_dp, _c_dp, _c_dp]
        L.akm_set_profiling.argtypes = [_H, ctypes.c_int]
        L.akm_get_profile.argtypes = [_H, _c_dp, ctypes.c_int]
        L.akm_blobs_append.argtypes = [_H, ctypes.c_uint64, ctypes.c_int, ctypes.c_double, ctypes.c_double,
                                       ctypes.c_int64, ctypes.c_int64]
        L.akm_blobs_rows.argtypes = [_H, ctypes.c_uint64, ctypes.c_int, ctypes.c_double, ctypes.c_double, _c_lp,
                                     ctypes.c_int64, _c_fp]
        _lib = L
    return _lib


def _dp(a):
    return a.ctypes.data_as(_c_dp) if a is not None else None


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


class PinnedArray:
    """A numpy view of page-locked host memory from akm_host_alloc: uploads from it skip the staging copy."""

    def __init__(self, shape, dtype=np.float64):
        self.shape = tuple(int(x) for x in shape)
        self.dtype = np.dtype(dtype)
        nbytes = int(np.prod(self.shape)) * self.dtype.itemsize
        self._p = ctypes.c_void_p()
        rc = lib().akm_host_alloc(nbytes, ctypes.byref(self._p))
        if rc:
            raise AkmError(rc, lib().akm_last_error(None).decode())
        buf = (ctypes.c_char * max(nbytes, 1)).from_address(self._p.value)
        self.array = np.frombuffer(buf, dtype=self.dtype, count=int(np.prod(self.shape))).reshape(self.shape)

    def free(self):
        if self._p:
            self.array = None
            lib().akm_host_free(self._p)
            self._p = ctypes.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.free()


def unique_id():
    buf = ctypes.create_string_buffer(UNIQUE_ID_BYTES)
    rc = lib().akm_comm_unique_id(buf)
    if rc:
        raise AkmError(rc, lib().akm_last_error(None).decode())
    return buf.raw


class Handle:
    """One partition of samples cached on one GPU (one Flink subtask's share of the ComQueue state)."""

    def __init__(self, d, metric=EUCLIDEAN, device=0):
        self._h = _H()
        self.d = int(d)
        self.metric = int(metric)
        L = lib()
        rc = L.akm_create(int(device), self.d, self.metric, ctypes.byref(self._h))
        if rc:
            raise AkmError(rc, L.akm_last_error(None).decode())

    def _ck(self, rc):
        if rc:
            raise AkmError(rc, lib().akm_last_error(self._h).decode())

    def close(self):
        if self._h:
            lib().akm_destroy(self._h)
            self._h = _H()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # ---- configuration
    def set_stream(self, cuda_stream_ptr):
        self._ck(lib().akm_set_stream(self._h, ctypes.c_void_p(cuda_stream_ptr)))

    def set_engine(self, engine):
        self._ck(lib().akm_set_engine(self._h, int(engine)))

    # ---- rows
    def reserve_rows(self, n):
        self._ck(lib().akm_reserve_rows(self._h, int(n)))

    def upload_rows(self, X):
        X = np.asarray(X)
        if X.ndim != 2 or X.shape[1] != self.d:
            raise ValueError("rows must be n x %d" % self.d)
        if X.dtype == np.float32:
            X = np.ascontiguousarray(X)
            self._ck(lib().akm_upload_rows_f32(self._h, X.ctypes.data_as(_c_fp), X.shape[0]))
        else:
            X = _f64(X)
            self._ck(lib().akm_upload_rows(self._h, _dp(X), X.shape[0]))

    def upload_rows_prepared(self, X):
        """Rows that already went through prepareVectorData on the caller's side (COSINE: unit vectors, stored as given)."""
        X = _f64(X)
        if X.ndim != 2 or X.shape[1] != self.d:
            raise ValueError("rows must be n x %d" % self.d)
        self._ck(lib().akm_upload_rows_prepared(self._h, _dp(X), X.shape[0]))

    def upload_rows_csr(self, indptr, indices, values):
        """Sparse samples in CSR form (scipy.sparse.csr_matrix fields): expanded to dense rows on the device."""
        indptr = np.ascontiguousarray(indptr, dtype=np.int64)
        indices = np.ascontiguousarray(indices, dtype=np.int32)
        values = np.ascontiguousarray(values, dtype=np.float64)
        self._ck(lib().akm_upload_rows_csr(self._h, indptr.ctypes.data_as(_c_lp), indices.ctypes.data_as(_c_ip), _dp(values),
                                           indptr.size - 1))

    def upload_wait(self):
        self._ck(lib().akm_upload_wait(self._h))

    def upload_rows_device_f32(self, dev_ptr, n):
        self._ck(lib().akm_upload_rows_device_f32(self._h, ctypes.c_void_p(int(dev_ptr)), int(n)))

    def clear_rows(self):
        self._ck(lib().akm_clear_rows(self._h))

    def num_rows(self):
        n = ctypes.c_int64()
        b = ctypes.c_int()
        self._ck(lib().akm_num_rows(self._h, ctypes.byref(n), ctypes.byref(b)))
        return n.value, b.value

    # ---- host-buffer supersteps
    def assign_accumulate(self, C, k=None, want_labels=False):
        """KMeansAssignCluster.calc. Returns (sum[k0*(d+1)], objective, labels|None)."""
        C = _f64(C)
        k0 = C.shape[0]
        k = k0 if k is None else int(k)
        s = np.empty(k0 * (self.d + 1), dtype=np.float64)
        obj = ctypes.c_double()
        n, _ = self.num_rows()
        labels = np.empty(n, dtype=np.int32) if want_labels else None
        self._ck(lib().akm_assign_accumulate(self._h, _dp(C), k0, k, _dp(s), ctypes.byref(obj),
                                             labels.ctypes.data_as(_c_ip) if want_labels else None))
        return s, obj.value, labels

    def update_centroids(self, sum_, k0, k, step_no=1, prev=None):
        """KMeansUpdateCentroids.updateCentroids (+ shift diag). Returns (C[k0,d], new_k, shift|None)."""
        sum_ = _f64(sum_)
        C = np.empty((k0, self.d), dtype=np.float64)
        nk = ctypes.c_int()
        prev = _f64(prev) if prev is not None else None
        shift = np.zeros(k0, dtype=np.float64) if prev is not None else None
        self._ck(lib().akm_update_centroids(self._h, _dp(sum_), k0, int(k), int(step_no), _dp(prev), _dp(C),
                                            ctypes.byref(nk), _dp(shift)))
        return C, nk.value, (shift[:nk.value] if shift is not None else None)

    def superstep(self, C, k=None, step_no=1, tol=0.0, want_sum=False, want_shift=False, out=None):
        """akm_superstep: assign + all-reduce + update + criterion with host buffers.
        Returns dict(C[k0,d], k, converged, objective, sum|None, shift|None).  `out`: a k0 x d float64 array to receive
        the new centroids (e.g. a PinnedArray's view: page-locked buffers skip the library's bounce copies)."""
        if not (isinstance(C, np.ndarray) and C.dtype == np.float64 and C.flags.c_contiguous):
            C = _f64(C)
        k0 = C.shape[0]
        k = k0 if k is None else int(k)
        Cn = out if out is not None else np.empty((k0, self.d), dtype=np.float64)
        nk, conv, obj = ctypes.c_int(), ctypes.c_int(), ctypes.c_double()
        s = np.empty(k0 * (self.d + 1), dtype=np.float64) if want_sum else None
        shift = np.zeros(k0, dtype=np.float64) if want_shift else None
        self._ck(lib().akm_superstep(self._h, _dp(C), k0, k, int(step_no), float(tol), _dp(Cn), ctypes.byref(nk),
                                     ctypes.byref(conv), _dp(s), _dp(shift), ctypes.byref(obj)))
        return dict(C=Cn, k=nk.value, converged=bool(conv.value), objective=obj.value, sum=s,
                    shift=(shift[:nk.value] if shift is not None else None))

    # ---- device-resident loop
    def set_centroids(self, C0):
        C0 = _f64(C0)
        self._ck(lib().akm_set_centroids(self._h, _dp(C0), C0.shape[0]))
        self.k0 = C0.shape[0]

    def step(self, tol=1e-4):
        """One superstep. Returns (converged, k, objective)."""
        conv, k, obj = ctypes.c_int(), ctypes.c_int(), ctypes.c_double()
        self._ck(lib().akm_step(self._h, float(tol), ctypes.byref(conv), ctypes.byref(k), ctypes.byref(obj)))
        return bool(conv.value), k.value, obj.value

    def get_model(self):
        """KMeansOutputModel.calc. Returns dict(C[k0,d], k, weights[k0], steps)."""
        C = np.empty((self.k0, self.d), dtype=np.float64)
        w = np.empty(self.k0, dtype=np.float64)
        k, st = ctypes.c_int(), ctypes.c_int()
        self._ck(lib().akm_get_model(self._h, _dp(C), ctypes.byref(k), _dp(w), ctypes.byref(st)))
        return dict(C=C, k=k.value, weights=w, steps=st.value)

    def get_labels(self):
        n, _ = self.num_rows()
        labels = np.empty(n, dtype=np.int32)
        self._ck(lib().akm_get_labels(self._h, labels.ctypes.data_as(_c_ip)))
        return labels

    def train(self, C0, max_iter=20, tol=1e-4):
        """KMeansTrainBatchOp.iterateICQ: set centroids, iterate until the criterion or maxIter."""
        self.set_centroids(C0)
        objective = []
        for _ in range(int(max_iter)):
            conv, k, obj = self.step(tol)
            objective.append(obj)
            if conv:
                break
        m = self.get_model()
        m["objective"] = np.array(objective)
        m["iters"] = m["steps"]
        return m

    # ---- NCCL
    def comm_init(self, uid_bytes, rank, world):
        buf = ctypes.create_string_buffer(uid_bytes, UNIQUE_ID_BYTES)
        self._ck(lib().akm_comm_init(self._h, buf, int(rank), int(world)))

    def allreduce_sum(self, a):
        a = _f64(a).copy()
        self._ck(lib().akm_allreduce_sum(self._h, _dp(a), a.size))
        return a

    # ---- prediction
    def predict(self, X, C, want_dist=True, want_probs=False):
        """KMeansModelMapper.map over a batch. Returns (idx int64[n], dist|None, probs|None)."""
        X, C = _f64(X), _f64(C)
        n, k = X.shape[0], C.shape[0]
        idx = np.empty(n, dtype=np.int64)
        dist = np.empty(n, dtype=np.float64) if want_dist else None
        probs = np.empty((n, k), dtype=np.float64) if want_probs else None
        self._ck(lib().akm_predict(self._h, _dp(X), n, _dp(C), k, idx.ctypes.data_as(_c_lp), _dp(dist), _dp(probs)))
        return idx, dist, probs

    def predict_cached(self, C, want_dist=True):
        C = _f64(C)
        n, _ = self.num_rows()
        idx = np.empty(n, dtype=np.int64)
        dist = np.empty(n, dtype=np.float64) if want_dist else None
        self._ck(lib().akm_predict_cached(self._h, _dp(C), C.shape[0], idx.ctypes.data_as(_c_lp), _dp(dist)))
        return idx, dist

    # ---- introspection
    def stats(self):
        st = Stats()
        self._ck(lib().akm_get_stats(self._h, ctypes.byref(st)))
        return {f: getattr(st, f) for f, _ in Stats._fields_}

    def set_option(self, name, value):
        self._ck(lib().akm_set_option(self._h, name.encode(), float(value)))

    def set_profiling(self, on=True):
        self._ck(lib().akm_set_profiling(self._h, 1 if on else 0))

    def get_profile(self, reset=False):
        out = np.zeros(8, dtype=np.float64)
        self._ck(lib().akm_get_profile(self._h, _dp(out), 1 if reset else 0))
        return dict(search_ms=out[0], refine_ms=out[1], accumulate_ms=out[2], update_ms=out[3], steps=int(out[4]),
                    reorder_ms=out[5], reorders=int(out[6]))

    # ---- synthetic benchmark data
    def blobs_append(self, seed, k_true, row0, n, spread=10.0, sigma=1.0):
        self._ck(lib().akm_blobs_append(self._h, int(seed), int(k_true), float(spread), float(sigma), int(row0), int(n)))

    # ---- initial centroids (KMeansInitCentroids.java:89-165), see include/akm.h
    def init_reset(self):
        self._ck(lib().akm_init_reset(self._h))

    def init_update(self, centers, id_base):
        C = np.ascontiguousarray(centers, dtype=np.float64)
        out = np.zeros(1, dtype=np.float64)
        self._ck(lib().akm_init_update(self._h, C.ctypes.data_as(_c_dp), C.shape[0], int(id_base), out.ctypes.data_as(_c_dp)))
        return float(out[0])

    def init_pick(self, scale, seed, round_no, row_base, cap):
        rows = np.empty(max(int(cap), 1), dtype=np.int64)
        cnt = ctypes.c_int(0)
        self._ck(lib().akm_init_pick(self._h, float(scale), int(seed) & 0xFFFFFFFFFFFFFFFF, int(round_no), int(row_base),
                                     rows.size, rows.ctypes.data_as(_c_lp), ctypes.byref(cnt)))
        if cnt.value > rows.size:
            raise RuntimeError("k-means|| picked %d rows, more than the %d the caller allowed for" % (cnt.value, rows.size))
        return rows[:cnt.value].copy()

    def init_counts(self, m_total):
        out = np.zeros(int(m_total), dtype=np.float64)
        self._ck(lib().akm_init_counts(self._h, int(m_total), out.ctypes.data_as(_c_dp)))
        return out

    def get_rows(self, rows_sorted):
        rows = np.ascontiguousarray(rows_sorted, dtype=np.int64)
        out = np.zeros((rows.size, self.d), dtype=np.float64)
        if rows.size:
            self._ck(lib().akm_get_rows(self._h, rows.ctypes.data_as(_c_lp), rows.size, out.ctypes.data_as(_c_dp)))
        return out

    def eval_clusters(self, labels, k):
        """EvalClusterBatchOp's per-row passes (akm_eval_clusters). Returns dict of per-cluster arrays + silhouette sum."""
        lab = np.ascontiguousarray(labels, dtype=np.int32)
        k = int(k)
        cnt = np.zeros(k); mean = np.zeros((k, self.d)); ds = np.zeros(k); dss = np.zeros(k); ns = np.zeros(k)
        sil = np.zeros(1)
        self._ck(lib().akm_eval_clusters(self._h, lab.ctypes.data_as(_c_ip), k, cnt.ctypes.data_as(_c_dp),
                                         mean.ctypes.data_as(_c_dp), ds.ctypes.data_as(_c_dp), dss.ctypes.data_as(_c_dp),
                                         ns.ctypes.data_as(_c_dp), sil.ctypes.data_as(_c_dp)))
        return dict(count=cnt, mean=mean, dist_sum=ds, dist_sq_sum=dss, norm_sq_sum=ns, silhouette_sum=float(sil[0]))

    def blobs_rows(self, seed, k_true, rows, spread=10.0, sigma=1.0):
        rows = np.ascontiguousarray(rows, dtype=np.int64)
        out = np.empty((rows.size, self.d), dtype=np.float32)
        self._ck(lib().akm_blobs_rows(self._h, int(seed), int(k_true), float(spread), float(sigma),
                                      rows.ctypes.data_as(_c_lp), rows.size, out.ctypes.data_as(_c_fp)))
        return out
