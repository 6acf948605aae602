"""Host-side mirror of Alink's KMeans operator API on top of the C ABI (no JVM in this image, so the
host layer the north_star asks for in Java is mirrored here in Python; the Java shim sources are
under java/ and INTEGRATION.md).  Names, parameters, defaults and error behaviour follow the reference:

  KMeansTrainBatchOp    core/src/main/java/com/alibaba/alink/operator/batch/clustering/KMeansTrainBatchOp.java:31-119
  KMeansPredictBatchOp  .../KMeansPredictBatchOp.java:12-28  (+ ModelMapBatchOp.java:34-54, KMeansModelMapper.java:34-105)
  KMeans / KMeansModel  core/src/main/java/com/alibaba/alink/pipeline/clustering/KMeans.java:17-33, KMeansModel.java
  params                params/clustering/BaseKMeansTrainParams.java:20-35 (maxIter=20 alias numIter, initMode
                        K_MEANS_PARALLEL, initSteps 2), HasKDefaultAs2 (k=2), HasEpsilonDv00001 (1e-4),
                        HasKMeansDistanceType (EUCLIDEAN|COSINE, alias metric), KMeansPredictParams, RichModelMapperParams
"""
import numpy as np

from . import model as model_io
from . import native
from . import init as init_mod
from .javafmt import java_vector_to_string
from .vector import is_sparse_text, rows_to_csr, rows_to_matrix

_DIST = {"EUCLIDEAN": native.EUCLIDEAN, "COSINE": native.COSINE}


def shard_bounds(task, parallelism, total):
    """DefaultDistributedInfo.startPos / localRowCnt (common/io/directreader/DefaultDistributedInfo.java:10-35)."""
    div, mod = divmod(int(total), int(parallelism))
    if mod == 0:
        return div * task, div
    if task >= mod:
        return div * task + mod, div
    return div * task + task, div + 1


class Table:
    """Minimal stand-in for a Flink Table / MemSourceBatchOp: rows + column names."""

    def __init__(self, rows, colNames):
        self.rows = [tuple(r) if isinstance(r, (list, tuple)) else (r,) for r in rows]
        self.colNames = list(colNames)

    def col(self, name):
        if name not in self.colNames:
            raise ValueError("Can not find column: " + str(name))        # TableUtil.findColIndexWithAssert
        i = self.colNames.index(name)
        return [r[i] for r in self.rows]

    def collect(self):
        return list(self.rows)

    def getColNames(self):
        return list(self.colNames)


MemSourceBatchOp = Table


class _WithParams:
    _defaults = {}

    def __init__(self, params=None):
        self.params = dict(self._defaults)
        if params:
            self.params.update(params)

    def _set(self, key, value):
        self.params[key] = value
        return self

    def get(self, key):
        return self.params.get(key)


class _TrainParams(_WithParams):
    _defaults = {"k": 2, "maxIter": 20, "epsilon": 1.0e-4, "distanceType": "EUCLIDEAN",
                 "initMode": "K_MEANS_PARALLEL", "initSteps": 2, "vectorCol": None, "initCentroids": None,
                 "randomSeed": 0}

    def setK(self, v): return self._set("k", int(v))
    def setMaxIter(self, v): return self._set("maxIter", int(v))
    setNumIter = setMaxIter                                       # alias "numIter"
    def setEpsilon(self, v): return self._set("epsilon", float(v))
    def setVectorCol(self, v): return self._set("vectorCol", v)
    def setInitMode(self, v): return self._set("initMode", _enum(v, ("RANDOM", "K_MEANS_PARALLEL"), "initMode"))
    def setInitSteps(self, v): return self._set("initSteps", int(v))

    def setDistanceType(self, v):
        return self._set("distanceType", _enum(v, ("EUCLIDEAN", "COSINE"), "distanceType"))
    setMetric = setDistanceType                                   # alias "metric"

    def setInitCentroids(self, C):
        """Not a reference parameter: fixes the initial centroids (Alink's samplers depend on Flink's
        DataSetUtils and are not reproducible; parity runs pass the centroids in)."""
        return self._set("initCentroids", np.asarray(C, dtype=np.float64))

    def setRandomSeed(self, v): return self._set("randomSeed", int(v))


class _PredictParams(_WithParams):
    _defaults = {"predictionCol": None, "predictionDetailCol": None, "predictionDistanceCol": None,
                 "reservedCols": None, "vectorCol": None}

    def setPredictionCol(self, v): return self._set("predictionCol", v)
    def setPredictionDetailCol(self, v): return self._set("predictionDetailCol", v)
    def setPredictionDistanceCol(self, v): return self._set("predictionDistanceCol", v)
    def setReservedCols(self, v): return self._set("reservedCols", list(v) if v is not None else None)


def _enum(v, allowed, name):
    s = str(v).upper()                                            # ParamUtil.searchEnum is case-insensitive
    if s not in allowed:
        raise ValueError("%s must be one of %s, got %r" % (name, allowed, v))
    return s


# ----------------------------------------------------------------------------------------------- iteration
def iterate_icq(partition, C0, max_iter, tol, allreduce=None):
    """KMeansTrainBatchOp.iterateICQ (:58-81) with host buffers, the way the Java ComputeFunctions drive it:
    KMeansAssignCluster -> AllReduce -> KMeansUpdateCentroids -> KMeansIterTermination, ping-pong included.
    `partition` provides assign_accumulate / update_centroids (native.Handle does); `allreduce` sums the
    k0*(d+1) buffer over ranks (identity when None, like one Flink subtask).
    Returns dict(C, k, weights, iters, objective)."""
    C0 = np.ascontiguousarray(C0, dtype=np.float64)
    k0, d = C0.shape
    # KMeansPreallocateCentroid: CENTROID1 = init (label-updated), CENTROID2 = copy, K = k0
    cur, _, _ = partition.update_centroids(_as_sum(C0), k0, k0, step_no=0, prev=None)
    k = k0
    objective = []
    sum_ = np.zeros(k0 * (d + 1))
    iters = 0
    for s in range(1, int(max_iter) + 1):
        sum_, obj, _ = partition.assign_accumulate(cur, k=k)
        if allreduce is not None:
            buf = np.concatenate([sum_, [obj]])
            buf = allreduce(buf)
            sum_, obj = buf[:-1], float(buf[-1])
        objective.append(obj)
        new, k, shift = partition.update_centroids(sum_, k0, k, step_no=s, prev=cur)
        cur = new
        iters = s
        if shift is not None and not np.any(shift >= tol):         # "d >= tol -> continue"; NaN counts as converged
            break
    weights = sum_.reshape(k0, d + 1)[:, d].copy()
    return dict(C=cur, k=k, weights=weights, iters=iters, objective=np.array(objective))


def _as_sum(C):
    """A sum buffer whose update reproduces `updateLabel(C)` on the raw initial centroids: weight 1 per row."""
    k0, d = C.shape
    s = np.ones((k0, d + 1))
    s[:, :d] = C
    return s.ravel()


# ----------------------------------------------------------------------------------------------- operators
class KMeansTrainBatchOp(_TrainParams):
    """new KMeansTrainBatchOp().setVectorCol(..).setK(..).linkFrom(in)"""

    def __init__(self, params=None, device=0, engine=native.ENGINE_AUTO):
        super().__init__(params)
        self.device = device
        self.engine = engine
        self.model_rows = None
        self.train_info = None

    def linkFrom(self, in_table, distributed=None):
        vec_col = self.get("vectorCol")
        if vec_col is None:
            raise ValueError("vectorCol is required")
        col = in_table.col(vec_col)
        csr, X = None, None
        if len(col) > 0 and all(is_sparse_text(v) for v in col):
            # SparseVector column: CSR arrays go to the device as they are (akm_upload_rows_csr)
            csr = rows_to_csr(col)
            n_local, d = len(col), csr[3]
        else:
            X, null_mask = rows_to_matrix(col)
            if null_mask.any():
                raise ValueError("null vector in training data")
            n_local, d = X.shape
        dist = distributed
        n_total = n_local
        if dist is not None:
            n_total = int(dist.all_sum(np.array([float(n_local)]))[0])
        if n_total == 0:
            raise RuntimeError("The train dataset is empty!")            # KMeansTrainBatchOp.java:106
        if dist is not None:
            d_all = int(dist.all_max(np.array([float(d)]))[0])           # vector summary -> vectorSize
            if X is not None and X.shape[1] != d_all:
                X = np.pad(X, ((0, 0), (0, d_all - X.shape[1])))
            d = d_all
        metric = _DIST[self.get("distanceType")]
        with native.Handle(d, metric, device=self.device) as h:
            h.set_engine(self.engine)
            if csr is not None:
                h.upload_rows_csr(csr[0], csr[1], csr[2])
            else:
                h.upload_rows(X)
            if dist is not None:
                dist.attach(h)
            init_info = {}
            C0 = self._initial_centroids(h, d, metric, dist, init_info)
            m = h.train(C0, max_iter=self.get("maxIter"), tol=self.get("epsilon"))
            self.train_info = dict(iters=m["iters"], objective=m["objective"], stats=h.stats(), init=init_info)
        k = m["k"]
        self.model_rows = model_io.save_model(m["C"][:k], m["weights"][:k], k, d, self.get("distanceType"), vec_col)
        return self

    def _initial_centroids(self, h, d, metric, dist, info):
        """KMeansInitCentroids.initKmeansCentroids (:40-66) over the rows cached in `h` (alink_b200/init.py)."""
        C0 = self.get("initCentroids")
        if C0 is not None:
            C0 = np.asarray(C0, dtype=np.float64)
            if C0.ndim != 2 or C0.shape[1] != d:
                raise ValueError("initCentroids must be k x %d" % d)
            info["mode"] = "given"
            return C0
        k, seed, mode = self.get("k"), self.get("randomSeed"), self.get("initMode")
        info["mode"] = mode
        if mode == "RANDOM":
            return init_mod.random_init(h, k, seed, dist)
        return init_mod.kmeans_parallel_init(h, k, self.get("initSteps"), seed, metric, dist, info)

    def getOutputTable(self):
        return Table(self.model_rows, ["model_id", "model_info"])

    def collect(self):
        return list(self.model_rows)


class KMeansPredictBatchOp(_PredictParams):
    """new KMeansPredictBatchOp().setPredictionCol(..).linkFrom(model, data)"""

    def __init__(self, params=None, device=0, engine=native.ENGINE_AUTO):
        super().__init__(params)
        self.device = device
        self.engine = engine
        self.output = None

    def linkFrom(self, model_op, data_table):
        rows = model_op.collect() if hasattr(model_op, "collect") else list(model_op)
        self.output = _predict_table(rows, data_table, self.params, self.device, self.engine)
        return self

    def collect(self):
        return self.output.collect()

    def getOutputTable(self):
        return self.output

    def getColNames(self):
        return self.output.getColNames()


def _predict_table(model_rows, data_table, params, device, engine):
    """KMeansModelMapper: loadModel (:100-105) + map per row (:62-97), micro-batched onto the GPU."""
    md = model_io.load_model(model_rows)
    pred_col = params.get("predictionCol")
    if pred_col is None:
        raise ValueError("predictionCol is required")
    detail_col, dist_col = params.get("predictionDetailCol"), params.get("predictionDistanceCol")
    reserved = params.get("reservedCols")
    in_cols = data_table.getColNames()
    keep = in_cols if reserved is None else list(reserved)
    out_cols = [c for c in keep if c not in (pred_col, detail_col, dist_col)] + [pred_col]
    if detail_col:
        out_cols.append(detail_col)
    if dist_col:
        out_cols.append(dist_col)
    vec_col = md["vectorCol"]
    d, k = md["vectorSize"], md["k"]
    X, null_mask = rows_to_matrix(data_table.col(vec_col), d)
    if X.shape[1] != d:
        raise ValueError("vector size %d does not match the model's vectorSize %d" % (X.shape[1], d))
    n = X.shape[0]
    idx = np.full(n, -1, dtype=np.int64)
    dist = np.full(n, np.nan)
    probs = None
    ok = ~null_mask
    if ok.any():
        with native.Handle(d, _DIST[md["distanceType"]], device=device) as h:
            h.set_engine(engine)
            i2, d2, p2 = h.predict(X[ok], md["centroids"][:k], want_dist=bool(dist_col), want_probs=bool(detail_col))
        idx[ok] = i2
        if d2 is not None:
            dist[ok] = d2
        if p2 is not None:
            probs = np.zeros((n, k))
            probs[ok] = p2
    keep_idx = [in_cols.index(c) for c in out_cols if c in in_cols and c not in (pred_col, detail_col, dist_col)]
    out_rows = []
    for r in range(n):
        row = [data_table.rows[r][i] for i in keep_idx]
        if null_mask[r]:                                              # null vector -> null outputs (:66-73)
            row.append(None)
            if detail_col:
                row.append(None)
            if dist_col:
                row.append(None)
        else:
            row.append(int(idx[r]))                                   # (long) index
            if detail_col:
                vec = np.zeros(k)
                vec[md["clusterIds"][:k]] = probs[r]                  # vec.set(clusterId(i), probs[i]) (:85-88)
                row.append(java_vector_to_string(vec))
            if dist_col:
                row.append(float(dist[r]))
        out_rows.append(tuple(row))
    return Table(out_rows, out_cols)


class KMeansModel(_PredictParams):
    """MapModel(KMeansModelMapper::new): transform(in) (pipeline/MapModel.java:39-44)."""

    def __init__(self, model_rows, params=None, device=0, engine=native.ENGINE_AUTO):
        super().__init__(params)
        self.model_rows = list(model_rows)
        self.device, self.engine = device, engine

    def transform(self, data_table):
        return _predict_table(self.model_rows, data_table, self.params, self.device, self.engine)

    def getModelData(self):
        return Table(self.model_rows, ["model_id", "model_info"])


class KMeans(_TrainParams, _PredictParams):
    """Pipeline stage: fit(in) = new KMeansTrainBatchOp(params).linkFrom(in) -> KMeansModel (KMeans.java:30-32)."""
    _defaults = dict(_TrainParams._defaults, **_PredictParams._defaults)

    def __init__(self, params=None, device=0, engine=native.ENGINE_AUTO):
        _WithParams.__init__(self, params)
        self.device, self.engine = device, engine

    def fit(self, in_table, distributed=None):
        tp = {k: self.params[k] for k in _TrainParams._defaults}
        op = KMeansTrainBatchOp(tp, device=self.device, engine=self.engine).linkFrom(in_table, distributed)
        pp = {k: self.params[k] for k in _PredictParams._defaults}
        mdl = KMeansModel(op.collect(), pp, device=self.device, engine=self.engine)
        mdl.train_info = op.train_info
        return mdl


class Pipeline:
    """new Pipeline().add(stage).fit(data) -> PipelineModel.transform(data) (pipeline/Pipeline.java:113-143)."""

    def __init__(self):
        self.stages = []

    def add(self, stage):
        self.stages.append(stage)
        return self

    def fit(self, table, distributed=None):
        models = []
        cur = table
        for i, st in enumerate(self.stages):
            if hasattr(st, "fit"):
                mdl = st.fit(cur, distributed) if isinstance(st, KMeans) else st.fit(cur)
                models.append(mdl)
                if i + 1 < len(self.stages):
                    cur = mdl.transform(cur)
            else:
                models.append(st)
                if i + 1 < len(self.stages):
                    cur = st.transform(cur)
        return PipelineModel(models)


class PipelineModel:
    def __init__(self, stages):
        self.stages = stages

    def transform(self, table):
        cur = table
        for st in self.stages:
            cur = st.transform(cur)
        return cur


class VectorAssembler(_WithParams):
    """VectorAssemblerMapper (operator/common/dataproc/vector/VectorAssemblerMapper.java:50-86) for numeric
    columns: concatenates the selected columns into one space-separated dense vector string."""
    _defaults = {"selectedCols": None, "outputCol": None}

    def setSelectedCols(self, v): return self._set("selectedCols", list(v))
    def setOutputCol(self, v): return self._set("outputCol", v)

    def transform(self, table):
        cols = self.get("selectedCols")
        idx = [table.getColNames().index(c) for c in cols]
        out = []
        for r in table.rows:
            vals = []
            for i in idx:
                v = r[i]
                if isinstance(v, str):
                    from .vector import parse_vector
                    vals.extend(parse_vector(v).tolist())
                else:
                    vals.append(float(v))
            from .javafmt import java_double_to_string
            out.append(tuple(r) + (" ".join(java_double_to_string(v) for v in vals),))
        return Table(out, table.getColNames() + [self.get("outputCol")])
