"""EvalClusterBatchOp: clustering metrics of a prediction table.

Mirror of operator/batch/evaluation/EvalClusterBatchOp.java:57-139 and its helpers
(operator/common/evaluation/ClusterEvaluationUtil.java, ClusterMetricsSummary.java, ClusterMetrics.java):
  * vector metrics (needs vectorCol): the O(N d) and O(N k d) passes -- per-cluster mean / compactness / SSW
    and the silhouette sum -- run on the GPU behind akm_eval_clusters; the k x k remainder (SSB, separation,
    Davies-Bouldin, Calinski-Harabasz, ClusterMetricsSummary.toMetrics :109-150) is numpy on the host;
  * label metrics (needs labelCol): purity, NMI, RI, ARI from the contingency table
    (ClusterEvaluationUtil.extractParamsFromConfusionMatrix :41-88), host only -- it never touches the vectors;
  * neither: cluster ids and counts (getBasicClusterStatistics :147-169).
Rows with a null in any selected field are skipped (EvaluationUtil.checkRowFieldNotNull).
"""
import json
import math

import numpy as np

from . import native
from .kmeans import _DIST, _WithParams, _enum
from .vector import rows_to_matrix


def _pair_dist(metric, a, b):
    """ContinuousDistance.calc(Vector, Vector): EuclideanDistance.java:45-47 / CosineDistance.java:46-50."""
    if metric == native.EUCLIDEAN:
        return float(np.sqrt(((a - b) ** 2).sum()))
    cross = float(np.sqrt((a * a).sum()) * np.sqrt((b * b).sum()))
    return 1.0 - (float((a * b).sum()) / cross if cross > 0.0 else 0.0)


def summary_to_metrics(metric, ids, count, mean, dist_sum, dist_sq_sum, silhouette_sum):
    """ClusterMetricsSummary.toMetrics (:109-150) + SaveDataAsParams (:220-232) on the non-empty clusters."""
    keep = count > 0
    ids = [ids[i] for i in np.nonzero(keep)[0]]
    count, mean, dist_sum, dist_sq_sum = count[keep], mean[keep], dist_sum[keep], dist_sq_sum[keep]
    k = int(count.size)
    total = int(count.sum())
    compact = dist_sum / count
    mean_all = (mean * count[:, None]).sum(axis=0) * (1.0 / total)
    ssb = float(sum(_pair_dist(metric, mean[i], mean_all) ** 2 * count[i] for i in range(k)))
    ssw = float(dist_sq_sum.sum())
    db = np.zeros(k)
    sep = 0.0
    if k > 1:
        iu = np.triu_indices(k, 1)
        if metric == native.EUCLIDEAN:
            D = np.zeros((k, k))
            step = max(1, (1 << 22) // max(1, k * mean.shape[1]))        # ~32 MB of differences at a time
            for i0 in range(0, k, step):
                diff = mean[i0:i0 + step, None, :] - mean[None, :, :]
                D[i0:i0 + step] = np.sqrt((diff * diff).sum(axis=2))
        else:
            nrm = np.sqrt((mean * mean).sum(axis=1))
            cross = nrm[:, None] * nrm[None, :]
            D = 1.0 - np.divide(mean @ mean.T, cross, out=np.zeros((k, k)), where=cross > 0)
        d_pairs = D[iu]
        sep = float(d_pairs.sum())
        with np.errstate(divide="ignore", invalid="ignore"):
            ratio = (compact[iu[0]] + compact[iu[1]]) / d_pairs
        np.maximum.at(db, iu[0], ratio)
        np.maximum.at(db, iu[1], ratio)
    with np.errstate(divide="ignore", invalid="ignore"):
        out = {
            "SSB": ssb, "SSW": ssw, "compactness": float(compact.sum() / k), "k": k, "count": total,
            "seperation": float(np.float64(2.0 * sep) / np.float64(k * k - k)),
            "daviesBouldin": float(db.sum() / k),
            "calinskiHarabaz": float(np.float64(ssb) * (total - k) / np.float64(ssw) / np.float64(k - 1)),
            "clusterArray": [str(i) for i in ids], "countArray": [float(c) for c in count],
            "silhouetteCoefficient": float(silhouette_sum / total),
        }
    return out


def label_metrics(actual, predicted):
    """ClusterEvaluationUtil.extractParamsFromConfusionMatrix (:41-88): rows = predictions, columns = labels."""
    a_ids = {v: i for i, v in enumerate(dict.fromkeys(actual))}
    p_ids = {v: i for i, v in enumerate(dict.fromkeys(predicted))}
    m = np.zeros((len(p_ids), len(a_ids)), dtype=np.int64)
    for a, p in zip(actual, predicted):
        m[p_ids[p], a_ids[a]] += 1
    act, pred, total = m.sum(axis=0), m.sum(axis=1), int(m.sum())

    def comb(x):
        return x * (x - 1) // 2

    def ent(f):
        r = 1.0 * f / total
        return 0.0 if f == 0 else r * math.log(r)

    e_act = sum(ent(int(f)) for f in act) / -math.log(2)
    e_pred = sum(ent(int(f)) for f in pred) / -math.log(2)
    tp_fp = sum(comb(int(f)) for f in act)
    tp_fn = sum(comb(int(f)) for f in pred)
    mi, purity, tp = 0.0, 0.0, 0
    for i in range(m.shape[0]):
        for j in range(m.shape[1]):
            v = int(m[i, j])
            if v:
                mi += 1.0 * v / total * math.log(1.0 * total * v / int(pred[i]) / int(act[j]))
            tp += comb(v)
        purity += int(m[i].max())
    purity /= total
    mi /= math.log(2)
    fp, fn = tp_fp - tp, tp_fn - tp
    tot_c = comb(total)
    tn = tot_c - tp - fn - fp

    def div(a, b):                                   # Java double division: x/0 -> Infinity / NaN, no exception
        return float(np.float64(a) / np.float64(b))

    with np.errstate(divide="ignore", invalid="ignore"):
        expected = div(1.0 * tp_fp * tp_fn, tot_c)
        max_index = 1.0 * (tp_fp + tp_fn) / 2
        return {"NMI": div(2.0 * mi, e_act + e_pred), "purity": purity,
                "RI": div(1.0 * (tp + tn), tp + tn + fp + fn), "ARI": div(tp - expected, max_index - expected)}


class ClusterMetrics:
    """Getter names of operator/common/evaluation/ClusterMetrics.java."""

    def __init__(self, params):
        self.params = params

    def _g(self, key):
        return self.params.get(key)

    def getSsb(self): return self._g("SSB")
    def getSsw(self): return self._g("SSW")
    def getCompactness(self): return self._g("compactness")
    def getSeperation(self): return self._g("seperation")
    def getDaviesBouldin(self): return self._g("daviesBouldin")
    def getCalinskiHarabaz(self): return self._g("calinskiHarabaz")
    def getSilhouetteCoefficient(self): return self._g("silhouetteCoefficient")
    def getK(self): return self._g("k")
    def getCount(self): return self._g("count")
    def getClusterArray(self): return self._g("clusterArray")
    def getCountArray(self): return self._g("countArray")
    def getPurity(self): return self._g("purity")
    def getNmi(self): return self._g("NMI")
    def getRi(self): return self._g("RI")
    def getAri(self): return self._g("ARI")

    def serialize(self):
        return json.dumps(self.params)


class EvalClusterBatchOp(_WithParams):
    """new EvalClusterBatchOp().setPredictionCol(..).setVectorCol(..).setLabelCol(..).linkFrom(in).collectMetrics()"""
    _defaults = {"labelCol": None, "predictionCol": None, "vectorCol": None, "distanceType": "EUCLIDEAN"}

    def __init__(self, params=None, device=0, engine=native.ENGINE_AUTO):
        super().__init__(params)
        self.device = device
        self.engine = engine
        self.metrics = None

    def setLabelCol(self, v): return self._set("labelCol", v)
    def setPredictionCol(self, v): return self._set("predictionCol", v)
    def setVectorCol(self, v): return self._set("vectorCol", v)
    def setDistanceType(self, v): return self._set("distanceType", _enum(v, ("EUCLIDEAN", "COSINE"), "distanceType"))

    def linkFrom(self, in_table):
        pred_col = self.get("predictionCol")
        if pred_col is None:
            raise ValueError("predictionCol is required")
        preds = in_table.col(pred_col)
        params = {}
        label_col, vec_col = self.get("labelCol"), self.get("vectorCol")
        if label_col is not None:
            labels = in_table.col(label_col)
            ok = [i for i in range(len(preds)) if preds[i] is not None and labels[i] is not None]
            params.update(label_metrics([str(labels[i]) for i in ok], [str(preds[i]) for i in ok]))
        if vec_col is not None:
            X, null_mask = rows_to_matrix(in_table.col(vec_col))
            ok = np.array([preds[i] is not None and not null_mask[i] for i in range(len(preds))], dtype=bool)
            ids = list(dict.fromkeys(str(preds[i]) for i in np.nonzero(ok)[0]))
            dense = {v: i for i, v in enumerate(ids)}
            lab = np.array([dense[str(preds[i])] if ok[i] else -1 for i in range(len(preds))], dtype=np.int32)
            if not ids:
                raise RuntimeError("no row with a prediction and a vector")
            if X.shape[1] < 1:
                raise ValueError("Vector Size must be at least 1!")          # ClusterEvaluationUtil.java:186
            metric = _DIST[self.get("distanceType")]
            with native.Handle(X.shape[1], metric, device=self.device) as h:
                h.set_engine(self.engine)
                h.upload_rows(X)
                r = h.eval_clusters(lab, len(ids))
            params.update(summary_to_metrics(metric, ids, r["count"], r["mean"], r["dist_sum"], r["dist_sq_sum"],
                                             r["silhouette_sum"]))
        else:
            ok = [p for p in preds if p is not None]
            cnt = {}
            for p in ok:
                cnt[str(p)] = cnt.get(str(p), 0.0) + 1.0
            params.update({"count": len(ok), "k": len(cnt), "clusterArray": list(cnt), "countArray": list(cnt.values())})
        self.metrics = ClusterMetrics(params)
        return self

    def collectMetrics(self):
        return self.metrics

    def collect(self):
        return [(self.metrics.serialize(),)]
