"""KMeans model rows in Alink's wire format, so GPU-trained models load in stock Alink and vice versa.

Model table = (model_id BIGINT, model_info STRING) (reference: common/model/SimpleModelDataConverter.java:36-59):
  row 0                         params JSON: every value is itself JSON-encoded ("k":"2", "vectorCol":"\\"Y\\"")
                                (KMeansTrainModelData.ParamSummary.toParams, KMeansTrainModelData.java:104-112)
  rows 1048576*(i+1) + slice    one ClusterSummary JSON per centroid {"clusterId","weight","vec":{"data":[...]}}
                                (KMeansModelDataConverter.java:20-32; id = 1048576*stringIndex + sliceIndex,
                                ModelConverterUtils.java:24,261-281; strings longer than 32 KiB are sliced)
Legacy centroid formats (KMeansUtil.java:282-310) are accepted on load.
"""
import json

import numpy as np

from .javafmt import java_double_to_string

MAX_NUM_SLICES = 1024 * 1024
SEGMENT_SIZE = 32 * 1024


def _jstr(s):
    return "null" if s is None else json.dumps(json.dumps(s))


def save_model(centroids, weights, k, vector_size, distance_type, vector_col):
    """KMeansOutputModel.calc's serialisation. Returns a list of (model_id, model_info)."""
    meta = ('{"vectorCol":%s,"latitudeCol":null,"longitudeCol":null,"distanceType":%s,"k":"%d","vectorSize":"%d"}'
            % (_jstr(vector_col), _jstr(distance_type), k, vector_size))
    strings = [meta]
    for i in range(k):
        data = ",".join(java_double_to_string(v) for v in centroids[i])
        strings.append('{"clusterId":%d,"weight":%s,"vec":{"data":[%s]}}' % (i, java_double_to_string(weights[i]), data))
    rows = []
    for si, st in enumerate(strings):
        nslice = max(1, (len(st) + SEGMENT_SIZE - 1) // SEGMENT_SIZE)
        for j in range(nslice):
            rows.append((MAX_NUM_SLICES * si + j, st[j * SEGMENT_SIZE:(j + 1) * SEGMENT_SIZE]))
    return rows


def load_model(rows):
    """KMeansModelDataConverter.load. Returns dict(k, vectorSize, distanceType, vectorCol, centroids[k,d],
    clusterIds[k], weights[k])."""
    pieces = {}
    for mid, info in sorted(((int(r[0]), r[1]) for r in rows), key=lambda t: t[0]):
        pieces.setdefault(mid // MAX_NUM_SLICES, []).append(info)
    strings = {si: "".join(v) for si, v in pieces.items()}
    if 0 not in strings:
        raise ValueError("model has no meta row")
    meta_raw = json.loads(strings[0])
    meta = {}
    for key, v in meta_raw.items():
        meta[key] = None if v is None else json.loads(v)
    if "metric" in meta and "distanceType" not in meta:
        meta["distanceType"] = meta["metric"]                  # alias (HasKMeansDistanceType.java:19-24)
    cents, ids, weights = [], [], []
    for si in sorted(k for k in strings if k != 0):
        obj = json.loads(strings[si])
        if "vec" in obj and obj["vec"] is not None:
            vec = obj["vec"]["data"] if isinstance(obj["vec"], dict) else obj["vec"]
        else:                                                  # OldClusterSummary: "center" is itself a JSON string
            c = obj["center"]
            c = json.loads(c) if isinstance(c, str) else c
            vec = c["data"] if isinstance(c, dict) else c
        cents.append(np.asarray(vec, dtype=np.float64))
        ids.append(int(obj["clusterId"]))
        weights.append(float(obj["weight"]))
    k = int(meta["k"])
    return dict(k=k, vectorSize=int(meta["vectorSize"]), distanceType=str(meta["distanceType"]).upper(),
                vectorCol=meta.get("vectorCol"), centroids=np.vstack(cents) if cents else np.zeros((0, 0)),
                clusterIds=np.array(ids, dtype=np.int64), weights=np.array(weights, dtype=np.float64))
