"""Timing experiments on the search kernel: dense (PRUNE=0) or pruned passes over ROWS x D rows, K centroids (METRIC=euclidean|cosine), with the
epilogue switched off in stages (AKM_EPI_DBG = option "epi_dbg": 1 = accumulators read but not evaluated, 2 = not even
read -- the MMA/TMA floor --, 3 = evaluated but not merged; results are invalid in those modes; they exist only in a
library built with `make -C alink_b200/csrc clean all EXTRA=-DAKM_EPI_DBG`)."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from alink_b200 import native
n, d, k = int(float(os.environ.get("ROWS", 20e6))), int(os.environ.get("D", 128)), int(os.environ.get("K", 1000))
metric = native.COSINE if os.environ.get("METRIC", "euclidean") == "cosine" else native.EUCLIDEAN
with native.Handle(d, metric) as h:
    h.reserve_rows(n)
    h.blobs_append(20260101, k, 0, n)
    C0 = h.blobs_rows(20260101, k, (np.arange(k, dtype=np.int64) * n) // k).astype(np.float64)
    h.set_option("tile_pruning", int(os.environ.get("PRUNE", "0")))
    h.set_option("delta_accumulate", 1)
    h.set_centroids(C0)
    h.set_profiling(True)
    for _ in range(6):
        h.step(0.0)
    h.get_profile(reset=True)
    h.set_option("epi_dbg", int(os.environ.get("AKM_EPI_DBG", "0")))
    for _ in range(5):
        h.step(0.0)
    p = h.get_profile()
    st = h.stats()
    print("prune=%s dbg=%s rows=%d search %.3f ms refine %.3f acc %.3f (tiles %d/%d)" % (os.environ.get("PRUNE", "0"), os.environ.get("AKM_EPI_DBG", "0"), n,
          p["search_ms"] / p["steps"], p["refine_ms"] / p["steps"], p["accumulate_ms"] / p["steps"], st["tiles_computed"], st["tiles_total"]))
