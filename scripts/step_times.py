"""Per-step phase timings of the device-resident loop (debugging / tuning aid)."""
import sys, os
import numpy as np
sys.path.insert(0, '.')
from alink_b200 import native
if os.environ.get("AKM_LIB"):
    native.LIB_PATH = os.environ["AKM_LIB"]          # A/B against another build of the library
N = int(sys.argv[1]) if len(sys.argv) > 1 else 20_000_000
d = int(sys.argv[4]) if len(sys.argv) > 4 else 128
metric = native.COSINE if (len(sys.argv) > 5 and sys.argv[5] == "cosine") else native.EUCLIDEAN
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 8
k = int(sys.argv[3]) if len(sys.argv) > 3 else 1000
h = native.Handle(d, metric)
if os.environ.get("AKM_PRUNE"):
    h.set_option("tile_pruning", int(os.environ["AKM_PRUNE"]))
h.reserve_rows(N)
h.blobs_append(20260101, k, 0, N)
C0 = h.blobs_rows(20260101, k, (np.arange(k, dtype=np.int64) * N) // k).astype(np.float64)
h.set_centroids(C0)
h.set_profiling(True)
for s in range(steps):
    h.get_profile(reset=True)
    conv, kk, obj = h.step(0.0)
    p = h.get_profile()
    st = h.stats()
    print("step %d: search %.2f refine %.2f acc %.2f upd %.2f ms | k=%d refined=%.2f%% tiles=%.1f%% obj=%.6g" % (
        s + 1, p["search_ms"], p["refine_ms"], p["accumulate_ms"], p["update_ms"], kk,
        100.0 * st["rows_refined"] / max(1, st["rows"]), 100.0 * st["tiles_computed"] / max(1, st["tiles_total"]), obj), flush=True)
