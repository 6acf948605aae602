"""KMeansPredictBatchOp-equivalent on the cached partition (BASELINE config 5's second half): rows/s of
akm_predict_cached with predictionDistanceCol, host arrays out.  Tuning aid."""
import sys, time
import numpy as np
sys.path.insert(0, '.')
from alink_b200 import native
N = int(sys.argv[1]) if len(sys.argv) > 1 else 20_000_000
d = int(sys.argv[2]) if len(sys.argv) > 2 else 256
k = int(sys.argv[3]) if len(sys.argv) > 3 else 500
metric = native.COSINE if (len(sys.argv) > 4 and sys.argv[4] == "cosine") else native.EUCLIDEAN
h = native.Handle(d, metric)
h.reserve_rows(N)
h.blobs_append(20260101, k, 0, N)
C0 = h.blobs_rows(20260101, k, (np.arange(k, dtype=np.int64) * N) // k).astype(np.float64)
m = h.train(C0, max_iter=3, tol=0.0)
C = m["C"][: m["k"]]
for want_dist in (False, True):
    for rep in range(3):
        t0 = time.perf_counter()
        idx, dist = h.predict_cached(C, want_dist=want_dist)
        dt = time.perf_counter() - t0
        print("predict_cached dist=%s rep %d: %.1f ms  %.3g rows/s" % (want_dist, rep, dt * 1e3, N / dt), flush=True)
