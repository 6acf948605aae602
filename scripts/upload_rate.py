"""Host -> device ingest rate of akm_upload_rows (pageable doubles in, prepareVectorData done): GB/s of host bytes.
usage: AKM_UPLOAD_THREADS=8 python scripts/upload_rate.py [GB] [d]"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from alink_b200 import native

gb = float(sys.argv[1]) if len(sys.argv) > 1 else 4.0
d = int(sys.argv[2]) if len(sys.argv) > 2 else 128
rng = np.random.default_rng(0)
blk = rng.normal(size=(65536, d)).astype(np.float32).astype(np.float64)
reps = max(1, int(gb * 1e9 / blk.nbytes))
X = np.empty((reps * blk.shape[0], d))
for i in range(reps):
    X[i * blk.shape[0]:(i + 1) * blk.shape[0]] = blk
for kind in ("fp32-representable", "general doubles"):
    if kind == "general doubles":
        X += 1e-9
    with native.Handle(d, native.EUCLIDEAN) as h:
        h.reserve_rows(X.shape[0])
        h.upload_rows(X[:1000]); h.upload_wait(); h.clear_rows(); h.reserve_rows(X.shape[0])
        t0 = time.perf_counter()
        h.upload_rows(X)
        t1 = time.perf_counter()
        h.upload_wait()
        t2 = time.perf_counter()
    print("threads=%s %s: %.2f GB in %.3f s (returned after %.3f s) = %.1f GB/s" % (
        os.environ.get("AKM_UPLOAD_THREADS", "default"), kind, X.nbytes / 1e9, t2 - t0, t1 - t0, X.nbytes / (t2 - t0) / 1e9))
