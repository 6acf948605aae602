import sys, time
import numpy as np
sys.path.insert(0, '.')
from alink_b200 import native
import oracle as orc

def run(n, d, k, kc, seed, metric=native.EUCLIDEAN):
    rng = np.random.default_rng(seed)
    centers = rng.uniform(-10, 10, size=(kc, d))
    X = (centers[rng.integers(0, kc, size=n)] + rng.normal(size=(n, d))).astype(np.float32).astype(np.float64)
    C = X[rng.choice(n, k, replace=False)].copy()
    Xp, xl = orc.prepare_rows(metric, X)
    Cp, cl = orc.prepare_centroids(metric, C)
    t0 = time.time()
    s_ref, l_ref, _ = orc.assign_accumulate(metric, Xp, xl, Cp, cl, threads=16)
    t1 = time.time()
    with native.Handle(d, metric) as h:
        h.set_engine(native.ENGINE_TENSOR)
        h.upload_rows(X)
        s, obj, labels = h.assign_accumulate(Cp, want_labels=True)
        st = h.stats()
    bad = int((labels != l_ref).sum())
    print("n=%d d=%d k=%d metric=%d: mismatches=%d certified=%d refined=%d engine=%d (oracle %.1fs)" % (
        n, d, k, metric, bad, st["rows_certified"], st["rows_refined"], st["engine_used"], t1 - t0), flush=True)
    return bad

bad = 0
bad += run(5000, 32, 16, 16, 1)
bad += run(300, 64, 100, 50, 2)
bad += run(20000, 128, 1000, 1000, 3)
bad += run(20000, 128, 1000, 300, 4)
bad += run(3000, 100, 257, 60, 5, native.COSINE)
bad += run(70000, 128, 300, 100, 6)
print("TOTAL MISMATCHES", bad)
