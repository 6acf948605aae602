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
# d > 128: K-streaming pair kernel
bad += run(20000, 256, 500, 500, 7, native.COSINE)
bad += run(20000, 256, 500, 200, 8)
bad += run(10000, 784, 10, 10, 9)
bad += run(5000, 200, 300, 80, 10)
bad += run(3000, 1000, 70, 70, 14)
bad += run(40000, 136, 2000, 2000, 15)
print("TOTAL MISMATCHES", bad)

# multi-pass training: exercises the seeded thresholds (chain of akm_step passes)
def train_check(n, d, k, kc, seed, metric=native.EUCLIDEAN, iters=8):
    rng = np.random.default_rng(seed)
    centers = rng.uniform(-10, 10, size=(kc, d))
    X = (centers[rng.integers(0, kc, size=n)] + rng.normal(size=(n, d))).astype(np.float32).astype(np.float64)
    C0 = X[rng.choice(n, k, replace=False)].copy()
    ref = orc.train(metric, X, C0, max_iter=iters, tol=0.0, threads=16)
    with native.Handle(d, metric) as h:
        h.set_engine(native.ENGINE_TENSOR)
        h.upload_rows(X)
        m = h.train(C0, max_iter=iters, tol=0.0)
        labels = h.get_labels()
        st = h.stats()
    bad = int((labels != ref["labels"]).sum())
    cerr = float(np.abs(m["C"][:m["k"]] - ref["C"][:ref["k"]]).max()) if m["k"] == ref["k"] else -1
    print("train n=%d d=%d k=%d metric=%d: mismatches=%d k=%d/%d max|dC|=%.3g certified=%d refined=%d" % (
        n, d, k, metric, bad, m["k"], ref["k"], cerr, st["rows_certified"], st["rows_refined"]), flush=True)
    return bad

bad = 0
bad += train_check(30000, 128, 1000, 300, 11)
bad += train_check(30000, 64, 100, 100, 12)
bad += train_check(20000, 100, 200, 50, 13, native.COSINE)
bad += train_check(20000, 256, 500, 200, 16, native.COSINE)
bad += train_check(20000, 784, 10, 10, 17)
bad += train_check(20000, 320, 600, 300, 18)
print("TRAIN MISMATCHES", bad)
