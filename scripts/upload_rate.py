"""Host -> device upload rate of akm_upload_rows (pageable numpy doubles / floats), tuning aid."""
import sys, time
import numpy as np
sys.path.insert(0, '.')
from alink_b200 import native
n, d = int(sys.argv[1]) if len(sys.argv) > 1 else 4_000_000, 128
X = np.random.default_rng(0).normal(size=(n, d)).astype(np.float32)
for name, A in (("f64", X.astype(np.float64)), ("f32", X)):
    for rep in range(3):
        with native.Handle(d, native.EUCLIDEAN) as h:
            h.reserve_rows(n)
            t0 = time.perf_counter()
            h.upload_rows(A)
            dt = time.perf_counter() - t0
        print("%s rep %d: %.3f s  %.1f GB/s host bytes" % (name, rep, dt, A.nbytes / dt / 1e9), flush=True)
