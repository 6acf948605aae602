"""Diagnostic: how far do the centroids still move in the steady state of the bench workload?
(decides how tight the per-row second-nearest bound can stay)   python scripts/shift_probe.py [rows]"""
import sys
import numpy as np
sys.path.insert(0, ".")
from alink_b200 import native

n = int(sys.argv[1]) if len(sys.argv) > 1 else 20_000_000
d, k, seed = 128, 1000, 20260101
with native.Handle(d, native.EUCLIDEAN) as h:
    h.reserve_rows(n)
    h.blobs_append(seed, k, 0, n)
    C = h.blobs_rows(seed, k, (np.arange(k, dtype=np.int64) * n) // k).astype(np.float64)
    kk = k
    for it in range(16):
        r = h.superstep(C, k=kk, step_no=it + 1, tol=0.0, want_shift=True)
        C, kk = r["C"], r["k"]
        sh = np.sort(np.asarray(r["shift"][:kk]))
        st = h.stats()
        print(it, "k", kk, "shift max %.4f p99 %.4f p90 %.4f median %.5f" % (sh[-1], sh[int(.99 * kk)], sh[int(.9 * kk)], sh[kk // 2]),
              "refined", st["rows_refined"], "tiles", round(st["tiles_computed"] / max(1, st["tiles_total"]), 3), "moved", st["rows_moved"])
