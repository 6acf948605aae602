"""Workload for the `ncu --set full` captures under profiles/: one BASELINE shape brought to its steady state
untimed, then -- between cudaProfilerStart/Stop, so run ncu with `--profile-from-start off` -- two steady supersteps
(bounds and pruning on) and two supersteps with every tile evaluated and every row re-accumulated.

  ncu --set full --clock-control none --import-source on --profile-from-start off \
      -k regex:'search_pair|rowprep|group_list|refine_cand|accumulate' -o /tmp/cfg4 \
      python scripts/ncu_capture.py 100000000 128 1000 euclidean
  ncu -i /tmp/cfg4.ncu-rep --page raw --csv > gpurun_out/r2_ncu_full_cfg4_raw.csv
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from alink_b200 import native  # noqa: E402

SEED = 20260101


def main():
    n, d, k = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
    metric = native.COSINE if (len(sys.argv) > 4 and sys.argv[4].lower().startswith("cos")) else native.EUCLIDEAN
    torch.cuda.init()
    with native.Handle(d, metric) as h:
        h.reserve_rows(n)
        h.blobs_append(SEED, k, 0, n)
        C0 = h.blobs_rows(SEED, k, (np.arange(k, dtype=np.int64) * n) // k).astype(np.float64)
        h.set_centroids(C0)
        for _ in range(14):                      # bench.py's from_init (10) + warm-up + first timed steps
            h.step(0.0)
        torch.cuda.synchronize()
        torch.cuda.profiler.start()
        for _ in range(2):
            h.step(0.0)
        h.set_option("tile_pruning", 0)
        h.set_option("delta_accumulate", 0)
        for _ in range(2):
            h.step(0.0)
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
        st = h.stats()
        print("k", h.get_model()["k"], "tiles", st["tiles_computed"], "/", st["tiles_total"])


if __name__ == "__main__":
    main()
