"""profiles/traffic.json from `ncu --set full ... --page raw --csv` exports: DRAM bytes per launch of the hot kernels,
keyed by the hash of the CUDA sources they were captured from (bench.py fills `roofline.traffic` only when that hash
matches the build it runs).
usage: python scripts/ncu_traffic.py profiles/r2_ncu_full_cfg4_raw.csv:100000000:128:1000 [more csv:rows:d:k ...]
Every capture is one bench.py run of the shape given after the file name; per kernel name the launches are averaged
(the search kernel is listed separately for passes with and without tile pruning, told apart by their duration)."""
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import source_hash  # noqa: E402

UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}


def main():
    out = {"source_hash": source_hash(), "kernels": []}
    for arg in sys.argv[1:]:
        path, rows, d, k = arg.split(":")
        rows, d, k = int(rows), int(d), int(k)
        with open(path) as f:
            table = list(csv.reader(f))
        hdr, units = table[0], table[1]
        col = {name: i for i, name in enumerate(hdr)}
        per = {}
        for r in table[2:]:
            name = r[col["Kernel Name"]].split("(")[0].replace("void ", "")
            rd = float(r[col["dram__bytes_read.sum"]]) * UNIT[units[col["dram__bytes_read.sum"]]]
            wr = float(r[col["dram__bytes_write.sum"]]) * UNIT[units[col["dram__bytes_write.sum"]]]
            ms = float(r[col["gpu__time_duration.sum"]]) * {"ms": 1.0, "us": 1e-3, "s": 1e3, "ns": 1e-6}[units[col["gpu__time_duration.sum"]]]
            tens = float(r[col["sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"]])
            per.setdefault(name, []).append((rd + wr, ms, tens))
        for name, v in sorted(per.items()):
            out["kernels"].append({"kernel": name, "rows": rows, "d": d, "k": k, "launches": len(v),
                                   "dram_bytes": sum(x[0] for x in v) / len(v), "ms": sum(x[1] for x in v) / len(v),
                                   "tensor_pipe_active_pct": sum(x[2] for x in v) / len(v),
                                   "file": os.path.relpath(path, ROOT)})
    with open(os.path.join(ROOT, "profiles", "traffic.json"), "w") as f:
        json.dump(out, f, indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
