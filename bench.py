#!/usr/bin/env python
"""bench.py -- KMeans Lloyd-iteration throughput (sample-iters/s) of the B200 path, per BASELINE.json.

Workload (config.workload): synthetic Gaussian blobs 100M x 128, k = 1000, EUCLIDEAN -- the
configuration the north_star's target is quoted on; it fits one B200 (51.2 GB fp32 + 25.6 GB fp16
search copy).  With --gpus N the same 100M rows are row-sharded over N ranks by Alink's
DefaultDistributedInfo formula (strong scaling) and the k x (d+1) accumulator is all-reduced with NCCL.

One "step" = one Lloyd superstep of KMeansTrainBatchOp.iterateICQ over the cached partition:
KMeansAssignCluster -> AllReduce -> KMeansUpdateCentroids -> KMeansIterTermination.
  value     : device-resident loop (akm_step) in steady state (after the from_init run and W more warm-up steps),
              CUDA events on the launching stream, max over ranks.
  from_init : SURVEY 8(d)'s throughput run -- akm_set_centroids(C0) then maxIter = 10 steps with epsilon = 0,
              timed from step 1 (the two cold passes and the regrouping included), per-step times listed.
  e2e       : the same superstep through the reference-facing host-buffer call (akm_superstep: centroids
              host->device, new centroids device->host every step, one synchronisation), wall clock, max over
              ranks; `e2e.fine_grained` = the three separate calls a host makes when it keeps AllReduce and
              KMeansUpdateCentroids in Java.  (The sample matrix is cached on the GPU at superstep 1 exactly like
              Alink caches it on the JVM heap, BaseComQueue.java:497-527; its upload is reported as `upload`.)
  parity    : the CPU leg's rows (the first n_cpu rows of the same data) evaluated from C0 by the oracle and by the
              GPU: labels, k x (d+1) sums and objective compared in the line.
  configs   : BASELINE configs 2, 3 and 5 (and config 5's KMeansPredictBatchOp pass) run time-boxed after the headline,
              each with its own roofline object (HBM-bound shapes against the HBM roof).
--impl reference times the CPU restatement of Alink's Java arithmetic (oracle/, all host threads, the
duplicated distance evaluation of KMeansUtil.java:73,124 included) -- the JVM itself cannot run here.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SEED = 20260101
METRIC = "KMeans sample-iters/sec"


def shard(task, par, total):
    """DefaultDistributedInfo.startPos / localRowCnt (reference: common/io/directreader/DefaultDistributedInfo.java:10-35)."""
    div, mod = divmod(total, par)
    if mod == 0:
        return div * task, div
    if task >= mod:
        return div * task + mod, div
    return div * task + task, div + 1


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            j = json.load(f)
        return dict(hbm=j["hbm_gbs"], tf_burst=j["bf16_tflops"], tf_sust=j.get("bf16_tflops_sustained", j["bf16_tflops"]),
                    source="measured (MEASURED_PEAKS.json)")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sust=1400.0, source="fallback (B200_PROFILING.md)")


class ClockSampler:
    """SM clock, power and throttle reasons sampled DURING the timed region (profiling recipe's clocks line).
    NVML is polled from a thread every ~2 ms (the timed region of a multi-GPU run is a few tens of ms, shorter
    than one `nvidia-smi -lms` period); nvidia-smi is the fallback when the NVML binding is missing."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.proc = None
        self.lines = []
        self.samples = []          # (sm_mhz, power_w, reasons_bitmask)
        self.stop_flag = False
        self.thread = None
        self.nvml = None
        self.sm_max = None

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nvml = pynvml
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[self.idx]) if vis and vis.split(",")[self.idx].isdigit() else self.idx
            self.dev = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.sm_max = float(pynvml.nvmlDeviceGetMaxClockInfo(self.dev, pynvml.NVML_CLOCK_SM))
            self.thread = threading.Thread(target=self._poll, daemon=True)
            self.thread.start()
            return
        except Exception:
            self.nvml = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _poll(self):
        n = self.nvml
        while not self.stop_flag:
            try:
                sm = float(n.nvmlDeviceGetClockInfo(self.dev, n.NVML_CLOCK_SM))
                pw = n.nvmlDeviceGetPowerUsage(self.dev) / 1000.0
                try:
                    rs = int(n.nvmlDeviceGetCurrentClocksEventReasons(self.dev))
                except Exception:
                    rs = int(n.nvmlDeviceGetCurrentClocksThrottleReasons(self.dev))
                self.samples.append((sm, pw, rs))
            except Exception:
                pass
            time.sleep(0.002)

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.nvml is not None:
            self.stop_flag = True
            self.thread.join(timeout=2)
            if not self.samples:
                return {"sm_mhz": None, "sm_max_mhz": self.sm_max, "reasons": ["no samples"]}
            n = self.nvml
            names = (("hw_slowdown", getattr(n, "nvmlClocksThrottleReasonHwSlowdown", 0x8)),
                     ("hw_thermal_slowdown", getattr(n, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40)),
                     ("sw_thermal_slowdown", getattr(n, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20)),
                     ("sw_power_cap", getattr(n, "nvmlClocksThrottleReasonSwPowerCap", 0x4)))
            reasons = sorted({nm for _, _, rs in self.samples for nm, bit in names if rs & bit})
            return {"sm_mhz": float(np.median([s[0] for s in self.samples])), "sm_max_mhz": self.sm_max,
                    "power_w_max": float(max(s[1] for s in self.samples)), "samples": len(self.samples),
                    "reasons": reasons, "source": "NVML, 2 ms period, timed region only"}
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, pw = [], [], set(), []
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); pw.append(float(f[3]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "power_w_max": float(max(pw)),
                "samples": len(sm), "reasons": sorted(reasons), "source": "nvidia-smi -lms 100"}


def cpu_rows(n, d, k, seed=SEED):
    """Gaussian blobs with the bench recipe (centres U(-10,10)^d, sigma 1, fp32-rounded) from numpy;
    used by --impl reference, which must not touch the GPU library."""
    rng = np.random.default_rng(seed)
    centres = rng.uniform(-10, 10, size=(k, d))
    blob = rng.integers(0, k, size=n)
    X = centres[blob] + rng.normal(size=(n, d))
    return X.astype(np.float32).astype(np.float64)


def cpu_lloyd_rate(X, C0, threads, steps, warmup, metric=0, keep_first=False):
    """sample-iters/s of the Java-order CPU restatement: assign (distances evaluated twice, like
    KMeansUtil.updateSumMatrix) + all-reduce in rank order + update, `threads` Flink subtasks."""
    import oracle as orc
    n, d = X.shape
    k0 = C0.shape[0]
    Xp, xl = orc.prepare_rows(metric, X)
    C, cl = orc.prepare_centroids(metric, C0)
    k = k0
    times = []
    first = None
    for s in range(warmup + steps):
        t0 = time.perf_counter()
        sm, labels, mind = orc.assign_accumulate(metric, Xp, xl, C, cl, k=k, dup=True, threads=threads)
        if s == 0 and keep_first:
            first = dict(sum=sm.copy(), labels=labels.copy(),
                         objective=float((mind ** 2).sum()) if metric == 0 else float(mind.sum()))
        C, cl, k = orc.update_centroids(metric, sm, k, k0, d)
        t1 = time.perf_counter()
        if s >= warmup:
            times.append(t1 - t0)
    total = sum(times)
    return n * len(times) / total, total / len(times), first


def size_cpu_sample(d, k, threads, target_s, metric=0):
    """Rows per CPU step so that one step costs about target_s seconds (probe, then scale)."""
    probe = max(threads * 200, 1000)
    X = cpu_rows(probe, d, k, seed=1)
    rate, _, _ = cpu_lloyd_rate(X, X[:k].copy() if probe >= k else cpu_rows(k, d, k, 2), threads, 1, 0, metric)
    return int(max(probe, min(rate * target_s, 4_000_000)))


def probe_jvm():
    """the Java path is timed only if a JVM and an Alink jar exist on this box."""
    out = {"java": None, "alink_jar": os.environ.get("ALINK_JAR") or None}
    try:
        r = subprocess.run(["java", "-version"], capture_output=True, text=True, timeout=20)
        out["java"] = (r.stderr or r.stdout).strip().splitlines()[0] if r.returncode == 0 else None
    except Exception as e:                                   # FileNotFoundError on this image
        out["java_probe"] = "%s: %s" % (type(e).__name__, e)
    jar_ok = bool(out["alink_jar"]) and os.path.exists(out["alink_jar"])
    out["runnable"] = bool(out["java"]) and jar_ok
    out["verdict"] = ("JVM + Alink jar present" if out["runnable"] else
                      "Alink's Java path is not runnable on this box (%s%s)" % (
                          "no java on PATH" if not out["java"] else "java present",
                          "" if jar_ok else ", no $ALINK_JAR"))
    return out


def run_reference(args, rank, world, emit):
    """Reference arm: Alink's own CPU implementation of the path = the oracle port (no JVM on this image)."""
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    n_s = args.cpu_rows or size_cpu_sample(args.d, args.k, threads, 4.0, args.metric)
    X = cpu_rows(n_s, args.d, args.k)
    C0 = X[(np.arange(args.k) * n_s) // args.k].copy()
    rate, sec, _ = cpu_lloyd_rate(X, C0, threads, args.steps, args.warmup, args.metric)
    jvm = probe_jvm()
    sample = "oracle port of Alink's Java order (2 dgemv/sample, no FMA), %d threads = Flink subtasks, %d of %d rows per step" % (
        threads, n_s, args.rows)
    out = {
        "impl": "reference", "metric": METRIC, "value": rate, "unit": "sample-iters/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args, 1),
        "cpu_baseline": {"value": rate, "unit": "sample-iters/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": rate, "unit": "sample-iters/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "jvm": jvm,
        "note": jvm["verdict"] + "; the C restatement of the Java arithmetic stands in for it",
    }
    emit(out)


def workload_config(args, world):
    return {"workload": "gaussian blobs %dx%d fp32-exact, k=%d, %s, init = rows floor(i*N/k)" % (
                args.rows, args.d, args.k, args.distance.upper()),
            "rows": args.rows, "d": args.d, "k": args.k, "distance": args.distance.upper(), "seed": SEED,
            "sharding": "rows by DefaultDistributedInfo over %d rank(s); allreduce k*(d+1) f64 via NCCL" % world,
            "l2": "inputs (%.1f GB per rank) far exceed the 126 MB L2; no flush needed" % (args.rows * args.d * 4 / world / 1e9),
            "epsilon": 0.0}


def source_hash():
    """Hash of the CUDA sources: keys the ncu traffic figures committed under profiles/ to the build they came from."""
    import hashlib
    hs = hashlib.sha1()
    csrc = os.path.join(ROOT, "alink_b200", "csrc")
    for name in sorted(os.listdir(csrc)):
        if name.endswith((".cu", ".cuh")):
            with open(os.path.join(csrc, name), "rb") as f:
                hs.update(name.encode()); hs.update(f.read())
    return hs.hexdigest()[:12]


def measured_traffic(kernel, rows, d, k):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch from a committed `ncu --set full` capture of THIS
    build at this shape (profiles/traffic.json, written by scripts/ncu_traffic.py); None when no capture matches."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if not os.path.exists(p):
        return None
    try:
        with open(p) as f:
            t = json.load(f)
    except Exception:
        return None
    if t.get("source_hash") != source_hash():
        return None
    for e in t.get("kernels", []):
        if e["kernel"].startswith(kernel) and e["d"] == d and e["k"] == k and e["rows"] > 0:
            return {"bytes_per_launch": e["dram_bytes"] / e["rows"] * rows, "capture": e.get("file"),
                    "capture_rows": e["rows"]}
    return None


def roofline_objects(pk, rows, d, k_nominal, k_real, iter_ms, search_ms, tiles_frac, engine_tensor, dense=None,
                     dense_search_ms=None, dense_iter_ms=None):
    """2*k*d flop and 4*d+4 bytes per sample-iteration; the bound follows from flop/byte against the
    measured ridge.  The utilisation figures (achieved / frac) are taken where the algorithmic work is what the
    kernels execute: the full-evaluation leg (every centroid tile of every row group, every row re-accumulated).
    In the steady state exact bounds skip most of that work, so its time is reported beside it (`steady_state`)
    together with the flops that were executed -- a speed-up, not a utilisation."""
    fpb = 2.0 * k_nominal * d / (4.0 * d + 4.0)
    ridge = pk["tf_sust"] * 1e12 / (pk["hbm"] * 1e9)
    bound = "hbm" if fpb < 0.8 * ridge else "tensor"
    flops_alg = 2.0 * rows * k_real * d
    flops_exec = flops_alg * tiles_frac
    byts = rows * (4.0 * d + 4.0)
    have_dense = dense_search_ms is not None and dense_search_ms > 0 and dense_iter_ms is not None and dense_iter_ms > 0
    k_ms = dense_search_ms if have_dense else search_ms
    k_flops = flops_alg if have_dense else flops_exec
    i_ms = dense_iter_ms if have_dense else iter_ms
    out = {"flop_per_byte": fpb, "ridge_flop_per_byte": ridge, "peak_source": pk["source"],
           "measured_on": "full-evaluation leg: every (row group, centroid tile) pair evaluated, every row re-accumulated"
           if have_dense else "timed steps (no full-evaluation leg in this run)"}
    steady = {"ms_per_step": iter_ms, "search_ms": search_ms, "tiles_evaluated_frac": tiles_frac,
              "executed_tflops": flops_exec / (search_ms / 1e3) / 1e12 if search_ms > 0 else 0.0,
              "nominal_tflops": 2.0 * rows * k_nominal * d / (iter_ms / 1e3) / 1e12,
              "note": "timed steps: row groups whose rows provably keep their centroid are not searched; nominal_tflops "
                      "(every sample x every centroid over the step time) is a speed-up figure, NOT a utilisation"}
    tens = {"bound": "tensor", "kernel": "search_pair_kernel (tcgen05 nearest-centroid search)",
            "achieved": k_flops / (k_ms / 1e3) / 1e12 if k_ms > 0 else 0.0, "peak": pk["tf_sust"],
            "unit": "TFLOP/s", "kernel_ms": k_ms, "k_real": k_real,
            "algorithmic": "2*k*d flop per sample x %d samples, real k%s" % (rows, "" if have_dense else " x (evaluated tiles / all tiles)"),
            "iteration": {"achieved": k_flops / (i_ms / 1e3) / 1e12, "frac": k_flops / (i_ms / 1e3) / 1e12 / pk["tf_sust"],
                          "ms": i_ms, "note": "same flops over the whole superstep"}}
    tens["frac"] = tens["achieved"] / pk["tf_sust"]
    tens["frac_of_burst"] = tens["achieved"] / pk["tf_burst"]
    hbm = {"bound": "hbm", "kernel": "whole iteration (search + refine + accumulate + update)",
           "achieved": byts / (i_ms / 1e3) / 1e9, "peak": pk["hbm"], "unit": "GB/s", "kernel_ms": i_ms,
           "algorithmic": "(4*d+4) B per sample x %d samples over the iteration time" % rows}
    hbm["frac"] = hbm["achieved"] / pk["hbm"]
    if not engine_tensor:
        tens["kernel"] = "assign_exact_kernel (fp64 CUDA cores)"
    main = dict(hbm if bound == "hbm" else tens)
    main.update(out)
    main["steady_state"] = steady
    if bound == "hbm":
        main["search_kernel"] = tens
    elif fpb < 1.25 * ridge:
        main["hbm_reading"] = hbm                        # at the ridge (config 5): both roofs, as SURVEY 8(d) asks
    if dense:
        main["dense"] = dense
    return main


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--rows", type=int, default=100_000_000)
    ap.add_argument("--d", type=int, default=128)
    ap.add_argument("--k", type=int, default=1000)
    ap.add_argument("--engine", default="auto", choices=["auto", "exact", "tensor"])
    ap.add_argument("--distance", default="euclidean", choices=["euclidean", "cosine"])
    ap.add_argument("--cpu-rows", type=int, default=0)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the BASELINE config 2 / 3 / 5 sub-lines")
    ap.add_argument("--upload-gb", type=float, default=8.0, help="host doubles pushed through akm_upload_rows for `upload`")
    ap.add_argument("--max-iter", type=int, default=10, help="steps of the from_init leg (SURVEY 8d: maxIter = 10)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)
    args.metric = 1 if args.distance == "cosine" else 0     # akm.h AKM_EUCLIDEAN / AKM_COSINE, same in oracle/

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    # The contract is ONE JSON line on stdout.  Native libraries (NCCL's version banner, for one) write to
    # fd 1 directly, so everything except that line is sent to stderr.
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)

    def emit(obj):
        os.write(json_fd, (json.dumps(obj) + "\n").encode())

    if args.impl == "reference":
        run_reference(args, rank, world, emit)
        return

    import torch
    import torch.distributed as dist
    from alink_b200 import native

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    pk = peaks()
    N, d, k = args.rows, args.d, args.k
    start, cnt = shard(rank, world, N)
    stream = torch.cuda.Stream()
    ENG = {"auto": native.ENGINE_AUTO, "exact": native.ENGINE_EXACT, "tensor": native.ENGINE_TENSOR}[args.engine]

    def timed_steps(h, steps):
        """`steps` supersteps of akm_step, CUDA events on the launching stream; returns (ms max over ranks, stats)."""
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        acc = dict(certified=0, refined=0, tiles_done=0, tiles_all=0, objective=None, k=None)
        with torch.cuda.stream(stream):
            e0.record(stream)
            for _ in range(steps):
                _, kk, obj = h.step(0.0)
                st = h.stats()
                acc["certified"] += st["rows_certified"]; acc["refined"] += st["rows_refined"]
                acc["tiles_done"] += st["tiles_computed"]; acc["tiles_all"] += st["tiles_total"]
                acc["objective"], acc["k"] = obj, kk
            e1.record(stream)
        barrier()
        return max_over_ranks(e0.elapsed_time(e1)), acc

    def phases(prof):
        sp = max(1, prof["steps"])
        return {"search": prof["search_ms"] / sp, "refine": prof["refine_ms"] / sp, "accumulate": prof["accumulate_ms"] / sp,
                "allreduce_update": prof["update_ms"] / sp}

    h = native.Handle(d, args.metric, device=local_rank)
    h.set_stream(stream.cuda_stream)
    h.set_engine(ENG)
    t0 = time.perf_counter()
    h.reserve_rows(cnt)
    h.blobs_append(SEED, k, start, cnt)
    torch.cuda.synchronize()
    gen_s = time.perf_counter() - t0
    init_rows = (np.arange(k, dtype=np.int64) * N) // k
    C0 = h.blobs_rows(SEED, k, init_rows).astype(np.float64)

    if world > 1:
        obj = [native.unique_id() if rank == 0 else None]
        dist.broadcast_object_list(obj, src=0)
        h.comm_init(obj[0], rank, world)

    # ---------------- from_init: maxIter supersteps from the initial centroids, nothing warmed up on this data
    for opt in filter(None, os.environ.get("AKM_BENCH_OPTIONS", "").split(",")):    # experiments: "regroup_after=1,..."
        h.set_option(opt.split("=")[0], int(opt.split("=")[1]))
    h.set_centroids(C0)
    h.set_profiling(True)
    launches_a = h.stats()["kernel_launches"]
    step_ms, step_info = [], []
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        ev0.record(stream)
        for _ in range(args.max_iter):
            t1 = time.perf_counter()
            h.step(0.0)                                       # returns after its own stream synchronisation
            step_ms.append((time.perf_counter() - t1) * 1e3)
            st_i = h.stats()
            step_info.append({"moved": int(st_i["rows_moved"]), "delta": int(st_i["accumulate_mode"]),
                              "tiles": round(st_i["tiles_computed"] / max(1, st_i["tiles_total"]), 3)})
        ev1.record(stream)
    barrier()
    fi_ms = max_over_ranks(ev0.elapsed_time(ev1))
    fi_prof = h.get_profile(reset=True)
    from_init = {"value": N * args.max_iter / (fi_ms / 1e3), "unit": "sample-iters/s", "max_iter": args.max_iter,
                 "ms_total": fi_ms, "ms_per_step_mean": fi_ms / max(1, args.max_iter),
                 "ms_per_step": [round(x, 3) for x in step_ms], "per_step": step_info,
                 "regroup": {"count": fi_prof["reorders"], "ms": fi_prof["reorder_ms"]},
                 "phases_ms_per_step_mean": phases(fi_prof),
                 "gpu_launches": int(h.stats()["kernel_launches"] - launches_a),
                 "note": "akm_set_centroids(C0) then %d akm_step calls with epsilon = 0, CUDA events around the loop (rank 0's "
                         "per-step wall times listed); includes the unordered first pass(es) and the regrouping of the rows"
                         % args.max_iter}

    # ---------------- value: the loop in steady state
    for _ in range(args.warmup):
        h.step(0.0)
    h.get_profile(reset=True)
    launches0 = h.stats()["kernel_launches"]
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ms_total, acc = timed_steps(h, args.steps)
    clocks = sampler.stop() if rank == 0 else None
    st = h.stats()
    launches = st["kernel_launches"] - launches0
    prof = h.get_profile(reset=True)
    value = N * args.steps / (ms_total / 1e3)
    ms_per_step = ms_total / args.steps
    objective, kk = acc["objective"], acc["k"]
    engine_used = st["engine_used"]

    # ---------------- the same supersteps the way the reference does them: every sample against every centroid
    # (tile pruning off) and the accumulator refilled from every row (delta accumulation off)
    h.set_option("tile_pruning", 0)
    h.set_option("delta_accumulate", 0)
    h.step(0.0)
    h.get_profile(reset=True)
    ms_full, acc_full = timed_steps(h, args.steps)
    prof_full = h.get_profile(reset=True)
    h.set_option("tile_pruning", 1)
    h.set_option("delta_accumulate", 1)
    h.set_profiling(False)
    ph_full = phases(prof_full)
    dense = {"ms_per_step": ms_full / args.steps, "value": N * args.steps / (ms_full / 1e3), "phases_ms_per_step": ph_full,
             "search_tflops": 2.0 * cnt * kk * d / (ph_full["search"] / 1e3) / 1e12 if ph_full["search"] > 0 else None,
             "search_frac": (2.0 * cnt * kk * d / (ph_full["search"] / 1e3) / 1e12 / pk["tf_sust"]) if ph_full["search"] > 0 else None,
             "iteration_frac": 2.0 * cnt * kk * d / (ms_full / args.steps / 1e3) / 1e12 / pk["tf_sust"],
             "iteration_frac_of_burst": 2.0 * cnt * kk * d / (ms_full / args.steps / 1e3) / 1e12 / pk["tf_burst"],
             "note": "tile_pruning = 0 and delta_accumulate = 0: same results, every (row group, centroid tile) pair "
                     "evaluated and every row re-accumulated each step; flops use the real k = %d" % kk}

    # ---------------- e2e: host-buffer supersteps (what the Java ComputeFunctions would call)
    e2e = None
    if not args.no_e2e:
        m = h.get_model()
        k0 = C0.shape[0]
        steps_e = max(1, args.steps)
        # host buffers in page-locked memory (akm_host_alloc), as the contract asks: the centroids go host->device from
        # one pinned buffer and come back into the other every step
        pa, pb = native.PinnedArray((k0, d)), native.PinnedArray((k0, d))
        np.copyto(pa.array, m["C"])
        bufs = [pa.array, pb.array]
        k_cur = m["k"]
        for i in range(2 + steps_e):                          # two untimed supersteps (seeds of the host-buffer chain)
            if i == 2:
                barrier()
                t0 = time.perf_counter()
            r = h.superstep(bufs[i & 1], k=k_cur, step_no=i + 1, tol=0.0, out=bufs[(i + 1) & 1])
            k_cur = r["k"]
        barrier()
        sec = max_over_ranks(time.perf_counter() - t0)
        pa.free(); pb.free()
        C_host, k_cur = m["C"], m["k"]
        for i in range(2 + steps_e):
            if i == 2:
                barrier()
                t0 = time.perf_counter()
            s, _, _ = h.assign_accumulate(C_host, k=k_cur)
            if world > 1:
                s = h.allreduce_sum(s)
            C_host, k_cur, _ = h.update_centroids(s, k0, k_cur, step_no=i + 1, prev=C_host)
        barrier()
        sec3 = max_over_ranks(time.perf_counter() - t0)
        e2e = {"value": N * steps_e / sec, "unit": "sample-iters/s", "h2d_bytes_per_step": k0 * d * 8,
               "d2h_bytes_per_step": k0 * d * 8 + 9 * 4 + 8, "ms_per_step": sec / steps_e * 1e3,
               "api": "akm_superstep (centroids in, new centroids + k + criterion + objective out; all-reduce on the device "
                      "buffer); host buffers page-locked (akm_host_alloc)",
               "fine_grained": {"value": N * steps_e / sec3, "ms_per_step": sec3 / steps_e * 1e3,
                                "h2d_bytes_per_step": k0 * d * 8 + (k0 * (d + 1) * 8 + k0 * d * 8) + (k0 * (d + 1) * 8 if world > 1 else 0),
                                "d2h_bytes_per_step": k0 * (d + 1) * 8 + 8 + k0 * d * 8 + k0 * 8 + (k0 * (d + 1) * 8 if world > 1 else 0),
                                "api": "akm_assign_accumulate + akm_allreduce_sum + akm_update_centroids"}}

    # ---------------- parity gate (rank 0): the CPU leg's rows from C0, oracle vs GPU
    cpu = None
    parity = None
    if rank == 0 and world == 1 and not args.no_cpu:
        import oracle as orc
        threads = os.cpu_count() or 1
        n_cpu = args.cpu_rows or size_cpu_sample(d, k, threads, 12.0, args.metric)
        n_cpu = min(n_cpu, cnt)
        Xc = h.blobs_rows(SEED, k, np.arange(n_cpu, dtype=np.int64)).astype(np.float64)
        rate, sec, first = cpu_lloyd_rate(Xc, C0, threads, 1, 0, args.metric, keep_first=True)
        jvm = probe_jvm()
        cpu = {"value": rate, "unit": "sample-iters/s", "cores": threads, "kind": "port", "jvm": jvm,
               "sample": "first %d of %d rows, 1 Lloyd iteration (%.1f s), C restatement of Alink's Java order incl. the duplicated "
                         "distance evaluation, %d threads = Flink subtasks; %s" % (n_cpu, N, sec, threads, jvm["verdict"])}
        Cp, _ = orc.prepare_centroids(args.metric, C0)
        # (a) the big partition's pass against C0: its labels for those rows
        _, _, lab_all = h.assign_accumulate(Cp, want_labels=True)
        mism_big = int((lab_all[:n_cpu] != first["labels"]).sum())
        del lab_all
        # (b) the same rows alone on a second handle: sums and objective as well
        with native.Handle(d, args.metric, device=local_rank) as h2:
            h2.set_engine(ENG)
            h2.upload_rows(Xc)
            s2, obj2, lab2 = h2.assign_accumulate(Cp, want_labels=True)
        ref_s = first["sum"].reshape(k, d + 1)
        got_s = s2.reshape(k, d + 1)
        scale = np.maximum(np.abs(ref_s).max(axis=1, keepdims=True), 1e-300)
        parity = {"rows": int(n_cpu), "label_mismatches": mism_big + int((lab2 != first["labels"]).sum()),
                  "label_mismatches_full_partition_pass": mism_big,
                  "count_mismatches": int((got_s[:, d] != ref_s[:, d]).sum()),
                  "sum_rel_err": float((np.abs(got_s - ref_s) / scale).max()),
                  "objective_rel_err": abs(obj2 - first["objective"]) / abs(first["objective"]),
                  "against": "oracle (C restatement of the Java order) from the initial centroids on the same rows; "
                             "bars: 0 label mismatches, sums / objective within 1e-6 relative"}
        del Xc

    # ---------------- upload path (one-off at superstep 1): host doubles through akm_upload_rows
    upload = None
    if rank == 0 and args.upload_gb > 0:
        blk = h.blobs_rows(SEED, k, np.arange(65536, dtype=np.int64)).astype(np.float64)
        reps = max(1, int(args.upload_gb * 1e9 / blk.nbytes))
        Xs = np.empty((reps * blk.shape[0], d), dtype=np.float64)
        for i in range(reps):
            Xs[i * blk.shape[0]:(i + 1) * blk.shape[0]] = blk
        with native.Handle(d, args.metric, device=local_rank) as h2:
            h2.reserve_rows(Xs.shape[0])
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            h2.upload_rows(Xs)
            t_ret = time.perf_counter() - t0
            h2.upload_wait()
            up_s = time.perf_counter() - t0
        # the same rows from page-locked caller memory (akm_host_alloc): no staging copy, PCIe-bound
        pinned = None
        try:
            rows_p = min(Xs.shape[0], int(2e9 / (d * 8)))
            with native.PinnedArray((rows_p, d)) as pa, native.Handle(d, args.metric, device=local_rank) as h2:
                np.copyto(pa.array, Xs[:rows_p])
                h2.reserve_rows(4 * rows_p)
                h2.upload_rows(pa.array[:1024]); h2.upload_wait(); h2.clear_rows(); h2.reserve_rows(4 * rows_p)
                t0 = time.perf_counter()
                for _ in range(4):
                    h2.upload_rows(pa.array)
                h2.upload_wait()
                dt = time.perf_counter() - t0
                pinned = {"host_bytes": int(4 * pa.array.nbytes), "seconds": dt, "host_gbps": 4 * pa.array.nbytes / dt / 1e9,
                          "note": "source in akm_host_alloc memory: DMA straight from the caller's buffer (doubles cross PCIe)"}
        except Exception as e:
            pinned = {"error": "%s: %s" % (type(e).__name__, e)}
        upload = {"rows": int(Xs.shape[0]), "host_bytes": int(Xs.nbytes), "seconds": up_s, "seconds_until_return": t_ret,
                  "host_gbps": Xs.nbytes / up_s / 1e9, "page_locked_source": pinned,
                  "note": "akm_upload_rows from pageable host doubles incl. prepareVectorData: copy threads narrow fp32-representable "
                          "chunks to fp32 into pinned double buffers, DMA and kernel overlap; one-off, outside the iteration"}
        del Xs

    # ---------------- roofline of the dominant kernel
    ph = phases(prof)
    tile_frac = (acc["tiles_done"] / acc["tiles_all"]) if acc["tiles_all"] else 1.0
    tens_used = engine_used == native.ENGINE_TENSOR
    roofline = roofline_objects(pk, cnt, d, k, kk, ms_per_step, ph["search"] if tens_used else ph["refine"],
                                tile_frac, tens_used, dense,
                                dense_search_ms=ph_full["search"] if tens_used else ph_full["refine"],
                                dense_iter_ms=ms_full / args.steps)
    tr = measured_traffic("search_pair_kernel", cnt, d, k)
    roofline["traffic"] = tr["bytes_per_launch"] if tr else None
    roofline["traffic_source"] = tr if tr else "no ncu capture of this build (source hash %s) under profiles/" % source_hash()
    acc_bytes = cnt * (d * 4 + 4)
    acc_full_ms = ph_full["accumulate"]
    tra = measured_traffic("accumulate_kernel", cnt, d, k)
    roofline_accumulate = {"bound": "hbm", "kernel": "accumulate_kernel (full pass, delta_accumulate = 0 leg)",
                           "achieved": acc_bytes / (acc_full_ms / 1e3) / 1e9 if acc_full_ms > 0 else 0.0, "peak": pk["hbm"], "unit": "GB/s",
                           "frac": (acc_bytes / (acc_full_ms / 1e3) / 1e9 / pk["hbm"]) if acc_full_ms > 0 else 0.0,
                           "traffic": tra["bytes_per_launch"] if tra else None, "kernel_ms": acc_full_ms}

    certified = sum_over_ranks(acc["certified"])
    refined = sum_over_ranks(acc["refined"])
    ranks_agree = None
    if world > 1:
        m = h.get_model()
        sig = float(np.abs(m["C"]).sum())
        lo = -max_over_ranks(-sig)
        ranks_agree = {"model_checksum_equal": bool(max_over_ranks(sig) == lo), "objective_equal": bool(max_over_ranks(objective) == -max_over_ranks(-objective))}
    h.close()
    del h
    torch.cuda.empty_cache()

    # ---------------- the other BASELINE shapes, time-boxed (rank 0, 1 GPU)
    configs = None
    if rank == 0 and world == 1 and not args.no_configs and (N, d, k, args.metric) == (100_000_000, 128, 1000, 0):
        configs = []
        for name, (n_c, d_c, k_c, met) in (("cfg2", (1_000_000, 784, 10, 0)), ("cfg3", (10_000_000, 64, 100, 0)),
                                           ("cfg5", (20_000_000, 256, 500, 1))):
            try:
                configs.append(run_config(native, torch, stream, pk, name, n_c, d_c, k_c, met, args.max_iter))
            except Exception as e:                            # a sub-line must never cost the headline
                configs.append({"config": name, "error": "%s: %s" % (type(e).__name__, e)})

    if rank == 0:
        out = {
            "metric": METRIC, "value": value, "unit": "sample-iters/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64 (search in fp16 on tensor cores, every decision certified or re-evaluated in fp64)"
            if engine_used == native.ENGINE_TENSOR else "f64",
            "data": "synthetic", "config": workload_config(args, world),
            "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
            "roofline": roofline, "roofline_accumulate": roofline_accumulate, "cpu_baseline": cpu, "parity": parity,
            "from_init": from_init,
            "phases_ms_per_step": ph,
            "full_evaluation": dense,
            "rows_certified_frac": certified / max(1.0, certified + refined),
            "objective_last": objective, "k_last": kk, "generate_s": gen_s, "upload": upload,
            "ranks_agree": ranks_agree, "configs": configs, "source_hash": source_hash(),
            "steady_state_note": "value = steps %d..%d of the run (after from_init and %d warm-up steps, epsilon = 0): labels have "
                                 "mostly stopped moving, which is when the exact bounds (seeds, tile lemma, second-nearest "
                                 "bounds: roofline.steady_state.tiles_evaluated_frac of the tiles are searched) and delta "
                                 "accumulation are at their best -- same results as an exhaustive step, less work; "
                                 "from_init is the run a user gets (maxIter = 10 from the initial centroids), "
                                 "full_evaluation the step with nothing skipped, and roofline is taken on the latter"
                                 % (args.max_iter + args.warmup + 1, args.max_iter + args.warmup + args.steps, args.warmup),
            "engine": {native.ENGINE_TENSOR: "tensor", native.ENGINE_EXACT: "exact"}.get(engine_used, "?"),
        }
        emit(out)
    if world > 1:
        dist.destroy_process_group()


def run_config(native, torch, stream, pk, name, n, d, k, metric, max_iter):
    """One BASELINE shape on one GPU: from_init (max_iter steps) + 5 steady steps, with its own roofline object;
    config 5 adds KMeansPredictBatchOp over the cached rows (ids + distances to host arrays)."""
    with native.Handle(d, metric, device=torch.cuda.current_device()) as h:
        h.set_stream(stream.cuda_stream)
        h.reserve_rows(n)
        h.blobs_append(SEED, k, 0, n)
        C0 = h.blobs_rows(SEED, k, (np.arange(k, dtype=np.int64) * n) // k).astype(np.float64)
        h.set_centroids(C0)
        h.set_profiling(True)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        step_ms = []
        with torch.cuda.stream(stream):
            e0.record(stream)
            for _ in range(max_iter):
                t1 = time.perf_counter()
                h.step(0.0)
                step_ms.append(round((time.perf_counter() - t1) * 1e3, 3))
            e1.record(stream)
        torch.cuda.synchronize()
        fi_ms = e0.elapsed_time(e1)
        for _ in range(3):
            h.step(0.0)
        h.get_profile(reset=True)
        tiles_done = tiles_all = 0
        with torch.cuda.stream(stream):
            e0.record(stream)
            for _ in range(5):
                _, kk, obj = h.step(0.0)
                st = h.stats()
                tiles_done += st["tiles_computed"]; tiles_all += st["tiles_total"]
            e1.record(stream)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 5
        prof = h.get_profile(reset=True)
        sp = max(1, prof["steps"])
        ph = {"search": prof["search_ms"] / sp, "refine": prof["refine_ms"] / sp, "accumulate": prof["accumulate_ms"] / sp,
              "allreduce_update": prof["update_ms"] / sp}
        # the reference's reading of an iteration: everything evaluated, everything re-accumulated
        h.set_option("tile_pruning", 0); h.set_option("delta_accumulate", 0)
        h.step(0.0)
        h.get_profile(reset=True)
        with torch.cuda.stream(stream):
            e0.record(stream)
            for _ in range(5):
                h.step(0.0)
            e1.record(stream)
        torch.cuda.synchronize()
        ms_dense = e0.elapsed_time(e1) / 5
        prof_d = h.get_profile(reset=True)
        spd = max(1, prof_d["steps"])
        ph_dense = {"search": prof_d["search_ms"] / spd, "refine": prof_d["refine_ms"] / spd, "accumulate": prof_d["accumulate_ms"] / spd,
                    "allreduce_update": prof_d["update_ms"] / spd}
        h.set_option("tile_pruning", 1); h.set_option("delta_accumulate", 1)
        engine_tensor = st["engine_used"] == native.ENGINE_TENSOR
        tf = (tiles_done / tiles_all) if tiles_all else 1.0
        out = {"config": name, "rows": n, "d": d, "k": k, "distance": "COSINE" if metric else "EUCLIDEAN",
               "value": n / (ms / 1e3), "unit": "sample-iters/s", "ms_per_step": ms, "phases_ms_per_step": ph,
               "k_last": kk, "objective_last": obj,
               "from_init": {"value": n * max_iter / (fi_ms / 1e3), "ms_total": fi_ms, "ms_per_step": step_ms, "max_iter": max_iter},
               "full_evaluation": {"ms_per_step": ms_dense, "value": n / (ms_dense / 1e3), "phases_ms_per_step": ph_dense,
                                   "hbm_frac": n * (4.0 * d + 4.0) / (ms_dense / 1e3) / 1e9 / pk["hbm"],
                                   "tensor_frac": 2.0 * n * kk * d / (ms_dense / 1e3) / 1e12 / pk["tf_sust"]},
               "roofline": roofline_objects(pk, n, d, k, kk, ms, ph["search"] if engine_tensor else ph["refine"], tf, engine_tensor,
                                            dense_search_ms=ph_dense["search"] if engine_tensor else ph_dense["refine"],
                                            dense_iter_ms=ms_dense)}
        tr = measured_traffic("search_pair_kernel", n, d, k)
        out["roofline"]["traffic"] = tr["bytes_per_launch"] if tr else None
        out["roofline"]["traffic_source"] = tr if tr else "no ncu capture of this build and shape under profiles/"
        out["roofline"]["traffic_kernel"] = "search_pair_kernel" 
        if name == "cfg5":
            m = h.get_model()
            C = m["C"][: m["k"]]
            h.predict_cached(C, want_dist=True)               # warm-up (allocations)
            t0 = time.perf_counter()
            h.predict_cached(C, want_dist=False)
            t_ids = time.perf_counter() - t0
            t0 = time.perf_counter()
            h.predict_cached(C, want_dist=True)
            t_dist = time.perf_counter() - t0
            out["predict_cached"] = {"rows_per_s_ids": n / t_ids, "rows_per_s_ids_and_distance": n / t_dist,
                                     "ms_ids": t_ids * 1e3, "ms_ids_and_distance": t_dist * 1e3, "k": m["k"],
                                     "note": "akm_predict_cached over the cached 20M rows, results in host arrays (wall clock)"}
        return out


if __name__ == "__main__":
    main()
