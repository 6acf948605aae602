#!/usr/bin/env python
"""bench.py -- KMeans Lloyd-iteration throughput (sample-iters/s) of the B200 path, per BASELINE.json.

Workload (config.workload): synthetic Gaussian blobs 100M x 128, k = 1000, EUCLIDEAN -- the
configuration the north_star's target is quoted on; it fits one B200 (51.2 GB fp32 + 25.6 GB fp16
search copy).  With --gpus N the same 100M rows are row-sharded over N ranks by Alink's
DefaultDistributedInfo formula (strong scaling) and the k x (d+1) accumulator is all-reduced with NCCL.

One "step" = one Lloyd superstep of KMeansTrainBatchOp.iterateICQ over the cached partition:
KMeansAssignCluster -> AllReduce -> KMeansUpdateCentroids -> KMeansIterTermination.
  value : device-resident loop (akm_step), CUDA events on the launching stream, max over ranks.
  e2e   : the same superstep through the reference-facing host-buffer calls (akm_assign_accumulate +
          akm_allreduce_sum + akm_update_centroids): centroids go host->device and the accumulator comes
          device->host every step, wall clock with synchronisation on both sides, max over ranks.
          (The sample matrix is cached on the GPU at superstep 1 exactly like Alink caches it on the JVM
          heap, BaseComQueue.java:497-527; its one-off upload is reported separately as `upload`.)
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


def cpu_lloyd_rate(X, C0, threads, steps, warmup, metric=0):
    """sample-iters/s of the Java-order CPU restatement: assign (distances evaluated twice, like
    KMeansUtil.updateSumMatrix) + all-reduce in rank order + update, `threads` Flink subtasks."""
    import oracle as orc
    n, d = X.shape
    k0 = C0.shape[0]
    Xp, xl = orc.prepare_rows(metric, X)
    C, cl = orc.prepare_centroids(metric, C0)
    k = k0
    times = []
    for s in range(warmup + steps):
        t0 = time.perf_counter()
        sm, _, _ = orc.assign_accumulate(metric, Xp, xl, C, cl, k=k, dup=True, threads=threads)
        C, cl, k = orc.update_centroids(metric, sm, k, k0, d)
        t1 = time.perf_counter()
        if s >= warmup:
            times.append(t1 - t0)
    total = sum(times)
    return n * len(times) / total, total / len(times)


def size_cpu_sample(d, k, threads, target_s, metric=0):
    """Rows per CPU step so that one step costs about target_s seconds (probe, then scale)."""
    probe = max(threads * 200, 1000)
    X = cpu_rows(probe, d, k, seed=1)
    rate, _ = cpu_lloyd_rate(X, X[:k].copy() if probe >= k else cpu_rows(k, d, k, 2), threads, 1, 0, metric)
    return int(max(probe, min(rate * target_s, 4_000_000)))


def run_reference(args, rank, world, emit):
    """Reference arm: Alink's own CPU implementation of the path = the oracle port (no JVM on this image)."""
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    n_s = args.cpu_rows or size_cpu_sample(args.d, args.k, threads, 4.0, args.metric)
    X = cpu_rows(n_s, args.d, args.k)
    C0 = X[(np.arange(args.k) * n_s) // args.k].copy()
    rate, sec = cpu_lloyd_rate(X, C0, threads, args.steps, args.warmup, args.metric)
    sample = "oracle port of Alink's Java order (2 dgemv/sample, no FMA), %d threads = Flink subtasks, %d of %d rows per step" % (
        threads, n_s, args.rows)
    out = {
        "impl": "reference", "metric": METRIC, "value": rate, "unit": "sample-iters/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args, 1),
        "cpu_baseline": {"value": rate, "unit": "sample-iters/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": rate, "unit": "sample-iters/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "JVM/Alink jars are not runnable on this image (no java); the C restatement stands in for the Java path",
    }
    emit(out)


def workload_config(args, world):
    return {"workload": "gaussian blobs %dx%d fp32-exact, k=%d, %s, init = rows floor(i*N/k)" % (
                args.rows, args.d, args.k, args.distance.upper()),
            "rows": args.rows, "d": args.d, "k": args.k, "distance": args.distance.upper(), "seed": SEED,
            "sharding": "rows by DefaultDistributedInfo over %d rank(s); allreduce k*(d+1) f64 via NCCL" % world,
            "l2": "inputs (%.1f GB per rank) far exceed the 126 MB L2; no flush needed" % (args.rows * args.d * 4 / world / 1e9),
            "epsilon": 0.0}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
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

    N, d, k = args.rows, args.d, args.k
    start, cnt = shard(rank, world, N)
    stream = torch.cuda.Stream()
    h = native.Handle(d, args.metric, device=local_rank)
    h.set_stream(stream.cuda_stream)
    h.set_engine({"auto": native.ENGINE_AUTO, "exact": native.ENGINE_EXACT, "tensor": native.ENGINE_TENSOR}[args.engine])
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

    # ---------------- value: device-resident loop
    h.set_centroids(C0)
    h.set_profiling(True)
    for _ in range(args.warmup):
        h.step(0.0)
    warm_prof = h.get_profile(reset=True)          # carries the one-off regrouping done during warm-up
    launches0 = h.stats()["kernel_launches"]
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    objective = None
    certified = refined = 0
    tiles_done = tiles_all = 0
    with torch.cuda.stream(stream):
        ev0.record(stream)
        for _ in range(args.steps):
            conv, kk, objective = h.step(0.0)
            st = h.stats()
            certified += st["rows_certified"]; refined += st["rows_refined"]
            tiles_done += st["tiles_computed"]; tiles_all += st["tiles_total"]
        ev1.record(stream)
    barrier()
    ms_total = max_over_ranks(ev0.elapsed_time(ev1))
    clocks = sampler.stop() if rank == 0 else None
    st = h.stats()
    launches = st["kernel_launches"] - launches0
    prof = h.get_profile()
    value = N * args.steps / (ms_total / 1e3)
    ms_per_step = ms_total / args.steps

    # ---------------- the same supersteps with tile pruning switched off (every sample against every centroid on the
    # tensor cores): reported beside `value` so that both readings of "one Lloyd iteration" are in the line
    h.set_option("tile_pruning", 0)
    h.get_profile(reset=True)
    barrier()
    ev2, ev3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        ev2.record(stream)
        for _ in range(args.steps):
            h.step(0.0)
        ev3.record(stream)
    barrier()
    ms_full = max_over_ranks(ev2.elapsed_time(ev3))
    prof_full = h.get_profile()
    h.set_option("tile_pruning", 1)
    h.set_profiling(False)
    sf = max(1, prof_full["steps"])
    full_eval = {"value": N * args.steps / (ms_full / 1e3), "ms_per_step": ms_full / args.steps,
                 "phases_ms_per_step": {"search": prof_full["search_ms"] / sf, "refine": prof_full["refine_ms"] / sf,
                                        "accumulate": prof_full["accumulate_ms"] / sf,
                                        "allreduce_update": prof_full["update_ms"] / sf},
                 "search_frac_of_peak": (2.0 * cnt * k * d / (prof_full["search_ms"] / sf / 1e3) / 1e12 / peaks()["tf_sust"])
                 if prof_full["search_ms"] > 0 else None,
                 "note": "tile_pruning = 0: same results, every (row group, centroid tile) pair evaluated"}

    # ---------------- e2e: host-buffer supersteps (what the Java ComputeFunctions would call)
    e2e = None
    if not args.no_e2e:
        m = h.get_model()
        C_host, k_cur, k0 = m["C"], m["k"], C0.shape[0]
        steps_e = max(1, args.steps)
        for i in range(1 + steps_e):
            if i == 1:
                barrier()
                t0 = time.perf_counter()
            s, _, _ = h.assign_accumulate(C_host, k=k_cur)
            if world > 1:
                s = h.allreduce_sum(s)
            C_host, k_cur, _ = h.update_centroids(s, k0, k_cur, step_no=i + 1, prev=C_host)
        barrier()
        sec = max_over_ranks(time.perf_counter() - t0)
        h2d = k0 * d * 8 + (k0 * (d + 1) * 8 + k0 * d * 8) + (k0 * (d + 1) * 8 if world > 1 else 0)
        d2h = k0 * (d + 1) * 8 + 8 + k0 * d * 8 + k0 * 8 + (k0 * (d + 1) * 8 if world > 1 else 0)
        e2e = {"value": N * steps_e / sec, "unit": "sample-iters/s", "h2d_bytes_per_step": h2d,
               "d2h_bytes_per_step": d2h, "ms_per_step": sec / steps_e * 1e3,
               "api": "akm_assign_accumulate + akm_allreduce_sum + akm_update_centroids (host buffers)"}

    # ---------------- upload path (one-off at superstep 1), measured on a bounded slice
    upload = None
    if rank == 0:
        m_up = min(cnt, 1_000_000)
        Xs = h.blobs_rows(SEED, k, np.arange(m_up, dtype=np.int64)).astype(np.float64)
        with native.Handle(d, args.metric, device=local_rank) as h2:
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            h2.upload_rows(Xs)
            torch.cuda.synchronize()
            up_s = time.perf_counter() - t0
        upload = {"rows": m_up, "seconds": up_s, "host_gbps": m_up * d * 8 / up_s / 1e9,
                  "note": "akm_upload_rows from pageable host doubles incl. prepareVectorData (pinned double-buffered staging, "
                          "4 copy threads); one-off, outside the iteration; ~25 GB/s on multi-GB uploads"}

    # ---------------- roofline of the dominant kernel (nearest-centroid search)
    pk = peaks()
    steps_p = max(1, prof["steps"])
    engine_used = st["engine_used"]
    search_ms = prof["search_ms"] / steps_p
    refine_ms = prof["refine_ms"] / steps_p
    dom_ms = search_ms if engine_used == native.ENGINE_TENSOR else refine_ms
    flops = 2.0 * cnt * k * d
    achieved = flops / (dom_ms / 1e3) / 1e12 if dom_ms > 0 else 0.0
    tile_frac = (tiles_done / tiles_all) if tiles_all else 1.0
    # what the tensor cores actually ran: evaluated (256-row group, 256-centroid tile) pairs, padded shapes included
    exec_flops = (tiles_done / max(1, args.steps)) * 2.0 * 256 * 256 * (((d + 63) // 64) * 64) if tiles_all else flops
    executed = exec_flops / (dom_ms / 1e3) / 1e12 if dom_ms > 0 else 0.0
    roofline = {"bound": "tensor", "kernel": "assign_search_tcgen05" if engine_used == native.ENGINE_TENSOR else "assign_exact_kernel (fp64 CUDA cores)",
                "achieved": achieved, "peak": pk["tf_sust"], "unit": "TFLOP/s", "frac": achieved / pk["tf_sust"],
                "peak_source": "bf16 dense sustained, " + pk["source"],
                # dram__bytes_read.sum + dram__bytes_write.sum of search_pair_kernel<2> from one `ncu --set full`
                # capture at 100M x 128 rows, k = 1000 (profiles/r1_ncu_full_100M_rows_final.csv): 26.81 + 1.24 GB per
                # launch = 280.4 B per row (the fp16 operand copy is 256 B per row), scaled to this launch's rows
                "traffic": (280.4 * cnt if (engine_used == native.ENGINE_TENSOR and d == 128 and k == 1000) else None),
                "traffic_unit": "bytes per launch (ncu dram read+write)",
                "algorithmic": "2*k*d flop per sample (SURVEY 8d) x %d samples per launch" % cnt,
                "kernel_ms": dom_ms,
                "executed": {"tflops": executed, "frac": executed / pk["tf_sust"], "tiles_computed_frac": tile_frac},
                "note": "achieved/frac count the ALGORITHMIC flops (every sample against every centroid); the kernel "
                        "evaluates only the (row group, centroid tile) pairs the triangle inequality cannot rule out "
                        "(`executed`), with identical results, so frac may exceed what the tensor pipe ran. The kernel "
                        "is bound by its epilogue (128 KB of accumulators per evaluated tile and CTA to read out of TMEM "
                        "and reduce: ~2000 cycles against ~1024 of MMA issue), not by the tensor pipe."}
    acc_ms = prof["accumulate_ms"] / steps_p
    acc_bytes = cnt * (d * 4 + 4)
    roofline_accumulate = {"bound": "hbm", "kernel": "accumulate_kernel", "achieved": acc_bytes / (acc_ms / 1e3) / 1e9 if acc_ms > 0 else 0.0,
                           "peak": pk["hbm"], "unit": "GB/s", "frac": (acc_bytes / (acc_ms / 1e3) / 1e9 / pk["hbm"]) if acc_ms > 0 else 0.0,
                           # accumulate_kernel: 52.89 + 0.05 GB per launch at 100M rows = 529.4 B per row
                           "traffic": (529.4 * cnt if d == 128 else None), "kernel_ms": acc_ms}

    # ---------------- CPU baseline beside it (rank 0, N = 1 only)
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        threads = os.cpu_count() or 1
        n_cpu = args.cpu_rows or size_cpu_sample(d, k, threads, 12.0, args.metric)
        n_cpu = min(n_cpu, cnt)
        Xc = h.blobs_rows(SEED, k, np.arange(n_cpu, dtype=np.int64)).astype(np.float64)
        rate, sec = cpu_lloyd_rate(Xc, C0, threads, 1, 0, args.metric)
        cpu = {"value": rate, "unit": "sample-iters/s", "cores": threads, "kind": "port",
               "sample": "first %d of %d rows, 1 Lloyd iteration (%.1f s), C restatement of Alink's Java order incl. the duplicated "
                         "distance evaluation, %d threads = Flink subtasks; JVM not runnable on this box (no java)" % (n_cpu, N, sec, threads)}

    certified = sum_over_ranks(certified)
    refined = sum_over_ranks(refined)
    h.close()
    if rank == 0:
        out = {
            "metric": METRIC, "value": value, "unit": "sample-iters/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64 (search in fp16 on tensor cores, every decision certified or re-evaluated in fp64)"
            if engine_used == native.ENGINE_TENSOR else "f64",
            "data": "synthetic", "config": workload_config(args, world),
            "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
            "roofline": roofline, "roofline_accumulate": roofline_accumulate, "cpu_baseline": cpu,
            "phases_ms_per_step": {"search": search_ms, "refine": refine_ms, "accumulate": acc_ms,
                                   "allreduce_update": prof["update_ms"] / steps_p},
            "full_evaluation": full_eval,
            "rows_certified_frac": certified / max(1.0, certified + refined),
            "objective_last": objective, "k_last": kk, "generate_s": gen_s, "upload": upload,
            "regroup": {"in_warmup": {"count": warm_prof["reorders"], "ms": warm_prof["reorder_ms"]},
                        "in_timed_steps": {"count": prof["reorders"] - warm_prof["reorders"], "ms": prof["reorder_ms"]},
                        "note": "rows are regrouped by cluster once after the 2nd pass (and again only if >25% of the "
                                "rows change cluster in one pass); its cost is one-off per training run"},
            "engine": {native.ENGINE_TENSOR: "tensor", native.ENGINE_EXACT: "exact"}.get(engine_used, "?"),
        }
        emit(out)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
