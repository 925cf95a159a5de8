#!/usr/bin/env python
"""bench.py — RF fit throughput (rows*trees/sec) on BASELINE.json's configs[1]:
RF 100 trees, depth 10, maxBins 32, 1M synthetic draw rows x 64 features, 50-class label.

A "step" is one complete fit (findSplits, binning, bagging, every group's histogram + split scan +
partition update, forest on the host).  `value` is timed with the raw f64 features already resident
in HBM (CUDA events on the engine's stream); `e2e` is the same fit through the reference-facing API
(RandomForestClassifier.fit on a host Dataset): pinned host buffers, H2D of features inside the
timed region, forest back on the host.

N > 1 (one process per GPU, torchrun): ONE forest of 100 trees on N x 1M rows, ROW-sharded across the
ranks (weak scaling: 1M rows per GPU; north_star's mode for large shapes): each rank bins, bags and
histograms its own rows and the ranks all-reduce the node histograms with NCCL every group.
`--mode trees` instead shards the trees of one 100*N-tree forest (all rows on every rank).  torch is
used for the rendezvous / barrier only.

--impl reference: times the CPU oracle (the restated Spark MLlib algorithm; no JVM/Spark exists on
this box, so there is no oracle/_ref) on the host cores, on a bounded sample of the same workload.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_ROWS, N_FEATURES, N_CLASSES = 1_000_000, 64, 50
TREES_PER_GPU, MAX_DEPTH, MAX_BINS, N_PARTS = 100, 10, 32, 1024
METRIC, UNIT = "rf_fit_rows_trees_per_sec", "rows*trees/s"
CPU_SAMPLE_TREES = 100   # the whole workload: ~8 s on 16 cores


def workload_name(n_gpus, mode="rows"):
    trees = TREES_PER_GPU * (n_gpus if mode == "trees" else 1)
    rows = N_ROWS * (n_gpus if mode == "rows" else 1)
    parts = N_PARTS * (n_gpus if mode == "rows" else 1)
    return (f"BASELINE.json configs[1]: RF {trees} trees, depth {MAX_DEPTH}, maxBins {MAX_BINS}, "
            f"{rows} synthetic draw rows x {N_FEATURES} features, {N_CLASSES}-class label, {parts} partitions")


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled every 200 ms during the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu, self.lines, self.proc = gpu_index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.gpu)], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        self.t.join(timeout=2)
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def oracle_params(ob, trees, seed=207336481):
    return ob.make_params(num_trees=trees, max_depth=MAX_DEPTH, max_bins=MAX_BINS, num_classes=N_CLASSES, seed=seed)


def time_oracle(trees):
    """CPU port timed on all host cores, on the same rows/features with `trees` trees."""
    from euromillioner_b200 import synth
    from oracle import binding as ob
    X, y = synth.flat(N_ROWS, N_FEATURES, N_CLASSES)
    po = ob.partition_offsets(N_ROWS, N_PARTS)
    t0 = time.perf_counter()
    fo = ob.fit(oracle_params(ob, trees), X, y, po)
    dt = time.perf_counter() - t0
    nodes = int(fo.export()["tree_off"][-1])
    return N_ROWS * trees / dt, dt, ob.lib().orc_num_threads(), nodes


def run_reference(args, rank):
    if rank != 0:
        return
    vals, dts = [], []
    for i in range(args.warmup + args.steps):
        v, dt, cores, _ = time_oracle(CPU_SAMPLE_TREES)
        if i >= args.warmup:
            vals.append(v); dts.append(dt)
    value = float(np.mean(vals))
    sample = (f"oracle port of Spark MLlib 2.4.6 RF (C++/OpenMP), same {N_ROWS} rows x {N_FEATURES} features, depth "
              f"{MAX_DEPTH}, {CPU_SAMPLE_TREES} of {TREES_PER_GPU} trees per step")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": float(np.mean(dts)) * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": {"workload": workload_name(1), "note": "Spark local[N] itself is unavailable on this box (no JVM)"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--mode", default="rows", choices=["rows", "trees"], help="how N > 1 GPUs share one forest")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("--gpus N > 1 must be launched with torchrun (one process per GPU)")

    from euromillioner_b200 import _lib, synth
    from euromillioner_b200.classifier import Context, Dataset, RandomForestClassifier
    L = _lib.lib()

    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        from euromillioner_b200.dist import make_context
        ctx = make_context(dist)
    else:
        ctx = Context(local_rank)

    def barrier():
        if dist is not None:
            import torch
            dist.barrier()
            torch.cuda.synchronize()

    mode = args.mode if world > 1 else "rows"
    T = TREES_PER_GPU * (world if mode == "trees" else 1)
    n_global = N_ROWS * (world if mode == "rows" else 1)
    # rows mode: this rank owns global rows [rank*1M, (rank+1)*1M) = global partitions [rank*1024, (rank+1)*1024)
    X, y = synth.flat(N_ROWS, N_FEATURES, N_CLASSES, row_start=rank * N_ROWS if mode == "rows" else 0)
    # pinned host copies (for the e2e leg) and device-resident copies (for `value`)
    hX, hy, dX, dy = C.c_void_p(), C.c_void_p(), C.c_void_p(), C.c_void_p()
    _lib.check(L.emrf_host_alloc_pinned(X.nbytes, C.byref(hX)))
    _lib.check(L.emrf_host_alloc_pinned(y.nbytes, C.byref(hy)))
    pX = np.ctypeslib.as_array(C.cast(hX, C.POINTER(C.c_double)), shape=X.shape)
    py = np.ctypeslib.as_array(C.cast(hy, C.POINTER(C.c_double)), shape=y.shape)
    pX[...] = X
    py[...] = y
    _lib.check(L.emrf_dev_alloc(ctx.h, X.nbytes, C.byref(dX)))
    _lib.check(L.emrf_dev_alloc(ctx.h, y.nbytes, C.byref(dy)))
    _lib.check(L.emrf_memcpy_h2d(ctx.h, dX, hX, X.nbytes))
    _lib.check(L.emrf_memcpy_h2d(ctx.h, dy, hy, y.nbytes))

    rf = (RandomForestClassifier(ctx).setNumTrees(T).setMaxDepth(MAX_DEPTH).setMaxBins(MAX_BINS)
          .setNumClasses(N_CLASSES).setSeed(207336481))
    if world > 1 and mode == "rows":
        goff = np.array([(i * n_global) // (N_PARTS * world) for i in range(N_PARTS * world + 1)], np.int64)
        assert goff[rank * N_PARTS] == rank * N_ROWS and goff[(rank + 1) * N_PARTS] == (rank + 1) * N_ROWS
        local = goff[rank * N_PARTS:(rank + 1) * N_PARTS + 1] - goff[rank * N_PARTS]
        shard = dict(partition_offsets=local, global_partition_offsets=goff, first_partition=rank * N_PARTS)
        ds_dev = Dataset(dX.value, dy.value, n_rows=N_ROWS, n_features=N_FEATURES, **shard)
        ds_host = Dataset(pX, py, **shard)
    else:
        ds_dev = Dataset(dX.value, dy.value, num_partitions=N_PARTS, n_rows=N_ROWS, n_features=N_FEATURES)
        ds_host = Dataset(pX, py, num_partitions=N_PARTS)
    # Dataset() copies non-contiguous input only; pX is contiguous, so ds_host.features aliases the pinned buffer
    assert ds_host.features.ctypes.data == hX.value

    for _ in range(args.warmup):
        rf.fit(ds_dev)
    # ---- timed region 1: inputs resident in HBM
    sampler = ClockSampler(local_rank)
    if rank == 0 and not os.environ.get("BENCH_NO_CLOCKS"):
        sampler.start()
    barrier()
    ev_ms, tms = [], []
    t0 = time.perf_counter()
    for _ in range(args.steps):
        model = rf.fit(ds_dev)
        tm = ctx.timings()
        ev_ms.append(tm["total_ms"]); tms.append(tm)
    barrier()
    wall_ms = (time.perf_counter() - t0) * 1e3
    # ---- timed region 2: end to end through the plugin API, host buffers
    barrier()
    t0 = time.perf_counter()
    e2e_tm, e2e_each = [], []
    for _ in range(args.steps):
        t1 = time.perf_counter()
        model = rf.fit(ds_host)
        e2e_each.append(round((time.perf_counter() - t1) * 1e3, 2))
        e2e_tm.append(ctx.timings())
    barrier()
    e2e_ms = (time.perf_counter() - t0) * 1e3
    clocks = sampler.stop() if rank == 0 else None

    step_ms = float(np.sum(ev_ms))
    if dist is not None:
        import torch
        t = torch.tensor([step_ms, e2e_ms, wall_ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        step_ms, e2e_ms, wall_ms = [float(v) for v in t.tolist()]
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    units = n_global * T * args.steps
    value = units / (step_ms / 1e3)
    hist_ms = float(np.sum([t["hist_kernel_ms"] for t in tms]))
    hist_bytes = float(np.sum([t["hist_alg_bytes"] for t in tms]))
    hist_launches = int(np.sum([t["hist_launches"] for t in tms]))
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = json.load(open(peaks_path))["hbm_gbs"], "measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    achieved = hist_bytes / (hist_ms / 1e3) / 1e9 if hist_ms > 0 else 0.0
    traffic = None
    tp = os.path.join(ROOT, "profiles", "hist_kernel_traffic.json")
    if os.path.exists(tp):
        traffic = json.load(open(tp)).get("dram_bytes_per_launch")
    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": step_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u32", "data": "synthetic",
        "config": {"workload": workload_name(world, mode),
                   "parallelism": ("row-sharded x%d (NCCL all-reduce of node histograms per group)" % world if mode == "rows"
                                   else "tree-sharded x%d" % world) if world > 1 else "1 GPU",
                   "l2": "inputs (512 MB f64 features, 1 GB entry lists) are larger than the 126 MB L2; no explicit flush",
                   "seed": 207336481, "timing": "CUDA events on the engine stream per fit (emrf_timings_t.total_ms)",
                   "wall_ms_per_step": wall_ms / args.steps, "ms_each_step": [round(v, 2) for v in ev_ms]},
        "clocks": clocks,
        "e2e": {"value": units / (e2e_ms / 1e3), "unit": UNIT,
                "h2d_bytes_per_step": int(np.mean([t["h2d_bytes"] for t in e2e_tm])),
                "d2h_bytes_per_step": int(np.mean([t["d2h_bytes"] for t in e2e_tm])),
                "ms_per_step": e2e_ms / args.steps, "ms_each_step": e2e_each,
                "phases_ms_per_step": {k: round(float(np.mean([t[k] for t in e2e_tm])), 3) for k in (
                    "total_ms", "h2d_ms", "find_splits_ms", "bin_ms", "bag_ms", "hist_kernel_ms", "partition_ms")}},
        "gpu_launches": int(np.sum([t["kernel_launches"] for t in tms])),
        "roofline": {"bound": "hbm", "kernel": "hist_scan_kernel (DTStatsAggregator update + binsToBestSplit)",
                     "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                     "peak_source": peak_src, "launches": hist_launches,
                     "avg_launch_ms": hist_ms / max(hist_launches, 1),
                     "algorithmic_bytes_per_launch": hist_bytes / max(hist_launches, 1),
                     "share_of_step": hist_ms / step_ms},
        "phases_ms_per_step": {k: float(np.mean([t[k] for t in tms])) for k in (
            "find_splits_ms", "bin_ms", "bag_ms", "hist_kernel_ms", "partition_ms", "host_scheduler_ms",
            "host_select_ms", "host_update_ms", "host_export_ms", "bag_kernel_ms")},
        "groups_per_step": int(tms[0]["num_groups"]), "total_nodes": int(model.totalNumNodes),
        "gain_mismatches": int(np.sum([t["gain_mismatches"] for t in tms])),
    }
    if world == 1 and not args.no_cpu_baseline:
        v, dt, cores, _ = time_oracle(CPU_SAMPLE_TREES)
        out["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                               "sample": f"oracle port (C++/OpenMP) on the same {N_ROWS} rows x {N_FEATURES} features, depth "
                                         f"{MAX_DEPTH}, {CPU_SAMPLE_TREES} of {TREES_PER_GPU} trees: {dt:.1f} s"}
    print(json.dumps(out), flush=True)
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
