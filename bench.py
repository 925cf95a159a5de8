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
histograms its own rows and the ranks reduce the node histograms with NCCL every group.
`--mode trees` instead shards the trees of one 100*N-tree forest (all rows on every rank).  torch is
used for the rendezvous / barrier only.

Besides the headline the line carries
  * parity evidence: `parity_vs_oracle` (N = 1: the timed forest against the CPU oracle's, field by field) and
    `parity_vs_1gpu` (N > 1: the sharded forest against a one-GPU fit of the same global rows on rank 0, plus a
    forest digest agreed by every rank);
  * `configs2`: one fit of BASELINE configs[2] (RF 500 trees, depth 12, 10M x 128, 8192 partitions) at this N —
    one GPU, row-sharded and tree-sharded — with the histogram kernel's roofline fraction, the collectives' time
    and the strong-scaling efficiency against the one-GPU fit of the same run;
  * `configs3`: BASELINE configs[3], the seven per-ball models dealt to the GPUs (model-parallel, no collective);
  * `predict`: BASELINE configs[4], 100M rows through a 500-tree depth-12 forest, rows/s.

--impl reference: times the CPU oracle (the restated Spark MLlib algorithm; no JVM/Spark exists on
this box, so there is no oracle/_ref) on the host cores, on the same workload.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_ROWS, N_FEATURES, N_CLASSES = 1_000_000, 64, 50
TREES_PER_GPU, MAX_DEPTH, MAX_BINS, N_PARTS = 100, 10, 32, 1024
METRIC, UNIT = "rf_fit_rows_trees_per_sec", "rows*trees/s"
CPU_SAMPLE_TREES = 100   # the whole workload: ~8 s on 16 cores
SEED = 207336481
# BASELINE configs[2]
C2_ROWS, C2_FEATURES, C2_TREES, C2_DEPTH, C2_PARTS = 10_000_000, 128, 500, 12, 8192
# BASELINE configs[3]: the 7 per-ball models (5 mains, 50 classes; 2 stars, 12 classes), 200 trees each, on the draw history
C3_TREES, C3_DEPTH, C3_DRAWS, C3_LAG, C3_PARTS = 200, 6, 1500, 10, 8
# BASELINE configs[4]: 100M rows through a 500-tree depth-12 forest, generated and predicted in resident chunks
C4_ROWS_TOTAL, C4_ROWS_CHUNK = 100_000_000, 10_000_000


def workload_name(n_gpus, mode="rows"):
    trees = TREES_PER_GPU * (n_gpus if mode == "trees" else 1)
    rows = N_ROWS * (n_gpus if mode == "rows" else 1)
    parts = N_PARTS * (n_gpus if mode == "rows" else 1)
    return (f"BASELINE.json configs[1]: RF {trees} trees, depth {MAX_DEPTH}, maxBins {MAX_BINS}, "
            f"{rows} synthetic draw rows x {N_FEATURES} features, {N_CLASSES}-class label, {parts} partitions")


def make_config(n_gpus, mode, parallelism, timing):
    """Both arms print the same keys (the driver compares the two lines' `config`)."""
    return {"workload": workload_name(n_gpus, mode), "seed": SEED,
            "l2": "inputs (512 MB f64 features, 1 GB entry lists) are larger than the 126 MB L2; no explicit flush",
            "parallelism": parallelism, "timing": timing}


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


def oracle_params(ob, trees, seed=SEED):
    return ob.make_params(num_trees=trees, max_depth=MAX_DEPTH, max_bins=MAX_BINS, num_classes=N_CLASSES, seed=seed)


def host_threads():
    """All the host threads this process may use.  torchrun exports OMP_NUM_THREADS=1 to its workers, which would turn
    the CPU arm into a single-threaded run: the affinity mask is what really bounds us."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def time_oracle(trees):
    """CPU port timed on all host cores, on the same rows/features with `trees` trees; returns the forest too."""
    from euromillioner_b200 import synth
    from oracle import binding as ob
    ob.lib().orc_set_num_threads(host_threads())
    X, y = synth.flat(N_ROWS, N_FEATURES, N_CLASSES)
    po = ob.partition_offsets(N_ROWS, N_PARTS)
    t0 = time.perf_counter()
    fo = ob.fit(oracle_params(ob, trees), X, y, po)
    dt = time.perf_counter() - t0
    return N_ROWS * trees / dt, dt, ob.lib().orc_num_threads(), fo.export()


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
              f"{MAX_DEPTH}, {CPU_SAMPLE_TREES} of {TREES_PER_GPU} trees per step; Spark local[N] itself is unavailable on "
              f"this box (no JVM)")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": float(np.mean(dts)) * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": make_config(1, "rows", f"CPU, {cores} OpenMP threads (all host cores this process may use)",
                              "wall clock around each fit (time.perf_counter)"),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }), flush=True)


def forests_equal(a, b):
    """Field-by-field equality of two exported forests (dicts of arrays); returns (equal, first differing field)."""
    for k in a:
        if k not in b or not np.array_equal(a[k], b[k]):
            return False, k
    return True, None


def forest_digest(exp):
    d = 0
    for k in sorted(exp):
        d = zlib.crc32(np.ascontiguousarray(exp[k]).tobytes(), d)
    return d


def peak_hbm():
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        return json.load(open(peaks_path))["hbm_gbs"], "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def hist_roofline(tms, peak):
    hist_ms = float(np.sum([t["hist_kernel_ms"] for t in tms]))
    hist_bytes = float(np.sum([t["hist_alg_bytes"] for t in tms]))
    achieved = hist_bytes / (hist_ms / 1e3) / 1e9 if hist_ms > 0 else 0.0
    return hist_ms, hist_bytes, achieved, achieved / peak


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-configs2", action="store_true", help="skip the BASELINE configs[2] block (10M x 128, 500 trees)")
    ap.add_argument("--no-configs3", action="store_true", help="skip the BASELINE configs[3] block (7 per-ball models)")
    ap.add_argument("--no-predict", action="store_true", help="skip the batch-predict block (BASELINE configs[4])")
    ap.add_argument("--no-parity", action="store_true", help="skip the N > 1 one-GPU parity fit")
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
    if args.warmup < 3:
        args.warmup = 3      # timing rules: at least three untimed fits

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
        # every context of this process (the communicator's and the plain ones of the parity / configs blocks) gets this
        # rank's share of the host cores: ranks that wait in a barrier spin on theirs, and an OpenMP team wider than its
        # share then loses its barriers to them (measured: a 1.8 s fit took 38 s)
        os.environ.setdefault("EMRF_HOST_THREADS", str(max(1, host_threads() // world)))
        ctx = make_context(dist)
    else:
        ctx = Context(local_rank)

    def barrier():
        if dist is not None:
            import torch
            dist.barrier()
            torch.cuda.synchronize()

    def allreduce_max(vals):
        if dist is None:
            return [float(v) for v in vals]
        import torch
        t = torch.tensor(list(vals), device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return [float(v) for v in t.tolist()]

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

    def estimator(c, trees, depth=MAX_DEPTH):
        return (RandomForestClassifier(c).setNumTrees(trees).setMaxDepth(depth).setMaxBins(MAX_BINS)
                .setNumClasses(N_CLASSES).setSeed(SEED))

    def shard_kw(rows_per_rank, parts_per_rank):
        """Row-sharded Dataset arguments for this rank: contiguous rows at partition boundaries."""
        ng, pg = rows_per_rank * world, parts_per_rank * world
        goff = np.array([(i * ng) // pg for i in range(pg + 1)], np.int64)
        assert goff[rank * parts_per_rank] == rank * rows_per_rank and goff[(rank + 1) * parts_per_rank] == (rank + 1) * rows_per_rank
        local = goff[rank * parts_per_rank:(rank + 1) * parts_per_rank + 1] - goff[rank * parts_per_rank]
        return dict(partition_offsets=local, global_partition_offsets=goff, first_partition=rank * parts_per_rank)

    rf = estimator(ctx, T)
    if world > 1 and mode == "rows":
        shard = shard_kw(N_ROWS, N_PARTS)
        ds_dev = Dataset(dX.value, dy.value, n_rows=N_ROWS, n_features=N_FEATURES, **shard)
        ds_host = Dataset(pX, py, **shard)
    else:
        ds_dev = Dataset(dX.value, dy.value, num_partitions=N_PARTS, n_rows=N_ROWS, n_features=N_FEATURES)
        ds_host = Dataset(pX, py, num_partitions=N_PARTS)
    # Dataset() copies non-contiguous input only; pX is contiguous, so ds_host.features aliases the pinned buffer
    assert ds_host.features.ctypes.data == hX.value

    for _ in range(args.warmup):
        model = rf.fit(ds_dev)   # held like in the timed loop: two models are alive while the next one is fitted
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
    # ---- timed region 2: end to end through the plugin API, host buffers (its own warm-up: the first fits from host
    # memory size the staging buffers of that path)
    for _ in range(args.warmup):
        model = rf.fit(ds_host)
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

    step_ms, e2e_ms, wall_ms = allreduce_max([float(np.sum(ev_ms)), e2e_ms, wall_ms])
    peak, peak_src = peak_hbm()

    # ---- parity evidence for the timed forest
    parity = {}
    exp = model.export()
    if world > 1:
        dig = float(forest_digest(exp))
        lo = -allreduce_max([-dig])[0]
        hi = allreduce_max([dig])[0]
        parity["forest_digest_same_on_all_ranks"] = bool(lo == hi)
        if not args.no_parity:
            # rank 0 fits the same global Dataset on its own GPU alone; the sharded forest must equal it bit for bit
            ok = 1.0
            if rank == 0:
                c1 = Context(local_rank)
                n1, p1, t1 = (n_global, N_PARTS * (world if mode == "rows" else 1), T)
                g = synth.DeviceFlat(c1, n1, N_FEATURES, N_CLASSES)
                m1 = estimator(c1, t1).fit(Dataset(g.X.value, g.y.value, num_partitions=p1, n_rows=n1, n_features=N_FEATURES))
                eq, field = forests_equal(exp, m1.export())
                parity["parity_vs_1gpu"] = bool(eq)
                if not eq:
                    parity["parity_first_difference"] = field
                parity["one_gpu_fit_ms"] = round(c1.timings()["total_ms"], 2)
                ok = 1.0 if eq else 0.0
                g.free()
                del m1
                c1.close()
            barrier()
            _ = ok

    # ---- BASELINE configs[2]: one fit of the shape the north star quotes its targets on, at this N
    configs2 = None
    if not args.no_configs2:
        configs2 = run_configs2(ctx, world, rank, local_rank, barrier, allreduce_max, estimator, shard_kw, peak)

    # ---- BASELINE configs[3]: the seven per-ball models, one fit per model, models dealt to the GPUs
    configs3 = None
    if not args.no_configs3:
        configs3 = run_configs3(ctx, world, rank, barrier, allreduce_max)

    # ---- batch prediction (BASELINE configs[4])
    predict = None
    if not args.no_predict:
        predict = run_predict(ctx, world, rank, barrier, allreduce_max, estimator)

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    units = n_global * T * args.steps
    value = units / (step_ms / 1e3)
    hist_ms, hist_bytes, achieved, frac = hist_roofline(tms, peak)
    hist_launches = int(np.sum([t["hist_launches"] for t in tms]))
    traffic = None
    tp = os.path.join(ROOT, "profiles", "hist_kernel_traffic.json")
    if world == 1 and os.path.exists(tp):   # one ncu capture of this round's kernel on this workload (per launch, mean)
        traffic = json.load(open(tp)).get("dram_bytes_per_launch")
    cfg = make_config(world, mode,
                      ("row-sharded x%d (NCCL reduce-scatter of node histograms per group)" % world if mode == "rows"
                       else "tree-sharded x%d" % world) if world > 1 else "1 GPU",
                      "CUDA events on the engine stream per fit (emrf_timings_t.total_ms)")
    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": step_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u32", "data": "synthetic",
        "config": cfg,
        "wall_ms_per_step": wall_ms / args.steps, "ms_each_step": [round(v, 2) for v in ev_ms],
        "clocks": clocks,
        "e2e": {"value": units / (e2e_ms / 1e3), "unit": UNIT,
                "h2d_bytes_per_step": int(np.mean([t["h2d_bytes"] for t in e2e_tm])),
                "d2h_bytes_per_step": int(np.mean([t["d2h_bytes"] for t in e2e_tm])),
                "ms_per_step": e2e_ms / args.steps, "ms_each_step": e2e_each,
                "phases_ms_per_step": {k: round(float(np.mean([t[k] for t in e2e_tm])), 3) for k in (
                    "total_ms", "h2d_ms", "find_splits_ms", "bin_ms", "bag_ms", "hist_kernel_ms", "partition_ms")}},
        "gpu_launches": int(np.sum([t["kernel_launches"] for t in tms])),
        "roofline": {"bound": "hbm", "kernel": "hist_scan_kernel (DTStatsAggregator update + binsToBestSplit)",
                     "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": frac, "traffic": traffic,
                     "peak_source": peak_src, "launches": hist_launches,
                     "avg_launch_ms": hist_ms / max(hist_launches, 1),
                     "algorithmic_bytes_per_launch": hist_bytes / max(hist_launches, 1),
                     "algorithmic_bytes": "(F_sub + 6) B per in-bag (row, tree, level) + F_sub*B*C*4 B per node (SURVEY 8d)",
                     "entry_bytes_only_frac": float(np.sum([t["hist_units"] for t in tms])) * (8 + 6) / (hist_ms / 1e3) / 1e9 / peak,
                     "share_of_step": hist_ms / step_ms},
        "phases_ms_per_step": {k: float(np.mean([t[k] for t in tms])) for k in (
            "find_splits_ms", "bin_ms", "bag_ms", "hist_kernel_ms", "partition_ms", "host_scheduler_ms",
            "host_select_ms", "host_update_ms", "host_export_ms", "bag_kernel_ms", "comm_ms")},
        "comm_bytes_per_step": int(np.mean([t["comm_bytes"] for t in tms])),
        "groups_per_step": int(tms[0]["num_groups"]), "total_nodes": int(model.totalNumNodes),
    }
    out.update(parity)
    if world == 1 and not args.no_cpu_baseline:
        v, dt, cores, oexp = time_oracle(CPU_SAMPLE_TREES)
        out["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                               "sample": f"oracle port (C++/OpenMP) on the same {N_ROWS} rows x {N_FEATURES} features, depth "
                                         f"{MAX_DEPTH}, {CPU_SAMPLE_TREES} of {TREES_PER_GPU} trees: {dt:.1f} s"}
        eq, field = forests_equal(exp, oexp)     # the timed forest (all 100 trees, 1M rows) against the oracle's
        out["parity_vs_oracle"] = bool(eq)
        if not eq:
            out["parity_first_difference"] = field
    if configs2 is not None:
        out["configs2"] = configs2
    if configs3 is not None:
        out["configs3"] = configs3
    if predict is not None:
        out["predict"] = predict
    print(json.dumps(out), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def run_configs2(ctx, world, rank, local_rank, barrier, allreduce_max, estimator, shard_kw, peak):
    """BASELINE configs[2]: RF 500 trees, depth 12, 10M x 128, 8192 partitions.  One warm-up and one timed fit per mode
    (a fit is seconds long); features generated in HBM by emrf_synth_flat.  N > 1: the rows (then the trees) of the SAME
    forest are sharded (strong scaling), and rank 0 fits it alone as well: efficiency = t_1 / (N * t_N), forests compared."""
    from euromillioner_b200 import synth
    from euromillioner_b200.classifier import Context, Dataset

    def timed_fit(c, ds, trees, reps=1):
        rf2 = estimator(c, trees, C2_DEPTH)
        m = rf2.fit(ds)                      # warm-up: buffers of this shape are allocated and faulted in
        del m                                # a model of this shape is ~1 GB: its host blocks go back to the engine's cache
        barrier()
        best, tm_best = None, None
        for _ in range(reps):
            m = rf2.fit(ds)
            tm = c.timings()
            ms = allreduce_max([tm["total_ms"]])[0]
            if best is None or ms < best:
                best, tm_best = ms, tm
        barrier()
        return m, best, tm_best

    def block(ms, tm, n_gpus):
        hist_ms, _, achieved, frac = hist_roofline([tm], peak)
        return {"ms_per_fit": round(ms, 2), "rows_trees_per_sec": C2_ROWS * C2_TREES / (ms / 1e3), "n_gpus": n_gpus,
                "hist_kernel_ms": round(hist_ms, 2), "hist_achieved_gbs": round(achieved, 1), "hist_frac_of_hbm_peak": round(frac, 4),
                "partition_ms": round(tm["partition_ms"], 2), "comm_ms": round(tm["comm_ms"], 2), "comm_bytes": int(tm["comm_bytes"]),
                "host_export_ms": round(tm["host_export_ms"], 2),
                "groups": int(tm["num_groups"]), "host_scheduler_ms": round(tm["host_scheduler_ms"], 1)}

    out = {"workload": f"BASELINE.json configs[2]: RF {C2_TREES} trees, depth {C2_DEPTH}, maxBins {MAX_BINS}, {C2_ROWS} rows x "
                       f"{C2_FEATURES} features, {N_CLASSES}-class label, {C2_PARTS} partitions; one warm-up + one timed fit per mode"}
    one_exp, t1 = None, None
    if rank == 0:
        c1 = ctx if world == 1 else Context(local_rank)
        g = synth.DeviceFlat(c1, C2_ROWS, C2_FEATURES, N_CLASSES)
        ds = Dataset(g.X.value, g.y.value, num_partitions=C2_PARTS, n_rows=C2_ROWS, n_features=C2_FEATURES)
        # no barrier inside: the other ranks wait below
        rf2 = estimator(c1, C2_TREES, C2_DEPTH)
        rf2.fit(ds)
        m = rf2.fit(ds)
        tm = c1.timings()
        t1 = tm["total_ms"]
        out["one_gpu"] = block(t1, tm, 1)
        out["one_gpu"]["total_nodes"] = int(m.totalNumNodes)
        if world > 1:
            one_exp = m.export()
        del m
        g.free()
        if world > 1:
            c1.close()
    if world == 1:
        return out
    barrier()
    t1 = allreduce_max([t1 if rank == 0 else 0.0])[0]
    # ---- row-sharded: contiguous rows at partition boundaries
    rows_r, parts_r = C2_ROWS // world, C2_PARTS // world
    g = synth.DeviceFlat(ctx, rows_r, C2_FEATURES, N_CLASSES, row_start=rank * rows_r)
    ds = Dataset(g.X.value, g.y.value, n_rows=rows_r, n_features=C2_FEATURES, **shard_kw(rows_r, parts_r))
    m, ms, tm = timed_fit(ctx, ds, C2_TREES)
    if rank == 0:
        out["row_sharded"] = block(ms, tm, world)
        out["row_sharded"]["efficiency_vs_one_gpu"] = round(t1 / (world * ms), 3)
        out["row_sharded"]["parity_vs_1gpu"] = bool(forests_equal(m.export(), one_exp)[0])
    del m
    g.free()
    barrier()
    # ---- tree-sharded: every rank holds all rows, trees t % world == rank are built here
    g = synth.DeviceFlat(ctx, C2_ROWS, C2_FEATURES, N_CLASSES)
    ds = Dataset(g.X.value, g.y.value, num_partitions=C2_PARTS, n_rows=C2_ROWS, n_features=C2_FEATURES)
    m, ms, tm = timed_fit(ctx, ds, C2_TREES)
    if rank == 0:
        out["tree_sharded"] = block(ms, tm, world)
        out["tree_sharded"]["efficiency_vs_one_gpu"] = round(t1 / (world * ms), 3)
        out["tree_sharded"]["parity_vs_1gpu"] = bool(forests_equal(m.export(), one_exp)[0])
    del m
    g.free()
    barrier()
    return out


def run_configs3(ctx, world, rank, barrier, allreduce_max):
    """BASELINE configs[3]: all 7 per-ball models (balls 0-4: 50 classes, balls 5-6: 12 classes), 200 trees each, depth 6, on
    the 1 490 lag-window rows of 1 500 synthetic draws.  Many-model / small-data: model-parallel, ball b is fitted by rank
    b % N with its own context and RNG, no collective.  Features are built on the device from the draws
    table (emrf_build_lag_features); rank 0 checks its first model against the CPU oracle."""
    from euromillioner_b200 import _lib, synth
    from euromillioner_b200.classifier import Context, Dataset, RandomForestClassifier
    c = Context(ctx.device) if world > 1 else ctx
    L = _lib.lib()
    draws = np.ascontiguousarray(synth.draws(C3_DRAWS), np.int32)
    n, F = C3_DRAWS - C3_LAG, 7 * C3_LAG
    dX, dy = C.c_void_p(), C.c_void_p()
    _lib.check(L.emrf_dev_alloc(c.h, n * F * 8, C.byref(dX)))
    _lib.check(L.emrf_dev_alloc(c.h, n * 8, C.byref(dy)))
    mine = [b for b in range(7) if b % world == rank]

    def fit_all():
        models = []
        for b in mine:
            _lib.check(L.emrf_build_lag_features(c.h, draws.ctypes.data, C3_DRAWS, C3_LAG, b, dX, dy))
            rf = (RandomForestClassifier(c).setNumTrees(C3_TREES).setMaxDepth(C3_DEPTH).setMaxBins(MAX_BINS)
                  .setNumClasses(50 if b < 5 else 12).setSeed(SEED + b))
            models.append(rf.fit(Dataset(dX.value, dy.value, num_partitions=C3_PARTS, n_rows=n, n_features=F)))
        return models

    fit_all()
    barrier()
    t0 = time.perf_counter()
    models = fit_all()
    ms = allreduce_max([(time.perf_counter() - t0) * 1e3])[0]
    barrier()
    out = {"workload": f"BASELINE.json configs[3]: 7 per-ball models x {C3_TREES} trees, depth {C3_DEPTH}, {n} rows x {F} lag features "
                       f"({C3_DRAWS} draws, {C3_LAG}-draw window), ball b on GPU b % {world}",
           "n_gpus": world, "ms_all_models": round(ms, 2), "models_per_sec": 7 / (ms / 1e3),
           "rows_trees_per_sec": 7 * n * C3_TREES / (ms / 1e3)}
    if rank == 0:
        from oracle import binding as ob
        Xh, yh, Cn = synth.draw_history(C3_DRAWS, C3_LAG, mine[0])
        o = ob.fit(ob.make_params(num_trees=C3_TREES, max_depth=C3_DEPTH, max_bins=MAX_BINS, num_classes=Cn, seed=SEED + mine[0]),
                   Xh, yh, ob.partition_offsets(len(Xh), C3_PARTS)).export()
        out["parity_vs_oracle_first_model"] = bool(forests_equal(models[0].export(), o)[0])
        out["nodes_first_model"] = int(models[0].totalNumNodes)
    del models
    L.emrf_dev_free(c.h, dX)
    L.emrf_dev_free(c.h, dy)
    if world > 1:
        c.close()
    return out


def run_predict(ctx, world, rank, barrier, allreduce_max, estimator):
    """Batch prediction, BASELINE configs[4]: 100M rows x 128 features through a 500-tree depth-12 forest; every rank takes
    100M / N rows (rows are independent: no collective) in chunks of 10M generated in HBM (a chunk's f64 features are
    10 GB), labels copied to the host.  The forest is fitted on 200k rows (what bounds predict is the number of node
    visits and leaf records read, not how the forest was made).  Timed: the predict calls only, not the generator."""
    from euromillioner_b200 import synth
    from euromillioner_b200.classifier import Context, Dataset
    c = Context(ctx.device) if world > 1 else ctx      # predict needs no communicator
    gfit = synth.DeviceFlat(c, 200_000, C2_FEATURES, N_CLASSES)
    m = estimator(c, C2_TREES, C2_DEPTH).fit(Dataset(gfit.X.value, gfit.y.value, num_partitions=64, n_rows=200_000,
                                                     n_features=C2_FEATURES))
    gfit.free()
    rows_rank = C4_ROWS_TOTAL // world
    ms_local, done, first = 0.0, 0, True
    while done < rows_rank:
        n = min(C4_ROWS_CHUNK, rows_rank - done)
        g = synth.DeviceFlat(c, n, C2_FEATURES, N_CLASSES, row_start=rank * rows_rank + done)
        ds = Dataset(g.X.value, None, n_rows=n, n_features=C2_FEATURES)
        if first:
            m.transform(ds, raw=False, probability=False)   # warm-up (forest upload, buffers)
            barrier()
            first = False
        t0 = time.perf_counter()
        m.transform(ds, raw=False, probability=False)
        ms_local += (time.perf_counter() - t0) * 1e3
        g.free()
        done += n
    ms = allreduce_max([ms_local])[0]
    barrier()
    total = rows_rank * world
    out = {"workload": f"BASELINE.json configs[4]: {total} rows x {C2_FEATURES} features through a {C2_TREES}-tree depth-{C2_DEPTH} forest "
                       f"({int(m.totalNumNodes)} nodes, fitted on 200000 rows); {rows_rank} rows per GPU in resident chunks of "
                       f"{C4_ROWS_CHUNK}, predicted labels copied to the host",
           "n_gpus": world, "ms": round(ms, 2), "rows_per_sec": total / (ms / 1e3),
           "tree_walks_per_sec": total * C2_TREES / (ms / 1e3),
           "input_gbs": total * C2_FEATURES * 8 / (ms / 1e3) / 1e9}
    if world > 1:
        c.close()
    return out


if __name__ == "__main__":
    main()
