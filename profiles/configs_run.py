"""Runs the non-bench BASELINE configs once for the record (profiles/r1_configs.md is written from its output):
  cfg1  20 trees depth 6 on 1 490 draw-history rows (ball 1)            fit
  cfg4  7 per-ball models x 200 trees depth 6 on the same rows           fit (sequential on one GPU)
  cfg3' 10M x 128, depth 12, T trees (default 100)                      fit, then predict 10M rows (cfg5 shape)
Usage: python profiles/configs_run.py [trees_for_cfg3]"""
import ctypes as C
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from euromillioner_b200 import _lib, synth  # noqa: E402
from euromillioner_b200.classifier import Context, Dataset, RandomForestClassifier  # noqa: E402

ctx = Context(0)
L = _lib.lib()
T3 = int(sys.argv[1]) if len(sys.argv) > 1 else 100

# ---- cfg1
X, y, Cn = synth.draw_history()
rf = RandomForestClassifier(ctx).setNumTrees(20).setMaxDepth(6).setNumClasses(Cn)
for _ in range(3):
    t0 = time.perf_counter(); m = rf.fit(Dataset(X, y, num_partitions=8)); dt = time.perf_counter() - t0
# (parity of this config against the oracle: tests/test_gpu_fit.py::test_cfg1_draw_history_ball1 — the oracle is
# test infrastructure and is not used from here)
print(f"cfg1: GPU fit {dt * 1e3:.2f} ms ({len(X) * 20 / dt:.3e} rows*trees/s), groups {ctx.timings()['num_groups']}, "
      f"nodes {m.totalNumNodes}", flush=True)

# ---- cfg4
t0 = time.perf_counter()
nodes = 0
for ball in range(7):
    Xb, yb, Cb = synth.draw_history(ball=ball)
    mb = RandomForestClassifier(ctx).setNumTrees(200).setMaxDepth(6).setNumClasses(Cb).setSeed(ball).fit(Dataset(Xb, yb, num_partitions=8))
    nodes += mb.totalNumNodes
dt = time.perf_counter() - t0
print(f"cfg4: 7 models x 200 trees: {dt * 1e3:.1f} ms incl. synthetic data ({7 * 200 * len(X) / dt:.3e} rows*trees/s), {nodes} nodes", flush=True)

# ---- cfg3-shaped fit + cfg5-shaped predict
N, F = 10_000_000, 128
dX, dy, dP = C.c_void_p(), C.c_void_p(), C.c_void_p()
_lib.check(L.emrf_dev_alloc(ctx.h, N * F * 8, C.byref(dX)))
_lib.check(L.emrf_dev_alloc(ctx.h, N * 8, C.byref(dy)))
_lib.check(L.emrf_dev_alloc(ctx.h, N * 8, C.byref(dP)))
for a in range(0, N, 500_000):
    b = min(N, a + 500_000)
    Xc, yc = synth.flat(b - a, F, 50, row_start=a)
    _lib.check(L.emrf_memcpy_h2d(ctx.h, C.c_void_p(dX.value + a * F * 8), Xc.ctypes.data, Xc.nbytes))
    _lib.check(L.emrf_memcpy_h2d(ctx.h, C.c_void_p(dy.value + a * 8), yc.ctypes.data, yc.nbytes))
rf = RandomForestClassifier(ctx).setNumTrees(T3).setMaxDepth(12).setNumClasses(50)
t0 = time.perf_counter()
m = rf.fit(Dataset(dX.value, dy.value, num_partitions=8192, n_rows=N, n_features=F))
dt = time.perf_counter() - t0
tm = ctx.timings()
print(f"cfg3 shape, {T3} trees: fit {tm['total_ms']:.0f} ms (wall {dt * 1e3:.0f}), {N * T3 / (tm['total_ms'] / 1e3):.3e} rows*trees/s, nodes {m.totalNumNodes}, "
      f"hist {tm['hist_kernel_ms']:.0f} ms ({tm['hist_alg_bytes'] / tm['hist_kernel_ms'] / 1e6:.0f} GB/s algorithmic), partition {tm['partition_ms']:.0f}, "
      f"bag kernel {tm['bag_kernel_ms']:.0f}, host {tm['host_scheduler_ms']:.0f}, groups {tm['num_groups']}", flush=True)
for rep in range(2):
    t0 = time.perf_counter()
    _lib.check(L.emrf_predict(ctx.h, m._h, dX, N, None, None, dP))
    dt = time.perf_counter() - t0
print(f"cfg5 shape: predict {N} rows x {T3} trees depth 12: {dt * 1e3:.0f} ms, {N / dt:.3e} rows/s, {N * T3 / dt:.3e} tree-walks/s", flush=True)
