"""Batch prediction timing (BASELINE configs[4] shape and the 1M x 100-tree case), features resident in HBM.
Usage: python profiles/predict_bench.py   — prints rows/s for both; the forest is fitted on 200k generated rows."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from euromillioner_b200 import synth  # noqa: E402
from euromillioner_b200.classifier import Context, Dataset, RandomForestClassifier  # noqa: E402

ctx = Context(0)
for (n, F, T, D) in ((1_000_000, 64, 100, 10), (10_000_000, 128, 500, 12)):
    gfit = synth.DeviceFlat(ctx, 200_000, F, 50)
    m = RandomForestClassifier(ctx).setNumTrees(T).setMaxDepth(D).setNumClasses(50).fit(
        Dataset(gfit.X.value, gfit.y.value, num_partitions=64, n_rows=200_000, n_features=F))
    gfit.free()
    g = synth.DeviceFlat(ctx, n, F, 50, row_start=200_000)
    ds = Dataset(g.X.value, None, n_rows=n, n_features=F)
    m.transform(ds, raw=False, probability=False)
    best = 1e9
    for _ in range(3):
        t0 = time.perf_counter()
        m.transform(ds, raw=False, probability=False)
        best = min(best, time.perf_counter() - t0)
    print(f"{n} rows x {F} features, {T} trees depth {D} ({m.totalNumNodes} nodes): {best * 1e3:.1f} ms, {n / best:.3e} rows/s, "
          f"{n * T / best:.3e} tree walks/s", flush=True)
    g.free()
