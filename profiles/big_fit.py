"""Large-shape fits on one GPU (BASELINE configs[2] shape: 10M x 128, depth 12) — size-independent checks
and phase timers.  Usage: python profiles/big_fit.py [rows] [features] [trees] [depth] [partitions]"""
import ctypes as C
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from euromillioner_b200 import _lib, synth  # noqa: E402
from euromillioner_b200.classifier import Context, Dataset, RandomForestClassifier  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
F = int(sys.argv[2]) if len(sys.argv) > 2 else 128
T = int(sys.argv[3]) if len(sys.argv) > 3 else 50
D = int(sys.argv[4]) if len(sys.argv) > 4 else 12
P = int(sys.argv[5]) if len(sys.argv) > 5 else 8192
ctx = Context(0)
L = _lib.lib()
t0 = time.time()
g = synth.DeviceFlat(ctx, N, F, 50)        # generated in HBM by the engine (emrf_synth_flat)
dX, dy = g.X, g.y
print(f"data ready in {time.time() - t0:.1f} s", flush=True)
rf = RandomForestClassifier(ctx).setNumTrees(T).setMaxDepth(D).setNumClasses(50)
for rep in range(2):
    t0 = time.time()
    m = rf.fit(Dataset(dX.value, dy.value, num_partitions=P, n_rows=N, n_features=F))
    tm = ctx.timings()
    print(f"fit {rep}: wall {time.time() - t0:.3f} s, rows*trees/s {N * T / (tm['total_ms'] / 1e3):.3e}, nodes {m.totalNumNodes}")
    print({k: (round(v, 2) if isinstance(v, float) else v) for k, v in tm.items()}, flush=True)
e = m.export()
for t in range(m.numTrees):
    a, b = e["tree_off"][t], e["tree_off"][t + 1]
    Lc, Rc, st = e["left"][a:b], e["right"][a:b], e["stats"][a:b]
    internal = np.where(Lc >= 0)[0]
    assert np.array_equal(st[internal], st[Lc[internal]] + st[Rc[internal]]), "class counts not conserved"
roots = e["stats"][e["tree_off"][:-1]].sum(axis=1)
print("root weight sums / N:", float(roots.min() / N), float(roots.max() / N), "counts conserved in every tree: OK")
