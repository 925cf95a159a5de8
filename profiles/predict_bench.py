"""Batch-predict throughput (BASELINE configs[4] shape scaled to one fit here): rows/s through emrf_predict
with device-resident features.  Usage: python profiles/predict_bench.py [rows] [trees] [depth] [features]"""
import ctypes as C
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from euromillioner_b200 import _lib, synth  # noqa: E402
from euromillioner_b200.classifier import Context, Dataset, RandomForestClassifier  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
T = int(sys.argv[2]) if len(sys.argv) > 2 else 100
D = int(sys.argv[3]) if len(sys.argv) > 3 else 10
F = int(sys.argv[4]) if len(sys.argv) > 4 else 64
ctx = Context(0)
L = _lib.lib()
X, y = synth.flat(N, F, 50)
model = RandomForestClassifier(ctx).setNumTrees(T).setMaxDepth(D).setNumClasses(50).fit(Dataset(X, y, num_partitions=1024))
print("forest:", model.totalNumNodes, "nodes", flush=True)
dX, dP, dQ = C.c_void_p(), C.c_void_p(), C.c_void_p()
_lib.check(L.emrf_dev_alloc(ctx.h, X.nbytes, C.byref(dX)))
_lib.check(L.emrf_dev_alloc(ctx.h, N * 8, C.byref(dP)))
_lib.check(L.emrf_dev_alloc(ctx.h, N * 50 * 8, C.byref(dQ)))
_lib.check(L.emrf_memcpy_h2d(ctx.h, dX, X.ctypes.data, X.nbytes))
for name, prob in (("labels only", None), ("labels + probability", dQ)):
    for rep in range(3):
        t0 = time.perf_counter()
        _lib.check(L.emrf_predict(ctx.h, model._h, dX, N, None, prob, dP))
        dt = time.perf_counter() - t0
    print(f"{name}: {dt * 1e3:.2f} ms, {N / dt:.3e} rows/s, {N * T / dt:.3e} tree-walks/s", flush=True)
pred = np.zeros(N)
_lib.check(L.emrf_memcpy_d2h(ctx.h, pred.ctypes.data, dP, N * 8))
print("train accuracy", float((pred == y).mean()))
