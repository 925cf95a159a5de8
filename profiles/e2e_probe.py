"""Fits of the bench workload from PINNED host buffers (the e2e leg of bench.py) with the engine's debug timers:
`EMRF_DEBUG_TIMERS=1 python profiles/e2e_probe.py [fits]`."""
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from euromillioner_b200 import _lib, synth  # noqa: E402
from euromillioner_b200.classifier import Context, Dataset, RandomForestClassifier  # noqa: E402

fits = int(sys.argv[1]) if len(sys.argv) > 1 else 5
ctx = Context(0)
L = _lib.lib()
X, y = synth.flat(bench.N_ROWS, bench.N_FEATURES, bench.N_CLASSES)
hX, hy = C.c_void_p(), C.c_void_p()
_lib.check(L.emrf_host_alloc_pinned(X.nbytes, C.byref(hX)))
_lib.check(L.emrf_host_alloc_pinned(y.nbytes, C.byref(hy)))
pX = np.ctypeslib.as_array(C.cast(hX, C.POINTER(C.c_double)), shape=X.shape)
py = np.ctypeslib.as_array(C.cast(hy, C.POINTER(C.c_double)), shape=y.shape)
pX[...] = X
py[...] = y
rf = (RandomForestClassifier(ctx).setNumTrees(bench.TREES_PER_GPU).setMaxDepth(bench.MAX_DEPTH).setMaxBins(bench.MAX_BINS)
      .setNumClasses(bench.N_CLASSES))
ds = Dataset(pX, py, num_partitions=bench.N_PARTS)
for _ in range(fits):
    m = rf.fit(ds)
    t = ctx.timings()
    print({k: round(t[k], 3) for k in ("total_ms", "h2d_ms", "find_splits_ms", "bin_ms", "bag_ms", "bag_kernel_ms",
                                        "hist_kernel_ms", "partition_ms", "host_scheduler_ms")}, file=sys.stderr, flush=True)
