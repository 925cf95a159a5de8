"""Host->device rate of the configs[1] feature matrix (512 MB) on this box, to place the e2e leg.  A: torch copy from the
engine's pinned buffer, idle.  B: emrf_memcpy_h2d.  C: the same torch copy while a device-resident fit runs in another
thread.  D: while only the host threads of this process spin."""
import ctypes as C
import os
import sys
import threading
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from euromillioner_b200 import _lib, synth  # noqa: E402
from euromillioner_b200.classifier import Context, Dataset, RandomForestClassifier  # noqa: E402

ctx = Context(0)
L = _lib.lib()
X, y = synth.flat(bench.N_ROWS, bench.N_FEATURES, bench.N_CLASSES)
hX, dX, dy = C.c_void_p(), C.c_void_p(), C.c_void_p()
_lib.check(L.emrf_host_alloc_pinned(X.nbytes, C.byref(hX)))
pX = np.ctypeslib.as_array(C.cast(hX, C.POINTER(C.c_double)), shape=X.shape)
pX[...] = X
_lib.check(L.emrf_dev_alloc(ctx.h, X.nbytes, C.byref(dX)))
_lib.check(L.emrf_dev_alloc(ctx.h, y.nbytes, C.byref(dy)))
hy = np.ascontiguousarray(y)
_lib.check(L.emrf_memcpy_h2d(ctx.h, dX, hX, X.nbytes))
_lib.check(L.emrf_memcpy_h2d(ctx.h, dy, hy.ctypes.data_as(C.c_void_p), y.nbytes))
tX = torch.from_numpy(pX).view(-1)
print("torch sees the buffer as pinned:", tX.is_pinned())
dst = torch.empty_like(tX, device="cuda")
s_copy = torch.cuda.Stream()
n = tX.numel()


def copy_ms(chunks=12):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(s_copy):
        e0.record()
        per = (n + chunks - 1) // chunks
        for i in range(0, n, per):
            dst[i:i + per].copy_(tX[i:i + per], non_blocking=True)
        e1.record()
    s_copy.synchronize()
    return e0.elapsed_time(e1)


for _ in range(3):
    ms = copy_ms()
print(f"A idle, torch copy of the engine's pinned buffer: {ms:.3f} ms {X.nbytes / ms / 1e6:.1f} GB/s")
t0 = time.perf_counter()
_lib.check(L.emrf_memcpy_h2d(ctx.h, dX, hX, X.nbytes))
ms = (time.perf_counter() - t0) * 1e3
print(f"B emrf_memcpy_h2d: {ms:.3f} ms {X.nbytes / ms / 1e6:.1f} GB/s")

rf = (RandomForestClassifier(ctx).setNumTrees(bench.TREES_PER_GPU).setMaxDepth(bench.MAX_DEPTH).setMaxBins(bench.MAX_BINS)
      .setNumClasses(bench.N_CLASSES))
ds = Dataset(dX.value, dy.value, num_partitions=bench.N_PARTS, n_rows=bench.N_ROWS, n_features=bench.N_FEATURES)
rf.fit(ds); rf.fit(ds)
stop = False


def fits():
    while not stop:
        rf.fit(ds)


th = threading.Thread(target=fits)
th.start()
time.sleep(0.2)
res = [copy_ms() for _ in range(20)]
stop = True
th.join()
print("C during device-resident fits: ms", [round(r, 2) for r in res])
res = [copy_ms() for _ in range(5)]
print("D after the fits (host threads idle/spinning): ms", [round(r, 2) for r in res])
