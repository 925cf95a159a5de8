"""Where the cycles of hist_scan_kernel go, phase by phase (debug build: make -C euromillioner_b200/csrc prof).
Usage: EMRF_LIB_PATH=euromillioner_b200/libemrf_prof.so python profiles/hist_phases.py
Thread 0 of every CTA adds clock64() deltas per phase into a device array; the sums are printed per fit."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from euromillioner_b200 import _lib, synth  # noqa: E402
from euromillioner_b200.classifier import Context, Dataset, RandomForestClassifier  # noqa: E402

ctx = Context(0)
L = _lib.lib()
X, y = synth.flat(bench.N_ROWS, bench.N_FEATURES, bench.N_CLASSES)
rf = (RandomForestClassifier(ctx).setNumTrees(bench.TREES_PER_GPU).setMaxDepth(bench.MAX_DEPTH).setMaxBins(bench.MAX_BINS)
      .setNumClasses(bench.N_CLASSES))
names = ["prologue", "zero", "accumulate", "flush/ticket", "prefix", "totals+parent", "candidates", "argmax+record"]
buf = (C.c_ulonglong * 16)()
for it in range(3):
    rf.fit(Dataset(X, y, num_partitions=bench.N_PARTS))
    L.emrf_debug_hist_prof(buf)
    tot = sum(buf[i] for i in range(8))
    print(f"fit {it}: CTAs {buf[8]}, scanning CTAs {buf[9]}, cycles/CTA {tot / max(buf[8], 1):.0f}")
    print("  " + " | ".join(f"{n} {100.0 * buf[i] / tot:.1f}% ({buf[i] / max(buf[9] if i >= 4 else buf[8], 1):.0f} cyc)"
                            for i, n in enumerate(names)))
    bb = (C.c_ulonglong * 8)()
    L.emrf_debug_bag_prof(bb)
    nb = max(bb[5], 1)
    print("  bag kernel, cycles per block of 32 uniforms: producer wait %.0f, well_block %.0f | consumer wait %.0f, poisson_block %.0f, "
          "emit+flush %.0f  (%d blocks)" % (bb[0] / nb, bb[1] / nb, bb[2] / nb, bb[3] / nb, bb[4] / nb, bb[5]))

