"""One fit of the bench workload (BASELINE.json configs[1]) for ncu: `python profiles/profile_fit.py [fits]`.
Prints the engine's phase timers so the plain run can be compared with the profiled one."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from euromillioner_b200 import synth  # noqa: E402
from euromillioner_b200.classifier import Context, Dataset, RandomForestClassifier  # noqa: E402

fits = int(sys.argv[1]) if len(sys.argv) > 1 else 1
trees = int(sys.argv[2]) if len(sys.argv) > 2 else bench.TREES_PER_GPU
ctx = Context(0)
X, y = synth.flat(bench.N_ROWS, bench.N_FEATURES, bench.N_CLASSES)
rf = (RandomForestClassifier(ctx).setNumTrees(trees).setMaxDepth(bench.MAX_DEPTH).setMaxBins(bench.MAX_BINS)
      .setNumClasses(bench.N_CLASSES))
for _ in range(fits):
    m = rf.fit(Dataset(X, y, num_partitions=bench.N_PARTS))
    t = ctx.timings()
    print({k: (round(v, 3) if isinstance(v, float) else v) for k, v in t.items()})
