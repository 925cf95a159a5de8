"""Small end-to-end run for compute-sanitizer (memcheck / racecheck): fit + transform on modest shapes that
exercise every kernel (multi-tile and single-tile nodes, feature chunking, generic and row-tile gathers)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from euromillioner_b200 import stages, synth  # noqa: E402
from euromillioner_b200.classifier import Context, Dataset, RandomForestClassifier  # noqa: E402

ctx = Context(0)
for n, F, T, D, kw in ((30000, 64, 6, 6, {}), (12000, 150, 3, 4, {}), (9000, 40, 2, 5, {"subset": "all"})):
    X, y = synth.flat(n, F, 50)
    rf = RandomForestClassifier(ctx).setNumTrees(T).setMaxDepth(D).setNumClasses(50).setFeatureSubsetStrategy(kw.get("subset", "auto"))
    m = rf.fit(Dataset(X, y, num_partitions=16))
    p = m.transform(Dataset(X[:3000])).columns["prediction"]
    print("fit", n, F, T, D, m.totalNumNodes, float((p == y[:3000]).mean()), flush=True)
d = synth.draws(200)
Xl, yl = stages.build_lag_features(d, 10, 0, ctx)
print("lag", Xl.shape, flush=True)
