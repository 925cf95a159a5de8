#!/usr/bin/env python
"""Diff a directory written by tools/SparkGolden.java (a real Spark 2.4.6 run) against this engine — the step that would
turn "parity unpinned" into "parity pinned".  Needs a B200 for the engine side; `--oracle` compares the CPU oracle instead.

    python tools/compare_spark_golden.py /tmp/golden_cfg1 [--oracle]

Reads params.txt (the SparkGolden command line), rebuilds the same synthetic Dataset (euromillioner_b200/synth.py), fits
with the same parameters, seed and partition count, and compares: toDebugString byte for byte (after substituting
Spark's uid), every node's prediction / impurity / gain / split / class counts, the first 64 rows' rawPrediction /
probability / prediction, and the feature importances.  Exit status 0 = identical (impurities, gains and importances may
differ in the last ulp only if the impurity is entropy — Math.log is not StrictMath)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from euromillioner_b200 import synth  # noqa: E402
from euromillioner_b200.classifier import debug_string  # noqa: E402


def main():
    d = sys.argv[1]
    use_oracle = "--oracle" in sys.argv[2:]
    with open(os.path.join(d, "params.txt")) as f:
        args = f.readline().split()
        uid = f.readline().split()[1]
    if args[0] == "draws":
        n_draws, lag, ball, trees, depth, parts, seed = [int(v) for v in args[1:8]]
        X, y, n_classes = synth.draw_history(n_draws, lag, ball)
    else:
        n, F, n_classes, trees, depth, parts, seed = [int(v) for v in args[1:8]]
        X, y = synth.flat(n, F, n_classes)
    if use_oracle:
        from oracle import binding as ob
        fo = ob.fit(ob.make_params(num_trees=trees, max_depth=depth, num_classes=n_classes, seed=seed), X, y,
                    ob.partition_offsets(len(X), parts))
        e = fo.export()
        raw, prob, pred = fo.predict(X[:64])
        imp = None
    else:
        from euromillioner_b200.classifier import Dataset, RandomForestClassifier
        m = (RandomForestClassifier().setNumTrees(trees).setMaxDepth(depth).setMaxBins(32).setSeed(seed).setNumClasses(n_classes)
             .fit(Dataset(X, y, num_partitions=parts)))
        e = m.export()
        cols = m.transform(Dataset(X[:64])).columns
        raw, prob, pred = cols["rawPrediction"], cols["probability"], cols["prediction"]
        imp = m.featureImportances
    bad = 0
    with open(os.path.join(d, "debug_string.txt")) as f:
        want = f.read()
    got = debug_string(e, trees, uid)
    if got != want:
        bad += 1
        gl, wl = got.splitlines(), want.splitlines()
        k = next((i for i in range(min(len(gl), len(wl))) if gl[i] != wl[i]), min(len(gl), len(wl)))
        print(f"toDebugString differs at line {k + 1}:\n  spark : {wl[k] if k < len(wl) else '<end>'}\n  engine: {gl[k] if k < len(gl) else '<end>'}")
    else:
        print(f"toDebugString identical ({len(want)} bytes)")
    rows = [ln.split() for ln in open(os.path.join(d, "nodes.txt"))]
    if len(rows) != len(e["left"]):
        bad += 1
        print(f"node count differs: spark {len(rows)}, engine {len(e['left'])}")
    else:
        fields = {"prediction": 2, "impurity": 3, "gain": 4, "feature": 5, "threshold": 6}
        for name, col in fields.items():
            sp = np.array([float(r[col]) for r in rows])
            mine = np.asarray(e[name], np.float64)
            if name == "feature":
                mine = np.where(e["left"] < 0, -1, mine)
            if name in ("gain", "threshold"):
                mine = np.where(e["left"] < 0, 0.0, mine)
            nd = int(np.sum(sp != mine))
            bad += nd > 0
            print(f"{name}: {nd} of {len(sp)} nodes differ")
        st = np.array([[float(v) for v in r[7:]] for r in rows])
        nd = int(np.sum(np.any(st != e["stats"], axis=1)))
        bad += nd > 0
        print(f"class counts: {nd} of {len(rows)} nodes differ")
    for i, ln in enumerate(open(os.path.join(d, "predictions.txt"))):
        a, b, c = ln.split("|")
        ok = (np.array_equal(np.array(a.split(), float), raw[i]) and np.array_equal(np.array(b.split(), float), prob[i])
              and float(c) == pred[i])
        if not ok:
            bad += 1
            print(f"row {i}: rawPrediction / probability / prediction differ")
    if imp is not None:
        sp = np.array(open(os.path.join(d, "feature_importances.txt")).read().split(), float)
        print("featureImportances:", "identical" if np.array_equal(sp, imp) else f"max abs diff {np.max(np.abs(sp - imp)):.3e}")
        bad += not np.allclose(sp, imp, rtol=1e-12, atol=0)
    print("PARITY PINNED: identical to the Spark run" if bad == 0 else f"{bad} differences")
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
