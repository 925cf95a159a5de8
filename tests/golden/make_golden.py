"""Generates tests/golden/cfg1_forest.json from the CPU oracle (run from the repo root).
No reference implementation of this path can run here (no JVM), so the fixture pins
the oracle's own output for BASELINE.json configs[0]."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from euromillioner_b200 import synth  # noqa: E402
from oracle import binding as ob  # noqa: E402

X, y, Cn = synth.draw_history()
p = ob.make_params(num_trees=20, max_depth=6, max_bins=32, num_classes=Cn)
fo = ob.fit(p, X, y, ob.partition_offsets(len(X), 8))
e = fo.export()
raw, prob, pred = fo.predict(X[:64])
out = {"tree_off": [int(v) for v in e["tree_off"]], "group_sizes": [int(v) for v in fo.group_sizes()],
       "feature_head": [int(v) for v in e["feature"][:200]], "threshold_head": [float(v) for v in e["threshold"][:200]],
       "gain_sum": float(e["gain"].sum()), "pred_head": [int(v) for v in pred]}
with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "cfg1_forest.json"), "w") as f:
    json.dump(out, f)
print("wrote cfg1_forest.json", out["tree_off"][-1], "nodes")


# ---- toDebugString of the first three trees, rendered here straight from the oracle's export with its own small
# recursive printer (NOT the product's renderer): Spark 2.4's Node.subtreeToString / TreeEnsembleModel.toDebugString
def jstr(v):
    v = float(v)
    assert (1e-3 <= abs(v) < 1e7 or v == 0.0) and v == v   # every number of this forest prints in plain decimal notation
    r = repr(v)
    return r if "." in r else r + ".0"


def subtree(base, i, indent):
    pad = " " * indent
    if e["left"][base + i] < 0:
        return f"{pad}Predict: {jstr(e['prediction'][base + i])}\n"
    f, t = int(e["feature"][base + i]), jstr(e["threshold"][base + i])
    return (f"{pad}If (feature {f} <= {t})\n" + subtree(base, int(e["left"][base + i]), indent + 1) +
            f"{pad}Else (feature {f} > {t})\n" + subtree(base, int(e["right"][base + i]), indent + 1))


txt = "RandomForestClassificationModel (uid=rfc_golden) with 3 trees\n"
for t in range(3):
    txt += f"  Tree {t} (weight 1.0):\n" + subtree(int(e["tree_off"][t]), 0, 4)
with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "cfg1_debug_string_3trees.txt"), "w") as f:
    f.write(txt)
print("wrote cfg1_debug_string_3trees.txt", len(txt), "bytes")
