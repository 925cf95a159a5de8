"""tools/compare_spark_golden.py reads what tools/SparkGolden.java writes.  No JVM exists here, so this test writes a
directory in SparkGolden's format from the CPU oracle's forest (numbers in java.lang.Double.toString form) and checks
that the comparison tool parses it and reports identity — and a difference once a threshold is edited."""
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _write_golden_dir(oracle, d, n_draws=400, trees=3, depth=4, parts=4, seed=207336481):
    from euromillioner_b200 import synth
    from euromillioner_b200.classifier import debug_string, java_double_to_string as j
    X, y, C = synth.draw_history(n_draws, 10, 0)
    fo = oracle.fit(oracle.make_params(num_trees=trees, max_depth=depth, num_classes=C, seed=seed), X, y,
                    oracle.partition_offsets(len(X), parts))
    e = fo.export()
    os.makedirs(d, exist_ok=True)
    with open(os.path.join(d, "params.txt"), "w") as f:
        f.write(f"draws {n_draws} 10 0 {trees} {depth} {parts} {seed} {d}\nuid rfc_0123456789ab\n")
    with open(os.path.join(d, "debug_string.txt"), "w") as f:
        f.write(debug_string(e, trees, "rfc_0123456789ab"))
    with open(os.path.join(d, "nodes.txt"), "w") as f:
        for t in range(trees):
            for i in range(int(e["tree_off"][t]), int(e["tree_off"][t + 1])):
                leaf = e["left"][i] < 0
                head = (f"{t} {'L' if leaf else 'I'} {j(e['prediction'][i])} {j(e['impurity'][i])} " +
                        ("0.0 -1 0.0" if leaf else f"{j(e['gain'][i])} {int(e['feature'][i])} {j(e['threshold'][i])}"))
                f.write(head + " " + " ".join(j(v) for v in e["stats"][i]) + "\n")
    raw, prob, pred = fo.predict(X[:64])
    with open(os.path.join(d, "predictions.txt"), "w") as f:
        for i in range(len(pred)):
            f.write(" ".join(j(v) for v in raw[i]) + " | " + " ".join(j(v) for v in prob[i]) + " | " + j(pred[i]) + "\n")
    with open(os.path.join(d, "feature_importances.txt"), "w") as f:
        f.write("\n")


def test_compare_tool_reads_the_spark_golden_format(oracle, tmp_path):
    d = str(tmp_path / "golden")
    _write_golden_dir(oracle, d)
    tool = os.path.join(ROOT, "tools", "compare_spark_golden.py")
    r = subprocess.run([sys.executable, tool, d, "--oracle"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "PARITY PINNED" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
    # one edited threshold must be reported
    p = os.path.join(d, "nodes.txt")
    lines = open(p).read().splitlines()
    k = next(i for i, ln in enumerate(lines) if ln.split()[1] == "I")
    parts = lines[k].split()
    parts[6] = "123.5"
    lines[k] = " ".join(parts)
    open(p, "w").write("\n".join(lines) + "\n")
    r = subprocess.run([sys.executable, tool, d, "--oracle"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 1 and "threshold: 1 of" in r.stdout, r.stdout[-2000:]


def test_java_host_and_jni_shim_agree():
    """The Java host is source only (no JDK here): every `native` method of Native.java must have its JNI function in
    emrf_jni.cpp, and the shim must type-check against include/emrf.h (compiled with a stand-in jni.h)."""
    import re
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    java = open(os.path.join(root, "euromillioner_b200/host/java/com/euromillioner/gpu/Native.java")).read()
    shim_path = os.path.join(root, "euromillioner_b200/host/jni/emrf_jni.cpp")
    shim = open(shim_path).read()
    natives = re.findall(r"static native [\w\[\]]+ (\w+)\(", java)
    assert len(natives) >= 15
    for name in natives:
        assert f"Java_com_euromillioner_gpu_Native_{name}(" in shim, name
    defined = re.findall(r"Java_com_euromillioner_gpu_Native_(\w+)\(", shim)
    assert sorted(set(defined)) == sorted(set(natives))
    # every Native.* call in the other classes names a declared method
    for cls in ("GpuContext", "GpuRandomForestClassifier", "GpuRandomForestClassificationModel"):
        src = open(os.path.join(root, f"euromillioner_b200/host/java/com/euromillioner/gpu/{cls}.java")).read()
        for used in set(re.findall(r"Native\.(\w+)\(", src)):
            assert used in natives, (cls, used)
    stub = os.path.join(root, "tests", "golden", "jni_stub")
    r = subprocess.run(["g++", "-std=c++17", "-fsyntax-only", "-I", stub, "-I", os.path.join(root, "include"), shim_path],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]
