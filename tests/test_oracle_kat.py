"""Pins the CPU oracle against every known answer available for this path.
The reference holds no tests or golden vectors, so these are upstream known
answers: JDK java.util.Random, MurmurHash3_x86_32 standard vectors, Scala 2.12 HashMap iteration
orders, Spark's own RandomForestSuite vectors for findSplitsForContinuousFeature, and the
BaggedPointSuite statistical bounds.  Parity against a real Spark run stays UNPINNED."""
import ctypes as C
import json
import os

import numpy as np
import pytest

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def _i32p(a):
    return a.ctypes.data_as(C.POINTER(C.c_int32))


def test_java_util_random(oracle):
    L = oracle.lib()
    i, l, d = C.c_int32(), C.c_int64(), C.c_double()
    L.orc_java_random(C.c_int64(42), C.byref(i), C.byref(l), C.byref(d))
    assert (i.value, l.value, d.value) == (-1170105035, -5025562857975149833, 0.7275636800328681)
    L.orc_java_random(C.c_int64(0), C.byref(i), C.byref(l), C.byref(d))
    assert (i.value, l.value) == (-1155484576, -4962768465676381896)


@pytest.mark.parametrize("data,seed,want", [
    (b"", 0, 0x00000000), (b"", 1, 0x514E28B7), (b"", 0xFFFFFFFF, 0x81F16F39), (b"\0\0\0\0", 0, 0x2362F9DE),
    (b"aaaa", 0x9747B28C, 0x5A97808A), (b"Hello, world!", 0x9747B28C, 0x24884CBA),
    (b"The quick brown fox jumps over the lazy dog", 0x9747B28C, 0x2FA826CD)])
def test_murmur3(oracle, data, seed, want):
    assert oracle.lib().orc_murmur3(data, len(data), C.c_uint32(seed)) == want


@pytest.mark.parametrize("seed,hs,first_int,first_double", [
    (0, 4790425454436209838, -278860881, 0.8446490682263027),
    (1, -1498111745827213969, -796323371, 0.13385709732307427),
    (42, -3557431703312098865, 1655315539, 0.6661236774413726),
    (207336481, 2719335625380177937, -1525896691, 0.26236745326918753)])
def test_xorshift_random_v24(oracle, seed, hs, first_int, first_double):
    # derived from the Appendix A.6 restatement: cross-check, not a JVM-verified KAT
    L = oracle.lib()
    assert L.orc_xorshift_hash_seed(seed, 64) == hs
    i, d = C.c_int32(), C.c_double()
    L.orc_xorshift(C.c_int64(seed), C.byref(i), C.byref(d))
    assert (i.value, d.value) == (first_int, first_double)


@pytest.mark.parametrize("n,want", [
    (9, [0, 5, 1, 6, 2, 7, 3, 8, 4]), (7, [0, 5, 1, 6, 2, 3, 4]),
    (20, [0, 5, 10, 14, 1, 6, 9, 13, 2, 17, 12, 7, 3, 18, 16, 11, 8, 19, 4, 15])])
def test_scala_immutable_hashmap_order(oracle, n, want):
    k = np.arange(n, dtype=np.int32)
    o = np.zeros(n, np.int32)
    oracle.lib().orc_scala_immutable_order(_i32p(k), n, _i32p(o))
    assert list(o) == want


def test_scala_small_maps_keep_mutable_order(oracle):
    # <= 4 trees: Map1..Map4 keep the mutable.HashMap iteration order (bucket high→low)
    k = np.array([3, 2, 1, 0], np.int32)
    o = np.zeros(4, np.int32)
    m = np.zeros(4, np.int32)
    oracle.lib().orc_group_tree_order(_i32p(k), 4, _i32p(o))
    oracle.lib().orc_scala_mutable_order(_i32p(k), 4, _i32p(m))
    assert list(o) == list(m) and sorted(o) == [0, 1, 2, 3]


@pytest.mark.parametrize("samples,num_splits,want", [
    ([0, 1, 0, 0, 1, 0, 1, 1], 2, [0.5]),
    ([0, 0, 1, 1, 2, 2, 3, 3], 2, [0.5, 2.5]),
    ([1, 1] + [2] * 22 + [3], 4, [1.5, 2.5]),
    ([2] * 19 + [3, 4, 5], 2, [2.5, 3.5]),
    ([0, 1] + [2] * 17, 1, [1.5])])
@pytest.mark.parametrize("policy", [0, 1])
def test_find_splits_for_continuous_feature_upstream_vectors(oracle, samples, num_splits, want, policy):
    """Spark ml/tree/impl/RandomForestSuite vectors; numExamples == samples.length (fraction 1)."""
    s = np.array(samples, np.float64)
    n_expected = len(s)
    if policy == 0:
        s = s[s != 0.0]  # 2.4 filters zeros and re-adds them as (numSamples - partNumSamples)
    out = np.zeros(64)
    n = oracle.lib().orc_find_splits_for_continuous_feature(
        s.ctypes.data_as(C.POINTER(C.c_double)), len(s), num_splits, C.c_int64(n_expected), policy,
        out.ctypes.data_as(C.POINTER(C.c_double)))
    assert list(out[:n]) == want


@pytest.mark.parametrize("seed", [123, 5354, 230, 349867, 23987])
def test_poisson_bagging_statistics(oracle, seed):
    """BaggedPointSuite: mean ~ rate, stddev ~ sqrt(rate) within 1 % (Well19937c has no known KAT)."""
    p = oracle.make_params(num_trees=100, num_classes=2, seed=seed)
    w, mw = oracle.bag_weights(p, 10000, oracle.partition_offsets(10000, 2))
    assert abs(w.mean() - 1.0) < 0.01 and abs(w.std() - 1.0) < 0.01 and mw < 20


def test_well19937c_two_implementations_agree(oracle):
    """Cross-check the oracle's Well19937c against an independent pure-Python transcription."""
    def well(seed, n):
        M = 0xFFFFFFFF
        v = [0] * 624
        v[0], v[1] = (seed >> 32) & M, seed & M
        for i in range(2, 624):
            l = v[i - 2] - (1 << 32) if v[i - 2] & 0x80000000 else v[i - 2]
            v[i] = (1812433253 * (l ^ (l >> 30)) + i) & M
        idx, out = 0, []
        for _ in range(n):
            i1, i2 = (idx + 623) % 624, (idx + 622) % 624
            v0, m1, m2, m3 = v[idx], v[(idx + 70) % 624], v[(idx + 179) % 624], v[(idx + 449) % 624]
            z0 = (0x80000000 & v[i1]) ^ (0x7FFFFFFF & v[i2])
            z1 = (v0 ^ (v0 << 25) ^ m1 ^ (m1 >> 27)) & M
            z2 = ((m2 >> 9) ^ m3 ^ (m3 >> 1)) & M
            z3 = z1 ^ z2
            z4 = (z0 ^ z1 ^ (z1 << 9) ^ z2 ^ (z2 << 21) ^ z3 ^ (z3 >> 21)) & M
            v[idx], v[i1] = z3, z4
            v[i2] &= 0x80000000
            idx = i1
            z4 ^= (z4 << 7) & 0xE46E1700
            z4 ^= (z4 << 15) & 0x9B868000
            out.append(z4 & M)
        return out
    got = np.zeros(2000, np.uint32)
    oracle.lib().orc_well19937c_bits32(C.c_int64(207336482), 2000, got.ctypes.data_as(C.POINTER(C.c_uint32)))
    assert list(got) == well(207336482, 2000)


def test_default_seed_is_class_name_hash():
    def java_hash(s):
        h = 0
        for ch in s:
            h = (31 * h + ord(ch)) & 0xFFFFFFFF
        return h - (1 << 32) if h & 0x80000000 else h
    assert java_hash("org.apache.spark.ml.classification.RandomForestClassifier") == 207336481


def test_cfg1_forest_matches_committed_golden(oracle):
    """BASELINE.json configs[0] (the reference's CPU-runnable case): 20 trees, depth 6, 1 500 draws.
    The golden file was produced by tests/golden/make_golden.py from this same oracle — it pins the
    oracle against accidental change, not against Spark."""
    from euromillioner_b200 import synth
    with open(os.path.join(GOLDEN, "cfg1_forest.json")) as f:
        g = json.load(f)
    X, y, Cn = synth.draw_history()
    p = oracle.make_params(num_trees=20, max_depth=6, max_bins=32, num_classes=Cn)
    fo = oracle.fit(p, X, y, oracle.partition_offsets(len(X), 8))
    e = fo.export()
    assert list(e["tree_off"]) == g["tree_off"]
    assert list(fo.group_sizes()) == g["group_sizes"]
    assert list(e["feature"][:200]) == g["feature_head"]
    assert [float(x) for x in e["threshold"][:200]] == g["threshold_head"]
    assert float(e["gain"].sum()) == g["gain_sum"]
    raw, prob, pred = fo.predict(X[:64])
    assert [int(v) for v in pred] == g["pred_head"]


def test_oracle_learns_like_an_independent_random_forest(oracle):
    """Secondary sanity oracle (not a parity claim): on a learnable problem the restated
    Spark forest must reach about the accuracy scikit-learn's RandomForestClassifier reaches."""
    sklearn = pytest.importorskip("sklearn.ensemble")
    rng = np.random.default_rng(11)
    X = rng.integers(1, 51, size=(6000, 20)).astype(np.float64)
    y = ((X[:, 0] > 25).astype(int) * 2 + (X[:, 3] + X[:, 7] > 50).astype(int)).astype(np.float64)
    flip = rng.random(len(y)) < 0.05
    y[flip] = rng.integers(0, 4, size=flip.sum())
    Xtr, ytr, Xte, yte = X[:4000], y[:4000], X[4000:], y[4000:]
    p = oracle.make_params(num_trees=30, max_depth=8, num_classes=4, seed=7)
    fo = oracle.fit(p, Xtr, ytr, oracle.partition_offsets(len(Xtr), 4))
    acc = float((fo.predict(Xte)[2] == yte).mean())
    sk = sklearn.RandomForestClassifier(n_estimators=30, max_depth=8, max_features="sqrt", random_state=7).fit(Xtr, ytr)
    acc_sk = float((sk.predict(Xte) == yte).mean())
    assert acc > 0.85 and abs(acc - acc_sk) < 0.05, (acc, acc_sk)


def test_debug_string_matches_committed_spark_format_text(oracle):
    """toDebugString in Spark 2.4's exact text: the product's renderer (classifier.debug_string) applied
    to the oracle's export of BASELINE configs[0] must reproduce tests/golden/cfg1_debug_string_3trees.txt, which
    make_golden.py rendered with its own recursive printer — header with (uid=...), `If (feature f <= t)` /
    `Else (feature f > t)` / `Predict: p`, one more space per level starting at four."""
    from euromillioner_b200 import synth
    from euromillioner_b200.classifier import debug_string
    X, y, Cn = synth.draw_history()
    p = oracle.make_params(num_trees=20, max_depth=6, max_bins=32, num_classes=Cn)
    e = oracle.fit(p, X, y, oracle.partition_offsets(len(X), 8)).export()
    with open(os.path.join(GOLDEN, "cfg1_debug_string_3trees.txt")) as f:
        want = f.read()
    assert debug_string(e, 3, "rfc_golden") == want


def test_java_double_to_string_known_answers():
    """java.lang.Double.toString known answers (JLS / JDK javadoc examples and boundary cases of the two notations)."""
    from euromillioner_b200.classifier import java_double_to_string as j
    for v, want in [(24.5, "24.5"), (12.0, "12.0"), (1e-5, "1.0E-5"), (1e7, "1.0E7"), (9999999.0, "9999999.0"),
                    (0.001, "0.001"), (0.00099, "9.9E-4"), (123456789.125, "1.23456789125E8"), (-3.25, "-3.25"),
                    (100.0, "100.0"), (1.5e-7, "1.5E-7"), (0.0, "0.0"), (-0.0, "-0.0"), (1e22, "1.0E22"), (0.1, "0.1"),
                    (12345678.9, "1.23456789E7"), (float("nan"), "NaN"), (float("inf"), "Infinity"),
                    (float("-inf"), "-Infinity"), (1.7976931348623157e308, "1.7976931348623157E308"), (0.012, "0.012")]:
        assert j(v) == want, (v, j(v), want)
