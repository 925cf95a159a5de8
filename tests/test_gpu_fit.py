"""End-to-end GPU parity through the reference-facing API (RandomForestClassifier.fit / Model.transform)
against the CPU oracle: scheduler trace, tree structure, thresholds, class counts and predicted labels
bit-exact; gains/impurities/probabilities bit-exact for gini (tolerance 1e-6 relative for entropy)."""
import numpy as np
import pytest

from euromillioner_b200 import synth
from euromillioner_b200.classifier import Dataset, RandomForestClassificationModel, RandomForestClassifier

pytestmark = pytest.mark.gpu


def _fit_both(ctx, oracle, X, y, parts, n_classes, **kw):
    impurity = kw.pop("impurity", "gini")
    subset = kw.pop("feature_subset", "auto")
    rf = RandomForestClassifier(ctx).setNumClasses(n_classes).setImpurity(impurity).setFeatureSubsetStrategy(subset)
    for k, v in kw.items():
        getattr(rf, "set" + k[0].upper() + k[1:])(v)
    model = rf.fit(Dataset(X, y, num_partitions=parts))
    po = oracle.make_params(num_trees=rf.numTrees, max_depth=rf.maxDepth, max_bins=rf.maxBins, num_classes=n_classes,
                            seed=rf.seed, impurity=impurity, feature_subset=subset, subsampling_rate=rf.subsamplingRate,
                            min_instances_per_node=rf.minInstancesPerNode, min_info_gain=rf.minInfoGain,
                            max_memory_mb=rf.maxMemoryInMB)
    ofo = oracle.fit(po, X, y, oracle.partition_offsets(len(X), parts))
    return model, ofo


def _assert_same_forest(model, ofo, exact=True):
    gs, tr = model.trace()
    assert np.array_equal(gs, ofo.group_sizes())
    assert np.array_equal(tr, ofo.trace())
    e, o = model.export(), ofo.export()
    for k in ("tree_off", "node_id", "left", "right", "feature", "prediction", "stats", "threshold"):
        assert np.array_equal(e[k], o[k]), k
    for k in ("gain", "impurity"):
        if exact:
            assert np.array_equal(e[k], o[k]), k
        else:
            assert np.allclose(e[k], o[k], rtol=1e-6, atol=1e-12), k


def test_cfg1_draw_history_ball1(ctx, oracle):
    """BASELINE.json configs[0]: 20 trees, depth 6, maxBins 32, 1 500 draws, 10-draw lags, ball 1."""
    X, y, C = synth.draw_history()
    model, ofo = _fit_both(ctx, oracle, X, y, 8, C, numTrees=20, maxDepth=6)
    _assert_same_forest(model, ofo)
    ds = model.transform(Dataset(X))
    raw, prob, pred = ofo.predict(X)
    assert np.array_equal(ds.columns["prediction"], pred)
    assert np.array_equal(ds.columns["rawPrediction"], raw)
    assert np.array_equal(ds.columns["probability"], prob)


@pytest.mark.parametrize("ball", [4, 6])
def test_cfg4_other_balls(ctx, oracle, ball):
    X, y, C = synth.draw_history(ball=ball)
    model, ofo = _fit_both(ctx, oracle, X, y, 8, C, numTrees=12, maxDepth=6, seed=ball)
    _assert_same_forest(model, ofo)


@pytest.mark.parametrize("n,F,T,D,parts,kw", [
    (60000, 64, 12, 8, 64, {}),                                   # cfg2-shaped, scaled down
    (30000, 32, 6, 10, 16, {"feature_subset": "all"}),            # forest independent of group order
    (30000, 32, 1, 6, 16, {}),                                    # single tree: no bagging, all features
    (20000, 24, 8, 7, 5, {"subsamplingRate": 0.7, "minInstancesPerNode": 25}),
    (20000, 24, 8, 7, 5, {"minInfoGain": 0.002, "feature_subset": "log2"}),
    (40000, 64, 30, 7, 32, {"maxMemoryInMB": 2}),                 # tiny groups: the memory check redraws the RNG
    (3000, 10, 40, 12, 3, {"feature_subset": "onethird"}),        # deep, small: pure children, many leaves
    (25000, 16, 5, 6, 4, {"maxBins": 100, "feature_subset": "0.5"}),
    (12000, 100, 4, 6, 6, {}),                                    # 100 features: rows padded to 128 bytes
    (9000, 150, 3, 5, 4, {}),                                     # rows wider than 128 B: generic byte-gather path
    (5000, 3, 4, 5, 1, {"maxBins": 2}),                           # one split per feature, one partition
    (4000, 1, 3, 4, 7, {}),                                       # a single feature
])
def test_flat_fits_match_oracle(ctx, oracle, n, F, T, D, parts, kw):
    X, y = synth.flat(n, F, 50)
    model, ofo = _fit_both(ctx, oracle, X, y, parts, 50, numTrees=T, maxDepth=D, **kw)
    _assert_same_forest(model, ofo)
    Xt, _ = synth.flat(5000, F, 50, seed=7)
    ds = model.transform(Dataset(Xt))
    raw, prob, pred = ofo.predict(Xt)
    assert np.array_equal(ds.columns["prediction"], pred)
    assert np.array_equal(ds.columns["rawPrediction"], raw)
    assert np.array_equal(ds.columns["probability"], prob)


def test_entropy_fit_within_tolerance(ctx, oracle):
    X, y = synth.flat(20000, 16, 10)
    y = y % 10
    model, ofo = _fit_both(ctx, oracle, X, y, 8, 10, numTrees=6, maxDepth=5, impurity="entropy")
    _assert_same_forest(model, ofo, exact=False)


def test_large_matrix_and_l2_blocked_tile_order_match_oracle(oracle):
    """A TreePoint matrix above 80 MB; with EMRF_L2_BLOCK=1 the histogram launch uses small tiles in row-range
    order.  Runs in a subprocess because the switch is read once per process."""
    import os, subprocess, sys
    code = ("import numpy as np, sys; sys.path.insert(0, %r)\n"
            "from euromillioner_b200 import synth\n"
            "from euromillioner_b200.classifier import Dataset, RandomForestClassifier\n"
            "from oracle import binding as ob\n"
            "X, y = synth.flat(700000, 128, 50)\n"
            "m = RandomForestClassifier().setNumTrees(3).setMaxDepth(5).setNumClasses(50).fit(Dataset(X, y, num_partitions=64))\n"
            "o = ob.fit(ob.make_params(num_trees=3, max_depth=5, num_classes=50), X, y, ob.partition_offsets(len(X), 64)).export()\n"
            "assert all(np.array_equal(v, o[k]) for k, v in m.export().items())\n"
            "print('BIG_OK')\n") % os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for flag in ("0", "1"):
        r = subprocess.run([sys.executable, "-c", code], env=dict(os.environ, EMRF_L2_BLOCK=flag), capture_output=True,
                           text=True, timeout=600)
        assert r.returncode == 0 and "BIG_OK" in r.stdout, r.stdout[-1500:] + r.stderr[-1500:]


def test_kernel_variants_match_oracle(oracle):
    """The histogram kernel's variants are chosen per launch / per node: rows staged by TMA gathers or by LDG/STS,
    16-bit counters when every node of a group is light and single-tile, column-major gathers for nodes holding
    >= 1/4 of the rows, tiled entry-list kernels.  A 200 000-row fit passes through all of them (u32 at the top, u16 below; dense roots, sparse deep
    nodes); the same fit with every variant switched off must give the same forest, and both must equal the
    oracle's.  Subprocesses: the switches are read once per process."""
    import os, subprocess, sys
    code = ("import numpy as np, sys; sys.path.insert(0, %r)\n"
            "from euromillioner_b200 import synth\n"
            "from euromillioner_b200.classifier import Dataset, RandomForestClassifier\n"
            "from oracle import binding as ob\n"
            "X, y = synth.flat(200000, 64, 50)\n"
            "m = RandomForestClassifier().setNumTrees(6).setMaxDepth(8).setNumClasses(50).fit(Dataset(X, y, num_partitions=96))\n"
            "o = ob.fit(ob.make_params(num_trees=6, max_depth=8, num_classes=50), X, y, ob.partition_offsets(len(X), 96)).export()\n"
            "assert all(np.array_equal(v, o[k]) for k, v in m.export().items())\n"
            "print('VARIANT_OK')\n") % os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    # {}: TMA-gathered row staging (the default); EMRF_NO_TMA=1: the LDG/STS staging it replaced, alone and with the
    # other variants off; EMRF_CM_DIV=0 with TMA: every node, dense roots included, goes through the gathers;
    # EMRF_NO_U16_MIXED=1: multi-tile nodes keep 32-bit launches (default: 16-bit tiles + piecewise scan of the sums);
    # EMRF_PART_3PASS=1: count / scan / scatter partition (default: one pass with look-back, right children reversed)
    for env in ({}, {"EMRF_NO_TMA": "1"}, {"EMRF_NO_TMA": "1", "EMRF_HIST_U32": "1", "EMRF_CM_DIV": "0", "EMRF_LIST_SIMPLE": "1"},
                {"EMRF_CM_DIV": "0"}, {"EMRF_CM_DIV": "64"}, {"EMRF_NO_U16_MIXED": "1"}, {"EMRF_NO_U16_MIXED": "1", "EMRF_CM_DIV": "0"},
                {"EMRF_PART_3PASS": "1"}):
        r = subprocess.run([sys.executable, "-c", code], env=dict(os.environ, **env), capture_output=True, text=True, timeout=600)
        assert r.returncode == 0 and "VARIANT_OK" in r.stdout, str(env) + r.stdout[-1500:] + r.stderr[-1500:]


def test_thresholds_scale_the_prediction_like_spark(ctx, oracle):
    """setThresholds: prediction = first maximum of probability / threshold (probability2prediction); a zero
    threshold wins wherever its class has non-zero probability; raw and probability columns do not change."""
    X, y = synth.flat(5000, 10, 4)
    rf = RandomForestClassifier(ctx).setNumTrees(7).setMaxDepth(5).setNumClasses(4)
    model = rf.fit(Dataset(X, y, num_partitions=4))
    ofo = oracle.fit(oracle.make_params(num_trees=7, max_depth=5, num_classes=4), X, y, oracle.partition_offsets(5000, 4))
    base = model.transform(Dataset(X)).columns
    for thr in ([0.1, 0.4, 0.3, 0.2], [1.0, 0.0, 1.0, 1.0], [0.25, 0.25, 0.25, 0.25]):
        got = model.setThresholds(thr).transform(Dataset(X)).columns
        _, oprob, opred = ofo.predict(X, thresholds=thr)
        assert np.array_equal(got["prediction"], opred)
        assert np.array_equal(got["probability"], oprob) and np.array_equal(got["rawPrediction"], base["rawPrediction"])
    assert not np.array_equal(model.setThresholds([0.1, 0.4, 0.3, 0.2]).transform(Dataset(X)).columns["prediction"], base["prediction"])
    assert np.array_equal(model.setThresholds(None).transform(Dataset(X)).columns["prediction"], base["prediction"])
    with pytest.raises(ValueError):
        model.setThresholds([0.5, 0.5]).transform(Dataset(X))            # length != numClasses
    with pytest.raises(ValueError):
        RandomForestClassifier(ctx).setThresholds([0.0, 0.0, 1.0, 1.0])  # more than one zero
    with pytest.raises(ValueError):
        RandomForestClassifier(ctx).setThresholds([-1.0, 1.0, 1.0, 1.0])
    with pytest.raises(ValueError, match="thresholds"):
        RandomForestClassifier(ctx).setNumTrees(2).setNumClasses(4).setThresholds([0.5, 0.5]).fit(Dataset(X, y))


def test_json_round_trip_is_exact(ctx):
    """toJson / fromJson: same flat arrays, same predictions, bit for bit."""
    X, y = synth.flat(3000, 9, 5)
    model = RandomForestClassifier(ctx).setNumTrees(5).setMaxDepth(6).setNumClasses(5).fit(Dataset(X, y, num_partitions=3))
    text = model.toJson()
    back = RandomForestClassificationModel.fromJson(text, ctx)
    a, b = model.export(), back.export()
    assert all(np.array_equal(a[k], b[k]) for k in a)
    pa, pb = model.transform(Dataset(X)).columns, back.transform(Dataset(X)).columns
    assert all(np.array_equal(pa[k], pb[k]) for k in ("rawPrediction", "probability", "prediction"))
    assert '"leftChild"' in text and '"impurityStats"' in text


def test_two_classes_and_empty_partitions(ctx, oracle):
    """Binary labels; a partition layout with empty partitions (their RNG streams are simply unused)."""
    X, y = synth.flat(6000, 12, 2)
    off = np.array([0, 0, 1500, 1500, 1500, 4000, 6000, 6000], np.int64)
    rf = RandomForestClassifier(ctx).setNumTrees(9).setMaxDepth(6).setNumClasses(2)
    model = rf.fit(Dataset(X, y, partition_offsets=off))
    po = oracle.make_params(num_trees=9, max_depth=6, num_classes=2)
    _assert_same_forest(model, oracle.fit(po, X, y, off))


def test_few_rows_and_inferred_classes(ctx, oracle):
    X, y = synth.flat(20, 5, 3)                  # rows < maxBins: numBins = min(maxBins, N)
    rf = RandomForestClassifier(ctx).setNumTrees(3).setMaxDepth(4)
    model = rf.fit(Dataset(X, y, num_partitions=2))
    assert model.numClasses == int(y.max()) + 1
    po = oracle.make_params(num_trees=3, max_depth=4, num_classes=model.numClasses)
    _assert_same_forest(model, oracle.fit(po, X, y, oracle.partition_offsets(20, 2)))


def test_invalid_inputs_raise_like_spark(ctx):
    X, y = synth.flat(100, 5, 3)
    with pytest.raises(ValueError, match="invalid label"):
        RandomForestClassifier(ctx).fit(Dataset(X, y + 0.5))
    with pytest.raises(ValueError, match="invalid label"):
        RandomForestClassifier(ctx).setNumClasses(2).fit(Dataset(X, y))
    y2 = y.copy()
    y2[0] = 150
    with pytest.raises(ValueError, match="100"):
        RandomForestClassifier(ctx).fit(Dataset(X, y2))


def test_device_resident_inputs_and_model_roundtrip(ctx, oracle):
    """fit() on features already in HBM gives the same forest; export → import → same predictions."""
    import ctypes as C
    from euromillioner_b200 import _lib
    L = _lib.lib()
    X, y = synth.flat(30000, 32, 50)
    dX, dy = C.c_void_p(), C.c_void_p()
    _lib.check(L.emrf_dev_alloc(ctx.h, X.nbytes, C.byref(dX)))
    _lib.check(L.emrf_dev_alloc(ctx.h, y.nbytes, C.byref(dy)))
    _lib.check(L.emrf_memcpy_h2d(ctx.h, dX, X.ctypes.data, X.nbytes))
    _lib.check(L.emrf_memcpy_h2d(ctx.h, dy, y.ctypes.data, y.nbytes))
    rf = RandomForestClassifier(ctx).setNumTrees(5).setMaxDepth(6).setNumClasses(50)
    m_dev = rf.fit(Dataset(dX.value, dy.value, num_partitions=8, n_rows=30000, n_features=32))
    m_host = rf.fit(Dataset(X, y, num_partitions=8))
    for k, v in m_host.export().items():
        assert np.array_equal(v, m_dev.export()[k]), k
    _lib.check(L.emrf_dev_free(ctx.h, dX))
    _lib.check(L.emrf_dev_free(ctx.h, dy))
    m2 = RandomForestClassificationModel.from_arrays(m_host.export(), 32, ctx)
    a = m_host.transform(Dataset(X[:2000])).columns
    b = m2.transform(Dataset(X[:2000])).columns
    for k in ("prediction", "rawPrediction", "probability"):
        assert np.array_equal(a[k], b[k])
    imp = m_host.featureImportances
    assert imp.shape == (32,) and abs(imp.sum() - 1.0) < 1e-12 and imp[0] > imp[10] and imp[3] > imp[10]
    assert m_host.toDebugString.startswith(f"RandomForestClassificationModel (uid={m_host.uid}) with 5 trees\n  Tree 0 (weight 1.0):\n    ")


def test_spark_ml_persistence_layout_round_trip(ctx, tmp_path):
    """save() writes Spark 2.4's model directory layout; load() rebuilds the same model."""
    import json
    import pyarrow.parquet as pq
    X, y = synth.flat(8000, 12, 5)
    y = y % 5
    m = RandomForestClassifier(ctx).setNumTrees(4).setMaxDepth(5).setNumClasses(5).fit(Dataset(X, y, num_partitions=4))
    path = str(tmp_path / "model")
    m.save(path, params={"numTrees": 4, "maxDepth": 5})
    meta = json.loads(open(path + "/metadata/part-00000").readline())
    assert meta["class"].endswith("RandomForestClassificationModel") and meta["numTrees"] == 4 and meta["numClasses"] == 5
    data = pq.read_table(path + "/data/part-00000.parquet")
    assert data.schema.names == ["treeID", "nodeData"] and data.num_rows == m.totalNumNodes
    assert [f.name for f in data.schema.field("nodeData").type] == ["id", "prediction", "impurity", "impurityStats", "gain",
                                                                     "leftChild", "rightChild", "split"]
    m2 = RandomForestClassificationModel.load(path, ctx)
    a, b = m.export(), m2.export()
    for k in ("tree_off", "left", "right", "feature", "threshold", "prediction", "impurity", "stats"):
        assert np.array_equal(a[k], b[k]), k
    assert np.array_equal(m.transform(Dataset(X[:500])).columns["probability"], m2.transform(Dataset(X[:500])).columns["probability"])


def test_fit_is_deterministic_and_counts_are_conserved(ctx):
    """Size-independent properties at a larger size than the oracle is run on."""
    X, y = synth.flat(300000, 64, 50)
    rf = RandomForestClassifier(ctx).setNumTrees(10).setMaxDepth(8).setNumClasses(50)
    m1 = rf.fit(Dataset(X, y, num_partitions=256))
    m2 = rf.fit(Dataset(X, y, num_partitions=256))
    e1, e2 = m1.export(), m2.export()
    for k in e1:
        assert np.array_equal(e1[k], e2[k]), k
    # children's class counts sum to the parent's, for every internal node of every tree
    for t in range(m1.numTrees):
        a, b = e1["tree_off"][t], e1["tree_off"][t + 1]
        L_, R_, st = e1["left"][a:b], e1["right"][a:b], e1["stats"][a:b]
        internal = np.where(L_ >= 0)[0]
        # pruned leaves keep the parent's counts, so conservation holds over the exported tree
        assert np.array_equal(st[internal], st[L_[internal]] + st[R_[internal]])
    # root count = sum of Poisson weights ~ N
    roots = e1["stats"][e1["tree_off"][:-1]].sum(axis=1)
    assert np.all(np.abs(roots - 300000) < 5000)
    # the context hands its device buffers back (a long-lived executor between jobs) and fits the same forest afterwards
    freed = ctx.release_buffers()
    assert freed > 50 * 1000 * 1000, freed       # TreePoint matrices, entry lists, weights of a 300k x 64 x 10-tree fit
    e3 = rf.fit(Dataset(X, y, num_partitions=256)).export()
    for k in e1:
        assert np.array_equal(e1[k], e3[k]), k


def test_gain_verifier_runs_and_finds_no_mismatch(oracle):
    """EMRF_VERIFY_GAINS=1 recomputes every split record on the host with plain IEEE arithmetic (the device uses a
    reciprocal + FMA division, see div_by): the counter must show that records were checked and that none differed.
    The switch is read once per process, hence the subprocess."""
    import os, subprocess, sys
    code = ("import numpy as np, sys; sys.path.insert(0, %r)\n"
            "from euromillioner_b200 import synth\n"
            "from euromillioner_b200.classifier import Context, Dataset, RandomForestClassifier\n"
            "ctx = Context(0)\n"
            "X, y = synth.flat(300000, 64, 50)\n"
            "m = RandomForestClassifier(ctx).setNumTrees(10).setMaxDepth(9).setNumClasses(50).fit(Dataset(X, y, num_partitions=128))\n"
            "t = ctx.timings()\n"
            "print('VERIFIED', t['gains_verified'], t['gain_mismatches'], t['total_nodes_split'])\n") % os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-c", code], env=dict(os.environ, EMRF_VERIFY_GAINS="1"), capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
    verified, mism, split = [int(v) for v in r.stdout.split("VERIFIED")[1].split()[:3]]
    assert verified == split and verified > 1000 and mism == 0


def test_predict_variants_chunks_and_threads(ctx, oracle):
    """emrf_predict: (a) the compact leaf records (integer counts, probabilities recomputed on the device) give the same
    raw votes / probabilities / labels as the oracle on more rows than one staging chunk holds (the chunks alternate
    between two streams); (b) two threads calling transform on one model and one context get those same answers
    (the header declares emrf_predict safe for concurrent callers); (c) the dense f64 records (EMRF_PREDICT_DENSE_LEAVES=1)
    agree, in a subprocess."""
    import os, subprocess, sys, threading
    X, y = synth.flat(60000, 64, 50)
    model, ofo = _fit_both(ctx, oracle, X, y, 16, 50, numTrees=12, maxDepth=7)
    Xb, _ = synth.flat(700000, 64, 50, seed=7)           # 358 MB of features: two staging chunks
    raw, prob, pred = ofo.predict(Xb)
    got = model.transform(Dataset(Xb)).columns
    assert np.array_equal(got["prediction"], pred) and np.array_equal(got["rawPrediction"], raw) and np.array_equal(got["probability"], prob)
    out = [None, None]

    def work(i):
        out[i] = model.transform(Dataset(Xb[i * 100000:(i + 3) * 100000])).columns
    th = [threading.Thread(target=work, args=(i,)) for i in range(2)]
    [t.start() for t in th]
    [t.join() for t in th]
    for i in range(2):
        sl = slice(i * 100000, (i + 3) * 100000)
        assert np.array_equal(out[i]["prediction"], pred[sl]) and np.array_equal(out[i]["rawPrediction"], raw[sl])
    code = ("import numpy as np, sys; sys.path.insert(0, %r)\n"
            "from euromillioner_b200 import synth\n"
            "from euromillioner_b200.classifier import Dataset, RandomForestClassifier\n"
            "from oracle import binding as ob\n"
            "X, y = synth.flat(60000, 64, 50)\n"
            "m = RandomForestClassifier().setNumTrees(12).setMaxDepth(7).setNumClasses(50).fit(Dataset(X, y, num_partitions=16))\n"
            "o = ob.fit(ob.make_params(num_trees=12, max_depth=7, num_classes=50), X, y, ob.partition_offsets(len(X), 16))\n"
            "raw, prob, pred = o.predict(X)\n"
            "g = m.transform(Dataset(X)).columns\n"
            "assert np.array_equal(g['prediction'], pred) and np.array_equal(g['rawPrediction'], raw) and np.array_equal(g['probability'], prob)\n"
            "print('DENSE_OK')\n") % os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-c", code], env=dict(os.environ, EMRF_PREDICT_DENSE_LEAVES="1"), capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "DENSE_OK" in r.stdout, r.stdout[-1500:] + r.stderr[-1500:]


def test_select_nodes_loop_policy_matches_oracle_when_a_group_fills_memory_exactly(ctx, oracle):
    """SURVEY.md A.9 policy: two classes x 4 features x 32 bins = 2048 B per node and maxMemoryInMB = 1 make every full
    group land EXACTLY on the limit (512 nodes), the one case where the `groupDone` loop (default) and the older
    `while (memUsage < maxMemoryUsage || memUsage == 0)` loop differ (the former spends one more draw of the driver
    RNG).  The engine must follow the oracle under either policy, and the two policies must give different forests."""
    rng = np.random.default_rng(3)
    X = rng.integers(1, 33, size=(40000, 16)).astype(np.float64)      # 32 distinct values: 32 bins per feature
    y = ((X[:, 0] + X[:, 5] + rng.integers(0, 3, size=len(X))) % 2).astype(np.float64)
    exports = []
    for pol in (0, 1):
        rf = (RandomForestClassifier(ctx).setNumTrees(40).setMaxDepth(9).setNumClasses(2).setMaxMemoryInMB(1))
        rf.selectNodesLoopPolicy = pol
        model = rf.fit(Dataset(X, y, num_partitions=8))
        po = oracle.make_params(num_trees=40, max_depth=9, num_classes=2, max_memory_mb=1, select_nodes_loop_policy=pol)
        ofo = oracle.fit(po, X, y, oracle.partition_offsets(len(X), 8))
        _assert_same_forest(model, ofo)
        assert 512 in list(ofo.group_sizes())
        exports.append(model.export())
    assert not np.array_equal(exports[0]["feature"], exports[1]["feature"])
