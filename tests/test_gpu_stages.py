"""GPU parity, stage by stage, through the C ABI.
Integer/byte/index work must be bit-exact against the oracle; gains bit-exact for gini."""
import numpy as np
import pytest

from euromillioner_b200 import stages, synth
from euromillioner_b200._lib import EmrfParams

pytestmark = pytest.mark.gpu


def _params(oracle, **kw):
    return oracle.make_params(cls=EmrfParams, **kw), oracle.make_params(**kw)


def _data(kind, n, F, seed=42):
    if kind == "flat":
        return synth.flat(n, F, 50, seed)
    rng = np.random.default_rng(seed)
    if kind == "real":      # continuous values: every sample distinct → stride walk over 10k distincts
        X = rng.normal(size=(n, F))
    elif kind == "sparse":  # mostly zeros: exercises the 2.4 zero filter / phantom zero count
        X = rng.integers(0, 4, size=(n, F)).astype(np.float64) * (rng.random((n, F)) < 0.3)
        X[:, 0] = 0.0       # all-zero feature: key absent → 0 splits
        X[:, 1] = 7.0       # constant feature
    elif kind == "ties":    # few distinct values, heavy ties, negative and -0.0
        X = rng.integers(-3, 4, size=(n, F)).astype(np.float64)
        X[X == 0] = rng.choice([0.0, -0.0], size=(X == 0).sum())
    y = rng.integers(0, 5, size=n).astype(np.float64)
    return X, y


@pytest.mark.parametrize("kind,n,F,parts,max_bins,policy", [
    ("flat", 1490, 70, 8, 32, 0), ("flat", 50000, 64, 64, 32, 0), ("flat", 50000, 64, 64, 32, 1),
    ("real", 30000, 16, 7, 32, 0), ("real", 5000, 9, 3, 16, 0), ("sparse", 20000, 12, 5, 32, 0),
    ("sparse", 20000, 12, 5, 32, 1), ("ties", 8000, 10, 4, 32, 0), ("ties", 8000, 10, 4, 4, 1),
    ("real", 40, 5, 2, 64, 0), ("real", 70000, 6, 9, 200, 0)])
def test_find_splits_matches_oracle(ctx, oracle, kind, n, F, parts, max_bins, policy):
    X, _ = _data(kind, n, F)
    pe, po = _params(oracle, num_trees=10, max_bins=max_bins, num_classes=5, find_splits_policy=policy)
    off = oracle.partition_offsets(n, parts)
    thr, ns = stages.find_splits(pe, X, off, ctx)
    thr_o, ns_o = oracle.find_splits(po, X, off)
    assert np.array_equal(ns, ns_o)
    for f in range(F):
        assert np.array_equal(thr[f, :ns[f]], thr_o[f, :ns_o[f]]), f


@pytest.mark.parametrize("kind,n,F", [("flat", 1490, 70), ("flat", 100003, 64), ("real", 20001, 33), ("ties", 9999, 10),
                                      ("sparse", 5000, 130)])
def test_binning_matches_oracle(ctx, oracle, kind, n, F):
    X, _ = _data(kind, n, F)
    if kind == "real":
        X[::97, 3] = np.nan   # NaN sorts above every threshold (Arrays.binarySearch)
    pe, po = _params(oracle, num_trees=10, max_bins=32, num_classes=5)
    thr, ns = oracle.find_splits(po, np.nan_to_num(X), oracle.partition_offsets(n, 4))
    if kind == "real":
        X[5, 0] = thr[0, 3]   # exact hit on a threshold goes to that threshold's bin
    cm, rm = stages.bin_features(X, thr, ns, ctx)
    want = oracle.bin_features(X, thr, ns)
    assert np.array_equal(rm, want)
    assert np.array_equal(cm, want.T)


@pytest.mark.parametrize("n,parts,T,rate,seed", [(1490, 8, 20, 1.0, 207336481), (20000, 64, 30, 1.0, 1),
                                                 (5000, 3, 7, 0.6, 99), (3000, 5, 1, 1.0, 5), (3000, 5, 1, 0.5, 5),
                                                 (100, 100, 4, 1.0, 3)])
def test_bag_weights_match_oracle(ctx, oracle, n, parts, T, rate, seed):
    pe, po = _params(oracle, num_trees=T, num_classes=2, subsampling_rate=rate, seed=seed)
    off = oracle.partition_offsets(n, parts)
    w, mw = stages.bag_weights(pe, n, off, ctx)
    w_o, mw_o = oracle.bag_weights(po, n, off)
    assert np.array_equal(w, w_o)
    assert mw == mw_o


def _random_nodes(rng, n_rows, T_nodes, sizes):
    rows, seg = [], [0]
    for s in sizes:
        r = np.sort(rng.choice(n_rows, size=s, replace=False))
        rows.append(r)
        seg.append(seg[-1] + s)
    return np.concatenate(rows), np.array(seg, np.int64)


@pytest.mark.parametrize("F,f_sub,C,max_bins,sizes", [
    (64, 8, 50, 32, [30000, 5, 1200, 0, 9000, 64]),      # multi-tile, tiny, empty and mid nodes
    (70, 9, 12, 32, [1490, 700, 1]),
    (20, 20, 3, 16, [20000, 333]),                       # "all" features
    (40, 40, 50, 32, [25000, 100]),                      # needs feature chunking (40*32*50*4 B > smem)
    (16, 4, 200, 8, [5000])])
def test_histogram_matches_numpy(ctx, oracle, F, f_sub, C, max_bins, sizes):
    rng = np.random.default_rng(F * 1000 + C)
    n = 40000
    nb = rng.integers(1, max_bins + 1, size=F).astype(np.int32)
    nb[0] = max_bins
    bins = (rng.integers(0, 1 << 30, size=(n, F)) % nb[None, :]).astype(np.uint8)
    labels = rng.integers(0, C, size=n)
    rows, seg = _random_nodes(rng, n, len(sizes), sizes)
    w = rng.integers(1, 9, size=len(rows))
    ent = stages.pack_entries(rows, w, labels[rows])
    subsets = np.stack([rng.permutation(F)[:f_sub] for _ in sizes]).astype(np.int32)
    got = stages.histogram(bins, ent, seg, subsets, nb, C, ctx)
    want = np.zeros_like(got)
    for g in range(len(sizes)):
        r, ww = rows[seg[g]:seg[g + 1]], w[seg[g]:seg[g + 1]]
        for k in range(f_sub):
            np.add.at(want[g, k], (bins[r, subsets[g, k]], labels[r]), ww.astype(np.uint32))
    assert np.array_equal(got, want)


@pytest.mark.parametrize("impurity", [0, 1])
@pytest.mark.parametrize("F,f_sub,C,max_bins,min_inst,min_gain", [
    (64, 8, 50, 32, 1, 0.0), (70, 9, 12, 32, 1, 0.0), (30, 30, 50, 32, 1, 0.0), (16, 5, 3, 8, 40, 0.0),
    (16, 5, 3, 8, 1, 0.05)])
def test_best_splits_match_oracle(ctx, oracle, impurity, F, f_sub, C, max_bins, min_inst, min_gain):
    rng = np.random.default_rng(F + 7 * C + impurity)
    n_nodes = 24
    nb = rng.integers(2, max_bins + 1, size=F).astype(np.int32)
    nb[0], nb[1] = max_bins, 1     # a feature with no splits
    B = int(nb.max())
    subsets = np.stack([rng.permutation(F)[:f_sub] for _ in range(n_nodes)]).astype(np.int32)
    subsets[3, 0] = 1              # the no-split feature leads a subset
    hist = np.zeros((n_nodes, f_sub, B, C), np.uint32)
    for g in range(n_nodes):
        m = int(rng.choice([3, 40, 1500, 60000]))
        lab = rng.integers(0, C, size=m) if g % 5 else np.full(m, 2)      # pure nodes: every gain is 0
        wt = rng.integers(1, 6, size=m)
        for k in range(f_sub):
            # ties on purpose: low-cardinality bins so several splits share the same gain
            b = rng.integers(0, max(1, nb[subsets[g, k]] // (1 + g % 3)), size=m)
            np.add.at(hist[g, k], (b, lab), wt.astype(np.uint32))
    got = stages.best_splits(hist, subsets, nb, impurity, min_inst, min_gain, ctx)
    for g in range(n_nodes):
        flat = np.concatenate([hist[g, k, :nb[subsets[g, k]], :].ravel() for k in range(f_sub)]).astype(np.float64)
        w = oracle.best_split(flat, subsets[g], nb, C, impurity, min_inst, min_gain)
        assert got["k"][g] == w["k"], g
        assert got["valid"][g] == int(w["valid"]), g
        if w["k"] >= 0:
            assert got["split"][g] == w["split"], g
            if impurity == 0:
                assert got["gain"][g] == w["gain"], g            # bit-exact (no FMA, same op order)
            else:
                assert got["gain"][g] == pytest.approx(w["gain"], rel=1e-6, abs=1e-12), g
            assert np.array_equal(got["left"][g].astype(np.float64), w["left"])
            assert np.array_equal((got["total"][g] - got["left"][g]).astype(np.float64), w["right"])


@pytest.mark.parametrize("sizes", [[30000, 5, 1200, 0, 9000, 2048, 2049], [1], [700000]])
def test_update_partition_is_stable_and_exact(ctx, sizes):
    rng = np.random.default_rng(len(sizes))
    n, F = 800000 if sizes[0] > 100000 else 40000, 6
    bins_cm = rng.integers(0, 32, size=(F, n)).astype(np.uint8)
    rows, seg = _random_nodes(rng, n, len(sizes), sizes)
    ent = stages.pack_entries(rows, rng.integers(1, 9, size=len(rows)), rng.integers(0, 50, size=len(rows)))
    sf = rng.integers(0, F, size=len(sizes)).astype(np.int32)
    sb = rng.integers(0, 31, size=len(sizes)).astype(np.int32)
    out, n_left = stages.update_partition(bins_cm, ent, seg, sf, sb, ctx)
    for g in range(len(sizes)):
        e = ent[seg[g]:seg[g + 1]]
        left = bins_cm[sf[g], (e & np.uint64(0xFFFFFFFF)).astype(np.int64)] <= sb[g]
        assert n_left[g] == left.sum()
        assert np.array_equal(out[seg[g]:seg[g + 1]], np.concatenate([e[left], e[~left]]))


@pytest.mark.parametrize("ball,lag", [(0, 10), (6, 10), (3, 1), (2, 25)])
def test_lag_feature_builder_matches_host_construction(ctx, ball, lag):
    """the draw-history feature builder (reference analogue Main.java:86-102)."""
    d = synth.draws(400)
    X, y = stages.build_lag_features(d, lag, ball, ctx)
    Xw, yw, _ = synth.draw_history(400, lag, ball)
    assert np.array_equal(X, Xw) and np.array_equal(y, yw)


@pytest.mark.gpu
@pytest.mark.parametrize("n,F,C_,start", [(3000, 10, 50, 0), (1000, 64, 50, 123456), (257, 3, 12, 7), (2048, 128, 50, 9_999_000)])
def test_device_generator_matches_numpy_generator(ctx, n, F, C_, start):
    """emrf_synth_flat (App. C flat mode, generated in HBM for the configs[2] bench shape) == synth.flat, bit for bit."""
    from euromillioner_b200 import synth
    d = synth.DeviceFlat(ctx, n, F, C_, seed=42, row_start=start)
    try:
        X, y = d.to_host()
    finally:
        d.free()
    Xn, yn = synth.flat(n, F, C_, seed=42, row_start=start)
    assert np.array_equal(X, Xn) and np.array_equal(y, yn)
