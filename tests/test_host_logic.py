"""CPU-side checks of the product: libemrf.so loads, exports every symbol include/emrf.h declares,
its host-side scheduler primitives (java.util.Random, XORShiftRandom, reservoir sampling, Scala
map orders, findSplits row sampler) agree with the oracle, and it fails loudly without a GPU."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import HAS_GPU, ROOT


def _lib():
    from euromillioner_b200 import _lib
    return _lib.lib()


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "emrf.h")).read()
    names = sorted(set(re.findall(r"\b(emrf_[a-z0-9_]+)\s*\(", hdr)))
    assert len(names) >= 30
    L = _lib()
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing
    assert b"sm_100a" in L.emrf_version()


@pytest.mark.skipif(HAS_GPU, reason="checks the no-GPU failure mode")
def test_no_cpu_fallback_without_gpu():
    from euromillioner_b200._lib import EmrfError
    from euromillioner_b200.classifier import Context
    with pytest.raises(EmrfError, match="no CPU fallback"):
        Context(0)


def test_java_random_and_xorshift_match_oracle(oracle):
    L, O = _lib(), oracle.lib()
    for seed in (0, 1, 42, 207336481, -77, 2**40 + 3):
        want = np.zeros(5, np.int64)
        O.orc_java_random_longs(C.c_int64(seed), 5, want.ctypes.data_as(C.POINTER(C.c_int64)))
        assert [L.emrf_host_java_random_next_long(seed, i) for i in range(5)] == list(want)
        wd = np.zeros(5)
        O.orc_xorshift_doubles(C.c_int64(seed), 5, wd.ctypes.data_as(C.POINTER(C.c_double)))
        assert [L.emrf_host_xorshift_next_double(seed, i) for i in range(5)] == list(wd)
    assert L.emrf_host_java_random_next_long(42, 0) == -5025562857975149833
    assert L.emrf_host_xorshift_next_double(207336481, 0) == 0.26236745326918753


@pytest.mark.parametrize("n_items,k", [(70, 9), (64, 8), (128, 12), (5, 9), (9, 9), (1000, 32)])
def test_reservoir_sample_matches_oracle(oracle, n_items, k):
    L, O = _lib(), oracle.lib()
    for seed in (1, -5025562857975149833, 987654321987):
        a, b = np.zeros(k, np.int32), np.zeros(k, np.int32)
        na = L.emrf_host_reservoir_sample(n_items, k, seed, a.ctypes.data)
        nb = O.orc_reservoir(n_items, k, C.c_int64(seed), b.ctypes.data_as(C.POINTER(C.c_int32)))
        assert na == nb == min(n_items, k)
        assert list(a[:na]) == list(b[:nb])


def test_group_tree_order_matches_oracle_and_known_answers(oracle):
    L, O = _lib(), oracle.lib()
    rng = np.random.default_rng(7)
    cases = [np.arange(n, dtype=np.int32) for n in (1, 3, 4, 5, 9, 13, 14, 20, 100, 500)]
    cases += [np.arange(n - 1, -1, -1, dtype=np.int32) for n in (4, 20, 100, 500)]
    cases += [rng.permutation(500)[:n].astype(np.int32) for n in (2, 4, 7, 30, 200)]
    for k in cases:
        a, b = np.zeros(len(k), np.int32), np.zeros(len(k), np.int32)
        assert L.emrf_host_group_tree_order(k.ctypes.data, len(k), a.ctypes.data) == len(k)
        O.orc_group_tree_order(k.ctypes.data_as(C.POINTER(C.c_int32)), len(k), b.ctypes.data_as(C.POINTER(C.c_int32)))
        assert list(a) == list(b)
    k = np.arange(9, dtype=np.int32)
    a = np.zeros(9, np.int32)
    L.emrf_host_group_tree_order(k.ctypes.data, 9, a.ctypes.data)
    assert list(a) == [0, 5, 1, 6, 2, 7, 3, 8, 4]


@pytest.mark.parametrize("n_rows,n_parts,max_bins", [(1490, 8, 32), (25000, 16, 32), (200000, 64, 32), (30000, 7, 128)])
def test_find_splits_row_sampler_matches_oracle(oracle, n_rows, n_parts, max_bins):
    from euromillioner_b200 import stages
    from euromillioner_b200._lib import EmrfParams
    po = oracle.partition_offsets(n_rows, n_parts)
    for seed in (207336481, 5):
        pe = oracle.make_params(num_trees=10, max_bins=max_bins, num_classes=2, seed=seed, cls=EmrfParams)
        po_ = oracle.make_params(num_trees=10, max_bins=max_bins, num_classes=2, seed=seed)
        got = stages.find_splits_sample_rows(pe, n_rows, po)
        want = oracle.find_splits_sample_rows(po_, n_rows, po)
        assert np.array_equal(got, want)
        if n_rows <= 10000:
            assert len(got) == n_rows


@pytest.mark.parametrize("F,kind,want", [(70, "auto", 9), (64, "auto", 8), (128, "sqrt", 12), (128, "log2", 7),
                                         (128, "onethird", 43), (64, "all", 64), (64, "10", 10), (64, "0.5", 32)])
def test_num_features_per_node(oracle, F, kind, want):
    from euromillioner_b200._lib import EmrfParams
    pe = oracle.make_params(num_trees=5, num_classes=2, feature_subset=kind, cls=EmrfParams)
    po = oracle.make_params(num_trees=5, num_classes=2, feature_subset=kind)
    assert _lib().emrf_host_num_features_per_node(C.byref(pe), F) == want
    assert oracle.lib().orc_num_features_per_node(C.byref(po), F) == want


def test_estimator_setters_mirror_spark_requires():
    from euromillioner_b200.classifier import RandomForestClassifier
    rf = RandomForestClassifier()
    assert (rf.numTrees, rf.maxDepth, rf.maxBins, rf.seed, rf.featureSubsetStrategy) == (20, 5, 32, 207336481, "auto")
    assert rf.setNumTrees(100).setMaxDepth(10).setMaxBins(32).setSeed(1) is rf
    for bad in (lambda: rf.setNumTrees(0), lambda: rf.setMaxDepth(31), lambda: rf.setMaxBins(1),
                lambda: rf.setImpurity("variance"), lambda: rf.setSubsamplingRate(0.0),
                lambda: rf.setFeatureSubsetStrategy("half")):
        with pytest.raises(ValueError):
            bad()


def _import_forest(tree_off, left, right, feature, F=4, C_=2):
    """emrf_forest_import on a raw description (no GPU needed: the check is host code). Returns the status code."""
    L = _lib()
    n = len(left)
    a = [np.asarray(tree_off, np.int64), np.arange(n, dtype=np.int32), np.asarray(left, np.int32), np.asarray(right, np.int32),
         np.asarray(feature, np.int32), np.zeros(n), np.zeros(n), np.zeros(n), np.zeros(n), np.ones(n * C_)]
    h = C.c_void_p()
    rc = L.emrf_forest_import(None, len(tree_off) - 1, C_, F, *[x.ctypes.data for x in a], C.byref(h))
    if rc == 0:
        L.emrf_forest_free(h)
    return rc


def test_forest_import_rejects_corrupted_descriptions():
    """RandomForestClassificationModel.load / fromJson feed emrf_forest_import from files: a description that is not a
    set of proper pre-order trees (cycle, shared child, child outside its tree, bad feature) must be refused, not walked."""
    ok = dict(tree_off=[0, 3, 4], left=[1, -1, -1, -1], right=[2, -1, -1, -1], feature=[0, -1, -1, -1])
    assert _import_forest(**ok) == 0
    bad = [
        dict(ok, tree_off=[1, 3, 4]),                      # does not start at 0
        dict(ok, tree_off=[0, 3, 3]),                      # empty tree
        dict(ok, left=[0, -1, -1, -1]),                    # a node is its own child (cycle)
        dict(ok, left=[1, -1, -1, -1], right=[1, -1, -1, -1]),   # both children the same node
        dict(ok, right=[3, -1, -1, -1]),                   # child outside the tree
        dict(ok, feature=[4, -1, -1, -1]),                 # feature out of range
        dict(tree_off=[0, 5], left=[1, 3, 3, -1, -1], right=[2, 4, 4, -1, -1], feature=[0, 1, 1, -1, -1]),  # shared children
        dict(tree_off=[0, 4], left=[1, -1, -1, -1], right=[2, -1, -1, -1], feature=[0, -1, -1, -1]),        # unreachable node
        dict(tree_off=[0, 3], left=[1, -1, -1], right=[-1, -1, -1], feature=[0, -1, -1]),                   # one child only
    ]
    for d in bad:
        assert _import_forest(**d) == -1, d
        assert b"forest description" in _lib().emrf_last_error()


@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_packed_slot_layout_of_the_reduce_scattered_histograms(world):
    """Every rank's share is one contiguous run of half slots, heavy nodes take two and light nodes one, nothing overlaps,
    and a node is owned by the rank whose share holds it (what ncclReduceScatter hands that rank)."""
    L = _lib()
    rng = np.random.default_rng(world)
    for n in (1, 2, 7, 100, 2542):
        for p_light in (0.0, 0.5, 0.93, 1.0):
            w = np.where(rng.random(n) < p_light, rng.integers(1, 65536, n), rng.integers(65536, 2**32, n)).astype(np.uint32)
            slot, owner = np.zeros(n, np.int32), np.zeros(n, np.int32)
            share = L.emrf_host_packed_slot_layout(w.ctypes.data, n, world, slot.ctypes.data, owner.ctypes.data)
            size = np.where(w <= 65535, 1, 2)
            assert share >= 1 and np.all(slot >= 0) and np.all(slot + size <= share * world)
            assert np.array_equal(owner, slot // share) and np.array_equal(owner, (slot + size - 1) // share)
            used = np.zeros(share * world, np.int32)
            for s_, z in zip(slot, size):
                used[s_:s_ + z] += 1
            assert used.max() == 1                                    # no two nodes share a half slot
            assert np.all(slot[w > 65535] % 2 == (share % 2) * (owner[w > 65535] % 2))   # whole slots stay word-aligned
            # light nodes sit behind the heavy ones of the same rank
            for r in range(world):
                hv, lt = slot[(owner == r) & (w > 65535)], slot[(owner == r) & (w <= 65535)]
                if len(hv) and len(lt):
                    assert hv.max() + 2 <= lt.min()
            assert np.bincount(owner, minlength=world).max() <= -(-np.sum(w > 65535) // world) + -(-np.sum(w <= 65535) // world)


def test_group_tiles_bound_is_never_below_the_exact_tile_count():
    """Rows sharded: launches are provisioned from bounds and tiled on the device from exact lengths.  Random forests of
    disjoint nodes: exact lengths partition (part of) each tree's entries, bounds are anything >= the exact length."""
    L = _lib()
    rng = np.random.default_rng(11)
    for case in range(200):
        n_trees = int(rng.integers(1, 6))
        tree_total = rng.integers(1, 2_000_000, n_trees).astype(np.uint64)
        n = int(rng.integers(1, 400))
        tree = rng.integers(0, n_trees, n).astype(np.int32)
        exact = np.zeros(n, np.uint32)
        for t in range(n_trees):                      # split (part of) the tree's entries among its nodes; zeros allowed
            idx = np.where(tree == t)[0]
            if len(idx):
                cuts = np.sort(rng.integers(0, int(tree_total[t]) + 1, len(idx) + 1))
                exact[idx] = np.diff(cuts).astype(np.uint32)
        ub = (exact + rng.integers(0, 3_000_000, n) * rng.integers(0, 2, n)).astype(np.uint32)    # loose for about half
        for tile_len in (256, 2048, 8192, 65536):
            got = L.emrf_host_group_tiles_bound(ub.ctypes.data, tree.ctypes.data, n, tree_total.ctypes.data, n_trees, tile_len)
            need = int(np.sum(np.maximum(1, -(-exact.astype(np.int64) // tile_len))))
            assert got >= need, (case, tile_len, got, need)
            assert got <= int(tree_total.sum()) // tile_len + n           # and no looser than the trees allow
