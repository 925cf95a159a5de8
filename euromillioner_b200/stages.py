"""Stage-level entry points of the C ABI (one kernel each), used by the parity tests.

Each function is a thin ctypes call into libemrf.so; arrays in, arrays out.  Names follow the
Spark routine the stage replaces.
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import check
from .classifier import _ptr, default_context


def pack_entries(rows, weights, labels):
    """entries word = row | weight << 32 | label << 40 (kernels.cuh pack_entry)."""
    return (np.asarray(rows, np.uint64) | (np.asarray(weights, np.uint64) << np.uint64(32))
            | (np.asarray(labels, np.uint64) << np.uint64(40)))


def find_splits(params, X, part_off, ctx=None):
    ctx = ctx or default_context()
    X = np.ascontiguousarray(X, np.float64)
    po = np.ascontiguousarray(part_off, np.int64)
    F = X.shape[1]
    thr = np.zeros((F, params.max_bins))
    ns = np.zeros(F, np.int32)
    check(_lib.lib().emrf_find_splits(ctx.h, C.byref(params), _ptr(X), X.shape[0], F,
                                      po.ctypes.data_as(C.POINTER(C.c_int64)), len(po) - 1, _ptr(thr), _ptr(ns)))
    return thr, ns


def find_splits_sample_rows(params, n_rows, part_off):
    po = np.ascontiguousarray(part_off, np.int64)
    rows = np.zeros(n_rows, np.int64)
    n = _lib.lib().emrf_find_splits_sample_rows(C.byref(params), n_rows, po.ctypes.data_as(C.POINTER(C.c_int64)),
                                                len(po) - 1, _ptr(rows), n_rows)
    if n < 0:
        raise ValueError(_lib.lib().emrf_last_error().decode())
    return rows[:n].copy()


def bin_features(X, thr, ns, ctx=None):
    ctx = ctx or default_context()
    X = np.ascontiguousarray(X, np.float64)
    thr = np.ascontiguousarray(thr, np.float64)
    ns = np.ascontiguousarray(ns, np.int32)
    N, F = X.shape
    cm = np.zeros((F, N), np.uint8)
    rm = np.zeros((N, F), np.uint8)
    check(_lib.lib().emrf_bin(ctx.h, _ptr(X), N, F, _ptr(thr), _ptr(ns), thr.shape[1], _ptr(cm), _ptr(rm)))
    return cm, rm


def bag_weights(params, n_rows, part_off, ctx=None):
    ctx = ctx or default_context()
    po = np.ascontiguousarray(part_off, np.int64)
    w = np.zeros((n_rows, params.num_trees), np.uint8)
    mw = C.c_int32()
    check(_lib.lib().emrf_bag_weights(ctx.h, C.byref(params), n_rows, po.ctypes.data_as(C.POINTER(C.c_int64)),
                                      len(po) - 1, _ptr(w), C.byref(mw)))
    return w, mw.value


def histogram(bins_rm, entries, seg_off, subsets, num_bins_all, n_classes, ctx=None):
    ctx = ctx or default_context()
    bins_rm = np.ascontiguousarray(bins_rm, np.uint8)
    entries = np.ascontiguousarray(entries, np.uint64)
    seg_off = np.ascontiguousarray(seg_off, np.int64)
    subsets = np.ascontiguousarray(subsets, np.int32)
    nb = np.ascontiguousarray(num_bins_all, np.int32)
    N, F = bins_rm.shape
    n_nodes, f_sub = subsets.shape
    B = int(nb.max())
    hist = np.zeros((n_nodes, f_sub, B, n_classes), np.uint32)
    check(_lib.lib().emrf_histogram(ctx.h, _ptr(bins_rm), N, F, _ptr(entries), _ptr(seg_off), n_nodes, _ptr(subsets),
                                    f_sub, _ptr(nb), n_classes, _ptr(hist)))
    return hist


def best_splits(hist, subsets, num_bins_all, impurity=0, min_instances=1, min_info_gain=0.0, ctx=None):
    ctx = ctx or default_context()
    hist = np.ascontiguousarray(hist, np.uint32)
    subsets = np.ascontiguousarray(subsets, np.int32)
    nb = np.ascontiguousarray(num_bins_all, np.int32)
    n_nodes, f_sub, B, Cn = hist.shape
    assert B == int(nb.max())
    k = np.zeros(n_nodes, np.int32)
    s = np.zeros(n_nodes, np.int32)
    gain = np.zeros(n_nodes)
    valid = np.zeros(n_nodes, np.int32)
    left = np.zeros((n_nodes, Cn), np.uint32)
    total = np.zeros((n_nodes, Cn), np.uint32)
    check(_lib.lib().emrf_best_splits(ctx.h, _ptr(hist), n_nodes, _ptr(subsets), f_sub, _ptr(nb), len(nb), Cn, impurity,
                                      min_instances, min_info_gain, _ptr(k), _ptr(s), _ptr(gain), _ptr(valid),
                                      _ptr(left), _ptr(total)))
    return {"k": k, "split": s, "gain": gain, "valid": valid, "left": left, "total": total}


def update_partition(bins_cm, entries, seg_off, split_feature, split_bin, ctx=None):
    ctx = ctx or default_context()
    bins_cm = np.ascontiguousarray(bins_cm, np.uint8)
    entries = np.ascontiguousarray(entries, np.uint64)
    seg_off = np.ascontiguousarray(seg_off, np.int64)
    sf = np.ascontiguousarray(split_feature, np.int32)
    sb = np.ascontiguousarray(split_bin, np.int32)
    F, N = bins_cm.shape
    out = np.zeros_like(entries)
    n_left = np.zeros(len(sf), np.int64)
    check(_lib.lib().emrf_update_partition(ctx.h, _ptr(bins_cm), N, F, _ptr(entries), _ptr(seg_off), len(sf), _ptr(sf),
                                           _ptr(sb), _ptr(out), _ptr(n_left)))
    return out, n_left


def build_lag_features(draws, lag=10, ball=0, ctx=None):
    """Draw table (n_draws x 7) -> (X: N x 7*lag float64, y: N float64) built on the device."""
    ctx = ctx or default_context()
    draws = np.ascontiguousarray(draws, np.int32)
    n = draws.shape[0] - lag
    X = np.zeros((n, 7 * lag))
    y = np.zeros(n)
    check(_lib.lib().emrf_build_lag_features(ctx.h, _ptr(draws), draws.shape[0], lag, ball, _ptr(X), _ptr(y)))
    return X, y
