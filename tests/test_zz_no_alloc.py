"""emrf_update_partition_noalloc — the partition update for hosts that cannot allocate on the GPU.

Who needs it: a process that gave all device memory to a pool allocator when it started (so a
cudaMalloc made by a library would fail), or a latency-bound loop that cannot take the device-wide
synchronisation cudaMalloc/cudaFree bring.  The entry point works only in buffers the caller owns
plus one fixed device array inside libemrf.so.  This test drives it through the C ABI with device
buffers the TEST allocates, and checks entries and left counts bit-for-bit against the regular
emrf_update_partition, on two calls in a row (the second smaller, so anything stale left in the
shared array by the first would show).  Collected last on purpose: a -x run has then already run
every other test."""
import ctypes as C

import numpy as np
import pytest

from euromillioner_b200 import _lib, stages

pytestmark = pytest.mark.gpu


def _dev(ctx, arr):
    L = _lib.lib()
    p = C.c_void_p()
    _lib.check(L.emrf_dev_alloc(ctx.h, max(arr.nbytes, 16), C.byref(p)))
    if arr.nbytes:
        _lib.check(L.emrf_memcpy_h2d(ctx.h, p, arr.ctypes.data, arr.nbytes))
    return p


def _run_noalloc(ctx, bins_cm, ent, seg, sf, sb):
    L = _lib.lib()
    bufs = []
    try:
        d_bins = _dev(ctx, bins_cm); bufs.append(d_bins)
        d_in = _dev(ctx, ent); bufs.append(d_in)
        d_out = _dev(ctx, np.zeros_like(ent)); bufs.append(d_out)
        d_nl = _dev(ctx, np.full(len(sf), 0xFFFFFFFFFFFFFFFF, np.uint64)); bufs.append(d_nl)
        seg = np.ascontiguousarray(seg, np.int64)
        sf, sb = np.ascontiguousarray(sf, np.int32), np.ascontiguousarray(sb, np.int32)
        _lib.check(L.emrf_update_partition_noalloc(ctx.h, d_bins, bins_cm.shape[1], bins_cm.shape[0], d_in, seg.ctypes.data,
                                                   len(sf), sf.ctypes.data, sb.ctypes.data, d_out, d_nl))
        out, nl = np.zeros_like(ent), np.zeros(len(sf), np.uint64)
        if ent.nbytes:
            _lib.check(L.emrf_memcpy_d2h(ctx.h, out.ctypes.data, d_out, ent.nbytes))
        _lib.check(L.emrf_memcpy_d2h(ctx.h, nl.ctypes.data, d_nl, nl.nbytes))
        return out, nl.astype(np.int64)
    finally:                       # also when the call raises (the oversized cases below)
        for p in bufs:
            L.emrf_dev_free(ctx.h, p)


def _numpy_partition(bins_cm, ent, seg, sf, sb):
    """The partition as plain numpy: inside every node's segment the entries whose row has bin(feature) <= split_bin come
    first, both sides in their original order (LearningNode.predictImpl on binned features)."""
    out, nl = ent.copy(), np.zeros(len(sf), np.int64)
    for i in range(len(sf)):
        a, b = int(seg[i]), int(seg[i + 1])
        rows = (ent[a:b] & np.uint64(0xFFFFFFFF)).astype(np.int64)
        left = bins_cm[sf[i], rows] <= sb[i]
        out[a:b] = np.concatenate([ent[a:b][left], ent[a:b][~left]])
        nl[i] = int(left.sum())
    return out, nl


def _case(rng, n_rows, sizes, F=7):
    bins_cm = rng.integers(0, 32, size=(F, n_rows)).astype(np.uint8)
    rows, seg = [], [0]
    for s in sizes:
        rows.append(np.sort(rng.choice(n_rows, size=s, replace=False)))
        seg.append(seg[-1] + s)
    rows = np.concatenate(rows) if rows else np.zeros(0, np.int64)
    ent = stages.pack_entries(rows, rng.integers(1, 9, size=len(rows)), rng.integers(0, 50, size=len(rows)))
    sf = rng.integers(0, F, size=len(sizes)).astype(np.int32)
    sb = rng.integers(0, 31, size=len(sizes)).astype(np.int32)
    return bins_cm, ent, np.array(seg, np.int64), sf, sb


def test_noalloc_partition_matches_regular_entry_point_twice(ctx):
    rng = np.random.default_rng(20200614)
    for sizes in ([150000, 3, 0, 2048, 2049, 40000], [5000, 17]):   # two calls in a row, second smaller
        bins_cm, ent, seg, sf, sb = _case(rng, 200000, sizes)
        want_out, want_nl = stages.update_partition(bins_cm, ent, seg, sf, sb, ctx)
        got_out, got_nl = _run_noalloc(ctx, bins_cm, ent, seg, sf, sb)
        assert np.array_equal(got_nl, want_nl)
        assert np.array_equal(got_out, want_out)
        np_out, np_nl = _numpy_partition(bins_cm, ent, seg, sf, sb)      # and against the definition itself
        assert np.array_equal(got_nl, np_nl)
        assert np.array_equal(got_out, np_out)


def test_noalloc_partition_rejects_oversized_requests(ctx):
    rng = np.random.default_rng(1)
    bins_cm, ent, seg, sf, sb = _case(rng, 100, [1] * 5000)          # more than 4096 nodes
    with pytest.raises(_lib.EmrfError, match="at most 4096 nodes"):
        _run_noalloc(ctx, bins_cm, ent, seg, sf, sb)
    L = _lib.lib()
    seg = np.array([0, 600_000_000], np.int64)                       # more than 262144 tiles
    one = np.zeros(1, np.int32)
    d = _dev(ctx, np.zeros(16, np.uint8))
    rc = L.emrf_update_partition_noalloc(ctx.h, d, 1, 1, d, seg.ctypes.data, 1, one.ctypes.data, one.ctypes.data, d, d)
    assert rc == _lib.EMRF_ERR_LIMIT and b"262144 tiles" in L.emrf_last_error()
    _lib.check(L.emrf_dev_free(ctx.h, d))


def test_noalloc_partition_rejects_aliased_and_host_buffers(ctx):
    """The partition is not done in place and works on device memory only: the same buffer for input and output, an
    output that overlaps the input, and a host pointer are refused (before anything is enqueued) instead of racing."""
    rng = np.random.default_rng(2)
    bins_cm, ent, seg, sf, sb = _case(rng, 5000, [3000, 1000])
    L = _lib.lib()
    d_bins, d_in, d_nl = _dev(ctx, bins_cm), _dev(ctx, np.concatenate([ent, ent])), _dev(ctx, np.zeros(2, np.uint64))
    try:
        args = lambda src, dst: (ctx.h, d_bins, bins_cm.shape[1], bins_cm.shape[0], src, seg.ctypes.data, 2, sf.ctypes.data,
                                 sb.ctypes.data, dst, d_nl)
        assert L.emrf_update_partition_noalloc(*args(d_in, d_in)) == _lib.EMRF_ERR_ARG
        assert b"must not alias" in L.emrf_last_error()
        overlapping = C.c_void_p(d_in.value + 8 * 100)
        assert L.emrf_update_partition_noalloc(*args(d_in, overlapping)) == _lib.EMRF_ERR_ARG
        host_out = np.zeros_like(ent)
        assert L.emrf_update_partition_noalloc(*args(d_in, C.c_void_p(host_out.ctypes.data))) == _lib.EMRF_ERR_ARG
        assert b"device pointers" in L.emrf_last_error()
        seg_bad = np.array([-1, 10, 20], np.int64)
        assert L.emrf_update_partition_noalloc(ctx.h, d_bins, bins_cm.shape[1], bins_cm.shape[0], d_in, seg_bad.ctypes.data, 2,
                                               sf.ctypes.data, sb.ctypes.data, C.c_void_p(d_in.value + 8 * 4000), d_nl) == _lib.EMRF_ERR_ARG
    finally:
        for p in (d_bins, d_in, d_nl):
            L.emrf_dev_free(ctx.h, p)
