"""ctypes loader for libemrf.so (the C ABI declared in include/emrf.h).

The product path has NO CPU fallback: if the CUDA library is missing or no B200 is visible,
loading / context creation raises.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("EMRF_LIB_PATH") or os.path.join(_HERE, "libemrf.so")  # override: A/B builds only

EMRF_OK, EMRF_ERR_ARG, EMRF_ERR_CUDA, EMRF_ERR_LIMIT, EMRF_ERR_INTERNAL = 0, -1, -2, -3, -4


class EmrfParams(C.Structure):
    _fields_ = [
        ("num_trees", C.c_int32), ("max_depth", C.c_int32), ("max_bins", C.c_int32),
        ("min_instances_per_node", C.c_int32), ("num_classes", C.c_int32),
        ("impurity", C.c_int32), ("feature_subset_kind", C.c_int32),
        ("feature_subset_value", C.c_double), ("min_info_gain", C.c_double),
        ("subsampling_rate", C.c_double),
        ("max_memory_mb", C.c_int32), ("find_splits_policy", C.c_int32),
        ("seed", C.c_int64),
        ("shard_mode", C.c_int32), ("policy_flags", C.c_int32),
    ]


class EmrfTimings(C.Structure):
    _fields_ = [(n, C.c_double) for n in (
        "total_ms", "h2d_ms", "find_splits_ms", "bin_ms", "bag_ms", "find_best_splits_ms",
        "hist_kernel_ms", "partition_ms", "host_scheduler_ms")] + [(n, C.c_int64) for n in (
            "hist_launches", "kernel_launches", "hist_alg_bytes", "hist_units", "num_groups",
            "total_nodes_split", "gain_mismatches", "h2d_bytes", "d2h_bytes")] + [
        (n, C.c_double) for n in ("host_select_ms", "host_update_ms", "host_export_ms", "bag_kernel_ms", "comm_ms")] + [
        (n, C.c_int64) for n in ("comm_bytes", "gains_verified")]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class EmrfError(RuntimeError):
    """Non-argument failure reported by the engine (CUDA error, unsupported shape …)."""


_lib = None


def lib():
    """Load libemrf.so; raises (never falls back) when it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise EmrfError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(make -C euromillioner_b200/csrc). There is no CPU fallback.")
    L = C.CDLL(LIB_PATH)
    vp, i32, i64, f64 = C.c_void_p, C.c_int32, C.c_int64, C.c_double
    P = C.POINTER
    L.emrf_last_error.restype = C.c_char_p
    L.emrf_version.restype = C.c_char_p
    L.emrf_ctx_create.argtypes = [i32, P(vp)]
    L.emrf_ctx_create_dist.argtypes = [i32, i32, i32, P(C.c_uint8), P(vp)]
    L.emrf_nccl_unique_id.argtypes = [P(C.c_uint8)]
    L.emrf_ctx_destroy.argtypes = [vp]
    L.emrf_ctx_destroy.restype = None
    L.emrf_get_timings.argtypes = [vp, P(EmrfTimings)]
    L.emrf_ctx_release_buffers.argtypes = [vp, P(i64)]
    L.emrf_dev_alloc.argtypes = [vp, i64, P(vp)]
    L.emrf_dev_free.argtypes = [vp, vp]
    L.emrf_host_alloc_pinned.argtypes = [i64, P(vp)]
    L.emrf_host_free_pinned.argtypes = [vp]
    L.emrf_memcpy_h2d.argtypes = [vp, vp, vp, i64]
    L.emrf_memcpy_d2h.argtypes = [vp, vp, vp, i64]
    L.emrf_fit.argtypes = [vp, P(EmrfParams), vp, vp, i64, i32, P(i64), i32, P(vp)]
    L.emrf_fit_sharded.argtypes = [vp, P(EmrfParams), vp, vp, i64, i32, P(i64), i32, i32, i32, P(vp)]
    L.emrf_forest_free.argtypes = [vp]
    L.emrf_forest_free.restype = None
    L.emrf_forest_info.argtypes = [vp, P(i32), P(i32), P(i32), P(i64)]
    L.emrf_forest_export.argtypes = [vp] + [vp] * 10
    L.emrf_forest_import.argtypes = [vp, i32, i32, i32] + [vp] * 10 + [P(vp)]
    L.emrf_forest_trace_info.argtypes = [vp, P(i32), P(i64), P(i32)]
    L.emrf_forest_trace.argtypes = [vp, vp, vp]
    L.emrf_forest_feature_importances.argtypes = [vp, vp]
    L.emrf_predict.argtypes = [vp, vp, vp, i64, vp, vp, vp]
    L.emrf_predict_thresholds.argtypes = [vp, vp, vp, i64, vp, vp, vp, vp]
    L.emrf_find_splits.argtypes = [vp, P(EmrfParams), vp, i64, i32, P(i64), i32, vp, vp]
    L.emrf_find_splits_sample_rows.argtypes = [P(EmrfParams), i64, P(i64), i32, vp, i64]
    L.emrf_find_splits_sample_rows.restype = i64
    L.emrf_bin.argtypes = [vp, vp, i64, i32, vp, vp, i32, vp, vp]
    L.emrf_bag_weights.argtypes = [vp, P(EmrfParams), i64, P(i64), i32, vp, P(i32)]
    L.emrf_histogram.argtypes = [vp, vp, i64, i32, vp, vp, i32, vp, i32, vp, i32, vp]
    L.emrf_best_splits.argtypes = [vp, vp, i32, vp, i32, vp, i32, i32, i32, i32, f64, vp, vp, vp, vp, vp, vp]
    L.emrf_update_partition.argtypes = [vp, vp, i64, i32, vp, vp, i32, vp, vp, vp, vp]
    L.emrf_build_lag_features.argtypes = [vp, vp, i32, i32, i32, vp, vp]
    L.emrf_synth_flat.argtypes = [vp, i64, i32, i32, i64, i64, vp, vp]
    L.emrf_update_partition_noalloc.argtypes = [vp, vp, i64, i32, vp, vp, i32, vp, vp, vp, vp]
    L.emrf_host_java_random_next_long.argtypes = [i64, i32]
    L.emrf_host_java_random_next_long.restype = i64
    L.emrf_host_xorshift_next_double.argtypes = [i64, i32]
    L.emrf_host_xorshift_next_double.restype = f64
    L.emrf_host_reservoir_sample.argtypes = [i32, i32, i64, vp]
    L.emrf_host_group_tree_order.argtypes = [vp, i32, vp]
    L.emrf_host_num_features_per_node.argtypes = [P(EmrfParams), i32]
    L.emrf_host_packed_slot_layout.argtypes = [vp, i32, i32, vp, vp]
    L.emrf_host_group_tiles_bound.argtypes = [vp, vp, i32, vp, i32, C.c_uint32]
    L.emrf_host_group_tiles_bound.restype = i64
    _lib = L
    return L


def check(rc):
    """Map a status code to the exception the JVM host would throw."""
    if rc == EMRF_OK:
        return
    msg = lib().emrf_last_error().decode(errors="replace")
    if rc == EMRF_ERR_ARG:
        raise ValueError(msg)          # IllegalArgumentException
    raise EmrfError(msg)               # RuntimeException / SparkException
