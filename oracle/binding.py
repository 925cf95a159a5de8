"""ctypes binding of the CPU oracle (oracle/liboracle.so).

TEST INFRASTRUCTURE ONLY — see the header of oracle/spark_rf_oracle.cpp.  Only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.
The product package (euromillioner_b200) must never import it.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


class OrcParams(C.Structure):
    _fields_ = [
        ("num_trees", C.c_int32), ("max_depth", C.c_int32), ("max_bins", C.c_int32),
        ("min_instances_per_node", C.c_int32), ("num_classes", C.c_int32),
        ("impurity", C.c_int32), ("feature_subset_kind", C.c_int32),
        ("feature_subset_value", C.c_double), ("min_info_gain", C.c_double),
        ("subsampling_rate", C.c_double),
        ("max_memory_mb", C.c_int32), ("find_splits_policy", C.c_int32),
        ("seed", C.c_int64),
        ("select_nodes_loop_policy", C.c_int32), ("pad", C.c_int32),
    ]


FEATURE_SUBSET = {"auto": 0, "all": 1, "sqrt": 2, "log2": 3, "onethird": 4}


def make_params(num_trees=20, max_depth=5, max_bins=32, num_classes=2, seed=207336481,
                impurity="gini", feature_subset="auto", subsampling_rate=1.0,
                min_instances_per_node=1, min_info_gain=0.0, max_memory_mb=256,
                find_splits_policy=0, select_nodes_loop_policy=0, cls=OrcParams):
    p = cls()
    p.num_trees, p.max_depth, p.max_bins = num_trees, max_depth, max_bins
    p.min_instances_per_node, p.num_classes = min_instances_per_node, num_classes
    p.impurity = {"gini": 0, "entropy": 1}[impurity]
    if isinstance(feature_subset, str) and feature_subset in FEATURE_SUBSET:
        p.feature_subset_kind, p.feature_subset_value = FEATURE_SUBSET[feature_subset], 0.0
    else:
        v = float(feature_subset)
        if float(feature_subset) == int(v) and "." not in str(feature_subset):
            p.feature_subset_kind, p.feature_subset_value = 5, v
        else:
            p.feature_subset_kind, p.feature_subset_value = 6, v
    p.min_info_gain, p.subsampling_rate = min_info_gain, subsampling_rate
    p.max_memory_mb, p.find_splits_policy, p.seed = max_memory_mb, find_splits_policy, seed
    if hasattr(p, "select_nodes_loop_policy"):
        p.select_nodes_loop_policy = select_nodes_loop_policy
    else:
        p.policy_flags = 1 if select_nodes_loop_policy else 0      # EmrfParams: bit 0
    return p


def build(force=False):
    so = os.path.join(_HERE, "liboracle.so")
    src = os.path.join(_HERE, "spark_rf_oracle.cpp")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(build())
        L = _LIB
        L.orc_last_error.restype = C.c_char_p
        L.orc_murmur3.restype = C.c_uint32
        L.orc_xorshift_hash_seed.restype = C.c_int64
        L.orc_xorshift_hash_seed.argtypes = [C.c_int64, C.c_int32]
        L.orc_fit.restype = C.c_void_p
        L.orc_forest_total_nodes.restype = C.c_int64
        L.orc_forest_total_nodes.argtypes = [C.c_void_p]
        L.orc_forest_trace_len.restype = C.c_int64
        L.orc_forest_trace_len.argtypes = [C.c_void_p]
        L.orc_forest_num_groups.argtypes = [C.c_void_p]
        L.orc_forest_f_sub.argtypes = [C.c_void_p]
        L.orc_forest_free.argtypes = [C.c_void_p]
        L.orc_find_splits_sample_rows.restype = C.c_int64
    return _LIB


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


def partition_offsets(n_rows, n_parts):
    """POLICY partition_slicing: P contiguous slices [floor(iN/P), floor((i+1)N/P))."""
    return np.array([(i * n_rows) // n_parts for i in range(n_parts + 1)], dtype=np.int64)


class OracleForest:
    def __init__(self, handle, num_trees, num_classes, num_features, max_bins):
        self.h, self.T, self.C, self.F, self.max_bins = handle, num_trees, num_classes, num_features, max_bins

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_forest_free(C.c_void_p(self.h))
            self.h = None

    def export(self):
        L = lib()
        n = L.orc_forest_total_nodes(C.c_void_p(self.h))
        out = {
            "tree_off": np.zeros(self.T + 1, np.int64), "node_id": np.zeros(n, np.int32),
            "left": np.zeros(n, np.int32), "right": np.zeros(n, np.int32), "feature": np.zeros(n, np.int32),
            "threshold": np.zeros(n, np.float64), "gain": np.zeros(n, np.float64),
            "impurity": np.zeros(n, np.float64), "prediction": np.zeros(n, np.float64),
            "stats": np.zeros((n, self.C), np.float64),
        }
        L.orc_forest_export(C.c_void_p(self.h), _p(out["tree_off"], C.c_int64), _p(out["node_id"], C.c_int32),
                            _p(out["left"], C.c_int32), _p(out["right"], C.c_int32), _p(out["feature"], C.c_int32),
                            _p(out["threshold"], C.c_double), _p(out["gain"], C.c_double),
                            _p(out["impurity"], C.c_double), _p(out["prediction"], C.c_double),
                            _p(out["stats"], C.c_double))
        return out

    def group_sizes(self):
        L = lib()
        g = np.zeros(L.orc_forest_num_groups(C.c_void_p(self.h)), np.int32)
        L.orc_forest_group_sizes(C.c_void_p(self.h), _p(g, C.c_int32))
        return g

    def trace(self):
        L = lib()
        tr = np.zeros(L.orc_forest_trace_len(C.c_void_p(self.h)), np.int32)
        L.orc_forest_trace(C.c_void_p(self.h), _p(tr, C.c_int32))
        return tr.reshape(-1, 2 + L.orc_forest_f_sub(C.c_void_p(self.h)))

    def thresholds(self):
        L = lib()
        thr = np.zeros((self.F, self.max_bins), np.float64)
        ns = np.zeros(self.F, np.int32)
        L.orc_forest_thresholds(C.c_void_p(self.h), C.c_int32(self.max_bins), _p(thr, C.c_double), _p(ns, C.c_int32))
        return thr, ns

    def predict(self, X, want_raw=True, want_prob=True, thresholds=None):
        L = lib()
        X = np.ascontiguousarray(X, np.float64)
        n = X.shape[0]
        raw = np.zeros((n, self.C)) if want_raw else None
        prob = np.zeros((n, self.C)) if want_prob else None
        pred = np.zeros(n)
        outs = (_p(raw, C.c_double) if want_raw else None, _p(prob, C.c_double) if want_prob else None, _p(pred, C.c_double))
        if thresholds is None:
            L.orc_predict(C.c_void_p(self.h), _p(X, C.c_double), C.c_int64(n), *outs)
        else:
            t = np.ascontiguousarray(thresholds, np.float64)
            L.orc_predict_thresholds(C.c_void_p(self.h), _p(X, C.c_double), C.c_int64(n), _p(t, C.c_double), *outs)
        return raw, prob, pred


def fit(params, X, y, part_off):
    L = lib()
    X = np.ascontiguousarray(X, np.float64)
    y = np.ascontiguousarray(y, np.float64)
    part_off = np.ascontiguousarray(part_off, np.int64)
    h = L.orc_fit(C.byref(params), _p(X, C.c_double), _p(y, C.c_double), C.c_int64(X.shape[0]),
                  C.c_int32(X.shape[1]), _p(part_off, C.c_int64), C.c_int32(len(part_off) - 1))
    if not h:
        raise ValueError(L.orc_last_error().decode())
    return OracleForest(h, params.num_trees, params.num_classes, X.shape[1], params.max_bins)


def find_splits(params, X, part_off):
    L = lib()
    X = np.ascontiguousarray(X, np.float64)
    part_off = np.ascontiguousarray(part_off, np.int64)
    F = X.shape[1]
    thr = np.zeros((F, params.max_bins), np.float64)
    ns = np.zeros(F, np.int32)
    rc = L.orc_find_splits(C.byref(params), _p(X, C.c_double), C.c_int64(X.shape[0]), C.c_int32(F),
                           _p(part_off, C.c_int64), C.c_int32(len(part_off) - 1), _p(thr, C.c_double),
                           _p(ns, C.c_int32))
    if rc != 0:
        raise ValueError(L.orc_last_error().decode())
    return thr, ns


def find_splits_sample_rows(params, n_rows, part_off):
    L = lib()
    part_off = np.ascontiguousarray(part_off, np.int64)
    cap = n_rows
    rows = np.zeros(cap, np.int64)
    n = L.orc_find_splits_sample_rows(C.byref(params), C.c_int64(n_rows), _p(part_off, C.c_int64),
                                      C.c_int32(len(part_off) - 1), _p(rows, C.c_int64), C.c_int64(cap))
    return rows[:n].copy()


def bin_features(X, thr, ns):
    L = lib()
    X = np.ascontiguousarray(X, np.float64)
    out = np.zeros(X.shape, np.uint8)
    L.orc_bin(_p(X, C.c_double), C.c_int64(X.shape[0]), C.c_int32(X.shape[1]), _p(thr, C.c_double),
              _p(ns, C.c_int32), C.c_int32(thr.shape[1]), _p(out, C.c_uint8))
    return out


def bag_weights(params, n_rows, part_off):
    L = lib()
    part_off = np.ascontiguousarray(part_off, np.int64)
    w = np.zeros((n_rows, params.num_trees), np.uint8)
    maxw = L.orc_bag_weights(C.byref(params), C.c_int64(n_rows), _p(part_off, C.c_int64),
                             C.c_int32(len(part_off) - 1), _p(w, C.c_uint8))
    return w, maxw


def best_split(hist, subset, num_bins_all, n_classes, impurity=0, min_instances=1, min_info_gain=0.0,
               node_calc=None):
    """hist: flat float64 [k][bin][class] with per-feature numBins.  Returns dict."""
    L = lib()
    hist = np.ascontiguousarray(hist, np.float64).ravel()
    subset = np.ascontiguousarray(subset, np.int32)
    nb = np.ascontiguousarray(num_bins_all, np.int32)
    k, s, valid = C.c_int32(), C.c_int32(), C.c_int32()
    gain, imp = C.c_double(), C.c_double()
    left = np.zeros(n_classes)
    right = np.zeros(n_classes)
    calc = None if node_calc is None else np.ascontiguousarray(node_calc, np.float64)
    L.orc_best_split(_p(hist, C.c_double), C.c_int64(hist.size), _p(subset, C.c_int32), C.c_int32(len(subset)),
                     _p(nb, C.c_int32), C.c_int32(n_classes), C.c_int32(impurity), C.c_int32(min_instances),
                     C.c_double(min_info_gain), _p(calc, C.c_double) if calc is not None else None,
                     C.byref(k), C.byref(s), C.byref(gain), C.byref(imp), _p(left, C.c_double),
                     _p(right, C.c_double), C.byref(valid))
    return {"k": k.value, "split": s.value, "gain": gain.value, "impurity": imp.value, "left": left,
            "right": right, "valid": bool(valid.value)}
