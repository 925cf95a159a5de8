"""Host-side mirror of Spark ML's RandomForestClassifier / RandomForestClassificationModel.

Same names, argument meaning and error behaviour as the Estimator/Model pair the reference would
call for this stage (Spark 2.4.6 ml/classification/RandomForestClassifier.scala; dependency pinned
at pom.xml:56-61; the stage it replaces in the reference program is the model train/predict block
Main.java:110-141).  Everything below the setters goes through the C ABI of include/emrf.h — there
is no Python or CPU compute path here.  (The Java 8 host + JNI shim for the same ABI are under
euromillioner_b200/host/; no JDK exists in this image, so they ship as source.)
"""
import ctypes as C
import threading

import numpy as np

from . import _lib
from ._lib import EmrfParams, EmrfTimings, check

DEFAULT_SEED = 207336481  # "org.apache.spark.ml.classification.RandomForestClassifier".hashCode
_SUBSET_KINDS = {"auto": 0, "all": 1, "sqrt": 2, "log2": 3, "onethird": 4}


def _ptr(a):
    return None if a is None else C.c_void_p(a.ctypes.data)


def java_double_to_string(x):
    """java.lang.Double.toString: what Scala string interpolation prints for a Double (`Predict: 12.0`, `<= 24.5`,
    `1.0E-5`, `1.0E7`).  Digits are the shortest that round-trip (Python's repr); Java switches to computerised scientific
    notation outside [1e-3, 1e7) and always shows at least one fractional digit.  (JDKs before 19 print a few doubles
    with one digit more than the shortest, JDK-4511638; none of the values a forest over integer-valued features
    holds — midpoints x.5 and class indices — is among them.)"""
    x = float(x)
    if x != x:
        return "NaN"
    if x in (float("inf"), float("-inf")):
        return "Infinity" if x > 0 else "-Infinity"
    if x == 0.0:
        return "-0.0" if str(x).startswith("-") else "0.0"
    sign = "-" if x < 0 else ""
    mant, _, exp = repr(abs(x)).partition("e")
    ip, _, fp = mant.partition(".")
    digits = (ip + fp).lstrip("0")
    # value = 0.DIGITS * 10^point
    point = len(ip) + (int(exp) if exp else 0)
    if ip.strip("0") == "":                                  # 0.000123 style mantissa
        lead = len(fp) - len(fp.lstrip("0"))
        point = -lead + (int(exp) if exp else 0)
    digits = digits.rstrip("0") or "0"
    if 1e-3 <= abs(x) < 1e7:
        if point <= 0:
            return sign + "0." + "0" * (-point) + digits
        if len(digits) <= point:
            return sign + digits + "0" * (point - len(digits)) + ".0"
        return sign + digits[:point] + "." + digits[point:]
    return sign + digits[0] + "." + (digits[1:] or "0") + "E" + str(point - 1)


def debug_string(e, num_trees, uid):
    """TreeEnsembleModel.toDebugString of Spark 2.4 for a flat export `e` (see emrf_forest_export): the header is
    RandomForestClassificationModel.toString, every tree is `  Tree i (weight w):` followed by
    Node.subtreeToString(indentFactor = 4): `If (feature f <= t)` / left subtree / `Else (feature f > t)` / right
    subtree, one more space per level, leaves `Predict: p`."""
    out = [f"RandomForestClassificationModel (uid={uid}) with {num_trees} trees\n"]
    for t in range(num_trees):
        out.append(f"  Tree {t} (weight 1.0):\n")
        base = int(e["tree_off"][t])
        stack = [(0, 4, None)]            # (node, indent, pending "Else" line to print first)
        while stack:
            i, ind, pre = stack.pop()
            if pre is not None:
                out.append(pre)
            pad = " " * ind
            if e["left"][base + i] < 0:
                out.append(f"{pad}Predict: {java_double_to_string(e['prediction'][base + i])}\n")
                continue
            f, thr = int(e["feature"][base + i]), java_double_to_string(e["threshold"][base + i])
            out.append(f"{pad}If (feature {f} <= {thr})\n")
            stack.append((int(e["right"][base + i]), ind + 1, f"{pad}Else (feature {f} > {thr})\n"))
            stack.append((int(e["left"][base + i]), ind + 1, None))
    return "".join(out)


class Context:
    """Owns one GPU (device, stream, scratch) — emrf_ctx_t."""

    def __init__(self, device=0, rank=0, world=1, nccl_uid=None):
        L = _lib.lib()
        h = C.c_void_p()
        if world > 1:
            uid = (C.c_uint8 * 128).from_buffer_copy(bytes(nccl_uid))
            check(L.emrf_ctx_create_dist(device, rank, world, uid, C.byref(h)))
        else:
            check(L.emrf_ctx_create(device, C.byref(h)))
        self.h, self.device, self.rank, self.world = h, device, rank, world

    @staticmethod
    def nccl_unique_id():
        uid = (C.c_uint8 * 128)()
        check(_lib.lib().emrf_nccl_unique_id(uid))
        return bytes(uid)

    def timings(self):
        t = EmrfTimings()
        check(_lib.lib().emrf_get_timings(self.h, C.byref(t)))
        return t.as_dict()

    def release_buffers(self):
        """Give the context's device buffers and the library's cached blocks back to the driver (the next fit
        re-allocates); returns the bytes freed on the device."""
        n = C.c_int64(0)
        check(_lib.lib().emrf_ctx_release_buffers(self.h, C.byref(n)))
        return int(n.value)

    def close(self):
        if getattr(self, "h", None):
            _lib.lib().emrf_ctx_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


_default_ctx = None
_default_lock = threading.Lock()


def default_context():
    global _default_ctx
    with _default_lock:
        if _default_ctx is None:
            _default_ctx = Context(0)
        return _default_ctx


class Dataset:
    """The slice of Dataset[Row] this stage reads: a dense `features` column (N x F float64), a
    `label` column (N float64) and the partition layout (Spark seeds bagging and split sampling per
    partition, so it is part of the input).  `features`/`label` may also be raw device pointers
    (int) when the data is already resident in HBM; then n_rows/n_features must be given."""

    def __init__(self, features, label=None, num_partitions=1, partition_offsets=None, n_rows=None, n_features=None,
                 global_partition_offsets=None, first_partition=0):
        if isinstance(features, (int, np.integer)):
            self.features, self.n_rows, self.n_features = int(features), int(n_rows), int(n_features)
        else:
            self.features = np.ascontiguousarray(features, np.float64)
            if self.features.ndim != 2:
                raise ValueError("features must be a 2-D array (rows x features)")
            self.n_rows, self.n_features = self.features.shape
        if label is None or isinstance(label, (int, np.integer)):
            self.label = None if label is None else int(label)
        else:
            self.label = np.ascontiguousarray(label, np.float64)
            if self.label.shape != (self.n_rows,):
                raise ValueError("label must have one value per row")
        if partition_offsets is None:
            # SparkContext.parallelize slicing: [floor(i*N/P), floor((i+1)*N/P))
            partition_offsets = [(i * self.n_rows) // num_partitions for i in range(num_partitions + 1)]
        self.partition_offsets = np.ascontiguousarray(partition_offsets, np.int64)
        # row-sharded fit: this object holds only the rows of global partitions
        # [first_partition, first_partition + n_local) of a Dataset described by global_partition_offsets
        self.global_partition_offsets = (None if global_partition_offsets is None
                                         else np.ascontiguousarray(global_partition_offsets, np.int64))
        self.first_partition = int(first_partition)
        self.columns = {}

    @classmethod
    def shard(cls, features_local, label_local, global_partition_offsets, first_partition, n_partitions_local):
        """The slice of a partitioned Dataset owned by one rank (emrf_fit_sharded)."""
        g = np.ascontiguousarray(global_partition_offsets, np.int64)
        local = g[first_partition:first_partition + n_partitions_local + 1] - g[first_partition]
        return cls(features_local, label_local, partition_offsets=local, global_partition_offsets=g,
                   first_partition=first_partition)

    def _fptr(self):
        return C.c_void_p(self.features) if isinstance(self.features, int) else _ptr(self.features)

    def _lptr(self):
        return C.c_void_p(self.label) if isinstance(self.label, int) else _ptr(self.label)


class RandomForestClassifier:
    """Estimator.  Setters return self, defaults are Spark's (ml/tree/treeParams.scala)."""

    def __init__(self, context=None):
        self._ctx = context
        self.numTrees, self.maxDepth, self.maxBins = 20, 5, 32
        self.minInstancesPerNode, self.minInfoGain = 1, 0.0
        self.impurity, self.featureSubsetStrategy = "gini", "auto"
        self.subsamplingRate, self.maxMemoryInMB, self.seed = 1.0, 256, DEFAULT_SEED
        self.cacheNodeIds, self.checkpointInterval = False, 10
        self.numClasses = 0  # 0 = infer from labels (Predictor.getNumClasses)
        self.findSplitsPolicy = 0
        self.selectNodesLoopPolicy = 0   # 1: pre-`groupDone` loop condition of selectNodesToSplit
        self.shardMode = 0

    # --- Spark ML setters -----------------------------------------------------------------
    def setNumTrees(self, v):
        if int(v) < 1:
            raise ValueError(f"numTrees must be >= 1 (got {v})")
        self.numTrees = int(v); return self

    def setMaxDepth(self, v):
        if not 0 <= int(v) <= 30:
            raise ValueError(f"maxDepth must be in [0, 30] (got {v})")
        self.maxDepth = int(v); return self

    def setMaxBins(self, v):
        if int(v) < 2:
            raise ValueError(f"maxBins must be >= 2 (got {v})")
        self.maxBins = int(v); return self

    def setMinInstancesPerNode(self, v):
        if int(v) < 1:
            raise ValueError("minInstancesPerNode must be >= 1")
        self.minInstancesPerNode = int(v); return self

    def setMinInfoGain(self, v):
        if float(v) < 0:
            raise ValueError("minInfoGain must be >= 0")
        self.minInfoGain = float(v); return self

    def setImpurity(self, v):
        if str(v).lower() not in ("gini", "entropy"):
            raise ValueError(f"impurity must be gini or entropy (got {v})")
        self.impurity = str(v).lower(); return self

    def setFeatureSubsetStrategy(self, v):
        s = str(v).lower()
        if s not in _SUBSET_KINDS:
            try:
                f = float(s)
            except ValueError:
                raise ValueError(f"unsupported featureSubsetStrategy: {v}")
            if not (f > 0):
                raise ValueError(f"unsupported featureSubsetStrategy: {v}")
        self.featureSubsetStrategy = s; return self

    def setSubsamplingRate(self, v):
        if not 0 < float(v) <= 1:
            raise ValueError("subsamplingRate must be in (0, 1]")
        self.subsamplingRate = float(v); return self

    def setMaxMemoryInMB(self, v):
        if int(v) < 0:
            raise ValueError("maxMemoryInMB must be >= 0")
        self.maxMemoryInMB = int(v); return self

    def setSeed(self, v):
        self.seed = int(v); return self

    def setCacheNodeIds(self, v):      # accepted; results do not depend on it
        self.cacheNodeIds = bool(v); return self

    def setCheckpointInterval(self, v):  # accepted; only meaningful with cacheNodeIds in Spark
        self.checkpointInterval = int(v); return self

    def setThresholds(self, v):
        """ProbabilisticClassifierParams.thresholds: one value per class, all >= 0, at most one 0; the model then
        predicts the class with the largest probability / threshold."""
        t = None if v is None else [float(x) for x in v]
        if t is not None:
            if any(not (x >= 0.0) for x in t):
                raise ValueError("thresholds must all be >= 0")
            if sum(1 for x in t if x == 0.0) > 1:
                raise ValueError("at most one threshold may be 0")
        self._thresholds = t
        return self

    def setFeaturesCol(self, v):       # column names are accepted for source compatibility; a Dataset here has one
        return self                    # features matrix and one label vector

    def setLabelCol(self, v):
        return self

    def setPredictionCol(self, v):
        return self

    def setRawPredictionCol(self, v):
        return self

    def setProbabilityCol(self, v):
        return self

    def setNumClasses(self, v):        # stands in for label-column NominalAttribute metadata
        self.numClasses = int(v); return self

    def _params(self):
        p = EmrfParams()
        p.num_trees, p.max_depth, p.max_bins = self.numTrees, self.maxDepth, self.maxBins
        p.min_instances_per_node, p.num_classes = self.minInstancesPerNode, self.numClasses
        p.impurity = 0 if self.impurity == "gini" else 1
        s = self.featureSubsetStrategy
        if s in _SUBSET_KINDS:
            p.feature_subset_kind, p.feature_subset_value = _SUBSET_KINDS[s], 0.0
        elif s.isdigit():
            p.feature_subset_kind, p.feature_subset_value = 5, float(int(s))
        else:
            p.feature_subset_kind, p.feature_subset_value = 6, float(s)
        p.min_info_gain, p.subsampling_rate = self.minInfoGain, self.subsamplingRate
        p.max_memory_mb, p.find_splits_policy, p.seed = self.maxMemoryInMB, self.findSplitsPolicy, self.seed
        p.shard_mode = self.shardMode
        p.policy_flags = 1 if self.selectNodesLoopPolicy else 0
        return p

    def fit(self, dataset):
        """Predictor.fit → RandomForestClassifier.train → RandomForest.run.  Synchronous."""
        ctx = self._ctx or default_context()
        L = _lib.lib()
        if dataset.label is None:
            raise ValueError("dataset has no label column")
        p = self._params()
        h = C.c_void_p()
        po = dataset.partition_offsets
        if dataset.global_partition_offsets is not None:
            g = dataset.global_partition_offsets
            p.shard_mode = 1
            check(L.emrf_fit_sharded(ctx.h, C.byref(p), dataset._fptr(), dataset._lptr(), dataset.n_rows, dataset.n_features,
                                     g.ctypes.data_as(C.POINTER(C.c_int64)), len(g) - 1, dataset.first_partition,
                                     len(po) - 1, C.byref(h)))
        else:
            check(L.emrf_fit(ctx.h, C.byref(p), dataset._fptr(), dataset._lptr(), dataset.n_rows, dataset.n_features,
                             po.ctypes.data_as(C.POINTER(C.c_int64)), len(po) - 1, C.byref(h)))
        model = RandomForestClassificationModel(ctx, h)
        thr = getattr(self, "_thresholds", None)
        if thr is not None:          # copyValues(model): the estimator's thresholds travel to the model
            if len(thr) != model.numClasses:   # Spark checks this in train(): require(thresholds.length == numClasses)
                raise ValueError(f"RandomForestClassifier.train() called with non-matching numClasses and thresholds.length. "
                                 f"numClasses={model.numClasses}, but thresholds has length {len(thr)}")
            model._thresholds = list(thr)
        return model


class RandomForestClassificationModel:
    """Model.  transform adds rawPrediction / probability / prediction (first-max argmax)."""

    def __init__(self, ctx, handle):
        self._ctx, self._h = ctx, handle
        L = _lib.lib()
        t, c, f, n = C.c_int32(), C.c_int32(), C.c_int32(), C.c_int64()
        check(L.emrf_forest_info(handle, C.byref(t), C.byref(c), C.byref(f), C.byref(n)))
        self.numTrees, self.numClasses, self.numFeatures, self.totalNumNodes = t.value, c.value, f.value, n.value
        self.treeWeights = [1.0] * self.numTrees
        self._export = None
        self._thresholds = None
        self.uid = "rfc_" + "%012x" % (abs(hash((self.numTrees, self.totalNumNodes, id(self)))) & 0xffffffffffff)   # Identifiable.randomUID("rfc")

    def setThresholds(self, v):
        """ProbabilisticClassificationModel.setThresholds (validated like the estimator's)."""
        RandomForestClassifier.setThresholds(self, v)
        return self

    def __del__(self):
        try:
            if self._h:
                _lib.lib().emrf_forest_free(self._h)
                self._h = None
        except Exception:
            pass

    @classmethod
    def from_arrays(cls, e, num_features, ctx=None):
        """Rebuild a model from the flat export (the persistence hook)."""
        ctx = ctx or default_context()
        L = _lib.lib()
        T = len(e["tree_off"]) - 1
        Cn = e["stats"].shape[1]
        arrs = [np.ascontiguousarray(e[k], dt) for k, dt in (
            ("tree_off", np.int64), ("node_id", np.int32), ("left", np.int32), ("right", np.int32),
            ("feature", np.int32), ("threshold", np.float64), ("gain", np.float64), ("impurity", np.float64),
            ("prediction", np.float64), ("stats", np.float64))]
        h = C.c_void_p()
        check(L.emrf_forest_import(ctx.h, T, Cn, num_features, *[_ptr(a) for a in arrs], C.byref(h)))
        return cls(ctx, h)

    def export(self):
        """Flat arrays: trees in index order, nodes pre-order (see emrf_forest_export)."""
        if self._export is None:
            n, Cn = self.totalNumNodes, self.numClasses
            e = {"tree_off": np.zeros(self.numTrees + 1, np.int64), "node_id": np.zeros(n, np.int32),
                 "left": np.zeros(n, np.int32), "right": np.zeros(n, np.int32), "feature": np.zeros(n, np.int32),
                 "threshold": np.zeros(n), "gain": np.zeros(n), "impurity": np.zeros(n), "prediction": np.zeros(n),
                 "stats": np.zeros((n, Cn))}
            check(_lib.lib().emrf_forest_export(self._h, *[_ptr(e[k]) for k in (
                "tree_off", "node_id", "left", "right", "feature", "threshold", "gain", "impurity", "prediction", "stats")]))
            self._export = e
        return self._export

    def toJson(self):
        """The forest as a JSON document: per tree a nested node object in the vocabulary of
        Spark's `NodeData` (`id`, `prediction`, `impurity`, `impurityStats`, `gain`, `split.featureIndex`,
        `split.threshold`, `leftChild`/`rightChild`), doubles written with repr() so that the round trip is exact."""
        import json
        e = self.export()

        def node(base, i):
            d = {"id": int(e["node_id"][base + i]), "prediction": float(e["prediction"][base + i]),
                 "impurity": float(e["impurity"][base + i]), "impurityStats": [float(v) for v in e["stats"][base + i]]}
            if e["left"][base + i] >= 0:
                d["gain"] = float(e["gain"][base + i])
                d["split"] = {"featureIndex": int(e["feature"][base + i]), "threshold": float(e["threshold"][base + i])}
                d["leftChild"] = node(base, int(e["left"][base + i]))
                d["rightChild"] = node(base, int(e["right"][base + i]))
            return d

        import sys
        sys.setrecursionlimit(max(sys.getrecursionlimit(), 10000))
        return json.dumps({"class": "org.apache.spark.ml.classification.RandomForestClassificationModel",
                           "numFeatures": self.numFeatures, "numClasses": self.numClasses,
                           "trees": [node(int(e["tree_off"][t]), 0) for t in range(self.numTrees)]})

    @classmethod
    def fromJson(cls, text, ctx=None):
        """Inverse of toJson (nodes are laid out pre-order, left subtree first, as emrf_forest_export does)."""
        import json
        doc = json.loads(text)
        cols = {k: [] for k in ("node_id", "left", "right", "feature", "threshold", "gain", "impurity", "prediction", "stats")}
        tree_off = [0]

        def put(d, base):
            i = len(cols["node_id"]) - base
            for k, v in (("node_id", d["id"]), ("left", -1), ("right", -1), ("feature", -1), ("threshold", 0.0),
                         ("gain", d.get("gain", 0.0)), ("impurity", d["impurity"]), ("prediction", d["prediction"]),
                         ("stats", d["impurityStats"])):
                cols[k].append(v)
            if "split" in d:
                at = base + i
                cols["feature"][at] = d["split"]["featureIndex"]
                cols["threshold"][at] = d["split"]["threshold"]
                cols["left"][at] = put(d["leftChild"], base)
                cols["right"][at] = put(d["rightChild"], base)
            return i

        import sys
        sys.setrecursionlimit(max(sys.getrecursionlimit(), 10000))
        for tree in doc["trees"]:
            put(tree, tree_off[-1])
            tree_off.append(len(cols["node_id"]))
        e = {k: np.asarray(v) for k, v in cols.items()}
        e["stats"] = np.asarray(cols["stats"], np.float64).reshape(len(cols["node_id"]), doc["numClasses"])
        e["tree_off"] = np.asarray(tree_off, np.int64)
        return cls.from_arrays(e, doc["numFeatures"], ctx)

    def trace(self):
        """(group_sizes, per-popped-node rows [tree, node id, F_sub features]) of the fit."""
        L = _lib.lib()
        g, n, fs = C.c_int32(), C.c_int64(), C.c_int32()
        check(L.emrf_forest_trace_info(self._h, C.byref(g), C.byref(n), C.byref(fs)))
        gs, tr = np.zeros(g.value, np.int32), np.zeros(n.value, np.int32)
        check(L.emrf_forest_trace(self._h, _ptr(gs), _ptr(tr)))
        return gs, tr.reshape(-1, 2 + fs.value) if fs.value >= 0 and n.value else tr

    @property
    def featureImportances(self):
        out = np.zeros(self.numFeatures)
        check(_lib.lib().emrf_forest_feature_importances(self._h, _ptr(out)))
        return out

    def toString(self):
        """RandomForestClassificationModel.toString (Spark 2.4: the uid is part of it)."""
        return f"RandomForestClassificationModel (uid={self.uid}) with {self.numTrees} trees"

    __str__ = toString

    @property
    def toDebugString(self):
        """Spark 2.4's text form, byte for byte (given the same uid): diffing it against a real Spark run of
        tools/SparkGolden.java is the way to pin parity."""
        return debug_string(self.export(), self.numTrees, self.uid)

    # ---- Spark ML persistence --------------------------------------------------------
    def save(self, path, uid="rfc_b200", params=None):
        """Writes the directory layout of Spark 2.4.x `RandomForestClassificationModel.save`
        (ml/tree/treeModels.scala EnsembleModelReadWrite.saveImpl): metadata/part-00000 (one JSON line),
        treesMetadata/ (treeID, metadata, weights) and data/ (treeID, nodeData) as Parquet.  NodeData ids are
        pre-order indices inside each tree, leaves carry leftChild = rightChild = -1, gain -1.0 and a
        SplitData(-1, [], -1).  Restated from the upstream source; NOT checked against a JVM (none here)."""
        import json
        import os
        import time
        import pyarrow as pa
        import pyarrow.parquet as pq
        e = self.export()
        os.makedirs(os.path.join(path, "metadata"), exist_ok=True)
        os.makedirs(os.path.join(path, "treesMetadata"), exist_ok=True)
        os.makedirs(os.path.join(path, "data"), exist_ok=True)
        now = int(time.time() * 1000)
        meta = {"class": "org.apache.spark.ml.classification.RandomForestClassificationModel", "timestamp": now,
                "sparkVersion": "2.4.6", "uid": uid, "paramMap": dict(params or {}), "defaultParamMap": {},
                "numFeatures": int(self.numFeatures), "numClasses": int(self.numClasses), "numTrees": int(self.numTrees)}
        with open(os.path.join(path, "metadata", "part-00000"), "w") as f:
            f.write(json.dumps(meta) + "\n")
        tree_meta = [json.dumps({"class": "org.apache.spark.ml.classification.DecisionTreeClassificationModel",
                                 "timestamp": now, "sparkVersion": "2.4.6", "uid": f"dtc_b200_{t}", "paramMap": {},
                                 "defaultParamMap": {}, "numFeatures": int(self.numFeatures),
                                 "numClasses": int(self.numClasses)}) for t in range(self.numTrees)]
        pq.write_table(pa.table({"treeID": pa.array(range(self.numTrees), pa.int32()), "metadata": pa.array(tree_meta),
                                 "weights": pa.array([1.0] * self.numTrees, pa.float64())}),
                       os.path.join(path, "treesMetadata", "part-00000.parquet"))
        split_t = pa.struct([("featureIndex", pa.int32()), ("leftCategoriesOrThreshold", pa.list_(pa.float64())),
                             ("numCategories", pa.int32())])
        node_t = pa.struct([("id", pa.int32()), ("prediction", pa.float64()), ("impurity", pa.float64()),
                            ("impurityStats", pa.list_(pa.float64())), ("gain", pa.float64()), ("leftChild", pa.int32()),
                            ("rightChild", pa.int32()), ("split", split_t)])
        tree_ids, nodes = [], []
        for t in range(self.numTrees):
            a, b = int(e["tree_off"][t]), int(e["tree_off"][t + 1])
            for i in range(a, b):
                leaf = e["left"][i] < 0
                nodes.append({"id": i - a, "prediction": float(e["prediction"][i]), "impurity": float(e["impurity"][i]),
                              "impurityStats": [float(v) for v in e["stats"][i]],
                              "gain": -1.0 if leaf else float(e["gain"][i]),
                              "leftChild": -1 if leaf else int(e["left"][i]), "rightChild": -1 if leaf else int(e["right"][i]),
                              "split": {"featureIndex": -1, "leftCategoriesOrThreshold": [], "numCategories": -1} if leaf else
                                       {"featureIndex": int(e["feature"][i]), "leftCategoriesOrThreshold": [float(e["threshold"][i])],
                                        "numCategories": -1}})
                tree_ids.append(t)
        pq.write_table(pa.table({"treeID": pa.array(tree_ids, pa.int32()), "nodeData": pa.array(nodes, node_t)}),
                       os.path.join(path, "data", "part-00000.parquet"))

    @classmethod
    def load(cls, path, ctx=None):
        """Reads a directory written by `save` (or by Spark 2.4.x, as far as the restated layout is right)."""
        import glob
        import json
        import os
        import pyarrow.parquet as pq
        with open(os.path.join(path, "metadata", "part-00000")) as f:
            meta = json.loads(f.readline())
        rows = []
        for fn in sorted(glob.glob(os.path.join(path, "data", "*.parquet"))):
            tb = pq.read_table(fn).to_pydict()
            rows += list(zip(tb["treeID"], tb["nodeData"]))
        rows.sort(key=lambda r: (r[0], r[1]["id"]))
        T, Cn = int(meta["numTrees"]), int(meta["numClasses"])
        n = len(rows)
        e = {"tree_off": np.zeros(T + 1, np.int64), "node_id": np.zeros(n, np.int32), "left": np.full(n, -1, np.int32),
             "right": np.full(n, -1, np.int32), "feature": np.full(n, -1, np.int32), "threshold": np.zeros(n),
             "gain": np.zeros(n), "impurity": np.zeros(n), "prediction": np.zeros(n), "stats": np.zeros((n, Cn))}
        for i, (t, nd) in enumerate(rows):
            e["tree_off"][t + 1] += 1
            e["prediction"][i], e["impurity"][i] = nd["prediction"], nd["impurity"]
            e["stats"][i] = nd["impurityStats"]
            if nd["leftChild"] >= 0:
                e["left"][i], e["right"][i] = nd["leftChild"], nd["rightChild"]
                e["feature"][i] = nd["split"]["featureIndex"]
                e["threshold"][i] = nd["split"]["leftCategoriesOrThreshold"][0]
                e["gain"][i] = nd["gain"]
        e["tree_off"] = np.cumsum(e["tree_off"])
        return cls.from_arrays(e, int(meta["numFeatures"]), ctx)

    def transform(self, dataset, raw=True, probability=True):
        """Returns the dataset with rawPrediction (N x C), probability (N x C), prediction (N)."""
        if dataset.n_features != self.numFeatures:
            raise ValueError(f"features column has {dataset.n_features} values per row, model expects {self.numFeatures}")
        n, Cn = dataset.n_rows, self.numClasses
        rawp = np.zeros((n, Cn)) if raw else None
        prob = np.zeros((n, Cn)) if probability else None
        pred = np.zeros(n)
        thr = getattr(self, "_thresholds", None)
        if thr is None:
            check(_lib.lib().emrf_predict(self._ctx.h, self._h, dataset._fptr(), n, _ptr(rawp), _ptr(prob), _ptr(pred)))
        else:
            if len(thr) != Cn:       # Spark: require(thresholds.length == numClasses)
                raise ValueError(f"thresholds has length {len(thr)}, but the model has {Cn} classes")
            t = np.ascontiguousarray(thr, np.float64)
            check(_lib.lib().emrf_predict_thresholds(self._ctx.h, self._h, dataset._fptr(), n, _ptr(t), _ptr(rawp), _ptr(prob),
                                                     _ptr(pred)))
        dataset.columns.update({"rawPrediction": rawp, "probability": prob, "prediction": pred})
        return dataset
