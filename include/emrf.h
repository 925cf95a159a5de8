/*
 * emrf.h — C ABI of the B200-native random-forest engine (libemrf.so).
 *
 * This is the drop-in boundary for the ONE hot path of mareksagan/Euromillioner that this
 * repository accelerates: the Spark MLlib 2.4.6 RandomForestClassifier fit/transform stage
 * (SURVEY.md §8).  The reference exposes no FFI/plugin interface of its own (its only native
 * boundary is inside the undeclared xgboost4j jar, Main.java:110,137,140 — the model stage these
 * entry points replace; the dependency that holds the RF arithmetic is pinned at pom.xml:56-61).
 * Each entry point below therefore cites the Spark ML / MLlib interface it stands in for
 * (upstream path under spark/mllib/src/main/scala/org/apache/spark/).
 *
 * Conventions: plain C, POD structs, caller-allocated outputs or opaque handles with a matching
 * _free/_destroy; every function returns int status (0 = OK, <0 = error) and never throws;
 * emrf_last_error() returns the calling thread's last message.  Pointers named *_hd may point
 * to host or device memory (detected with cudaPointerGetAttributes); all others are host.
 * There is NO CPU fallback: every compute entry point fails with EMRF_ERR_CUDA when no usable
 * sm_100 device is present.
 */
#ifndef EMRF_H_
#define EMRF_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define EMRF_OK 0
#define EMRF_ERR_ARG (-1)     /* IllegalArgumentException on the JVM side (Spark `require`) */
#define EMRF_ERR_CUDA (-2)    /* RuntimeException / SparkException on the JVM side */
#define EMRF_ERR_LIMIT (-3)   /* shape outside what this build supports */
#define EMRF_ERR_INTERNAL (-4)

typedef struct emrf_ctx emrf_ctx_t;       /* device, stream, scratch; one fit at a time per ctx */
typedef struct emrf_forest emrf_forest_t; /* host + device copy of a trained forest; immutable */

/* Spark ML RandomForestClassifier params (ml/classification/RandomForestClassifier.scala,
 * ml/tree/treeParams.scala): setNumTrees/setMaxDepth/setMaxBins/setMinInstancesPerNode/
 * setImpurity/setFeatureSubsetStrategy/setMinInfoGain/setSubsamplingRate/setMaxMemoryInMB/setSeed. */
typedef struct {
  int32_t num_trees;               /* default 20 */
  int32_t max_depth;               /* default 5, <= 30 */
  int32_t max_bins;                /* default 32, 2..256 */
  int32_t min_instances_per_node;  /* default 1 */
  int32_t num_classes;             /* 0 = infer max(label)+1 (<= 100, Predictor.getNumClasses) */
  int32_t impurity;                /* 0 gini, 1 entropy (entropy gains are not bit-exact) */
  int32_t feature_subset_kind;     /* 0 auto, 1 all, 2 sqrt, 3 log2, 4 onethird, 5 n, 6 fraction */
  double feature_subset_value;     /* for kinds 5 and 6 */
  double min_info_gain;            /* default 0.0 */
  double subsampling_rate;         /* default 1.0, (0,1] */
  int32_t max_memory_mb;           /* default 256: bounds nodes per group exactly as Spark does */
  int32_t find_splits_policy;      /* 0 = Spark 2.4 (zero filter + expected sample count); 1 = pre-2.4 */
  int64_t seed;                    /* default 207336481 = "…RandomForestClassifier".hashCode */
  int32_t shard_mode;              /* multi-GPU: 0 tree-sharded, 1 row-sharded (ignored when world==1) */
  int32_t policy_flags;            /* bit 0: RandomForest.selectNodesToSplit ends a group with the pre-`groupDone` loop condition
                                      (`while (memUsage < maxMemoryUsage || memUsage == 0)`): differs from the default only when a
                                      group's memory lands exactly on maxMemoryInMB */
} emrf_params_t;

/* Phase timers named after Spark's TimeTracker phases (ml/tree/impl/TimeTracker.scala), device
 * time measured with CUDA events on the ctx stream, plus the counters bench.py needs. */
typedef struct {
  double total_ms;          /* whole emrf_fit: first byte on device → forest on host */
  double h2d_ms;            /* host→device copy of features/labels (0 when inputs were device-resident) */
  double find_splits_ms;    /* findSplits */
  double bin_ms;            /* TreePoint conversion (K1) */
  double bag_ms;            /* remaining wait for the bag kernel after binning + per-tree row lists */
  double find_best_splits_ms; /* all groups: histogram + split scan kernels (K2+K3), device time */
  double hist_kernel_ms;    /* of which: the histogram(+fused scan) kernel, summed over launches */
  double partition_ms;      /* row→node partition update kernels (K4), summed */
  double host_scheduler_ms; /* selectNodesToSplit + driver-side tree updates, wall time */
  int64_t hist_launches;    /* number of histogram kernel launches (= number of groups) */
  int64_t kernel_launches;  /* all kernels launched by the fit */
  int64_t hist_alg_bytes;   /* algorithmic bytes of the histogram kernel: sum (F_sub+6)*rows + hist out */
  int64_t hist_units;       /* in-bag (row,tree) visits summed over groups */
  int64_t num_groups;
  int64_t total_nodes_split; /* nodes that went through findBestSplits */
  int64_t gain_mismatches;  /* device-vs-host recomputed gain differences (must be 0) */
  int64_t h2d_bytes;        /* bytes copied host→device inside the fit (inputs + group descriptors) */
  int64_t d2h_bytes;        /* bytes copied device→host inside the fit (split records, segment sizes …) */
  double host_select_ms;    /* host_scheduler_ms split: selectNodesToSplit + descriptor staging */
  double host_update_ms;    /*   driver-side node updates after each group */
  double host_export_ms;    /*   LearningNode.toNode(prune) + flat export */
  double bag_kernel_ms;     /* the Poisson kernel alone (second stream, overlaps h2d/findSplits/binning) */
  double comm_ms;           /* multi-GPU: device time of the per-group collectives (reduceByKey's merge of the node
                               histograms: reduce-scatter, and the all-reduce of the split records), CUDA events */
  int64_t comm_bytes;       /*   bytes this rank handed to those collectives (send buffers, summed over groups) */
  int64_t gains_verified;   /* EMRF_VERIFY_GAINS=1: split records recomputed on the host (gain_mismatches counts the differences) */
} emrf_timings_t;

const char* emrf_last_error(void);
/* build/version string; also proves the library loaded */
const char* emrf_version(void);

/* ---- context ------------------------------------------------------------------------------ */
int emrf_ctx_create(int32_t device, emrf_ctx_t** out);
/* Multi-GPU: one process per GPU.  `nccl_uid` is the 128-byte ncclUniqueId made by rank 0 with
 * emrf_nccl_unique_id and broadcast by the launcher (torch.distributed / any side channel). */
int emrf_nccl_unique_id(uint8_t* uid128);
int emrf_ctx_create_dist(int32_t device, int32_t rank, int32_t world, const uint8_t* nccl_uid,
                         emrf_ctx_t** out);
void emrf_ctx_destroy(emrf_ctx_t* ctx);
int emrf_get_timings(const emrf_ctx_t* ctx, emrf_timings_t* out);
/* A context keeps its fit-sized device buffers (TreePoint matrices, entry lists, histogram slots: ~1.4 GB after a
 * configs[1] fit, tens of GB at 10M x 128) and the library keeps freed blocks in a per-device cache, so that the next fit
 * allocates nothing.  A long-lived host (a Spark executor) calls this between jobs to give all of it back to the driver;
 * the next fit re-allocates.  Not to be called while a fit or predict on this context is running. */
int emrf_ctx_release_buffers(emrf_ctx_t* ctx, int64_t* bytes_released);

/* Device/pinned memory helpers for hosts without a CUDA binding of their own (the JNI shim). */
int emrf_dev_alloc(emrf_ctx_t* ctx, int64_t bytes, void** out);
int emrf_dev_free(emrf_ctx_t* ctx, void* p);
int emrf_host_alloc_pinned(int64_t bytes, void** out);
int emrf_host_free_pinned(void* p);
int emrf_memcpy_h2d(emrf_ctx_t* ctx, void* dst_dev, const void* src_host, int64_t bytes);
int emrf_memcpy_d2h(emrf_ctx_t* ctx, void* dst_host, const void* src_dev, int64_t bytes);

/* ---- Estimator: RandomForestClassifier.fit → RandomForest.run ------------------------------
 * Replaces ml/classification/RandomForestClassifier.train + ml/tree/impl/RandomForest.run.
 * features: row-major N×F float64 (a Dataset's dense `features` vectors), labels: N float64
 * (integral, 0 <= label < num_classes, as Predictor.extractLabeledPoints validates).
 * partition_offsets[P+1]: the Dataset's partition layout (rows of partition p are
 * [off[p], off[p+1])) — part of the contract because Spark seeds bagging and split-sampling per
 * partition (BaggedPoint.convertToBaggedRDD, RDD.sample). */
int emrf_fit(emrf_ctx_t* ctx, const emrf_params_t* params, const double* features_hd,
             const double* labels_hd, int64_t n_rows, int32_t n_features,
             const int64_t* partition_offsets, int32_t n_partitions, emrf_forest_t** out);

/* Row-sharded fit (one process per GPU; ctx made by emrf_ctx_create_dist).  The Dataset's rows are
 * split across ranks at partition boundaries: this rank passes only the rows of global partitions
 * [first_partition, first_partition + n_partitions_local), every rank passes the same global
 * partition_offsets.  Per group each rank builds partial node histograms for its rows and the ranks
 * sum them with an NCCL all-reduce — the role of reduceByKey(_.merge(_)) over the per-partition
 * DTStatsAggregators in RandomForest.findBestSplits.  Every rank returns the same forest, identical
 * to the single-GPU forest on the concatenated rows (integer histograms: reduction order is moot). */
int emrf_fit_sharded(emrf_ctx_t* ctx, const emrf_params_t* params, const double* features_hd,
                     const double* labels_hd, int64_t n_rows_local, int32_t n_features,
                     const int64_t* partition_offsets_global, int32_t n_partitions_global,
                     int32_t first_partition, int32_t n_partitions_local, emrf_forest_t** out);

/* ---- Model: RandomForestClassificationModel ------------------------------------------------ */
void emrf_forest_free(emrf_forest_t* f);
int emrf_forest_info(const emrf_forest_t* f, int32_t* num_trees, int32_t* num_classes,
                     int32_t* num_features, int64_t* total_nodes);
/* Flat export, trees in index order, nodes pre-order per tree (what Model.trees / toDebugString
 * walk): tree_off[T+1]; per node: Spark node id, left/right (index within the tree, -1 = leaf),
 * split feature (-1 = leaf), threshold, gain, impurity, prediction, class counts stats[C]. */
int emrf_forest_export(const emrf_forest_t* f, int64_t* tree_off, int32_t* node_id, int32_t* left,
                       int32_t* right, int32_t* feature, double* threshold, double* gain,
                       double* impurity, double* prediction, double* stats);
int emrf_forest_import(emrf_ctx_t* ctx, int32_t num_trees, int32_t num_classes, int32_t num_features,
                       const int64_t* tree_off, const int32_t* node_id, const int32_t* left,
                       const int32_t* right, const int32_t* feature, const double* threshold,
                       const double* gain, const double* impurity, const double* prediction,
                       const double* stats, emrf_forest_t** out);
/* Scheduler trace of the fit that made this forest: group sizes, and per popped node
 * (tree, node id, F_sub feature ids) in pop order — what selectNodesToSplit decided. */
int emrf_forest_trace_info(const emrf_forest_t* f, int32_t* num_groups, int64_t* trace_len, int32_t* f_sub);
int emrf_forest_trace(const emrf_forest_t* f, int32_t* group_sizes, int32_t* trace);
/* featureImportances (ml/tree/treeModels.scala TreeEnsembleModel.featureImportances) */
int emrf_forest_feature_importances(const emrf_forest_t* f, double* out /* F */);

/* RandomForestClassificationModel.transform: predictRaw / raw2probabilityInPlace / raw2prediction.
 * raw/prob: N×C or NULL; pred: N or NULL.  Safe for concurrent callers on one ctx. */
int emrf_predict(emrf_ctx_t* ctx, const emrf_forest_t* f, const double* features_hd, int64_t n_rows,
                 double* raw_hd, double* prob_hd, double* pred_hd);

/* Same, with ProbabilisticClassifierParams.thresholds set (Spark ML `setThresholds`; `[S]`
 * ProbabilisticClassificationModel.raw2prediction -> probability2prediction): prediction = first index of the
 * largest probability(c) / thresholds[c] (a class with threshold 0 and non-zero probability always wins; NaN
 * never does).  `thresholds` is a HOST array of num_classes doubles, all >= 0 and at most one 0 (EMRF_ERR_ARG
 * otherwise, as Spark's `require`); NULL = emrf_predict.  raw / probability outputs are unaffected. */
int emrf_predict_thresholds(emrf_ctx_t* ctx, const emrf_forest_t* f, const double* features_hd, int64_t n_rows,
                            const double* thresholds, double* raw_hd, double* prob_hd, double* pred_hd);

/* ---- stage-level entry points (each kernel parity-checked alone) ---------------------------- */
/* RandomForest.findSplits: thresholds_out[F*max_bins] (row f holds n_splits_out[f] values). */
int emrf_find_splits(emrf_ctx_t* ctx, const emrf_params_t* params, const double* features_hd, int64_t n_rows,
                     int32_t n_features, const int64_t* partition_offsets, int32_t n_partitions,
                     double* thresholds_out, int32_t* n_splits_out);
/* the rows RDD.sample keeps for findSplits (host logic; no GPU needed) */
int64_t emrf_find_splits_sample_rows(const emrf_params_t* params, int64_t n_rows, const int64_t* partition_offsets,
                                     int32_t n_partitions, int64_t* rows_out, int64_t cap);
/* TreePoint.convertToTreeRDD (K1): bins_cm_out[F*N] column-major, bins_rm_out[N*F] row-major
 * (either may be NULL). */
int emrf_bin(emrf_ctx_t* ctx, const double* features_hd, int64_t n_rows, int32_t n_features,
             const double* thresholds, const int32_t* n_splits, int32_t thr_stride, uint8_t* bins_cm_out,
             uint8_t* bins_rm_out);
/* BaggedPoint.convertToBaggedRDD: weights_out[N*T] row-major uint8 */
int emrf_bag_weights(emrf_ctx_t* ctx, const emrf_params_t* params, int64_t n_rows, const int64_t* partition_offsets,
                     int32_t n_partitions, uint8_t* weights_out, int32_t* max_weight_out);
/* DTStatsAggregator (K2): nodes are given as segments of `entries` (packed row | weight<<32 |
 * label<<40); subsets[n_nodes*f_sub]; hist_out[n_nodes][f_sub][max_nb][C] uint32 where max_nb =
 * max over features of num_bins_all. */
int emrf_histogram(emrf_ctx_t* ctx, const uint8_t* bins_rm, int64_t n_rows, int32_t n_features,
                   const uint64_t* entries, const int64_t* node_seg_off, int32_t n_nodes, const int32_t* subsets,
                   int32_t f_sub, const int32_t* num_bins_all, int32_t n_classes, uint32_t* hist_out);
/* binsToBestSplit (K3) on histograms laid out as emrf_histogram writes them.  Outputs per node:
 * k (index into the subset, -1 = no feature with splits), split index, gain, valid flag,
 * left[C] and total[C] class counts. */
int emrf_best_splits(emrf_ctx_t* ctx, const uint32_t* hist, int32_t n_nodes, const int32_t* subsets, int32_t f_sub,
                     const int32_t* num_bins_all, int32_t n_features, int32_t n_classes, int32_t impurity,
                     int32_t min_instances_per_node, double min_info_gain, int32_t* k_out, int32_t* split_out,
                     double* gain_out, int32_t* valid_out, uint32_t* left_out, uint32_t* total_out);
/* LearningNode.predictImpl one level (K4): stable partition of each node's entries by
 * bin(feature) <= split_bin.  entries_out has the same segments; n_left_out[n_nodes]. */
int emrf_update_partition(emrf_ctx_t* ctx, const uint8_t* bins_cm, int64_t n_rows, int32_t n_features,
                          const uint64_t* entries, const int64_t* node_seg_off, int32_t n_nodes,
                          const int32_t* split_feature, const int32_t* split_bin, uint64_t* entries_out,
                          int64_t* n_left_out);

/* Draw-history feature builder (reference analogue: the row -> features loop of
 * Main.java:86-102).  draws: n_draws x 7 int32 (5 mains, 2 stars).  Output rows t = lag..n_draws-1:
 * features_hd[(t-lag)*7*lag + (l-1)*7 + q] = draws[t-l][q], labels_hd[t-lag] = draws[t][ball] - 1
 * (labels_hd may be NULL).  Outputs may be host or device buffers. */
int emrf_build_lag_features(emrf_ctx_t* ctx, const int32_t* draws, int32_t n_draws, int32_t lag, int32_t ball,
                            double* features_hd, double* labels_hd);

/* Synthetic flat-mode inputs of SURVEY.md Appendix C, generated in place on the device (the reference's own inputs come
 * from a web scrape, Main.java:37-67, which has no network here): rows [row_start, row_start + n_rows) of
 * x[i][j] = 1 + splitmix64(seed, i*F + j) mod 50 and label[i] = (x[i][0] + x[i][3] + splitmix64(seed ^ 0xA5A5A5A5, i) mod 5)
 * mod n_classes, bit-identical to euromillioner_b200/synth.py flat().  labels_dev may be NULL.  Used by bench.py for the
 * 10M x 128 shape of BASELINE configs[2], whose 10 GB should not cross PCIe. */
int emrf_synth_flat(emrf_ctx_t* ctx, int64_t n_rows, int32_t n_features, int32_t n_classes, int64_t seed,
                    int64_t row_start, double* features_dev, double* labels_dev);

/* Partition update for callers that cannot allocate device memory (a process whose GPU memory is
 * wholly owned by a pool allocator, or a loop that must avoid cudaMalloc's device-wide sync).
 * No counterpart in the reference or in Spark: same computation as emrf_update_partition, but
 * bins_cm_dev (column-major uint8, `col_stride` bytes between feature columns), entries_dev,
 * entries_out_dev (distinct from entries_dev) and n_left_dev[n_nodes] are DEVICE buffers owned by
 * the caller, and all working storage is a fixed 2 293 760-byte device array inside the library
 * (held by every GPU that loads it, until exit).  Limits per call: 4096 nodes, 262 144 tiles of
 * 2048 entries; larger requests fail with EMRF_ERR_LIMIT.  The first call on a device looks up
 * the array's address (do it before any stream capture); concurrent calls on one GPU race.
 * Refused with EMRF_ERR_ARG before anything is enqueued: a host pointer for any of the four buffers, an output that is
 * or overlaps the input (the partition is not done in place), a negative node_seg_off[0].  The call returns after its
 * GPU work has completed, and no C++ exception leaves it. */
int emrf_update_partition_noalloc(emrf_ctx_t* ctx, const uint8_t* bins_cm_dev, int64_t col_stride,
                                  int32_t n_features, const uint64_t* entries_dev,
                                  const int64_t* node_seg_off, int32_t n_nodes,
                                  const int32_t* split_feature, const int32_t* split_bin,
                                  uint64_t* entries_out_dev, uint64_t* n_left_dev);

/* ---- host-logic probes (no GPU needed): the sequential, parity-critical pieces of
 * RandomForest.selectNodesToSplit that run on the host inside emrf_fit ------------------------- */
int64_t emrf_host_java_random_next_long(int64_t seed, int32_t nth);           /* java.util.Random */
double emrf_host_xorshift_next_double(int64_t seed, int32_t nth);             /* XORShiftRandom (2.4 hashSeed) */
int32_t emrf_host_reservoir_sample(int32_t n_items, int32_t k, int64_t seed, int32_t* out); /* SamplingUtils */
int32_t emrf_host_group_tree_order(const int32_t* trees_first_seen, int32_t n, int32_t* out); /* Scala map order */
int32_t emrf_host_num_features_per_node(const emrf_params_t* params, int32_t n_features);
/* rows sharded: half-slot position and owner rank of every node of a group in the reduce-scattered histogram buffer
 * (nodes of weight <= 65535 take one half slot and travel as packed 16-bit counters); returns half slots per rank */
/* rows sharded, groups tiled on the device: the number of tiles a group's launch is provisioned with, from the nodes'
 * entry bounds, their trees and the trees' local in-bag totals (never below the exact count) */
int64_t emrf_host_group_tiles_bound(const uint32_t* ub, const int32_t* tree, int32_t n, const uint64_t* tree_total,
                                    int32_t n_trees, uint32_t tile_len);
int32_t emrf_host_packed_slot_layout(const uint32_t* weight, int32_t n, int32_t world, int32_t* slot_out, int32_t* owner_out);

#ifdef __cplusplus
}
#endif
#endif /* EMRF_H_ */
