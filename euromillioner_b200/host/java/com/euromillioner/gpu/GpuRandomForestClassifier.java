package com.euromillioner.gpu;

import java.nio.ByteBuffer;
import java.nio.ByteOrder;

/**
 * Drop-in for org.apache.spark.ml.classification.RandomForestClassifier on the Euromillions stage
 * (the model block of the reference, Main.java:110-141).  Same setter names and defaults as Spark 2.4.6
 * (ml/tree/treeParams.scala); fit() hands dense rows to the B200 engine through Native.
 * SOURCE ONLY (no JDK in the build image).  To keep this file free of a Spark compile dependency the
 * data arrives as arrays; a Spark adapter would collect `(label, features)` per partition with
 * rdd.mapPartitionsWithIndex and call fitPartitions (one array per partition, any total size).
 */
public final class GpuRandomForestClassifier {
    private int numTrees = 20, maxDepth = 5, maxBins = 32, minInstancesPerNode = 1, maxMemoryInMB = 256, numClasses = 0;
    private double minInfoGain = 0.0, subsamplingRate = 1.0, subsetValue = 0.0;
    private int impurity = 0, subsetKind = 0;
    private long seed = "org.apache.spark.ml.classification.RandomForestClassifier".hashCode(); // 207336481
    private double[] thresholds;   // ProbabilisticClassifierParams.thresholds, handed to the fitted model
    private final int device;
    private final GpuContext distributed;   // non-null: this estimator is one rank of a multi-GPU job

    public GpuRandomForestClassifier() { this(0); }
    public GpuRandomForestClassifier(int device) { this.device = device; this.distributed = null; }
    /** One rank of a multi-GPU job (GpuContext.distributed): use fitSharded. */
    public GpuRandomForestClassifier(GpuContext distributedContext) { this.device = -1; this.distributed = distributedContext; }

    public GpuRandomForestClassifier setNumTrees(int v) { require(v >= 1, "numTrees must be >= 1"); numTrees = v; return this; }
    public GpuRandomForestClassifier setMaxDepth(int v) { require(v >= 0 && v <= 30, "maxDepth must be in [0,30]"); maxDepth = v; return this; }
    public GpuRandomForestClassifier setMaxBins(int v) { require(v >= 2, "maxBins must be >= 2"); maxBins = v; return this; }
    public GpuRandomForestClassifier setMinInstancesPerNode(int v) { require(v >= 1, "minInstancesPerNode must be >= 1"); minInstancesPerNode = v; return this; }
    public GpuRandomForestClassifier setMinInfoGain(double v) { require(v >= 0, "minInfoGain must be >= 0"); minInfoGain = v; return this; }
    public GpuRandomForestClassifier setSubsamplingRate(double v) { require(v > 0 && v <= 1, "subsamplingRate must be in (0,1]"); subsamplingRate = v; return this; }
    public GpuRandomForestClassifier setMaxMemoryInMB(int v) { require(v >= 0, "maxMemoryInMB must be >= 0"); maxMemoryInMB = v; return this; }
    public GpuRandomForestClassifier setSeed(long v) { seed = v; return this; }
    public GpuRandomForestClassifier setNumClasses(int v) { numClasses = v; return this; }
    /** ProbabilisticClassifierParams.setThresholds: all >= 0, at most one 0 (checked natively at transform, as Spark's require). */
    public GpuRandomForestClassifier setThresholds(double[] v) { thresholds = v == null ? null : v.clone(); return this; }
    public GpuRandomForestClassifier setImpurity(String v) {
        String s = v.toLowerCase();
        require(s.equals("gini") || s.equals("entropy"), "impurity must be gini or entropy");
        impurity = s.equals("gini") ? 0 : 1; return this;
    }
    public GpuRandomForestClassifier setFeatureSubsetStrategy(String v) {
        String s = v.toLowerCase();
        switch (s) {
            case "auto": subsetKind = 0; break; case "all": subsetKind = 1; break; case "sqrt": subsetKind = 2; break;
            case "log2": subsetKind = 3; break; case "onethird": subsetKind = 4; break;
            default:
                if (s.matches("\\d+")) { subsetKind = 5; subsetValue = Integer.parseInt(s); }
                else { subsetKind = 6; subsetValue = Double.parseDouble(s); require(subsetValue > 0 && subsetValue <= 1, "bad featureSubsetStrategy"); }
        }
        return this;
    }

    private int[] intParams() { return new int[] {numTrees, maxDepth, maxBins, minInstancesPerNode, numClasses, impurity, subsetKind, maxMemoryInMB, 0}; }
    private double[] doubleParams() { return new double[] {subsetValue, minInfoGain, subsamplingRate}; }
    private GpuRandomForestClassificationModel model(GpuContext ctx, long forest) {
        GpuRandomForestClassificationModel m = new GpuRandomForestClassificationModel(ctx, forest);
        if (thresholds != null) m.setThresholds(thresholds);
        return m;
    }

    /** features: N x F row-major (below 2 GiB; larger inputs: fitPartitions); labels: N; partitionOffsets: P+1 row
     *  offsets of the Dataset's partitions. */
    public GpuRandomForestClassificationModel fit(double[] features, double[] labels, int nFeatures, long[] partitionOffsets) {
        long n = labels.length;
        require((long) features.length == n * nFeatures, "features must hold N x F values");
        require((long) features.length * 8L <= Integer.MAX_VALUE, "more than 2 GiB of features: use fitPartitions");
        ByteBuffer fb = ByteBuffer.allocateDirect((int) ((long) features.length * 8L)).order(ByteOrder.nativeOrder());
        fb.asDoubleBuffer().put(features);
        ByteBuffer lb = ByteBuffer.allocateDirect(labels.length * 8).order(ByteOrder.nativeOrder());
        lb.asDoubleBuffer().put(labels);
        GpuContext ctx = distributed != null ? distributed.retain() : GpuContext.forDevice(device);
        try {
            return model(ctx, Native.fit(ctx.handle, intParams(), doubleParams(), seed, fb, lb, n, nFeatures, partitionOffsets));
        } catch (RuntimeException e) { ctx.close(); throw e; }
    }

    /** One (features, labels) pair per Spark partition, in partition order; any total size.  The rows are gathered into
     *  page-locked native memory (full-rate PCIe copy, no 2 GiB limit). */
    public GpuRandomForestClassificationModel fitPartitions(double[][] features, double[][] labels, int nFeatures) {
        return fitBlocks(features, labels, nFeatures, null, 0);
    }

    /** Row-sharded multi-GPU fit: this rank's partitions are global partitions [firstPartition, firstPartition +
     *  features.length) of a Dataset whose global partition offsets are `partitionOffsetsGlobal` (the same on every rank). */
    public GpuRandomForestClassificationModel fitSharded(double[][] features, double[][] labels, int nFeatures,
                                                        long[] partitionOffsetsGlobal, int firstPartition) {
        require(distributed != null, "fitSharded needs an estimator made from GpuContext.distributed(...)");
        return fitBlocks(features, labels, nFeatures, partitionOffsetsGlobal, firstPartition);
    }

    private GpuRandomForestClassificationModel fitBlocks(double[][] features, double[][] labels, int nFeatures, long[] globalOffsets,
                                                        int firstPartition) {
        final int parts = features.length;
        require(labels.length == parts, "one label array per partition");
        long[] off = new long[parts + 1];
        for (int p = 0; p < parts; p++) {
            require((long) features[p].length == (long) labels[p].length * nFeatures, "partition " + p + ": features must hold rows x F values");
            off[p + 1] = off[p] + labels[p].length;
        }
        final long n = off[parts];
        final long fAddr = Native.hostAllocPinned(Math.max(n * nFeatures * 8L, 8L));
        long lAddr = 0;
        GpuContext ctx = null;
        try {
            lAddr = Native.hostAllocPinned(Math.max(n * 8L, 8L));
            for (int p = 0; p < parts; p++) {
                Native.putDoubles(fAddr, off[p] * nFeatures, features[p], features[p].length);
                Native.putDoubles(lAddr, off[p], labels[p], labels[p].length);
            }
            ctx = distributed != null ? distributed.retain() : GpuContext.forDevice(device);
            final long forest = globalOffsets == null
                ? Native.fitAt(ctx.handle, intParams(), doubleParams(), seed, fAddr, lAddr, n, nFeatures, off)
                : Native.fitShardedAt(ctx.handle, intParams(), doubleParams(), seed, fAddr, lAddr, n, nFeatures, globalOffsets,
                                      firstPartition, parts);
            return model(ctx, forest);
        } catch (RuntimeException e) {
            if (ctx != null) ctx.close();
            throw e;
        } finally {
            Native.hostFreePinned(fAddr);
            if (lAddr != 0) Native.hostFreePinned(lAddr);
        }
    }

    private static void require(boolean ok, String msg) { if (!ok) throw new IllegalArgumentException("requirement failed: " + msg); }
}
