package com.euromillioner.gpu;

import java.nio.ByteBuffer;
import java.nio.ByteOrder;

/**
 * Drop-in for org.apache.spark.ml.classification.RandomForestClassifier on the Euromillions stage
 * (the model block of the reference, Main.java:110-141).  Same setter names and defaults as Spark 2.4.6
 * (ml/tree/treeParams.scala); fit() hands dense rows to the B200 engine through Native.
 * SOURCE ONLY (no JDK in the build image).  To keep this file free of a Spark compile dependency the
 * data arrives as arrays; a Spark adapter would collect `(label, features)` per partition with
 * rdd.mapPartitionsWithIndex and pass the partition sizes as `partitionOffsets`.
 */
public final class GpuRandomForestClassifier {
    private int numTrees = 20, maxDepth = 5, maxBins = 32, minInstancesPerNode = 1, maxMemoryInMB = 256, numClasses = 0;
    private double minInfoGain = 0.0, subsamplingRate = 1.0, subsetValue = 0.0;
    private int impurity = 0, subsetKind = 0;
    private long seed = "org.apache.spark.ml.classification.RandomForestClassifier".hashCode(); // 207336481
    private final int device;

    public GpuRandomForestClassifier() { this(0); }
    public GpuRandomForestClassifier(int device) { this.device = device; }

    public GpuRandomForestClassifier setNumTrees(int v) { require(v >= 1, "numTrees must be >= 1"); numTrees = v; return this; }
    public GpuRandomForestClassifier setMaxDepth(int v) { require(v >= 0 && v <= 30, "maxDepth must be in [0,30]"); maxDepth = v; return this; }
    public GpuRandomForestClassifier setMaxBins(int v) { require(v >= 2, "maxBins must be >= 2"); maxBins = v; return this; }
    public GpuRandomForestClassifier setMinInstancesPerNode(int v) { require(v >= 1, "minInstancesPerNode must be >= 1"); minInstancesPerNode = v; return this; }
    public GpuRandomForestClassifier setMinInfoGain(double v) { require(v >= 0, "minInfoGain must be >= 0"); minInfoGain = v; return this; }
    public GpuRandomForestClassifier setSubsamplingRate(double v) { require(v > 0 && v <= 1, "subsamplingRate must be in (0,1]"); subsamplingRate = v; return this; }
    public GpuRandomForestClassifier setMaxMemoryInMB(int v) { require(v >= 0, "maxMemoryInMB must be >= 0"); maxMemoryInMB = v; return this; }
    public GpuRandomForestClassifier setSeed(long v) { seed = v; return this; }
    public GpuRandomForestClassifier setNumClasses(int v) { numClasses = v; return this; }
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

    /** features: N x F row-major; labels: N; partitionOffsets: P+1 row offsets of the Dataset's partitions. */
    public GpuRandomForestClassificationModel fit(double[] features, double[] labels, int nFeatures, long[] partitionOffsets) {
        long n = labels.length;
        require(features.length == n * nFeatures, "features must hold N x F values");
        ByteBuffer fb = ByteBuffer.allocateDirect(features.length * 8).order(ByteOrder.nativeOrder());
        fb.asDoubleBuffer().put(features);
        ByteBuffer lb = ByteBuffer.allocateDirect(labels.length * 8).order(ByteOrder.nativeOrder());
        lb.asDoubleBuffer().put(labels);
        long ctx = Native.ctxCreate(device);
        int[] p = {numTrees, maxDepth, maxBins, minInstancesPerNode, numClasses, impurity, subsetKind, maxMemoryInMB, 0};
        double[] d = {subsetValue, minInfoGain, subsamplingRate};
        long forest = Native.fit(ctx, p, d, seed, fb, lb, n, nFeatures, partitionOffsets);
        return new GpuRandomForestClassificationModel(ctx, forest);
    }

    private static void require(boolean ok, String msg) { if (!ok) throw new IllegalArgumentException("requirement failed: " + msg); }
}
