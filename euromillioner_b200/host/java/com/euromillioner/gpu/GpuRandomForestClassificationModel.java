package com.euromillioner.gpu;

import java.nio.ByteBuffer;
import java.nio.ByteOrder;

/** Mirror of RandomForestClassificationModel: numTrees/numClasses/numFeatures/totalNumNodes,
 *  featureImportances, transform (rawPrediction, probability, prediction).  SOURCE ONLY. */
public final class GpuRandomForestClassificationModel implements AutoCloseable {
    private final long ctx;
    private long forest;
    public final int numTrees, numClasses, numFeatures;
    public final long totalNumNodes;

    GpuRandomForestClassificationModel(long ctx, long forest) {
        this.ctx = ctx; this.forest = forest;
        long[] i = Native.forestInfo(forest);
        numTrees = (int) i[0]; numClasses = (int) i[1]; numFeatures = (int) i[2]; totalNumNodes = i[3];
    }
    public double[] featureImportances() { return Native.featureImportances(forest); }
    private double[] thresholds;   // ProbabilisticClassificationModel.thresholds (null = unset)
    public GpuRandomForestClassificationModel setThresholds(double[] t) {
        if (t != null && t.length != numClasses)
            throw new IllegalArgumentException("thresholds has length " + t.length + ", but the model has " + numClasses + " classes");
        thresholds = t == null ? null : t.clone();
        return this;
    }

    /** Returns {rawPrediction (N*C), probability (N*C), prediction (N)}. */
    public double[][] transform(double[] features) {
        long n = features.length / numFeatures;
        ByteBuffer fb = ByteBuffer.allocateDirect(features.length * 8).order(ByteOrder.nativeOrder());
        fb.asDoubleBuffer().put(features);
        ByteBuffer raw = ByteBuffer.allocateDirect((int) (n * numClasses * 8)).order(ByteOrder.nativeOrder());
        ByteBuffer prob = ByteBuffer.allocateDirect((int) (n * numClasses * 8)).order(ByteOrder.nativeOrder());
        ByteBuffer pred = ByteBuffer.allocateDirect((int) (n * 8)).order(ByteOrder.nativeOrder());
        if (thresholds == null) Native.predict(ctx, forest, fb, n, raw, prob, pred);
        else Native.predictThresholds(ctx, forest, fb, n, thresholds, raw, prob, pred);
        double[][] out = {new double[(int) (n * numClasses)], new double[(int) (n * numClasses)], new double[(int) n]};
        raw.asDoubleBuffer().get(out[0]); prob.asDoubleBuffer().get(out[1]); pred.asDoubleBuffer().get(out[2]);
        return out;
    }
    @Override public void close() { if (forest != 0) { Native.forestFree(forest); forest = 0; Native.ctxDestroy(ctx); } }
}
