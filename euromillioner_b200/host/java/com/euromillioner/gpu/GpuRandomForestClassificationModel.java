package com.euromillioner.gpu;

import java.nio.ByteBuffer;
import java.nio.ByteOrder;
import java.util.ArrayDeque;
import java.util.UUID;

/** Mirror of RandomForestClassificationModel: numTrees/numClasses/numFeatures/totalNumNodes, featureImportances,
 *  transform (rawPrediction, probability, prediction), toDebugString, and the flat arrays a host persists
 *  (exportArrays / fromArrays — what Spark's NodeData rows are built from).  SOURCE ONLY. */
public final class GpuRandomForestClassificationModel implements AutoCloseable {
    /** The model as flat arrays: trees in index order, nodes pre-order per tree (emrf_forest_export). */
    public static final class Arrays {
        public long[] treeOff;
        public int[] nodeId, left, right, feature;
        public double[] threshold, gain, impurity, prediction, stats;   // stats: totalNumNodes x numClasses class counts
    }

    private final GpuContext ctx;
    private long forest;
    public final int numTrees, numClasses, numFeatures;
    public final long totalNumNodes;
    /** Identifiable.randomUID("rfc"): "rfc_" + 12 hex digits. */
    public final String uid = "rfc_" + UUID.randomUUID().toString().replace("-", "").substring(0, 12);
    private double[] thresholds;   // ProbabilisticClassificationModel.thresholds (null = unset)

    GpuRandomForestClassificationModel(GpuContext ctx, long forest) {
        this.ctx = ctx; this.forest = forest;
        long[] i = Native.forestInfo(forest);
        numTrees = (int) i[0]; numClasses = (int) i[1]; numFeatures = (int) i[2]; totalNumNodes = i[3];
    }

    public double[] featureImportances() { return Native.featureImportances(forest); }
    public GpuRandomForestClassificationModel setThresholds(double[] t) {
        if (t != null && t.length != numClasses)
            throw new IllegalArgumentException("thresholds has length " + t.length + ", but the model has " + numClasses + " classes");
        thresholds = t == null ? null : t.clone();
        return this;
    }

    public Arrays exportArrays() {
        final int n = Math.toIntExact(totalNumNodes);
        Arrays a = new Arrays();
        a.treeOff = new long[numTrees + 1];
        a.nodeId = new int[n]; a.left = new int[n]; a.right = new int[n]; a.feature = new int[n];
        a.threshold = new double[n]; a.gain = new double[n]; a.impurity = new double[n]; a.prediction = new double[n];
        a.stats = new double[Math.multiplyExact(n, numClasses)];
        Native.forestExport(forest, a.treeOff, a.nodeId, a.left, a.right, a.feature, a.threshold, a.gain, a.impurity, a.prediction, a.stats);
        return a;
    }
    /** A model read back from storage (validated natively: tree offsets, child indices, one parent per node). */
    public static GpuRandomForestClassificationModel fromArrays(int device, int numClasses, int numFeatures, Arrays a) {
        GpuContext ctx = GpuContext.forDevice(device);
        try {
            long f = Native.forestImport(ctx.handle, a.treeOff.length - 1, numClasses, numFeatures, a.treeOff, a.nodeId, a.left, a.right,
                                         a.feature, a.threshold, a.gain, a.impurity, a.prediction, a.stats);
            return new GpuRandomForestClassificationModel(ctx, f);
        } catch (RuntimeException e) { ctx.close(); throw e; }
    }

    /** RandomForestClassificationModel.toString */
    @Override public String toString() {
        return "RandomForestClassificationModel (uid=" + uid + ") with " + numTrees + " trees";
    }
    /** TreeEnsembleModel.toDebugString of Spark 2.4: the header, then per tree `  Tree i (weight 1.0):` and
     *  Node.subtreeToString(indentFactor = 4).  Doubles print through Double.toString, as Scala's string interpolation does. */
    public String toDebugString() {
        Arrays a = exportArrays();
        StringBuilder sb = new StringBuilder(toString()).append('\n');
        ArrayDeque<long[]> stack = new ArrayDeque<>();     // {node, indent, 1 if an "Else" line precedes the node, parent node}
        for (int t = 0; t < numTrees; t++) {
            sb.append("  Tree ").append(t).append(" (weight 1.0):\n");
            final int base = (int) a.treeOff[t];
            stack.push(new long[] {0, 4, 0, 0});
            while (!stack.isEmpty()) {
                long[] e = stack.pop();
                final int i = base + (int) e[0], ind = (int) e[1];
                if (e[2] != 0) {
                    final int p = base + (int) e[3];
                    pad(sb, ind - 1).append("Else (feature ").append(a.feature[p]).append(" > ").append(a.threshold[p]).append(")\n");
                }
                if (a.left[i] < 0) { pad(sb, ind).append("Predict: ").append(a.prediction[i]).append('\n'); continue; }
                pad(sb, ind).append("If (feature ").append(a.feature[i]).append(" <= ").append(a.threshold[i]).append(")\n");
                stack.push(new long[] {a.right[i], ind + 1, 1, e[0]});
                stack.push(new long[] {a.left[i], ind + 1, 0, e[0]});
            }
        }
        return sb.toString();
    }
    private static StringBuilder pad(StringBuilder sb, int n) { for (int k = 0; k < n; k++) sb.append(' '); return sb; }

    /** Returns {rawPrediction (N*C), probability (N*C), prediction (N)}.  Rows go to the engine in slices whose buffers stay
     *  below 2 GiB (a direct ByteBuffer's limit). */
    public double[][] transform(double[] features) {
        final long n = features.length / numFeatures;
        final double[][] out = {new double[Math.toIntExact(n * numClasses)], new double[Math.toIntExact(n * numClasses)], new double[(int) n]};
        final long rowBytes = 8L * Math.max(numFeatures, numClasses);
        final int slice = (int) Math.max(1L, Math.min(n, (1L << 30) / rowBytes));
        ByteBuffer fb = ByteBuffer.allocateDirect(slice * numFeatures * 8).order(ByteOrder.nativeOrder());
        ByteBuffer raw = ByteBuffer.allocateDirect(slice * numClasses * 8).order(ByteOrder.nativeOrder());
        ByteBuffer prob = ByteBuffer.allocateDirect(slice * numClasses * 8).order(ByteOrder.nativeOrder());
        ByteBuffer pred = ByteBuffer.allocateDirect(slice * 8).order(ByteOrder.nativeOrder());
        for (long r0 = 0; r0 < n; r0 += slice) {
            final int rows = (int) Math.min(slice, n - r0);
            fb.clear();
            fb.asDoubleBuffer().put(features, Math.toIntExact(r0 * numFeatures), rows * numFeatures);
            if (thresholds == null) Native.predict(ctx.handle, forest, fb, rows, raw, prob, pred);
            else Native.predictThresholds(ctx.handle, forest, fb, rows, thresholds, raw, prob, pred);
            raw.clear(); prob.clear(); pred.clear();
            raw.asDoubleBuffer().get(out[0], Math.toIntExact(r0 * numClasses), rows * numClasses);
            prob.asDoubleBuffer().get(out[1], Math.toIntExact(r0 * numClasses), rows * numClasses);
            pred.asDoubleBuffer().get(out[2], (int) r0, rows);
        }
        return out;
    }
    @Override public void close() { if (forest != 0) { Native.forestFree(forest); forest = 0; ctx.close(); } }
}
