package com.euromillioner.gpu;

import java.nio.ByteBuffer;

/**
 * JNI declarations, one per entry point of include/emrf.h (libemrf.so + libemrf_jni.so).
 * SOURCE ONLY: no JDK exists in the build image, so this class has never been compiled or run there.
 * Bulk data travels in direct ByteBuffers (native order); status codes become exceptions in the shim:
 * EMRF_ERR_ARG -> IllegalArgumentException (Spark's `require`), everything else -> RuntimeException.
 */
final class Native {
    static { System.loadLibrary("emrf_jni"); }
    private Native() {}

    static native long ctxCreate(int device);
    static native void ctxDestroy(long ctx);

    /** params: {numTrees,maxDepth,maxBins,minInstancesPerNode,numClasses,impurity,subsetKind,maxMemoryInMB,findSplitsPolicy};
     *  dparams: {subsetValue,minInfoGain,subsamplingRate}. features: N*F doubles row-major; labels: N doubles. */
    static native long fit(long ctx, int[] params, double[] dparams, long seed, ByteBuffer features, ByteBuffer labels,
                           long nRows, int nFeatures, long[] partitionOffsets);
    static native void forestFree(long forest);
    static native long[] forestInfo(long forest);              // {numTrees, numClasses, numFeatures, totalNodes}
    static native void forestExport(long forest, long[] treeOff, int[] nodeId, int[] left, int[] right, int[] feature,
                                    double[] threshold, double[] gain, double[] impurity, double[] prediction, double[] stats);
    static native double[] featureImportances(long forest);
    /** raw/prob: N*C doubles or null; pred: N doubles. Safe to call from several executor task threads. */
    static native void predict(long ctx, long forest, ByteBuffer features, long nRows, ByteBuffer raw, ByteBuffer prob,
                               ByteBuffer pred);
    /** Same with ProbabilisticClassifierParams.thresholds (numClasses values, all >= 0, at most one 0). */
    static native void predictThresholds(long ctx, long forest, ByteBuffer features, long nRows, double[] thresholds,
                                         ByteBuffer raw, ByteBuffer prob, ByteBuffer pred);
}
