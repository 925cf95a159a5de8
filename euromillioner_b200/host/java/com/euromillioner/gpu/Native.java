package com.euromillioner.gpu;

import java.nio.ByteBuffer;

/**
 * JNI declarations, one per entry point of include/emrf.h (libemrf.so + libemrf_jni.so).
 * SOURCE ONLY: no JDK exists in the build image, so this class has never been compiled or run there.
 * Bulk data travels either in direct ByteBuffers (native order) or, above 2 GiB, in page-locked native blocks
 * addressed by a long (hostAllocPinned / putDoubles); status codes become exceptions in the shim:
 * EMRF_ERR_ARG -> IllegalArgumentException (Spark's `require`), everything else -> RuntimeException.
 */
final class Native {
    static { System.loadLibrary("emrf_jni"); }
    private Native() {}

    static native long ctxCreate(int device);
    /** One process per GPU: `uid` is the 128-byte id made by rank 0 with ncclUniqueId() and sent to the others. */
    static native long ctxCreateDist(int device, int rank, int world, byte[] uid);
    static native byte[] ncclUniqueId();
    static native void ctxDestroy(long ctx);
    /** Gives the context's device buffers and the library's cached blocks back to the driver; returns bytes freed. */
    static native long ctxReleaseBuffers(long ctx);

    /** Page-locked host memory for inputs larger than a ByteBuffer can hold (and for full-rate PCIe copies). */
    static native long hostAllocPinned(long bytes);
    static native void hostFreePinned(long addr);
    /** Copies src[0, len) to the native address dst + dstOffsetDoubles * 8. */
    static native void putDoubles(long dst, long dstOffsetDoubles, double[] src, int len);

    /** params: {numTrees,maxDepth,maxBins,minInstancesPerNode,numClasses,impurity,subsetKind,maxMemoryInMB,findSplitsPolicy};
     *  dparams: {subsetValue,minInfoGain,subsamplingRate}. features: N*F doubles row-major; labels: N doubles. */
    static native long fit(long ctx, int[] params, double[] dparams, long seed, ByteBuffer features, ByteBuffer labels,
                           long nRows, int nFeatures, long[] partitionOffsets);
    /** The same with the rows at native addresses (hostAllocPinned blocks or device memory). */
    static native long fitAt(long ctx, int[] params, double[] dparams, long seed, long featuresAddr, long labelsAddr,
                             long nRows, int nFeatures, long[] partitionOffsets);
    /** Row-sharded fit (emrf_fit_sharded): this rank passes the rows of global partitions
     *  [firstPartition, firstPartition + nPartitionsLocal); every rank passes the same global offsets. */
    static native long fitShardedAt(long ctx, int[] params, double[] dparams, long seed, long featuresAddr, long labelsAddr,
                                    long nRowsLocal, int nFeatures, long[] partitionOffsetsGlobal, int firstPartition,
                                    int nPartitionsLocal);
    static native void forestFree(long forest);
    static native long[] forestInfo(long forest);              // {numTrees, numClasses, numFeatures, totalNodes}
    static native void forestExport(long forest, long[] treeOff, int[] nodeId, int[] left, int[] right, int[] feature,
                                    double[] threshold, double[] gain, double[] impurity, double[] prediction, double[] stats);
    /** emrf_forest_import: a model from its flat arrays (a saved model read back by the host). */
    static native long forestImport(long ctx, int numTrees, int numClasses, int numFeatures, long[] treeOff, int[] nodeId,
                                    int[] left, int[] right, int[] feature, double[] threshold, double[] gain,
                                    double[] impurity, double[] prediction, double[] stats);
    static native double[] featureImportances(long forest);
    /** raw/prob: N*C doubles or null; pred: N doubles. Safe to call from several executor task threads. */
    static native void predict(long ctx, long forest, ByteBuffer features, long nRows, ByteBuffer raw, ByteBuffer prob,
                               ByteBuffer pred);
    /** Same with ProbabilisticClassifierParams.thresholds (numClasses values, all >= 0, at most one 0). */
    static native void predictThresholds(long ctx, long forest, ByteBuffer features, long nRows, double[] thresholds,
                                         ByteBuffer raw, ByteBuffer prob, ByteBuffer pred);
}
