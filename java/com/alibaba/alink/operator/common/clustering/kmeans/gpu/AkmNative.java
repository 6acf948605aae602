package com.alibaba.alink.operator.common.clustering.kmeans.gpu;

/**
 * JNI view of include/akm.h (libakm.so).  Not compiled in this repository (the build image has no JDK);
 * the C side of these natives is java/akm_jni.c.  Every method returns akm's int status; non-zero is turned
 * into a RuntimeException by {@link #check}, which fails the Flink job exactly like the unchecked exceptions
 * of the stock ComputeFunctions do.
 */
public final class AkmNative {
    static {
        System.loadLibrary("akm_jni");   // links against libakm.so
    }

    public static final int EUCLIDEAN = 0;
    public static final int COSINE = 1;

    private AkmNative() {}

    public static native int create(int device, int vectorSize, int metric, long[] handleOut);
    public static native int destroy(long handle);
    public static native String lastError(long handle);
    public static native int uploadRows(long handle, double[] rowsRowMajor, long n);
    /** rows that already went through FastDistance.prepareVectorData (the ComQueue's "trainData") */
    public static native int uploadRowsPrepared(long handle, double[] rowsRowMajor, long n);
    /** SparseVector rows as CSR: indptr[n + 1], the concatenated SparseVector.getIndices() / getValues() */
    public static native int uploadRowsCsr(long handle, long[] indptr, int[] indices, double[] values, long n);
    public static native int uploadWait(long handle);
    /** assign + all-reduce (NCCL) + update + criterion in one native call; out arrays may be null except cOut, kOut */
    public static native int superstep(long handle, double[] centroids, int k0, int k, int stepNo, double tol,
                                       double[] cOut, int[] kOut, int[] convergedOut, double[] sumOut,
                                       double[] shiftOut, double[] objectiveOut);
    public static native int assignAccumulate(long handle, double[] centroids, int k0, int k,
                                              double[] sumOut, double[] objectiveOut);
    public static native int updateCentroids(long handle, double[] sum, int k0, int k, int stepNo,
                                             double[] prev, double[] cOut, int[] kOut, double[] shiftOut);
    public static native int commUniqueId(byte[] idOut128);
    public static native int commInit(long handle, byte[] id128, int rank, int world);
    public static native int allReduceSum(long handle, double[] buf, long count);
    public static native int predict(long handle, double[] rows, long n, double[] centroids, int k,
                                     long[] idxOut, double[] distOut, double[] probsOut);

    // device-resident superstep loop (the whole queue of KMeansTrainBatchOp.iterateICQ on the GPU)
    public static native int setCentroids(long handle, double[] c0, int k0);
    public static native int step(long handle, double tol, int[] convergedOut, int[] kOut, double[] objectiveOut);
    public static native int getModel(long handle, double[] cOut, int[] kOut, double[] weightsOut, int[] stepsOut);

    // KMeansInitCentroids.kMeansPlusPlusInit passes (include/akm.h, "Initial centroids on the device")
    public static native int initReset(long handle);
    public static native int initUpdate(long handle, double[] centers, int m, int idBase, double[] sumCostOut);
    public static native int initPick(long handle, double scale, long seed, int round, long rowBase, int cap,
                                      long[] rowsOut, int[] countOut);
    public static native int initCounts(long handle, int mTotal, double[] countsOut);
    public static native int getRows(long handle, long[] rowsSorted, int m, double[] out);

    // EvalClusterBatchOp: per-cluster statistics + silhouette sum of the cached rows
    public static native int evalClusters(long handle, int[] labels, int k, double[] countOut, double[] meanOut,
                                          double[] distSumOut, double[] distSqSumOut, double[] normSqSumOut,
                                          double[] silhouetteSumOut);

    public static void check(long handle, int rc) {
        if (rc != 0) {
            throw new RuntimeException("akm error " + rc + ": " + lastError(handle));
        }
    }
}
