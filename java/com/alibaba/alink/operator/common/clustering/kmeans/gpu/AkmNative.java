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
    public static native int assignAccumulate(long handle, double[] centroids, int k0, int k,
                                              double[] sumOut, double[] objectiveOut);
    public static native int updateCentroids(long handle, double[] sum, int k0, int k, int stepNo,
                                             double[] prev, double[] cOut, int[] kOut, double[] shiftOut);
    public static native int commUniqueId(byte[] idOut128);
    public static native int commInit(long handle, byte[] id128, int rank, int world);
    public static native int allReduceSum(long handle, double[] buf, long count);
    public static native int predict(long handle, double[] rows, long n, double[] centroids, int k,
                                     long[] idxOut, double[] distOut, double[] probsOut);

    public static void check(long handle, int rc) {
        if (rc != 0) {
            throw new RuntimeException("akm error " + rc + ": " + lastError(handle));
        }
    }
}
