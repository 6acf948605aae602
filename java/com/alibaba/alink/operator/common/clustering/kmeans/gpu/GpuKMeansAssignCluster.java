package com.alibaba.alink.operator.common.clustering.kmeans.gpu;

import com.alibaba.alink.common.comqueue.ComContext;
import com.alibaba.alink.common.comqueue.ComputeFunction;
import com.alibaba.alink.operator.batch.clustering.KMeansTrainBatchOp;
import com.alibaba.alink.operator.common.distance.FastDistanceMatrixData;
import com.alibaba.alink.operator.common.distance.FastDistanceVectorData;
import org.apache.flink.api.java.tuple.Tuple2;

/**
 * Drop-in for KMeansAssignCluster (same ComContext keys, same outputs).  Superstep 1 uploads the cached
 * partition once (the GPU analogue of the JVM-heap cache of BaseComQueue.java:497-527); every superstep then
 * costs one k0 x d upload and one k0 x (d+1) download.
 *
 * The cached samples were already label-updated by KMeansTrainBatchOp (prepareVectorData, :95-101): COSINE vectors
 * arrive normalised.  They are uploaded with akm_upload_rows_prepared, which stores them as given, so the
 * assignment is computed from exactly the doubles the stock KMeansAssignCluster would read.  The upload call
 * returns as soon as a chunk is staged, so filling the next chunk overlaps the transfer of the previous one.
 */
public class GpuKMeansAssignCluster extends ComputeFunction {
    static final String HANDLE = "akmHandle";
    private final int metric;
    private final int device;

    public GpuKMeansAssignCluster(int metric, int device) {
        this.metric = metric;
        this.device = device;
    }

    @Override
    public void calc(ComContext context) {
        Integer vectorSize = context.getObj(KMeansTrainBatchOp.VECTOR_SIZE);
        Integer k = context.getObj(KMeansTrainBatchOp.K);
        Tuple2<Integer, FastDistanceMatrixData> centroids = context.getStepNo() % 2 == 0
            ? context.getObj(KMeansTrainBatchOp.CENTROID1) : context.getObj(KMeansTrainBatchOp.CENTROID2);
        int k0 = centroids.f1.getVectors().numCols();

        double[] sumMatrixData = context.getObj(KMeansTrainBatchOp.CENTROID_ALL_REDUCE);
        if (sumMatrixData == null) {
            sumMatrixData = new double[k0 * (vectorSize + 1)];
            context.putObj(KMeansTrainBatchOp.CENTROID_ALL_REDUCE, sumMatrixData);
        }
        Iterable<FastDistanceVectorData> trainData = context.getObj(KMeansTrainBatchOp.TRAIN_DATA);
        if (trainData == null) {
            return;
        }
        Long handle = context.getObj(HANDLE);
        if (handle == null) {
            long[] h = new long[1];
            AkmNative.check(0, AkmNative.create(device + context.getTaskId(), vectorSize, metric, h));
            handle = h[0];
            context.putObj(HANDLE, handle);
            final int chunk = Math.max(1, (1 << 23) / vectorSize);      // 64 MB staging buffers
            double[] buf = new double[chunk * vectorSize];
            int filled = 0;
            for (FastDistanceVectorData sample : trainData) {
                com.alibaba.alink.common.linalg.Vector vec = sample.getVector();
                if (vec instanceof com.alibaba.alink.common.linalg.DenseVector) {
                    System.arraycopy(((com.alibaba.alink.common.linalg.DenseVector) vec).getData(), 0, buf,
                        filled * vectorSize, vectorSize);
                } else {
                    // a prepared SparseVector (already normalised under COSINE) is expanded into the 64 MB chunk only;
                    // raw sparse tables can go through AkmNative.uploadRowsCsr instead (INTEGRATION.md)
                    com.alibaba.alink.common.linalg.SparseVector sv = (com.alibaba.alink.common.linalg.SparseVector) vec;
                    java.util.Arrays.fill(buf, filled * vectorSize, (filled + 1) * vectorSize, 0.0);
                    int[] idx = sv.getIndices();
                    double[] val = sv.getValues();
                    for (int i = 0; i < idx.length; i++) {
                        buf[filled * vectorSize + idx[i]] = val[i];
                    }
                }
                if (++filled == chunk) {
                    AkmNative.check(handle, AkmNative.uploadRowsPrepared(handle, buf, filled));
                    filled = 0;
                }
            }
            if (filled > 0) {
                AkmNative.check(handle, AkmNative.uploadRowsPrepared(handle, buf, filled));
            }
            AkmNative.check(handle, AkmNative.uploadWait(handle));
        }
        // DenseMatrix is column-major d x k0 == row-major k0 x d: pass the backing array as is
        AkmNative.check(handle, AkmNative.assignAccumulate(handle, centroids.f1.getVectors().getData(), k0, k,
            sumMatrixData, null));
    }
}
