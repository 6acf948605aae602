package com.alibaba.alink.operator.common.clustering.kmeans.gpu;

import com.alibaba.alink.common.linalg.DenseVector;
import com.alibaba.alink.common.linalg.Vector;
import com.alibaba.alink.common.mapper.ModelMapper;
import com.alibaba.alink.common.model.ModelConverterUtils;
import com.alibaba.alink.common.utils.OutputColsHelper;
import com.alibaba.alink.operator.common.clustering.kmeans.KMeansTrainModelData;
import com.alibaba.alink.operator.common.clustering.kmeans.KMeansUtil;
import com.alibaba.alink.params.clustering.KMeansPredictParams;
import com.alibaba.alink.params.shared.clustering.HasKMeansWithHaversineDistanceType.DistanceType;
import org.apache.flink.api.common.typeinfo.TypeInformation;
import org.apache.flink.api.common.typeinfo.Types;
import org.apache.flink.api.java.tuple.Tuple2;
import org.apache.flink.ml.api.misc.param.Params;
import org.apache.flink.table.api.TableSchema;
import org.apache.flink.types.Row;

import java.util.ArrayList;
import java.util.List;

/**
 * GPU counterpart of KMeansModelMapper: same constructor, same output schema (prediction LONG [, detail STRING]
 * [, distance DOUBLE] appended through OutputColsHelper), same results -- but the unit of work is a micro-batch.
 * KMeansPredictBatchOp keeps its parameters; ModelMapBatchOp.linkFrom swaps
 * <pre>  .map(new ModelMapperAdapter(mapper, modelSource))</pre> for
 * <pre>  .mapPartition(new GpuModelMapperAdapter(mapper, modelSource, device))</pre>
 * ({@link GpuModelMapperAdapter} collects up to BATCH rows, calls {@link #mapBatch}, emits the results in input order).
 * {@link #map(Row)} still works (a batch of one) so LocalPredictor and the stream operator need no change.
 *
 * The model matrix handed to the library is the RAW centroid matrix of the model rows: akm_predict applies
 * FastDistance.updateLabel itself (KMeansUtil.transformTrainDataToPredictData does the same on the Java side), so
 * COSINE centroids are normalised exactly once.
 */
public class GpuKMeansModelMapper extends ModelMapper {
    /** rows per akm_predict call: 1 MB of doubles at d = 128, large enough to fill the GPU, small enough to pipeline */
    public static final int BATCH = 8192;

    private final OutputColsHelper outputColsHelper;
    private final boolean isPredDetail;
    private final boolean isPredDistance;
    private final int device;

    private transient KMeansTrainModelData model;      // raw cluster vectors, ids, weights
    private transient int[] colIdx;
    private transient double[] centroids;              // k x d row-major == DenseMatrix(d, k) column-major
    private transient int k, d;
    private transient long handle;

    public GpuKMeansModelMapper(TableSchema modelSchema, TableSchema dataSchema, Params params, int device) {
        super(modelSchema, dataSchema, params);
        this.device = device;
        String[] reserved = this.params.get(KMeansPredictParams.RESERVED_COLS);
        isPredDetail = params.contains(KMeansPredictParams.PREDICTION_DETAIL_COL);
        isPredDistance = params.contains(KMeansPredictParams.PREDICTION_DISTANCE_COL);
        List<String> cols = new ArrayList<>();
        List<TypeInformation> types = new ArrayList<>();
        cols.add(this.params.get(KMeansPredictParams.PREDICTION_COL));
        types.add(Types.LONG);
        if (isPredDetail) {
            cols.add(params.get(KMeansPredictParams.PREDICTION_DETAIL_COL));
            types.add(Types.STRING);
        }
        if (isPredDistance) {
            cols.add(params.get(KMeansPredictParams.PREDICTION_DISTANCE_COL));
            types.add(Types.DOUBLE);
        }
        outputColsHelper = new OutputColsHelper(dataSchema, cols.toArray(new String[0]),
            types.toArray(new TypeInformation[0]), reserved);
    }

    @Override
    public TableSchema getOutputSchema() {
        return outputColsHelper.getResultSchema();
    }

    @Override
    public void loadModel(List<Row> modelRows) {
        Tuple2<Params, Iterable<String>> metaAndData = ModelConverterUtils.extractModelMetaAndData(modelRows);
        model = KMeansUtil.loadModelForTrain(metaAndData.f0, metaAndData.f1);
        if (model.params.distanceType == DistanceType.HAVERSINE) {
            throw new IllegalArgumentException("the GPU mapper covers EUCLIDEAN and COSINE; use KMeansModelMapper for HAVERSINE");
        }
        k = model.params.k;
        d = model.params.vectorSize;
        centroids = new double[k * d];
        for (int i = 0; i < k; i++) {
            System.arraycopy(model.getClusterVector(i).getData(), 0, centroids, i * d, d);
        }
        colIdx = KMeansUtil.getKmeansPredictColIdxs(model.params, getDataSchema().getFieldNames());
        long[] h = new long[1];
        int metric = model.params.distanceType == DistanceType.COSINE ? AkmNative.COSINE : AkmNative.EUCLIDEAN;
        AkmNative.check(0, AkmNative.create(device, d, metric, h));
        handle = h[0];
    }

    @Override
    public void close() {
        if (handle != 0) {
            AkmNative.destroy(handle);
            handle = 0;
        }
    }

    @Override
    public Row map(Row row) {
        List<Row> one = new ArrayList<>(1);
        one.add(row);
        return mapBatch(one).get(0);
    }

    /**
     * KMeansModelMapper.map for a list of rows, results in the same order.  Rows without a vector keep null outputs
     * (KMeansModelMapper.java:66-73) and never reach the GPU.
     */
    public List<Row> mapBatch(List<Row> rows) {
        int n = rows.size();
        int[] slot = new int[n];                         // position of row i in the dense batch, -1 = null vector
        int m = 0;
        double[] x = new double[n * d];
        for (int i = 0; i < n; i++) {
            Vector v = KMeansUtil.getKMeansPredictVector(colIdx, rows.get(i));
            if (v == null) {
                slot[i] = -1;
                continue;
            }
            double[] dense = (v instanceof DenseVector) ? ((DenseVector) v).getData() : toDense(v);
            if (dense.length != d) {
                throw new IllegalArgumentException("vector size " + dense.length + " != model vectorSize " + d);
            }
            System.arraycopy(dense, 0, x, m * d, d);
            slot[i] = m++;
        }
        long[] idx = new long[Math.max(m, 1)];
        double[] dist = isPredDistance ? new double[Math.max(m, 1)] : null;
        double[] probs = isPredDetail ? new double[Math.max(m, 1) * k] : null;
        if (m > 0) {
            AkmNative.check(handle, AkmNative.predict(handle, x, m, centroids, k, idx, dist, probs));
        }
        List<Row> out = new ArrayList<>(n);
        for (int i = 0; i < n; i++) {
            List<Object> res = new ArrayList<>(3);
            int s = slot[i];
            if (s < 0) {
                res.add(null);
                if (isPredDetail) {
                    res.add(null);
                }
                if (isPredDistance) {
                    res.add(null);
                }
            } else {
                res.add(idx[s]);                         // (long) index, -1 included
                if (isPredDetail) {
                    DenseVector vec = new DenseVector(k);
                    for (int j = 0; j < k; j++) {         // scatter by clusterId like KMeansModelMapper.java:85-88
                        vec.set((int) model.getClusterId(j), probs[s * k + j]);
                    }
                    res.add(vec.toString());
                }
                if (isPredDistance) {
                    res.add(dist[s]);
                }
            }
            out.add(outputColsHelper.getResultRow(rows.get(i), Row.of(res.toArray(new Object[0]))));
        }
        return out;
    }

    private double[] toDense(Vector v) {
        double[] a = new double[d];
        com.alibaba.alink.common.linalg.SparseVector sv = (com.alibaba.alink.common.linalg.SparseVector) v;
        int[] ind = sv.getIndices();
        double[] val = sv.getValues();
        for (int i = 0; i < ind.length; i++) {
            a[ind[i]] = val[i];
        }
        return a;
    }
}
