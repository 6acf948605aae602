package com.alibaba.alink.operator.common.clustering.kmeans.gpu;

/**
 * Sketch of the prediction side.  KMeansModelMapper.map is row-at-a-time (ModelMapBatchOp.java:43-46); a GPU
 * mapper has to batch, so KMeansPredictBatchOp.linkFrom swaps `.map(new ModelMapperAdapter(..))` for
 * `.mapPartition(..)` that collects B rows, calls AkmNative.predict once, and emits the B output rows in
 * order through the same OutputColsHelper (schema unchanged: pred LONG [, detail STRING][, distance DOUBLE],
 * KMeansModelMapper.java:40-53).  idx = -1 / null vectors are handled on the Java side exactly as in
 * KMeansModelMapper.java:66-73.
 */
public final class GpuKMeansModelMapper {
    private GpuKMeansModelMapper() {}
}
