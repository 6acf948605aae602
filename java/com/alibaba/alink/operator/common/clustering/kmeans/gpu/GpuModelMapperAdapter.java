package com.alibaba.alink.operator.common.clustering.kmeans.gpu;

import com.alibaba.alink.common.model.ModelSource;
import org.apache.flink.api.common.functions.RichMapPartitionFunction;
import org.apache.flink.configuration.Configuration;
import org.apache.flink.types.Row;
import org.apache.flink.util.Collector;

import java.util.ArrayList;
import java.util.List;

/**
 * mapPartition counterpart of ModelMapperAdapter (common/mapper/ModelMapperAdapter.java:36-49): loads the model rows
 * from the broadcast set at open(), then feeds the partition to the mapper in micro-batches so that one native
 * call covers thousands of rows.  Output order within the partition equals input order.
 */
public class GpuModelMapperAdapter extends RichMapPartitionFunction<Row, Row> {
    private final GpuKMeansModelMapper mapper;
    private final ModelSource modelSource;

    public GpuModelMapperAdapter(GpuKMeansModelMapper mapper, ModelSource modelSource) {
        this.mapper = mapper;
        this.modelSource = modelSource;
    }

    @Override
    public void open(Configuration parameters) throws Exception {
        mapper.loadModel(modelSource.getModelRows(getRuntimeContext()));
    }

    @Override
    public void mapPartition(Iterable<Row> values, Collector<Row> out) {
        List<Row> batch = new ArrayList<>(GpuKMeansModelMapper.BATCH);
        for (Row row : values) {
            batch.add(row);
            if (batch.size() == GpuKMeansModelMapper.BATCH) {
                flush(batch, out);
            }
        }
        flush(batch, out);
    }

    private void flush(List<Row> batch, Collector<Row> out) {
        if (batch.isEmpty()) {
            return;
        }
        for (Row r : mapper.mapBatch(batch)) {
            out.collect(r);
        }
        batch.clear();
    }

    @Override
    public void close() {
        mapper.close();
    }
}
