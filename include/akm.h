/*
 * akm.h -- C ABI of the B200-native KMeans hot path for Alink ("akm" = Alink KMeans).
 *
 * One shared library (alink_b200/csrc/libakm.so), plain C types only, no C++
 * exceptions or abort() across the boundary.  Every function returns an int:
 * AKM_OK (0) or a negative AKM_ERR_* class; akm_last_error() gives the message.
 * A Java shim maps non-zero to RuntimeException (the reference fails the Flink
 * job through unchecked exceptions, e.g. KMeansUtil.java:68-71,
 * KMeansTrainBatchOp.java:106).
 *
 * Reference paths:  CORE = core/src/main/java/com/alibaba/alink
 *
 * Layout conventions (same bytes Alink already holds):
 *   - a block of n samples is n x d row-major doubles == the d x n column-major
 *     DenseMatrix Alink uses (CORE/common/linalg/DenseMatrix.java:26-33,202-204);
 *   - centroids are k0 x d row-major == FastDistanceMatrixData.vectors (d x k0
 *     column-major, CORE/operator/common/distance/FastDistanceMatrixData.java:18-26);
 *   - the accumulator is double[k0*(d+1)], row j = [sum of samples (d), count]
 *     == "centroidAllReduce" (CORE/operator/batch/clustering/KMeansTrainBatchOp.java:34-41).
 *
 * Threading: a handle belongs to one thread at a time (one Flink
 * subtask <-> one handle <-> one GPU + one CUDA stream); distinct handles may be
 * used concurrently.  The caller owns every host pointer; the library copies in
 * and out during the call and never retains host pointers.
 */
#ifndef AKM_H_
#define AKM_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct akm_handle_s *akm_handle;

/* HasKMeansDistanceType (CORE/params/shared/clustering/HasKMeansDistanceType.java:37-45) */
#define AKM_EUCLIDEAN 0
#define AKM_COSINE 1

#define AKM_OK 0
#define AKM_ERR_BADARG (-1)      /* IllegalArgumentException / NullPointerException analogue */
#define AKM_ERR_CUDA (-2)
#define AKM_ERR_NCCL (-3)
#define AKM_ERR_OOM (-4)
#define AKM_ERR_NOFINITE (-5)    /* some row has no distance < Double.MAX_VALUE: Java's
                                    getMinPointIndex returns -1 and updateSumMatrix then throws
                                    ArrayIndexOutOfBounds (KMeansUtil.java:162-173,75) */
#define AKM_ERR_STATE (-6)       /* IllegalStateException analogue (call order) */
#define AKM_ERR_UNSUPPORTED (-7)

/* assignment engines (akm_set_engine) */
#define AKM_ENGINE_AUTO 0        /* tensor-core search + exact certification when the shape allows */
#define AKM_ENGINE_EXACT 1       /* every distance evaluated in fp64 in the Java operation order */
#define AKM_ENGINE_TENSOR 2      /* force the tcgen05 path (error if the shape is unsupported) */

/* ------------------------------------------------------------------ lifecycle */

/* One handle per (sessionId, taskId): replaces the per-task objects the ComQueue keeps in
 * SessionSharedObjs (CORE/common/comqueue/SessionSharedObjs.java:50-73) for this path.
 * `d` = vectorSize (KMeansTrainBatchOp.VECTOR_SIZE), metric = AKM_EUCLIDEAN | AKM_COSINE. */
int akm_create(int device, int d, int metric, akm_handle *out);

/* Frees every device buffer; called from the ComQueue's clearObjs stage
 * (CORE/common/comqueue/BaseComQueue.java:420-442). */
int akm_destroy(akm_handle h);

/* Message of the last error on this handle (or of the calling thread when h == NULL). */
const char *akm_last_error(akm_handle h);

/* Version / build info string. */
const char *akm_version(void);

/* Use an existing CUDA stream (e.g. torch's current stream) for all work of this handle.
 * stream == NULL restores the handle's own stream.  No reference counterpart. */
int akm_set_stream(akm_handle h, void *cuda_stream);

/* Select the assignment engine (default AKM_ENGINE_AUTO).  Results are identical by
 * construction; EXACT exists for verification.  No reference counterpart. */
int akm_set_engine(akm_handle h, int engine);

/* ------------------------------------------------------------------ training data (superstep 1) */

/* Optional: pre-size device storage for n_total rows (avoids regrowth while appending). */
int akm_reserve_rows(akm_handle h, int64_t n_total);

/* Appends n samples (host, n x d doubles) to the partition cached on the GPU and performs
 * FastDistance.prepareVectorData on them (CORE/operator/common/distance/FastDistance.java:136-140,
 * called from KMeansTrainBatchOp.java:95-101): EUCLIDEAN label = sum x_i^2 in index order
 * (EuclideanDistance.java:59-65), COSINE x <- x * (1/||x||) when ||x|| > 0 (CosineDistance.java:60-65).
 * Replaces the List<FastDistanceVectorData> "trainData" cached by the ComQueue
 * (BaseComQueue.java:497-527).  Rows whose values are all exactly representable in fp32 are kept
 * as fp32 (4 B/value); the first value that is not switches the partition to fp64 losslessly.  (A COSINE
 * fp32 partition keeps the raw values and forms x * (1/||x||) wherever it reads them, which is the same
 * double Java stores.)  The host rows are staged through a pinned double buffer by a few copy threads, which
 * also narrow fp32-representable chunks to fp32 so that they cross PCIe at 4 bytes per value. */
int akm_upload_rows(akm_handle h, const double *x_host, int64_t n);

/* Same for fp32 host rows (values widened exactly). */
int akm_upload_rows_f32(akm_handle h, const float *x_host, int64_t n);

/* Appends n samples that ALREADY went through FastDistance.prepareVectorData on the Java side -- what the ComQueue
 * hands a ComputeFunction as "trainData" (KMeansTrainBatchOp.java:95-101 ran CosineDistance.updateLabel :58-65 on
 * every vector before the queue starts).  COSINE rows are stored exactly as given (no second normalisation, so the
 * assignment is bit-identical to the one Java computes from the same FastDistanceVectorData); the partition is
 * then held in fp64, since unit-length doubles are not fp32 values.  EUCLIDEAN: identical to akm_upload_rows
 * (prepareVectorData leaves the vector untouched and the label is recomputed here).  Cluster evaluation under
 * COSINE (akm_eval_clusters) wants the raw vectors: use akm_upload_rows for a partition that is evaluated. */
int akm_upload_rows_prepared(akm_handle h, const double *x_host, int64_t n);

/* Appends n sparse samples in CSR form (SparseVector training data: CORE/common/linalg/BLAS.java:217-240 and
 * KMeansUtil.java:79-82 walk the non-zeros; VectorUtil.java:52-114 parses "$size$i:v ..." into them).  The rows are
 * expanded to dense vectors on the device -- the terms a dense evaluation adds for the zeros are + c_i * 0.0 and
 * + 0.0, which change no finite sum, so every result equals the sparse evaluation bit for bit -- and then go
 * through prepareVectorData like any other upload.  Only the non-zeros cross PCIe.
 *   indptr   int64[n + 1], indptr[0] = 0, non-decreasing
 *   indices  int32[indptr[n]], 0 <= index < d, unique within a row (SparseVector keeps sorted unique indices)
 *   values   double[indptr[n]] */
int akm_upload_rows_csr(akm_handle h, const int64_t *indptr, const int32_t *indices, const double *values, int64_t n);

/* The upload calls above return once the caller's rows sit in the library's pinned staging buffers: the host
 * buffer may be refilled at once (KMeans' Java side fills its next chunk while this one is still crossing
 * PCIe), and every later call on the handle is ordered behind the copies.  akm_upload_wait blocks until all
 * uploaded rows are resident and reports any deferred CUDA error.  No reference counterpart. */
int akm_upload_wait(akm_handle h);

/* Page-locked host memory for upload buffers (e.g. behind a java.nio direct ByteBuffer that the Java side fills
 * instead of a double[]).  When the source of an upload call lies in such memory (or in memory the caller registered
 * with CUDA), the rows are read by the DMA engine straight from it: no staging copy, PCIe sets the rate.  The
 * buffer must then stay untouched until akm_upload_wait (or any later synchronising call on the handle) returns.
 * A partition whose doubles are all fp32-representable is still stored as fp32 (the check runs on the device).
 * No reference counterpart. */
int akm_host_alloc(size_t bytes, void **out);
int akm_host_free(void *p);

/* Same, but the source already lives in device memory on the handle's device (used to measure
 * the resident-data throughput and to feed data produced on the GPU). */
int akm_upload_rows_device_f32(akm_handle h, const float *x_dev, int64_t n);

/* Drops the cached partition (keeps the handle). */
int akm_clear_rows(akm_handle h);

/* Number of cached rows; storage kind: 4 = fp32, 8 = fp64. */
int akm_num_rows(akm_handle h, int64_t *n_out, int *bytes_per_value_out);

/* ------------------------------------------------------------------ per-superstep work, host buffers */

/* Replaces KMeansAssignCluster.calc (CORE/operator/common/clustering/kmeans/KMeansAssignCluster.java:30-67)
 * = KMeansUtil.updateSumMatrix for every cached sample (KMeansUtil.java:60-85) with
 * FastDistance.calc(vector, matrix) (EuclideanDistance.java:109-116, CosineDistance.java:107-110)
 * and getMinPointIndex (KMeansUtil.java:162-173).
 *   centroids  k0 x d, as stored in FastDistanceMatrixData.vectors (already label-updated:
 *              for COSINE the rows are unit vectors; stale rows >= k are zero).  COSINE rows that are NOT unit
 *              vectors get CosineDistance.calc's arithmetic on them as they are (1 - x.c), like in Java; the
 *              library then evaluates every centroid for every row in that pass (its bounds are unit-sphere geometry)
 *   k0         number of rows of the buffer (= the init matrix's numCols, KMeansPreallocateCentroid.java:22-39)
 *   k          current cluster count (context "k"), only the first k centroids are candidates
 *   sum_out    double[k0*(d+1)], fully overwritten ("Arrays.fill(sumMatrixData, 0.0)" then accumulate)
 *   objective_out  nullable; sum over samples of min distance^2 (EUCLIDEAN) / min distance (COSINE)
 *              against `centroids` (the reference train loop has no objective)
 *   labels_out nullable int32[n]: closest cluster index per cached sample (updateSumMatrix's return) */
int akm_assign_accumulate(akm_handle h, const double *centroids, int k0, int k,
                          double *sum_out, double *objective_out, int32_t *labels_out);

/* Replaces KMeansUpdateCentroids.updateCentroids (KMeansUpdateCentroids.java:52-69) followed by the
 * diagonal of KMeansIterTermination.calc (KMeansIterTermination.java:28-44).
 *   sum        double[k0*(d+1)] after the all-reduce
 *   step_no    ComContext.getStepNo(): the buffer written at an odd step is CENTROID1 (:38-42), which
 *              fixes the order of the two label additions in the shift distance
 *   prev       nullable k0 x d: the centroid buffer the assignment used (other ping-pong buffer);
 *              when given, shift_out[j] = distance(prev_j, new_j) as the matrix x matrix path computes it
 *   c_out      k0 x d, fully rewritten (zero fill, 1.0/weight scaling, empty clusters compacted away,
 *              updateLabel over all k0 rows: COSINE renormalises)
 *   k_out      new cluster count
 *   shift_out  nullable double[k0] (first *k_out entries valid) */
int akm_update_centroids(akm_handle h, const double *sum, int k0, int k, int step_no, const double *prev,
                         double *c_out, int *k_out, double *shift_out);

/* One whole superstep of KMeansTrainBatchOp.iterateICQ (KMeansTrainBatchOp.java:58-81) for hosts that keep the
 * centroid buffers in Java between supersteps: KMeansAssignCluster.calc (KMeansAssignCluster.java:30-67) ->
 * AllReduce(CENTROID_ALL_REDUCE) (AllReduce.java:85-120; ncclAllReduce on the device buffer when a communicator
 * is attached, identity otherwise) -> KMeansUpdateCentroids.updateCentroids (KMeansUpdateCentroids.java:52-69) ->
 * the diagonal of KMeansIterTermination.calc (KMeansIterTermination.java:28-44).  Same results as calling
 * akm_assign_accumulate, akm_allreduce_sum and akm_update_centroids(prev = centroids) one after the other, with one
 * host->device copy (the centroids), one device->host batch and one synchronisation; the accumulator stays on
 * the device.
 *   centroids      k0 x d as in akm_assign_accumulate (the buffer the assignment reads)
 *   step_no, tol   as in akm_update_centroids / akm_step
 *   c_out          k0 x d new centroid buffer;  k_out new cluster count;  converged_out nullable
 *   sum_out        nullable double[k0*(d+1)]: the all-reduced accumulator ("centroidAllReduce", whose counts
 *                  KMeansOutputModel reads)
 *   shift_out      nullable double[k0] (first *k_out valid);  objective_out nullable, over all ranks */
int akm_superstep(akm_handle h, const double *centroids, int k0, int k, int step_no, double tol, double *c_out,
                  int *k_out, int *converged_out, double *sum_out, double *shift_out, double *objective_out);

/* ------------------------------------------------------------------ device-resident iteration */

/* KMeansPreallocateCentroid.calc (KMeansPreallocateCentroid.java:22-39) + KMeansUtil.buildCentroidsMatrix's
 * updateLabel (KMeansUtil.java:35-45): publishes CENTROID1 = init, CENTROID2 = copy, K = k0, step = 0.
 * c0 = k0 x d raw initial centroids (for COSINE they are normalised here, like updateLabel does). */
int akm_set_centroids(akm_handle h, const double *c0, int k0);

/* One superstep of KMeansTrainBatchOp.iterateICQ (KMeansTrainBatchOp.java:58-81) entirely on the
 * device stream: KMeansAssignCluster -> AllReduce (NCCL when a communicator is attached, identity
 * otherwise) -> KMeansUpdateCentroids -> KMeansIterTermination.  Only three scalars come back.
 *   tol            epsilon (HasEpsilonDv00001.java:12-18)
 *   converged_out  1 when every centroid moved < tol (the ComQueue criterion)
 *   k_out          cluster count after the update
 *   objective_out  nullable, objective of this assignment pass over ALL ranks */
int akm_step(akm_handle h, double tol, int *converged_out, int *k_out, double *objective_out);

/* KMeansOutputModel.calc (KMeansOutputModel.java:37-80): newest centroid buffer (k0 x d, first *k_out
 * rows are the model), weights[j] = count of cluster j in the LAST assignment pass read at the
 * uncompacted index j (as the reference does), step counter. Any output may be NULL. */
int akm_get_model(akm_handle h, double *c_out, int *k_out, double *weights_out, int *steps_out);

/* Labels of the last assignment pass for the cached rows (int32[n]). */
int akm_get_labels(akm_handle h, int32_t *labels_out);

/* ------------------------------------------------------------------ AllReduce over NVLink (NCCL) */

/* Replaces AllReduce(CENTROID_ALL_REDUCE) (CORE/common/comqueue/communication/AllReduce.java:85-120,
 * SUM :125-133) inside akm_step.  id_out receives an ncclUniqueId (128 bytes) created on rank 0;
 * the host side broadcasts it (Flink broadcast set / torch.distributed) and every rank calls
 * akm_comm_init with it. */
#define AKM_UNIQUE_ID_BYTES 128
int akm_comm_unique_id(void *id_out);
int akm_comm_init(akm_handle h, const void *id, int rank, int world);
/* Stand-alone all-reduce (SUM, fp64) of a host buffer through the handle's communicator. */
int akm_allreduce_sum(akm_handle h, double *buf, int64_t count);

/* ------------------------------------------------------------------ prediction */

/* Replaces KMeansModelMapper.map (CORE/operator/common/clustering/kmeans/KMeansModelMapper.java:62-97)
 * for a micro-batch of n rows (the Java side batches inside mapPartition):
 * prepareVectorData -> getClusterDistances -> getMinPointIndex.
 *   x_host     n x d raw vectors; centroids = model matrix k x d as stored in the model rows
 *              (updateLabel is applied here like transformTrainDataToPredictData does, KMeansUtil.java:246-260)
 *   idx_out    int64[n]  = (long) index (:82); -1 when no distance is < Double.MAX_VALUE
 *   dist_out   nullable double[n]   = clusterDistances[index]   (predictionDistanceCol, :91-93)
 *   probs_out  nullable double[n*k] = getProbArrayFromDistanceArray (KMeansUtil.java:262-269), row-major,
 *              indexed by centroid position (the caller scatters by clusterId, :85-88) */
int akm_predict(akm_handle h, const double *x_host, int64_t n, const double *centroids, int k,
                int64_t *idx_out, double *dist_out, double *probs_out);

/* Prediction over the rows already cached on the device (KMeansPredictBatchOp over the training
 * table); outputs as above. */
int akm_predict_cached(akm_handle h, const double *centroids, int k,
                       int64_t *idx_out, double *dist_out);

/* ------------------------------------------------------------------ introspection (no reference counterpart) */

/* Counters of the last assignment pass: rows decided by the tensor-core search alone, rows that
 * were re-evaluated exactly, number of kernel launches issued by this handle so far. */
typedef struct {
    int64_t rows;
    int64_t rows_certified;
    int64_t rows_refined;
    int64_t kernel_launches;
    int32_t engine_used;      /* AKM_ENGINE_EXACT or AKM_ENGINE_TENSOR */
    int32_t storage_bytes;    /* 4 or 8 */
    int64_t rows_exact_all;   /* subset of rows_refined that was evaluated against all k centroids */
    int64_t tiles_total;      /* tensor engine: (256-row groups) x (256-centroid tiles) of the last pass */
    int64_t tiles_computed;   /* ... of which were evaluated; the rest were ruled out by bounds ("tile_pruning") */
    int64_t rows_moved;       /* rows whose label differs from the previous pass's (-1: no previous pass to compare with) */
    int32_t accumulate_mode;  /* 0: the accumulator was refilled from every row, 1: only the moved rows were applied */
    int32_t reserved0;
} akm_stats;
int akm_get_stats(akm_handle h, akm_stats *out);

/* Tuning options (no reference counterpart).  None of them changes a result.
 *   "reorder_min_rows"  partitions with at least this many rows are kept grouped by cluster inside the library
 *                       (outputs are always returned in the caller's row order); default 262144, a huge value
 *                       disables it.
 *   "regroup_after"     regroup once this many tensor-engine passes have produced labels (default 2: the labels of
 *                       the first pass come from the initial centroids and most of them move again).
 *   "tile_pruning"      (default 1) from the second pass of a training sequence on, the tensor engine skips every
 *                       (256-row group, 256-centroid tile) pair that the triangle inequality proves free of the
 *                       argmin and of every near-tie candidate, and -- on a partition that is kept grouped by
 *                       cluster -- every row group whose rows provably keep their centroid (a lower bound of each
 *                       row's distance to its runner-up is carried from pass to pass).  All of it is bounds from
 *                       the previous pass; akm_stats.tiles_computed / tiles_total reports how much was evaluated;
 *                       0 evaluates every pair.
 *   "delta_accumulate"  (default 1) once few rows change cluster per pass, the per-cluster sums are kept between
 *                       supersteps and only the rows whose label changed are moved (counts stay exact, sums pick
 *                       up one fp64 rounding per moved value; a full pass runs whenever more than 1/16 of the rows
 *                       moved in the previous pass, every 64 passes, and after any call that rewrites the labels);
 *                       0 refills the accumulator from every row in every superstep like the reference does. */
int akm_set_option(akm_handle h, const char *name, double value);

/* Per-phase device timing of akm_step with CUDA events on the handle's stream (off by default).
 * akm_get_profile fills out[0..6] = cumulative milliseconds of {nearest-centroid search, exact
 * refinement, per-cluster accumulation, all-reduce + centroid update}, the number of steps counted,
 * the milliseconds spent regrouping rows by cluster and the number of regroupings so far;
 * reset != 0 clears the counters.  No reference counterpart. */
int akm_set_profiling(akm_handle h, int on);
int akm_get_profile(akm_handle h, double *out7, int reset);

/* Synthetic Gaussian-blob rows of SURVEY 8(d), generated on the device by a counter-based hash RNG
 * (a pure function of seed and the global row index, so any shard can be produced independently):
 *   centre[j] ~ U(-spread, spread)^d,  row r = fp32( centre[blob(r)] + sigma * N(0,1)^d ),
 *   blob(r) = hash(seed, r) mod k_true.
 * akm_blobs_append appends rows [row0, row0+n) to the cached partition (prepareVectorData included);
 * akm_blobs_rows writes the m x d fp32 values of arbitrary global rows to host memory (used for the
 * initial centroids "rows floor(i*N/k)" and for the CPU baseline's sample).  No reference counterpart. */
int akm_blobs_append(akm_handle h, uint64_t seed, int k_true, double spread, double sigma, int64_t row0, int64_t n);
int akm_blobs_rows(akm_handle h, uint64_t seed, int k_true, double spread, double sigma, const int64_t *rows,
                   int64_t m, float *out_host);

/* ---- Initial centroids on the device --------------------------------------------------------
 * Replaces the distance passes of KMeansInitCentroids.kMeansPlusPlusInit
 * (operator/common/clustering/kmeans/KMeansInitCentroids.java:89-165) over the cached partition.  The library
 * keeps, per cached row, the Tuple5 fields f2 (nearest chosen centre id) and f3 (distance to it; the reference
 * uses the distance, not its square).  Centres are copies of rows (TransformToCenter :208-219 maps filtered
 * records); the rows themselves stay samples.  Call order, one handle per partition:
 *   akm_init_reset                       after the last upload: f2 = -1, f3 = Double.MAX_VALUE  (:97-105)
 *   akm_init_update(centers, m, base)    CalWeight (:221-248) against m new centres whose ids are base..base+m-1,
 *                                        then SUM(f3) over this partition (:120) -> *sum_cost_out
 *   akm_init_pick(scale, seed, round,..) FilterNewCenter (:185-206): a row is picked when u < f3 * scale
 *                                        (scale = 2k / all-reduced SUM).  u is a counter-based hash of
 *                                        (seed, round, row_base + row index), not java.util.Random(subtask): the
 *                                        pick does not depend on the sharding.  Writes the picked rows' indices
 *                                        (caller's row order, ascending) to rows_out[0..min(*count_out, cap));
 *                                        *count_out > cap means the list was cut: repeat the call with a larger
 *                                        cap (the call changes no state).
 *   akm_get_rows(rows, m, out)           the prepared vectors (COSINE: normalised) of the given rows, m x d
 *   akm_init_counts(m_total, counts)     number of rows whose nearest centre is j (:140-150), j < m_total
 * Row indices are positions in upload order, strictly increasing in every list.  The weighted k-means++ /
 * Lloyd on the ~2k candidates (LocalKmeansFunc.java:36-122) is host work in the reference and stays host work
 * (alink_b200/init.py). */
int akm_init_reset(akm_handle h);
int akm_init_update(akm_handle h, const double *centers, int m, int id_base, double *sum_cost_out);
int akm_init_pick(akm_handle h, double scale, uint64_t seed, int round, int64_t row_base, int cap, int64_t *rows_out,
                  int *count_out);
int akm_init_counts(akm_handle h, int m_total, double *counts_out);
int akm_get_rows(akm_handle h, const int64_t *rows_sorted, int m, double *out);

/* ---- Cluster evaluation over the cached partition -------------------------------------------
 * The per-row work of EvalClusterBatchOp (operator/batch/evaluation/EvalClusterBatchOp.java:84-104):
 * ClusterEvaluationUtil.getClusterStatistics (operator/common/evaluation/ClusterEvaluationUtil.java:171-218) per
 * cluster and calSilhouetteCoefficient (:106-145) per row.  `labels` are the predicted cluster ids of the cached
 * rows in upload order, as dense ints 0..k-1 (negative = row without a prediction: skipped, like a null field).
 * Outputs, all double, caller-allocated:
 *   count_out[k]        rows per cluster (0 for ids that never occur; the reference has no such group)
 *   mean_out[k*d]       sumVector.scale(1.0 / total); COSINE: mean of the normalised vectors
 *   dist_sum_out[k]     sum of distance.calc(mean, vec) -> compactness = dist_sum / count
 *   dist_sq_sum_out[k]  sum of the squared distances (SSW = sum over clusters)
 *   norm_sq_sum_out[k]  vectorNormL2Sum
 *   silhouette_sum_out  sum over rows of the silhouette value (the op divides by COUNT)
 * The k x k remainder (SSB, separation, Davies-Bouldin, Calinski-Harabasz: ClusterMetricsSummary.java:109-150)
 * is cheap host arithmetic on these arrays.  A COSINE handle evaluates with CosineDistance, a EUCLIDEAN one with
 * EuclideanDistance.  Overwrites the label buffer akm_get_labels reads. */
int akm_eval_clusters(akm_handle h, const int32_t *labels, int k, double *count_out, double *mean_out,
                      double *dist_sum_out, double *dist_sq_sum_out, double *norm_sq_sum_out, double *silhouette_sum_out);

/* Centroid update for callers that may not allocate device memory (no counterpart in the reference).
 * Meant for hosts that share a GPU under a fixed memory reservation or run on a latency-critical thread:
 * cudaMalloc/cudaFree can fail there and they stall every stream on the device.  The caller passes DEVICE
 * pointers for input and output; the kernel's working memory (the empty-cluster compaction map) is a fixed
 * array built into the library image (about 64 KB per GPU, held until the process exits), so this call
 * performs no allocation at all and needs no handle.
 *   sum_dev    double[k0*(d+1)] on `device`, the all-reduced accumulator
 *   c_out_dev  double[k0*d] on `device`, fully rewritten exactly like akm_update_centroids does
 *   k_out_dev  int on `device`, the new cluster count
 * Limits and rules: k0 <= AKM_PREALLOC_MAX_K, otherwise AKM_ERR_BADARG.  All calls on a device share the
 * built-in array and it keeps the previous call's contents (the call clears what it accumulates into), so
 * calls that overlap on the GPU race: serialise them.  The first call on each device looks up the array's
 * device address; make that call before any stream capture.  Work is enqueued on `cuda_stream` (may be
 * NULL); outputs are valid once that stream has reached the end of the call's work. */
#define AKM_PREALLOC_MAX_K 16384
int akm_update_centroids_prealloc(int device, void *cuda_stream, int metric, int d, const double *sum_dev,
                                  int k0, int k, double *c_out_dev, int *k_out_dev);

#ifdef __cplusplus
}
#endif
#endif /* AKM_H_ */
