// Internal declarations of libakm (not part of the C ABI).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <string>
#include <vector>

#include "../../include/akm.h"

// ---------------------------------------------------------------- row storage
// A block of samples resident in HBM after FastDistance.prepareVectorData.
//   x32 / x64 : n x d row-major, exactly the values Java holds (fp32 only when every
//               value is fp32-representable, so widening back to double is lossless)
//   xlabel    : EUCLIDEAN ||x||^2 summed in index order (EuclideanDistance.java:59-65); COSINE |x_raw|
//   COSINE    : an fp64 store holds the normalised rows.  An fp32 store (every RAW value fp32-representable)
//               holds the raw rows; the value Java holds, x * (1.0 / |x|), is recomputed on the fly with exactly
//               the rounding of CosineDistance.updateLabel (:60-65) -- see row_value().
//   x16       : n x d16 fp16 copy scaled by 2^x16_shift, operand of the tcgen05 search
//   xerr      : per row float2 {||x||_2 rounded up, ||x - x16/scale||_2 rounded up}
struct RowStore {
    int64_t n = 0, cap = 0;
    int store = 4;            // bytes per value of the exact copy (4 or 8)
    float *x32 = nullptr;
    double *x64 = nullptr;
    double *xlabel = nullptr;
    __half *x16 = nullptr;
    float2 *xerr = nullptr;
    int d16 = 0;              // row stride of x16 in elements (multiple of 8)
    int64_t x16_cap = 0;      // rows allocated for x16 / xerr
    bool want16 = false;
    int x16_shift = 0;        // x16 = fp16(x * 2^x16_shift)
    bool x16_valid = false;
    // Headroom for the regrouping (reorder.cu): the two wide arrays are allocated with `head` extra rows and the data
    // starts at row offset 0 or >= one wave inside its allocation; a regrouping moves every wave of rows into the
    // space the previous wave left, so the rows cross HBM once instead of twice.
    void *x_alloc = nullptr; int64_t x_head = 0, x_off = 0;        // allocation behind x32 / x64
    __half *x16_alloc = nullptr; int64_t x16_head = 0, x16_off = 0;
    uint64_t x16_gen = 0;     // bumped whenever x16 is rebuilt (the search seeds refer to one generation)
    // cluster-ordered layout (reorder.cu): orig[pos] = caller's row index of the row now stored at pos
    int32_t *orig = nullptr; int64_t orig_cap = 0, orig_n = 0;
    bool reordered = false;
};

// 1.0 / |x| the way CosineDistance.updateLabel forms it (untouched rows -- |x| not > 0 -- keep their values)
__device__ __forceinline__ double cos_inv(double nrm) { return nrm > 0.0 ? 1.0 / nrm : 1.0; }
// true when a partition of storage type T under METRIC keeps RAW rows that must be scaled on the fly
template <typename T, int METRIC> struct RawCosine { static constexpr bool value = (METRIC == AKM_COSINE) && sizeof(T) == 4; };

struct Scratch;   // akm_api.cu

struct akm_handle_s {
    int device = 0, d = 0, metric = 0, engine = AKM_ENGINE_AUTO;
    cudaStream_t own_stream = nullptr, stream = nullptr;
    int num_sms = 148;
    size_t smem_optin = 0;

    RowStore rows;            // cached training partition
    RowStore batch;           // scratch store for akm_predict micro-batches
    int32_t *labels = nullptr; int64_t labels_cap = 0;
    double *mind = nullptr; int64_t mind_cap = 0;   // per-row min distance (predict)

    // centroids: ping-pong (KMeansAssignCluster.java:38-42)
    int k0 = 0, k = 0, step = 0, cur = 0;
    double *cent[2] = {nullptr, nullptr};
    double *clabel[2] = {nullptr, nullptr};
    int cent_cap = 0;
    double *sum = nullptr;    // k0*(d+1) + 2 : [L] = objective, [L+1] = no-finite-distance flag (both all-reduced)
    // incremental accumulation (kernels_exact.cu, accumulate_step): sums of this partition and the label every
    // row was added under
    double *sum_local = nullptr; int sum_local_k0 = 0;
    double *sum_tmp = nullptr; int sum_tmp_k0 = 0;    // renumbering target (accumulate_renumber), swapped with sum_local
    int32_t *labels_acc = nullptr; int64_t labels_acc_cap = 0;
    bool acc_valid = false, acc_count_pending = false, xl_total_valid = false, delta_accumulate = true;
    int acc_k0 = 0, acc_delta_passes = 0, acc_mode_last = 0;
    int64_t acc_n = 0, acc_changed_total = 0, last_changed = -1;
    int *cmap = nullptr;      // k0 compaction map (old index -> new index)
    int *inv = nullptr; int inv_cap = 0;   // new index -> old index
    double *shift = nullptr;  // k0
    int kvec_cap = 0;
    Scratch *sc = nullptr;    // buffers of the host-buffer entry points
    double *dscal = nullptr;  // device scalars: [0] objective, [1] sum xlabel (local)
    int *iscal = nullptr;     // device ints: [0] new k, [1] converged, [2] error flag, [3] inexact flag, [4] worklist count
    double *h_dscal = nullptr; int *h_iscal = nullptr;   // pinned mirrors

    // staging for uploads
    void *stage = nullptr; size_t stage_bytes = 0;       // device staging (two halves, used alternately)
    void *csr_dev = nullptr; size_t csr_bytes = 0;       // device copy of one CSR chunk (akm_upload_rows_csr)
    void *hstage = nullptr; size_t hstage_bytes = 0;     // pinned host staging (two halves)
    cudaEvent_t stage_free[2] = {nullptr, nullptr};      // half i of both buffers may be refilled
    bool stage_used[2] = {false, false}; int stage_slot = 0;
    struct PendingChunk { bool active; bool prepared; int64_t row0, rows; };   // fp32-fit check still owed (akm_api.cu)
    PendingChunk pend[2] = {{false, false, 0, 0}, {false, false, 0, 0}};
    // pinned bounce arena for the few-MB host buffers of the per-superstep calls (centroids, accumulator):
    // copies from/to pageable memory would be staged by the driver and block the calling thread one by one
    char *pin = nullptr; size_t pin_bytes = 0, pin_off = 0;
    struct PendingOut { void *dst; const void *src; size_t bytes; };
    std::vector<PendingOut> pin_out;

    // tensor path scratch (kernels_tensor.cu)
    void *tc = nullptr;

    // NCCL
    void *comm = nullptr; int rank = 0, world = 1;
    double *ar_buf = nullptr; size_t ar_cap = 0;      // device staging of akm_allreduce_sum

    // k-means|| seeding state of the cached partition (init.cu): nearest chosen centre and the distance to it
    double *init_cost = nullptr; int32_t *init_near = nullptr; int64_t init_cap = 0;
    int64_t *init_rows = nullptr; int init_rows_cap = 0;      // device list of row indices (picks / gathers)
    double *init_out = nullptr; size_t init_out_count = 0;    // gathered rows / counts

    // cluster-ordered layout scratch and policy state (reorder.cu)
    void *reorder_scratch = nullptr; size_t reorder_scratch_bytes = 0;
    int32_t *reorder_dest = nullptr; int64_t reorder_dest_cap = 0;
    int *reorder_hist = nullptr; size_t reorder_hist_bytes = 0;
    int64_t changed_since_reorder = 0;
    int reorders = 0;
    double reorder_ms = 0;
    int64_t reorder_min_rows = 1 << 18;   // partitions smaller than this are left in input order
    int regroup_after = 2;                // regroup once this many tensor-engine passes have produced labels
    // the 256-row groups the tensor search of the running pass looked at (device list + length; nobody else's labels
    // can have changed): set by the search, consumed by accumulate_step of the same pass
    const int2 *searched_list = nullptr; const int *searched_n = nullptr; bool searched_valid = false;
    int epi_dbg = 0;                      // -DAKM_EPI_DBG builds: timing experiments on the search epilogue (results invalid when != 0)
    bool tile_pruning = true;             // skip (row group, centroid tile) pairs the triangle inequality rules out

    // optional per-phase profiling of akm_step
    bool profiling = false;
    cudaEvent_t pev[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
    double prof_ms[4] = {0, 0, 0, 0};
    int64_t prof_steps = 0;
    cudaEvent_t mid_event = nullptr;   // recorded by tensor_assign between search and refinement
    bool mid_recorded = false;

    akm_stats stats{};
    std::string err;
};

// ---------------------------------------------------------------- error helpers
void akm_set_tls_error(const std::string &s);
int akm_fail(akm_handle h, int code, const std::string &msg);

#define AKM_CUDA(h, call)                                                                  \
    do {                                                                                   \
        cudaError_t e__ = (call);                                                          \
        if (e__ != cudaSuccess)                                                            \
            return akm_fail((h), e__ == cudaErrorMemoryAllocation ? AKM_ERR_OOM : AKM_ERR_CUDA, \
                            std::string(#call) + ": " + cudaGetErrorString(e__));          \
    } while (0)

#define AKM_TRY(expr)                  \
    do {                               \
        int rc__ = (expr);             \
        if (rc__ != AKM_OK) return rc__; \
    } while (0)

// ---------------------------------------------------------------- akm_api.cu
// device -> pageable host copy through the pinned staging halves (synchronises the handle's stream)
int download_pipelined(akm_handle h, const void *dev, void *host, size_t bytes);

// ---------------------------------------------------------------- kernels_exact.cu
// prepareVectorData over `n` staged rows (src: double or float, row-major) appended at row `row0`.
// prepared: the caller already ran prepareVectorData (COSINE rows are stored as given, not normalised again)
int launch_prep_rows(akm_handle h, RowStore &rs, const void *src, bool src_is_f32, int64_t row0, int64_t n,
                     bool prepared = false, int *inexact_flag = nullptr);
// CSR rows [.., rows) -> dense doubles in dense_dev (indptr_dev holds rows + 1 entries whose first value is nnz0)
int launch_csr_expand(akm_handle h, const int64_t *indptr_dev, const int32_t *indices_dev, const double *values_dev,
                      int64_t nnz0, int64_t rows, double *dense_dev, int *bad_index_dev);
// widen rows [0,n) of x32 into x64 (storage migration)
int launch_widen_rows(akm_handle h, const float *x32, const double *xlabel, double *x64, int64_t n);
// updateLabel on k0 centroids in place (EUCL: labels; COS: normalise rows unless cos_normalise == false,
// which is what KMeansAssignCluster sees: a matrix that was label-updated already)
int launch_prepare_centroids(akm_handle h, double *C, double *clabel, int k0, bool cos_normalise);
// exact (Java order) assignment of rows [0,n) or of the rows in `rowlist`
int launch_assign_exact(akm_handle h, const RowStore &rs, const int32_t *rowlist, int64_t nrows,
                        const double *C, const double *clabel, int k0, int k, int32_t *labels,
                        double *mind, double *dist_all);
int launch_assign_exact_list(akm_handle h, const RowStore &rs, const int32_t *rowlist, const int *count_dev,
                             const double *C, const double *clabel, int k0, int k, int32_t *labels);
// per-cluster sums/counts (+ objective) from labels
int launch_accumulate(akm_handle h, const RowStore &rs, const int32_t *labels, const double *C, const double *clabel,
                      int k0, double *sum, double *objective);
// KMeansUpdateCentroids + KMeansIterTermination diag
// cmap / inv: k0 ints each owned by the caller's state (the device-resident loop and the host-buffer calls keep
// separate maps: the loop's map feeds the next pass's seeds); inv == nullptr selects the single-CTA kernel
int launch_update(akm_handle h, const double *sum, int k, int k0, double *Cnew, double *clnew,
                  const double *Cprev, const double *clprev, bool new_is_c1, double tol,
                  double *shift, int *iscal, int *cmap, int *inv);
// accumulation half of a superstep over h->rows / h->labels (full pass or moved rows only); sum_out: k0*(d+1)+2
// doubles = sums, objective of this partition, no-finite-distance flag.  accumulate_absorb() after h_iscal[8] is back.
int accumulate_step(akm_handle h, const double *C, const double *clabel, int k0, double *sum_out);
void accumulate_absorb(akm_handle h);
// after an update that dropped clusters: renumber the incremental sums and their label record by `cmap` (device)
int accumulate_renumber(akm_handle h, const int *cmap, int k_old, int k0);
int launch_flag_from_sum(akm_handle h, const double *flag);
// probs from all distances (getProbArrayFromDistanceArray)
int launch_probs(akm_handle h, const double *dist_all, int64_t n, int k, double *probs);
int launch_labels_to_i64(akm_handle h, const int32_t *labels, int64_t n, int64_t *out);

// ---------------------------------------------------------------- kernels_tensor.cu
bool tensor_supported(akm_handle h, int k0);
// approximate search on tensor cores + certification + exact refinement; fills labels (and mind when non-null)
int tensor_assign(akm_handle h, RowStore &rs, const double *C, const double *clabel, int k0, int k,
                  int32_t *labels, double *mind, int chain, const int *remap);
void tensor_invalidate_seeds(akm_handle h);
int tensor_prepare_x16(akm_handle h, RowStore &rs);
int tensor_warm(akm_handle h, RowStore &rs, int k0);
void tensor_free(akm_handle h);

// ---------------------------------------------------------------- akm_nccl.cu
int nccl_allreduce_device(akm_handle h, double *dbuf, int64_t count);
void nccl_destroy(akm_handle h);

// ---------------------------------------------------------------- reorder.cu
int reorder_rows_by_label(akm_handle h, RowStore &rs, int32_t *labels, float *useed, int k);
int reorder_reserve(akm_handle h, const RowStore &rs, int64_t n);
// rows of headroom the wide arrays of a partition of `cap` rows get (0: the partition is never regrouped)
int64_t reorder_head_rows(akm_handle h, const RowStore &rs, int store, int64_t cap);
int ensure_orig(akm_handle h, RowStore &rs);
int copy_rows_out_i32(akm_handle h, const RowStore &rs, const int32_t *dev, int32_t *host, int64_t n);
int copy_rows_out_i64(akm_handle h, const RowStore &rs, const int64_t *dev, int64_t *host, int64_t n);
int copy_rows_out_f64(akm_handle h, const RowStore &rs, const double *dev, double *host, int64_t n);
void tensor_rows_permuted(akm_handle h);   // kernels_tensor.cu: per-row state that does not move with a regrouping is dropped
float *tensor_useed(akm_handle h);   // kernels_tensor.cu: per-row seeds of the cached partition (may be null)

// ---------------------------------------------------------------- init.cu
int launch_init_reset(akm_handle h, int64_t n, double *cost, int32_t *near);
int launch_init_merge(akm_handle h, int64_t n, const int32_t *labels, const double *mind, int id_base, double *cost,
                      int32_t *near, double *sum_dev);
int launch_init_pick(akm_handle h, const RowStore &rs, const double *cost, int64_t row_base, double scale,
                     uint64_t seed, int round, int cap, int64_t *out_rows_dev, int *out_count_dev);
int launch_gather_rows(akm_handle h, const RowStore &rs, const int64_t *rows_sorted_dev, int m, double *out_dev);
int launch_init_counts(akm_handle h, int64_t n, const int32_t *near, int m, double *counts_dev);

// ---------------------------------------------------------------- eval.cu
int launch_labels_in(akm_handle h, const RowStore &rs, const int32_t *src_dev, int32_t *dst_dev);
int launch_eval_own(akm_handle h, const RowStore &rs, const int32_t *labels, int k, const double *mean,
                    const double *mean_norm, double *out3k);
int launch_eval_silhouette(akm_handle h, const RowStore &rs, const int32_t *labels, int k, const double *mean,
                           const double *cnt, const double *normsum, double *sil_sum);

// ---------------------------------------------------------------- blobs.cu
int launch_blobs(akm_handle h, uint64_t seed, int k_true, double spread, double sigma, const int64_t *rows_dev,
                 int64_t row0, int64_t m, float *out_dev);

static inline int64_t ceil_div64(int64_t a, int64_t b) { return (a + b - 1) / b; }
