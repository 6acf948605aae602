// k-means|| seeding passes over the cached partition.
//
// Reference: KMeansInitCentroids.kMeansPlusPlusInit (operator/common/clustering/kmeans/KMeansInitCentroids.java:89-165).
// There every sample carries a Tuple5 (id, vector, nearestCenterId f2, nearestCenterDist f3, isCenter f4); one
// round = SUM(f3) over the samples that are not centres, a Bernoulli pick with probability f3 * 2k / SUM
// (FilterNewCenter :185-206) and a CalWeight map (:221-248) that lowers (f2, f3) against the new centres.
// Centres are COPIES there (TransformToCenter maps the filtered records; `dataOnly` keeps every sample), so a
// picked row stays a data row: the next CalWeight finds it at distance 0 from its own copy, which both stops it
// from being picked again and gives every centre a weight of at least one in the final groupBy.
// Here f2/f3 are two device arrays in the partition's current row order, CalWeight is the ordinary
// nearest-centroid pass of the library against the new centres, and the pick uses a counter-based hash of the
// GLOBAL row index instead of java.util.Random(subtask) so that the result does not depend on how the rows are
// sharded over GPUs.
#include <cfloat>

#include "akm_internal.cuh"

__global__ void init_reset_kernel(int64_t n, double *__restrict__ cost, int32_t *__restrict__ near) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) { cost[i] = DBL_MAX; near[i] = -1; }
}

// CalWeight: (near, cost) <- min over the new centres; SUM(cost) over the partition
__global__ void __launch_bounds__(256)
init_merge_kernel(int64_t n, const int32_t *__restrict__ labels, const double *__restrict__ mind, int id_base,
                  double *__restrict__ cost, int32_t *__restrict__ near, double *__restrict__ sum_out) {
    __shared__ double red[8];
    double s = 0.0;
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) {
        double c = cost[i];
        const double dn = mind[i];
        if (labels[i] >= 0 && dn < c) { c = dn; cost[i] = dn; near[i] = id_base + labels[i]; }
        s += c;
    }
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0.0;
        for (int w = 0; w < 8; ++w) a += red[w];
        atomicAdd(sum_out, a);
    }
}

__device__ __forceinline__ uint64_t mix64(uint64_t z) {      // splitmix64 finaliser
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

// FilterNewCenter: a copy of the row becomes a centre when u(seed, round, global row) < cost * scale
__global__ void init_pick_kernel(int64_t n, const double *__restrict__ cost, const int32_t *__restrict__ orig, int64_t row_base, double scale, uint64_t seed,
                                 int round, int cap, int64_t *__restrict__ out_rows, int *__restrict__ out_count) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) {
        const int64_t local = orig ? (int64_t)orig[i] : i;
        const uint64_t g = (uint64_t)(row_base + local);
        const uint64_t hsh = mix64(mix64(seed + 0x9E3779B97F4A7C15ull * (uint64_t)(round + 1)) ^ (g * 0xD1B54A32D192ED03ull));
        const double u = (double)(hsh >> 11) * (1.0 / 9007199254740992.0);      // 53 bits -> [0, 1)
        if (u < cost[i] * scale) {
            const int pos = atomicAdd(out_count, 1);
            if (pos < cap) out_rows[pos] = local;
        }
    }
}

__device__ __forceinline__ int find_sorted(const int64_t *__restrict__ a, int m, int64_t v) {
    int lo = 0, hi = m;
    while (lo < hi) { const int mid = (lo + hi) >> 1; if (a[mid] < v) lo = mid + 1; else hi = mid; }
    return (lo < m && a[lo] == v) ? lo : -1;
}

// one warp per stored row: copy it (as double) to out[p] when the caller asked for it
// (xnorm != nullptr: COSINE rows stored raw; the prepared vector is x * (1.0 / |x|))
template <typename T>
__global__ void gather_rows_kernel(int64_t n, const T *__restrict__ X, const double *__restrict__ xnorm, int d,
                                   const int32_t *__restrict__ orig, const int64_t *__restrict__ rows_sorted, int m,
                                   double *__restrict__ out) {
    const int lane = threadIdx.x & 31;
    int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t stride = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (; i < n; i += stride) {
        const int64_t local = orig ? (int64_t)orig[i] : i;
        const int p = find_sorted(rows_sorted, m, local);
        if (p < 0) continue;
        const double rinv = xnorm ? cos_inv(xnorm[i]) : 1.0;
        for (int c = lane; c < d; c += 32) {
            const double v = (double)X[i * d + c];
            out[(size_t)p * d + c] = xnorm ? __dmul_rn(rinv, v) : v;
        }
    }
}

// number of rows whose nearest centre is j (the groupBy/SUM of :140-150)
__global__ void init_counts_kernel(int64_t n, const int32_t *__restrict__ near, int m, double *__restrict__ counts) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) {
        const int j = near[i];
        // consecutive rows usually share the centre: one atomic per run inside a warp
        const unsigned peers = __match_any_sync(__activemask(), j);
        if (j >= 0 && j < m && (int)(__ffs(peers) - 1) == (int)(threadIdx.x & 31)) atomicAdd(counts + j, (double)__popc(peers));
    }
}

int launch_init_reset(akm_handle h, int64_t n, double *cost, int32_t *near) {
    if (n <= 0) return AKM_OK;
    init_reset_kernel<<<h->num_sms * 8, 256, 0, h->stream>>>(n, cost, near);
    h->stats.kernel_launches++;
    AKM_CUDA(h, cudaGetLastError());
    return AKM_OK;
}

int launch_init_merge(akm_handle h, int64_t n, const int32_t *labels, const double *mind, int id_base, double *cost,
                      int32_t *near, double *sum_dev) {
    AKM_CUDA(h, cudaMemsetAsync(sum_dev, 0, sizeof(double), h->stream));
    if (n <= 0) return AKM_OK;
    init_merge_kernel<<<h->num_sms * 8, 256, 0, h->stream>>>(n, labels, mind, id_base, cost, near, sum_dev);
    h->stats.kernel_launches++;
    AKM_CUDA(h, cudaGetLastError());
    return AKM_OK;
}

int launch_init_pick(akm_handle h, const RowStore &rs, const double *cost, int64_t row_base, double scale,
                     uint64_t seed, int round, int cap, int64_t *out_rows_dev, int *out_count_dev) {
    AKM_CUDA(h, cudaMemsetAsync(out_count_dev, 0, sizeof(int), h->stream));
    if (rs.n <= 0) return AKM_OK;
    init_pick_kernel<<<h->num_sms * 8, 256, 0, h->stream>>>(rs.n, cost, rs.reordered ? rs.orig : nullptr, row_base, scale,
                                                            seed, round, cap, out_rows_dev, out_count_dev);
    h->stats.kernel_launches++;
    AKM_CUDA(h, cudaGetLastError());
    return AKM_OK;
}

int launch_gather_rows(akm_handle h, const RowStore &rs, const int64_t *rows_sorted_dev, int m, double *out_dev) {
    if (rs.n <= 0 || m <= 0) return AKM_OK;
    const int32_t *orig = rs.reordered ? rs.orig : nullptr;
    const double *xnorm = (h->metric == AKM_COSINE && rs.store == 4) ? rs.xlabel : nullptr;
    if (rs.store == 4) gather_rows_kernel<float><<<h->num_sms * 8, 256, 0, h->stream>>>(rs.n, rs.x32, xnorm, h->d, orig, rows_sorted_dev, m, out_dev);
    else gather_rows_kernel<double><<<h->num_sms * 8, 256, 0, h->stream>>>(rs.n, rs.x64, nullptr, h->d, orig, rows_sorted_dev, m, out_dev);
    h->stats.kernel_launches++;
    AKM_CUDA(h, cudaGetLastError());
    return AKM_OK;
}

int launch_init_counts(akm_handle h, int64_t n, const int32_t *near, int m, double *counts_dev) {
    AKM_CUDA(h, cudaMemsetAsync(counts_dev, 0, sizeof(double) * (size_t)m, h->stream));
    if (n <= 0) return AKM_OK;
    init_counts_kernel<<<h->num_sms * 8, 256, 0, h->stream>>>(n, near, m, counts_dev);
    h->stats.kernel_launches++;
    AKM_CUDA(h, cudaGetLastError());
    return AKM_OK;
}
