// Cluster evaluation over the cached partition: the per-row passes of EvalClusterBatchOp.
//
// Reference: ClusterEvaluationUtil.getClusterStatistics (operator/common/evaluation/ClusterEvaluationUtil.java:171-218)
// -- per cluster: mean vector, sum of distances to the mean, sum of their squares, sum of squared norms -- and
// calSilhouetteCoefficient (:106-145), which evaluates every row against every cluster's summary.  The k x k
// part (SSB, separation, Davies-Bouldin, Calinski-Harabasz; ClusterMetricsSummary.java:109-150) is host work.
#include <cfloat>
#include <climits>

#include "akm_internal.cuh"

// labels in the caller's row order -> the partition's current row order
__global__ void labels_in_kernel(const int32_t *__restrict__ src, const int32_t *__restrict__ orig, int64_t n,
                                 int32_t *__restrict__ dst) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) dst[i] = src[orig ? orig[i] : i];
}

// One warp per row: distance of the row to the mean of its own cluster, the way ContinuousDistance.calc(Vector,
// Vector) evaluates it (EuclideanDistance.java:45-47: sqrt(sum (a-b)^2); CosineDistance.java:46-50:
// 1 - a.b / (|a| |b|)), accumulated per cluster: out[0][c] += d, out[1][c] += d*d, out[2][c] += |x|^2.
template <typename T, int METRIC>
__global__ void eval_own_kernel(const T *__restrict__ X, int64_t n, int d, const int32_t *__restrict__ labels, int k,
                                const double *__restrict__ mean, const double *__restrict__ mean_norm,
                                const double *__restrict__ xnorm, double *__restrict__ out) {
    const int lane = threadIdx.x & 31;
    int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t stride = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (; row < n; row += stride) {
        const int c = labels[row];
        if (c < 0 || c >= k) continue;
        const T *x = X + row * d;
        const double *m = mean + (size_t)c * d;
        double a = 0.0, nsq = 0.0;
        const double rinv = RawCosine<T, METRIC>::value ? cos_inv(xnorm[row]) : 1.0;
        for (int i = lane; i < d; i += 32) {
            double xv = (double)x[i];
            if (RawCosine<T, METRIC>::value) xv = __dmul_rn(rinv, xv);
            nsq = fma(xv, xv, nsq);
            if (METRIC == AKM_EUCLIDEAN) { const double df = m[i] - xv; a = fma(df, df, a); }
            else a = fma(m[i], xv, a);
        }
        for (int o = 16; o > 0; o >>= 1) {
            a += __shfl_xor_sync(0xffffffffu, a, o);
            nsq += __shfl_xor_sync(0xffffffffu, nsq, o);
        }
        if (lane == 0) {
            double dist;
            if (METRIC == AKM_EUCLIDEAN) dist = sqrt(a);
            else { const double cross = mean_norm[c] * sqrt(nsq); dist = 1.0 - (cross > 0.0 ? a / cross : 0.0); }
            atomicAdd(out + c, dist);
            atomicAdd(out + k + c, dist * dist);
            atomicAdd(out + 2 * k + c, nsq);
        }
    }
}

// Silhouette: every row against every cluster summary.  64 rows x 64 clusters per tile, 4 x 4 per thread
// (the tiling of assign_exact_kernel), fp64 FMA dot products.
//   EUCLIDEAN  diss_j = cnt_j |x|^2 - 2 cnt_j x.mean_j + normsum_j ; own: / (cnt_j - 1), others: min over j of / cnt_j
//   COSINE     diss_j = 1 - x_raw.mean_j                           ; own: * cnt_j / (cnt_j - 1), others: min over j
// (the reference takes the RAW vector here although the means are means of normalised vectors; rowscalar carries
// |x_raw| for a COSINE partition: an fp64 store holds normalised rows and multiplies back, an fp32 store holds the
// raw rows and uses them as they are)
constexpr int SV_TR = 64, SV_TC = 64, SV_KC = 32, SV_LD = 66;

template <typename T, int METRIC>
__global__ void __launch_bounds__(256)
eval_silhouette_kernel(const T *__restrict__ X, const double *__restrict__ rowscalar, int64_t n, int d,
                       const int32_t *__restrict__ labels, int k, const double *__restrict__ mean,
                       const double *__restrict__ cnt, const double *__restrict__ normsum, double *__restrict__ sil_sum) {
    __shared__ __align__(16) double xs[SV_KC][SV_LD];
    __shared__ __align__(16) double cs[SV_KC][SV_LD];
    __shared__ double red[8];
    const int tid = threadIdx.x;
    const int tr = tid >> 4, tc = tid & 15;
    const int lcol = tid & 31, lrow0 = tid >> 5;
    double total = 0.0;
    for (int64_t tile = blockIdx.x; tile * SV_TR < n; tile += gridDim.x) {
        const int64_t base = tile * SV_TR;
        double cur[4], nb[4], rs[4];
        int lab[4];
#pragma unroll
        for (int a = 0; a < 4; ++a) {
            const int64_t r = base + 4 * tr + a;
            cur[a] = 0.0; nb[a] = DBL_MAX;
            lab[a] = (r < n) ? labels[r] : -1;
            rs[a] = (r < n) ? rowscalar[r] : 0.0;
        }
        for (int c0 = 0; c0 < k; c0 += SV_TC) {
            double t[4][4];
#pragma unroll
            for (int a = 0; a < 4; ++a)
#pragma unroll
                for (int b = 0; b < 4; ++b) t[a][b] = 0.0;
            for (int i0 = 0; i0 < d; i0 += SV_KC) {
                __syncthreads();
#pragma unroll
                for (int it = 0; it < 8; ++it) {
                    const int r = lrow0 + 8 * it;
                    const int col = i0 + lcol;
                    const int64_t row = base + r;
                    double xv = 0.0, cv = 0.0;
                    if (row < n && col < d) xv = (double)X[row * d + col];
                    const int cj = c0 + r;
                    if (cj < k && col < d) cv = mean[(size_t)cj * d + col];
                    xs[lcol][r] = xv;
                    cs[lcol][r] = cv;
                }
                __syncthreads();
                const int lim = min(SV_KC, d - i0);
                for (int i = 0; i < lim; ++i) {
                    const double2 xa = *reinterpret_cast<const double2 *>(&xs[i][4 * tr]);
                    const double2 xb = *reinterpret_cast<const double2 *>(&xs[i][4 * tr + 2]);
                    const double2 ca = *reinterpret_cast<const double2 *>(&cs[i][4 * tc]);
                    const double2 cb = *reinterpret_cast<const double2 *>(&cs[i][4 * tc + 2]);
                    const double xv[4] = {xa.x, xa.y, xb.x, xb.y};
                    const double cv[4] = {ca.x, ca.y, cb.x, cb.y};
#pragma unroll
                    for (int a = 0; a < 4; ++a)
#pragma unroll
                        for (int b = 0; b < 4; ++b) t[a][b] = fma(cv[b], xv[a], t[a][b]);
                }
            }
#pragma unroll
            for (int b = 0; b < 4; ++b) {
                const int j = c0 + 4 * tc + b;
                if (j >= k) continue;
                const double cj = cnt[j];
                if (!(cj > 0.0)) continue;                    // empty clusters do not exist in the reference's summary
                const double nsj = normsum[j];
#pragma unroll
                for (int a = 0; a < 4; ++a) {
                    if (METRIC == AKM_EUCLIDEAN) {
                        const double diss = cj * rs[a] - 2.0 * cj * t[a][b] + nsj;
                        if (j == lab[a]) { if (cj > 1.0) cur[a] = diss / (cj - 1.0); }
                        else nb[a] = fmin(nb[a], diss / cj);
                    } else {
                        const double diss = 1.0 - (RawCosine<T, METRIC>::value ? t[a][b] : rs[a] * t[a][b]);
                        if (j == lab[a]) { if (cj > 1.0) cur[a] = diss * cj / (cj - 1.0); }
                        else nb[a] = fmin(nb[a], diss);
                    }
                }
            }
        }
#pragma unroll
        for (int a = 0; a < 4; ++a) {
#pragma unroll
            for (int o = 8; o > 0; o >>= 1) {
                nb[a] = fmin(nb[a], __shfl_xor_sync(0xffffffffu, nb[a], o));
                cur[a] += __shfl_xor_sync(0xffffffffu, cur[a], o);          // exactly one lane holds a non-zero value
            }
            if (tc == 0 && lab[a] >= 0 && lab[a] < k)
                total += (cur[a] < nb[a]) ? 1.0 - (cur[a] / nb[a]) : (nb[a] / cur[a]) - 1.0;
        }
        __syncthreads();
    }
    for (int o = 16; o > 0; o >>= 1) total += __shfl_xor_sync(0xffffffffu, total, o);
    if ((tid & 31) == 0) red[tid >> 5] = total;
    __syncthreads();
    if (tid == 0) {
        double a = 0.0;
        for (int w = 0; w < 8; ++w) a += red[w];
        atomicAdd(sil_sum, a);
    }
}

int launch_labels_in(akm_handle h, const RowStore &rs, const int32_t *src_dev, int32_t *dst_dev) {
    if (rs.n <= 0) return AKM_OK;
    labels_in_kernel<<<h->num_sms * 8, 256, 0, h->stream>>>(src_dev, rs.reordered ? rs.orig : nullptr, rs.n, dst_dev);
    h->stats.kernel_launches++;
    AKM_CUDA(h, cudaGetLastError());
    return AKM_OK;
}

int launch_eval_own(akm_handle h, const RowStore &rs, const int32_t *labels, int k, const double *mean,
                    const double *mean_norm, double *out3k) {
    AKM_CUDA(h, cudaMemsetAsync(out3k, 0, sizeof(double) * 3 * (size_t)k, h->stream));
    if (rs.n <= 0) return AKM_OK;
    const int blocks = h->num_sms * 8;
    if (h->metric == AKM_EUCLIDEAN) {
        if (rs.store == 4) eval_own_kernel<float, AKM_EUCLIDEAN><<<blocks, 256, 0, h->stream>>>(rs.x32, rs.n, h->d, labels, k, mean, mean_norm, rs.xlabel, out3k);
        else eval_own_kernel<double, AKM_EUCLIDEAN><<<blocks, 256, 0, h->stream>>>(rs.x64, rs.n, h->d, labels, k, mean, mean_norm, rs.xlabel, out3k);
    } else {
        if (rs.store == 4) eval_own_kernel<float, AKM_COSINE><<<blocks, 256, 0, h->stream>>>(rs.x32, rs.n, h->d, labels, k, mean, mean_norm, rs.xlabel, out3k);
        else eval_own_kernel<double, AKM_COSINE><<<blocks, 256, 0, h->stream>>>(rs.x64, rs.n, h->d, labels, k, mean, mean_norm, rs.xlabel, out3k);
    }
    h->stats.kernel_launches++;
    AKM_CUDA(h, cudaGetLastError());
    return AKM_OK;
}

int launch_eval_silhouette(akm_handle h, const RowStore &rs, const int32_t *labels, int k, const double *mean,
                           const double *cnt, const double *normsum, double *sil_sum) {
    AKM_CUDA(h, cudaMemsetAsync(sil_sum, 0, sizeof(double), h->stream));
    if (rs.n <= 0) return AKM_OK;
    const int blocks = (int)std::min<int64_t>(ceil_div64(rs.n, SV_TR), (int64_t)h->num_sms * 8);
    if (h->metric == AKM_EUCLIDEAN) {
        if (rs.store == 4) eval_silhouette_kernel<float, AKM_EUCLIDEAN><<<blocks, 256, 0, h->stream>>>(rs.x32, rs.xlabel, rs.n, h->d, labels, k, mean, cnt, normsum, sil_sum);
        else eval_silhouette_kernel<double, AKM_EUCLIDEAN><<<blocks, 256, 0, h->stream>>>(rs.x64, rs.xlabel, rs.n, h->d, labels, k, mean, cnt, normsum, sil_sum);
    } else {
        if (rs.store == 4) eval_silhouette_kernel<float, AKM_COSINE><<<blocks, 256, 0, h->stream>>>(rs.x32, rs.xlabel, rs.n, h->d, labels, k, mean, cnt, normsum, sil_sum);
        else eval_silhouette_kernel<double, AKM_COSINE><<<blocks, 256, 0, h->stream>>>(rs.x64, rs.xlabel, rs.n, h->d, labels, k, mean, cnt, normsum, sil_sum);
    }
    h->stats.kernel_launches++;
    AKM_CUDA(h, cudaGetLastError());
    return AKM_OK;
}
