// Exact (fp64, Java operation order) kernels of the KMeans hot path.
//
// "Java order" means: every product and every sum is rounded separately (no FMA:
// __dmul_rn / __dadd_rn are never contracted by nvcc) and dot products run in
// index order, exactly like the F2J reference BLAS loops the Alink code reaches
// (dgemv 'T': BLAS.java:204-212; daxpy: BLAS.java:96-98; own dot: BLAS.java:103-110).
// These kernels are the ground truth on the device: the tensor-core search
// (kernels_tensor.cu) only proposes candidates and falls back to them.
#include <algorithm>
#include <cfloat>
#include <climits>
#include <cstdlib>
#include <type_traits>

#include "akm_internal.cuh"

#define DMUL(a, b) __dmul_rn((a), (b))
#define DADD(a, b) __dadd_rn((a), (b))

// =====================================================================================
// prepareVectorData  (FastDistance.java:136-140 -> Euclidean/CosineDistance.updateLabel)
// =====================================================================================
constexpr int PREP_ROWS = 128;   // rows per CTA == threads per CTA
constexpr int PREP_COLS = 32;

template <typename SrcT, int METRIC, int STORE>
__global__ void __launch_bounds__(PREP_ROWS)
prep_rows_kernel(const SrcT *__restrict__ src, int64_t n, int d, int64_t row0, float *__restrict__ x32,
                 double *__restrict__ x64, double *__restrict__ xlabel, double *__restrict__ sum_xlabel,
                 int *__restrict__ inexact_flag, int prepared) {
    __shared__ double tile[PREP_ROWS][PREP_COLS + 1];
    __shared__ double inv_s[PREP_ROWS];
    __shared__ double red[PREP_ROWS / 32];
    const int tid = threadIdx.x;
    const int64_t base = (int64_t)blockIdx.x * PREP_ROWS;
    const int lc = tid % PREP_COLS, lr0 = tid / PREP_COLS;
    bool inexact = false;
    double s = 0.0;   // sequential sum of squares of row `tid`

    for (int c0 = 0; c0 < d; c0 += PREP_COLS) {
        __syncthreads();
#pragma unroll 8
        for (int it = 0; it < PREP_ROWS / (PREP_ROWS / PREP_COLS); ++it) {
            const int r = lr0 + it * (PREP_ROWS / PREP_COLS);
            const int64_t row = base + r;
            const int col = c0 + lc;
            double v = 0.0;
            if (row < n && col < d) {
                v = (double)src[row * d + col];
                if (METRIC == AKM_EUCLIDEAN) {
                    if (STORE == 4) {
                        const float f = (float)v;
                        if (!((double)f == v)) inexact = true;   // NaN/overflow also land here
                        x32[(row0 + row) * d + col] = f;
                    } else {
                        x64[(row0 + row) * d + col] = v;
                    }
                }
            }
            tile[r][lc] = v;
        }
        __syncthreads();
        const int lim = min(PREP_COLS, d - c0);
        for (int i = 0; i < lim; ++i) {
            const double v = tile[tid][i];
            s = DADD(s, DMUL(v, v));
        }
    }
    const int64_t myrow = base + tid;
    if (METRIC == AKM_EUCLIDEAN) {
        if (myrow < n) xlabel[row0 + myrow] = s;
        // block sum of labels (only used for diagnostics)
        double t = (myrow < n) ? s : 0.0;
        for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
        if ((tid & 31) == 0) red[tid >> 5] = t;
        __syncthreads();
        if (tid == 0) {
            double a = 0.0;
            for (int w = 0; w < PREP_ROWS / 32; ++w) a += red[w];
            atomicAdd(sum_xlabel, a);
        }
    } else {
        // CosineDistance.updateLabel :60-65: norm = sqrt(dot); if (norm > 0) x *= 1.0 / norm
        const double nrm = sqrt(s);
        inv_s[tid] = (nrm > 0.0 && !prepared) ? (1.0 / nrm) : -1.0;   // -1 == leave untouched
        if (myrow < n) xlabel[row0 + myrow] = nrm;          // |x_raw|: not a distance label; scales fp32-stored rows
        __syncthreads();
        // fp32 store: raw rows (the scaling happens where they are read); fp64 store: normalised rows
        for (int c0 = 0; c0 < d; c0 += PREP_COLS) {
#pragma unroll 8
            for (int it = 0; it < PREP_ROWS / (PREP_ROWS / PREP_COLS); ++it) {
                const int r = lr0 + it * (PREP_ROWS / PREP_COLS);
                const int64_t row = base + r;
                const int col = c0 + lc;
                if (row < n && col < d) {
                    double v = (double)src[row * d + col];
                    if (STORE == 4) {
                        const float f = (float)v;
                        if (!((double)f == v)) inexact = true;
                        x32[(row0 + row) * d + col] = f;
                    } else {
                        const double inv = inv_s[r];
                        if (inv >= 0.0) v = DMUL(inv, v);
                        x64[(row0 + row) * d + col] = v;
                    }
                }
            }
        }
    }
    if (STORE == 4 && inexact) atomicExch(inexact_flag, 1);
}

int launch_prep_rows(akm_handle h, RowStore &rs, const void *src, bool src_is_f32, int64_t row0, int64_t n, bool prepared,
                     int *inexact_flag) {
    if (n <= 0) return AKM_OK;
    const unsigned grid = (unsigned)ceil_div64(n, PREP_ROWS);
    double *sx = h->dscal + 1;
    int *flag = inexact_flag ? inexact_flag : h->iscal + 3;
#define PREP_LAUNCH(SRC, M, ST)                                                              \
    prep_rows_kernel<SRC, M, ST><<<grid, PREP_ROWS, 0, h->stream>>>((const SRC *)src, n, h->d, row0, rs.x32, \
                                                                    rs.x64, rs.xlabel, sx, flag, prepared ? 1 : 0)
    if (h->metric == AKM_EUCLIDEAN) {
        if (rs.store == 4) { if (src_is_f32) PREP_LAUNCH(float, AKM_EUCLIDEAN, 4); else PREP_LAUNCH(double, AKM_EUCLIDEAN, 4); }
        else               { if (src_is_f32) PREP_LAUNCH(float, AKM_EUCLIDEAN, 8); else PREP_LAUNCH(double, AKM_EUCLIDEAN, 8); }
    } else {
        if (rs.store == 4) { if (src_is_f32) PREP_LAUNCH(float, AKM_COSINE, 4); else PREP_LAUNCH(double, AKM_COSINE, 4); }
        else               { if (src_is_f32) PREP_LAUNCH(float, AKM_COSINE, 8); else PREP_LAUNCH(double, AKM_COSINE, 8); }
    }
#undef PREP_LAUNCH
    h->stats.kernel_launches++;
    AKM_CUDA(h, cudaGetLastError());
    return AKM_OK;
}

// CSR -> dense rows in a staging buffer (zero-filled first): one warp per row walks its non-zeros.
__global__ void csr_expand_kernel(const int64_t *__restrict__ indptr, const int32_t *__restrict__ indices,
                                  const double *__restrict__ values, int64_t nnz0, int64_t rows, int d,
                                  double *__restrict__ dense, int *__restrict__ bad_index) {
    const int lane = threadIdx.x & 31;
    int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t stride = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (; r < rows; r += stride) {
        const int64_t a = indptr[r] - nnz0, b = indptr[r + 1] - nnz0;
        for (int64_t i = a + lane; i < b; i += 32) {
            const int c = indices[i];
            if (c >= 0 && c < d) dense[r * d + c] = values[i];
            else atomicExch(bad_index, 1);
        }
    }
}

int launch_csr_expand(akm_handle h, const int64_t *indptr_dev, const int32_t *indices_dev, const double *values_dev,
                      int64_t nnz0, int64_t rows, double *dense_dev, int *bad_index_dev) {
    AKM_CUDA(h, cudaMemsetAsync(dense_dev, 0, sizeof(double) * (size_t)rows * (size_t)h->d, h->stream));
    if (rows <= 0) return AKM_OK;
    const int blocks = (int)std::min<int64_t>((rows + 7) / 8, (int64_t)h->num_sms * 16);
    csr_expand_kernel<<<blocks, 256, 0, h->stream>>>(indptr_dev, indices_dev, values_dev, nnz0, rows, h->d, dense_dev, bad_index_dev);
    h->stats.kernel_launches++;
    AKM_CUDA(h, cudaGetLastError());
    return AKM_OK;
}

// fp32 -> fp64 storage migration.  EUCLIDEAN: plain widening.  COSINE (xlabel != nullptr): the fp64 store holds
// the normalised rows, so the scaling a raw fp32 store defers is applied here.
__global__ void widen_rows_kernel(const float *__restrict__ a, const double *__restrict__ xlabel, int d,
                                  double *__restrict__ b, int64_t count) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < count; i += stride) {
        double v = (double)a[i];
        if (xlabel != nullptr) { const double nrm = xlabel[i / d]; if (nrm > 0.0) v = DMUL(1.0 / nrm, v); }
        b[i] = v;
    }
}

int launch_widen_rows(akm_handle h, const float *x32, const double *xlabel, double *x64, int64_t n) {
    if (n <= 0) return AKM_OK;
    widen_rows_kernel<<<h->num_sms * 8, 256, 0, h->stream>>>(x32, h->metric == AKM_COSINE ? xlabel : nullptr, h->d, x64,
                                                           n * (int64_t)h->d);
    h->stats.kernel_launches++;
    AKM_CUDA(h, cudaGetLastError());
    return AKM_OK;
}

// =====================================================================================
// updateLabel on a centroid matrix (EuclideanDistance.java:66-86, CosineDistance.java:66-84)
// =====================================================================================
__device__ __forceinline__ void centroid_label_row(int metric, double *c, int d, double *label_out,
                                                   bool cos_normalise = true) {
    double s = 0.0;
    for (int i = 0; i < d; ++i) s = DADD(s, DMUL(c[i], c[i]));
    if (metric == AKM_EUCLIDEAN) {
        *label_out = s;
    } else if (!cos_normalise) {
        *label_out = 0.0;
    } else {
        const double nrm = sqrt(s);
        if (nrm > 0.0) {
            const double inv = 1.0 / nrm;
            for (int i = 0; i < d; ++i) c[i] = DMUL(inv, c[i]);   // BLAS.scal
        }
        *label_out = 0.0;
    }
}

__global__ void prepare_centroids_kernel(int metric, double *C, double *clabel, int k0, int d, int cos_normalise) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < k0) centroid_label_row(metric, C + (size_t)j * d, d, clabel + j, cos_normalise != 0);
}

int launch_prepare_centroids(akm_handle h, double *C, double *clabel, int k0, bool cos_normalise) {
    if (k0 <= 0) return AKM_OK;
    prepare_centroids_kernel<<<(k0 + 127) / 128, 128, 0, h->stream>>>(h->metric, C, clabel, k0, h->d,
                                                                       cos_normalise ? 1 : 0);
    h->stats.kernel_launches++;
    AKM_CUDA(h, cudaGetLastError());
    return AKM_OK;
}

// =====================================================================================
// Exact assignment: KMeansUtil.updateSumMatrix's distance + argmin part
//   t_j = sum_i C[j][i]*x[i]  (index order, separate roundings)
//   EUCL: D_j = sqrt(|((0 + -2*t_j) + xlabel) + clabel_j|)      EuclideanDistance.java:109-116
//   COS : D_j = 1 + (-1)*t_j                                    CosineDistance.java:107-110
//   a   = first j < k with D_j < running min (start Double.MAX_VALUE)   KMeansUtil.java:162-173
// One CTA = 64 rows; loops over 64-centroid tiles; each thread owns a 4x4 (row, centroid) patch.
// =====================================================================================
constexpr int AE_TR = 64, AE_TC = 64, AE_KC = 32, AE_LD = 66;

template <typename T, int METRIC, bool INDIRECT>
__global__ void __launch_bounds__(256)
assign_exact_kernel(const T *__restrict__ X, const double *__restrict__ xlabel, const int32_t *__restrict__ rowlist,
                    int64_t nrows_arg, const int *__restrict__ nrows_dev, int d, const double *__restrict__ C,
                    const double *__restrict__ clabel, int k, int32_t *__restrict__ labels, double *__restrict__ mind,
                    double *__restrict__ dist_all, int kd, int *__restrict__ err_flag) {
    __shared__ __align__(16) double xs[AE_KC][AE_LD];
    __shared__ __align__(16) double cs[AE_KC][AE_LD];
    __shared__ int64_t srow[AE_TR];
    const int tid = threadIdx.x;
    const int tr = tid >> 4, tc = tid & 15;
    const int64_t nrows = nrows_dev ? (int64_t)(*nrows_dev) : nrows_arg;   // list length may live on the device
  for (int64_t tile = blockIdx.x; tile * AE_TR < nrows; tile += gridDim.x) {
    const int64_t base = tile * AE_TR;
    __syncthreads();
    if (tid < AE_TR) {
        const int64_t p = base + tid;
        int64_t r = -1;
        if (p < nrows) r = INDIRECT ? (int64_t)rowlist[p] : p;
        srow[tid] = r;
    }
    __syncthreads();

    double bestv[4];
    int besti[4];
    double xl[4];
#pragma unroll
    for (int a = 0; a < 4; ++a) {
        bestv[a] = DBL_MAX;
        besti[a] = INT_MAX;
        const int64_t r = srow[4 * tr + a];
        xl[a] = (METRIC == AKM_EUCLIDEAN && r >= 0) ? xlabel[r] : 0.0;
    }
    const int lcol = tid & 31, lrow0 = tid >> 5;   // loader mapping: warp = one row, 32 consecutive columns

    for (int c0 = 0; c0 < k; c0 += AE_TC) {
        double t[4][4];
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int b = 0; b < 4; ++b) t[a][b] = 0.0;

        for (int i0 = 0; i0 < d; i0 += AE_KC) {
            __syncthreads();
#pragma unroll
            for (int it = 0; it < 8; ++it) {
                const int r = lrow0 + 8 * it;
                const int col = i0 + lcol;
                const int64_t row = srow[r];
                double xv = 0.0, cv = 0.0;
                if (row >= 0 && col < d) {
                    xv = (double)X[row * d + col];
                    if (RawCosine<T, METRIC>::value) xv = DMUL(cos_inv(xlabel[row]), xv);
                }
                const int cj = c0 + r;
                if (cj < k && col < d) cv = C[(size_t)cj * d + col];
                xs[lcol][r] = xv;
                cs[lcol][r] = cv;
            }
            __syncthreads();
            const int lim = min(AE_KC, d - i0);
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
                    for (int b = 0; b < 4; ++b) t[a][b] = DADD(t[a][b], DMUL(cv[b], xv[a]));
            }
        }
        // distances + running first-minimum
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            const int j = c0 + 4 * tc + b;
            if (j < k) {
                const double cl = (METRIC == AKM_EUCLIDEAN) ? clabel[j] : 0.0;
#pragma unroll
                for (int a = 0; a < 4; ++a) {
                    double D;
                    if (METRIC == AKM_EUCLIDEAN) {
                        const double y = DADD(0.0, DMUL(-2.0, t[a][b]));     // y = 0 (beta == 0); y += alpha*temp
                        D = sqrt(fabs(DADD(DADD(y, xl[a]), cl)));
                    } else {
                        D = DADD(1.0, DMUL(-1.0, t[a][b]));                  // y = 1 (beta == 1); y += alpha*temp
                    }
                    if (dist_all != nullptr) {
                        const int64_t p = base + 4 * tr + a;
                        if (p < nrows) dist_all[p * kd + j] = D;
                    }
                    if (D < DBL_MAX && (D < bestv[a] || (D == bestv[a] && j < besti[a]))) {
                        bestv[a] = D;
                        besti[a] = j;
                    }
                }
            }
        }
    }
    // merge the 16 centroid lanes of each row: lowest distance, then lowest index (== first minimum)
#pragma unroll
    for (int a = 0; a < 4; ++a) {
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) {
            const double ov = __shfl_xor_sync(0xffffffffu, bestv[a], o);
            const int oi = __shfl_xor_sync(0xffffffffu, besti[a], o);
            if (oi != INT_MAX && (ov < bestv[a] || (ov == bestv[a] && oi < besti[a]))) {
                bestv[a] = ov;
                besti[a] = oi;
            }
        }
        if (tc == 0) {
            const int64_t p = base + 4 * tr + a;
            const int64_t r = srow[4 * tr + a];
            if (r >= 0) {
                const int64_t o = INDIRECT ? r : p;
                const int lab = (besti[a] == INT_MAX) ? -1 : besti[a];
                labels[o] = lab;
                if (mind != nullptr) mind[o] = (lab < 0) ? nan("") : bestv[a];
                if (lab < 0) atomicExch(err_flag, 1);
            }
        }
    }
  }
}

int launch_assign_exact(akm_handle h, const RowStore &rs, const int32_t *rowlist, int64_t nrows, const double *C,
                        const double *clabel, int k0, int k, int32_t *labels, double *mind, double *dist_all) {
    (void)k0;
    if (nrows <= 0) return AKM_OK;
    const unsigned grid = (unsigned)ceil_div64(nrows, AE_TR);
    int *err = h->iscal + 2;
#define AE_LAUNCH(T, PTR, M, IND)                                                                        \
    assign_exact_kernel<T, M, IND><<<grid, 256, 0, h->stream>>>(PTR, rs.xlabel, rowlist, nrows, nullptr, h->d, C, clabel, k, \
                                                                labels, mind, dist_all, k, err)
    const bool ind = rowlist != nullptr;
    if (h->metric == AKM_EUCLIDEAN) {
        if (rs.store == 4) { if (ind) AE_LAUNCH(float, rs.x32, AKM_EUCLIDEAN, true); else AE_LAUNCH(float, rs.x32, AKM_EUCLIDEAN, false); }
        else               { if (ind) AE_LAUNCH(double, rs.x64, AKM_EUCLIDEAN, true); else AE_LAUNCH(double, rs.x64, AKM_EUCLIDEAN, false); }
    } else {
        if (rs.store == 4) { if (ind) AE_LAUNCH(float, rs.x32, AKM_COSINE, true); else AE_LAUNCH(float, rs.x32, AKM_COSINE, false); }
        else               { if (ind) AE_LAUNCH(double, rs.x64, AKM_COSINE, true); else AE_LAUNCH(double, rs.x64, AKM_COSINE, false); }
    }
#undef AE_LAUNCH
    h->stats.kernel_launches++;
    AKM_CUDA(h, cudaGetLastError());
    return AKM_OK;
}

// Exact assignment of the rows in a device-resident list whose length also lives on the device
// (no host round trip between the tensor-core search and its exact fallback).
int launch_assign_exact_list(akm_handle h, const RowStore &rs, const int32_t *rowlist, const int *count_dev,
                             const double *C, const double *clabel, int k0, int k, int32_t *labels) {
    (void)k0;
    const unsigned grid = (unsigned)(h->num_sms * 4);
    int *err = h->iscal + 2;
#define AEL_LAUNCH(T, PTR, M)                                                                                  \
    assign_exact_kernel<T, M, true><<<grid, 256, 0, h->stream>>>(PTR, rs.xlabel, rowlist, 0, count_dev, h->d, C, clabel, k, \
                                                                 labels, nullptr, nullptr, k, err)
    if (h->metric == AKM_EUCLIDEAN) { if (rs.store == 4) AEL_LAUNCH(float, rs.x32, AKM_EUCLIDEAN); else AEL_LAUNCH(double, rs.x64, AKM_EUCLIDEAN); }
    else { if (rs.store == 4) AEL_LAUNCH(float, rs.x32, AKM_COSINE); else AEL_LAUNCH(double, rs.x64, AKM_COSINE); }
#undef AEL_LAUNCH
    h->stats.kernel_launches++;
    AKM_CUDA(h, cudaGetLastError());
    return AKM_OK;
}

// =====================================================================================
// Per-cluster sums and counts: the daxpy/count half of KMeansUtil.updateSumMatrix (:76-83)
// as a segmented reduction.  One CTA owns a chunk of rows: it counting-sorts the chunk's
// row offsets by label in shared memory, then each warp reduces whole (cluster, chunk)
// segments in registers (lane <-> columns lane + 32*m, coalesced 128-byte row reads, fp64
// accumulation) and flushes one fp64 atomic per (segment, column).
// The objective is accumulated on the way: EUCL sum (x-c)^2 (robust form of D^2),
// COS count - c.S.
// =====================================================================================
constexpr int ACC_THREADS = 256;
constexpr int ACC_MAX_CHUNK = 65536;   // row offsets are uint16

template <typename T> struct Vec16;
template <> struct Vec16<float> { using type = float4; static constexpr int N = 4;
    static __device__ __forceinline__ void unpack(const float4 &v, double (&o)[4]) { o[0] = v.x; o[1] = v.y; o[2] = v.z; o[3] = v.w; } };
template <> struct Vec16<double> { using type = double2; static constexpr int N = 2;
    static __device__ __forceinline__ void unpack(const double2 &v, double (&o)[2]) { o[0] = v.x; o[1] = v.y; } };

// VEC == true: each lane owns 16-byte column groups (lane + 32*m) -> one 128-bit load per row and group
// (requires d % (16/sizeof(T)) == 0); VEC == false: scalar columns lane + 32*m (any d).
// SUB > 1 (narrow rows, DC == 1): a warp reads SUB rows at a time, 32/SUB lanes each.
// After the sort the chunk's rows are cut into items of item_rows consecutive SORTED positions; warps take
// items round-robin and flush at every cluster boundary inside an item, so the work per warp does not depend
// on how many clusters the chunk holds (a cluster-ordered partition has one or two per chunk).
template <typename T, int DC, int METRIC, bool VEC, int SUB>
__global__ void __launch_bounds__(ACC_THREADS, 3)
accumulate_kernel(const T *__restrict__ X, const int32_t *__restrict__ labels, int64_t n, int d, int k0,
                  int chunk_rows, int item_rows, const double *__restrict__ C, double *__restrict__ sum,
                  double *__restrict__ objective, const double *__restrict__ xlabel,
                  const double *__restrict__ clabel) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    int *off = reinterpret_cast<int *>(smem_raw);           // k0 + 1
    int *cursor = off + (k0 + 1);                           // k0
    unsigned short *sorted = reinterpret_cast<unsigned short *>(cursor + k0 + ((k0 & 1) ? 1 : 0));
    __shared__ int scan_tmp[ACC_THREADS];
    __shared__ double obj_red[ACC_THREADS / 32];
    constexpr int VN = VEC ? Vec16<T>::N : 1;               // columns per lane per group
    constexpr int L = 32 / SUB;                             // lanes per row
    constexpr int COLS_PER_PASS = DC * L * VN;
    static_assert(SUB == 1 || DC == 1, "narrow-row mode covers the whole row with one group");

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int sub = lane / L, cl = lane % L;
    const int64_t base = (int64_t)blockIdx.x * chunk_rows;
    const int rows = (int)min((int64_t)chunk_rows, n - base);
    const int cb0 = blockIdx.y * COLS_PER_PASS;

    for (int j = tid; j <= k0; j += ACC_THREADS) off[j] = 0;
    __syncthreads();
    for (int r = tid; r < rows; r += ACC_THREADS) {
        const int lab = labels[base + r];
        if (lab >= 0 && lab < k0) atomicAdd(&off[lab + 1], 1);
    }
    __syncthreads();
    // inclusive scan of off[1..k0] -> off[j] = start of cluster j
    {
        const int per = (k0 + ACC_THREADS - 1) / ACC_THREADS;
        const int lo = min(k0, tid * per), hi = min(k0, lo + per);
        int s = 0;
        for (int j = lo; j < hi; ++j) s += off[j + 1];
        scan_tmp[tid] = s;
        __syncthreads();
        for (int o = 1; o < ACC_THREADS; o <<= 1) {
            int v = (tid >= o) ? scan_tmp[tid - o] : 0;
            __syncthreads();
            scan_tmp[tid] += v;
            __syncthreads();
        }
        int run = (tid > 0) ? scan_tmp[tid - 1] : 0;
        for (int j = lo; j < hi; ++j) {
            const int c = off[j + 1];
            cursor[j] = run;
            run += c;
            off[j + 1] = run;   // safe: off[j+1] of this range is only read by this thread
        }
    }
    __syncthreads();
    for (int r = tid; r < rows; r += ACC_THREADS) {
        const int lab = labels[base + r];
        if (lab >= 0 && lab < k0) {
            const int p = atomicAdd(&cursor[lab], 1);
            sorted[p] = (unsigned short)r;
        }
    }
    __syncthreads();

    double obj = 0.0;
    const int total = off[k0];
    for (int p0 = warp * item_rows; p0 < total; p0 += (ACC_THREADS / 32) * item_rows) {
        const int p1 = min(total, p0 + item_rows);
        int j;
        {   // largest j with off[j] <= p0 (then off[j+1] > p0)
            int lo = 0, hi = k0;
            while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (off[mid] <= p0) lo = mid; else hi = mid; }
            j = lo;
        }
        int p = p0;
        while (p < p1) {
            const int s = p, e = min(off[j + 1], p1);
            double acc[DC][VN];
#pragma unroll
            for (int m = 0; m < DC; ++m)
#pragma unroll
                for (int v = 0; v < VN; ++v) acc[m][v] = 0.0;
            constexpr int U = (DC * VN <= 4) ? 6 : ((DC * VN <= 8) ? 3 : 2);   // row loads in flight per lane (register budget)
            for (; p + U * SUB <= e; p += U * SUB) {
                using RawT = typename std::conditional<VEC, typename Vec16<T>::type, T>::type;
                RawT raw[U][DC];                      // loads stay in their storage type until they are consumed
                double rsc[U];                        // COSINE rows stored raw: 1 / |x| per row
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    const int64_t ridx = base + sorted[p + u * SUB + sub];
                    const T *row = X + ridx * (int64_t)d + cb0;
                    rsc[u] = RawCosine<T, METRIC>::value ? cos_inv(xlabel[ridx]) : 1.0;
#pragma unroll
                    for (int m = 0; m < DC; ++m) {
                        const int c0 = (cl + L * m) * VN;
                        if (cb0 + c0 < d) raw[u][m] = *reinterpret_cast<const RawT *>(row + c0);
                        else raw[u][m] = RawT{};
                    }
                }
#pragma unroll
                for (int u = 0; u < U; ++u)
#pragma unroll
                    for (int m = 0; m < DC; ++m) {
                        double val[VN];
                        if constexpr (VEC) Vec16<T>::unpack(raw[u][m], val);
                        else val[0] = (double)raw[u][m];
#pragma unroll
                        for (int v = 0; v < VN; ++v) acc[m][v] += RawCosine<T, METRIC>::value ? DMUL(rsc[u], val[v]) : val[v];
                    }
            }
            if (p < e) {                              // last, partial batch: same loads, predicated per row
                using RawT = typename std::conditional<VEC, typename Vec16<T>::type, T>::type;
                RawT raw[U][DC];
                double rsc[U];
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    const int q = p + u * SUB + sub;
                    const bool ok = q < e;
                    const int64_t ridx = base + sorted[ok ? q : s];
                    const T *row = X + ridx * (int64_t)d + cb0;
                    rsc[u] = RawCosine<T, METRIC>::value ? cos_inv(xlabel[ridx]) : 1.0;
#pragma unroll
                    for (int m = 0; m < DC; ++m) {
                        const int c0 = (cl + L * m) * VN;
                        if (ok && cb0 + c0 < d) raw[u][m] = *reinterpret_cast<const RawT *>(row + c0);
                        else raw[u][m] = RawT{};
                    }
                }
#pragma unroll
                for (int u = 0; u < U; ++u)
#pragma unroll
                    for (int m = 0; m < DC; ++m) {
                        double val[VN];
                        if constexpr (VEC) Vec16<T>::unpack(raw[u][m], val);
                        else val[0] = (double)raw[u][m];
#pragma unroll
                        for (int v = 0; v < VN; ++v) acc[m][v] += RawCosine<T, METRIC>::value ? DMUL(rsc[u], val[v]) : val[v];
                    }
            }
            p = e;
            if constexpr (SUB > 1) {
#pragma unroll
                for (int o = L; o < 32; o <<= 1)
#pragma unroll
                    for (int v = 0; v < VN; ++v) acc[0][v] += __shfl_xor_sync(0xffffffffu, acc[0][v], o);
            }
            // objective of the piece, the way Java's distance is put together:
            //   EUCL  sum ||x||^2 - 2 c.S + n ||c||^2  (||x||^2 = the stored label, S = the sums just formed)
            //   COS   n - c.S
            if (METRIC == AKM_EUCLIDEAN && blockIdx.y == 0) {
                double xs = 0.0;
                for (int q = s + lane; q < e; q += 32) xs += xlabel[base + sorted[q]];
                obj += xs;
                if (lane == 0) obj += (double)(e - s) * clabel[j];
            }
            double *out = sum + (size_t)j * (d + 1);
            if (sub == 0) {
                double cen[DC][VN];                   // only needed here: keeps the row loop's registers free
#pragma unroll
                for (int m = 0; m < DC; ++m)
#pragma unroll
                    for (int v = 0; v < VN; ++v) {
                        const int col = cb0 + (cl + L * m) * VN + v;
                        cen[m][v] = (col < d && objective != nullptr) ? C[(size_t)j * d + col] : 0.0;
                    }
#pragma unroll
                for (int m = 0; m < DC; ++m)
#pragma unroll
                    for (int v = 0; v < VN; ++v) {
                        const int col = cb0 + (cl + L * m) * VN + v;
                        if (col < d) {
                            atomicAdd(out + col, acc[m][v]);
                            if (METRIC == AKM_EUCLIDEAN) obj -= 2.0 * cen[m][v] * acc[m][v];
                            else obj -= cen[m][v] * acc[m][v];
                        }
                    }
            }
            if (lane == 0 && blockIdx.y == 0) {
                atomicAdd(out + d, (double)(e - s));
                if (METRIC == AKM_COSINE) obj += (double)(e - s);
            }
            if (p < p1) { ++j; while (off[j + 1] <= p) ++j; }     // next non-empty cluster
        }
    }
    if (objective != nullptr) {
        for (int o = 16; o > 0; o >>= 1) obj += __shfl_xor_sync(0xffffffffu, obj, o);
        if (lane == 0) obj_red[warp] = obj;
        __syncthreads();
        if (tid == 0) {
            double a = 0.0;
            for (int w = 0; w < ACC_THREADS / 32; ++w) a += obj_red[w];
            atomicAdd(objective, a);
        }
    }
}

// General fallback (k0 too large for the shared-memory sort): one warp per row, fp64 atomics.
template <typename T, int METRIC>
__global__ void accumulate_atomic_kernel(const T *__restrict__ X, const int32_t *__restrict__ labels, int64_t n, int d,
                                         int k0, const double *__restrict__ C, double *__restrict__ sum,
                                         double *__restrict__ objective, const double *__restrict__ xlabel) {
    const int lane = threadIdx.x & 31;
    int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t stride = ((int64_t)gridDim.x * blockDim.x) >> 5;
    double obj = 0.0;
    for (; row < n; row += stride) {
        const int lab = labels[row];
        if (lab < 0 || lab >= k0) continue;
        double *out = sum + (size_t)lab * (d + 1);
        const double *c = C + (size_t)lab * d;
        for (int col = lane; col < d; col += 32) {
            double v = (double)X[row * d + col];
            if (RawCosine<T, METRIC>::value) v = DMUL(cos_inv(xlabel[row]), v);
            atomicAdd(out + col, v);
            if (METRIC == AKM_EUCLIDEAN) { const double df = v - c[col]; obj = fma(df, df, obj); }
            else obj -= c[col] * v;
        }
        if (lane == 0) { atomicAdd(out + d, 1.0); if (METRIC == AKM_COSINE) obj += 1.0; }
    }
    if (objective != nullptr) {
        for (int o = 16; o > 0; o >>= 1) obj += __shfl_xor_sync(0xffffffffu, obj, o);
        if (lane == 0 && obj != 0.0) atomicAdd(objective, obj);
    }
}

template <typename T, int METRIC>
static int accumulate_dispatch(akm_handle h, const T *X, int64_t n, const int32_t *labels, const double *C, int k0,
                               double *sum, double *objective, const double *xlabel, const double *clabel) {
    const int d = h->d;
    const size_t fixed = (size_t)(2 * k0 + 2) * sizeof(int);
    const size_t budget = h->smem_optin > 8192 ? h->smem_optin - 8192 : 0;
    if (fixed + 2048 * 2 > budget) {
        const int blocks = h->num_sms * 8;
        accumulate_atomic_kernel<T, METRIC><<<blocks, 256, 0, h->stream>>>(X, labels, n, d, k0, C, sum, objective, xlabel);
        h->stats.kernel_launches++;
        AKM_CUDA(h, cudaGetLastError());
        return AKM_OK;
    }
    // chunk of rows per CTA: large enough that one flush per (cluster, chunk) amortises, small enough that three
    // CTAs fit the shared memory of an SM; row offsets inside a chunk are uint16
    static const int env_chunk = getenv("AKM_ACC_CHUNK") ? atoi(getenv("AKM_ACC_CHUNK")) : 0;   // tuning aid
    int64_t chunk = ceil_div64(n, (int64_t)h->num_sms * 6);
    chunk = ((chunk + 1023) / 1024) * 1024;
    const int64_t cap3 = (int64_t)(((size_t)(h->smem_optin + 1024) / 3 - 4096 - fixed) / 2);
    int64_t chunk_max = std::min<int64_t>(32768, std::max<int64_t>(2048, (cap3 / 1024) * 1024));   // = the regroup window
    if (env_chunk > 0) chunk_max = env_chunk;
    if (chunk > chunk_max) chunk = chunk_max;
    if (chunk > ACC_MAX_CHUNK) chunk = ACC_MAX_CHUNK;
    const int64_t max_by_smem = (int64_t)((budget - fixed) / 2);
    if (chunk > max_by_smem) chunk = (max_by_smem / 1024) * 1024;
    if (chunk < 1024) chunk = 1024;
    const size_t smem = fixed + (size_t)chunk * 2;
    constexpr int VN = 16 / (int)sizeof(T);
    const bool vec = (d % VN) == 0;
    const int groups = vec ? (d / VN + 31) / 32 : (d + 31) / 32;       // 32-lane column groups needed
    const int maxdc = vec ? (VN == 4 ? 2 : 4) : 8;                      // register budget: DC*VN <= 8
    const int DCsel = groups >= maxdc ? maxdc : groups;
    // narrow rows: several rows per warp-wide load
    const int subsel = (vec && d / VN <= 8) ? 4 : ((vec && d / VN <= 16) ? 2 : 1);
    const int cols_per_pass = DCsel * (32 / subsel) * (vec ? VN : 1);
    const dim3 grid((unsigned)ceil_div64(n, chunk), (unsigned)((d + cols_per_pass - 1) / cols_per_pass));
    // items: about four per warp, a multiple of the unrolled row step, at least 48 rows
    int item = (int)((chunk + 31) / 32);
    item = std::max(48, ((item + 47) / 48) * 48);
#define ACC_CASE(DCV, VECV, SUBV)                                                                              \
    {                                                                                                          \
        auto kern = accumulate_kernel<T, DCV, METRIC, VECV, SUBV>;                                             \
        AKM_CUDA(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));        \
        kern<<<grid, ACC_THREADS, smem, h->stream>>>(X, labels, n, d, k0, (int)chunk, item, C, sum, objective, xlabel, clabel); \
    }
    if (vec) {
        if (subsel == 4) ACC_CASE(1, true, 4)
        else if (subsel == 2) ACC_CASE(1, true, 2)
        else switch (DCsel) {
            case 1: ACC_CASE(1, true, 1) break;
            case 2: ACC_CASE(2, true, 1) break;
            case 3: ACC_CASE(3, true, 1) break;
            default: ACC_CASE(4, true, 1) break;
        }
    } else {
        switch (DCsel) {
            case 1: ACC_CASE(1, false, 1) break;
            case 2: ACC_CASE(2, false, 1) break;
            case 3: ACC_CASE(3, false, 1) break;
            case 4: ACC_CASE(4, false, 1) break;
            case 5: ACC_CASE(5, false, 1) break;
            case 6: ACC_CASE(6, false, 1) break;
            case 7: ACC_CASE(7, false, 1) break;
            default: ACC_CASE(8, false, 1) break;
        }
    }
#undef ACC_CASE
    h->stats.kernel_launches++;
    AKM_CUDA(h, cudaGetLastError());
    return AKM_OK;
}

int launch_accumulate(akm_handle h, const RowStore &rs, const int32_t *labels, const double *C, const double *clabel,
                      int k0, double *sum, double *objective) {
    AKM_CUDA(h, cudaMemsetAsync(sum, 0, sizeof(double) * (size_t)k0 * (h->d + 1), h->stream));   // Arrays.fill :59
    if (objective) AKM_CUDA(h, cudaMemsetAsync(objective, 0, sizeof(double), h->stream));
    if (rs.n <= 0) return AKM_OK;
    if (h->metric == AKM_EUCLIDEAN) {
        if (rs.store == 4) return accumulate_dispatch<float, AKM_EUCLIDEAN>(h, rs.x32, rs.n, labels, C, k0, sum, objective, rs.xlabel, clabel);
        return accumulate_dispatch<double, AKM_EUCLIDEAN>(h, rs.x64, rs.n, labels, C, k0, sum, objective, rs.xlabel, clabel);
    }
    if (rs.store == 4) return accumulate_dispatch<float, AKM_COSINE>(h, rs.x32, rs.n, labels, C, k0, sum, objective, rs.xlabel, clabel);
    return accumulate_dispatch<double, AKM_COSINE>(h, rs.x64, rs.n, labels, C, k0, sum, objective, rs.xlabel, clabel);
}

// =====================================================================================
// Incremental accumulation.  The reference refills the k x (d+1) accumulator from every sample in every
// superstep (KMeansAssignCluster.java:59-63).  From the third pass of a run on, only a small share of
// the samples changes cluster, so the handle keeps the sums of its partition (sum_local) together with
// the label each row was added under (labels_acc) and moves just the rows whose label changed:
//     S[old] -= x,  S[new] += x,  count[old] -= 1,  count[new] += 1.
// Counts stay exact.  The sums pick up one fp64 rounding per moved value, so the policy in
// accumulate_step() falls back to a full pass when more than 1/16 of the rows moved in the previous pass,
// after 64 incremental passes, and once the rows moved since the last full pass exceed 1/16 of the
// partition (at most n/16 roundings of 2^-53 relative each: many orders below the 1e-6 parity bar;
// clusters that end up empty are reset to exact zeros).
// =====================================================================================
template <typename T, int METRIC>
__global__ void __launch_bounds__(256)
accumulate_delta_kernel(const T *__restrict__ X, const double *__restrict__ xlabel, const int32_t *__restrict__ labels,
                        int32_t *__restrict__ labels_acc, int64_t n, int d, int k0, double *__restrict__ sum,
                        int *__restrict__ changed_out, const int2 *__restrict__ glist, const int *__restrict__ glist_n) {
    // glist: the 256-row groups the search of this pass looked at -- no other row can have changed its label
    const int lane = threadIdx.x & 31;
    const int64_t nblk = glist ? (int64_t)*glist_n * 8 : (n + 31) / 32;
    int64_t it = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t stride = ((int64_t)gridDim.x * blockDim.x) >> 5;
    int moved = 0;
    for (; it < nblk; it += stride) {
        const int64_t blk = glist ? (int64_t)glist[it >> 3].x * 8 + (it & 7) : it;
        const int64_t r = blk * 32 + lane;
        int lb = -1, la = -1;
        if (r < n) { lb = labels[r]; la = labels_acc[r]; }
        if (lb < 0 || lb >= k0) lb = -1;
        if (la < 0 || la >= k0) la = -1;
        unsigned m = __ballot_sync(0xffffffffu, la != lb);
        if (m == 0) continue;
        moved += __popc(m);
        if (la != lb) labels_acc[r] = lb;
        while (m) {
            const int i = __ffs(m) - 1;
            m &= m - 1;
            const int64_t row = blk * 32 + i;
            const int from = __shfl_sync(0xffffffffu, la, i), to = __shfl_sync(0xffffffffu, lb, i);
            const T *x = X + row * (int64_t)d;
            const double rinv = RawCosine<T, METRIC>::value ? cos_inv(xlabel[row]) : 1.0;
            double *sf = sum + (size_t)(from < 0 ? 0 : from) * (d + 1);
            double *st = sum + (size_t)(to < 0 ? 0 : to) * (d + 1);
            for (int c = lane; c < d; c += 32) {
                double v = (double)x[c];
                if (RawCosine<T, METRIC>::value) v = DMUL(rinv, v);
                if (from >= 0) atomicAdd(sf + c, -v);
                if (to >= 0) atomicAdd(st + c, v);
            }
            if (lane == 0) {
                if (from >= 0) atomicAdd(sf + d, -1.0);
                if (to >= 0) atomicAdd(st + d, 1.0);
            }
        }
    }
    if (lane == 0 && moved) atomicAdd(changed_out, moved);
}

// labels_acc <- labels; *changed_out += number of rows whose label differs from the one recorded before
__global__ void labels_sync_kernel(const int32_t *__restrict__ labels, int32_t *__restrict__ labels_acc, int64_t n,
                                   int *__restrict__ changed_out) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    int moved = 0;
    for (; i < n; i += stride) {
        const int lb = labels[i];
        if (labels_acc[i] != lb) { ++moved; labels_acc[i] = lb; }
    }
    for (int o = 16; o > 0; o >>= 1) moved += __shfl_xor_sync(0xffffffffu, moved, o);
    if ((threadIdx.x & 31) == 0 && moved) atomicAdd(changed_out, moved);
}

__global__ void sum_f64_kernel(const double *__restrict__ a, int64_t n, double *__restrict__ out) {
    __shared__ double red[8];
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    double s = 0.0;
    for (; i < n; i += stride) s += a[i];
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += red[w];
        atomicAdd(out, t);
    }
}

// sum_out[0..L) <- sum_local (rows of empty clusters as exact zeros, in both buffers); sum_out[L] += objective of the
// partition put together the way Java's distance is:  EUCL  sum ||x||^2 - 2 c.S + n ||c||^2,  COS  n - c.S
// (sum_out[L] was zeroed by the caller); sum_out[L+1] = the no-finite-distance flag of this pass (all-reduced with
// the sums so that every rank of a multi-GPU step fails together).  One warp per cluster.
__global__ void __launch_bounds__(256)
finish_sums_kernel(int metric, double *__restrict__ sum_local, const double *__restrict__ C,
                   const double *__restrict__ clabel, int k0, int d, const double *__restrict__ xl_total,
                   const int *__restrict__ err_flag, double *__restrict__ sum_out) {
    const int lane = threadIdx.x & 31;
    const int j = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const size_t L = (size_t)k0 * (d + 1);
    if (j == 0 && lane == 0) {
        if (metric == AKM_EUCLIDEAN) atomicAdd(sum_out + L, *xl_total);
        sum_out[L + 1] = (*err_flag != 0) ? 1.0 : 0.0;
    }
    if (j >= k0) return;
    double *row = sum_local + (size_t)j * (d + 1);
    double *out = sum_out + (size_t)j * (d + 1);
    const double cnt = row[d];
    double dot = 0.0;
    if (cnt == 0.0) {
        for (int c = lane; c <= d; c += 32) { row[c] = 0.0; out[c] = 0.0; }
        return;
    }
    for (int c = lane; c < d; c += 32) {
        const double v = row[c];
        out[c] = v;
        dot = fma(C[(size_t)j * d + c], v, dot);
    }
    for (int o = 16; o > 0; o >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, o);
    if (lane == 0) {
        out[d] = cnt;
        const double part = (metric == AKM_EUCLIDEAN) ? (cnt * clabel[j] - 2.0 * dot) : (cnt - dot);
        atomicAdd(sum_out + L, part);
    }
}

// The accumulation half of one superstep over the cached partition and h->labels.  sum_out has k0*(d+1)+2 slots.
int accumulate_step(akm_handle h, const double *C, const double *clabel, int k0, double *sum_out) {
    RowStore &rs = h->rows;
    const int d = h->d;
    const int64_t n = rs.n;
    const size_t L = (size_t)k0 * (size_t)(d + 1);
    AKM_CUDA(h, cudaMemsetAsync(sum_out + L, 0, 2 * sizeof(double), h->stream));
    AKM_CUDA(h, cudaMemsetAsync(h->iscal + 8, 0, sizeof(int), h->stream));
    if (n <= 0) {
        AKM_CUDA(h, cudaMemsetAsync(sum_out, 0, sizeof(double) * L, h->stream));
        return AKM_OK;
    }
    if (h->sum_local_k0 < k0) {
        if (h->sum_local) cudaFree(h->sum_local);
        h->sum_local = nullptr; h->sum_local_k0 = 0;
        AKM_CUDA(h, cudaMalloc((void **)&h->sum_local, sizeof(double) * (L + 2)));
        h->sum_local_k0 = k0;
        h->acc_valid = false;
    }
    if (h->labels_acc_cap < n) {
        if (h->labels_acc) cudaFree(h->labels_acc);
        h->labels_acc = nullptr; h->labels_acc_cap = 0;
        AKM_CUDA(h, cudaMalloc((void **)&h->labels_acc, sizeof(int32_t) * (size_t)std::max<int64_t>(rs.cap, n)));
        h->labels_acc_cap = std::max<int64_t>(rs.cap, n);
        h->acc_valid = false;
    }
    if (!h->xl_total_valid) {
        AKM_CUDA(h, cudaMemsetAsync(h->dscal + 2, 0, sizeof(double), h->stream));
        if (h->metric == AKM_EUCLIDEAN) {
            sum_f64_kernel<<<h->num_sms * 4, 256, 0, h->stream>>>(rs.xlabel, n, h->dscal + 2);
            h->stats.kernel_launches++;
        }
        h->xl_total_valid = true;
    }
    const bool known = h->acc_valid && h->acc_n == n && h->acc_k0 == k0;
    const bool delta = known && h->delta_accumulate && h->last_changed >= 0 && h->last_changed * 16 <= n &&
                       h->acc_delta_passes < 64 && h->acc_changed_total * 16 <= n;
    if (delta) {
        const int blocks = (int)std::min<int64_t>((n + 255) / 256, (int64_t)h->num_sms * 16);
        // the tensor search of this pass says which groups it looked at (consumed here: valid for one pass)
        const int2 *gl = h->searched_valid ? h->searched_list : nullptr;
        const int *gln = h->searched_valid ? h->searched_n : nullptr;
#define DELTA(T, PTR, M) accumulate_delta_kernel<T, M><<<blocks, 256, 0, h->stream>>>(PTR, rs.xlabel, h->labels, h->labels_acc, n, d, k0, h->sum_local, h->iscal + 8, gl, gln)
        if (h->metric == AKM_EUCLIDEAN) { if (rs.store == 4) DELTA(float, rs.x32, AKM_EUCLIDEAN); else DELTA(double, rs.x64, AKM_EUCLIDEAN); }
        else { if (rs.store == 4) DELTA(float, rs.x32, AKM_COSINE); else DELTA(double, rs.x64, AKM_COSINE); }
#undef DELTA
        h->stats.kernel_launches++;
        AKM_CUDA(h, cudaGetLastError());
        h->acc_delta_passes++;
        h->acc_mode_last = 1;
        h->acc_count_pending = true;
    } else {
        // full pass; the label copy doubles as the count of rows that moved since the previous pass
        labels_sync_kernel<<<h->num_sms * 8, 256, 0, h->stream>>>(h->labels, h->labels_acc, n, h->iscal + 8);
        h->stats.kernel_launches++;
        AKM_CUDA(h, cudaGetLastError());
        AKM_TRY(launch_accumulate(h, rs, h->labels, C, clabel, k0, h->sum_local, nullptr));
        h->acc_delta_passes = 0;
        h->acc_changed_total = 0;
        h->acc_mode_last = 0;
        h->acc_count_pending = known;       // without a previous record the count means nothing
        if (!known) h->last_changed = -1;
    }
    h->searched_valid = false;
    finish_sums_kernel<<<(k0 + 7) / 8, 256, 0, h->stream>>>(h->metric, h->sum_local, C, clabel, k0, d, h->dscal + 2,
                                                          h->iscal + 2, sum_out);
    h->stats.kernel_launches++;
    AKM_CUDA(h, cudaGetLastError());
    h->acc_valid = true; h->acc_n = n; h->acc_k0 = k0;
    return AKM_OK;
}

// after the caller copied iscal[8] back (h_iscal[8] = rows moved in the pass accumulate_step just ran)
void accumulate_absorb(akm_handle h) {
    h->stats.accumulate_mode = h->acc_mode_last;
    h->stats.rows_moved = h->acc_count_pending ? (int64_t)h->h_iscal[8] : -1;
    if (!h->acc_count_pending) return;
    h->acc_count_pending = false;
    h->last_changed = h->h_iscal[8];
    if (h->acc_mode_last == 1) h->acc_changed_total += h->h_iscal[8];
}

// KMeansUpdateCentroids drops empty clusters and renumbers the rest (cmap: old index -> new index, -1 = dropped).
// The next pass's labels use the new numbering, so the record behind the incremental sums follows: otherwise
// nearly every row would look "moved" and the next pass would fall back to a full accumulation.
__global__ void remap_labels_kernel(int32_t *__restrict__ labels_acc, int64_t n, const int *__restrict__ cmap, int k_old) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) {
        const int la = labels_acc[i];
        if (la >= 0 && la < k_old) labels_acc[i] = cmap[la];
    }
}
__global__ void remap_sums_kernel(const double *__restrict__ src, double *__restrict__ dst, const int *__restrict__ cmap,
                                  int k_old, int d) {
    const int j = blockIdx.x;
    if (j >= k_old) return;
    const int m = cmap[j];
    if (m < 0) return;
    for (int c = threadIdx.x; c <= d; c += blockDim.x) dst[(size_t)m * (d + 1) + c] = src[(size_t)j * (d + 1) + c];
}
int accumulate_renumber(akm_handle h, const int *cmap, int k_old, int k0) {
    if (!h->acc_valid || !h->sum_local || !h->labels_acc || h->acc_k0 != k0 || h->rows.n <= 0) return AKM_OK;
    const size_t L = (size_t)k0 * (size_t)(h->d + 1);
    if (h->sum_tmp_k0 < k0) {
        if (h->sum_tmp) cudaFree(h->sum_tmp);
        h->sum_tmp = nullptr; h->sum_tmp_k0 = 0;
        AKM_CUDA(h, cudaMalloc((void **)&h->sum_tmp, sizeof(double) * (L + 2)));
        h->sum_tmp_k0 = k0;
    }
    AKM_CUDA(h, cudaMemsetAsync(h->sum_tmp, 0, sizeof(double) * L, h->stream));
    remap_sums_kernel<<<k_old, 128, 0, h->stream>>>(h->sum_local, h->sum_tmp, cmap, k_old, h->d);
    remap_labels_kernel<<<h->num_sms * 8, 256, 0, h->stream>>>(h->labels_acc, h->rows.n, cmap, k_old);
    h->stats.kernel_launches += 2;
    AKM_CUDA(h, cudaGetLastError());
    std::swap(h->sum_local, h->sum_tmp);
    std::swap(h->sum_local_k0, h->sum_tmp_k0);
    return AKM_OK;
}

// if some rank flagged a row without a finite distance, every rank raises the error (sum[L+1] was all-reduced)
__global__ void flag_from_sum_kernel(const double *__restrict__ flag, int *__restrict__ err_flag) {
    if (*flag != 0.0) *err_flag = 1;
}
int launch_flag_from_sum(akm_handle h, const double *flag) {
    flag_from_sum_kernel<<<1, 1, 0, h->stream>>>(flag, h->iscal + 2);
    h->stats.kernel_launches++;
    AKM_CUDA(h, cudaGetLastError());
    return AKM_OK;
}

// =====================================================================================
// KMeansUpdateCentroids.updateCentroids (:52-69) + KMeansIterTermination diagonal (:28-44)
// Single CTA: k0*d is at most a few hundred thousand values.
//   iscal[0] = new k, iscal[1] = converged
// =====================================================================================
constexpr int UPD_THREADS = 1024;

__global__ void __launch_bounds__(UPD_THREADS)
update_kernel(int metric, const double *__restrict__ sum, int k, int k0, int d, double *__restrict__ Cnew,
              double *__restrict__ clnew, const double *__restrict__ Cprev, const double *__restrict__ clprev,
              int new_is_c1, double tol, double *__restrict__ shift, int *__restrict__ cmap, int *__restrict__ iscal) {
    __shared__ int scan_tmp[UPD_THREADS];
    __shared__ int s_newk, s_notconv;
    const int tid = threadIdx.x;
    const int per = (k + UPD_THREADS - 1) / UPD_THREADS;
    const int lo = min(k, tid * per), hi = min(k, lo + per);
    int cnt = 0;
    for (int j = lo; j < hi; ++j) cnt += (sum[(size_t)j * (d + 1) + d] != 0.0) ? 1 : 0;   // "if (weight == 0) continue"
    scan_tmp[tid] = cnt;
    if (tid == 0) s_notconv = 0;
    __syncthreads();
    for (int o = 1; o < UPD_THREADS; o <<= 1) {
        const int v = (tid >= o) ? scan_tmp[tid - o] : 0;
        __syncthreads();
        scan_tmp[tid] += v;
        __syncthreads();
    }
    int run = (tid > 0) ? scan_tmp[tid - 1] : 0;
    for (int j = lo; j < hi; ++j) {
        const bool ne = sum[(size_t)j * (d + 1) + d] != 0.0;
        cmap[j] = ne ? run : -1;
        run += ne ? 1 : 0;
    }
    if (tid == UPD_THREADS - 1) s_newk = scan_tmp[UPD_THREADS - 1];
    // Arrays.fill(data, 0.0)
    for (size_t i = tid; i < (size_t)k0 * d; i += UPD_THREADS) Cnew[i] = 0.0;
    __syncthreads();
    // daxpy(d, 1.0/weight, buffer, start, data, index*d):  y = 0 + (1/w)*x
    for (size_t e = tid; e < (size_t)k * d; e += UPD_THREADS) {
        const int j = (int)(e / d), i = (int)(e % d);
        const int m = cmap[j];
        if (m >= 0) {
            const double w = sum[(size_t)j * (d + 1) + d];
            const double a = 1.0 / w;
            Cnew[(size_t)m * d + i] = DADD(0.0, DMUL(a, sum[(size_t)j * (d + 1) + i]));
        }
    }
    __syncthreads();
    // distance.updateLabel over ALL k0 columns
    for (int j = tid; j < k0; j += UPD_THREADS) centroid_label_row(metric, Cnew + (size_t)j * d, d, clnew + j);
    __syncthreads();
    const int newk = s_newk;
    if (Cprev != nullptr) {
        // left = CENTROID1, right = CENTROID2 (KMeansIterTermination.java:33):
        // EUCL sqrt(|(-2*dot(right,left) + label_right) + label_left|), COS (-1*dot) + 1*1
        const double *c1 = new_is_c1 ? Cnew : Cprev, *l1 = new_is_c1 ? clnew : clprev;
        const double *c2 = new_is_c1 ? Cprev : Cnew, *l2 = new_is_c1 ? clprev : clnew;
        int notconv = 0;
        for (int j = tid; j < newk; j += UPD_THREADS) {
            const double *a = c2 + (size_t)j * d, *b = c1 + (size_t)j * d;
            double t = 0.0;
            for (int l = 0; l < d; ++l) t = DADD(t, DMUL(a[l], b[l]));
            double v;
            if (metric == AKM_EUCLIDEAN) {
                v = DMUL(-2.0, t);
                v = sqrt(fabs(DADD(DADD(v, l2[j]), l1[j])));
            } else {
                v = DADD(DMUL(-1.0, t), DMUL(1.0, 1.0));
            }
            if (shift) shift[j] = v;
            if (v >= tol) notconv = 1;
        }
        if (notconv) atomicExch(&s_notconv, 1);
        __syncthreads();
    }
    if (tid == 0) {
        iscal[0] = newk;
        iscal[1] = (Cprev != nullptr && s_notconv == 0) ? 1 : 0;
    }
}

// ---- multi-CTA version: a one-block scan builds the compaction maps, then one warp per centroid row does
// the scaling, updateLabel and the shift distance.  Java's index-order sums are kept: the row is staged in
// shared memory and every lane walks it in order (a 128-long dependent chain per row instead of a 128-long
// chain of uncoalesced global loads per thread, which is what made the single-CTA kernel take 0.3 ms).
__global__ void __launch_bounds__(UPD_THREADS)
update_map_kernel(const double *__restrict__ sum, int k, int k0, int d, int *__restrict__ cmap, int *__restrict__ inv,
                  int *__restrict__ iscal) {
    __shared__ int scan_tmp[UPD_THREADS];
    const int tid = threadIdx.x;
    const int per = (k + UPD_THREADS - 1) / UPD_THREADS;
    const int lo = min(k, tid * per), hi = min(k, lo + per);
    int cnt = 0;
    for (int j = lo; j < hi; ++j) cnt += (sum[(size_t)j * (d + 1) + d] != 0.0) ? 1 : 0;
    scan_tmp[tid] = cnt;
    for (int m = tid; m < k0; m += UPD_THREADS) inv[m] = -1;
    __syncthreads();
    for (int o = 1; o < UPD_THREADS; o <<= 1) {
        const int v = (tid >= o) ? scan_tmp[tid - o] : 0;
        __syncthreads();
        scan_tmp[tid] += v;
        __syncthreads();
    }
    int run = (tid > 0) ? scan_tmp[tid - 1] : 0;
    for (int j = lo; j < hi; ++j) {
        const bool ne = sum[(size_t)j * (d + 1) + d] != 0.0;
        cmap[j] = ne ? run : -1;
        if (ne) inv[run] = j;
        run += ne ? 1 : 0;
    }
    if (tid == UPD_THREADS - 1) { iscal[0] = scan_tmp[UPD_THREADS - 1]; iscal[1] = 1; }   // new k; "converged" until a row objects
}

__global__ void __launch_bounds__(256)
update_rows_kernel(int metric, const double *__restrict__ sum, int k0, int d, double *__restrict__ Cnew,
                   double *__restrict__ clnew, const double *__restrict__ Cprev, const double *__restrict__ clprev,
                   int new_is_c1, double tol, double *__restrict__ shift, const int *__restrict__ inv,
                   int *__restrict__ iscal) {
    extern __shared__ double upd_smem[];                 // [8 warps][2][d]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int m = blockIdx.x * 8 + warp;                 // destination row
    if (m >= k0) return;
    double *rowbuf = upd_smem + (size_t)warp * 2 * d;
    double *prevbuf = rowbuf + d;
    const int newk = iscal[0];
    const int src = inv[m];
    // Arrays.fill(data, 0) then daxpy(d, 1.0/weight, buffer, start, data, index*d): y = 0 + (1/w)*x
    double a = 0.0;
    if (src >= 0) a = 1.0 / sum[(size_t)src * (d + 1) + d];
    for (int c = lane; c < d; c += 32) {
        double v = 0.0;
        if (src >= 0) v = DADD(0.0, DMUL(a, sum[(size_t)src * (d + 1) + c]));
        rowbuf[c] = v;
        if (Cprev != nullptr) prevbuf[c] = Cprev[(size_t)m * d + c];
    }
    __syncwarp();
    // distance.updateLabel (all k0 rows): sum of squares in index order
    double s = 0.0;
    for (int i = 0; i < d; ++i) { const double v = rowbuf[i]; s = DADD(s, DMUL(v, v)); }
    double label = s;
    if (metric == AKM_COSINE) {
        label = 0.0;
        const double nrm = sqrt(s);
        if (nrm > 0.0) {
            const double invn = 1.0 / nrm;
            __syncwarp();
            for (int c = lane; c < d; c += 32) rowbuf[c] = DMUL(invn, rowbuf[c]);
            __syncwarp();
        }
    }
    for (int c = lane; c < d; c += 32) Cnew[(size_t)m * d + c] = rowbuf[c];
    if (lane == 0) clnew[m] = label;
    // KMeansIterTermination: left = CENTROID1, right = CENTROID2
    if (Cprev != nullptr && m < newk) {
        const double *c1 = new_is_c1 ? rowbuf : prevbuf, *c2 = new_is_c1 ? prevbuf : rowbuf;
        const double l1 = new_is_c1 ? label : clprev[m], l2 = new_is_c1 ? clprev[m] : label;
        double t = 0.0;
        for (int l = 0; l < d; ++l) t = DADD(t, DMUL(c2[l], c1[l]));
        double v;
        if (metric == AKM_EUCLIDEAN) {
            v = DMUL(-2.0, t);
            v = sqrt(fabs(DADD(DADD(v, l2), l1)));
        } else {
            v = DADD(DMUL(-1.0, t), DMUL(1.0, 1.0));
        }
        if (lane == 0) {
            if (shift) shift[m] = v;
            if (v >= tol) iscal[1] = 0;                  // not converged (benign race: everybody writes 0)
        }
    }
}

int launch_update(akm_handle h, const double *sum, int k, int k0, double *Cnew, double *clnew, const double *Cprev,
                  const double *clprev, bool new_is_c1, double tol, double *shift, int *iscal, int *cmap, int *inv) {
    const size_t smem = (size_t)8 * 2 * h->d * sizeof(double);
    if (smem > 96 * 1024 || inv == nullptr) {
        update_kernel<<<1, UPD_THREADS, 0, h->stream>>>(h->metric, sum, k, k0, h->d, Cnew, clnew, Cprev, clprev,
                                                        new_is_c1 ? 1 : 0, tol, shift, cmap, iscal);
    } else {
        update_map_kernel<<<1, UPD_THREADS, 0, h->stream>>>(sum, k, k0, h->d, cmap, inv, iscal);
        if (smem > 48 * 1024)
            AKM_CUDA(h, cudaFuncSetAttribute(update_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        update_rows_kernel<<<(k0 + 7) / 8, 256, smem, h->stream>>>(h->metric, sum, k0, h->d, Cnew, clnew, Cprev, clprev,
                                                                  new_is_c1 ? 1 : 0, tol, shift, inv, iscal);
        h->stats.kernel_launches++;
    }
    h->stats.kernel_launches++;
    AKM_CUDA(h, cudaGetLastError());
    return AKM_OK;
}

// =====================================================================================
// KMeansUtil.getProbArrayFromDistanceArray (:262-269): sequential sum (StatUtils.sum),
// ratio = 1.0 / sum / (k-1), probs = 1/(k-1) + (-ratio)*d_i
// =====================================================================================
__global__ void probs_kernel(const double *__restrict__ dist_all, int64_t n, int k, double *__restrict__ probs) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n) return;
    const double *dr = dist_all + r * k;
    double s = 0.0;
    for (int i = 0; i < k; ++i) s = DADD(s, dr[i]);
    const double ratio = (1.0 / s) / (double)(k - 1);
    const double base = 1.0 / (double)(k - 1);
    double *pr = probs + r * k;
    for (int i = 0; i < k; ++i) pr[i] = DADD(base, DMUL(-ratio, dr[i]));
}

int launch_probs(akm_handle h, const double *dist_all, int64_t n, int k, double *probs) {
    if (n <= 0) return AKM_OK;
    probs_kernel<<<(unsigned)ceil_div64(n, 128), 128, 0, h->stream>>>(dist_all, n, k, probs);
    h->stats.kernel_launches++;
    AKM_CUDA(h, cudaGetLastError());
    return AKM_OK;
}

__global__ void labels_to_i64_kernel(const int32_t *__restrict__ a, int64_t n, int64_t *__restrict__ b) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) b[i] = (int64_t)a[i];
}

int launch_labels_to_i64(akm_handle h, const int32_t *labels, int64_t n, int64_t *out) {
    if (n <= 0) return AKM_OK;
    labels_to_i64_kernel<<<h->num_sms * 4, 256, 0, h->stream>>>(labels, n, out);
    h->stats.kernel_launches++;
    AKM_CUDA(h, cudaGetLastError());
    return AKM_OK;
}
