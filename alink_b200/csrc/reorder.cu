// Cluster-ordered partitions.
//
// The reference never reorders samples, and nothing it computes depends on their order except the order of
// the fp64 additions (which the GPU path does not reproduce bit for bit anyway).  Keeping the
// rows of a partition grouped by their current cluster inside waves of a few million rows makes the hot
// kernels coherent: the 32 rows a warp of the search epilogue owns share their nearest centroids (the rare
// candidate path stops diverging), neighbouring worklist entries gather the same centroid rows, and the
// accumulate kernel reads long contiguous runs.  Every per-row array moves together (exact copy, fp16 copy,
// norms, labels, seeds); `orig` remembers the caller's row index so outputs are returned in input order.
#include <algorithm>

#include "akm_internal.cuh"

constexpr int REORDER_CHUNK = 32768;
constexpr int REORDER_THREADS = 512;

// Counting sort of the rows of one WAVE (as many 32K-row chunks as the scratch buffer holds, ~8M rows) by label:
//   reorder_hist_kernel    one CTA per chunk: bucket counts of the chunk            -> hist[chunk][bucket]
//   reorder_scan_kernel    one CTA per wave: exclusive offsets of every (bucket, chunk) pair inside the wave
//   reorder_scatter_kernel one CTA per chunk: dest[row] = new global row index
// A cluster's rows end up contiguous inside each wave (thousands of rows per run at the bench shape), in chunk
// order; the order inside a (chunk, bucket) cell is whatever the atomics produce.
__global__ void __launch_bounds__(REORDER_THREADS)
reorder_hist_kernel(const int32_t *__restrict__ labels, int64_t n, int k, int *__restrict__ hist) {
    extern __shared__ int rk_smem[];
    const int tid = threadIdx.x;
    const int64_t base = (int64_t)blockIdx.x * REORDER_CHUNK;
    const int rows = (int)min((int64_t)REORDER_CHUNK, n - base);
    const int nb = k + 1;                                   // bucket k = rows without a valid label
    for (int j = tid; j < nb; j += REORDER_THREADS) rk_smem[j] = 0;
    __syncthreads();
    for (int r = tid; r < rows; r += REORDER_THREADS) {
        int lab = labels[base + r];
        if (lab < 0 || lab >= k) lab = k;
        atomicAdd(&rk_smem[lab], 1);
    }
    __syncthreads();
    for (int j = tid; j < nb; j += REORDER_THREADS) hist[(size_t)blockIdx.x * nb + j] = rk_smem[j];
}

__global__ void __launch_bounds__(1024)
reorder_scan_kernel(int *__restrict__ hist, int nchunks, int wave_chunks, int k, int *__restrict__ totals) {
    __shared__ int scan_tmp[1024];
    const int tid = threadIdx.x;
    const int nb = k + 1;
    const int c0 = blockIdx.x * wave_chunks, c1 = min(nchunks, c0 + wave_chunks);
    int *tot = totals + (size_t)blockIdx.x * nb;
    for (int b = tid; b < nb; b += 1024) {                  // per bucket: running offset over the wave's chunks
        int run = 0;
        for (int c = c0; c < c1; ++c) { const int t = hist[(size_t)c * nb + b]; hist[(size_t)c * nb + b] = run; run += t; }
        tot[b] = run;
    }
    __syncthreads();
    const int per = (nb + 1023) / 1024;                      // exclusive scan of the bucket totals
    const int lo = min(nb, tid * per), hi = min(nb, lo + per);
    int s = 0;
    for (int j = lo; j < hi; ++j) s += tot[j];
    scan_tmp[tid] = s;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {
        const int v = (tid >= o) ? scan_tmp[tid - o] : 0;
        __syncthreads();
        scan_tmp[tid] += v;
        __syncthreads();
    }
    int run = (tid > 0) ? scan_tmp[tid - 1] : 0;
    for (int j = lo; j < hi; ++j) { const int c = tot[j]; tot[j] = run; run += c; }
    __syncthreads();
    for (int b = tid; b < nb; b += 1024) {
        const int off = tot[b];
        for (int c = c0; c < c1; ++c) hist[(size_t)c * nb + b] += off;
    }
}

__global__ void __launch_bounds__(REORDER_THREADS)
reorder_scatter_kernel(const int32_t *__restrict__ labels, int64_t n, int k, const int *__restrict__ hist,
                       int wave_chunks, int32_t *__restrict__ dest) {
    extern __shared__ int rk_smem[];
    const int tid = threadIdx.x;
    const int64_t base = (int64_t)blockIdx.x * REORDER_CHUNK;
    const int rows = (int)min((int64_t)REORDER_CHUNK, n - base);
    const int nb = k + 1;
    const int64_t wave_row0 = (int64_t)(blockIdx.x / wave_chunks) * wave_chunks * REORDER_CHUNK;
    for (int j = tid; j < nb; j += REORDER_THREADS) rk_smem[j] = hist[(size_t)blockIdx.x * nb + j];
    __syncthreads();
    for (int r = tid; r < rows; r += REORDER_THREADS) {
        int lab = labels[base + r];
        if (lab < 0 || lab >= k) lab = k;
        dest[base + r] = (int32_t)(wave_row0 + atomicAdd(&rk_smem[lab], 1));
    }
}

// scratch[(dest[r] - row0)] = src[r] for rows r in [row0, row0 + rows); one warp per row, 16-byte or 4-byte units
template <typename U>
__global__ void permute_rows_kernel(const U *__restrict__ src, U *__restrict__ scratch, const int32_t *__restrict__ dest,
                                    int64_t row0, int64_t rows, int units_per_row) {
    const int lane = threadIdx.x & 31;
    int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t stride = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (; w < rows; w += stride) {
        const int64_t r = row0 + w;
        const int64_t t = (int64_t)dest[r] - row0;
        const U *s = src + r * units_per_row;
        U *d = scratch + t * units_per_row;
        for (int u = lane; u < units_per_row; u += 32) d[u] = s[u];
    }
}

__global__ void iota_kernel(int32_t *a, int64_t from, int64_t n) {
    int64_t i = from + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) a[i] = (int32_t)i;
}

// out[orig[i]] = in[i]
template <typename V>
__global__ void unpermute_kernel(const V *__restrict__ in, const int32_t *__restrict__ orig, int64_t n, V *__restrict__ out) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) out[orig[i]] = in[i];
}

static int permute_array(akm_handle h, void *arr, size_t row_bytes, const int32_t *dest, int64_t row0, int64_t rows,
                         void *scratch) {
    if (!arr || row_bytes == 0) return AKM_OK;
    const int blocks = h->num_sms * 16;
    char *base = (char *)arr;
    if (row_bytes % 16 == 0 && ((uintptr_t)arr % 16) == 0)
        permute_rows_kernel<int4><<<blocks, 256, 0, h->stream>>>((const int4 *)arr, (int4 *)scratch, dest, row0, rows, (int)(row_bytes / 16));
    else
        permute_rows_kernel<int><<<blocks, 256, 0, h->stream>>>((const int *)arr, (int *)scratch, dest, row0, rows, (int)(row_bytes / 4));
    h->stats.kernel_launches++;
    AKM_CUDA(h, cudaGetLastError());
    AKM_CUDA(h, cudaMemcpyAsync(base + (size_t)row0 * row_bytes, scratch, (size_t)rows * row_bytes, cudaMemcpyDeviceToDevice, h->stream));
    return AKM_OK;
}

int ensure_orig(akm_handle h, RowStore &rs) {
    if (rs.orig_cap < rs.cap || !rs.orig) {
        int32_t *no = nullptr;
        AKM_CUDA(h, cudaMalloc((void **)&no, sizeof(int32_t) * (size_t)std::max<int64_t>(rs.cap, 1)));
        if (rs.orig && rs.orig_n > 0)
            AKM_CUDA(h, cudaMemcpyAsync(no, rs.orig, sizeof(int32_t) * (size_t)rs.orig_n, cudaMemcpyDeviceToDevice, h->stream));
        AKM_CUDA(h, cudaStreamSynchronize(h->stream));
        if (rs.orig) cudaFree(rs.orig);
        rs.orig = no;
        rs.orig_cap = rs.cap;
    }
    if (rs.orig_n < rs.n) {   // rows appended since the last reorder keep their own index
        iota_kernel<<<h->num_sms * 4, 256, 0, h->stream>>>(rs.orig, rs.orig_n, rs.n);
        h->stats.kernel_launches++;
        AKM_CUDA(h, cudaGetLastError());
        rs.orig_n = rs.n;
    }
    return AKM_OK;
}

// Scratch of the regrouping for a partition of n rows: one wave of rows (as many 32K-row chunks as fit 4 GB) and the
// destination index of every row.  akm_reserve_rows calls this up front so that no cudaMalloc (which stalls the
// device) lands inside the first supersteps of a training run.
int reorder_reserve(akm_handle h, const RowStore &rs, int64_t n) {
    if (n <= 0 || n >= ((int64_t)1 << 31)) return AKM_OK;
    const size_t xbytes = (size_t)h->d * (rs.store == 4 ? 4 : 8);
    const size_t max_row = std::max(xbytes, std::max<size_t>((size_t)((h->d + 7) / 8) * 16, 8));   // rows, fp16 copy, side arrays
    const size_t budget = (size_t)4 << 30;
    const int64_t nchunks = ceil_div64(n, REORDER_CHUNK);
    const int64_t wave_chunks = std::min<int64_t>(std::max<int64_t>(1, (int64_t)(budget / (max_row * REORDER_CHUNK))), nchunks);
    const size_t need = (size_t)wave_chunks * REORDER_CHUNK * max_row;
    if (h->reorder_scratch_bytes < need) {
        if (h->reorder_scratch) cudaFree(h->reorder_scratch);
        h->reorder_scratch = nullptr; h->reorder_scratch_bytes = 0;
        AKM_CUDA(h, cudaMalloc(&h->reorder_scratch, need));
        h->reorder_scratch_bytes = need;
    }
    if (h->reorder_dest_cap < n) {
        if (h->reorder_dest) cudaFree(h->reorder_dest);
        h->reorder_dest = nullptr; h->reorder_dest_cap = 0;
        AKM_CUDA(h, cudaMalloc((void **)&h->reorder_dest, sizeof(int32_t) * (size_t)n));
        h->reorder_dest_cap = n;
    }
    return AKM_OK;
}

// Regroup the rows of `rs` by `labels` inside chunks.  labels / useed are permuted with the rows.
int reorder_rows_by_label(akm_handle h, RowStore &rs, int32_t *labels, float *useed, int k) {
    const int64_t n = rs.n;
    if (n <= 0 || n >= ((int64_t)1 << 31)) return AKM_OK;
    if ((size_t)(k + 3) * sizeof(int) > 160 * 1024) return AKM_OK;       // bucket counters live in shared memory
    AKM_TRY(ensure_orig(h, rs));
    const int d = h->d;
    const size_t xbytes = (size_t)d * (rs.store == 4 ? 4 : 8);
    const size_t x16bytes = (rs.x16 && rs.x16_valid && rs.x16_cap >= n) ? (size_t)rs.d16 * 2 : 0;
    const size_t max_row = std::max(xbytes, std::max<size_t>((size_t)((d + 7) / 8) * 16, 8));      // as in reorder_reserve
    AKM_TRY(reorder_reserve(h, rs, n));
    const size_t budget = (size_t)4 << 30;
    const int64_t nchunks = ceil_div64(n, REORDER_CHUNK);
    const int64_t wave_chunks = std::min<int64_t>(std::max<int64_t>(1, (int64_t)(budget / (max_row * REORDER_CHUNK))), nchunks);
    const size_t smem = (size_t)(k + 3) * sizeof(int);
    const int64_t nwaves = ceil_div64(nchunks, wave_chunks);
    const size_t hist_bytes = ((size_t)nchunks + (size_t)nwaves) * (size_t)(k + 1) * sizeof(int);
    if (h->reorder_hist_bytes < hist_bytes) {
        if (h->reorder_hist) cudaFree(h->reorder_hist);
        h->reorder_hist = nullptr; h->reorder_hist_bytes = 0;
        AKM_CUDA(h, cudaMalloc((void **)&h->reorder_hist, hist_bytes));
        h->reorder_hist_bytes = hist_bytes;
    }
    int *hist = h->reorder_hist, *totals = hist + (size_t)nchunks * (k + 1);
    AKM_CUDA(h, cudaFuncSetAttribute(reorder_hist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    AKM_CUDA(h, cudaFuncSetAttribute(reorder_scatter_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    reorder_hist_kernel<<<(unsigned)nchunks, REORDER_THREADS, smem, h->stream>>>(labels, n, k, hist);
    reorder_scan_kernel<<<(unsigned)nwaves, 1024, 0, h->stream>>>(hist, (int)nchunks, (int)wave_chunks, k, totals);
    reorder_scatter_kernel<<<(unsigned)nchunks, REORDER_THREADS, smem, h->stream>>>(labels, n, k, hist, (int)wave_chunks,
                                                                                  h->reorder_dest);
    h->stats.kernel_launches += 3;
    AKM_CUDA(h, cudaGetLastError());
    for (int64_t c0 = 0; c0 < nchunks; c0 += wave_chunks) {
        const int64_t row0 = c0 * REORDER_CHUNK;
        const int64_t rows = std::min<int64_t>(n - row0, wave_chunks * REORDER_CHUNK);
        void *sc = h->reorder_scratch;
        const int32_t *dest = h->reorder_dest;
        AKM_TRY(permute_array(h, rs.store == 4 ? (void *)rs.x32 : (void *)rs.x64, xbytes, dest, row0, rows, sc));
        AKM_TRY(permute_array(h, rs.xlabel, 8, dest, row0, rows, sc));
        if (x16bytes) {
            AKM_TRY(permute_array(h, rs.x16, x16bytes, dest, row0, rows, sc));
            AKM_TRY(permute_array(h, rs.xerr, 8, dest, row0, rows, sc));
        }
        AKM_TRY(permute_array(h, rs.orig, 4, dest, row0, rows, sc));
        if (useed) AKM_TRY(permute_array(h, useed, 4, dest, row0, rows, sc));
        AKM_TRY(permute_array(h, labels, 4, dest, row0, rows, sc));   // last: the ranks were computed from it
    }
    rs.reordered = true;
    return AKM_OK;
}

// Device array in the partition's current row order -> host array in the caller's row order.
template <typename V>
static int copy_out_impl(akm_handle h, const RowStore &rs, const V *dev, V *host, int64_t n) {
    if (n <= 0) return AKM_OK;
    const size_t need = sizeof(V) * (size_t)n;
    if (!rs.reordered || !rs.orig) return download_pipelined(h, dev, host, need);
    if (h->reorder_scratch_bytes < need) {
        if (h->reorder_scratch) cudaFree(h->reorder_scratch);
        h->reorder_scratch = nullptr; h->reorder_scratch_bytes = 0;
        AKM_CUDA(h, cudaMalloc(&h->reorder_scratch, need));
        h->reorder_scratch_bytes = need;
    }
    unpermute_kernel<V><<<h->num_sms * 8, 256, 0, h->stream>>>(dev, rs.orig, n, (V *)h->reorder_scratch);
    h->stats.kernel_launches++;
    AKM_CUDA(h, cudaGetLastError());
    return download_pipelined(h, h->reorder_scratch, host, need);
}

int copy_rows_out_i32(akm_handle h, const RowStore &rs, const int32_t *dev, int32_t *host, int64_t n) {
    return copy_out_impl<int32_t>(h, rs, dev, host, n);
}
int copy_rows_out_i64(akm_handle h, const RowStore &rs, const int64_t *dev, int64_t *host, int64_t n) {
    return copy_out_impl<int64_t>(h, rs, dev, host, n);
}
int copy_rows_out_f64(akm_handle h, const RowStore &rs, const double *dev, double *host, int64_t n) {
    return copy_out_impl<double>(h, rs, dev, host, n);
}
