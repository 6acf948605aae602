// Cluster-ordered partitions.
//
// The reference never reorders samples, and nothing it computes depends on their order except the order of
// the fp64 additions (which the GPU path does not reproduce bit for bit anyway).  Keeping the
// rows of a partition grouped by their current cluster inside chunks of REORDER_CHUNK rows makes the hot
// kernels coherent: the 32 rows a warp of the search epilogue owns share their nearest centroids (the rare
// candidate path stops diverging), neighbouring worklist entries gather the same centroid rows, and the
// accumulate kernel reads long contiguous runs.  Every per-row array moves together (exact copy, fp16 copy,
// norms, labels, seeds); `orig` remembers the caller's row index so outputs are returned in input order.
#include <algorithm>

#include "akm_internal.cuh"

constexpr int REORDER_CHUNK = 32768;
constexpr int REORDER_THREADS = 512;

// One CTA per chunk: counting sort of the chunk's rows by label -> dest[row] = new global row index.
__global__ void __launch_bounds__(REORDER_THREADS)
reorder_rank_kernel(const int32_t *__restrict__ labels, int64_t n, int k, int32_t *__restrict__ dest) {
    extern __shared__ int rk_smem[];
    int *cursor = rk_smem;   // k + 2 buckets (k = invalid labels)
    __shared__ int scan_tmp[REORDER_THREADS];
    const int tid = threadIdx.x;
    const int64_t base = (int64_t)blockIdx.x * REORDER_CHUNK;
    const int rows = (int)min((int64_t)REORDER_CHUNK, n - base);
    const int nb = k + 1;
    for (int j = tid; j <= nb; j += REORDER_THREADS) cursor[j] = 0;
    __syncthreads();
    for (int r = tid; r < rows; r += REORDER_THREADS) {
        int lab = labels[base + r];
        if (lab < 0 || lab >= k) lab = k;
        atomicAdd(&cursor[lab], 1);
    }
    __syncthreads();
    const int per = (nb + REORDER_THREADS - 1) / REORDER_THREADS;
    const int lo = min(nb, tid * per), hi = min(nb, lo + per);
    int s = 0;
    for (int j = lo; j < hi; ++j) s += cursor[j];
    scan_tmp[tid] = s;
    __syncthreads();
    for (int o = 1; o < REORDER_THREADS; o <<= 1) {
        const int v = (tid >= o) ? scan_tmp[tid - o] : 0;
        __syncthreads();
        scan_tmp[tid] += v;
        __syncthreads();
    }
    int run = (tid > 0) ? scan_tmp[tid - 1] : 0;
    for (int j = lo; j < hi; ++j) { const int c = cursor[j]; cursor[j] = run; run += c; }
    __syncthreads();
    for (int r = tid; r < rows; r += REORDER_THREADS) {
        int lab = labels[base + r];
        if (lab < 0 || lab >= k) lab = k;
        dest[base + r] = (int32_t)(base + atomicAdd(&cursor[lab], 1));
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

// Regroup the rows of `rs` by `labels` inside chunks.  labels / useed are permuted with the rows.
int reorder_rows_by_label(akm_handle h, RowStore &rs, int32_t *labels, float *useed, int k) {
    const int64_t n = rs.n;
    if (n <= 0 || n >= ((int64_t)1 << 31)) return AKM_OK;
    if ((size_t)(k + 3) * sizeof(int) > 160 * 1024) return AKM_OK;       // bucket counters live in shared memory
    AKM_TRY(ensure_orig(h, rs));
    const int d = h->d;
    const size_t xbytes = (size_t)d * (rs.store == 4 ? 4 : 8);
    const size_t x16bytes = (rs.x16 && rs.x16_valid && rs.x16_cap >= n) ? (size_t)rs.d16 * 2 : 0;
    const size_t max_row = std::max(xbytes, std::max<size_t>(x16bytes, 8));
    // wave = as many chunks as fit a bounded scratch buffer
    const size_t budget = (size_t)4 << 30;
    int64_t wave_chunks = std::max<int64_t>(1, (int64_t)(budget / (max_row * REORDER_CHUNK)));
    const int64_t nchunks = ceil_div64(n, REORDER_CHUNK);
    wave_chunks = std::min(wave_chunks, nchunks);
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
    const size_t smem = (size_t)(k + 3) * sizeof(int);
    AKM_CUDA(h, cudaFuncSetAttribute(reorder_rank_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    reorder_rank_kernel<<<(unsigned)nchunks, REORDER_THREADS, smem, h->stream>>>(labels, n, k, h->reorder_dest);
    h->stats.kernel_launches++;
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
    if (!rs.reordered || !rs.orig) {
        AKM_CUDA(h, cudaMemcpyAsync(host, dev, sizeof(V) * (size_t)n, cudaMemcpyDeviceToHost, h->stream));
        return AKM_OK;
    }
    const size_t need = sizeof(V) * (size_t)n;
    if (h->reorder_scratch_bytes < need) {
        if (h->reorder_scratch) cudaFree(h->reorder_scratch);
        h->reorder_scratch = nullptr; h->reorder_scratch_bytes = 0;
        AKM_CUDA(h, cudaMalloc(&h->reorder_scratch, need));
        h->reorder_scratch_bytes = need;
    }
    unpermute_kernel<V><<<h->num_sms * 8, 256, 0, h->stream>>>(dev, rs.orig, n, (V *)h->reorder_scratch);
    h->stats.kernel_launches++;
    AKM_CUDA(h, cudaGetLastError());
    AKM_CUDA(h, cudaMemcpyAsync(host, h->reorder_scratch, need, cudaMemcpyDeviceToHost, h->stream));
    return AKM_OK;
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
