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

// A wave = the rows regrouped together (a cluster's rows end up contiguous inside each wave): as many 32K-row
// chunks as make ~4 GB of the widest per-row array.
static int64_t wave_chunks_for(int d, int store, int64_t nchunks) {
    const size_t xbytes = (size_t)d * (size_t)store;
    const size_t max_row = std::max(xbytes, std::max<size_t>((size_t)((d + 7) / 8) * 16, 8));   // rows, fp16 copy, side arrays
    const size_t budget = (size_t)4 << 30;
    return std::min<int64_t>(std::max<int64_t>(1, (int64_t)(budget / (max_row * REORDER_CHUNK))), std::max<int64_t>(nchunks, 1));
}

int64_t reorder_head_rows(akm_handle h, const RowStore &rs, int store, int64_t cap) {
    if (&rs != &h->rows || h->engine == AKM_ENGINE_EXACT || cap < h->reorder_min_rows || cap >= ((int64_t)1 << 31)) return 0;
    return wave_chunks_for(h->d, store, ceil_div64(cap, REORDER_CHUNK)) * REORDER_CHUNK;
}

// Scratch of the regrouping for a partition of n rows: one wave of the narrow per-row arrays (8 bytes per row; the
// wide arrays move inside their own allocations, see RowStore) or, for a store without headroom, of the widest
// array; plus the destination index of every row.  akm_reserve_rows calls this up front so that no cudaMalloc
// (which stalls the device) lands inside the first supersteps of a training run.
int reorder_reserve(akm_handle h, const RowStore &rs, int64_t n) {
    if (n <= 0 || n >= ((int64_t)1 << 31)) return AKM_OK;
    const int64_t nchunks = ceil_div64(n, REORDER_CHUNK);
    const int64_t wave_rows = wave_chunks_for(h->d, rs.store, nchunks) * REORDER_CHUNK;
    const bool shifting = rs.x_head >= wave_rows && (!rs.x16_alloc || rs.x16_head >= wave_rows);
    const size_t xbytes = (size_t)h->d * (size_t)rs.store;
    const size_t max_row = shifting ? 32 : std::max(xbytes, std::max<size_t>((size_t)((h->d + 7) / 8) * 16, 32));   // 28 B of narrow arrays
    const size_t need = (size_t)wave_rows * max_row;
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

// dst[dest[r]] = src[r] for rows r in [row0, row0 + rows) (src and dst are the data pointers before / after the move).
// A warp takes four rows at a time so that four row loads (and their destination indices) are in flight per lane:
// the kernel is a scattered copy and lives on memory-level parallelism.
template <typename U>
__global__ void __launch_bounds__(256)
move_rows_kernel(const U *__restrict__ src, U *__restrict__ dst, const int32_t *__restrict__ dest,
                 int64_t row0, int64_t rows, int units_per_row) {
    const int lane = threadIdx.x & 31;
    int64_t w = (((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5) * 4;
    const int64_t stride = (((int64_t)gridDim.x * blockDim.x) >> 5) * 4;
    for (; w < rows; w += stride) {
        const int cnt = (int)min((int64_t)4, rows - w);
        int64_t t[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) t[i] = (i < cnt) ? (int64_t)dest[row0 + w + i] : 0;
        for (int u = lane; u < units_per_row; u += 32) {
            U v[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) if (i < cnt) v[i] = src[(row0 + w + i) * units_per_row + u];
#pragma unroll
            for (int i = 0; i < 4; ++i) if (i < cnt) dst[t[i] * units_per_row + u] = v[i];
        }
    }
}

// The narrow per-row arrays of one wave in one launch, one thread per row: scatter into the scratch (laid out as
// five consecutive arrays of `rows` entries), then copy back.  (A warp per 4- or 8-byte row would leave 30 lanes idle.)
__global__ void __launch_bounds__(256)
move_narrow_kernel(const int32_t *__restrict__ dest, int64_t row0, int64_t rows, const double *__restrict__ xlabel,
                   const float2 *__restrict__ xerr, const int32_t *__restrict__ orig, const float *__restrict__ useed,
                   const int32_t *__restrict__ labels, char *__restrict__ scratch) {
    double *s_xl = reinterpret_cast<double *>(scratch);
    float2 *s_xe = reinterpret_cast<float2 *>(scratch + (size_t)rows * 8);
    int32_t *s_or = reinterpret_cast<int32_t *>(scratch + (size_t)rows * 16);
    float *s_us = reinterpret_cast<float *>(scratch + (size_t)rows * 20);
    int32_t *s_lb = reinterpret_cast<int32_t *>(scratch + (size_t)rows * 24);
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < rows; i += stride) {
        const int64_t r = row0 + i;
        const int64_t t = (int64_t)dest[r] - row0;
        s_xl[t] = xlabel[r];
        if (xerr) s_xe[t] = xerr[r];
        s_or[t] = orig[r];
        if (useed) s_us[t] = useed[r];
        s_lb[t] = labels[r];
    }
}
__global__ void __launch_bounds__(256)
copy_narrow_back_kernel(int64_t row0, int64_t rows, double *__restrict__ xlabel, float2 *__restrict__ xerr,
                        int32_t *__restrict__ orig, float *__restrict__ useed, int32_t *__restrict__ labels,
                        const char *__restrict__ scratch) {
    const double *s_xl = reinterpret_cast<const double *>(scratch);
    const float2 *s_xe = reinterpret_cast<const float2 *>(scratch + (size_t)rows * 8);
    const int32_t *s_or = reinterpret_cast<const int32_t *>(scratch + (size_t)rows * 16);
    const float *s_us = reinterpret_cast<const float *>(scratch + (size_t)rows * 20);
    const int32_t *s_lb = reinterpret_cast<const int32_t *>(scratch + (size_t)rows * 24);
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < rows; i += stride) {
        const int64_t r = row0 + i;
        xlabel[r] = s_xl[i];
        if (xerr) xerr[r] = s_xe[i];
        orig[r] = s_or[i];
        if (useed) useed[r] = s_us[i];
        labels[r] = s_lb[i];
    }
}

// One wide array (rows of row_bytes, data at row offset `off` of an allocation with `head` spare rows): every wave
// moves into the space next to it -- towards the front when the data sits behind the headroom (waves in ascending
// order: wave i lands where wave i-1 was), towards the back otherwise (descending order).  Returns the new offset.
static int shift_array(akm_handle h, char *alloc, size_t row_bytes, int64_t off, int64_t wave_rows, int64_t n,
                       const int32_t *dest, int64_t *new_off) {
    const bool down = off > 0;                               // off >= wave_rows was checked by the caller
    const int64_t noff = down ? 0 : wave_rows;
    const char *src = alloc + (size_t)off * row_bytes;
    char *dst = alloc + (size_t)noff * row_bytes;
    const int64_t nwaves = ceil_div64(n, wave_rows);
    const int blocks = h->num_sms * 8;
    for (int64_t i = 0; i < nwaves; ++i) {
        const int64_t w = down ? i : nwaves - 1 - i;
        const int64_t row0 = w * wave_rows, rows = std::min<int64_t>(wave_rows, n - row0);
        if (row_bytes % 16 == 0)
            move_rows_kernel<int4><<<blocks, 256, 0, h->stream>>>((const int4 *)src, (int4 *)dst, dest, row0, rows, (int)(row_bytes / 16));
        else
            move_rows_kernel<int><<<blocks, 256, 0, h->stream>>>((const int *)src, (int *)dst, dest, row0, rows, (int)(row_bytes / 4));
        h->stats.kernel_launches++;
    }
    AKM_CUDA(h, cudaGetLastError());
    *new_off = noff;
    return AKM_OK;
}

// Regroup the rows of `rs` by `labels` inside waves.  labels / useed are permuted with the rows.
int reorder_rows_by_label(akm_handle h, RowStore &rs, int32_t *labels, float *useed, int k) {
    const int64_t n = rs.n;
    if (n <= 0 || n >= ((int64_t)1 << 31)) return AKM_OK;
    if ((size_t)(k + 3) * sizeof(int) > 160 * 1024) return AKM_OK;       // bucket counters live in shared memory
    AKM_TRY(ensure_orig(h, rs));
    const int d = h->d;
    const size_t xbytes = (size_t)d * (rs.store == 4 ? 4 : 8);
    const bool have16 = rs.x16 && rs.x16_valid && rs.x16_cap >= n;
    const size_t x16bytes = have16 ? (size_t)rs.d16 * 2 : 0;
    AKM_TRY(reorder_reserve(h, rs, n));
    const int64_t nchunks = ceil_div64(n, REORDER_CHUNK);
    const int64_t wave_chunks = wave_chunks_for(d, rs.store, nchunks);
    const int64_t wave_rows = wave_chunks * REORDER_CHUNK;
    // the wide arrays shift inside their allocations when both have the headroom and sit at offset 0 or >= one wave
    const bool shift_x = rs.x_alloc && rs.x_head >= wave_rows && (rs.x_off == 0 || rs.x_off >= wave_rows) && xbytes % 4 == 0;
    const bool shift_16 = !have16 || (rs.x16_alloc && rs.x16_head >= wave_rows && (rs.x16_off == 0 || rs.x16_off >= wave_rows));
    const bool shifting = shift_x && shift_16;
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
    const int32_t *dest = h->reorder_dest;
    if (shifting) {
        int64_t noff = 0;
        AKM_TRY(shift_array(h, (char *)rs.x_alloc, xbytes, rs.x_off, wave_rows, n, dest, &noff));
        rs.x_off = noff;
        if (rs.store == 4) rs.x32 = (float *)rs.x_alloc + (size_t)noff * d; else rs.x64 = (double *)rs.x_alloc + (size_t)noff * d;
        if (have16) {
            AKM_TRY(shift_array(h, (char *)rs.x16_alloc, x16bytes, rs.x16_off, wave_rows, n, dest, &noff));
            rs.x16_off = noff;
            rs.x16 = rs.x16_alloc + (size_t)noff * rs.d16;
        }
    }
    for (int64_t c0 = 0; c0 < nchunks; c0 += wave_chunks) {
        const int64_t row0 = c0 * REORDER_CHUNK;
        const int64_t rows = std::min<int64_t>(n - row0, wave_rows);
        void *sc = h->reorder_scratch;
        if (!shifting) {
            AKM_TRY(permute_array(h, rs.store == 4 ? (void *)rs.x32 : (void *)rs.x64, xbytes, dest, row0, rows, sc));
            if (x16bytes) AKM_TRY(permute_array(h, rs.x16, x16bytes, dest, row0, rows, sc));
        }
        // rows % 4 == 0 for every wave but the last, whose scratch sub-arrays only need 4-byte alignment after the
        // two 8-byte ones: pad the layout to the wave size instead of the row count
        move_narrow_kernel<<<h->num_sms * 8, 256, 0, h->stream>>>(dest, row0, rows, rs.xlabel, x16bytes ? rs.xerr : nullptr,
                                                                  rs.orig, useed, labels, (char *)sc);
        copy_narrow_back_kernel<<<h->num_sms * 8, 256, 0, h->stream>>>(row0, rows, rs.xlabel, x16bytes ? rs.xerr : nullptr,
                                                                       rs.orig, useed, labels, (const char *)sc);
        h->stats.kernel_launches += 2;
        AKM_CUDA(h, cudaGetLastError());
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
