// C ABI of libakm (include/akm.h): handle management and the host-side sequencing of the
// per-partition work Alink's ComQueue stages perform (KMeansAssignCluster, AllReduce,
// KMeansUpdateCentroids, KMeansIterTermination, KMeansOutputModel, KMeansModelMapper.map).
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <new>
#include <thread>

#include "akm_internal.cuh"

static thread_local std::string g_tls_error;

void akm_set_tls_error(const std::string &s) { g_tls_error = s; }

int akm_fail(akm_handle h, int code, const std::string &msg) {
    if (h) h->err = msg;
    g_tls_error = msg;
    return code;
}

#define REQUIRE(h, cond, msg) \
    do { if (!(cond)) return akm_fail((h), AKM_ERR_BADARG, (msg)); } while (0)

static int upload_resolve(akm_handle h);
static int bind_device(akm_handle h, bool uploading = false) {
    if (!h) return akm_fail(nullptr, AKM_ERR_BADARG, "null handle");
    AKM_CUDA(h, cudaSetDevice(h->device));
    // anything but a further upload wants the rows as they will finally be stored
    if (!uploading && (h->pend[0].active || h->pend[1].active)) return upload_resolve(h);
    return AKM_OK;
}

template <typename T>
static int dev_alloc(akm_handle h, T **p, size_t count) {
    *p = nullptr;
    if (count == 0) count = 1;
    AKM_CUDA(h, cudaMalloc((void **)p, count * sizeof(T)));
    return AKM_OK;
}

template <typename T>
static void dev_free(T *&p) {
    if (p) cudaFree(p);
    p = nullptr;
}

static void rowstore_free(RowStore &rs) {
    if (rs.x_alloc) cudaFree(rs.x_alloc);
    if (rs.x16_alloc) cudaFree(rs.x16_alloc);
    rs.x_alloc = nullptr; rs.x16_alloc = nullptr; rs.x32 = nullptr; rs.x64 = nullptr; rs.x16 = nullptr;
    rs.x_head = rs.x_off = rs.x16_head = rs.x16_off = 0;
    dev_free(rs.xlabel); dev_free(rs.xerr); dev_free(rs.orig);
    rs.n = rs.cap = 0;
    rs.x16_valid = false;
    rs.x16_cap = 0; rs.orig_cap = rs.orig_n = 0; rs.reordered = false;
}

// allocate the exact copy for `cap` rows of `store` bytes per value (+ regrouping headroom), data at offset 0
static int rowstore_alloc_x(akm_handle h, RowStore &rs, int store, int64_t cap, void **out, int64_t *head_out) {
    const int64_t head = reorder_head_rows(h, rs, store, cap);
    *out = nullptr;
    AKM_CUDA(h, cudaMalloc(out, std::max<size_t>(1, (size_t)(cap + head) * (size_t)h->d * (size_t)store)));
    *head_out = head;
    return AKM_OK;
}

// Grow (or create) the exact store so that `need` rows fit; keeps existing rows.
static int rowstore_reserve(akm_handle h, RowStore &rs, int64_t need) {
    if (need <= rs.cap) return AKM_OK;
    int64_t cap = rs.cap == 0 ? need : std::max<int64_t>(need, rs.cap + rs.cap / 2);
    const size_t d = (size_t)h->d;
    double *nl = nullptr;
    AKM_TRY(dev_alloc(h, &nl, (size_t)cap));
    if (rs.n > 0) AKM_CUDA(h, cudaMemcpyAsync(nl, rs.xlabel, sizeof(double) * (size_t)rs.n, cudaMemcpyDeviceToDevice, h->stream));
    void *nx = nullptr;
    int64_t head = 0;
    AKM_TRY(rowstore_alloc_x(h, rs, rs.store, cap, &nx, &head));
    const void *old = rs.store == 4 ? (const void *)rs.x32 : (const void *)rs.x64;
    if (rs.n > 0) AKM_CUDA(h, cudaMemcpyAsync(nx, old, (size_t)rs.store * (size_t)rs.n * d, cudaMemcpyDeviceToDevice, h->stream));
    AKM_CUDA(h, cudaStreamSynchronize(h->stream));
    if (rs.x_alloc) cudaFree(rs.x_alloc);
    rs.x_alloc = nx; rs.x_head = head; rs.x_off = 0;
    if (rs.store == 4) { rs.x32 = (float *)nx; rs.x64 = nullptr; } else { rs.x64 = (double *)nx; rs.x32 = nullptr; }
    dev_free(rs.xlabel);
    rs.xlabel = nl;
    rs.cap = cap;
    rs.x16_valid = false;
    return AKM_OK;
}

// fp32 -> fp64 storage migration (lossless: every stored value was fp32-exact; COSINE rows get normalised here)
static int rowstore_widen(akm_handle h, RowStore &rs) {
    void *nx = nullptr;
    int64_t head = 0;
    AKM_TRY(rowstore_alloc_x(h, rs, 8, rs.cap, &nx, &head));
    AKM_TRY(launch_widen_rows(h, rs.x32, rs.xlabel, (double *)nx, rs.n));
    AKM_CUDA(h, cudaStreamSynchronize(h->stream));
    if (rs.x_alloc) cudaFree(rs.x_alloc);
    rs.x_alloc = nx; rs.x_head = head; rs.x_off = 0;
    rs.x32 = nullptr;
    rs.x64 = (double *)nx;
    rs.store = 8;
    rs.x16_valid = false;
    return AKM_OK;
}

// Upload pipeline: the caller's (pageable) rows are copied into one half of a pinned buffer by a few host threads
// while the other half is on its way to the device and through prepareVectorData, so the host copy, the DMA and
// the kernel overlap.  (A cudaMemcpyAsync straight from pageable memory is staged by the driver at ~11 GB/s and
// serialises with everything else.)  The copy threads narrow a chunk of doubles to fp32 when every value of the
// chunk survives the round trip -- the usual case for feature vectors -- so it crosses PCIe at 4 bytes per value;
// a chunk with any other value is sent as doubles and switches the partition to fp64 storage.
constexpr size_t UPLOAD_HALF_BYTES = (size_t)64 << 20;
constexpr int UPLOAD_MAX_THREADS = 32;

static int upload_threads() {
    static const int n = [] {
        if (const char *e = getenv("AKM_UPLOAD_THREADS")) return std::max(1, std::min(UPLOAD_MAX_THREADS, atoi(e)));
        const unsigned hc = std::thread::hardware_concurrency();
        return (int)std::max(2u, std::min(16u, hc));          // measured on a 16-vCPU host: 4 / 8 / 16 / 32 threads -> 18.5 / 19.2 / 24.0 / 23.6 GB/s
    }();
    return n;
}

static int ensure_dev_stage(akm_handle h, size_t bytes) {
    if (h->stage_bytes >= bytes) return AKM_OK;
    if (h->stage) cudaFree(h->stage);
    if (h->csr_dev) cudaFree(h->csr_dev);
    h->stage = nullptr; h->stage_bytes = 0;
    AKM_CUDA(h, cudaMalloc(&h->stage, bytes));
    h->stage_bytes = bytes;
    return AKM_OK;
}

static int ensure_upload_stage(akm_handle h, size_t half_bytes) {
    AKM_TRY(ensure_dev_stage(h, 2 * half_bytes));
    if (h->hstage_bytes < 2 * half_bytes) {
        AKM_CUDA(h, cudaStreamSynchronize(h->stream));  // a chunk of an earlier upload may still be in flight
        h->stage_used[0] = h->stage_used[1] = false;
        if (h->hstage) cudaFreeHost(h->hstage);       // (the pin arena belongs to pin_begin / akm_destroy alone)
        h->hstage = nullptr; h->hstage_bytes = 0;
        AKM_CUDA(h, cudaHostAlloc(&h->hstage, 2 * half_bytes, cudaHostAllocDefault));
        h->hstage_bytes = 2 * half_bytes;
    }
    for (int i = 0; i < 2; ++i)
        if (!h->stage_free[i]) AKM_CUDA(h, cudaEventCreateWithFlags(&h->stage_free[i], cudaEventDisableTiming));
    return AKM_OK;
}

template <typename F>
static void run_parts(size_t count, size_t align, size_t min_parallel, F &&fn) {     // fn(thread, begin, end)
    const int T = upload_threads();
    if (count < min_parallel || T == 1) { fn(0, (size_t)0, count); return; }
    std::thread workers[UPLOAD_MAX_THREADS - 1];
    const size_t part = ((count / T) + align - 1) / align * align;
    for (int t = 1; t < T; ++t) {
        const size_t lo = std::min(count, part * t), hi = std::min(count, lo + part);
        workers[t - 1] = std::thread([&fn, t, lo, hi]() { if (hi > lo) fn(t, lo, hi); });
    }
    fn(0, (size_t)0, std::min(count, part));
    for (int t = 1; t < T; ++t) workers[t - 1].join();
}

static void host_copy_parallel(void *dst, const void *src, size_t bytes) {
    run_parts(bytes, 4096, (size_t)4 << 20, [=](int, size_t lo, size_t hi) { memcpy((char *)dst + lo, (const char *)src + lo, hi - lo); });
}

// dst[i] = (float)src[i]; false when some value does not survive the round trip (NaN included): the caller then
// sends the chunk as doubles
static bool host_narrow_parallel(float *dst, const double *src, size_t count) {
    bool ok[UPLOAD_MAX_THREADS];
    for (bool &o : ok) o = true;
    run_parts(count, 1024, (size_t)1 << 19, [&ok, dst, src](int t, size_t lo, size_t hi) {
        bool good = true;
        for (size_t i0 = lo; i0 < hi && good; i0 += 65536) {       // give up early on the first bad block
            const size_t i1 = std::min(hi, i0 + 65536);
            int bad = 0;
            for (size_t i = i0; i < i1; ++i) {
                const float f = (float)src[i];
                dst[i] = f;
                bad |= ((double)f != src[i]);
            }
            good = bad == 0;
        }
        ok[t] = good;
    });
    for (bool o : ok) if (!o) return false;
    return true;
}

// Device -> pageable host memory through the same pinned halves: the DMA of chunk i+1 overlaps the host copy of
// chunk i.  Returns with the data in `host` (the stream is synchronised).
int download_pipelined(akm_handle h, const void *dev, void *host, size_t bytes) {
    if (bytes == 0) return AKM_OK;
    if (bytes < ((size_t)8 << 20)) {
        AKM_CUDA(h, cudaMemcpyAsync(host, dev, bytes, cudaMemcpyDeviceToHost, h->stream));
        AKM_CUDA(h, cudaStreamSynchronize(h->stream));
        return AKM_OK;
    }
    const size_t half = (size_t)32 << 20;
    AKM_TRY(ensure_upload_stage(h, half));
    const size_t hh = h->hstage_bytes / 2;
    size_t prev_off = 0, prev_len = 0;
    int slot = 0;
    for (size_t off = 0; off < bytes; off += half, slot ^= 1) {
        const size_t len = std::min(half, bytes - off);
        char *hbuf = (char *)h->hstage + (size_t)slot * hh;
        AKM_CUDA(h, cudaMemcpyAsync(hbuf, (const char *)dev + off, len, cudaMemcpyDeviceToHost, h->stream));
        AKM_CUDA(h, cudaEventRecord(h->stage_free[slot], h->stream));
        if (prev_len) {
            AKM_CUDA(h, cudaEventSynchronize(h->stage_free[slot ^ 1]));
            host_copy_parallel((char *)host + prev_off, (char *)h->hstage + (size_t)(slot ^ 1) * hh, prev_len);
        }
        prev_off = off; prev_len = len;
    }
    slot ^= 1;                                            // the slot of the last chunk
    AKM_CUDA(h, cudaEventSynchronize(h->stage_free[slot]));
    host_copy_parallel((char *)host + prev_off, (char *)h->hstage + (size_t)slot * hh, prev_len);
    return AKM_OK;
}

// ---- small host buffers of the per-superstep calls go through a pinned arena: every copy is truly asynchronous
// and the call synchronises once; results reach the caller's memory in pin_finish().
static int pin_begin(akm_handle h, size_t total_bytes) {
    h->pin_off = 0;
    h->pin_out.clear();
    total_bytes += 1024;
    if (h->pin_bytes >= total_bytes) return AKM_OK;
    if (h->pin) cudaFreeHost(h->pin);
    h->pin = nullptr; h->pin_bytes = 0;
    AKM_CUDA(h, cudaHostAlloc((void **)&h->pin, total_bytes, cudaHostAllocDefault));
    h->pin_bytes = total_bytes;
    return AKM_OK;
}
static char *pin_take(akm_handle h, size_t bytes) {
    const size_t off = (h->pin_off + 63) & ~(size_t)63;
    if (off + bytes > h->pin_bytes) return nullptr;
    h->pin_off = off + bytes;
    return h->pin + off;
}
// host buffers that are page-locked already (akm_host_alloc, or registered by the caller) need no bounce copy
static bool host_is_pinned(const void *p) {
    cudaPointerAttributes pa{};
    if (cudaPointerGetAttributes(&pa, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return pa.type == cudaMemoryTypeHost;
}
static int pin_h2d(akm_handle h, void *dev, const void *host, size_t bytes) {
    char *b = host_is_pinned(host) ? nullptr : pin_take(h, bytes);
    if (!b) { AKM_CUDA(h, cudaMemcpyAsync(dev, host, bytes, cudaMemcpyHostToDevice, h->stream)); return AKM_OK; }
    memcpy(b, host, bytes);
    AKM_CUDA(h, cudaMemcpyAsync(dev, b, bytes, cudaMemcpyHostToDevice, h->stream));
    return AKM_OK;
}
static int pin_d2h(akm_handle h, void *host, const void *dev, size_t bytes) {
    char *b = host_is_pinned(host) ? nullptr : pin_take(h, bytes);
    if (!b) { AKM_CUDA(h, cudaMemcpyAsync(host, dev, bytes, cudaMemcpyDeviceToHost, h->stream)); return AKM_OK; }
    AKM_CUDA(h, cudaMemcpyAsync(b, dev, bytes, cudaMemcpyDeviceToHost, h->stream));
    h->pin_out.push_back({host, b, bytes});
    return AKM_OK;
}
// after the stream has been synchronised
static void pin_finish(akm_handle h) {
    for (const auto &o : h->pin_out) memcpy(o.dst, o.src, o.bytes);
    h->pin_out.clear();
}

enum SrcKind { SRC_HOST_F64, SRC_HOST_F32, SRC_DEV_F32 };

// Chunks of doubles that went from page-locked caller memory straight into fp32 storage are checked after the fact
// (the device flags values that are not fp32 values).  Their source doubles are still in the device staging halves,
// so a misfit is repaired without the caller's buffer: the partition is widened and the chunk prepared again.
static int upload_resolve(akm_handle h) {
    if (!h->pend[0].active && !h->pend[1].active) return AKM_OK;
    AKM_CUDA(h, cudaStreamSynchronize(h->stream));
    const bool bad0 = h->pend[0].active && h->h_iscal[10] != 0, bad1 = h->pend[1].active && h->h_iscal[11] != 0;
    if ((bad0 || bad1) && h->rows.store == 4) {
        const size_t dhalf = h->stage_bytes / 2;
        const int64_t n_keep = h->rows.n;
        AKM_TRY(rowstore_widen(h, h->rows));
        for (int slot = 0; slot < 2; ++slot) {
            if (!((slot == 0) ? bad0 : bad1)) continue;
            AKM_TRY(launch_prep_rows(h, h->rows, (char *)h->stage + (size_t)slot * dhalf, false, h->pend[slot].row0,
                                     h->pend[slot].rows, h->pend[slot].prepared, h->iscal + 10 + slot));
        }
        AKM_CUDA(h, cudaStreamSynchronize(h->stream));
        h->rows.n = n_keep;
    }
    h->pend[0].active = h->pend[1].active = false;
    return AKM_OK;
}

// Appends the rows; returns when the last chunk is staged (copies and kernels still in flight on h->stream).
static int upload_rows(akm_handle h, RowStore &rs, const void *src, SrcKind kind, int64_t n, bool prepared = false) {
    if (n == 0) return AKM_OK;
    // COSINE rows the caller has already normalised are kept as the doubles they are (an fp32 store would hold raw
    // rows and scale them wherever they are read)
    if (prepared && h->metric == AKM_COSINE && rs.store == 4) AKM_TRY(rowstore_widen(h, rs));
    AKM_TRY(rowstore_reserve(h, rs, rs.n + n));
    if (&rs == &h->rows) { h->xl_total_valid = false; h->acc_valid = false; h->last_changed = -1; }
    rs.x16_valid = false;
    const size_t d = (size_t)h->d;
    if (kind == SRC_DEV_F32) {
        AKM_TRY(launch_prep_rows(h, rs, src, true, rs.n, n, prepared));
        rs.n += n;
    } else {
        const size_t elt = kind == SRC_HOST_F64 ? 8 : 4;
        const size_t row_bytes = d * elt;
        int64_t chunk = (int64_t)std::max<size_t>(1, UPLOAD_HALF_BYTES / row_bytes);
        chunk = std::min<int64_t>(chunk, n);
        const size_t half = (size_t)chunk * row_bytes;
        if (h->stage_bytes < 2 * half) AKM_TRY(upload_resolve(h));      // the halves are about to be reallocated
        AKM_TRY(ensure_upload_stage(h, half));
        const size_t hhalf = h->hstage_bytes / 2, dhalf = h->stage_bytes / 2;
        // a source in page-locked memory (akm_host_alloc, or registered by the caller) is read by the DMA engine
        // directly: no staging copy, no narrowing -- PCIe, not the host's memory bandwidth, sets the rate
        cudaPointerAttributes pa{};
        const bool pinned_src = &rs == &h->rows && cudaPointerGetAttributes(&pa, src) == cudaSuccess && pa.type == cudaMemoryTypeHost;
        if (!pinned_src) cudaGetLastError();
        for (int64_t off = 0; pinned_src && off < n; off += chunk) {
            const int slot = h->stage_slot;
            h->stage_slot ^= 1;
            const int64_t m = std::min<int64_t>(chunk, n - off);
            char *dbuf = (char *)h->stage + (size_t)slot * dhalf;
            // the chunk that used this half two chunks ago: done by now -- was every value of it an fp32 value?
            if (h->pend[slot].active) {
                AKM_CUDA(h, cudaEventSynchronize(h->stage_free[slot]));
                if (h->h_iscal[10 + slot] != 0) AKM_TRY(upload_resolve(h));
                h->pend[slot].active = false;
            }
            const bool as_f32 = kind == SRC_HOST_F32;
            const bool check = !as_f32 && rs.store == 4;     // doubles into fp32 storage: the device reports misfits
            if (check) AKM_CUDA(h, cudaMemsetAsync(h->iscal + 10 + slot, 0, sizeof(int), h->stream));
            AKM_CUDA(h, cudaMemcpyAsync(dbuf, (const char *)src + (size_t)off * row_bytes, (size_t)m * row_bytes, cudaMemcpyHostToDevice, h->stream));
            AKM_TRY(launch_prep_rows(h, rs, dbuf, as_f32, rs.n, m, prepared, h->iscal + 10 + slot));
            if (check) {
                AKM_CUDA(h, cudaMemcpyAsync(h->h_iscal + 10 + slot, h->iscal + 10 + slot, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
                h->pend[slot] = {true, prepared, rs.n, m};
            }
            AKM_CUDA(h, cudaEventRecord(h->stage_free[slot], h->stream));
            h->stage_used[slot] = true;
            rs.n += m;
        }
        for (int64_t off = 0; !pinned_src && off < n; off += chunk) {
            const int slot = h->stage_slot;
            h->stage_slot ^= 1;
            const int64_t m = std::min<int64_t>(chunk, n - off);
            char *hbuf = (char *)h->hstage + (size_t)slot * hhalf;
            char *dbuf = (char *)h->stage + (size_t)slot * dhalf;
            if (h->pend[slot].active) AKM_TRY(upload_resolve(h));
            if (h->stage_used[slot]) AKM_CUDA(h, cudaEventSynchronize(h->stage_free[slot]));   // DMA + kernel of two chunks ago are done
            const char *sp = (const char *)src + (size_t)off * row_bytes;
            bool as_f32 = kind == SRC_HOST_F32;
            if (kind == SRC_HOST_F64 && host_narrow_parallel((float *)hbuf, (const double *)sp, (size_t)m * d)) as_f32 = true;
            else host_copy_parallel(hbuf, sp, (size_t)m * row_bytes);
            if (!as_f32 && rs.store == 4) AKM_TRY(rowstore_widen(h, rs));      // first value that is not an fp32 value
            AKM_CUDA(h, cudaMemcpyAsync(dbuf, hbuf, (size_t)m * d * (as_f32 ? 4 : 8), cudaMemcpyHostToDevice, h->stream));
            AKM_TRY(launch_prep_rows(h, rs, dbuf, as_f32, rs.n, m, prepared));
            AKM_CUDA(h, cudaEventRecord(h->stage_free[slot], h->stream));
            h->stage_used[slot] = true;
            rs.n += m;
        }
    }
    if (rs.reordered) AKM_TRY(ensure_orig(h, rs));   // appended rows keep their own index
    return AKM_OK;
}

static int ensure_labels(akm_handle h, int64_t n) {
    if (h->labels_cap >= n) return AKM_OK;
    dev_free(h->labels);
    AKM_TRY(dev_alloc(h, &h->labels, (size_t)n));
    h->labels_cap = n;
    return AKM_OK;
}

static int ensure_mind(akm_handle h, int64_t n) {
    if (h->mind_cap >= n) return AKM_OK;
    dev_free(h->mind);
    AKM_TRY(dev_alloc(h, &h->mind, (size_t)n));
    h->mind_cap = n;
    return AKM_OK;
}

// cent[0] = CENTROID1, cent[1] = CENTROID2 (the ping-pong pair of the device-resident loop)
static int ensure_centroids(akm_handle h, int k0) {
    if (h->cent_cap >= k0) return AKM_OK;
    // preserve nothing: akm_set_centroids re-uploads after growth
    for (int i = 0; i < 2; ++i) { dev_free(h->cent[i]); dev_free(h->clabel[i]); }
    dev_free(h->sum);
    const size_t kd = (size_t)k0 * (size_t)h->d;
    for (int i = 0; i < 2; ++i) {
        AKM_TRY(dev_alloc(h, &h->cent[i], kd));
        AKM_TRY(dev_alloc(h, &h->clabel[i], (size_t)k0));
    }
    AKM_TRY(dev_alloc(h, &h->sum, (size_t)k0 * (size_t)(h->d + 1) + 2));
    h->cent_cap = k0;
    return AKM_OK;
}

static int ensure_kvec(akm_handle h, int k0) {
    if (h->kvec_cap >= k0) return AKM_OK;
    dev_free(h->cmap); dev_free(h->shift); dev_free(h->inv);
    h->inv_cap = 0;
    AKM_TRY(dev_alloc(h, &h->cmap, (size_t)k0));
    AKM_TRY(dev_alloc(h, &h->inv, (size_t)k0));
    h->inv_cap = k0;
    AKM_TRY(dev_alloc(h, &h->shift, (size_t)k0));
    h->kvec_cap = k0;
    return AKM_OK;
}

// scratch centroid buffers for the host-buffer entry points (they do not disturb the ping-pong state)
struct Scratch {
    double *c[2] = {nullptr, nullptr};
    double *l[2] = {nullptr, nullptr};
    double *sum = nullptr;
    int *cmap = nullptr, *inv = nullptr;    // compaction maps of akm_update_centroids / akm_superstep (not the loop's)
    int cap = 0;
    double *dist_all = nullptr; size_t dist_all_count = 0;
    double *probs = nullptr; size_t probs_count = 0;
    int64_t *idx64 = nullptr; int64_t idx64_cap = 0;
};

static Scratch *scratch_of(akm_handle h) {
    if (!h->sc) h->sc = new Scratch();
    return h->sc;
}

static void scratch_free(akm_handle h) {
    Scratch *e = h->sc;
    if (!e) return;
    for (int i = 0; i < 2; ++i) { dev_free(e->c[i]); dev_free(e->l[i]); }
    dev_free(e->cmap); dev_free(e->inv);
    dev_free(e->sum); dev_free(e->dist_all); dev_free(e->probs); dev_free(e->idx64);
    delete e;
    h->sc = nullptr;
}

static int ensure_scratch(akm_handle h, int k0) {
    Scratch *s = scratch_of(h);
    if (s->cap >= k0) return AKM_OK;
    for (int i = 0; i < 2; ++i) { dev_free(s->c[i]); dev_free(s->l[i]); }
    dev_free(s->sum); dev_free(s->cmap); dev_free(s->inv);
    const size_t kd = (size_t)k0 * (size_t)h->d;
    for (int i = 0; i < 2; ++i) {
        AKM_TRY(dev_alloc(h, &s->c[i], kd));
        AKM_TRY(dev_alloc(h, &s->l[i], (size_t)k0));
    }
    AKM_TRY(dev_alloc(h, &s->sum, (size_t)k0 * (size_t)(h->d + 1) + 2));
    AKM_TRY(dev_alloc(h, &s->cmap, (size_t)k0));
    AKM_TRY(dev_alloc(h, &s->inv, (size_t)k0));
    s->cap = k0;
    return ensure_kvec(h, k0);
}

// h->labels is about to hold something other than the last training assignment of h->rows (prediction labels,
// caller-provided ids, ...): neither the search seeds nor the incremental sums describe it any more
static void invalidate_label_state(akm_handle h) {
    tensor_invalidate_seeds(h);
    h->acc_valid = false;
    h->last_changed = -1;
}

// ------------------------------------------------------------------ assignment dispatch
static int run_assign(akm_handle h, RowStore &rs, const double *C, const double *clabel, int k0, int k,
                      int32_t *labels, double *mind, int chain = 0, const int *remap = nullptr) {
    h->stats.rows = rs.n;
    h->stats.storage_bytes = rs.store;
    h->searched_valid = false;                 // only a tensor search that skipped groups says which ones it looked at
    bool use_tensor = false;
    if (h->engine != AKM_ENGINE_EXACT) {
        use_tensor = tensor_supported(h, k0) && rs.n > 0;
        if (h->engine == AKM_ENGINE_TENSOR && !use_tensor && rs.n > 0)
            return akm_fail(h, AKM_ERR_UNSUPPORTED, "tensor engine does not support this shape");
    }
    if (use_tensor) {
        h->stats.engine_used = AKM_ENGINE_TENSOR;
        return tensor_assign(h, rs, C, clabel, k0, k, labels, mind, chain, remap);
    }
    if (labels == h->labels) tensor_invalidate_seeds(h);
    h->stats.engine_used = AKM_ENGINE_EXACT;
    h->stats.rows_certified = 0;
    h->stats.rows_refined = rs.n;
    h->stats.rows_exact_all = rs.n;
    return launch_assign_exact(h, rs, nullptr, rs.n, C, clabel, k0, k, labels, mind, nullptr);
}

// h_iscal[4] (rows handed to the exact refinement by the tensor search) was copied back by the caller
static void absorb_tensor_stats(akm_handle h) {
    if (h->stats.engine_used == AKM_ENGINE_TENSOR) {
        h->stats.rows_refined = h->h_iscal[4];
        h->stats.rows_certified = h->stats.rows - h->h_iscal[4];
        h->stats.rows_exact_all = h->h_iscal[5];
        h->stats.tiles_computed = h->h_iscal[7];
    } else {
        h->stats.tiles_total = h->stats.tiles_computed = 0;
    }
}

// Keep the cached partition grouped by cluster (reorder.cu).  Policy: regroup after the first tensor-engine
// pass, once more two passes later (the first labels come from the initial centroids and move a lot), and
// afterwards only when a single pass moved more than a quarter of the rows.
static int maybe_reorder(akm_handle h, int k) {
    static const int env = getenv("AKM_REORDER") ? atoi(getenv("AKM_REORDER")) : -1;   // 0 = never
    if (env == 0) return AKM_OK;
    if (h->stats.engine_used != AKM_ENGINE_TENSOR || h->rows.n < h->reorder_min_rows) return AKM_OK;
    h->changed_since_reorder++;                       // tensor passes since the last regrouping (or since upload)
    bool go = !h->rows.reordered && h->changed_since_reorder >= h->regroup_after;   // default 2: the first labels move a lot
    if (!go && h->rows.reordered && (int64_t)h->h_iscal[6] * 4 > h->rows.n) go = true;
    if (!go) return AKM_OK;
    if (h->profiling) cudaEventRecord(h->pev[3], h->stream);
    AKM_TRY(reorder_rows_by_label(h, h->rows, h->labels, tensor_useed(h), k));
    tensor_rows_permuted(h);
    // the incremental sums were formed from these very labels: their record moves with the rows
    if (h->acc_valid && h->labels_acc && h->labels_acc_cap >= h->rows.n)
        AKM_CUDA(h, cudaMemcpyAsync(h->labels_acc, h->labels, sizeof(int32_t) * (size_t)h->rows.n, cudaMemcpyDeviceToDevice, h->stream));
    if (h->profiling) {
        cudaEventRecord(h->pev[4], h->stream);
        cudaEventSynchronize(h->pev[4]);
        float ms = 0.f;
        cudaEventElapsedTime(&ms, h->pev[3], h->pev[4]);
        h->reorder_ms += ms;
    }
    h->changed_since_reorder = 0;
    h->reorders++;
    return AKM_OK;
}

static int check_error_flag(akm_handle h) {
    // h_iscal[2] was copied back by the caller
    if (h->h_iscal[2] != 0) {
        cudaMemsetAsync(h->iscal + 2, 0, sizeof(int), h->stream);
        return akm_fail(h, AKM_ERR_NOFINITE,
                        "a sample has no distance < Double.MAX_VALUE (Java: getMinPointIndex returns -1)");
    }
    return AKM_OK;
}

// ====================================================================== C ABI
extern "C" {

const char *akm_version(void) { return "akm 0.1.0 (sm_100a; exact fp64 + tcgen05 search)"; }

const char *akm_last_error(akm_handle h) { return h ? h->err.c_str() : g_tls_error.c_str(); }

int akm_create(int device, int d, int metric, akm_handle *out) {
    if (!out) return akm_fail(nullptr, AKM_ERR_BADARG, "out is null");
    *out = nullptr;
    if (d <= 0) return akm_fail(nullptr, AKM_ERR_BADARG, "vector size must be positive");
    if (metric != AKM_EUCLIDEAN && metric != AKM_COSINE)
        return akm_fail(nullptr, AKM_ERR_BADARG, "distanceType must be EUCLIDEAN or COSINE");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev <= 0)
        return akm_fail(nullptr, AKM_ERR_CUDA, std::string("no CUDA device: ") + cudaGetErrorString(e));
    if (device < 0 || device >= ndev) return akm_fail(nullptr, AKM_ERR_BADARG, "bad device ordinal");
    akm_handle h = new (std::nothrow) akm_handle_s();
    if (!h) return akm_fail(nullptr, AKM_ERR_OOM, "host allocation failed");
    h->device = device; h->d = d; h->metric = metric;
    h->rows.store = 4;
    h->batch.store = h->rows.store;
    int rc = AKM_OK;
    do {
        if ((rc = bind_device(h))) break;
        cudaDeviceProp prop;
        if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) { rc = akm_fail(h, AKM_ERR_CUDA, "cudaGetDeviceProperties failed"); break; }
        h->num_sms = prop.multiProcessorCount;
        h->smem_optin = prop.sharedMemPerBlockOptin;
        if (cudaStreamCreateWithFlags(&h->own_stream, cudaStreamNonBlocking) != cudaSuccess) { rc = akm_fail(h, AKM_ERR_CUDA, "cudaStreamCreate failed"); break; }
        h->stream = h->own_stream;
        if (cudaMalloc((void **)&h->dscal, 8 * sizeof(double)) != cudaSuccess ||
            cudaMalloc((void **)&h->iscal, 16 * sizeof(int)) != cudaSuccess ||
            cudaMallocHost((void **)&h->h_dscal, 8 * sizeof(double)) != cudaSuccess ||
            cudaMallocHost((void **)&h->h_iscal, 16 * sizeof(int)) != cudaSuccess) {
            rc = akm_fail(h, AKM_ERR_OOM, "scalar buffers allocation failed"); break;
        }
        cudaMemsetAsync(h->dscal, 0, 8 * sizeof(double), h->stream);
        cudaMemsetAsync(h->iscal, 0, 16 * sizeof(int), h->stream);
        memset(h->h_dscal, 0, 8 * sizeof(double));
        memset(h->h_iscal, 0, 16 * sizeof(int));
        if (cudaStreamSynchronize(h->stream) != cudaSuccess) { rc = akm_fail(h, AKM_ERR_CUDA, "stream sync failed"); break; }
    } while (0);
    if (rc != AKM_OK) { std::string m = h->err; akm_destroy(h); g_tls_error = m; return rc; }
    *out = h;
    return AKM_OK;
}

int akm_destroy(akm_handle h) {
    if (!h) return AKM_OK;
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    nccl_destroy(h);
    tensor_free(h);
    scratch_free(h);
    rowstore_free(h->rows);
    rowstore_free(h->batch);
    dev_free(h->labels); dev_free(h->mind); dev_free(h->labels_acc); dev_free(h->sum_local); dev_free(h->sum_tmp);
    for (int i = 0; i < 2; ++i) { dev_free(h->cent[i]); dev_free(h->clabel[i]); }
    dev_free(h->sum); dev_free(h->cmap); dev_free(h->inv); dev_free(h->shift); dev_free(h->dscal); dev_free(h->iscal);
    dev_free(h->init_cost); dev_free(h->init_near); dev_free(h->init_rows); dev_free(h->init_out);
    if (h->h_dscal) cudaFreeHost(h->h_dscal);
    if (h->h_iscal) cudaFreeHost(h->h_iscal);
    if (h->stage) cudaFree(h->stage);
    if (h->hstage) cudaFreeHost(h->hstage);
    if (h->pin) cudaFreeHost(h->pin);
    for (int i = 0; i < 2; ++i) if (h->stage_free[i]) cudaEventDestroy(h->stage_free[i]);
    if (h->reorder_scratch) cudaFree(h->reorder_scratch);
    if (h->reorder_dest) cudaFree(h->reorder_dest);
    if (h->reorder_hist) cudaFree(h->reorder_hist);
    for (int i = 0; i < 5; ++i) if (h->pev[i]) cudaEventDestroy(h->pev[i]);
    if (h->mid_event) cudaEventDestroy(h->mid_event);
    if (h->own_stream) cudaStreamDestroy(h->own_stream);
    delete h;
    return AKM_OK;
}

int akm_set_stream(akm_handle h, void *cuda_stream) {
    AKM_TRY(bind_device(h));
    AKM_CUDA(h, cudaStreamSynchronize(h->stream));
    h->stream = cuda_stream ? (cudaStream_t)cuda_stream : h->own_stream;
    return AKM_OK;
}

int akm_set_engine(akm_handle h, int engine) {
    if (!h) return akm_fail(nullptr, AKM_ERR_BADARG, "null handle");
    REQUIRE(h, engine == AKM_ENGINE_AUTO || engine == AKM_ENGINE_EXACT || engine == AKM_ENGINE_TENSOR, "bad engine");
    h->engine = engine;
    return AKM_OK;
}

int akm_reserve_rows(akm_handle h, int64_t n_total) {
    AKM_TRY(bind_device(h));
    REQUIRE(h, n_total >= 0, "negative row count");
    AKM_TRY(rowstore_reserve(h, h->rows, n_total));
    // everything else whose size follows from the row count, so that the supersteps allocate nothing
    if (n_total >= h->reorder_min_rows && h->engine != AKM_ENGINE_EXACT) AKM_TRY(reorder_reserve(h, h->rows, n_total));
    return AKM_OK;
}

int akm_upload_rows(akm_handle h, const double *x_host, int64_t n) {
    AKM_TRY(bind_device(h, /*uploading=*/true));
    REQUIRE(h, n >= 0, "negative row count");
    REQUIRE(h, n == 0 || x_host, "rows pointer is null");
    return upload_rows(h, h->rows, x_host, SRC_HOST_F64, n);
}

int akm_upload_rows_prepared(akm_handle h, const double *x_host, int64_t n) {
    AKM_TRY(bind_device(h, /*uploading=*/true));
    REQUIRE(h, n >= 0, "negative row count");
    REQUIRE(h, n == 0 || x_host, "rows pointer is null");
    return upload_rows(h, h->rows, x_host, SRC_HOST_F64, n, /*prepared=*/true);
}

int akm_host_alloc(size_t bytes, void **out) {
    if (!out) return akm_fail(nullptr, AKM_ERR_BADARG, "out is null");
    *out = nullptr;
    cudaError_t e = cudaHostAlloc(out, std::max<size_t>(bytes, 1), cudaHostAllocPortable);
    if (e != cudaSuccess) return akm_fail(nullptr, e == cudaErrorMemoryAllocation ? AKM_ERR_OOM : AKM_ERR_CUDA,
                                          std::string("cudaHostAlloc: ") + cudaGetErrorString(e));
    return AKM_OK;
}

int akm_host_free(void *p) {
    if (p) cudaFreeHost(p);
    return AKM_OK;
}

int akm_upload_rows_csr(akm_handle h, const int64_t *indptr, const int32_t *indices, const double *values, int64_t n) {
    AKM_TRY(bind_device(h));
    REQUIRE(h, n >= 0, "negative row count");
    if (n == 0) return AKM_OK;
    REQUIRE(h, indptr && indptr[0] == 0, "indptr is null or does not start at 0");
    REQUIRE(h, indptr[n] == 0 || (indices && values), "indices / values is null");
    RowStore &rs = h->rows;
    const size_t d = (size_t)h->d;
    // every value an fp32 value?  (the non-zeros only: cheap on the host, and it settles the storage before any row lands)
    bool exact = true;
    for (int64_t i = 0; i < indptr[n] && exact; ++i) exact = (double)(float)values[i] == values[i];
    if (!exact && rs.store == 4) AKM_TRY(rowstore_widen(h, rs));
    AKM_TRY(rowstore_reserve(h, rs, rs.n + n));
    h->xl_total_valid = false; h->acc_valid = false; h->last_changed = -1;
    rs.x16_valid = false;
    int64_t chunk = (int64_t)std::max<size_t>(1, UPLOAD_HALF_BYTES / (d * 8));      // dense doubles of a chunk fill one staging half
    chunk = std::min<int64_t>(chunk, n);
    AKM_TRY(ensure_dev_stage(h, 2 * (size_t)chunk * d * 8));
    const size_t dhalf = h->stage_bytes / 2;
    for (int64_t off = 0; off < n; off += chunk) {
        const int64_t m = std::min<int64_t>(chunk, n - off);
        for (int64_t r = off; r < off + m; ++r)
            if (indptr[r + 1] < indptr[r]) return akm_fail(h, AKM_ERR_BADARG, "indptr must be non-decreasing");
        const int64_t nnz0 = indptr[off], nnz = indptr[off + m] - nnz0;
        const size_t need = sizeof(int64_t) * (size_t)(m + 1) + (size_t)nnz * 12 + 64;
        if (h->csr_bytes < need) {
            AKM_CUDA(h, cudaStreamSynchronize(h->stream));
            if (h->csr_dev) cudaFree(h->csr_dev);
            h->csr_dev = nullptr; h->csr_bytes = 0;
            AKM_CUDA(h, cudaMalloc(&h->csr_dev, need + need / 2));
            h->csr_bytes = need + need / 2;
        }
        // (pageable sources: these copies return once the driver has staged them, so the caller's arrays are free
        // again when the call returns)
        int64_t *ip = (int64_t *)h->csr_dev;
        double *vp = (double *)(ip + ((m + 1 + 1) & ~(int64_t)1));
        int32_t *xp = (int32_t *)(vp + nnz);
        const int slot = h->stage_slot;
        h->stage_slot ^= 1;
        double *dense = (double *)((char *)h->stage + (size_t)slot * dhalf);
        AKM_CUDA(h, cudaMemcpyAsync(ip, indptr + off, sizeof(int64_t) * (size_t)(m + 1), cudaMemcpyHostToDevice, h->stream));
        if (nnz > 0) {
            AKM_CUDA(h, cudaMemcpyAsync(vp, values + nnz0, sizeof(double) * (size_t)nnz, cudaMemcpyHostToDevice, h->stream));
            AKM_CUDA(h, cudaMemcpyAsync(xp, indices + nnz0, sizeof(int32_t) * (size_t)nnz, cudaMemcpyHostToDevice, h->stream));
        }
        AKM_CUDA(h, cudaMemsetAsync(h->iscal + 9, 0, sizeof(int), h->stream));
        AKM_TRY(launch_csr_expand(h, ip, xp, vp, nnz0, m, dense, h->iscal + 9));
        AKM_TRY(launch_prep_rows(h, rs, dense, false, rs.n, m, false));
        AKM_CUDA(h, cudaMemcpyAsync(h->h_iscal + 9, h->iscal + 9, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
        AKM_CUDA(h, cudaStreamSynchronize(h->stream));            // the CSR chunk buffer is reused by the next chunk
        if (h->h_iscal[9] != 0) return akm_fail(h, AKM_ERR_BADARG, "a column index lies outside [0, vectorSize)");
        rs.n += m;
    }
    if (rs.reordered) AKM_TRY(ensure_orig(h, rs));
    return AKM_OK;
}

int akm_upload_wait(akm_handle h) {
    AKM_TRY(bind_device(h));
    AKM_CUDA(h, cudaStreamSynchronize(h->stream));
    return AKM_OK;
}

int akm_upload_rows_f32(akm_handle h, const float *x_host, int64_t n) {
    AKM_TRY(bind_device(h, /*uploading=*/true));
    REQUIRE(h, n >= 0, "negative row count");
    REQUIRE(h, n == 0 || x_host, "rows pointer is null");
    return upload_rows(h, h->rows, x_host, SRC_HOST_F32, n);
}

int akm_upload_rows_device_f32(akm_handle h, const float *x_dev, int64_t n) {
    AKM_TRY(bind_device(h));
    REQUIRE(h, n >= 0, "negative row count");
    REQUIRE(h, n == 0 || x_dev, "rows pointer is null");
    return upload_rows(h, h->rows, x_dev, SRC_DEV_F32, n);
}

int akm_clear_rows(akm_handle h) {
    AKM_TRY(bind_device(h));
    AKM_CUDA(h, cudaStreamSynchronize(h->stream));
    rowstore_free(h->rows);
    h->rows.store = 4;
    invalidate_label_state(h);
    h->xl_total_valid = false;
    dev_free(h->init_cost); dev_free(h->init_near);
    h->init_cap = 0;
    return AKM_OK;
}

int akm_num_rows(akm_handle h, int64_t *n_out, int *bytes_per_value_out) {
    if (!h) return akm_fail(nullptr, AKM_ERR_BADARG, "null handle");
    if (n_out) *n_out = h->rows.n;
    if (bytes_per_value_out) *bytes_per_value_out = h->rows.store;
    return AKM_OK;
}

int akm_assign_accumulate(akm_handle h, const double *centroids, int k0, int k, double *sum_out,
                          double *objective_out, int32_t *labels_out) {
    AKM_TRY(bind_device(h));
    REQUIRE(h, centroids && sum_out, "centroids / sum_out is null");        // Preconditions.checkNotNull(sumMatrix)
    REQUIRE(h, k0 > 0 && k >= 0 && k <= k0, "need 0 <= k <= k0 and k0 > 0");
    const int d = h->d;
    const size_t L = (size_t)k0 * (size_t)(d + 1);
    AKM_TRY(ensure_scratch(h, k0));
    Scratch *s = scratch_of(h);
    if (h->rows.n == 0) {
        // "if (trainData == null) return;" happens before the fill in Java; an empty list zero-fills
        memset(sum_out, 0, L * sizeof(double));
        if (objective_out) *objective_out = 0.0;
        return AKM_OK;
    }
    AKM_TRY(ensure_labels(h, h->rows.n));
    AKM_TRY(pin_begin(h, sizeof(double) * ((size_t)k0 * d + L)));
    AKM_TRY(pin_h2d(h, s->c[0], centroids, sizeof(double) * (size_t)k0 * d));
    AKM_TRY(launch_prepare_centroids(h, s->c[0], s->l[0], k0, /*cos_normalise=*/false));
    AKM_TRY(run_assign(h, h->rows, s->c[0], s->l[0], k0, k, h->labels, nullptr, /*chain=*/2, nullptr));
    AKM_TRY(accumulate_step(h, s->c[0], s->l[0], k0, s->sum));
    AKM_TRY(pin_d2h(h, sum_out, s->sum, sizeof(double) * L));
    AKM_CUDA(h, cudaMemcpyAsync(h->h_dscal, s->sum + L, sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    AKM_CUDA(h, cudaMemcpyAsync(h->h_iscal + 2, h->iscal + 2, 7 * sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    if (labels_out) AKM_TRY(copy_rows_out_i32(h, h->rows, h->labels, labels_out, h->rows.n));
    AKM_CUDA(h, cudaStreamSynchronize(h->stream));
    pin_finish(h);
    absorb_tensor_stats(h);
    accumulate_absorb(h);
    if (objective_out) *objective_out = h->h_dscal[0];
    AKM_TRY(check_error_flag(h));
    return maybe_reorder(h, k);
}

int akm_update_centroids(akm_handle h, const double *sum, int k0, int k, int step_no, const double *prev,
                         double *c_out, int *k_out, double *shift_out) {
    AKM_TRY(bind_device(h));
    REQUIRE(h, sum && c_out && k_out, "sum / c_out / k_out is null");
    REQUIRE(h, k0 > 0 && k >= 0 && k <= k0, "need 0 <= k <= k0 and k0 > 0");
    const int d = h->d;
    const size_t L = (size_t)k0 * (size_t)(d + 1);
    AKM_TRY(ensure_scratch(h, k0));
    Scratch *s = scratch_of(h);
    AKM_TRY(pin_begin(h, sizeof(double) * (L + 2 * (size_t)k0 * d + (size_t)k0)));
    AKM_TRY(pin_h2d(h, s->sum, sum, sizeof(double) * L));
    const double *dprev = nullptr, *lprev = nullptr;
    if (prev) {
        AKM_TRY(pin_h2d(h, s->c[0], prev, sizeof(double) * (size_t)k0 * d));
        AKM_TRY(launch_prepare_centroids(h, s->c[0], s->l[0], k0, false));
        dprev = s->c[0]; lprev = s->l[0];
    }
    // the buffer written at step s is CENTROID1 when s is odd (KMeansUpdateCentroids.java:38-42)
    AKM_TRY(launch_update(h, s->sum, k, k0, s->c[1], s->l[1], dprev, lprev, (step_no % 2) != 0, 0.0, h->shift, h->iscal,
                          s->cmap, s->inv));
    AKM_TRY(pin_d2h(h, c_out, s->c[1], sizeof(double) * (size_t)k0 * d));
    // the first *k_out entries of shift are meaningful; k_out is not known before the sync, so fetch k entries
    // (k_out <= k) into the arena and hand over the valid prefix afterwards
    char *shift_pin = (shift_out && prev && k > 0) ? pin_take(h, sizeof(double) * (size_t)k) : nullptr;
    if (shift_pin) AKM_CUDA(h, cudaMemcpyAsync(shift_pin, h->shift, sizeof(double) * (size_t)k, cudaMemcpyDeviceToHost, h->stream));
    AKM_CUDA(h, cudaMemcpyAsync(h->h_iscal, h->iscal, 2 * sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    AKM_CUDA(h, cudaStreamSynchronize(h->stream));
    pin_finish(h);
    *k_out = h->h_iscal[0];
    if (shift_out && prev && *k_out > 0) {
        if (shift_pin) memcpy(shift_out, shift_pin, sizeof(double) * (size_t)(*k_out));
        else AKM_CUDA(h, cudaMemcpy(shift_out, h->shift, sizeof(double) * (size_t)(*k_out), cudaMemcpyDeviceToHost));
    }
    return AKM_OK;
}

// The three stages above in one call: one host->device copy (centroids), one device->host batch, one
// synchronisation; the accumulator never leaves the device between the stages.
int akm_superstep(akm_handle h, const double *centroids, int k0, int k, int step_no, double tol, double *c_out,
                  int *k_out, int *converged_out, double *sum_out, double *shift_out, double *objective_out) {
    AKM_TRY(bind_device(h));
    REQUIRE(h, centroids && c_out && k_out, "centroids / c_out / k_out is null");
    REQUIRE(h, k0 > 0 && k >= 0 && k <= k0, "need 0 <= k <= k0 and k0 > 0");
    const int d = h->d;
    const size_t L = (size_t)k0 * (size_t)(d + 1), kd = (size_t)k0 * d;
    AKM_TRY(ensure_scratch(h, k0));
    Scratch *s = scratch_of(h);
    AKM_TRY(ensure_labels(h, std::max<int64_t>(h->rows.n, 1)));
    AKM_TRY(pin_begin(h, sizeof(double) * (2 * kd + L + (size_t)k0 + 64)));
    AKM_TRY(pin_h2d(h, s->c[0], centroids, sizeof(double) * kd));
    AKM_TRY(launch_prepare_centroids(h, s->c[0], s->l[0], k0, /*cos_normalise=*/false));
    if (h->rows.n > 0) AKM_TRY(run_assign(h, h->rows, s->c[0], s->l[0], k0, k, h->labels, nullptr, /*chain=*/2, nullptr));
    AKM_TRY(accumulate_step(h, s->c[0], s->l[0], k0, s->sum));
    if (h->comm) {
        AKM_TRY(nccl_allreduce_device(h, s->sum, (int64_t)L + 2));
        AKM_TRY(launch_flag_from_sum(h, s->sum + L + 1));
    }
    AKM_TRY(launch_update(h, s->sum, k, k0, s->c[1], s->l[1], s->c[0], s->l[0], (step_no % 2) != 0, tol, h->shift, h->iscal,
                          s->cmap, s->inv));
    AKM_TRY(pin_d2h(h, c_out, s->c[1], sizeof(double) * kd));
    if (sum_out) AKM_TRY(pin_d2h(h, sum_out, s->sum, sizeof(double) * L));
    char *shift_pin = (shift_out && k > 0) ? pin_take(h, sizeof(double) * (size_t)k) : nullptr;
    if (shift_pin) AKM_CUDA(h, cudaMemcpyAsync(shift_pin, h->shift, sizeof(double) * (size_t)k, cudaMemcpyDeviceToHost, h->stream));
    AKM_CUDA(h, cudaMemcpyAsync(h->h_iscal, h->iscal, 9 * sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    AKM_CUDA(h, cudaMemcpyAsync(h->h_dscal, s->sum + L, sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    AKM_CUDA(h, cudaStreamSynchronize(h->stream));
    pin_finish(h);
    if (h->rows.n > 0) absorb_tensor_stats(h);
    accumulate_absorb(h);
    *k_out = h->h_iscal[0];
    if (converged_out) *converged_out = h->h_iscal[1];
    if (objective_out) *objective_out = h->h_dscal[0];
    if (shift_out && *k_out > 0) {
        if (shift_pin) memcpy(shift_out, shift_pin, sizeof(double) * (size_t)(*k_out));
        else AKM_CUDA(h, cudaMemcpy(shift_out, h->shift, sizeof(double) * (size_t)(*k_out), cudaMemcpyDeviceToHost));
    }
    AKM_TRY(check_error_flag(h));
    if (h->rows.n == 0) return AKM_OK;
    AKM_TRY(maybe_reorder(h, k));
    // the host will pass the compacted buffer back with the new k: the next labels use the new numbering
    if (*k_out != k) AKM_TRY(accumulate_renumber(h, s->cmap, k, k0));
    return AKM_OK;
}

int akm_set_centroids(akm_handle h, const double *c0, int k0) {
    AKM_TRY(bind_device(h));
    REQUIRE(h, c0, "initial centroids are null");            // KMeansPreallocateCentroid.java:31-32
    REQUIRE(h, k0 > 0, "need at least one initial centroid");
    AKM_TRY(ensure_centroids(h, k0));
    AKM_TRY(ensure_kvec(h, k0));
    const size_t kd = (size_t)k0 * (size_t)h->d;
    AKM_CUDA(h, cudaMemcpyAsync(h->cent[0], c0, sizeof(double) * kd, cudaMemcpyHostToDevice, h->stream));
    AKM_TRY(launch_prepare_centroids(h, h->cent[0], h->clabel[0], k0, true));
    AKM_CUDA(h, cudaMemcpyAsync(h->cent[1], h->cent[0], sizeof(double) * kd, cudaMemcpyDeviceToDevice, h->stream));
    AKM_CUDA(h, cudaMemcpyAsync(h->clabel[1], h->clabel[0], sizeof(double) * (size_t)k0, cudaMemcpyDeviceToDevice, h->stream));
    AKM_CUDA(h, cudaMemsetAsync(h->sum, 0, sizeof(double) * ((size_t)k0 * (h->d + 1) + 2), h->stream));
    // step 0 also prepares what the supersteps would otherwise build or allocate on their first pass
    if (h->rows.n > 0) {
        AKM_TRY(ensure_labels(h, h->rows.n));
        if (h->engine != AKM_ENGINE_EXACT) AKM_TRY(tensor_warm(h, h->rows, k0));
    }
    AKM_CUDA(h, cudaStreamSynchronize(h->stream));
    h->k0 = k0; h->k = k0; h->step = 0;
    tensor_invalidate_seeds(h);
    h->last_changed = -1;                            // new centroids: expect most rows to move (full accumulation)
    return AKM_OK;
}

int akm_step(akm_handle h, double tol, int *converged_out, int *k_out, double *objective_out) {
    AKM_TRY(bind_device(h));
    if (h->k0 <= 0) return akm_fail(h, AKM_ERR_STATE, "akm_set_centroids must be called first");
    const int d = h->d, k0 = h->k0;
    const size_t L = (size_t)k0 * (size_t)(d + 1);
    const int s = ++h->step;
    const int k_before = h->k;
    const int rd = (s % 2 == 0) ? 0 : 1;   // even: CENTROID1, odd: CENTROID2 (KMeansAssignCluster.java:38-42)
    const int wr = 1 - rd;
    AKM_TRY(ensure_labels(h, std::max<int64_t>(h->rows.n, 1)));
    const bool prof = h->profiling;
    h->mid_recorded = false;
    if (prof) cudaEventRecord(h->pev[0], h->stream);
    if (h->rows.n > 0)
        AKM_TRY(run_assign(h, h->rows, h->cent[rd], h->clabel[rd], k0, h->k, h->labels, nullptr, /*chain=*/1,
                           s > 1 ? h->cmap : nullptr));
    if (prof) cudaEventRecord(h->pev[1], h->stream);
    AKM_TRY(accumulate_step(h, h->cent[rd], h->clabel[rd], k0, h->sum));
    if (prof) cudaEventRecord(h->pev[2], h->stream);
    if (h->comm) {
        // sums, objective and the no-finite-distance flag travel together: every rank sees the same error
        AKM_TRY(nccl_allreduce_device(h, h->sum, (int64_t)L + 2));
        AKM_TRY(launch_flag_from_sum(h, h->sum + L + 1));
    }
    AKM_TRY(launch_update(h, h->sum, h->k, k0, h->cent[wr], h->clabel[wr], h->cent[rd], h->clabel[rd],
                          /*new_is_c1=*/wr == 0, tol, h->shift, h->iscal, h->cmap, h->inv_cap >= k0 ? h->inv : nullptr));
    if (prof) cudaEventRecord(h->pev[3], h->stream);
    AKM_CUDA(h, cudaMemcpyAsync(h->h_iscal, h->iscal, 9 * sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    AKM_CUDA(h, cudaMemcpyAsync(h->h_dscal, h->sum + L, sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    AKM_CUDA(h, cudaStreamSynchronize(h->stream));
    if (h->rows.n > 0) absorb_tensor_stats(h);
    accumulate_absorb(h);
    h->k = h->h_iscal[0];
    if (prof) {
        float a = 0.f, b = 0.f, c = 0.f, m1 = 0.f;
        cudaEventElapsedTime(&a, h->pev[0], h->pev[1]);
        cudaEventElapsedTime(&b, h->pev[1], h->pev[2]);
        cudaEventElapsedTime(&c, h->pev[2], h->pev[3]);
        if (h->mid_recorded) {
            cudaEventElapsedTime(&m1, h->pev[0], h->mid_event);
            h->prof_ms[0] += m1; h->prof_ms[1] += a - m1;
        } else if (h->stats.engine_used == AKM_ENGINE_EXACT) {
            h->prof_ms[1] += a;
        } else {
            h->prof_ms[0] += a;
        }
        h->prof_ms[2] += b; h->prof_ms[3] += c;
        h->prof_steps++;
    }
    if (converged_out) *converged_out = h->h_iscal[1];
    if (k_out) *k_out = h->k;
    if (objective_out) *objective_out = h->h_dscal[0];
    AKM_TRY(check_error_flag(h));
    if (h->rows.n == 0) return AKM_OK;
    // labels index the centroids this step assigned against (k_before clusters)
    AKM_TRY(maybe_reorder(h, k_before));
    // clusters were dropped: the next pass numbers them anew (after the regrouping, which copies the label record
    // in the numbering the sums still have)
    if (h->k != k_before) AKM_TRY(accumulate_renumber(h, h->cmap, k_before, k0));
    return AKM_OK;
}

int akm_get_model(akm_handle h, double *c_out, int *k_out, double *weights_out, int *steps_out) {
    AKM_TRY(bind_device(h));
    if (h->k0 <= 0) return akm_fail(h, AKM_ERR_STATE, "no centroids set");
    const int d = h->d, k0 = h->k0;
    // newest buffer = the one written by the last step (larger stamp, KMeansOutputModel.java:50-54)
    const int newest = (h->step == 0) ? 1 : ((h->step % 2 == 0) ? 1 : 0);
    if (c_out) AKM_CUDA(h, cudaMemcpyAsync(c_out, h->cent[newest], sizeof(double) * (size_t)k0 * d, cudaMemcpyDeviceToHost, h->stream));
    if (weights_out) {
        // weight_j = buffer[j*(d+1)+d] of the last all-reduced pass, uncompacted (KMeansOutputModel.java:60-68)
        AKM_CUDA(h, cudaMemcpy2DAsync(weights_out, sizeof(double), h->sum + d, sizeof(double) * (size_t)(d + 1),
                                      sizeof(double), (size_t)k0, cudaMemcpyDeviceToHost, h->stream));
    }
    AKM_CUDA(h, cudaStreamSynchronize(h->stream));
    if (k_out) *k_out = h->k;
    if (steps_out) *steps_out = h->step;
    return AKM_OK;
}

int akm_get_labels(akm_handle h, int32_t *labels_out) {
    AKM_TRY(bind_device(h));
    REQUIRE(h, labels_out, "labels_out is null");
    if (h->rows.n == 0) return AKM_OK;
    if (!h->labels || h->labels_cap < h->rows.n) return akm_fail(h, AKM_ERR_STATE, "no assignment pass has run");
    AKM_TRY(copy_rows_out_i32(h, h->rows, h->labels, labels_out, h->rows.n));
    AKM_CUDA(h, cudaStreamSynchronize(h->stream));
    return AKM_OK;
}

// ---------------------------------------------------------------- prediction
static int predict_store(akm_handle h, RowStore &rs, const double *dC, const double *dl, int k, int64_t *idx_out,
                         double *dist_out, double *probs_out) {
    Scratch *s = scratch_of(h);
    const int64_t n = rs.n;
    if (n == 0) return AKM_OK;
    invalidate_label_state(h);                       // h->labels is about to hold prediction labels
    AKM_TRY(ensure_labels(h, n));
    AKM_TRY(ensure_mind(h, n));
    if (s->idx64_cap < n) {
        dev_free(s->idx64);
        AKM_TRY(dev_alloc(h, &s->idx64, (size_t)n));
        s->idx64_cap = n;
    }
    if (probs_out) {
        // probabilities need every distance exactly as Java computes it: exact engine, in slabs
        const int64_t slab = std::max<int64_t>(1, std::min<int64_t>(n, ((int64_t)256 << 20) / ((int64_t)k * 8)));
        if (s->dist_all_count < (size_t)slab * k) {
            dev_free(s->dist_all); dev_free(s->probs);
            AKM_TRY(dev_alloc(h, &s->dist_all, (size_t)slab * k));
            AKM_TRY(dev_alloc(h, &s->probs, (size_t)slab * k));
            s->dist_all_count = s->probs_count = (size_t)slab * k;
        }
        for (int64_t off = 0; off < n; off += slab) {
            const int64_t m = std::min<int64_t>(slab, n - off);
            RowStore view = rs;
            if (rs.store == 4) view.x32 = rs.x32 + off * h->d; else view.x64 = rs.x64 + off * h->d;
            view.xlabel = rs.xlabel + off;
            view.n = m;
            AKM_TRY(launch_assign_exact(h, view, nullptr, m, dC, dl, k, k, h->labels + off, h->mind + off, s->dist_all));
            AKM_TRY(launch_probs(h, s->dist_all, m, k, s->probs));
            AKM_CUDA(h, cudaMemcpyAsync(probs_out + (size_t)off * k, s->probs, sizeof(double) * (size_t)m * k,
                                        cudaMemcpyDeviceToHost, h->stream));
            AKM_CUDA(h, cudaStreamSynchronize(h->stream));
        }
    } else {
        AKM_TRY(run_assign(h, rs, dC, dl, k, k, h->labels, dist_out ? h->mind : nullptr));
    }
    AKM_TRY(launch_labels_to_i64(h, h->labels, n, s->idx64));
    AKM_TRY(copy_rows_out_i64(h, rs, s->idx64, idx_out, n));
    if (dist_out) AKM_TRY(copy_rows_out_f64(h, rs, h->mind, dist_out, n));
    AKM_CUDA(h, cudaMemcpyAsync(h->h_iscal + 2, h->iscal + 2, 6 * sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    AKM_CUDA(h, cudaStreamSynchronize(h->stream));
    if (!probs_out) absorb_tensor_stats(h);
    if (h->h_iscal[2] != 0) {
        cudaMemsetAsync(h->iscal + 2, 0, sizeof(int), h->stream);
        // Java: (long) -1 is emitted; reading clusterDistances[-1] throws only when a distance/detail column is asked for
        if (dist_out || probs_out)
            return akm_fail(h, AKM_ERR_NOFINITE, "a row has no distance < Double.MAX_VALUE (index -1)");
    }
    return AKM_OK;
}

int akm_predict(akm_handle h, const double *x_host, int64_t n, const double *centroids, int k, int64_t *idx_out,
                double *dist_out, double *probs_out) {
    AKM_TRY(bind_device(h));
    REQUIRE(h, n >= 0, "negative row count");
    REQUIRE(h, centroids && k > 0, "model centroids missing");
    REQUIRE(h, n == 0 || (x_host && idx_out), "rows / idx_out is null");
    if (n == 0) return AKM_OK;
    AKM_TRY(ensure_scratch(h, k));
    Scratch *s = scratch_of(h);
    const int d = h->d;
    AKM_CUDA(h, cudaMemcpyAsync(s->c[0], centroids, sizeof(double) * (size_t)k * d, cudaMemcpyHostToDevice, h->stream));
    AKM_TRY(launch_prepare_centroids(h, s->c[0], s->l[0], k, /*cos_normalise=*/true));   // loadModel -> updateLabel
    const int64_t macro = std::max<int64_t>(1, std::min<int64_t>(n, ((int64_t)2 << 30) / ((int64_t)d * 8)));
    for (int64_t off = 0; off < n; off += macro) {
        const int64_t m = std::min<int64_t>(macro, n - off);
        h->batch.n = 0;
        h->batch.x16_valid = false;
        AKM_TRY(upload_rows(h, h->batch, x_host + (size_t)off * d, SRC_HOST_F64, m));
        AKM_TRY(predict_store(h, h->batch, s->c[0], s->l[0], k, idx_out + off, dist_out ? dist_out + off : nullptr,
                              probs_out ? probs_out + (size_t)off * k : nullptr));
    }
    return AKM_OK;
}

int akm_predict_cached(akm_handle h, const double *centroids, int k, int64_t *idx_out, double *dist_out) {
    AKM_TRY(bind_device(h));
    REQUIRE(h, centroids && k > 0, "model centroids missing");
    REQUIRE(h, h->rows.n == 0 || idx_out, "idx_out is null");
    AKM_TRY(ensure_scratch(h, k));
    Scratch *s = scratch_of(h);
    AKM_CUDA(h, cudaMemcpyAsync(s->c[0], centroids, sizeof(double) * (size_t)k * h->d, cudaMemcpyHostToDevice, h->stream));
    AKM_TRY(launch_prepare_centroids(h, s->c[0], s->l[0], k, true));
    return predict_store(h, h->rows, s->c[0], s->l[0], k, idx_out, dist_out, nullptr);
}

// ---------------------------------------------------------------- initial centroids (KMeansInitCentroids)
static int ensure_init(akm_handle h) {
    const int64_t n = std::max<int64_t>(h->rows.n, 1);
    if (h->init_cap >= n && h->init_cost) return AKM_OK;
    dev_free(h->init_cost); dev_free(h->init_near);
    h->init_cap = 0;
    AKM_TRY(dev_alloc(h, &h->init_cost, (size_t)n));
    AKM_TRY(dev_alloc(h, &h->init_near, (size_t)n));
    h->init_cap = n;
    return launch_init_reset(h, h->rows.n, h->init_cost, h->init_near);
}

static int ensure_init_lists(akm_handle h, int m, size_t out_count) {
    if (h->init_rows_cap < m) {
        dev_free(h->init_rows);
        h->init_rows_cap = 0;
        AKM_TRY(dev_alloc(h, &h->init_rows, (size_t)std::max(m, 1)));
        h->init_rows_cap = std::max(m, 1);
    }
    if (h->init_out_count < out_count) {
        dev_free(h->init_out);
        h->init_out_count = 0;
        AKM_TRY(dev_alloc(h, &h->init_out, std::max<size_t>(out_count, 1)));
        h->init_out_count = std::max<size_t>(out_count, 1);
    }
    return AKM_OK;
}

int akm_init_reset(akm_handle h) {
    AKM_TRY(bind_device(h));
    if (h->init_cap >= std::max<int64_t>(h->rows.n, 1) && h->init_cost)
        return launch_init_reset(h, h->rows.n, h->init_cost, h->init_near);
    return ensure_init(h);
}

int akm_init_update(akm_handle h, const double *centers, int m, int id_base, double *sum_cost_out) {
    AKM_TRY(bind_device(h));
    REQUIRE(h, centers && m > 0 && id_base >= 0 && sum_cost_out, "centers / sum_cost_out missing");
    AKM_TRY(ensure_init(h));
    AKM_TRY(ensure_scratch(h, m));
    Scratch *s = scratch_of(h);
    const int64_t n = h->rows.n;
    AKM_CUDA(h, cudaMemcpyAsync(s->c[0], centers, sizeof(double) * (size_t)m * h->d, cudaMemcpyHostToDevice, h->stream));
    // the centres are rows of the prepared partition: labels only, no second normalisation
    AKM_TRY(launch_prepare_centroids(h, s->c[0], s->l[0], m, /*cos_normalise=*/false));
    if (n > 0) {
        invalidate_label_state(h);
        AKM_TRY(ensure_labels(h, n));
        AKM_TRY(ensure_mind(h, n));
        AKM_TRY(run_assign(h, h->rows, s->c[0], s->l[0], m, m, h->labels, h->mind));
    }
    AKM_TRY(launch_init_merge(h, n, h->labels, h->mind, id_base, h->init_cost, h->init_near, h->dscal));
    AKM_CUDA(h, cudaMemcpyAsync(h->h_dscal, h->dscal, sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    AKM_CUDA(h, cudaMemcpyAsync(h->h_iscal + 2, h->iscal + 2, 6 * sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    AKM_CUDA(h, cudaStreamSynchronize(h->stream));
    absorb_tensor_stats(h);
    AKM_TRY(check_error_flag(h));
    *sum_cost_out = h->h_dscal[0];
    return AKM_OK;
}

int akm_init_pick(akm_handle h, double scale, uint64_t seed, int round, int64_t row_base, int cap, int64_t *rows_out,
                  int *count_out) {
    AKM_TRY(bind_device(h));
    REQUIRE(h, cap > 0 && rows_out && count_out, "rows_out / count_out missing");
    AKM_TRY(ensure_init(h));
    AKM_TRY(ensure_init_lists(h, cap, 0));
    AKM_TRY(launch_init_pick(h, h->rows, h->init_cost, row_base, scale, seed, round, cap, h->init_rows, h->iscal));
    AKM_CUDA(h, cudaMemcpyAsync(h->h_iscal, h->iscal, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    AKM_CUDA(h, cudaStreamSynchronize(h->stream));
    const int got = h->h_iscal[0];
    *count_out = got;                                    // may exceed cap: the caller sees that the list was cut
    const int take = std::min(got, cap);
    if (take > 0) {
        AKM_CUDA(h, cudaMemcpyAsync(rows_out, h->init_rows, sizeof(int64_t) * (size_t)take, cudaMemcpyDeviceToHost, h->stream));
        AKM_CUDA(h, cudaStreamSynchronize(h->stream));
        std::sort(rows_out, rows_out + take);            // the order atomics produced is not reproducible
    }
    return AKM_OK;
}

static int upload_row_list(akm_handle h, const int64_t *rows_sorted, int m, size_t out_count) {
    for (int i = 0; i < m; ++i) {
        if (rows_sorted[i] < 0 || rows_sorted[i] >= h->rows.n || (i > 0 && rows_sorted[i] <= rows_sorted[i - 1]))
            return akm_fail(h, AKM_ERR_BADARG, "row indices must be strictly increasing and inside the partition");
    }
    AKM_TRY(ensure_init_lists(h, m, out_count));
    AKM_CUDA(h, cudaMemcpyAsync(h->init_rows, rows_sorted, sizeof(int64_t) * (size_t)m, cudaMemcpyHostToDevice, h->stream));
    return AKM_OK;
}

int akm_init_counts(akm_handle h, int m_total, double *counts_out) {
    AKM_TRY(bind_device(h));
    REQUIRE(h, m_total > 0 && counts_out, "counts_out missing");
    AKM_TRY(ensure_init(h));
    AKM_TRY(ensure_init_lists(h, 1, (size_t)m_total));
    AKM_TRY(launch_init_counts(h, h->rows.n, h->init_near, m_total, h->init_out));
    AKM_CUDA(h, cudaMemcpyAsync(counts_out, h->init_out, sizeof(double) * (size_t)m_total, cudaMemcpyDeviceToHost, h->stream));
    AKM_CUDA(h, cudaStreamSynchronize(h->stream));
    return AKM_OK;
}

int akm_get_rows(akm_handle h, const int64_t *rows_sorted, int m, double *out) {
    AKM_TRY(bind_device(h));
    REQUIRE(h, m >= 0 && (m == 0 || (rows_sorted && out)), "rows / out missing");
    if (m == 0) return AKM_OK;
    AKM_TRY(upload_row_list(h, rows_sorted, m, (size_t)m * h->d));
    AKM_TRY(launch_gather_rows(h, h->rows, h->init_rows, m, h->init_out));
    AKM_CUDA(h, cudaMemcpyAsync(out, h->init_out, sizeof(double) * (size_t)m * h->d, cudaMemcpyDeviceToHost, h->stream));
    AKM_CUDA(h, cudaStreamSynchronize(h->stream));
    return AKM_OK;
}

// ---------------------------------------------------------------- cluster evaluation (EvalClusterBatchOp)
int akm_eval_clusters(akm_handle h, const int32_t *labels, int k, double *count_out, double *mean_out,
                      double *dist_sum_out, double *dist_sq_sum_out, double *norm_sq_sum_out, double *silhouette_sum_out) {
    AKM_TRY(bind_device(h));
    REQUIRE(h, k > 0 && count_out && mean_out && dist_sum_out && dist_sq_sum_out && norm_sq_sum_out && silhouette_sum_out,
            "output buffers missing");
    const int64_t n = h->rows.n;
    const int d = h->d;
    REQUIRE(h, n == 0 || labels, "labels missing");
    AKM_TRY(ensure_scratch(h, k));
    Scratch *s = scratch_of(h);
    const size_t L = (size_t)k * (d + 1);
    std::vector<double> sums(L, 0.0);
    if (n > 0) {
        invalidate_label_state(h);                       // h->labels is about to hold the caller's cluster ids
        AKM_TRY(ensure_labels(h, n));
        AKM_TRY(ensure_init_lists(h, 1, (size_t)(n + 1) / 2 + 1));        // staging for the int32 labels
        int32_t *stage = reinterpret_cast<int32_t *>(h->init_out);
        AKM_CUDA(h, cudaMemcpyAsync(stage, labels, sizeof(int32_t) * (size_t)n, cudaMemcpyHostToDevice, h->stream));
        AKM_TRY(launch_labels_in(h, h->rows, stage, h->labels));
        AKM_TRY(launch_accumulate(h, h->rows, h->labels, s->c[0], s->l[0], k, s->sum, nullptr));
        AKM_CUDA(h, cudaMemcpyAsync(sums.data(), s->sum, sizeof(double) * L, cudaMemcpyDeviceToHost, h->stream));
        AKM_CUDA(h, cudaStreamSynchronize(h->stream));
    }
    // mean = sum.scale(1.0 / total)  (ClusterEvaluationUtil.java:204); |mean| for the COSINE distance to the mean
    std::vector<double> mnorm((size_t)k, 0.0);
    for (int j = 0; j < k; ++j) {
        const double c = sums[(size_t)j * (d + 1) + d];
        count_out[j] = c;
        const double inv = 1.0 / c;
        double nn = 0.0;
        for (int i = 0; i < d; ++i) {
            const double m = (c > 0.0) ? sums[(size_t)j * (d + 1) + i] * inv : 0.0;
            mean_out[(size_t)j * d + i] = m;
            nn += m * m;
        }
        mnorm[j] = std::sqrt(nn);
    }
    if (n == 0) {
        for (int j = 0; j < k; ++j) dist_sum_out[j] = dist_sq_sum_out[j] = norm_sq_sum_out[j] = 0.0;
        *silhouette_sum_out = 0.0;
        return AKM_OK;
    }
    // device copies: means in c[0], |mean| in l[0], counts in l[1]; per-cluster outputs in c[1] (3k doubles, d >= 1)
    AKM_TRY(ensure_init_lists(h, 1, (size_t)3 * k + 1));
    double *out3k = h->init_out;
    AKM_CUDA(h, cudaMemcpyAsync(s->c[0], mean_out, sizeof(double) * (size_t)k * d, cudaMemcpyHostToDevice, h->stream));
    AKM_CUDA(h, cudaMemcpyAsync(s->l[0], mnorm.data(), sizeof(double) * (size_t)k, cudaMemcpyHostToDevice, h->stream));
    AKM_CUDA(h, cudaMemcpyAsync(s->l[1], count_out, sizeof(double) * (size_t)k, cudaMemcpyHostToDevice, h->stream));
    AKM_TRY(launch_eval_own(h, h->rows, h->labels, k, s->c[0], s->l[0], out3k));
    AKM_TRY(launch_eval_silhouette(h, h->rows, h->labels, k, s->c[0], s->l[1], out3k + 2 * (size_t)k, h->dscal));
    std::vector<double> host3((size_t)3 * k);
    AKM_CUDA(h, cudaMemcpyAsync(host3.data(), out3k, sizeof(double) * 3 * (size_t)k, cudaMemcpyDeviceToHost, h->stream));
    AKM_CUDA(h, cudaMemcpyAsync(h->h_dscal, h->dscal, sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    AKM_CUDA(h, cudaStreamSynchronize(h->stream));
    for (int j = 0; j < k; ++j) {
        dist_sum_out[j] = host3[j];
        dist_sq_sum_out[j] = host3[(size_t)k + j];
        norm_sq_sum_out[j] = host3[2 * (size_t)k + j];
    }
    *silhouette_sum_out = h->h_dscal[0];
    return AKM_OK;
}

int akm_get_stats(akm_handle h, akm_stats *out) {
    if (!h || !out) return akm_fail(h, AKM_ERR_BADARG, "null argument");
    *out = h->stats;
    return AKM_OK;
}

}  // extern "C"

extern "C" int akm_set_profiling(akm_handle h, int on) {
    AKM_TRY(bind_device(h));
    if (on && !h->pev[0]) {
        for (int i = 0; i < 5; ++i) AKM_CUDA(h, cudaEventCreate(&h->pev[i]));
        AKM_CUDA(h, cudaEventCreate(&h->mid_event));
    }
    h->profiling = on != 0;
    return AKM_OK;
}

extern "C" int akm_get_profile(akm_handle h, double *out5, int reset) {  // out5 has room for 7 values
    if (!h || !out5) return akm_fail(h, AKM_ERR_BADARG, "null argument");
    for (int i = 0; i < 4; ++i) out5[i] = h->prof_ms[i];
    out5[4] = (double)h->prof_steps;
    out5[5] = h->reorder_ms;
    out5[6] = (double)h->reorders;
    if (reset) { for (int i = 0; i < 4; ++i) h->prof_ms[i] = 0; h->prof_steps = 0; h->reorder_ms = 0; }
    return AKM_OK;
}

// ---------------------------------------------------------------- synthetic blobs (benchmark data)
extern "C" int akm_blobs_append(akm_handle h, uint64_t seed, int k_true, double spread, double sigma, int64_t row0,
                                int64_t n) {
    AKM_TRY(bind_device(h));
    REQUIRE(h, k_true > 0 && n >= 0 && row0 >= 0, "bad blob parameters");
    if (n == 0) return AKM_OK;
    AKM_TRY(rowstore_reserve(h, h->rows, h->rows.n + n));
    const int64_t chunk = std::max<int64_t>(1, std::min<int64_t>(n, ((int64_t)512 << 20) / ((int64_t)h->d * 4)));
    AKM_TRY(ensure_dev_stage(h, (size_t)chunk * h->d * 4));
    for (int64_t off = 0; off < n; off += chunk) {
        const int64_t m = std::min<int64_t>(chunk, n - off);
        AKM_TRY(launch_blobs(h, seed, k_true, spread, sigma, nullptr, row0 + off, m, (float *)h->stage));
        AKM_TRY(upload_rows(h, h->rows, h->stage, SRC_DEV_F32, m));
    }
    return AKM_OK;
}

extern "C" int akm_blobs_rows(akm_handle h, uint64_t seed, int k_true, double spread, double sigma,
                              const int64_t *rows, int64_t m, float *out_host) {
    AKM_TRY(bind_device(h));
    REQUIRE(h, k_true > 0 && m >= 0 && (m == 0 || (rows && out_host)), "bad blob parameters");
    if (m == 0) return AKM_OK;
    int64_t *drows = nullptr;
    float *dout = nullptr;
    AKM_CUDA(h, cudaMalloc((void **)&drows, sizeof(int64_t) * (size_t)m));
    cudaError_t e = cudaMalloc((void **)&dout, sizeof(float) * (size_t)m * h->d);
    if (e != cudaSuccess) { cudaFree(drows); return akm_fail(h, AKM_ERR_OOM, cudaGetErrorString(e)); }
    int rc = AKM_OK;
    do {
        if (cudaMemcpyAsync(drows, rows, sizeof(int64_t) * (size_t)m, cudaMemcpyHostToDevice, h->stream) != cudaSuccess) { rc = akm_fail(h, AKM_ERR_CUDA, "H2D failed"); break; }
        if ((rc = launch_blobs(h, seed, k_true, spread, sigma, drows, 0, m, dout))) break;
        if (cudaMemcpyAsync(out_host, dout, sizeof(float) * (size_t)m * h->d, cudaMemcpyDeviceToHost, h->stream) != cudaSuccess) { rc = akm_fail(h, AKM_ERR_CUDA, "D2H failed"); break; }
        if (cudaStreamSynchronize(h->stream) != cudaSuccess) { rc = akm_fail(h, AKM_ERR_CUDA, "sync failed"); break; }
    } while (0);
    cudaFree(drows); cudaFree(dout);
    return rc;
}

extern "C" int akm_set_option(akm_handle h, const char *name, double value) {
    if (!h || !name) return akm_fail(h, AKM_ERR_BADARG, "null argument");
    if (strcmp(name, "reorder_min_rows") == 0) { h->reorder_min_rows = (int64_t)value; return AKM_OK; }
    if (strcmp(name, "tile_pruning") == 0) { h->tile_pruning = value != 0.0; return AKM_OK; }
    if (strcmp(name, "delta_accumulate") == 0) { h->delta_accumulate = value != 0.0; h->acc_valid = false; h->last_changed = -1; return AKM_OK; }
    if (strcmp(name, "epi_dbg") == 0) { h->epi_dbg = (int)value; return AKM_OK; }
    if (strcmp(name, "regroup_after") == 0) { h->regroup_after = std::max(1, (int)value); return AKM_OK; }
    return akm_fail(h, AKM_ERR_BADARG, std::string("unknown option: ") + name);
}
