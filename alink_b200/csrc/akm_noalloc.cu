// akm_update_centroids_prealloc: the centroid update for callers that must not allocate device memory.
//
// Who needs this: a Flink TaskManager (or any long-running host) that shares its GPU with other jobs under a
// fixed memory reservation, or that runs the update on a latency-critical thread.  cudaMalloc / cudaFree can
// fail once the reservation is used up, and they synchronise the whole device, stalling every other stream
// of every other task in the process.  Such a caller owns its input and output device buffers already (from
// its own pool, carved at start-up) and only lacks the few kilobytes of working memory the update kernel
// needs for the empty-cluster compaction map.  Here that working memory is a fixed array inside the library
// image itself, so this entry point performs no allocation of any kind (no cudaMalloc, cudaFree,
// cudaMallocAsync, cudaHostAlloc, no handle).
//
// Consequences a user must know:
//   * the array is AKM_PREALLOC_MAX_K ints + a 256-byte control block (about 64 KB); every GPU that loads
//     libakm.so holds it until the process exits, whether or not this entry point is ever called;
//   * k0 > AKM_PREALLOC_MAX_K is rejected with AKM_ERR_BADARG;
//   * all calls on one device share the array, and it keeps whatever the last call left in it.  The
//     control block is cleared at the start of every call; the map is rewritten for the first k entries
//     before it is read.  Two calls that overlap on the GPU (different streams, or different threads
//     without ordering) race on it -- serialise them;
//   * the first call on each device resolves the array's device address (cudaGetSymbolAddress).  Make that
//     first call before starting any stream capture on that device.
#include <cuda_runtime.h>

#include <mutex>
#include <string>

#include "../../include/akm.h"

void akm_set_tls_error(const std::string &s);   // akm_api.cu

namespace {

constexpr int kCtrlInts = 64;                                   // 256-byte control block: [0] new k
constexpr int kScratchInts = kCtrlInts + AKM_PREALLOC_MAX_K;

// Working memory of prealloc_update_kernel.  In host code this name is only a stand-in for the symbol;
// the usable device address comes from cudaGetSymbolAddress (scratch_for_device below).
static __device__ __align__(256) int g_prealloc_scratch[kScratchInts];

constexpr int kMaxDevices = 64;
std::mutex g_mu;
int *g_scratch_addr[kMaxDevices] = {nullptr};

int fail(int code, const std::string &msg) {
    akm_set_tls_error(msg);
    return code;
}

// Once per device, with that device current: turn the symbol into a device pointer.
int scratch_for_device(int device, int **out) {
    std::lock_guard<std::mutex> lk(g_mu);
    if (!g_scratch_addr[device]) {
        int *ptr = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void **)&ptr, g_prealloc_scratch);
        if (e != cudaSuccess || ptr == nullptr)
            return fail(AKM_ERR_CUDA, std::string("cudaGetSymbolAddress(g_prealloc_scratch): ") + cudaGetErrorString(e));
        g_scratch_addr[device] = ptr;
    }
    *out = g_scratch_addr[device];
    return AKM_OK;
}

// Same arithmetic as KMeansUpdateCentroids.updateCentroids (reference: core/src/main/java/com/alibaba/alink/
// operator/common/clustering/kmeans/KMeansUpdateCentroids.java:52-69): zero fill, skip weight == 0, scale
// by the reciprocal (0 + (1.0/w) * x, separately rounded), compact, updateLabel over all k0 rows (COSINE:
// renormalise).  One CTA; `scratch` = control block followed by the compaction map.
__global__ void __launch_bounds__(1024)
prealloc_update_kernel(int metric, const double *__restrict__ sum, int k, int k0, int d, double *__restrict__ c_out,
                       int *__restrict__ k_out, int *__restrict__ scratch) {
    __shared__ int scan_tmp[1024];
    int *ctrl = scratch;
    int *cmap = scratch + kCtrlInts;
    const int tid = threadIdx.x;
    const int per = (k + 1023) / 1024;
    const int lo = min(k, tid * per), hi = min(k, lo + per);
    int cnt = 0;
    for (int j = lo; j < hi; ++j) cnt += (sum[(size_t)j * (d + 1) + d] != 0.0) ? 1 : 0;
    if (cnt) atomicAdd(&ctrl[0], cnt);                          // ctrl[0] was cleared by the host before launch
    scan_tmp[tid] = cnt;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {
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
    for (size_t i = tid; i < (size_t)k0 * d; i += 1024) c_out[i] = 0.0;
    __syncthreads();
    for (size_t e = tid; e < (size_t)k * d; e += 1024) {
        const int j = (int)(e / d), i = (int)(e % d);
        const int m = cmap[j];
        if (m >= 0) {
            const double a = 1.0 / sum[(size_t)j * (d + 1) + d];
            c_out[(size_t)m * d + i] = __dadd_rn(0.0, __dmul_rn(a, sum[(size_t)j * (d + 1) + i]));
        }
    }
    __syncthreads();
    if (metric == AKM_COSINE) {
        for (int j = tid; j < k0; j += 1024) {
            double *c = c_out + (size_t)j * d;
            double s = 0.0;
            for (int i = 0; i < d; ++i) s = __dadd_rn(s, __dmul_rn(c[i], c[i]));
            const double nrm = sqrt(s);
            if (nrm > 0.0) {
                const double inv = 1.0 / nrm;
                for (int i = 0; i < d; ++i) c[i] = __dmul_rn(inv, c[i]);
            }
        }
    }
    __syncthreads();
    if (tid == 0) *k_out = ctrl[0];
}

}  // namespace

extern "C" int akm_update_centroids_prealloc(int device, void *cuda_stream, int metric, int d, const double *sum_dev,
                                             int k0, int k, double *c_out_dev, int *k_out_dev) {
    if (device < 0 || device >= kMaxDevices) return fail(AKM_ERR_BADARG, "bad device ordinal");
    if (metric != AKM_EUCLIDEAN && metric != AKM_COSINE) return fail(AKM_ERR_BADARG, "distanceType must be EUCLIDEAN or COSINE");
    if (d <= 0 || k0 <= 0 || k < 0 || k > k0) return fail(AKM_ERR_BADARG, "need d > 0 and 0 <= k <= k0, k0 > 0");
    if (k0 > AKM_PREALLOC_MAX_K)
        return fail(AKM_ERR_BADARG, "k0 = " + std::to_string(k0) + " exceeds AKM_PREALLOC_MAX_K = " +
                                        std::to_string(AKM_PREALLOC_MAX_K) +
                                        " (the built-in working memory is fixed; use akm_update_centroids instead)");
    if (!sum_dev || !c_out_dev || !k_out_dev) return fail(AKM_ERR_BADARG, "sum_dev / c_out_dev / k_out_dev is null");
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return fail(AKM_ERR_CUDA, std::string("cudaSetDevice: ") + cudaGetErrorString(e));
    int *scratch = nullptr;
    int rc = scratch_for_device(device, &scratch);
    if (rc != AKM_OK) return rc;
    cudaStream_t stream = (cudaStream_t)cuda_stream;
    // the array still holds the previous call's data: clear the part that is accumulated into
    e = cudaMemsetAsync(scratch, 0, kCtrlInts * sizeof(int), stream);
    if (e != cudaSuccess) return fail(AKM_ERR_CUDA, std::string("cudaMemsetAsync: ") + cudaGetErrorString(e));
    prealloc_update_kernel<<<1, 1024, 0, stream>>>(metric, sum_dev, k, k0, d, c_out_dev, k_out_dev, scratch);
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail(AKM_ERR_CUDA, std::string("prealloc_update_kernel launch: ") + cudaGetErrorString(e));
    return AKM_OK;
}
