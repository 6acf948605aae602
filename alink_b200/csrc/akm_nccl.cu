// AllReduce(CENTROID_ALL_REDUCE) over NVLink: replaces the two Flink partitionCustom shuffles of
// CORE/common/comqueue/communication/AllReduce.java:85-120 with one ncclAllReduce(double, sum)
// enqueued on the handle's stream (no host round trip inside akm_step).
//
// libnccl is resolved with dlopen at first use: inside a PyTorch process this binds the NCCL that
// torch already loaded, inside a JVM it binds the system libnccl.so.2.  The few declarations below
// restate the public NCCL ABI (nccl.h: ncclUniqueId is 128 opaque bytes, ncclDouble = 8, ncclSum = 0).
#include <dlfcn.h>

#include <mutex>

#include "akm_internal.cuh"

namespace {
typedef struct ncclComm *ncclComm_t;
typedef struct { char internal[AKM_UNIQUE_ID_BYTES]; } ncclUniqueId;
typedef int ncclResult_t;
constexpr int kNcclDouble = 8;
constexpr int kNcclSum = 0;

struct Api {
    void *lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
    std::string why;
};

Api *api() {
    static Api a;
    static std::once_flag once;
    std::call_once(once, [] {
        const char *names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char *n : names) {
            a.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
            if (a.lib) break;
        }
        if (!a.lib) { a.why = std::string("dlopen(libnccl.so.2) failed: ") + dlerror(); return; }
        a.GetUniqueId = (decltype(a.GetUniqueId))dlsym(a.lib, "ncclGetUniqueId");
        a.CommInitRank = (decltype(a.CommInitRank))dlsym(a.lib, "ncclCommInitRank");
        a.AllReduce = (decltype(a.AllReduce))dlsym(a.lib, "ncclAllReduce");
        a.CommDestroy = (decltype(a.CommDestroy))dlsym(a.lib, "ncclCommDestroy");
        a.GetErrorString = (decltype(a.GetErrorString))dlsym(a.lib, "ncclGetErrorString");
        if (!a.GetUniqueId || !a.CommInitRank || !a.AllReduce || !a.CommDestroy) {
            a.why = "libnccl is missing a required symbol";
            a.lib = nullptr;
        }
    });
    return &a;
}

int nccl_fail(akm_handle h, const char *what, ncclResult_t r) {
    Api *a = api();
    std::string m = std::string(what) + ": " + (a->GetErrorString ? a->GetErrorString(r) : "nccl error");
    return akm_fail(h, AKM_ERR_NCCL, m);
}
}  // namespace

int nccl_allreduce_device(akm_handle h, double *dbuf, int64_t count) {
    if (!h->comm || h->world <= 1) return AKM_OK;   // P = 1: identity (AllReduce.java at one subtask)
    Api *a = api();
    ncclResult_t r = a->AllReduce(dbuf, dbuf, (size_t)count, kNcclDouble, kNcclSum, (ncclComm_t)h->comm, h->stream);
    if (r != 0) return nccl_fail(h, "ncclAllReduce", r);
    return AKM_OK;
}

void nccl_destroy(akm_handle h) {
    if (h->comm) {
        Api *a = api();
        if (a->lib) a->CommDestroy((ncclComm_t)h->comm);
        h->comm = nullptr;
    }
    if (h->ar_buf) { cudaFree(h->ar_buf); h->ar_buf = nullptr; h->ar_cap = 0; }
}

extern "C" {

int akm_comm_unique_id(void *id_out) {
    if (!id_out) return akm_fail(nullptr, AKM_ERR_BADARG, "id_out is null");
    Api *a = api();
    if (!a->lib) return akm_fail(nullptr, AKM_ERR_NCCL, a->why);
    ncclUniqueId id;
    ncclResult_t r = a->GetUniqueId(&id);
    if (r != 0) return nccl_fail(nullptr, "ncclGetUniqueId", r);
    memcpy(id_out, &id, sizeof(id));
    return AKM_OK;
}

int akm_comm_init(akm_handle h, const void *id, int rank, int world) {
    if (!h || !id) return akm_fail(h, AKM_ERR_BADARG, "null argument");
    if (world < 1 || rank < 0 || rank >= world) return akm_fail(h, AKM_ERR_BADARG, "bad rank / world size");
    cudaError_t e = cudaSetDevice(h->device);
    if (e != cudaSuccess) return akm_fail(h, AKM_ERR_CUDA, cudaGetErrorString(e));
    nccl_destroy(h);
    h->rank = rank; h->world = world;
    if (world == 1) return AKM_OK;
    Api *a = api();
    if (!a->lib) return akm_fail(h, AKM_ERR_NCCL, a->why);
    ncclUniqueId uid;
    memcpy(&uid, id, sizeof(uid));
    ncclComm_t comm = nullptr;
    ncclResult_t r = a->CommInitRank(&comm, world, uid, rank);
    if (r != 0) return nccl_fail(h, "ncclCommInitRank", r);
    h->comm = comm;
    // NCCL sets its channels up on the first collective (hundreds of milliseconds): do that here, once, instead of
    // inside the first superstep of a training run
    if (h->ar_cap < 8) {
        if (h->ar_buf) cudaFree(h->ar_buf);
        h->ar_buf = nullptr; h->ar_cap = 0;
        AKM_CUDA(h, cudaMalloc((void **)&h->ar_buf, 8 * sizeof(double)));
        h->ar_cap = 8;
    }
    AKM_CUDA(h, cudaMemsetAsync(h->ar_buf, 0, 8 * sizeof(double), h->stream));
    r = a->AllReduce(h->ar_buf, h->ar_buf, 8, kNcclDouble, kNcclSum, comm, h->stream);
    if (r != 0) return nccl_fail(h, "ncclAllReduce (warm-up)", r);
    AKM_CUDA(h, cudaStreamSynchronize(h->stream));
    return AKM_OK;
}

int akm_allreduce_sum(akm_handle h, double *buf, int64_t count) {
    if (!h || !buf || count < 0) return akm_fail(h, AKM_ERR_BADARG, "bad argument");
    cudaError_t e = cudaSetDevice(h->device);
    if (e != cudaSuccess) return akm_fail(h, AKM_ERR_CUDA, cudaGetErrorString(e));
    if (!h->comm || h->world <= 1 || count == 0) return AKM_OK;
    if (h->ar_cap < (size_t)count) {                     // grow-only device staging: no allocation on the per-superstep path
        if (h->ar_buf) cudaFree(h->ar_buf);
        h->ar_buf = nullptr; h->ar_cap = 0;
        AKM_CUDA(h, cudaMalloc((void **)&h->ar_buf, sizeof(double) * (size_t)count));
        h->ar_cap = (size_t)count;
    }
    double *d = h->ar_buf;
    AKM_CUDA(h, cudaMemcpyAsync(d, buf, sizeof(double) * (size_t)count, cudaMemcpyHostToDevice, h->stream));
    AKM_TRY(nccl_allreduce_device(h, d, count));
    AKM_CUDA(h, cudaMemcpyAsync(buf, d, sizeof(double) * (size_t)count, cudaMemcpyDeviceToHost, h->stream));
    AKM_CUDA(h, cudaStreamSynchronize(h->stream));
    return AKM_OK;
}

}  // extern "C"
