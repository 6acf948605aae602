// tcgen05 / TMA nearest-centroid search (placeholder until the tensor path lands).
#include "akm_internal.cuh"

bool tensor_supported(akm_handle, int) { return false; }
int tensor_assign(akm_handle h, RowStore &, const double *, const double *, int, int, int32_t *, double *) {
    return akm_fail(h, AKM_ERR_UNSUPPORTED, "tensor engine not built");
}
int tensor_prepare_x16(akm_handle, RowStore &) { return AKM_OK; }
void tensor_free(akm_handle) {}
