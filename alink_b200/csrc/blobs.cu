// Synthetic Gaussian-blob generator (benchmark data of SURVEY 8d; not on the reference's path).
// Counter-based: value(seed, row, col) is a pure function, so shards are independent of the GPU count.
#include "akm_internal.cuh"

__host__ __device__ __forceinline__ uint64_t mix64(uint64_t z) {   // splitmix64 finaliser
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

__device__ __forceinline__ float u01(uint32_t x) { return ((float)(x >> 8) + 0.5f) * (1.0f / 16777216.0f); }

__global__ void blobs_kernel(uint64_t seed, int k_true, float spread, float sigma, const int64_t *__restrict__ rows,
                             int64_t row0, int64_t m, int d, float *__restrict__ out) {
    const int64_t pairs_per_row = (d + 1) / 2;
    int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t total = m * pairs_per_row;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; t < total; t += stride) {
        const int64_t lr = t / pairs_per_row;
        const int c0 = (int)(t % pairs_per_row) * 2;
        const int64_t r = rows ? rows[lr] : row0 + lr;
        const int blob = (int)(mix64(seed ^ mix64((uint64_t)r * 0x2545F4914F6CDD1Dull + 1)) % (uint64_t)k_true);
        const uint64_t hn = mix64(mix64(seed + 0x1234567ull) ^ ((uint64_t)r * 0xD1342543DE82EF95ull + (uint64_t)c0));
        // Box-Muller: two normals from two uniforms
        const float u1 = u01((uint32_t)hn), u2 = u01((uint32_t)(hn >> 32));
        const float rad = sqrtf(-2.0f * __logf(u1));
        float sn, cs;
        __sincosf(6.28318530717958647692f * u2, &sn, &cs);
        const float z[2] = {rad * cs, rad * sn};
#pragma unroll
        for (int e = 0; e < 2; ++e) {
            const int c = c0 + e;
            if (c < d) {
                const uint64_t hc = mix64(mix64(seed + 0x77ull) ^ ((uint64_t)blob * 0x9E3779B97F4A7C15ull + (uint64_t)c));
                const float centre = (2.0f * u01((uint32_t)hc) - 1.0f) * spread;
                out[lr * d + c] = centre + sigma * z[e];
            }
        }
    }
}

int launch_blobs(akm_handle h, uint64_t seed, int k_true, double spread, double sigma, const int64_t *rows_dev,
                 int64_t row0, int64_t m, float *out_dev) {
    if (m <= 0) return AKM_OK;
    blobs_kernel<<<h->num_sms * 16, 256, 0, h->stream>>>(seed, k_true, (float)spread, (float)sigma, rows_dev, row0, m,
                                                         h->d, out_dev);
    h->stats.kernel_launches++;
    AKM_CUDA(h, cudaGetLastError());
    return AKM_OK;
}
