// Nearest-centroid search on the 5th-generation tensor cores (tcgen05 + TMEM + TMA), with a
// rigorous per-row certificate and exact re-evaluation of every row the certificate does not cover.
//
//   search_pair_kernel   S[n, j] = ||c_j||^2 - 2 x16_n . c16_j   (EUCLIDEAN)   /   - x16_n . c16_j   (COSINE)
//                        fp16 operands (scaled by powers of two), fp32 accumulation in TMEM, the running
//                        minimum and the set {j : S[n,j] <= min + W_n} fused in the epilogue.
//   W_n                  = 2 * (bound on |S - true score|) + 2 * (bound on Java's own fp64 rounding), from
//                        per-row ||x||, ||x - x16|| and per-pass max ||c||, max ||c - c16|| (Cauchy-Schwarz).
//                        If exactly one centroid is inside the window, Java's argmin (KMeansUtil.java:162-173)
//                        is that centroid -- no tie is possible -- and the row is "certified".
//   refine_candidates    rows with 2..8 candidates: fp64 evaluation of just those centroids; a clear winner
//                        (gap above the fp64 error bound of both evaluations) decides the row.
//   assign_exact (list)  everything else (true ties, > 8 candidates, non-finite data) is evaluated against
//                        all k centroids in the exact Java operation order (kernels_exact.cu).
// The result is therefore identical to the exact engine by construction; the tensor cores only prune.
//
// Data movement: X16 tiles (128 rows x 64 halfs, 128B-swizzled) and centroid tiles arrive by TMA into the shared
// memory of a CTA PAIR; one thread issues cta_group::2 MMAs (M = 256, N = 256) that read both CTAs' operands, so
// every centroid half-tile is fetched from L2 once per pair.  Accumulators are double buffered in TMEM so the
// epilogue of tile t overlaps the MMAs of tile t+1.
#include <cuda.h>
#include <cfloat>
#include <cmath>
#include <cstdio>
#include <cstdlib>

#include "akm_internal.cuh"

// ------------------------------------------------------------------------------------------ PTX helpers
namespace {

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    return ok != 0;
}
// Bounded wait: a protocol bug must end in a trap, never in a hung GPU.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t spins = 0;
    while (!mbar_try_wait(bar, parity)) {
        if (++spins > (1u << 24)) __trap();
    }
}
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
          "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
          "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
// K-major, 128B-swizzled operand tile (rows x 64 halfs, 128 B per row, 8-row atoms of 1024 B):
}  // namespace

// ------------------------------------------------------------------------------------------ constants
constexpr int TILE_M = 128;           // rows per MMA (TMEM lanes)
constexpr int KBLK = 64;              // halfs per 128-byte swizzled row
constexpr int BOX_BYTES = TILE_M * KBLK * 2;   // 16 KB
constexpr int MAX_CAND = 4;            // per column quarter (ring size)
constexpr int WL_CAND = 8;             // per worklist entry
constexpr int WL_BLOCK = 2048;         // worklist slots a warp reserves at a time
constexpr float PAD_SCORE = 1.0e30f;  // score offset of padding centroids (never candidates)

// per-pass constants produced by c16_rows_kernel
struct PassConst {
    float c16_max;     // max_j ||c16_j / scale||   (rounded up)
    float dc_max;      // max_j ||c_j - c16_j / scale||
    float c_max;       // max_j ||c_j||
    float score_scale; // -2 * 2^-(sx+sc) (EUCLIDEAN) or -2^-(sx+sc) (COSINE)
    int sc;            // centroid scale exponent
    int bad;           // non-finite centroid: tensor search disabled for this pass
    float c_dev;       // max_j | ||c_j|| - 1 | (rounded up): the COSINE bounds are chords between UNIT vectors
};

struct SearchParams {
    int64_t n;              // rows
    int ntiles;             // centroid tiles of 256
    int a_stages;           // K-streaming kernel: ring stages
    int kblocks;            // K-streaming kernel: 64-half blocks per row
    int d, metric, k;
    const float *cn;        // ntiles*128 score offsets
    const PassConst *pc;
    const float2 *xerr;     // per row {||x||, ||x - x16||}
    int32_t *labels;
    int32_t *wl_rows;       // worklist: rows that need exact work
    int32_t *wl_cand;       // WL_CAND candidates per entry (-1 = unused, first == -2: all centroids)
    int wl_cap;             // worklist slots allocated (a multiple of WL_BLOCK)
    int32_t *exact_rows;    // exact list: direct destination of rows when the worklist is full
    int *exact_count;
    int *counters;          // [0] reserved worklist slots, [2] labels changed, [3] worklist entries used, [4] tiles produced
    float eta;              // accumulation error factor
    // seeds: a rigorous upper bound of this pass's minimum score per row, derived from the previous pass
    float *useed;           // in: previous pass min score + its error bound; out: same for this pass (nullable)
    const float *sig_old;   // per OLD label: distance its centroid moved (INF if the cluster was dropped)
    int k_prev;             // 0: no seeds this pass
    const float2 *wm;       // per row {recording window (encodes W, see encode_window), seeded upper bound of the
                            // minimum score (INF: none)} from rowprep_kernel
    float2 *lb;             // per row {lower bound of the distance to every centroid but the row's own, r2lo}
                            // (nullable; rowprep_kernel presets .x to the bound for the tiles that are not evaluated)
    const int2 *glist;      // the row groups (256 rows) to search, ascending: {group, mask of the centroid tiles it needs};
    const int *glist_n;     // its length (on the device).  nullptr: every group, every tile
    int dbg;                // builds with -DAKM_EPI_DBG only (option "epi_dbg"; timing experiments, results invalid):
                            // 1 = accumulators read but not evaluated, 2 = not even read (the MMA/TMA floor),
                            // 3 = evaluated but not merged
};

// ------------------------------------------------------------------------------------------ per-row window / seed
// Runs once per pass before the search: everything the epilogue needs per row that lives in global memory
// (norms, previous label -> centroid movement, previous minimum) is folded into two floats, so the search
// kernel's per-group prologue is a single coalesced load.
//   W  = 2.004 * B + 2 * eps_J   (NaN when the centroids are not finite -> every row goes exact)
//   m0 = rigorous upper bound of this pass's minimum approximate score (INF when there is no seed):
//        the previous winner a had true score <= useed; its centroid moved by sig_old[a].
//
// Tile pruning (Elkan's lemma at the granularity the tensor cores work at).  With a seed, ub bounds the distance
// of the row to its previous centroid a at its new position; every centroid j of a 256-centroid tile T is at
// least dmin[a][T] away from c_a, hence at least dmin - ub away from the row.  If
//     dmin * (dmin - 2 ub) > 2 B + W            (EUCLIDEAN; twice that for COSINE, whose score is half the
//                                                squared chord between unit vectors)
// no score of the tile can come within the window of the row's minimum, so the tile holds neither the argmin nor
// a candidate.  A tile is skipped for a 256-row group when that holds for every row of the group; rows of a
// cluster-ordered partition share their previous centroid, so whole groups agree.  gmask[g] = OR of the rows' masks.
//
// Second-nearest bound (Hamerly's test, per row).  A searched row also learns a lower bound lb of its distance to
// every centroid other than its winner: the epilogue records scores up to a window W' >= W above the running minimum
// (only those within W are candidates; the others just bound lb from above), so whatever was not recorded lies more
// than W' above the minimum.  In the next pass every other centroid is at least lb - (largest movement of another
// centroid) away, the own one at most ub: when the gap is positive by more than Java's own rounding of the two
// distances, the row needs no tile at all.  W' = W * 2^e travels in the slot of W with e in the five low mantissa bits (encode_window), so the
// search kernel's per-row state does not grow.
// a window without an exponent must decode to itself: its low bits are cleared (rounding up)
__device__ __forceinline__ float plain_window(float W) {
    if (!(W == W)) return W;
    return __uint_as_float((__float_as_uint(W) + 31u) & ~31u);
}
__device__ __forceinline__ float encode_window(float W, float target) {
    if (!(W == W) || !(W > 1e-20f) || !(target > W) || !(target < 1e30f)) return plain_window(W);
    int e = 0;
    frexpf(target / W, &e);                       // target / W = f * 2^e, f in [0.5, 1)
    e = max(0, min(e - 1, 24));
    const uint32_t bits = ((__float_as_uint(ldexpf(W, e)) + 31u) & ~31u) | (uint32_t)e;
    return __uint_as_float(bits);
}
__device__ __forceinline__ float decode_window(float Wenc) {     // the window W that certifies (>= the encoded one)
    const uint32_t e = __float_as_uint(Wenc) & 31u;
    return Wenc * __uint_as_float((127u - e) << 23);
}

// largest and second largest movement among the surviving old clusters (dropped ones carry INF and no longer compete)
__global__ void __launch_bounds__(256) shift_top2_kernel(const float *__restrict__ sig_old, int k_prev, float *__restrict__ top) {
    __shared__ float s1[256], s2[256];
    __shared__ int sa[256];
    float m1 = 0.f, m2 = 0.f;
    int a1 = -1;
    for (int i = threadIdx.x; i < k_prev; i += 256) {
        const float v = sig_old[i];
        if (!(v < 3.0e38f)) { if (v != v) { m1 = NAN; } continue; }
        if (v > m1) { m2 = m1; m1 = v; a1 = i; } else if (v > m2) m2 = v;
    }
    s1[threadIdx.x] = m1; s2[threadIdx.x] = m2; sa[threadIdx.x] = a1;
    __syncthreads();
    if (threadIdx.x == 0) {
        float b1 = 0.f, b2 = 0.f; int ba = -1; bool nan = false;
        for (int i = 0; i < 256; ++i) {
            if (s1[i] != s1[i]) nan = true;
            if (s1[i] > b1) { b2 = fmaxf(b1, s2[i]); b1 = s1[i]; ba = sa[i]; }
            else b2 = fmaxf(b2, s1[i]);
        }
        top[0] = nan ? NAN : b1; top[1] = nan ? NAN : b2; top[2] = __int_as_float(ba);
    }
}

constexpr int ROWPREP_THREADS = 128;     // two rows of the 256-row group per thread: twice the loads in flight
__global__ void __launch_bounds__(ROWPREP_THREADS)
rowprep_kernel(int64_t n, int d, int metric, float eta, const PassConst *__restrict__ pc,
               const float2 *__restrict__ xerr, int32_t *__restrict__ labels,
               float *__restrict__ useed, const float *__restrict__ sig_old, int k_prev,
               const int *__restrict__ remap, float2 *__restrict__ wm,
               const float *__restrict__ dmin, int ntiles, int k, uint32_t *__restrict__ gmask,
               float2 *__restrict__ lbv, int lb_valid, const float *__restrict__ top) {
    const float c16_max = pc->c16_max, dc_max = pc->dc_max, c_max = pc->c_max;
    const bool bad = pc->bad != 0;
    // Alink's COSINE centroid buffers hold unit vectors (updateLabel normalises them); a host that passes anything
    // else gets Java's arithmetic on it, but no pruning: lemma and bounds are chord geometry on the unit sphere
    const bool geometry_ok = metric == AKM_EUCLIDEAN || pc->c_dev < 1e-12f;
    // movement of the fastest and second fastest surviving centroid (a row excludes its own)
    const float top1 = (lbv && k_prev > 0) ? top[0] : 0.f, top2 = (lbv && k_prev > 0) ? top[1] : 0.f;
    const int top_a = (lbv && k_prev > 0) ? __float_as_int(top[2]) : -1;
    const int64_t ngroups = (n + 255) / 256;
    const uint32_t all_tiles = ntiles >= 32 ? 0xFFFFFFFFu : ((1u << ntiles) - 1u);
    for (int64_t g = blockIdx.x; g < ngroups; g += gridDim.x) {
        uint32_t mask = 0, own = 0;
        // everything the two rows need from global memory first, then the arithmetic
        float2 xe2[2], lb2[2];
        int a2[2];
        float u02[2];
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
            const int64_t row = g * 256 + hh * ROWPREP_THREADS + threadIdx.x;
            xe2[hh] = make_float2(0.f, 0.f); lb2[hh] = make_float2(0.f, 0.f); a2[hh] = -1; u02[hh] = INFINITY;
            if (row < n) {
                xe2[hh] = xerr[row];
                if (k_prev > 0) {
                    a2[hh] = labels[row];
                    u02[hh] = useed[row];
                    if (lbv != nullptr && lb_valid) lb2[hh] = lbv[row];
                }
            }
        }
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
            const int64_t row = g * 256 + hh * ROWPREP_THREADS + threadIdx.x;
            if (row >= n) continue;
            const float r = xe2[hh].x, er = xe2[hh].y;
            const float rc = r + c_max;
            // |S - true| <= 2(|dx||c16| + |x||dc|) + 2 eta |x16||c16| + fp32 epilogue rounding
            const float B = 2.f * (er * c16_max + r * dc_max) + 2.f * eta * (r + er) * c16_max + 4.8e-7f * rc * rc;
            const float epsJ = (float)(d + 16) * 4.5e-16f * rc * rc;    // Java's own fp64 rounding
            float W = 2.004f * B + 2.f * epsJ + 1e-30f;
            if (bad) W = NAN;
            float m0 = INFINITY;
            float Wenc = plain_window(W);
            W = decode_window(Wenc);                        // the value the search certifies with
            // distance^2 (COSINE: chord^2) >= alpha * score + r2lo for a score that is short of the truth by at most W / 2
            // (-INF when the centroids of this pass are not unit vectors under COSINE: the epilogue's bound then comes out 0)
            const float r2lo = (metric == AKM_EUCLIDEAN) ? (r * r * 0.999998f - 0.5f * W) : (geometry_ok ? (2.f - W) : -INFINITY);
            float lb_out = INFINITY;
            uint32_t need = all_tiles;
            const int a = a2[hh];
            if (k_prev > 0 && a >= 0 && a < k_prev) {
                const float sg_a = sig_old[a], u0 = u02[hh];
                float u, ut, ub;                                // ub: distance (COSINE: chord) to c_a, rounded up
                if (metric == AKM_EUCLIDEAN) {
                    const float r2 = r * r;
                    const float t = sqrtf(fmaxf(u0 + r2 * 1.000001f, 0.f)) * 1.000001f + sg_a;
                    ut = t * t * 1.000001f - r2 * 0.999998f + 2e-6f * (t * t + r2);   // bounds the true score
                    u = ut + 0.5f * W + 8e-6f * (t * t + r2);                          // ... and the approximate one
                    ub = t;
                } else {
                    ut = u0 + sg_a * (r + er) * 1.000001f + 2e-7f;
                    u = ut + 0.5f * W + 8e-7f;
                    ub = sqrtf(fmaxf(2.f + 2.f * u, 0.f)) * 1.000001f + 1e-6f;
                    if (!(fabsf(r - 1.f) < 1e-5f)) ub = INFINITY;   // not a unit vector (zero row): no pruning
                }
                if (u == u) m0 = u;                                 // NaN: no seed
                if (!(ut == ut)) ut = INFINITY;
                const int ja = remap ? remap[a] : a;                // this pass's numbering (-1: cluster dropped)
                if (dmin != nullptr && geometry_ok && ja >= 0 && ja < k && u == u && ub < 3.0e37f && W == W) {
                    const float rhs = ((metric == AKM_EUCLIDEAN) ? 2.f : 4.f) * W * 1.01f + 1e-30f;
                    // (dmin of the row's own tile leaves c_a itself out: when that tile passes too, no other centroid
                    //  can take the row)
                    own |= 1u << (ja / 256);
                    need = 0u;
                    // second-nearest bound carried over from the last search of this row.  Nothing is evaluated for
                    // a row that passes, so the margin only has to cover Java's own rounding of the two distances
                    // (the tile lemma's margin also keeps would-be candidates of the fp16 window out of pruned tiles).
                    float lbp = 0.f;
                    if (lbv != nullptr && lb_valid) lbp = lb2[hh].x - ((a == top_a) ? top2 : top1);
                    if (!(lbp < 3.0e37f)) lbp = 0.f;            // NaN, or a bound nobody computed
                    const float rhs_row = 16.f * epsJ + 1e-30f;
                    if (lbp > ub && (lbp - ub) * (lbp + ub) * 0.999999f > rhs_row) {
                        lb_out = lbp * 0.999999f;                      // every other centroid is out of reach
                    } else {
                        float geo = INFINITY;                          // bound for the tiles this row lets go
                        for (int t = 0; t < ntiles; ++t) {
                            const float dm = sqrtf(dmin[(size_t)ja * ntiles + t]) * 0.999999f;    // table holds squares
                            const float lhs = dm * (dm - 2.f * ub) * 0.999999f;
                            if (!(dm > 2.f * ub && lhs > rhs)) need |= 1u << t;
                            else geo = fminf(geo, (dm - ub) * 0.999999f);
                        }
                        lb_out = fmaxf(geo, lbp * 0.999999f);
                    }
                    if (need == 0u) {
                        // The row keeps its centroid whatever the search finds.  If its group is searched for the
                        // sake of other rows it only rides along: a seed of -INF lets no score of its through, the
                        // epilogue leaves its label, seed and bound alone, and it never reaches the refinement.
                        m0 = -INFINITY;
                        useed[row] = ut;
                    } else if (lbv != nullptr) {
                        // record scores up to 1.5 x the distance of the own centroid: the next pass can then tell
                        // that far apart the second nearest is (alpha * window = (1.5^2 - 1) * ub^2)
                        Wenc = encode_window(W, W + ((metric == AKM_EUCLIDEAN) ? 1.25f : 0.625f) * ub * ub);
                    }
                }
                // express the old label in this pass's cluster numbering (empty clusters were compacted away),
                // so the search kernel can tell which rows really changed cluster
                if (remap) labels[row] = remap[a];
            }
            mask |= need;
            wm[row] = make_float2(Wenc, m0);
            // searched rows: the epilogue lowers .x to what it saw; rows that pass keep it for the next pass
            if (lbv != nullptr) lbv[row] = make_float2(lb_out, r2lo);
        }
        if (gmask != nullptr) {
            mask = __reduce_or_sync(0xffffffffu, mask);
            own = __reduce_or_sync(0xffffffffu, own);
            // A group none of whose rows needs any tile is not searched at all (its rows have seen to their seeds and
            // bounds above).  A group that is searched must see the tile of every row's previous centroid: that is
            // where the row's winner is.
            if (__syncthreads_or((int)mask) && (threadIdx.x & 31) == 0 && (mask | own)) atomicOr(gmask + g, mask | own);
        }
    }
}

// The groups that need a search with their tile masks: the search kernel's three roles walk this list, so groups
// that are skipped cost them nothing and every per-group prefetch is one searched group ahead.  A warp looks at 32
// consecutive groups and reserves their slots with one atomic (neighbouring groups stay together; the order of the
// warps' runs is the order they arrive in, which no result depends on).
__global__ void __launch_bounds__(256) group_list_kernel(const uint32_t *__restrict__ gmask, int ngroups, int2 *__restrict__ glist,
                                                         int *__restrict__ count) {
    const int lane = threadIdx.x & 31;
    const int nw = (ngroups + 31) / 32;
    for (int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; w < nw; w += (gridDim.x * blockDim.x) >> 5) {
        const int g = w * 32 + lane;
        const uint32_t m = g < ngroups ? gmask[g] : 0u;
        const unsigned b = __ballot_sync(0xffffffffu, m != 0u);
        if (b == 0u) continue;
        int base = 0;
        if (lane == 0) base = atomicAdd(count, __popc(b));
        base = __shfl_sync(0xffffffffu, base, 0);
        if (m) glist[base + __popc(b & ((1u << lane) - 1u))] = make_int2(g, (int)m);
    }
}

// dmin2[a][T] = min over the centroids j != a of tile T of |c_a - c_j|^2 (COSINE: C holds unit vectors, so this is
// the squared chord), as float rounded down.  32 x 32 (a, j) pairs per CTA, operands staged through shared memory,
// fp64 accumulation; the per-tile minimum is taken with an integer atomicMin on the bit pattern of the non-negative
// float (the table is preset to a huge value).  k*k*d flops, once per pass.
__global__ void __launch_bounds__(256)
centroid_tile_dist_kernel(const double *__restrict__ C, int k, int d, int ntiles, unsigned *__restrict__ dmin2) {
    __shared__ double sa[32][33], sb[32][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;          // 32 x 8
    const int a0 = blockIdx.y * 32, j0 = blockIdx.x * 32;
    double acc[4] = {0.0, 0.0, 0.0, 0.0};
    for (int c0 = 0; c0 < d; c0 += 32) {
        __syncthreads();
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int r = ty + 8 * i, col = c0 + tx;
            sa[r][tx] = (a0 + r < k && col < d) ? C[(size_t)(a0 + r) * d + col] : 0.0;
            sb[r][tx] = (j0 + r < k && col < d) ? C[(size_t)(j0 + r) * d + col] : 0.0;
        }
        __syncthreads();
#pragma unroll 8
        for (int c = 0; c < 32; ++c) {
            const double bj = sb[tx][c];
#pragma unroll
            for (int i = 0; i < 4; ++i) { const double df = sa[ty + 8 * i][c] - bj; acc[i] = fma(df, df, acc[i]); }
        }
    }
    const int j = j0 + tx;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int a = a0 + ty + 8 * i;                     // the same for the whole warp; its 32 j share one tile
        unsigned bits = 0x7F7F7F7Fu;                       // the table's preset: "no other centroid in this tile"
        if (a < k && j < k && a != j && acc[i] == acc[i]) bits = __float_as_uint(fmaxf(__double2float_rd(acc[i] * (1.0 - 1e-9)), 0.f));
        bits = __reduce_min_sync(0xffffffffu, bits);       // non-negative floats order like their bit patterns
        if (tx == 0 && a < k && bits != 0x7F7F7F7Fu) atomicMin(dmin2 + (size_t)a * ntiles + j0 / 256, bits);
    }
}

// ------------------------------------------------------------------------------------------ search kernel
// cta_group::2: one thread of the leader CTA issues M=256 x N=256 x K=16 MMAs that read A and B from both CTAs'
// shared memory and write each CTA's 128 rows x 256 columns into its own TMEM.
//   KB = 1, 2  the whole fp16 centroid matrix fits the shared memory of the pair (each CTA keeps its half of
//              every 256-centroid tile): centroids are loaded once and never move again; per CTA 2 stages of
//              X16 rows (128 rows each) + resident centroids + score offsets + candidate rings.
//   KB = 0     any other shape: rows and centroids stream through a ring of 64-half blocks.
namespace {
__device__ __forceinline__ void tma_load_2d_2sm(uint32_t dst, const CUtensorMap *map, uint32_t leader_bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(dst), "l"(map), "r"(leader_bar), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void tmem_alloc_2sm(uint32_t dst_smem, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_2sm(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void umma_f16_2sm(uint32_t d_tmem, uint32_t a_lo, uint32_t b_lo, uint32_t desc_hi, uint32_t idesc,
                                             uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\t"
        "setp.ne.b32 p, %5, 0;\n\t"
        "mov.b64 da, {%1, %3};\n\t"
        "mov.b64 db, {%2, %3};\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], da, db, %4, p;\n\t}"
        ::"r"(d_tmem), "r"(a_lo), "r"(b_lo), "r"(desc_hi), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void umma_commit_2sm(uint32_t bar, uint16_t mask) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(bar), "h"(mask) : "memory");
}
// arrive on the barrier at the same shared-memory offset in CTA `cta` of the cluster
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t bar, uint32_t cta) {
    uint32_t remote;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(bar), "r"(cta));
    asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(remote) : "memory");
}
// two fp32 FMAs in one instruction (FFMA2): {d0,d1} = {a0,a1} * {s,s} + {c0,c1}
__device__ __forceinline__ void ffma2(float &d0, float &d1, uint32_t a0, uint32_t a1, uint64_t s2, float c0, float c1) {
    uint64_t a, c, d;
    asm("mov.b64 %0, {%1, %2};" : "=l"(a) : "r"(a0), "r"(a1));
    asm("mov.b64 %0, {%1, %2};" : "=l"(c) : "f"(c0), "f"(c1));
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(s2), "l"(c));
    asm("mov.b64 {%0, %1}, %2;" : "=f"(d0), "=f"(d1) : "l"(d));
}
constexpr uint32_t PEER_BIT_MASK = 0xFEFFFFFFu;   // clears the CTA-pair bit: address of the leader CTA's copy
constexpr uint32_t IDESC_F16_256x256 = (1u << 4) | ((256u >> 3) << 17) | ((256u >> 4) << 24);
constexpr int PAIR_N = 256;                       // centroids per tile (128 per CTA)
}  // namespace

// SPLIT (one centroid tile, k <= 128): only two of the four 64-column quarters hold centroids, so the sixteen epilogue
// warps work as two teams of eight -- quarters {0,1} take the even row groups of the pair, quarters {2,3} the odd ones,
// each team with its own accumulator buffer and its own merge -- instead of leaving half the warps idle while the
// per-group merge and its barriers (the cost that dominates at one tile per group) run on the others.
// LB: the epilogue also records how far the runner-up of every row is (second-nearest bound, see rowprep_kernel).
template <int KB, bool SPLIT, bool LB>
__global__ void __launch_bounds__(576, 1)
search_pair_kernel(const __grid_constant__ CUtensorMap tmapX, const __grid_constant__ CUtensorMap tmapC,
                   const SearchParams p) {
    extern __shared__ unsigned char smem_raw[];
    unsigned char *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    // KB > 0: centroids resident, rows staged whole (2 stages).  KB == 0 (STREAM): a ring of p.a_stages stages,
    // each one 64-half block of this CTA's 128 rows + the same block of its 128 centroids of the current tile.
    constexpr bool STREAM = KB == 0;
    constexpr int A_STAGE = STREAM ? 2 * BOX_BYTES : KB * BOX_BYTES;   // 128 rows of this CTA (STREAM: rows + centroids block)
    constexpr int B_TILE = KB * BOX_BYTES;             // this CTA's 128 centroids of a 256-centroid tile
    constexpr int MAX_STAGES = 8;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const bool leader = rank == 0;
    const int pair_id = blockIdx.x >> 1, npairs = gridDim.x >> 1;
    const int nt = p.ntiles;                           // tiles of 256 centroids
    const int ns = STREAM ? p.a_stages : 2;
    const int nkb = STREAM ? p.kblocks : KB;

    unsigned char *a_smem = smem;
    unsigned char *b_smem = smem + (size_t)ns * A_STAGE;
    float *cn_s = reinterpret_cast<float *>(b_smem + (size_t)nt * B_TILE);
    uint64_t *bars = reinterpret_cast<uint64_t *>(cn_s + nt * PAIR_N);
    uint64_t *t_full = bars, *t_empty = bars + 2, *b_full = bars + 4, *a_full = bars + 5, *a_empty = a_full + MAX_STAGES;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(a_empty + MAX_STAGES);
    float *pc_s = reinterpret_cast<float *>(tmem_slot + 2);
    float2 *cand_s = reinterpret_cast<float2 *>(pc_s + 8);                 // [4 quarters][128 rows][MAX_CAND]
    float *xch_s = reinterpret_cast<float *>(cand_s + 4 * TILE_M * MAX_CAND);   // [4 quarters][128 rows][4]
    int *mstage_s = reinterpret_cast<int *>(xch_s + 4 * TILE_M * 4);            // [128 rows][WL_CAND] merged candidates

    if (threadIdx.x == 0) {
        for (int i = 0; i < ns; ++i) { mbar_init(smem_u32(a_full + i), 1); mbar_init(smem_u32(a_empty + i), 1); }
        mbar_init(smem_u32(b_full), 1);
        for (int i = 0; i < 2; ++i) { mbar_init(smem_u32(t_full + i), 1); mbar_init(smem_u32(t_empty + i), SPLIT ? 16 : 32); }
        fence_barrier_init();
        pc_s[0] = p.pc->c16_max; pc_s[1] = p.pc->dc_max; pc_s[2] = p.pc->c_max; pc_s[3] = p.pc->score_scale;
        pc_s[4] = p.pc->bad ? 1.f : 0.f;
    }
    for (int i = threadIdx.x; i < nt * PAIR_N; i += blockDim.x) cn_s[i] = p.cn[i];
    if (warp == 1) tmem_alloc_2sm(smem_u32(tmem_slot), 512);
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    // the launch walks `ng` entries: the groups rowprep_kernel / group_list_kernel found in need of a search, or all
    const int ng = p.glist ? *p.glist_n : (int)((p.n + 2 * TILE_M - 1) / (2 * TILE_M));     // 256 rows per pair and group
    auto entry = [&](int i) -> int2 { return p.glist ? p.glist[i] : make_int2(i, -1); };

    if (warp == 0) {
        // =============================== TMA producer (both CTAs; completion lands on the leader's barriers)
        if (lane == 0) {
            if constexpr (STREAM) {
                int s_i = 0, ph = 0;
                int2 e_next = pair_id < ng ? entry(pair_id) : make_int2(0, 0);
                for (int i = pair_id; i < ng; i += npairs) {
                    const int2 en = e_next;
                    if (i + npairs < ng) e_next = entry(i + npairs);
                    const int64_t row0 = (int64_t)en.x * (2 * TILE_M) + (int64_t)rank * TILE_M;
                    const uint32_t gm = (uint32_t)en.y;
                    for (int t = 0; t < nt; ++t)
                        for (int kb = 0; kb < nkb && ((gm >> t) & 1u); ++kb) {
                            mbar_wait(smem_u32(a_empty + s_i), ph ^ 1);
                            if (leader) mbar_expect_tx(smem_u32(a_full + s_i), 2u * A_STAGE);
                            const uint32_t full_leader = smem_u32(a_full + s_i) & PEER_BIT_MASK;
                            const uint32_t dst = smem_u32(a_smem + (size_t)s_i * A_STAGE);
                            tma_load_2d_2sm(dst, &tmapX, full_leader, kb * KBLK, (int)row0);
                            tma_load_2d_2sm(dst + BOX_BYTES, &tmapC, full_leader, kb * KBLK, t * PAIR_N + (int)rank * TILE_M);
                            if (++s_i == ns) { s_i = 0; ph ^= 1; }
                        }
                }
            } else {
                const uint32_t bfull_leader = smem_u32(b_full) & PEER_BIT_MASK;
                if (leader) mbar_expect_tx(smem_u32(b_full), 2u * (uint32_t)nt * B_TILE);
                for (int t = 0; t < nt; ++t)
                    for (int kb = 0; kb < KB; ++kb)
                        tma_load_2d_2sm(smem_u32(b_smem + (size_t)t * B_TILE + kb * BOX_BYTES), &tmapC, bfull_leader, kb * KBLK,
                                        t * PAIR_N + (int)rank * TILE_M);
                int it = 0;
                int2 e_next = pair_id < ng ? entry(pair_id) : make_int2(0, 0);
                for (int i = pair_id; i < ng; i += npairs, ++it) {
                    const int2 en = e_next;
                    if (i + npairs < ng) e_next = entry(i + npairs);
                    const int as = it & 1;
                    mbar_wait(smem_u32(a_empty + as), ((it >> 1) & 1) ^ 1);
                    if (leader) mbar_expect_tx(smem_u32(a_full + as), 2u * A_STAGE);
                    const uint32_t afull_leader = smem_u32(a_full + as) & PEER_BIT_MASK;
                    const int64_t row0 = (int64_t)en.x * (2 * TILE_M) + (int64_t)rank * TILE_M;
                    for (int kb = 0; kb < KB; ++kb)
                        tma_load_2d_2sm(smem_u32(a_smem + (size_t)as * A_STAGE + kb * BOX_BYTES), &tmapX, afull_leader, kb * KBLK,
                                        (int)row0);
                }
            }
        }
    } else if (warp == 1) {
        // =============================== MMA issuer: one thread of the leader CTA drives both tensor cores
        if (leader) {
            constexpr uint32_t DESC_HI = 64u | (1u << 14) | (2u << 29);
            const uint32_t a_lo0 = ((smem_u32(a_smem) & 0x3FFFFu) >> 4) | (1u << 16);
            if constexpr (STREAM) {
                int s_i = 0, ph = 0, t_it = 0;
                int2 e_next = pair_id < ng ? entry(pair_id) : make_int2(0, 0);
                for (int i = pair_id; i < ng; i += npairs) {
                    const uint32_t gm = (uint32_t)e_next.y;
                    if (i + npairs < ng) e_next = entry(i + npairs);
                    for (int t = 0; t < nt; ++t) {
                        if (!((gm >> t) & 1u)) continue;            // pruned for this group (rowprep_kernel)
                        const int tb = t_it & 1;
                        mbar_wait(smem_u32(t_empty + tb), ((t_it >> 1) & 1) ^ 1);
                        ++t_it;
                        tc_fence_after();
                        const uint32_t d0 = tmem_base + (uint32_t)(tb * PAIR_N);
                        for (int kb = 0; kb < nkb; ++kb) {
                            mbar_wait(smem_u32(a_full + s_i), ph);
                            tc_fence_after();
                            if (lane == 0) {
                                const uint32_t a_lo = a_lo0 + (uint32_t)s_i * (A_STAGE >> 4);
                                const uint32_t b_lo = a_lo + (BOX_BYTES >> 4);
#pragma unroll
                                for (int k4 = 0; k4 < KBLK / 16; ++k4)
                                    umma_f16_2sm(d0, a_lo + 2 * k4, b_lo + 2 * k4, DESC_HI, IDESC_F16_256x256, (kb | k4) != 0);
                                umma_commit_2sm(smem_u32(a_empty + s_i), 3);       // frees the stage in both CTAs
                                if (kb == nkb - 1) umma_commit_2sm(smem_u32(t_full + tb), 3);
                            }
                            __syncwarp();
                            if (++s_i == ns) { s_i = 0; ph ^= 1; }
                        }
                    }
                }
                if (lane == 0 && t_it) atomicAdd(p.counters + 4, t_it);   // accumulator tiles produced by this pair
            } else {
                const uint32_t b_lo0 = ((smem_u32(b_smem) & 0x3FFFFu) >> 4) | (1u << 16);
                mbar_wait(smem_u32(b_full), 0);
                int it = 0, t_it = 0;
                int2 e_next = pair_id < ng ? entry(pair_id) : make_int2(0, 0);
                for (int i = pair_id; i < ng; i += npairs, ++it) {
                    const uint32_t gm = (uint32_t)e_next.y;
                    if (i + npairs < ng) e_next = entry(i + npairs);
                    const int as = it & 1;
                    mbar_wait(smem_u32(a_full + as), (it >> 1) & 1);
                    for (int t = 0; t < nt; ++t) {
                        if (!((gm >> t) & 1u)) continue;            // pruned for this group (rowprep_kernel)
                        // (SPLIT: one tile per entry, so entries -- and with them the teams -- alternate between the buffers)
                        const int tb = t_it & 1;
                        mbar_wait(smem_u32(t_empty + tb), ((t_it >> 1) & 1) ^ 1);
                        ++t_it;
                        tc_fence_after();
                        if (lane == 0) {
                            const uint32_t a_lo = a_lo0 + (uint32_t)as * (A_STAGE >> 4);
                            const uint32_t b_lo = b_lo0 + (uint32_t)t * (B_TILE >> 4);
                            const uint32_t d0 = tmem_base + (uint32_t)(tb * PAIR_N);
#pragma unroll
                            for (int kb = 0; kb < KB; ++kb)
#pragma unroll
                                for (int k4 = 0; k4 < KBLK / 16; ++k4)
                                    umma_f16_2sm(d0, a_lo + (uint32_t)((kb * BOX_BYTES) >> 4) + 2 * k4,
                                                 b_lo + (uint32_t)((kb * BOX_BYTES) >> 4) + 2 * k4, DESC_HI, IDESC_F16_256x256,
                                                 (kb | k4) != 0);
                            umma_commit_2sm(smem_u32(t_full + tb), 3);
                        }
                        __syncwarp();
                    }
                    if (lane == 0) umma_commit_2sm(smem_u32(a_empty + as), 3);
                    __syncwarp();
                }
                if (lane == 0 && t_it) atomicAdd(p.counters + 4, t_it);   // accumulator tiles produced by this pair
            }
        }
    } else {
        // =============================== epilogue: 16 warps per CTA, one thread per (row, 64-column quarter)
        const int e = warp - 2;
        const int quarter = e >> 2;
        const int q = warp & 3;
        const float scale = pc_s[3];
        uint64_t scale2;
        asm("mov.b64 %0, {%1, %1};" : "=l"(scale2) : "f"(scale));
        constexpr int NQ = SPLIT ? 2 : 4;                          // column quarters merged per row
        const int team = SPLIT ? (quarter >> 1) : 0;
        const int qbase = SPLIT ? (quarter & 2) : 0;               // the quarter that owns the merge of this thread's row
        const int colq = SPLIT ? (quarter & 1) : quarter;          // 64-column block of a tile this thread scans
        const int ifirst = pair_id + team * npairs, istep = SPLIT ? 2 * npairs : npairs;   // entries of this thread's team
        const float *cn_q = cn_s + colq * 64;
        const int rt = q * 32 + lane;
        float2 *my_ring = cand_s + ((size_t)quarter * TILE_M + rt) * MAX_CAND;
        float *my_xch = xch_s + ((size_t)quarter * TILE_M + rt) * 4;
        const int bar_id = 1 + q + 4 * team;
        int t_it = 0;
        int wl_pos = 0, wl_end = 0, wl_used = 0;
        float2 wm_next = make_float2(0.f, INFINITY);
        int a_next = -1;
        int2 e_next = ifirst < ng ? entry(ifirst) : make_int2(0, 0);
        {
            const int64_t row0 = (int64_t)e_next.x * (2 * TILE_M) + (int64_t)rank * TILE_M + rt;
            if (ifirst < ng && row0 < p.n) { wm_next = p.wm[row0]; a_next = (p.k_prev > 0 && quarter == qbase) ? p.labels[row0] : -1; }
        }
        for (int i = ifirst; i < ng; i += istep) {
            const int64_t row = (int64_t)e_next.x * (2 * TILE_M) + (int64_t)rank * TILE_M + rt;
            const bool valid = row < p.n;
            const float2 wm = wm_next;
            const int a_old = a_next;
            const uint32_t gm = (uint32_t)e_next.y;
            if (i + istep < ng) {   // prefetch the next entry's two floats per row (and old label) while this one is processed
                e_next = entry(i + istep);
                const int64_t rown = (int64_t)e_next.x * (2 * TILE_M) + (int64_t)rank * TILE_M + rt;
                if (rown < p.n) { wm_next = p.wm[rown]; a_next = (p.k_prev > 0 && quarter == qbase) ? p.labels[rown] : -1; }
                if constexpr (LB) { if (quarter == qbase && rown < p.n) asm volatile("prefetch.global.L2 [%0];" ::"l"(p.lb + rown)); }
            }
            // W: recording window (scores up to W above the running minimum reach the slow path); the certifying
            // window Wc <= W is decoded from it there.  Padding rows: NaN thresholds never admit anything.
            const float W = valid ? wm.x : NAN;
            float m = wm.y;
            float thr = m + W;
            int ncand = 0;
            float dropped = INFINITY;
            if constexpr (LB) my_xch[1] = INFINITY;      // smallest recorded score that was no candidate

            for (int t = 0; t < nt; ++t) {
                if (!((gm >> t) & 1u)) continue;                    // pruned for this group: no accumulator was produced
                // (SPLIT: the pair's groups alternate between the two buffers, and so between the two teams)
                const int tb = SPLIT ? team : (t_it & 1);
                mbar_wait(smem_u32(t_full + tb), SPLIT ? (t_it & 1) : ((t_it >> 1) & 1));
                ++t_it;
                tc_fence_after();
                const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(tb * PAIR_N + colq * 64);
                // columns at or beyond k are padding (score offset PAD_SCORE): a quarter or chunk that holds nothing
                // else is not read at all -- small k (BASELINE configs 2 and 3) leaves most issue slots free
                const int live = p.k - (t * PAIR_N + colq * 64);         // real centroids in this quarter of the tile
#ifdef AKM_EPI_DBG
                if (p.dbg == 1 || p.dbg == 2) {
                    if (p.dbg == 1 && live > 0) {
                        uint32_t v[32];
                        tmem_ld32(taddr, v);
                        tmem_ld_wait();
                        if (v[0] == 0x12345678u && v[17] == 0x9abcdef0u) m = 0.f;
                        if (live > 32) {
                            tmem_ld32(taddr + 32, v);
                            tmem_ld_wait();
                            if (v[1] == 0x12345678u && v[13] == 0x9abcdef0u) m = 0.f;
                        }
                    }
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive_cluster(smem_u32(t_empty + tb), 0);
                    continue;
                }
#endif
                if (live <= 0) {
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive_cluster(smem_u32(t_empty + tb), 0);
                    continue;
                }
#pragma unroll
                for (int c = 0; c < 2; ++c) {
                    uint32_t v[32];
                    __syncwarp();
                    if (c == 1 && live <= 32) {
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) mbar_arrive_cluster(smem_u32(t_empty + tb), 0);
                        break;
                    }
                    tmem_ld32(taddr + c * 32, v);
                    tmem_ld_wait();
                    if (c == 1) {
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) mbar_arrive_cluster(smem_u32(t_empty + tb), 0);   // the leader's MMA thread waits on it
                    }
                    const float *cn_t = cn_q + t * PAIR_N + c * 32;
                    float sc[32];
#pragma unroll
                    for (int gq = 0; gq < 8; ++gq) {
                        const float4 o = *reinterpret_cast<const float4 *>(cn_t + gq * 4);
                        ffma2(sc[gq * 4 + 0], sc[gq * 4 + 1], v[gq * 4 + 0], v[gq * 4 + 1], scale2, o.x, o.y);
                        ffma2(sc[gq * 4 + 2], sc[gq * 4 + 3], v[gq * 4 + 2], v[gq * 4 + 3], scale2, o.z, o.w);
                    }
                    float g8[4];
#pragma unroll
                    for (int gq = 0; gq < 4; ++gq) {
                        const float a1 = fminf(fminf(sc[gq * 8 + 0], sc[gq * 8 + 1]), sc[gq * 8 + 2]);
                        const float a2 = fminf(fminf(sc[gq * 8 + 3], sc[gq * 8 + 4]), sc[gq * 8 + 5]);
                        g8[gq] = fminf(fminf(sc[gq * 8 + 6], sc[gq * 8 + 7]), fminf(a1, a2));
                    }
                    const float cmin = fminf(fminf(g8[0], g8[1]), fminf(g8[2], g8[3]));
                    m = fminf(m, cmin);
                    thr = m + W;                                   // padding rows: W = -INF (set above)
                    if (cmin <= thr) {
                        const int j0 = t * PAIR_N + colq * 64 + c * 32;
                        const float thr_c = LB ? m + decode_window(W) : thr;   // candidates; the rest only bounds the runner-up
                        float far = INFINITY;
                        if constexpr (LB) far = my_xch[1];
#pragma unroll
                        for (int gq = 0; gq < 4; ++gq) {
                            if (g8[gq] <= thr) {
#pragma unroll
                                for (int i = 0; i < 8; ++i) {
                                    const float v = sc[gq * 8 + i];
                                    if (v <= thr_c) {
                                        const int slot = ncand & (MAX_CAND - 1);
                                        if (ncand >= MAX_CAND) dropped = fminf(dropped, my_ring[slot].y);
                                        my_ring[slot] = make_float2(__int_as_float(j0 + gq * 8 + i), v);
                                        ++ncand;
                                    } else if (LB && v <= thr) {
                                        far = fminf(far, v);
                                    }
                                }
                            }
                        }
                        if constexpr (LB) my_xch[1] = far;
                    }
                }
            }
#ifdef AKM_EPI_DBG
            if (p.dbg != 0) continue;
#endif
            // ---- merge the four column quarters of each row (quarter 0 owns the row), then the certificate
            float far_own = INFINITY;
            if constexpr (LB) far_own = my_xch[1];
            if (quarter != qbase) {
                my_xch[0] = m; my_xch[1] = dropped; my_xch[2] = __int_as_float(ncand);
                if constexpr (LB) my_xch[3] = far_own;
            }
            asm volatile("bar.sync %0, %1;" ::"r"(bar_id), "n"(NQ * 32) : "memory");
            if (quarter == qbase) {
                bool need_exact = false, changed = false;
                // merged candidate list of this row: 8 ints in the exchange area of quarter 0 (unused otherwise) + 4 more
                int *stage = reinterpret_cast<int *>(mstage_s + ((size_t)team * TILE_M + rt) * WL_CAND);   // (SPLIT: one area per team)
                if (valid && wm.y != -INFINITY) {           // (-INF: the row keeps its centroid, rowprep_kernel has seen to it)
                    float2 lbr = make_float2(0.f, 0.f);
                    if constexpr (LB) lbr = p.lb[row];
                    float xm[NQ], xd[NQ];
                    int xn[NQ];
                    xm[0] = m; xd[0] = dropped; xn[0] = ncand;
                    float far = far_own;
#pragma unroll
                    for (int qq = 1; qq < NQ; ++qq) {
                        const float4 x4 = *reinterpret_cast<const float4 *>(xch_s + ((size_t)(qbase + qq) * TILE_M + rt) * 4);
                        xm[qq] = x4.x; xd[qq] = x4.y; xn[qq] = __float_as_int(x4.z);
                        if constexpr (LB) far = fminf(far, x4.w);
                    }
                    float mF = fminf(xm[0], xm[1]);
                    if constexpr (NQ == 4) mF = fminf(mF, fminf(xm[2], xm[3]));
                    const float Wc = LB ? decode_window(W) : W;
                    const float thrF = mF + Wc;
                    if constexpr (LB) far = fminf(far, mF + W);   // nothing that was not recorded lies below this
                    bool overflow = false;
                    int cnt = 0, best = -1, first = -1;
                    *reinterpret_cast<int4 *>(stage) = make_int4(-1, -1, -1, -1);
                    *reinterpret_cast<int4 *>(stage + 4) = make_int4(-1, -1, -1, -1);
#pragma unroll
                    for (int qq = 0; qq < NQ; ++qq) {
                        overflow = overflow || !(xd[qq] > thrF);
                        if constexpr (LB) far = fminf(far, xd[qq]);   // (an evicted candidate makes the row exact anyway)
                        const float2 *ring = cand_s + ((size_t)(qbase + qq) * TILE_M + rt) * MAX_CAND;
                        const int lim = min(xn[qq], MAX_CAND);
                        for (int i = 0; i < lim; ++i) {           // rings are nearly always empty or hold one entry
                            const float2 cv = ring[i];
                            if (cv.y <= thrF) {
                                const int j = __float_as_int(cv.x);
                                if (cnt < WL_CAND) stage[cnt] = (j < p.k) ? j : -1;
                                if (cnt == 0) first = j;
                                if (cv.y == mF) best = j;
                                ++cnt;
                            } else if (LB) {
                                far = fminf(far, cv.y);           // was a candidate of an earlier, larger minimum
                            }
                        }
                    }
                    overflow = overflow || cnt > WL_CAND;
                    if (best < 0 && cnt > 0) best = first;
                    const bool all = overflow || best < 0 || best >= p.k;
                    need_exact = all || cnt != 1;
                    if (all) stage[0] = -2;
                    if (p.k_prev > 0 && a_old != best) changed = true;
                    p.labels[row] = best;
                    if (p.useed) p.useed[row] = (best >= 0 && mF < 3.0e38f) ? (mF + 0.5f * Wc) : INFINITY;
                    if constexpr (LB) {
                        // every centroid but the candidates scored at least `far`: as a distance (COSINE: chord), rounded
                        // down; rows that are evaluated against all centroids learn nothing
                        float lbn = 0.f;
                        if (!all && far == far)                 // (several candidates: the refinement adds their exact distances)
                            lbn = sqrtf(fmaxf(((p.metric == AKM_EUCLIDEAN) ? far : 2.f * far) * (far > 0.f ? 0.999999f : 1.000001f) + lbr.y, 0.f)) * 0.999999f;
                        p.lb[row].x = fminf(lbr.x, lbn);
                    }
                }
                const unsigned chm = __ballot_sync(0xffffffffu, changed);
                if (chm && lane == 0) atomicAdd(p.counters + 2, __popc(chm));
                const unsigned mask = __ballot_sync(0xffffffffu, need_exact);
                if (mask) {
                    const int cnt = __popc(mask);
                    if (wl_pos + cnt > wl_end) {
                        for (int i = wl_pos + lane; i < wl_end; i += 32) p.wl_rows[i] = -1;
                        int base = 0;
                        if (lane == 0) base = atomicAdd(p.counters, WL_BLOCK);
                        wl_pos = __shfl_sync(0xffffffffu, base, 0);
                        wl_end = wl_pos + WL_BLOCK;
                    }
                    const int base = wl_pos;
                    wl_pos += cnt;
                    wl_used += cnt;
                    if (need_exact) {
                        const int pos = base + __popc(mask & ((1u << lane) - 1));
                        p.wl_rows[pos] = (int32_t)row;
                        int4 *dst = reinterpret_cast<int4 *>(p.wl_cand + (size_t)pos * WL_CAND);
                        dst[0] = *reinterpret_cast<const int4 *>(stage);
                        dst[1] = *reinterpret_cast<const int4 *>(stage + 4);
                    }
                }
            }
            asm volatile("bar.sync %0, %1;" ::"r"(bar_id), "n"(NQ * 32) : "memory");
        }
        if (quarter == qbase) {
            for (int i = wl_pos + lane; i < wl_end; i += 32) p.wl_rows[i] = -1;
            if (lane == 0 && wl_used) atomicAdd(p.counters + 3, wl_used);
        }
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    if (warp == 1) tmem_dealloc_2sm(tmem_base, 512);
}

// ------------------------------------------------------------------------------------------ operand preparation
__global__ void absmax_kernel_f32(const float *__restrict__ x, int64_t count, unsigned *__restrict__ out_bits) {
    float m = 0.f;
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < count; i += stride) { const float a = fabsf(x[i]); m = (a > m || a != a) ? a : m; }
    for (int o = 16; o > 0; o >>= 1) { const float t = __shfl_xor_sync(0xffffffffu, m, o); m = (t > m || t != t) ? t : m; }
    if ((threadIdx.x & 31) == 0) atomicMax(out_bits, __float_as_uint(m));   // NaN bits compare above +inf
}
__global__ void absmax_kernel_f64(const double *__restrict__ x, int64_t count, unsigned *__restrict__ out_bits) {
    float m = 0.f;
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < count; i += stride) {
        const double ad = fabs(x[i]);
        float a = __double2float_ru(ad);
        m = (a > m || a != a) ? a : m;
    }
    for (int o = 16; o > 0; o >>= 1) { const float t = __shfl_xor_sync(0xffffffffu, m, o); m = (t > m || t != t) ? t : m; }
    if ((threadIdx.x & 31) == 0) atomicMax(out_bits, __float_as_uint(m));
}

// one warp per row: x16 = fp16(x * 2^sx), xerr = {||x||, ||x - x16 * 2^-sx||} (rounded up)
// (xnorm != nullptr: COSINE rows stored raw, the operand is built from x * (1.0 / |x|))
template <typename T>
__global__ void x16_build_kernel(const T *__restrict__ X, const double *__restrict__ xnorm, int64_t n, int d, int d16, int sx,
                                 __half *__restrict__ x16, float2 *__restrict__ xerr) {
    const int lane = threadIdx.x & 31;
    int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t stride = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const double up = ldexp(1.0, sx), down = ldexp(1.0, -sx);
    for (; row < n; row += stride) {
        double s2 = 0.0, e2 = 0.0;
        const double rinv = xnorm ? cos_inv(xnorm[row]) : 1.0;
        for (int c = lane; c < d16; c += 32) {
            __half hv = __float2half_rn(0.f);
            if (c < d) {
                double v = (double)X[row * d + c];
                if (xnorm) v = __dmul_rn(rinv, v);
                hv = __double2half(v * up);
                const double back = (double)__half2float(hv) * down;
                const double dl = v - back;
                s2 = fma(v, v, s2);
                e2 = fma(dl, dl, e2);
            }
            x16[row * d16 + c] = hv;
        }
        for (int o = 16; o > 0; o >>= 1) {
            s2 += __shfl_xor_sync(0xffffffffu, s2, o);
            e2 += __shfl_xor_sync(0xffffffffu, e2, o);
        }
        if (lane == 0) {
            float2 o;
            o.x = __double2float_ru(sqrt(s2) * (1.0 + 1e-9));
            o.y = __double2float_ru(sqrt(e2) * (1.0 + 1e-9));
            xerr[row] = o;
        }
    }
}

// fp16 centroid operand (scaled), score offsets and per-pass constants, in two multi-CTA kernels:
// c16_max_kernel finds max |c| (-> scale exponent), c16_rows_kernel converts one centroid row per warp,
// records the norms the error bound needs (atomic max on the float bit patterns; PassConst is zeroed
// first) and the distance each OLD label's centroid moved since the previous pass.
__global__ void c16_max_kernel(const double *__restrict__ C, size_t count, unsigned *__restrict__ maxbits) {
    float m = 0.f;
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (; i < count; i += stride) {
        const float a = __double2float_ru(fabs(C[i]));
        m = (a > m || a != a) ? a : m;
    }
    for (int o = 16; o > 0; o >>= 1) { const float t = __shfl_xor_sync(0xffffffffu, m, o); m = (t > m || t != t) ? t : m; }
    if ((threadIdx.x & 31) == 0) atomicMax(maxbits, __float_as_uint(m));   // NaN bit patterns sort above +inf
}

__global__ void __launch_bounds__(256)
c16_rows_kernel(const double *__restrict__ C, const double *__restrict__ clabel, int k, int kpad, int d, int d16,
                int metric, int sx, const unsigned *__restrict__ maxbits, __half *__restrict__ c16,
                float *__restrict__ cn, PassConst *__restrict__ pc, const double *__restrict__ Cprev,
                const int *__restrict__ remap, int k_prev, float *__restrict__ sig_old) {
    const int lane = threadIdx.x & 31;
    const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const float mm = __uint_as_float(*maxbits);
    const bool bad = !(mm < 1.0e9f);                      // non-finite or absurdly large: the search is skipped
    int e = 0;
    if (!bad && mm > 0.f) frexpf(mm, &e);                 // mm = f * 2^e, f in [0.5, 1)
    const int sc = (bad || !(mm > 0.f)) ? 0 : 14 - e;     // max |c| * 2^sc in [2^13, 2^14)
    if (w == 0 && lane == 0) {
        const float sgn = (metric == AKM_EUCLIDEAN) ? -2.f : -1.f;
        pc->score_scale = sgn * ldexpf(1.f, -(sx + sc));
        pc->sc = sc;
        if (bad) pc->bad = 1;                             // (PassConst is zeroed before the launch)
    }
    if (w < kpad) {
        const int j = w;
        const double up = ldexp(1.0, sc), down = ldexp(1.0, -sc);
        double s2 = 0.0, e2 = 0.0, t2 = 0.0;
        for (int c = lane; c < d16; c += 32) {
            __half hv = __float2half_rn(0.f);
            if (j < k && c < d && !bad) {
                const double v = C[(size_t)j * d + c];
                hv = __double2half(v * up);
                const double back = (double)__half2float(hv) * down;
                const double dl = v - back;
                s2 = fma(v, v, s2); e2 = fma(dl, dl, e2); t2 = fma(back, back, t2);
            }
            c16[(size_t)j * d16 + c] = hv;
        }
        for (int o = 16; o > 0; o >>= 1) {
            s2 += __shfl_xor_sync(0xffffffffu, s2, o);
            e2 += __shfl_xor_sync(0xffffffffu, e2, o);
            t2 += __shfl_xor_sync(0xffffffffu, t2, o);
        }
        if (lane == 0) {
            if (j < k) {
                cn[j] = (metric == AKM_EUCLIDEAN) ? (float)clabel[j] : 0.f;
                atomicMax(reinterpret_cast<unsigned *>(&pc->c_max), __float_as_uint(__double2float_ru(sqrt(s2) * (1.0 + 1e-9))));
                atomicMax(reinterpret_cast<unsigned *>(&pc->dc_max), __float_as_uint(__double2float_ru(sqrt(e2) * (1.0 + 1e-9))));
                atomicMax(reinterpret_cast<unsigned *>(&pc->c16_max), __float_as_uint(__double2float_ru(sqrt(t2) * (1.0 + 1e-9))));
                atomicMax(reinterpret_cast<unsigned *>(&pc->c_dev), __float_as_uint(__double2float_ru(fabs(sqrt(s2) - 1.0))));
            } else {
                cn[j] = PAD_SCORE;
            }
        }
    } else if (w < kpad + k_prev) {
        // how far did the centroid of OLD label a move? (remap: old index -> new index, -1 = dropped)
        const int a = w - kpad;
        const int j = remap ? remap[a] : a;
        double s2 = 0.0;
        if (j >= 0 && j < k)
            for (int c = lane; c < d; c += 32) { const double df = C[(size_t)j * d + c] - Cprev[(size_t)a * d + c]; s2 = fma(df, df, s2); }
        for (int o = 16; o > 0; o >>= 1) s2 += __shfl_xor_sync(0xffffffffu, s2, o);
        if (lane == 0) sig_old[a] = (j >= 0 && j < k) ? __double2float_ru(sqrt(s2) * (1.0 + 1e-9)) : INFINITY;
    }
}

// ------------------------------------------------------------------------------------------ refinement of 2..4 candidates
// One warp per worklist entry.  fp64 FMA evaluation of the candidate scores; if the best is separated from the
// runner-up by more than the combined rounding bound of this evaluation and of Java's, it is Java's argmin.
// Otherwise (or when the entry asks for all centroids) the row goes to the exact list.
// One THREAD per worklist entry: the work per entry is tiny (a few 128-long fp64 dot products), so the
// kernel is all about memory behaviour.  A thread streams its own sample row and its candidates' centroid
// rows 16-32 bytes at a time; neighbouring entries come from neighbouring rows of a cluster-ordered
// partition, so they share candidates and the centroid rows are L1 hits for the rest of the warp.
#ifndef RF_MINB
#define RF_MINB 6            // 72 registers: the kernel is latency-bound, resident warps matter more than spill-free wide entries
#endif
template <typename T, int METRIC>
__global__ void __launch_bounds__(128, RF_MINB)
refine_candidates_kernel(const T *__restrict__ X, const double *__restrict__ xlabel, int d,
                         const double *__restrict__ C, const double *__restrict__ clabel, int k,
                         const int32_t *__restrict__ wl_rows, const int32_t *__restrict__ wl_cand,
                         const int *__restrict__ counters, int wl_cap, int32_t *__restrict__ labels,
                         int32_t *__restrict__ exact_rows, int *__restrict__ exact_count,
                         float *__restrict__ useed, float2 *__restrict__ lbv) {
    const int nwl = min(counters[0], wl_cap);             // blocks reserved past the capacity were never written
    int e = blockIdx.x * blockDim.x + threadIdx.x;
    const int stride = gridDim.x * blockDim.x;
    for (; e < nwl; e += stride) {
        const int row = wl_rows[e];
        if (row < 0) continue;                           // unused slot of a reserved block
        const int4 c0 = *reinterpret_cast<const int4 *>(wl_cand + (size_t)e * WL_CAND);
        const int4 c1 = *reinterpret_cast<const int4 *>(wl_cand + (size_t)e * WL_CAND + 4);
        const int cand[WL_CAND] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
        bool to_exact = cand[0] == -2;
        int winner = -1;
        if (!to_exact) {
            int nc = 0;
#pragma unroll
            for (int i = 0; i < WL_CAND; ++i) nc += (cand[i] >= 0) ? 1 : 0;
            const T *x = X + (int64_t)row * d;
            double acc[WL_CAND];
#pragma unroll
            for (int i = 0; i < WL_CAND; ++i) acc[i] = 0.0;
            const bool wide = nc > 3;
            const double rinv = RawCosine<T, METRIC>::value ? cos_inv(xlabel[row]) : 1.0;
            // one full 32-byte sector of the sample row per step (each sector is fetched exactly once even when
            // the L1 cannot hold the lines of all resident warps)
            constexpr int STEP = 32 / (int)sizeof(T);
            const bool vec = (d % STEP) == 0;            // rows and centroid rows are then 32-byte aligned
            for (int c = 0; c < d; c += STEP) {
                double xv[STEP];
                if (vec) {
                    // 128-bit loads: a warp-wide scalar load of 32 different rows costs one L1 tag cycle per
                    // row, so the number of load instructions (not bytes) is what bounds this kernel
                    const int4 *xp = reinterpret_cast<const int4 *>(x + c);
                    const int4 r0 = xp[0], r1 = xp[1];
                    if (sizeof(T) == 4) {
                        xv[0] = __int_as_float(r0.x); xv[1] = __int_as_float(r0.y); xv[2] = __int_as_float(r0.z); xv[3] = __int_as_float(r0.w);
                        xv[4 % STEP] = __int_as_float(r1.x); xv[5 % STEP] = __int_as_float(r1.y);
                        xv[6 % STEP] = __int_as_float(r1.z); xv[7 % STEP] = __int_as_float(r1.w);
                    } else {
                        xv[0] = __hiloint2double(r0.y, r0.x); xv[1] = __hiloint2double(r0.w, r0.z);
                        xv[2 % STEP] = __hiloint2double(r1.y, r1.x); xv[3 % STEP] = __hiloint2double(r1.w, r1.z);
                    }
                } else {
#pragma unroll
                    for (int u = 0; u < STEP; ++u) xv[u] = (c + u < d) ? (double)x[c + u] : 0.0;
                }
                if (RawCosine<T, METRIC>::value) {
#pragma unroll
                    for (int u = 0; u < STEP; ++u) xv[u] = __dmul_rn(rinv, xv[u]);
                }
#pragma unroll
                for (int i = 0; i < WL_CAND; ++i) {
                    if (i >= 3 && !wide) break;
                    if (cand[i] >= 0) {
                        const double *cr = C + (size_t)cand[i] * d + c;
                        if (vec) {
#pragma unroll
                            for (int u = 0; u < STEP; u += 2) {
                                const double2 cv = *reinterpret_cast<const double2 *>(cr + u);
                                acc[i] = fma(xv[u], cv.x, acc[i]);
                                acc[i] = fma(xv[u + 1], cv.y, acc[i]);
                            }
                        } else {
#pragma unroll
                            for (int u = 0; u < STEP; ++u) if (c + u < d) acc[i] = fma(xv[u], cr[u], acc[i]);
                        }
                    }
                }
            }
            const double xl = (METRIC == AKM_EUCLIDEAN) ? xlabel[row] : 1.0;
            double v1 = DBL_MAX, v2 = DBL_MAX, clmax = 0.0;
            int i1 = -1;
#pragma unroll
            for (int i = 0; i < WL_CAND; ++i) {
                if (cand[i] < 0) continue;
                const double cl = (METRIC == AKM_EUCLIDEAN) ? clabel[cand[i]] : 1.0;
                clmax = fmax(clmax, cl);
                const double v = (METRIC == AKM_EUCLIDEAN) ? (xl + cl - 2.0 * acc[i]) : (1.0 - acc[i]);
                if (v < v1 || (v == v1 && cand[i] < i1)) { v2 = v1; v1 = v; i1 = cand[i]; }
                else if (v < v2) v2 = v;
            }
            const double rc = sqrt(fabs(xl)) + sqrt(clmax);
            const double tau = 8.0 * (double)(d + 16) * 2.220446049250313e-16 * rc * rc;
            if (i1 >= 0 && (v2 - v1) > tau) winner = i1;   // false for NaN
            else to_exact = true;
            if (lbv != nullptr && !to_exact) {
                // The exact values tighten what the next pass knows about this row: the seed (an upper bound of the
                // winner's score, |c|^2 - 2 x.c or -x.c) and the second-nearest bound, which the search set from
                // everything that was no candidate and which the losing candidates now lower.
                const double s1 = (METRIC == AKM_EUCLIDEAN) ? (v1 - xl) : (v1 - 1.0);
                useed[row] = __double2float_ru(s1 + tau + 1e-7 * fabs(s1));
                const double d2 = fmax(((METRIC == AKM_EUCLIDEAN) ? 1.0 : 2.0) * (v2 - tau), 0.0);
                const float lb2 = (v2 < DBL_MAX) ? __double2float_rd(sqrt(d2) * (1.0 - 1e-7)) : INFINITY;
                lbv[row].x = fminf(lbv[row].x, lb2);
            }
        }
        if (to_exact && lbv != nullptr) lbv[row].x = 0.f;     // the exact list may pick any centroid
        if (to_exact) exact_rows[atomicAdd(exact_count, 1)] = row;
        else labels[row] = winner;
    }
}

// distance of each row to its final centroid in the exact Java order (predictionDistanceCol).  One thread per
// row; the dot product is summed strictly in index order (separate multiply and add, like F2J dgemv), but the
// operands arrive 32 bytes at a time with 128-bit loads (a thread-per-row kernel is bound by the number of load
// instructions, not by bytes).  Neighbouring rows of a cluster-ordered partition share the centroid row.
template <typename T, int METRIC>
__global__ void __launch_bounds__(128)
winner_distance_kernel(const T *__restrict__ X, const double *__restrict__ xlabel, int64_t n, int d,
                       const double *__restrict__ C, const double *__restrict__ clabel,
                       const int32_t *__restrict__ labels, double *__restrict__ mind) {
    int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    constexpr int STEP = 32 / (int)sizeof(T);
    const bool vec = (d % STEP) == 0;                    // rows 32-byte aligned, centroid rows 16-byte aligned
    for (; row < n; row += stride) {
        const int j = labels[row];
        if (j < 0) { mind[row] = nan(""); continue; }
        const T *x = X + row * d;
        const double *c = C + (size_t)j * d;
        const double rinv = RawCosine<T, METRIC>::value ? cos_inv(xlabel[row]) : 1.0;
        double t = 0.0;
        if (vec) {
            for (int i = 0; i < d; i += STEP) {
                const int4 r0 = reinterpret_cast<const int4 *>(x + i)[0], r1 = reinterpret_cast<const int4 *>(x + i)[1];
                double xv[STEP];
                if (sizeof(T) == 4) {
                    xv[0] = __int_as_float(r0.x); xv[1] = __int_as_float(r0.y); xv[2] = __int_as_float(r0.z); xv[3] = __int_as_float(r0.w);
                    xv[4 % STEP] = __int_as_float(r1.x); xv[5 % STEP] = __int_as_float(r1.y);
                    xv[6 % STEP] = __int_as_float(r1.z); xv[7 % STEP] = __int_as_float(r1.w);
                } else {
                    xv[0] = __hiloint2double(r0.y, r0.x); xv[1] = __hiloint2double(r0.w, r0.z);
                    xv[2 % STEP] = __hiloint2double(r1.y, r1.x); xv[3 % STEP] = __hiloint2double(r1.w, r1.z);
                }
#pragma unroll
                for (int u = 0; u < STEP; u += 2) {
                    const double2 cv = *reinterpret_cast<const double2 *>(c + i + u);
                    const double a0 = RawCosine<T, METRIC>::value ? __dmul_rn(rinv, xv[u]) : xv[u];
                    const double a1 = RawCosine<T, METRIC>::value ? __dmul_rn(rinv, xv[u + 1]) : xv[u + 1];
                    t = __dadd_rn(t, __dmul_rn(cv.x, a0));
                    t = __dadd_rn(t, __dmul_rn(cv.y, a1));
                }
            }
        } else {
            for (int i = 0; i < d; ++i) {
                const double a = RawCosine<T, METRIC>::value ? __dmul_rn(rinv, (double)x[i]) : (double)x[i];
                t = __dadd_rn(t, __dmul_rn(c[i], a));
            }
        }
        double D;
        if (METRIC == AKM_EUCLIDEAN) {
            const double y = __dadd_rn(0.0, __dmul_rn(-2.0, t));
            D = sqrt(fabs(__dadd_rn(__dadd_rn(y, xlabel[row]), clabel[j])));
        } else {
            D = __dadd_rn(1.0, __dmul_rn(-1.0, t));
        }
        mind[row] = D;
    }
}

// ------------------------------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

struct TensorCtx {
    EncodeTiledFn encode = nullptr;
    __half *c16 = nullptr; float *cn = nullptr; int kpad_cap = 0;
    PassConst *pc = nullptr;
    int32_t *wl_rows = nullptr, *wl_cand = nullptr, *exact_rows = nullptr; int64_t wl_cap = 0, wl_slots = 0;
    int *counters = nullptr;     // [0] worklist, [1] exact list
    unsigned *absmax = nullptr;
    float2 *wm = nullptr; int64_t wm_cap = 0;
    // tile pruning: per-centroid distance to the nearest other centroid of every tile, per-group tile masks
    float *dmin = nullptr; size_t dmin_cap = 0;
    uint32_t *gmask = nullptr; int2 *glist = nullptr; int64_t gmask_cap = 0;
    // seeds (only for the cached training partition)
    float *useed = nullptr; int64_t useed_cap = 0;
    float2 *lb = nullptr; bool lb_valid = false;     // second-nearest bounds (same capacity as useed), see rowprep_kernel
    float *top = nullptr;                            // shift_top2_kernel's three values
    double *Cprev = nullptr; float *sig_old = nullptr; int prev_cap = 0;
    int k_prev = 0; int64_t seed_rows = -1; uint64_t seed_x16 = 0;
    int chain = 0, step_prev = -1;
};

static TensorCtx *ctx_of(akm_handle h) {
    if (!h->tc) h->tc = new TensorCtx();
    return (TensorCtx *)h->tc;
}

void tensor_free(akm_handle h) {
    TensorCtx *c = (TensorCtx *)h->tc;
    if (!c) return;
    cudaFree(c->c16); cudaFree(c->cn); cudaFree(c->pc); cudaFree(c->wl_rows); cudaFree(c->wl_cand);
    cudaFree(c->exact_rows); cudaFree(c->counters); cudaFree(c->absmax);
    cudaFree(c->useed); cudaFree(c->Cprev); cudaFree(c->sig_old); cudaFree(c->wm); cudaFree(c->dmin); cudaFree(c->gmask);
    cudaFree(c->lb); cudaFree(c->top); cudaFree(c->glist);
    delete c;
    h->tc = nullptr;
}

static int round_up(int a, int b) { return (a + b - 1) / b * b; }

bool tensor_supported(akm_handle h, int k0) {
    if (h->d > 8192) return false;            // K is streamed in 64-half blocks; the bound keeps eta small
    if (k0 > 8064) return false;              // score offsets live in shared memory
    if (h->smem_optin < 200 * 1024) return false;
    return true;
}

static int encode_2d(akm_handle h, TensorCtx *c, CUtensorMap *map, void *base, uint64_t cols, uint64_t rows,
                     uint64_t row_stride_bytes, uint32_t box_cols, uint32_t box_rows) {
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {row_stride_bytes};
    cuuint32_t box[2] = {box_cols, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = c->encode(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, base, dims, strides, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return akm_fail(h, AKM_ERR_CUDA, "cuTensorMapEncodeTiled failed (" + std::to_string((int)r) + ")");
    return AKM_OK;
}

int tensor_prepare_x16(akm_handle h, RowStore &rs) {
    TensorCtx *c = ctx_of(h);
    if (rs.x16_valid) return AKM_OK;
    const int d = h->d, d16 = round_up(d, 8);
    if (!c->absmax) AKM_CUDA(h, cudaMalloc((void **)&c->absmax, sizeof(unsigned)));
    AKM_CUDA(h, cudaMemsetAsync(c->absmax, 0, sizeof(unsigned), h->stream));
    const int64_t count = rs.n * (int64_t)d;
    if (rs.store == 4) absmax_kernel_f32<<<h->num_sms * 8, 256, 0, h->stream>>>(rs.x32, count, c->absmax);
    else absmax_kernel_f64<<<h->num_sms * 8, 256, 0, h->stream>>>(rs.x64, count, c->absmax);
    h->stats.kernel_launches++;
    unsigned bits = 0;
    AKM_CUDA(h, cudaMemcpyAsync(&bits, c->absmax, sizeof(unsigned), cudaMemcpyDeviceToHost, h->stream));
    AKM_CUDA(h, cudaStreamSynchronize(h->stream));
    float mx;
    memcpy(&mx, &bits, sizeof(float));
    if (!(mx < 3.0e38f)) { rs.x16_shift = INT_MIN; rs.x16_valid = true; return AKM_OK; }   // non-finite data: exact engine only
    const bool raw_cosine = h->metric == AKM_COSINE && rs.store == 4;
    if (raw_cosine && mx > 0.f) mx = 1.0f;               // the operand is built from the normalised values, |x_hat_i| <= 1
    int e = 0;
    if (mx > 0.f) frexpf(mx, &e);
    rs.x16_shift = (mx > 0.f) ? 14 - e : 0;
    if (rs.d16 != d16 || !rs.x16 || rs.x16_cap < rs.n) {
        if (rs.x16_alloc) cudaFree(rs.x16_alloc);
        if (rs.xerr) cudaFree(rs.xerr);
        rs.x16 = nullptr; rs.x16_alloc = nullptr; rs.xerr = nullptr;
        const int64_t cap = std::max<int64_t>(rs.cap, 1);
        const int64_t head = (&rs == &h->rows) ? reorder_head_rows(h, rs, rs.store, cap) : 0;   // like the exact copy
        AKM_CUDA(h, cudaMalloc((void **)&rs.x16_alloc, sizeof(__half) * (size_t)(cap + head) * d16));
        AKM_CUDA(h, cudaMalloc((void **)&rs.xerr, sizeof(float2) * (size_t)cap));
        rs.x16 = rs.x16_alloc; rs.x16_head = head; rs.x16_off = 0;
        rs.d16 = d16;
        rs.x16_cap = cap;
        rs.want16 = true;
    }
    rs.x16_gen++;
    const int blocks = (int)std::min<int64_t>((rs.n + 7) / 8, (int64_t)h->num_sms * 32);
    if (rs.store == 4) x16_build_kernel<float><<<blocks, 256, 0, h->stream>>>(rs.x32, raw_cosine ? rs.xlabel : nullptr, rs.n, d, d16,
                                                                               rs.x16_shift, rs.x16, rs.xerr);
    else x16_build_kernel<double><<<blocks, 256, 0, h->stream>>>(rs.x64, nullptr, rs.n, d, d16, rs.x16_shift, rs.x16, rs.xerr);
    h->stats.kernel_launches++;
    AKM_CUDA(h, cudaGetLastError());
    rs.x16_valid = true;
    return AKM_OK;
}

float *tensor_useed(akm_handle h) {
    TensorCtx *c = (TensorCtx *)h->tc;
    return (c && c->k_prev > 0) ? c->useed : nullptr;
}

void tensor_invalidate_seeds(akm_handle h) {
    TensorCtx *c = (TensorCtx *)h->tc;
    if (c) { c->k_prev = 0; c->chain = 0; c->lb_valid = false; }
}

// the rows were regrouped: labels and seeds moved with them, the second-nearest bounds are rebuilt by the next search
void tensor_rows_permuted(akm_handle h) {
    TensorCtx *c = (TensorCtx *)h->tc;
    if (c) c->lb_valid = false;
}

// chain: 0 = stand-alone pass (no seeds), 1 = akm_step sequence (remap = compaction map of the last update),
//        2 = sequence of host-buffer passes (identity map, only while k is unchanged)
template <int KB, bool SPLIT, bool LB>
static int launch_pair(akm_handle h, const CUtensorMap &mx, const CUtensorMap &mc, const SearchParams &sp, size_t smem, int grid) {
    auto kern = search_pair_kernel<KB, SPLIT, LB>;
    AKM_CUDA(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(576);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = h->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    AKM_CUDA(h, cudaLaunchKernelEx(&cfg, kern, mx, mc, sp));
    h->stats.kernel_launches++;
    return AKM_OK;
}

// Every buffer of the search whose size follows from the partition (rows) and the centroid buffer (k0): allocated
// here so that akm_set_centroids can do it before the first superstep (tensor_warm) instead of inside it.
static int ensure_tensor_buffers(akm_handle h, TensorCtx *c, RowStore &rs, int k0, int kpad, bool train_rows, bool *fresh_seeds) {
    const int d = h->d, d16 = rs.d16;
    if (c->kpad_cap < kpad) {
        cudaFree(c->c16); cudaFree(c->cn);
        c->c16 = nullptr; c->cn = nullptr;
        AKM_CUDA(h, cudaMalloc((void **)&c->c16, sizeof(__half) * (size_t)kpad * d16));
        AKM_CUDA(h, cudaMalloc((void **)&c->cn, sizeof(float) * (size_t)kpad));
        c->kpad_cap = kpad;
    }
    if (!c->pc) AKM_CUDA(h, cudaMalloc((void **)&c->pc, sizeof(PassConst)));
    if (!c->counters) AKM_CUDA(h, cudaMalloc((void **)&c->counters, 8 * sizeof(int)));
    if (c->wl_cap < rs.n) {
        cudaFree(c->wl_rows); cudaFree(c->wl_cand); cudaFree(c->exact_rows);
        c->wl_rows = c->wl_cand = c->exact_rows = nullptr;
        // worst case: every row needs exact work; a warp abandons up to 31 slots of each 2048-slot block (a ballot that
        // does not fit starts a new block) and leaves one partial block at the end (at most 8 reserving warps per CTA)
        size_t slots = (size_t)rs.n + (size_t)rs.n / 64 + (size_t)h->num_sms * 16 * WL_BLOCK;
        slots = (slots + WL_BLOCK - 1) / WL_BLOCK * WL_BLOCK;
        c->wl_slots = (int64_t)std::min<size_t>(slots, (((size_t)1 << 31) - ((size_t)1 << 23)) / WL_BLOCK * WL_BLOCK);   // int counter, + slack for one last reservation per warp
        slots = (size_t)c->wl_slots;
        AKM_CUDA(h, cudaMalloc((void **)&c->wl_rows, sizeof(int32_t) * slots));
        AKM_CUDA(h, cudaMalloc((void **)&c->wl_cand, sizeof(int32_t) * slots * WL_CAND));
        AKM_CUDA(h, cudaMalloc((void **)&c->exact_rows, sizeof(int32_t) * (size_t)rs.n));
        c->wl_cap = rs.n;
    }
    if (c->wm_cap < rs.n) {
        cudaFree(c->wm); c->wm = nullptr;
        AKM_CUDA(h, cudaMalloc((void **)&c->wm, sizeof(float2) * (size_t)rs.n));
        c->wm_cap = rs.n;
    }
    if (train_rows) {
        if (c->useed_cap < rs.n) {
            cudaFree(c->useed); c->useed = nullptr;
            AKM_CUDA(h, cudaMalloc((void **)&c->useed, sizeof(float) * (size_t)rs.n));
            cudaFree(c->lb); c->lb = nullptr; c->lb_valid = false;
            AKM_CUDA(h, cudaMalloc((void **)&c->lb, sizeof(float2) * (size_t)rs.n));
            c->useed_cap = rs.n;
            *fresh_seeds = true;
        }
        if (!c->top) AKM_CUDA(h, cudaMalloc((void **)&c->top, 4 * sizeof(float)));
        // pruning: tile distance table (k <= k0 centroids x their tiles) and the per-group masks / list
        const size_t dneed = (size_t)k0 * (size_t)(kpad / PAIR_N);
        if (c->dmin_cap < dneed) {
            cudaFree(c->dmin); c->dmin = nullptr; c->dmin_cap = 0;
            AKM_CUDA(h, cudaMalloc((void **)&c->dmin, sizeof(float) * dneed));
            c->dmin_cap = dneed;
        }
        const int64_t gneed = ceil_div64(rs.n, 2 * TILE_M);
        if (c->gmask_cap < gneed) {
            cudaFree(c->gmask); cudaFree(c->glist); c->gmask = nullptr; c->glist = nullptr; c->gmask_cap = 0;
            AKM_CUDA(h, cudaMalloc((void **)&c->gmask, sizeof(uint32_t) * (size_t)gneed));
            AKM_CUDA(h, cudaMalloc((void **)&c->glist, sizeof(int2) * (size_t)gneed));
            c->gmask_cap = gneed;
        }
        if (c->prev_cap < k0) {
            cudaFree(c->Cprev); cudaFree(c->sig_old); c->Cprev = nullptr; c->sig_old = nullptr;
            AKM_CUDA(h, cudaMalloc((void **)&c->Cprev, sizeof(double) * (size_t)k0 * d));
            AKM_CUDA(h, cudaMalloc((void **)&c->sig_old, sizeof(float) * (size_t)k0));
            c->prev_cap = k0;
            *fresh_seeds = true;
        }
    }
    return AKM_OK;
}

// One-off preparation of the cached partition for the tensor engine (fp16 operand copy + scratch), so that the first
// superstep of a training run does neither.  Part of KMeansPreallocateCentroid's step-0 work in this library.
int tensor_warm(akm_handle h, RowStore &rs, int k0) {
    if (!tensor_supported(h, k0) || rs.n <= 2 * TILE_M || rs.n >= ((int64_t)1 << 31) - 256) return AKM_OK;
    TensorCtx *c = ctx_of(h);
    AKM_TRY(tensor_prepare_x16(h, rs));
    if (rs.x16_shift == INT_MIN) return AKM_OK;
    bool fresh = false;
    const int kpad = (k0 + PAIR_N - 1) / PAIR_N * PAIR_N;
    AKM_TRY(ensure_tensor_buffers(h, c, rs, k0, kpad, true, &fresh));
    if (fresh) c->k_prev = 0;
    return AKM_OK;
}

int tensor_assign(akm_handle h, RowStore &rs, const double *C, const double *clabel, int k0, int k, int32_t *labels,
                  double *mind, int chain, const int *remap) {
    TensorCtx *c = ctx_of(h);
    const int d = h->d;
    if (!c->encode) {
        void *fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
        if (e != cudaSuccess || !fn) return akm_fail(h, AKM_ERR_CUDA, "cuTensorMapEncodeTiled is unavailable");
        c->encode = (EncodeTiledFn)fn;
    }
    AKM_TRY(tensor_prepare_x16(h, rs));
    auto exact_all = [&]() {
        if (labels == h->labels) c->k_prev = 0;      // the seeds no longer describe the labels in that buffer
        h->stats.engine_used = AKM_ENGINE_EXACT;
        h->stats.rows_certified = 0; h->stats.rows_refined = rs.n; h->stats.rows_exact_all = rs.n;
        return launch_assign_exact(h, rs, nullptr, rs.n, C, clabel, k0, k, labels, mind, nullptr);
    };
    if (rs.x16_shift == INT_MIN || k <= 0) return exact_all();

    const int KB = (d + KBLK - 1) / KBLK;
    // shared-memory plan of a CTA: centroids resident (d <= 128 and they fit) or, for every other shape, both
    // operands streamed in 64-half blocks
    const size_t budget = std::min<size_t>(h->smem_optin, 227 * 1024);
    const int ntiles = (k + PAIR_N - 1) / PAIR_N;
    const int kpad = ntiles * PAIR_N;
    // one centroid tile of which at most half holds centroids: two teams of epilogue warps (see the kernel)
    static const int split_env = getenv("AKM_TC_SPLIT") ? atoi(getenv("AKM_TC_SPLIT")) : 1;      // tuning aid
    const bool split = split_env != 0 && ntiles == 1 && k <= PAIR_N / 2 && KB <= 2;
    const size_t pair_fixed = (size_t)kpad * 4 + 256 + (size_t)4 * TILE_M * MAX_CAND * 8 + (size_t)4 * TILE_M * 16 +
                              (size_t)(split ? 2 : 1) * TILE_M * WL_CAND * 4 + 1024;   // offsets, barriers, rings, exchange, merge staging (per team), alignment
    size_t pair_smem = (size_t)2 * KB * BOX_BYTES + (size_t)ntiles * KB * BOX_BYTES + pair_fixed;
    static const bool stream_env = getenv("AKM_TC_STREAM") && atoi(getenv("AKM_TC_STREAM")) != 0;   // tuning aid
    const bool pair_resident = KB <= 2 && pair_smem <= budget && !stream_env;
    int stream_stages = 0;
    if (!pair_resident && budget > pair_fixed) stream_stages = (int)std::min<size_t>(8, (budget - pair_fixed) / (2 * BOX_BYTES));
    const bool pair_stream = stream_stages >= 3;
    if (pair_stream) pair_smem = (size_t)stream_stages * 2 * BOX_BYTES + pair_fixed;
    // a partition of at most one row group (256 rows) is not worth a cluster launch
    if (!(pair_resident || pair_stream) || h->num_sms < 2 || rs.n <= 2 * TILE_M) return exact_all();
    const int d16 = rs.d16;
    bool fresh_seed_buffers = false;
    AKM_TRY(ensure_tensor_buffers(h, c, rs, k0, kpad, &rs == &h->rows, &fresh_seed_buffers));
    if (rs.n >= ((int64_t)1 << 31) - 256) return exact_all();

    // seeds are valid only when this pass continues the previous one on the same rows and label buffer
    const bool train_rows = (&rs == &h->rows) && labels == h->labels && chain != 0;
    bool use_seed = train_rows && c->k_prev > 0 && c->chain == chain && c->seed_rows == rs.n && c->seed_x16 == rs.x16_gen;
    if (use_seed && chain == 1 && c->step_prev != h->step - 1) use_seed = false;
    if (use_seed && chain == 2 && c->k_prev != k) use_seed = false;
    if (fresh_seed_buffers) use_seed = false;
    static const bool no_seed_env = getenv("AKM_NOSEED") != nullptr;      // debugging aid
    if (no_seed_env) use_seed = false;
    const int k_prev = use_seed ? c->k_prev : 0;
    AKM_CUDA(h, cudaMemsetAsync(c->counters, 0, 8 * sizeof(int), h->stream));
    AKM_CUDA(h, cudaMemsetAsync(c->pc, 0, sizeof(PassConst), h->stream));
    AKM_CUDA(h, cudaMemsetAsync(c->absmax, 0, sizeof(unsigned), h->stream));
    c16_max_kernel<<<std::min(h->num_sms, (k * d + 255) / 256), 256, 0, h->stream>>>(C, (size_t)k * d, c->absmax);
    c16_rows_kernel<<<(kpad + k_prev + 7) / 8, 256, 0, h->stream>>>(C, clabel, k, kpad, d, d16, h->metric, rs.x16_shift,
                                                                     c->absmax, c->c16, c->cn, c->pc, c->Cprev,
                                                                     (chain == 1) ? remap : nullptr, k_prev, c->sig_old);
    h->stats.kernel_launches += 2;
    AKM_CUDA(h, cudaGetLastError());

    CUtensorMap mapX, mapC;
    AKM_TRY(encode_2d(h, c, &mapX, rs.x16, (uint64_t)d, (uint64_t)rs.n, (uint64_t)d16 * 2, KBLK, TILE_M));
    AKM_TRY(encode_2d(h, c, &mapC, c->c16, (uint64_t)d, (uint64_t)kpad, (uint64_t)d16 * 2, KBLK, TILE_M));

    SearchParams sp{};
    sp.n = rs.n;
    sp.ntiles = ntiles; sp.a_stages = pair_stream ? stream_stages : 2;
    sp.kblocks = KB;
    sp.d = d; sp.metric = h->metric; sp.k = k;
    sp.cn = c->cn; sp.pc = c->pc; sp.xerr = rs.xerr; sp.labels = labels;
    sp.wl_rows = c->wl_rows; sp.wl_cand = c->wl_cand; sp.counters = c->counters;
    sp.wl_cap = (int)c->wl_slots; sp.exact_rows = c->exact_rows; sp.exact_count = c->counters + 1;
    sp.eta = (float)(KB * KBLK + 32) * 2.3841858e-7f;       // (d16 + 32) * 2^-22
    sp.useed = train_rows ? c->useed : nullptr;
    sp.sig_old = c->sig_old;
    sp.k_prev = k_prev;
    // tile pruning needs seeds (an upper bound of every row's distance to its previous centroid) and pays when a
    // group's rows share that centroid, i.e. on a cluster-ordered partition; a group that needs no tile at all (every
    // other centroid provably out of reach of every row) is skipped whole
    const int ngroups2 = (int)ceil_div64(rs.n, 2 * TILE_M);
    const bool prune = k_prev > 0 && h->tile_pruning;
    if (prune) {
        const size_t need = (size_t)k * ntiles;           // (buffers: ensure_tensor_buffers; pruning needs seeds, so train_rows)
        if (c->dmin_cap < need || c->gmask_cap < ngroups2) return akm_fail(h, AKM_ERR_STATE, "pruning buffers were not reserved");
        AKM_CUDA(h, cudaMemsetAsync(c->gmask, 0, sizeof(uint32_t) * (size_t)ngroups2, h->stream));
        AKM_CUDA(h, cudaMemsetAsync(c->dmin, 0x7F, sizeof(float) * need, h->stream));     // 3.4e38: "no other centroid in this tile"
        centroid_tile_dist_kernel<<<dim3((k + 31) / 32, (k + 31) / 32), 256, 0, h->stream>>>(C, k, d, ntiles,
                                                                                             reinterpret_cast<unsigned *>(c->dmin));
        h->stats.kernel_launches++;
    }
    // second-nearest bounds live with the seeds: only the cached training partition keeps them between passes
    // and they pay only where groups can be skipped: pruning on, rows cluster-ordered
    float2 *lbv = (train_rows && h->tile_pruning && rs.reordered) ? c->lb : nullptr;
    if (!use_seed || !lbv) c->lb_valid = false;
    if (lbv && k_prev > 0) {
        shift_top2_kernel<<<1, 256, 0, h->stream>>>(c->sig_old, k_prev, c->top);
        h->stats.kernel_launches++;
    }
    rowprep_kernel<<<h->num_sms * 16, ROWPREP_THREADS, 0, h->stream>>>(rs.n, d, h->metric, sp.eta, c->pc, rs.xerr, labels, c->useed,
                                                          c->sig_old, k_prev, (chain == 1) ? remap : nullptr, c->wm,
                                                          prune ? c->dmin : nullptr, ntiles, k, prune ? c->gmask : nullptr,
                                                          lbv, c->lb_valid ? 1 : 0, c->top);
    h->stats.kernel_launches++;
    sp.lb = lbv;
    AKM_CUDA(h, cudaGetLastError());
    sp.wm = c->wm;
    sp.dbg = h->epi_dbg;
    if (prune) {
        group_list_kernel<<<std::max(1, std::min(h->num_sms, (ngroups2 + 255) / 256)), 256, 0, h->stream>>>(c->gmask, ngroups2, c->glist, c->counters + 5);
        h->stats.kernel_launches++;
        AKM_CUDA(h, cudaGetLastError());
        sp.glist = c->glist; sp.glist_n = c->counters + 5;
    }
    // (counters are reused by the next pass only: the list stays readable for this pass's accumulation)
    h->searched_valid = false;
    if (prune && train_rows && labels == h->labels) { h->searched_list = c->glist; h->searched_n = c->counters + 5; h->searched_valid = true; }
    h->stats.tiles_total = (int64_t)ngroups2 * ntiles;
    const int pgrid = std::max(1, std::min(h->num_sms / 2, ngroups2)) * 2;
    int rc;
    const bool lbk = sp.lb != nullptr;
    if (pair_stream) rc = lbk ? launch_pair<0, false, true>(h, mapX, mapC, sp, pair_smem, pgrid) : launch_pair<0, false, false>(h, mapX, mapC, sp, pair_smem, pgrid);
    else if (split && KB == 1) rc = lbk ? launch_pair<1, true, true>(h, mapX, mapC, sp, pair_smem, pgrid) : launch_pair<1, true, false>(h, mapX, mapC, sp, pair_smem, pgrid);
    else if (split) rc = lbk ? launch_pair<2, true, true>(h, mapX, mapC, sp, pair_smem, pgrid) : launch_pair<2, true, false>(h, mapX, mapC, sp, pair_smem, pgrid);
    else if (KB == 1) rc = lbk ? launch_pair<1, false, true>(h, mapX, mapC, sp, pair_smem, pgrid) : launch_pair<1, false, false>(h, mapX, mapC, sp, pair_smem, pgrid);
    else rc = lbk ? launch_pair<2, false, true>(h, mapX, mapC, sp, pair_smem, pgrid) : launch_pair<2, false, false>(h, mapX, mapC, sp, pair_smem, pgrid);
    AKM_TRY(rc);
    if (train_rows) {
        AKM_CUDA(h, cudaMemcpyAsync(c->Cprev, C, sizeof(double) * (size_t)k * d, cudaMemcpyDeviceToDevice, h->stream));
        c->k_prev = k; c->chain = chain; c->step_prev = h->step; c->seed_rows = rs.n; c->seed_x16 = rs.x16_gen;
        c->lb_valid = lbv != nullptr;
    } else if (labels == h->labels) {
        c->k_prev = 0;                       // somebody else's labels now live in the buffer the seeds refer to
    }
    if (h->profiling && h->mid_event) { cudaEventRecord(h->mid_event, h->stream); h->mid_recorded = true; }

    // refinement: candidates first, then the exact list (lengths stay on the device)
    const int rblocks = h->num_sms * 32;
#define REFINE(T, PTR, M)                                                                                         \
    refine_candidates_kernel<T, M><<<rblocks, 128, 0, h->stream>>>(PTR, rs.xlabel, d, C, clabel, k, c->wl_rows, \
                                                                   c->wl_cand, c->counters, (int)c->wl_slots, labels, c->exact_rows, c->counters + 1, \
                                                                   c->useed, sp.lb)
    if (h->metric == AKM_EUCLIDEAN) { if (rs.store == 4) REFINE(float, rs.x32, AKM_EUCLIDEAN); else REFINE(double, rs.x64, AKM_EUCLIDEAN); }
    else { if (rs.store == 4) REFINE(float, rs.x32, AKM_COSINE); else REFINE(double, rs.x64, AKM_COSINE); }
#undef REFINE
    h->stats.kernel_launches++;
    AKM_CUDA(h, cudaGetLastError());
    AKM_TRY(launch_assign_exact_list(h, rs, c->exact_rows, c->counters + 1, C, clabel, k0, k, labels));
    if (mind) {
        const int wb = h->num_sms * 16;
        if (h->metric == AKM_EUCLIDEAN) {
            if (rs.store == 4) winner_distance_kernel<float, AKM_EUCLIDEAN><<<wb, 128, 0, h->stream>>>(rs.x32, rs.xlabel, rs.n, d, C, clabel, labels, mind);
            else winner_distance_kernel<double, AKM_EUCLIDEAN><<<wb, 128, 0, h->stream>>>(rs.x64, rs.xlabel, rs.n, d, C, clabel, labels, mind);
        } else {
            if (rs.store == 4) winner_distance_kernel<float, AKM_COSINE><<<wb, 128, 0, h->stream>>>(rs.x32, rs.xlabel, rs.n, d, C, clabel, labels, mind);
            else winner_distance_kernel<double, AKM_COSINE><<<wb, 128, 0, h->stream>>>(rs.x64, rs.xlabel, rs.n, d, C, clabel, labels, mind);
        }
        h->stats.kernel_launches++;
        AKM_CUDA(h, cudaGetLastError());
    }
    // counts for akm_get_stats (read back with the other scalars by the caller)
    // iscal[4] = rows refined, [5] = rows evaluated against all k, [6] = labels changed
    AKM_CUDA(h, cudaMemcpyAsync(h->iscal + 4, c->counters + 3, sizeof(int), cudaMemcpyDeviceToDevice, h->stream));
    AKM_CUDA(h, cudaMemcpyAsync(h->iscal + 5, c->counters + 1, 2 * sizeof(int), cudaMemcpyDeviceToDevice, h->stream));
    AKM_CUDA(h, cudaMemcpyAsync(h->iscal + 7, c->counters + 4, sizeof(int), cudaMemcpyDeviceToDevice, h->stream));
    h->stats.engine_used = AKM_ENGINE_TENSOR;
    return AKM_OK;
}
