"""Initial centroids: RANDOM and K_MEANS_PARALLEL (k-means||), SURVEY 8 f1.

Mirror of operator/common/clustering/kmeans/KMeansInitCentroids.java (:40-66 dispatch, :69-87 RANDOM,
:89-165 k-means||) and LocalKmeansFunc.java (:36-122, weighted k-means++ and at most 30 Lloyd iterations on the
candidates).  The passes over the N cached rows run on the GPU through the akm_init_* entry points of
include/akm.h; what the reference runs in ONE subtask on a few thousand candidates (`setParallelism(1)`,
KMeansInitCentroids.java:162-165) is host work here as well.

Not reproducible from the reference: which rows Flink's DataSetUtils.sampleWithSize draws and the
java.util.Random(subtask) stream of FilterNewCenter depend on Flink's partitioning.  Here the first centre and
the RANDOM sample come from numpy's generator seeded with `randomSeed`, and the Bernoulli picks from a
counter-based hash of (seed, round, global row index), so a run is reproducible and independent of the number
of GPUs.  LocalKmeansFunc's own generator is `new Random(0)`; JavaRandom below reproduces that stream.
"""
import numpy as np

EUCLIDEAN, COSINE = 0, 1
LOCAL_MAX_ITER = 30                                   # LocalKmeansFunc.java:25


class JavaRandom:
    """java.util.Random: 48-bit LCG, next(bits) = (int)(seed >>> (48 - bits))."""
    _MULT, _ADD, _MASK = 0x5DEECE66D, 0xB, (1 << 48) - 1

    def __init__(self, seed):
        self.seed = (int(seed) ^ self._MULT) & self._MASK

    def _next(self, bits):
        self.seed = (self.seed * self._MULT + self._ADD) & self._MASK
        v = self.seed >> (48 - bits)
        return v - (1 << 32) if v >= (1 << 31) else v           # (int) cast

    def next_double(self):
        return ((self._next(26) << 27) + self._next(27)) * (1.0 / (1 << 53))

    def next_int(self, bound):
        if bound <= 0:
            raise ValueError("bound must be positive")
        r = self._next(31)
        m = bound - 1
        if (bound & m) == 0:
            return (bound * r) >> 31
        u = r
        r = u % bound
        while u - r + m >= (1 << 31):                           # int overflow test of the original loop
            u = self._next(31)
            r = u % bound
        return r


def _pairwise(metric, X, xl, C, cl):
    """True distances (not squared) between prepared rows X and prepared centroids C."""
    t = X @ C.T
    if metric == EUCLIDEAN:
        return np.sqrt(np.abs(xl[:, None] + cl[None, :] - 2.0 * t))
    return 1.0 - t


def _update_label(metric, C):
    """FastDistance.updateLabel on a centroid matrix: EUCLIDEAN -> squared norms, COSINE -> normalise rows."""
    if metric == COSINE:
        nrm = np.sqrt((C * C).sum(axis=1))
        nz = nrm > 0
        C[nz] *= (1.0 / nrm[nz])[:, None]
        return np.zeros(C.shape[0])
    return (C * C).sum(axis=1)


def pick_weight(weights, costs, rnd):
    """LocalKmeansFunc.pickWeight (:107-121): inverse CDF over weight * cost, off-by-one of the original kept."""
    cum = np.concatenate([[0.0], np.cumsum(weights * costs)])
    r = rnd.next_double() * cum[-1]
    hit = np.nonzero(cum >= r)[0]
    res = int(hit[0]) if hit.size else len(costs)
    return 0 if res == 0 else res - 1


def local_kmeans(k, weights, samples, metric):
    """LocalKmeansFunc.kmeans: weighted k-means++ seeding, then Lloyd until no label changes (<= 30 rounds).
    Quirks of the original are kept: the sum matrix is never cleared between rounds (:48,:57-60), an empty
    cluster is re-seeded with a random sample (:73-75), costs are distances, not squared distances."""
    samples = np.ascontiguousarray(samples, dtype=np.float64)
    weights = np.asarray(weights, dtype=np.float64)
    m, d = samples.shape
    rnd = JavaRandom(0)
    xl = (samples * samples).sum(axis=1)
    # ---- sampleInitialCentroids (:88-105)
    costs = np.ones(m)
    chosen = []
    index = 0
    for i in range(k):
        if i > 0:
            c = samples[index:index + 1]
            dist = _pairwise(metric, samples, xl, c, xl[index:index + 1])[:, 0]
            costs = np.minimum(dist, costs) if i > 1 else dist
        index = pick_weight(weights, costs, rnd)
        chosen.append(index)
    C = samples[chosen].copy()
    cl = _update_label(metric, C)
    # ---- Lloyd (:52-80)
    sums = np.zeros((k, d + 1))
    labels = np.zeros(m, dtype=np.int64)
    for _ in range(LOCAL_MAX_ITER):
        D = _pairwise(metric, samples, xl, C, cl)
        cid = np.argmin(D, axis=1)                              # first minimum, like getMinPointIndex
        np.add.at(sums[:, :d], cid, samples * weights[:, None])
        np.add.at(sums[:, d], cid, weights)
        converged = bool((cid == labels).all())
        labels = cid
        C[:] = 0.0
        for j in range(k):
            w = sums[j, d]
            if w > 0:
                C[j] = sums[j, :d] * (1.0 / w)
            else:
                C[j] = samples[rnd.next_int(m)]
        cl = _update_label(metric, C)
        if converged:
            break
    return C


class _Solo:
    """Stand-in for alink_b200.distributed.TorchDistributed on a single process."""
    rank, world = 0, 1

    def all_sum(self, arr):
        return np.asarray(arr, dtype=np.float64)

    def gather(self, obj):
        return [obj]


def _shard_starts(dist, n_local):
    counts = dist.gather(int(n_local))
    return np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)


def _fetch_global_rows(handle, dist, starts, global_rows):
    """Prepared vectors of the given (sorted, distinct) global rows, identical on every rank."""
    lo, hi = starts[dist.rank], starts[dist.rank + 1]
    mine = global_rows[(global_rows >= lo) & (global_rows < hi)] - lo
    vecs = handle.get_rows(mine)
    parts = dist.gather(vecs)
    return np.vstack([p for p in parts if p.size]) if any(p.size for p in parts) else np.zeros((0, handle.d))


def random_init(handle, k, seed, dist=None):
    """KMeansInitCentroids.randomInit (:69-87): k distinct rows, uniformly."""
    dist = dist or _Solo()
    starts = _shard_starts(dist, handle.num_rows()[0])
    total = int(starts[-1])
    rng = np.random.default_rng(seed)
    pick = np.sort(rng.choice(total, size=min(k, total), replace=False)).astype(np.int64)
    return _fetch_global_rows(handle, dist, starts, pick)


def kmeans_parallel_init(handle, k, init_steps, seed, metric, dist=None, info=None):
    """KMeansInitCentroids.kMeansPlusPlusInit (:89-165) on the rows cached in `handle`."""
    dist = dist or _Solo()
    n_local = handle.num_rows()[0]
    starts = _shard_starts(dist, n_local)
    total = int(starts[-1])
    row_base = int(starts[dist.rank])
    rng = np.random.default_rng(seed)
    first = np.array([int(rng.integers(total))], dtype=np.int64)        # sampleWithSize(.., 1, 0L)  (:108-111)
    handle.init_reset()
    centers = _fetch_global_rows(handle, dist, starts, first)           # TransformToCenter: centres are copies
    sum_cost = float(dist.all_sum(np.array([handle.init_update(centers, 0)]))[0])     # CalWeight + SUM
    for rnd_no in range(1, max(int(init_steps), 1)):                    # iterate(initSteps - 1)  (:119)
        if not (sum_cost > 0.0):
            break                                                       # every row coincides with a centre
        scale = 2.0 * k / sum_cost                                      # FilterNewCenter.open (:197-201)
        cap = int(min(n_local, 8 * k + 4096)) + 1
        picked = handle.init_pick(scale, seed, rnd_no, row_base, cap)
        new = dist.gather(handle.get_rows(picked))
        new = np.vstack([p for p in new if p.size]) if any(p.size for p in new) else np.zeros((0, handle.d))
        if new.shape[0] == 0:
            continue
        base = centers.shape[0]
        centers = np.vstack([centers, new])
        sum_cost = float(dist.all_sum(np.array([handle.init_update(new, base)]))[0])
    counts = dist.all_sum(handle.init_counts(centers.shape[0]))
    keep = counts > 0                                                   # inner join with the per-centre counts (:140-150);
    samples, weights = centers[keep], counts[keep]                      # a centre loses its own row only to an identical earlier one
    if info is not None:
        info.update(candidates=int(centers.shape[0]), kept=int(samples.shape[0]), sum_cost=sum_cost)
    if samples.shape[0] <= k:                                           # LocalKmeans.mapPartition (:283-285)
        return samples.copy()
    return local_kmeans(k, weights, samples, metric)
