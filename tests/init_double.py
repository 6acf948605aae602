"""CPU stand-in for the akm_init_* entry points (include/akm.h), built on the oracle's Java-order distances.

Test infrastructure only: it lets the host logic of alink_b200/init.py run without a GPU and is the checker the
GPU seeding passes are compared with (same hash, same update rule)."""
import numpy as np

import oracle as orc

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def _mix64(z):
    z = np.asarray(z, dtype=np.uint64)
    with np.errstate(over="ignore"):
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def uniform(seed, round_no, global_rows):
    """The pick's uniform in [0,1): same integer arithmetic as init_pick_kernel (alink_b200/csrc/init.cu)."""
    with np.errstate(over="ignore"):
        a = _mix64(np.uint64(seed & 0xFFFFFFFFFFFFFFFF) + np.uint64(0x9E3779B97F4A7C15) * np.uint64(round_no + 1))
        g = np.asarray(global_rows, dtype=np.uint64) * np.uint64(0xD1B54A32D192ED03)
        h = _mix64(a ^ g)
    return (h >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)


class InitDouble:
    def __init__(self, X, metric):
        self.metric = metric
        self.Xp, self.xl = orc.prepare_rows(metric, X)
        self.n, self.d = self.Xp.shape
        self.init_reset()

    def num_rows(self):
        return self.n, 8

    def init_reset(self):
        self.cost = np.full(self.n, np.finfo(np.float64).max)
        self.near = np.full(self.n, -1, dtype=np.int64)

    def init_update(self, centers, id_base):
        C = np.ascontiguousarray(centers, dtype=np.float64)
        if self.metric == orc.EUCLIDEAN:
            C, cl = orc.prepare_centroids(self.metric, C)
        else:
            cl = np.zeros(C.shape[0])
        if self.n:
            _, labels, mind = orc.assign_accumulate(self.metric, self.Xp, self.xl, C, cl)
            better = mind < self.cost
            self.cost[better] = mind[better]
            self.near[better] = id_base + labels[better]
        return float(self.cost.sum())

    def init_pick(self, scale, seed, round_no, row_base, cap):
        u = uniform(seed, round_no, row_base + np.arange(self.n, dtype=np.int64))
        rows = np.nonzero(u < self.cost * scale)[0].astype(np.int64)
        assert rows.size <= cap
        return rows

    def init_counts(self, m_total):
        ok = self.near >= 0
        return np.bincount(self.near[ok], minlength=m_total).astype(np.float64)[:m_total]

    def get_rows(self, rows):
        return self.Xp[np.asarray(rows, dtype=np.int64)].copy()
