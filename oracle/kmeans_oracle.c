/*
 * kmeans_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement (plain C, IEEE fp64, no FMA, no re-association) of the
 * arithmetic Alink's Java KMeans hot path performs.  It exists so that the
 * CUDA path in alink_b200/csrc can be checked bit-for-bit (labels) and to
 * 1e-6 (centroids/objective).  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load it.
 *
 * The reference is pure Java (no JVM in this image: it cannot be compiled or
 * run here), and its inner loops live in a third-party dependency that is not
 * vendored in /root/reference:
 *     com.github.fommil.netlib:core:1.1.2 -> net.sourceforge.f2j:
 *     arpack_combined_all:0.1  (F2J = machine translation of the Fortran
 *     reference BLAS; reached through org.scalanlp:breeze_2.11:0.11.2,
 *     core/pom.xml:166-178, BLAS.java:15,26)
 * so the loops below restate the *published reference-BLAS algorithms*
 * (dgemv 'T', dgemm 'T','N', daxpy, dscal) and are anchored on Alink's own
 * call sites, JUnit literals and the one bit-exact pin of this path
 * (KMeansModelMapperTest.java:58).  See tests/test_oracle_golden.py.
 *
 * Compile: gcc -O2 -ffp-contract=off -fno-fast-math -shared -fPIC -pthread
 *
 * Path shorthands: CORE = core/src/main/java/com/alibaba/alink
 */
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <float.h>

#define ORC_EUCLIDEAN 0
#define ORC_COSINE 1

/* ---- CORE/common/linalg/BLAS.java:103-110  (own sequential loop) ---- */
double orc_dot(const double *x, const double *y, int n) {
    double s = 0.;
    for (int i = 0; i < n; i++) s += x[i] * y[i];
    return s;
}

/* ---- F2J daxpy, reached from BLAS.java:96-98: y[i] = y[i] + a*x[i] ---- */
void orc_daxpy(int n, double a, const double *x, double *y) {
    if (n <= 0 || a == 0.0) return;
    for (int i = 0; i < n; i++) y[i] = y[i] + a * x[i];
}

/* ---- F2J dscal, reached from BLAS.java:129-131 ---- */
void orc_dscal(int n, double a, double *x) {
    for (int i = 0; i < n; i++) x[i] = a * x[i];
}

/* ---- reference-BLAS dgemv('T', m, n, alpha, A, lda=m, x, 1, beta, y, 1),
 *      reached from BLAS.java:204-212.  A is column-major m x n; column j is
 *      the contiguous slice A[j*m .. j*m+m)  (DenseMatrix.java:202-204). ---- */
void orc_dgemv_t(int m, int n, double alpha, const double *A, const double *x,
                 double beta, double *y) {
    if (beta != 1.0) {
        if (beta == 0.0) for (int j = 0; j < n; j++) y[j] = 0.0;
        else for (int j = 0; j < n; j++) y[j] = beta * y[j];
    }
    if (alpha == 0.0) return;
    for (int j = 0; j < n; j++) {
        double temp = 0.0;
        const double *col = A + (size_t)j * m;
        for (int i = 0; i < m; i++) temp = temp + col[i] * x[i];
        y[j] = y[j] + alpha * temp;
    }
}

/* ---- reference-BLAS dgemv('N'), BLAS.java:204-212 with transA=false ---- */
void orc_dgemv_n(int m, int n, double alpha, const double *A, const double *x,
                 double beta, double *y) {
    if (beta != 1.0) {
        if (beta == 0.0) for (int i = 0; i < m; i++) y[i] = 0.0;
        else for (int i = 0; i < m; i++) y[i] = beta * y[i];
    }
    if (alpha == 0.0) return;
    for (int j = 0; j < n; j++) {
        if (x[j] != 0.0) {
            double temp = alpha * x[j];
            for (int i = 0; i < m; i++) y[i] = y[i] + temp * A[(size_t)j * m + i];
        }
    }
}

/* ---- EuclideanDistance.updateLabel :58-87 / CosineDistance.updateLabel :59-85
 *      for one vector (FastDistance.prepareVectorData :136-140).
 *      EUCL: returns label = sum x_i^2 (sequential).  COS: normalises x in
 *      place when the norm is > 0, label unused (returns 0). ---- */
double orc_prepare_vector(int metric, double *x, int d) {
    double s = orc_dot(x, x, d);
    if (metric == ORC_EUCLIDEAN) return s;
    double nrm = sqrt(s);
    if (nrm > 0) orc_dscal(d, 1.0 / nrm, x);     /* DenseVector.scaleEqual -> BLAS.scal */
    return 0.0;
}

void orc_prepare_rows(int metric, double *X, int64_t n, int d, double *xlabel) {
    for (int64_t r = 0; r < n; r++) {
        double l = orc_prepare_vector(metric, X + (size_t)r * d, d);
        if (xlabel) xlabel[r] = l;
    }
}

/* ---- updateLabel for a FastDistanceMatrixData holding k0 vectors (all k0
 *      columns, also stale zero ones): EuclideanDistance.java:66-86,
 *      CosineDistance.java:66-84 ---- */
void orc_prepare_centroids(int metric, double *C, int k0, int d, double *clabel) {
    for (int j = 0; j < k0; j++) {
        double *c = C + (size_t)j * d;
        double s = 0.0;
        for (int i = 0; i < d; i++) s += c[i] * c[i];
        if (metric == ORC_EUCLIDEAN) {
            clabel[j] = s;
        } else {
            double nrm = sqrt(s);
            if (nrm > 0) orc_dscal(d, 1.0 / nrm, c);
            if (clabel) clabel[j] = 0.0;
        }
    }
}

/* ---- vec x matrix distances: EuclideanDistance.java:109-116,
 *      CosineDistance.java:107-110.  res has k0 entries. ---- */
void orc_distances(int metric, const double *x, double xlabel, const double *C,
                   const double *clabel, int k0, int d, double *res) {
    if (metric == ORC_EUCLIDEAN) {
        orc_dgemv_t(d, k0, -2.0, C, x, 0.0, res);
        for (int j = 0; j < k0; j++) res[j] = sqrt(fabs(res[j] + xlabel + clabel[j]));
    } else {
        for (int j = 0; j < k0; j++) res[j] = 1.0;
        orc_dgemv_t(d, k0, -1.0, C, x, 1.0, res);
    }
}

/* ---- KMeansUtil.getMinPointIndex :162-173 ---- */
int orc_min_index(const double *data, int end) {
    int index = -1;
    double min = DBL_MAX;
    for (int i = 0; i < end; i++) {
        if (data[i] < min) { index = i; min = data[i]; }
    }
    return index;
}

/* ---- KMeansAssignCluster.calc :59-63 over rows [r0, r1) + KMeansUtil.
 *      updateSumMatrix :60-85.  `sum` (k0*(d+1)) must be zeroed by the caller.
 *      dup != 0 repeats the distance evaluation like the Java path does
 *      (KMeansUtil.java:73 and :124) -- only used when timing the baseline.
 *      Returns 0, or -1 if some row had no finite minimum (Java would throw
 *      ArrayIndexOutOfBounds on index -1). ---- */
static int assign_range(int metric, const double *X, const double *xlabel,
                        int64_t r0, int64_t r1, int d, const double *C,
                        const double *clabel, int k0, int k, double *sum,
                        int32_t *labels, double *mind, double *dm, int dup) {
    for (int64_t r = r0; r < r1; r++) {
        const double *x = X + (size_t)r * d;
        double xl = xlabel ? xlabel[r] : 0.0;
        orc_distances(metric, x, xl, C, clabel, k0, d, dm);
        if (dup) orc_distances(metric, x, xl, C, clabel, k0, d, dm);
        int a = orc_min_index(dm, k);
        if (a < 0) return -1;
        double *row = sum + (size_t)a * (d + 1);
        orc_daxpy(d, 1.0, x, row);            /* BLAS.axpy(vectorSize, sampleWeight=1, ...) */
        row[d] += 1.0;
        if (labels) labels[r] = a;
        if (mind) mind[r] = dm[a];
    }
    return 0;
}

int orc_assign_accumulate(int metric, const double *X, const double *xlabel,
                          int64_t n, int d, const double *C, const double *clabel,
                          int k0, int k, double *sum, int32_t *labels,
                          double *mind, int dup) {
    double *dm = (double *)malloc(sizeof(double) * (size_t)(k0 > 0 ? k0 : 1));
    memset(sum, 0, sizeof(double) * (size_t)k0 * (d + 1));   /* Arrays.fill :59 */
    int rc = assign_range(metric, X, xlabel, 0, n, d, C, clabel, k0, k, sum,
                          labels, mind, dm, dup);
    free(dm);
    return rc;
}

/* ---- KMeansUpdateCentroids.updateCentroids :52-69.  Cnew is k0*d, fully
 *      rewritten; returns the new k (non-empty clusters, compacted). ---- */
int orc_update_centroids(int metric, const double *sum, int k, int k0, int d,
                         double *Cnew, double *clabel_new) {
    int index = 0;
    memset(Cnew, 0, sizeof(double) * (size_t)k0 * d);
    for (int j = 0; j < k; j++) {
        const double *row = sum + (size_t)j * (d + 1);
        double w = row[d];
        if (w == 0) continue;
        orc_daxpy(d, 1.0 / w, row, Cnew + (size_t)index * d);
        index++;
    }
    orc_prepare_centroids(metric, Cnew, k0, d, clabel_new);
    return index;
}

/* ---- KMeansIterTermination.calc :28-44 (left = centroids1, right =
 *      centroids2; only the diagonal of the k0 x k0 matrix is inspected).
 *      Matrix x matrix formulas: EuclideanDistance.java:126-141 (dgemm 'T','N'
 *      alpha=-2 beta=0 then sqrt(|(v + rightNorm) + leftNorm|)),
 *      CosineDistance.java:120-123 (fill 1, dgemm alpha=-1 beta=1).
 *      Returns 1 when every d_jj < tol for j < k. ---- */
int orc_converged(int metric, const double *C1, const double *l1, const double *C2,
                  const double *l2, int k, int d, double tol, double *diag) {
    int conv = 1;
    for (int j = 0; j < k; j++) {
        const double *a = C2 + (size_t)j * d;   /* right^T row */
        const double *b = C1 + (size_t)j * d;   /* left column */
        double temp = 0.0;
        for (int l = 0; l < d; l++) temp = temp + a[l] * b[l];
        double v;
        if (metric == ORC_EUCLIDEAN) {
            v = -2.0 * temp;                       /* beta == 0: C = alpha*temp */
            v = sqrt(fabs(v + l2[j] + l1[j]));
        } else {
            v = -1.0 * temp + 1.0 * 1.0;           /* C = alpha*temp + beta*C   */
        }
        if (diag) diag[j] = v;
        if (conv && v >= tol) conv = 0;            /* first failing j ends the Java loop */
    }
    return conv;
}

/* objective a13: EUCL sum D^2, COS sum D over the assignment pass */
static double objective_of(int metric, const double *mind, int64_t n) {
    double s = 0.0;
    for (int64_t r = 0; r < n; r++) s += (metric == ORC_EUCLIDEAN) ? mind[r] * mind[r] : mind[r];
    return s;
}

/* ---- DefaultDistributedInfo.java:10-35 ---- */
int64_t orc_shard_start(int64_t task, int64_t par, int64_t total) {
    int64_t div = total / par, mod = total % par;
    if (mod == 0) return div * task;
    if (task >= mod) return div * task + mod;
    return div * task + task;
}
int64_t orc_shard_count(int64_t task, int64_t par, int64_t total) {
    int64_t div = total / par, mod = total % par;
    if (mod == 0) return div;
    return task >= mod ? div : div + 1;
}

typedef struct {
    int metric; const double *X; const double *xlabel; int64_t r0, r1; int d;
    const double *C; const double *clabel; int k0, k; double *sum;
    int32_t *labels; double *mind; int dup; int rc;
} shard_job;

static void *shard_main(void *p) {
    shard_job *j = (shard_job *)p;
    double *dm = (double *)malloc(sizeof(double) * (size_t)j->k0);
    memset(j->sum, 0, sizeof(double) * (size_t)j->k0 * (j->d + 1));
    j->rc = assign_range(j->metric, j->X, j->xlabel, j->r0, j->r1, j->d, j->C,
                         j->clabel, j->k0, j->k, j->sum, j->labels, j->mind, dm, j->dup);
    free(dm);
    return NULL;
}

/* One assignment pass with P "Flink subtasks" (threads), each owning the
 * contiguous shard DefaultDistributedInfo gives it, followed by the AllReduce
 * SUM (AllReduce.java:125-133).  The Java all-reduce adds in message-arrival
 * order (AllReduce.java:307-316), which is not reproducible; the oracle fixes
 * rank order 0..P-1 (documented deviation; identity at P = 1). */
int orc_assign_accumulate_mt(int metric, const double *X, const double *xlabel,
                             int64_t n, int d, const double *C, const double *clabel,
                             int k0, int k, double *sum, int32_t *labels,
                             double *mind, int dup, int P) {
    if (P <= 1) return orc_assign_accumulate(metric, X, xlabel, n, d, C, clabel, k0, k,
                                             sum, labels, mind, dup);
    size_t L = (size_t)k0 * (d + 1);
    double *part = (double *)malloc(sizeof(double) * L * P);
    shard_job *jobs = (shard_job *)malloc(sizeof(shard_job) * P);
    pthread_t *th = (pthread_t *)malloc(sizeof(pthread_t) * P);
    for (int t = 0; t < P; t++) {
        int64_t s = orc_shard_start(t, P, n), c = orc_shard_count(t, P, n);
        shard_job jb = {metric, X, xlabel, s, s + c, d, C, clabel, k0, k,
                        part + L * t, labels, mind, dup, 0};
        jobs[t] = jb;
        pthread_create(&th[t], NULL, shard_main, &jobs[t]);
    }
    int rc = 0;
    for (int t = 0; t < P; t++) { pthread_join(th[t], NULL); if (jobs[t].rc) rc = jobs[t].rc; }
    memcpy(sum, part, sizeof(double) * L);
    for (int t = 1; t < P; t++)
        for (size_t i = 0; i < L; i++) sum[i] += part[L * t + i];
    free(part); free(jobs); free(th);
    return rc;
}

/* ---- Whole training run (SURVEY 8a "compact restatement"), P subtasks.
 *   X         n x d row-major; MODIFIED IN PLACE for COSINE (normalised), like Java
 *   C0        k0 x d initial centroids (row-major == Alink d x k0 column-major)
 *   out_C     k0 x d  final centroid buffer (first *out_k rows valid, rest as Java leaves them)
 *   out_w     k0      weight_j = sum[j][d] of the LAST pass, uncompacted index (KMeansOutputModel :60-68)
 *   out_obj   max_iter objective per executed pass (may be NULL)
 *   labels    n labels of the LAST pass (may be NULL)
 * Returns executed iteration count, or -1 on the no-finite-distance error. ---- */
int orc_train(int metric, double *X, int64_t n, int d, const double *C0, int k0,
              int max_iter, double tol, int P, double *out_C, int *out_k,
              double *out_w, double *out_obj, int32_t *labels) {
    size_t L = (size_t)k0 * (d + 1);
    double *xlabel = (double *)malloc(sizeof(double) * (size_t)(n > 0 ? n : 1));
    double *cur = (double *)malloc(sizeof(double) * (size_t)k0 * d);
    double *nxt = (double *)malloc(sizeof(double) * (size_t)k0 * d);
    double *lcur = (double *)calloc((size_t)k0, sizeof(double));
    double *lnxt = (double *)calloc((size_t)k0, sizeof(double));
    double *sum = (double *)calloc(L, sizeof(double));
    double *mind = (double *)malloc(sizeof(double) * (size_t)(n > 0 ? n : 1));
    orc_prepare_rows(metric, X, n, d, xlabel);           /* KMeansTrainBatchOp.java:95-101 */
    memcpy(cur, C0, sizeof(double) * (size_t)k0 * d);
    orc_prepare_centroids(metric, cur, k0, d, lcur);     /* KMeansUtil.buildCentroidsMatrix :35-45 */
    memcpy(nxt, cur, sizeof(double) * (size_t)k0 * d);   /* KMeansPreallocateCentroid :22-39 */
    memcpy(lnxt, lcur, sizeof(double) * (size_t)k0);
    int k = k0, it = 0;
    for (int s = 1; s <= max_iter; s++) {
        int rc = orc_assign_accumulate_mt(metric, X, xlabel, n, d, cur, lcur, k0, k, sum,
                                          labels, mind, 0, P);
        if (rc) { it = -1; break; }
        if (out_obj) out_obj[s - 1] = objective_of(metric, mind, n);
        k = orc_update_centroids(metric, sum, k, k0, d, nxt, lnxt);
        double *t = cur; cur = nxt; nxt = t;             /* ping-pong :38-42 */
        t = lcur; lcur = lnxt; lnxt = t;
        it = s;
        /* criterion on subtask 0: left = CENTROID1, right = CENTROID2; the
         * diagonal is symmetric in the operands up to the order of the two
         * label additions, so keep Java's roles: at odd s CENTROID1 is new. */
        const double *c1 = (s % 2 == 1) ? cur : nxt, *l1 = (s % 2 == 1) ? lcur : lnxt;
        const double *c2 = (s % 2 == 1) ? nxt : cur, *l2 = (s % 2 == 1) ? lnxt : lcur;
        if (orc_converged(metric, c1, l1, c2, l2, k, d, tol, NULL)) break;
    }
    if (out_C) memcpy(out_C, cur, sizeof(double) * (size_t)k0 * d);
    if (out_k) *out_k = k;
    if (out_w) for (int j = 0; j < k0; j++) out_w[j] = sum[(size_t)j * (d + 1) + d];
    free(xlabel); free(cur); free(nxt); free(lcur); free(lnxt); free(sum); free(mind);
    return it;
}

/* ---- KMeansUtil.getProbArrayFromDistanceArray :262-269
 *      (StatUtils.sum = sequential sum; F2J daxpy) ---- */
void orc_probs(const double *dist, int len, double *probs) {
    double sum = 0.0;
    for (int i = 0; i < len; i++) sum += dist[i];
    double ratio = 1.0 / sum / (len - 1);
    for (int i = 0; i < len; i++) probs[i] = 1.0 / (len - 1);
    orc_daxpy(len, -ratio, dist, probs);
}

/* ---- KMeansModelMapper.map :62-97 for a batch of rows.  C is the model's
 *      k x d matrix as loaded (transformTrainDataToPredictData :246-260 runs
 *      updateLabel on it; done here on a private copy).  X is not modified.
 *      idx[n]; dist (nullable) n; probs (nullable) n*k. Returns 0 / -1. ---- */
int orc_predict(int metric, const double *X, int64_t n, int d, const double *C,
                int k, int64_t *idx, double *dist, double *probs) {
    double *Cm = (double *)malloc(sizeof(double) * (size_t)k * d);
    double *cl = (double *)calloc((size_t)k, sizeof(double));
    double *x = (double *)malloc(sizeof(double) * (size_t)d);
    double *dm = (double *)malloc(sizeof(double) * (size_t)k);
    memcpy(Cm, C, sizeof(double) * (size_t)k * d);
    orc_prepare_centroids(metric, Cm, k, d, cl);
    int rc = 0;
    for (int64_t r = 0; r < n; r++) {
        memcpy(x, X + (size_t)r * d, sizeof(double) * (size_t)d);
        double xl = orc_prepare_vector(metric, x, d);
        orc_distances(metric, x, xl, Cm, cl, k, d, dm);
        int a = orc_min_index(dm, k);
        idx[r] = a;
        if (a < 0) { rc = -1; if (dist) dist[r] = NAN; continue; }
        if (dist) dist[r] = dm[a];
        if (probs) orc_probs(dm, k, probs + (size_t)r * k);
    }
    free(Cm); free(cl); free(x); free(dm);
    return rc;
}

/* ------------------------------------------------------------------------------------------------
 * EvalClusterBatchOp, vector part.  Follows, without copying:
 *   ClusterEvaluationUtil.getClusterStatistics  operator/common/evaluation/ClusterEvaluationUtil.java:171-218
 *   ClusterMetricsSummary.toMetrics              operator/common/evaluation/ClusterMetricsSummary.java:109-150
 *   ClusterEvaluationUtil.calSilhouetteCoefficient                                             :106-145
 *   EuclideanDistance.calc(Vector,Vector) :45-47, CosineDistance.calc(Vector,Vector) :46-50
 * labels are dense cluster ids 0..k-1 (< 0: row skipped); clusters are visited in id order (Flink's group
 * order is unspecified; every aggregate below except the floating-point summation order is order-free).
 * out[0..8] = ssb, ssw, compactness, separation, daviesBouldin, calinskiHarabaz, silhouette, k (non-empty), count
 * ------------------------------------------------------------------------------------------------ */
static double eval_dist(int metric, const double *a, const double *b, int d) {
    if (metric == ORC_EUCLIDEAN) {
        double s = 0.0;
        for (int i = 0; i < d; ++i) { double df = a[i] - b[i]; s += df * df; }
        return sqrt(s);
    }
    double dot = orc_dot(a, b, d);
    double cross = sqrt(orc_dot(a, a, d)) * sqrt(orc_dot(b, b, d));
    return 1.0 - (cross > 0.0 ? dot / cross : 0.0);
}

int orc_eval_clusters(int metric, const double *X, int64_t n, int d, const int *labels, int k,
                      double *cnt, double *mean, double *compact, double *dss, double *nsum, double *out) {
    double *vec = (double *)malloc(sizeof(double) * (size_t)d);
    double *sumall = (double *)calloc((size_t)d, sizeof(double));
    if (!vec || !sumall) { free(vec); free(sumall); return -1; }
    int kk = 0;
    int64_t total = 0;
    for (int j = 0; j < k; ++j) {
        double *m = mean + (size_t)j * d;
        for (int i = 0; i < d; ++i) m[i] = 0.0;
        int64_t c = 0;
        for (int64_t r = 0; r < n; ++r) {
            if (labels[r] != j) continue;
            const double *x = X + (size_t)r * d;
            if (metric == ORC_EUCLIDEAN) {
                for (int i = 0; i < d; ++i) m[i] += x[i];
            } else {
                double s = 1.0 / sqrt(orc_dot(x, x, d));            /* vec.scaleEqual(1.0 / vec.normL2()) */
                for (int i = 0; i < d; ++i) m[i] += x[i] * s;
            }
            ++c;
        }
        cnt[j] = (double)c;
        compact[j] = dss[j] = nsum[j] = 0.0;
        if (c == 0) continue;
        double inv = 1.0 / (double)c;
        for (int i = 0; i < d; ++i) m[i] = m[i] * inv;               /* sumVector.scale(1.0 / total) */
        double dsum = 0.0;
        for (int64_t r = 0; r < n; ++r) {
            if (labels[r] != j) continue;
            const double *x = X + (size_t)r * d;
            const double *v = x;
            if (metric != ORC_EUCLIDEAN) {
                double s = 1.0 / sqrt(orc_dot(x, x, d));
                for (int i = 0; i < d; ++i) vec[i] = x[i] * s;
                v = vec;
            }
            double dd = eval_dist(metric, m, v, d);
            dsum += dd;
            dss[j] += dd * dd;
            nsum[j] += orc_dot(v, v, d);
        }
        compact[j] = dsum / (double)c;
        for (int i = 0; i < d; ++i) sumall[i] += m[i] * (double)c;    /* meanVector.scale(clusterCnt), merged */
        ++kk;
        total += c;
    }
    double ssb = 0.0, ssw = 0.0, comp = 0.0, sep = 0.0;
    double inv_total = 1.0 / (double)total;
    for (int i = 0; i < d; ++i) vec[i] = sumall[i] * inv_total;      /* overall mean */
    double *db = (double *)calloc((size_t)k, sizeof(double));
    if (!db) { free(vec); free(sumall); return -1; }
    for (int j = 0; j < k; ++j) {
        if (cnt[j] == 0.0) continue;
        double dm = eval_dist(metric, mean + (size_t)j * d, vec, d);
        ssb += pow(dm, 2) * cnt[j];
        ssw += dss[j];
        comp += compact[j];
    }
    for (int i = 0; i < k; ++i) {
        if (cnt[i] == 0.0) continue;
        for (int j = i + 1; j < k; ++j) {
            if (cnt[j] == 0.0) continue;
            double dm = eval_dist(metric, mean + (size_t)i * d, mean + (size_t)j * d, d);
            sep += dm;
            double tmp = (compact[i] + compact[j]) / dm;
            if (tmp > db[i]) db[i] = tmp;
            if (tmp > db[j]) db[j] = tmp;
        }
    }
    double dbsum = 0.0;
    for (int j = 0; j < k; ++j) if (cnt[j] != 0.0) dbsum += db[j];
    /* silhouette */
    double sil = 0.0;
    for (int64_t r = 0; r < n; ++r) {
        int lab = labels[r];
        if (lab < 0 || lab >= k) continue;
        const double *x = X + (size_t)r * d;
        double cur = 0.0, nb = DBL_MAX;
        double nsq = orc_dot(x, x, d);
        for (int j = 0; j < k; ++j) {
            if (cnt[j] == 0.0) continue;
            double dt = orc_dot(x, mean + (size_t)j * d, d);
            if (metric == ORC_EUCLIDEAN) {
                double diss = cnt[j] * nsq - 2.0 * cnt[j] * dt + nsum[j];
                if (j == lab) { if (cnt[j] > 1.0) cur = diss / (cnt[j] - 1.0); }
                else { double v = diss / cnt[j]; if (v < nb) nb = v; }
            } else {
                double diss = 1.0 - dt;
                if (j == lab) { if (cnt[j] > 1.0) cur = diss * cnt[j] / (cnt[j] - 1.0); }
                else if (diss < nb) nb = diss;
            }
        }
        sil += (cur < nb) ? 1.0 - (cur / nb) : (nb / cur) - 1.0;
    }
    out[0] = ssb; out[1] = ssw; out[2] = comp / kk; out[3] = 2.0 * sep / ((double)kk * kk - kk);
    out[4] = dbsum / kk; out[5] = ssb * (double)(total - kk) / ssw / (double)(kk - 1);
    out[6] = sil / (double)total; out[7] = (double)kk; out[8] = (double)total;
    free(vec); free(sumall); free(db);
    return 0;
}
