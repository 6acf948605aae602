/* JNI glue between AkmNative.java and libakm.so (include/akm.h).  Not built here: no JDK / jni.h in the image.
 * Build on a host with a JDK:  gcc -shared -fPIC -I$JAVA_HOME/include -I$JAVA_HOME/include/linux \
 *                                  -I../include akm_jni.c -L../alink_b200/csrc -lakm -o libakm_jni.so       */
#include <jni.h>
#include "akm.h"

#define H(h) ((akm_handle)(intptr_t)(h))

JNIEXPORT jint JNICALL Java_com_alibaba_alink_operator_common_clustering_kmeans_gpu_AkmNative_create(
    JNIEnv *env, jclass cls, jint device, jint d, jint metric, jlongArray out) {
    akm_handle h = NULL;
    int rc = akm_create(device, d, metric, &h);
    jlong v = (jlong)(intptr_t)h;
    (*env)->SetLongArrayRegion(env, out, 0, 1, &v);
    return rc;
}

JNIEXPORT jint JNICALL Java_com_alibaba_alink_operator_common_clustering_kmeans_gpu_AkmNative_destroy(
    JNIEnv *env, jclass cls, jlong h) { return akm_destroy(H(h)); }

JNIEXPORT jstring JNICALL Java_com_alibaba_alink_operator_common_clustering_kmeans_gpu_AkmNative_lastError(
    JNIEnv *env, jclass cls, jlong h) { return (*env)->NewStringUTF(env, akm_last_error(H(h))); }

JNIEXPORT jint JNICALL Java_com_alibaba_alink_operator_common_clustering_kmeans_gpu_AkmNative_uploadRows(
    JNIEnv *env, jclass cls, jlong h, jdoubleArray rows, jlong n) {
    jdouble *p = (*env)->GetPrimitiveArrayCritical(env, rows, NULL);
    int rc = akm_upload_rows(H(h), p, n);
    (*env)->ReleasePrimitiveArrayCritical(env, rows, p, JNI_ABORT);
    return rc;
}

JNIEXPORT jint JNICALL Java_com_alibaba_alink_operator_common_clustering_kmeans_gpu_AkmNative_uploadRowsPrepared(
    JNIEnv *env, jclass cls, jlong h, jdoubleArray rows, jlong n) {
    jdouble *p = (*env)->GetPrimitiveArrayCritical(env, rows, NULL);
    int rc = akm_upload_rows_prepared(H(h), p, n);       /* returns once the rows are staged: the array is free again */
    (*env)->ReleasePrimitiveArrayCritical(env, rows, p, JNI_ABORT);
    return rc;
}

JNIEXPORT jint JNICALL Java_com_alibaba_alink_operator_common_clustering_kmeans_gpu_AkmNative_uploadRowsCsr(
    JNIEnv *env, jclass cls, jlong h, jlongArray indptr, jintArray indices, jdoubleArray values, jlong n) {
    jlong *ip = (*env)->GetLongArrayElements(env, indptr, NULL);
    jint *ix = (*env)->GetIntArrayElements(env, indices, NULL);
    jdouble *v = (*env)->GetDoubleArrayElements(env, values, NULL);
    int rc = akm_upload_rows_csr(H(h), (const int64_t *)ip, (const int32_t *)ix, v, n);
    (*env)->ReleaseLongArrayElements(env, indptr, ip, JNI_ABORT);
    (*env)->ReleaseIntArrayElements(env, indices, ix, JNI_ABORT);
    (*env)->ReleaseDoubleArrayElements(env, values, v, JNI_ABORT);
    return rc;
}

JNIEXPORT jint JNICALL Java_com_alibaba_alink_operator_common_clustering_kmeans_gpu_AkmNative_uploadWait(
    JNIEnv *env, jclass cls, jlong h) { return akm_upload_wait(H(h)); }

JNIEXPORT jint JNICALL Java_com_alibaba_alink_operator_common_clustering_kmeans_gpu_AkmNative_superstep(
    JNIEnv *env, jclass cls, jlong h, jdoubleArray cent, jint k0, jint k, jint step, jdouble tol, jdoubleArray cout,
    jintArray kout, jintArray convout, jdoubleArray sumout, jdoubleArray shiftout, jdoubleArray objout) {
    jdouble *c = (*env)->GetDoubleArrayElements(env, cent, NULL);
    jdouble *co = (*env)->GetDoubleArrayElements(env, cout, NULL);
    jdouble *so = sumout ? (*env)->GetDoubleArrayElements(env, sumout, NULL) : NULL;
    jdouble *sh = shiftout ? (*env)->GetDoubleArrayElements(env, shiftout, NULL) : NULL;
    int nk = 0, conv = 0;
    double obj = 0.0;
    int rc = akm_superstep(H(h), c, k0, k, step, tol, co, &nk, &conv, so, sh, &obj);
    (*env)->ReleaseDoubleArrayElements(env, cent, c, JNI_ABORT);
    (*env)->ReleaseDoubleArrayElements(env, cout, co, 0);
    if (so) (*env)->ReleaseDoubleArrayElements(env, sumout, so, 0);
    if (sh) (*env)->ReleaseDoubleArrayElements(env, shiftout, sh, 0);
    jint v = nk, w = conv;
    (*env)->SetIntArrayRegion(env, kout, 0, 1, &v);
    if (convout) (*env)->SetIntArrayRegion(env, convout, 0, 1, &w);
    if (objout) (*env)->SetDoubleArrayRegion(env, objout, 0, 1, &obj);
    return rc;
}

JNIEXPORT jint JNICALL Java_com_alibaba_alink_operator_common_clustering_kmeans_gpu_AkmNative_assignAccumulate(
    JNIEnv *env, jclass cls, jlong h, jdoubleArray cent, jint k0, jint k, jdoubleArray sum, jdoubleArray obj) {
    jdouble *c = (*env)->GetDoubleArrayElements(env, cent, NULL);
    jdouble *s = (*env)->GetDoubleArrayElements(env, sum, NULL);
    double o = 0.0;
    int rc = akm_assign_accumulate(H(h), c, k0, k, s, &o, NULL);
    (*env)->ReleaseDoubleArrayElements(env, cent, c, JNI_ABORT);
    (*env)->ReleaseDoubleArrayElements(env, sum, s, 0);
    if (obj != NULL) (*env)->SetDoubleArrayRegion(env, obj, 0, 1, &o);
    return rc;
}

JNIEXPORT jint JNICALL Java_com_alibaba_alink_operator_common_clustering_kmeans_gpu_AkmNative_updateCentroids(
    JNIEnv *env, jclass cls, jlong h, jdoubleArray sum, jint k0, jint k, jint step, jdoubleArray prev,
    jdoubleArray cout, jintArray kout, jdoubleArray shift) {
    jdouble *s = (*env)->GetDoubleArrayElements(env, sum, NULL);
    jdouble *p = prev ? (*env)->GetDoubleArrayElements(env, prev, NULL) : NULL;
    jdouble *c = (*env)->GetDoubleArrayElements(env, cout, NULL);
    jdouble *sh = shift ? (*env)->GetDoubleArrayElements(env, shift, NULL) : NULL;
    int nk = 0;
    int rc = akm_update_centroids(H(h), s, k0, k, step, p, c, &nk, sh);
    (*env)->ReleaseDoubleArrayElements(env, sum, s, JNI_ABORT);
    if (p) (*env)->ReleaseDoubleArrayElements(env, prev, p, JNI_ABORT);
    (*env)->ReleaseDoubleArrayElements(env, cout, c, 0);
    if (sh) (*env)->ReleaseDoubleArrayElements(env, shift, sh, 0);
    jint v = nk;
    (*env)->SetIntArrayRegion(env, kout, 0, 1, &v);
    return rc;
}

JNIEXPORT jint JNICALL Java_com_alibaba_alink_operator_common_clustering_kmeans_gpu_AkmNative_commUniqueId(
    JNIEnv *env, jclass cls, jbyteArray id) {
    char buf[AKM_UNIQUE_ID_BYTES];
    int rc = akm_comm_unique_id(buf);
    (*env)->SetByteArrayRegion(env, id, 0, AKM_UNIQUE_ID_BYTES, (jbyte *)buf);
    return rc;
}

JNIEXPORT jint JNICALL Java_com_alibaba_alink_operator_common_clustering_kmeans_gpu_AkmNative_commInit(
    JNIEnv *env, jclass cls, jlong h, jbyteArray id, jint rank, jint world) {
    char buf[AKM_UNIQUE_ID_BYTES];
    (*env)->GetByteArrayRegion(env, id, 0, AKM_UNIQUE_ID_BYTES, (jbyte *)buf);
    return akm_comm_init(H(h), buf, rank, world);
}

JNIEXPORT jint JNICALL Java_com_alibaba_alink_operator_common_clustering_kmeans_gpu_AkmNative_allReduceSum(
    JNIEnv *env, jclass cls, jlong h, jdoubleArray buf, jlong count) {
    jdouble *b = (*env)->GetDoubleArrayElements(env, buf, NULL);
    int rc = akm_allreduce_sum(H(h), b, count);
    (*env)->ReleaseDoubleArrayElements(env, buf, b, 0);
    return rc;
}

JNIEXPORT jint JNICALL Java_com_alibaba_alink_operator_common_clustering_kmeans_gpu_AkmNative_predict(
    JNIEnv *env, jclass cls, jlong h, jdoubleArray rows, jlong n, jdoubleArray cent, jint k, jlongArray idx,
    jdoubleArray dist, jdoubleArray probs) {
    jdouble *x = (*env)->GetDoubleArrayElements(env, rows, NULL);
    jdouble *c = (*env)->GetDoubleArrayElements(env, cent, NULL);
    jlong *i = (*env)->GetLongArrayElements(env, idx, NULL);
    jdouble *d = dist ? (*env)->GetDoubleArrayElements(env, dist, NULL) : NULL;
    jdouble *p = probs ? (*env)->GetDoubleArrayElements(env, probs, NULL) : NULL;
    int rc = akm_predict(H(h), x, n, c, k, (int64_t *)i, d, p);
    (*env)->ReleaseDoubleArrayElements(env, rows, x, JNI_ABORT);
    (*env)->ReleaseDoubleArrayElements(env, cent, c, JNI_ABORT);
    (*env)->ReleaseLongArrayElements(env, idx, i, 0);
    if (d) (*env)->ReleaseDoubleArrayElements(env, dist, d, 0);
    if (p) (*env)->ReleaseDoubleArrayElements(env, probs, p, 0);
    return rc;
}

/* ---- device-resident loop ---------------------------------------------------------------------- */
JNIEXPORT jint JNICALL Java_com_alibaba_alink_operator_common_clustering_kmeans_gpu_AkmNative_setCentroids(JNIEnv *env, jclass cls, jlong h, jdoubleArray c0, jint k0) {
    jdouble *c = (*env)->GetDoubleArrayElements(env, c0, NULL);
    int rc = akm_set_centroids(H(h), c, k0);
    (*env)->ReleaseDoubleArrayElements(env, c0, c, JNI_ABORT);
    return rc;
}

JNIEXPORT jint JNICALL Java_com_alibaba_alink_operator_common_clustering_kmeans_gpu_AkmNative_step(JNIEnv *env, jclass cls, jlong h, jdouble tol, jintArray conv, jintArray k,
                                 jdoubleArray obj) {
    int cv = 0, kk = 0;
    double o = 0.0;
    int rc = akm_step(H(h), tol, &cv, &kk, &o);
    jint jc = cv, jk = kk;
    (*env)->SetIntArrayRegion(env, conv, 0, 1, &jc);
    (*env)->SetIntArrayRegion(env, k, 0, 1, &jk);
    (*env)->SetDoubleArrayRegion(env, obj, 0, 1, &o);
    return rc;
}

JNIEXPORT jint JNICALL Java_com_alibaba_alink_operator_common_clustering_kmeans_gpu_AkmNative_getModel(JNIEnv *env, jclass cls, jlong h, jdoubleArray cOut, jintArray kOut,
                                     jdoubleArray wOut, jintArray stepsOut) {
    jdouble *c = (*env)->GetDoubleArrayElements(env, cOut, NULL);
    jdouble *w = (*env)->GetDoubleArrayElements(env, wOut, NULL);
    int kk = 0, st = 0;
    int rc = akm_get_model(H(h), c, &kk, w, &st);
    jint jk = kk, js = st;
    (*env)->SetIntArrayRegion(env, kOut, 0, 1, &jk);
    (*env)->SetIntArrayRegion(env, stepsOut, 0, 1, &js);
    (*env)->ReleaseDoubleArrayElements(env, cOut, c, 0);
    (*env)->ReleaseDoubleArrayElements(env, wOut, w, 0);
    return rc;
}

/* ---- initial centroids -------------------------------------------------------------------------- */
JNIEXPORT jint JNICALL Java_com_alibaba_alink_operator_common_clustering_kmeans_gpu_AkmNative_initReset(JNIEnv *env, jclass cls, jlong h) { return akm_init_reset(H(h)); }

JNIEXPORT jint JNICALL Java_com_alibaba_alink_operator_common_clustering_kmeans_gpu_AkmNative_initUpdate(JNIEnv *env, jclass cls, jlong h, jdoubleArray centers, jint m, jint idBase,
                                       jdoubleArray sumOut) {
    jdouble *c = (*env)->GetDoubleArrayElements(env, centers, NULL);
    double s = 0.0;
    int rc = akm_init_update(H(h), c, m, idBase, &s);
    (*env)->ReleaseDoubleArrayElements(env, centers, c, JNI_ABORT);
    (*env)->SetDoubleArrayRegion(env, sumOut, 0, 1, &s);
    return rc;
}

JNIEXPORT jint JNICALL Java_com_alibaba_alink_operator_common_clustering_kmeans_gpu_AkmNative_initPick(JNIEnv *env, jclass cls, jlong h, jdouble scale, jlong seed, jint round,
                                     jlong rowBase, jint cap, jlongArray rowsOut, jintArray countOut) {
    jlong *r = (*env)->GetLongArrayElements(env, rowsOut, NULL);
    int cnt = 0;
    int rc = akm_init_pick(H(h), scale, (uint64_t)seed, round, rowBase, cap, (int64_t *)r, &cnt);
    jint jc = cnt;
    (*env)->SetIntArrayRegion(env, countOut, 0, 1, &jc);
    (*env)->ReleaseLongArrayElements(env, rowsOut, r, 0);
    return rc;
}

JNIEXPORT jint JNICALL Java_com_alibaba_alink_operator_common_clustering_kmeans_gpu_AkmNative_initCounts(JNIEnv *env, jclass cls, jlong h, jint mTotal, jdoubleArray countsOut) {
    jdouble *c = (*env)->GetDoubleArrayElements(env, countsOut, NULL);
    int rc = akm_init_counts(H(h), mTotal, c);
    (*env)->ReleaseDoubleArrayElements(env, countsOut, c, 0);
    return rc;
}

JNIEXPORT jint JNICALL Java_com_alibaba_alink_operator_common_clustering_kmeans_gpu_AkmNative_getRows(JNIEnv *env, jclass cls, jlong h, jlongArray rows, jint m, jdoubleArray out) {
    jlong *r = (*env)->GetLongArrayElements(env, rows, NULL);
    jdouble *o = (*env)->GetDoubleArrayElements(env, out, NULL);
    int rc = akm_get_rows(H(h), (const int64_t *)r, m, o);
    (*env)->ReleaseLongArrayElements(env, rows, r, JNI_ABORT);
    (*env)->ReleaseDoubleArrayElements(env, out, o, 0);
    return rc;
}

/* ---- cluster evaluation ------------------------------------------------------------------------- */
JNIEXPORT jint JNICALL Java_com_alibaba_alink_operator_common_clustering_kmeans_gpu_AkmNative_evalClusters(JNIEnv *env, jclass cls, jlong h, jintArray labels, jint k,
                                         jdoubleArray cnt, jdoubleArray mean, jdoubleArray dsum, jdoubleArray dsq,
                                         jdoubleArray nsq, jdoubleArray sil) {
    jint *l = (*env)->GetIntArrayElements(env, labels, NULL);
    jdouble *a = (*env)->GetDoubleArrayElements(env, cnt, NULL);
    jdouble *b = (*env)->GetDoubleArrayElements(env, mean, NULL);
    jdouble *c = (*env)->GetDoubleArrayElements(env, dsum, NULL);
    jdouble *d = (*env)->GetDoubleArrayElements(env, dsq, NULL);
    jdouble *e = (*env)->GetDoubleArrayElements(env, nsq, NULL);
    double s = 0.0;
    int rc = akm_eval_clusters(H(h), (const int32_t *)l, k, a, b, c, d, e, &s);
    (*env)->ReleaseIntArrayElements(env, labels, l, JNI_ABORT);
    (*env)->ReleaseDoubleArrayElements(env, cnt, a, 0);
    (*env)->ReleaseDoubleArrayElements(env, mean, b, 0);
    (*env)->ReleaseDoubleArrayElements(env, dsum, c, 0);
    (*env)->ReleaseDoubleArrayElements(env, dsq, d, 0);
    (*env)->ReleaseDoubleArrayElements(env, nsq, e, 0);
    (*env)->SetDoubleArrayRegion(env, sil, 0, 1, &s);
    return rc;
}
