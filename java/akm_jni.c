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
