/*
 * JNI shim: lets oni-ml's Scala post-LDA scorer call the GPU scoring kernels in place of the
 * Spark UDF over broadcast maps (SuspiciousConnectsScoreFunction.scala:11-23,
 * netflow/FlowPostLDA.scala:94-125, dns/DNSPostLDA.scala:75-92).
 *
 * NOT BUILT IN THIS IMAGE: there is no JDK / jni.h here (probed: java, javac, jni.h absent), so this
 * file is compiled only where JAVA_HOME is set:
 *
 *   gcc -O2 -fPIC -shared -I"$JAVA_HOME/include" -I"$JAVA_HOME/include/linux" -I../../include \
 *       onilda_jni.c -L../lib -lonilda -Wl,-rpath,'$ORIGIN/../lib' -o ../lib/libonilda_jni.so
 *
 * Scala side (strings -> indices stay in Scala, which already owns wordDictionary and
 * documentDictionary, OniLDACWrapper.scala:40-60; an unknown ip / word is passed as -1):
 *
 *   object OniGpuScorer {
 *     System.loadLibrary("onilda_jni")
 *     @native def create(theta: Array[Double], numDocs: Long, phi: Array[Double], numTerms: Int, k: Int, device: Int): Long
 *     @native def score(h: Long, doc: Array[Int], word: Array[Int]): Array[Double]
 *     @native def scoreFlowMin(h: Long, sdoc: Array[Int], sword: Array[Int], ddoc: Array[Int], dword: Array[Int]): Array[Double]
 *     @native def topK(h: Long, score: Array[Double], threshold: Double, k: Long): Array[Long]
 *     @native def destroy(h: Long): Unit
 *   }
 */
#include <jni.h>
#include <stdint.h>
#include <stdlib.h>

#include "onilda.h"

#define CLS "org_opennetworkinsight_OniGpuScorer_00024"
#define FN(name) Java_org_opennetworkinsight_OniGpuScorer_00024_##name

static void throw_last(JNIEnv* e)
{
    jclass c = (*e)->FindClass(e, "java/lang/RuntimeException");
    if (c) (*e)->ThrowNew(e, c, onilda_last_error());
}

JNIEXPORT jlong JNICALL FN(create)(JNIEnv* e, jobject self, jdoubleArray theta, jlong num_docs, jdoubleArray phi,
                                   jint num_terms, jint k, jint device)
{
    (void)self;
    jdouble* t = (*e)->GetPrimitiveArrayCritical(e, theta, NULL);
    jdouble* p = (*e)->GetPrimitiveArrayCritical(e, phi, NULL);
    onilda_scorer* s = NULL;
    int rc = (t && p) ? onilda_scorer_create(t, num_docs, p, num_terms, k, device, &s) : ONILDA_ENOMEM;
    if (p) (*e)->ReleasePrimitiveArrayCritical(e, phi, p, JNI_ABORT);
    if (t) (*e)->ReleasePrimitiveArrayCritical(e, theta, t, JNI_ABORT);
    if (rc != ONILDA_OK) { throw_last(e); return 0; }
    return (jlong)(intptr_t)s;
}

static jdoubleArray score_common(JNIEnv* e, jlong h, jintArray a, jintArray b, jintArray c, jintArray d)
{
    const jsize n = (*e)->GetArrayLength(e, a);
    jdoubleArray out = (*e)->NewDoubleArray(e, n);
    if (!out) return NULL;
    jint* pa = (*e)->GetPrimitiveArrayCritical(e, a, NULL);
    jint* pb = (*e)->GetPrimitiveArrayCritical(e, b, NULL);
    jint* pc = c ? (*e)->GetPrimitiveArrayCritical(e, c, NULL) : NULL;
    jint* pd = d ? (*e)->GetPrimitiveArrayCritical(e, d, NULL) : NULL;
    jdouble* po = (*e)->GetPrimitiveArrayCritical(e, out, NULL);
    int rc = ONILDA_ENOMEM;
    if (pa && pb && po && (!c || (pc && pd)))
        rc = c ? onilda_score_flow_min((onilda_scorer*)(intptr_t)h, (const int32_t*)pa, (const int32_t*)pb,
                                       (const int32_t*)pc, (const int32_t*)pd, n, po)
               : onilda_score((onilda_scorer*)(intptr_t)h, (const int32_t*)pa, (const int32_t*)pb, n, po);
    if (po) (*e)->ReleasePrimitiveArrayCritical(e, out, po, 0);
    if (pd) (*e)->ReleasePrimitiveArrayCritical(e, d, pd, JNI_ABORT);
    if (pc) (*e)->ReleasePrimitiveArrayCritical(e, c, pc, JNI_ABORT);
    if (pb) (*e)->ReleasePrimitiveArrayCritical(e, b, pb, JNI_ABORT);
    if (pa) (*e)->ReleasePrimitiveArrayCritical(e, a, pa, JNI_ABORT);
    if (rc != ONILDA_OK) { throw_last(e); return NULL; }
    return out;
}

JNIEXPORT jdoubleArray JNICALL FN(score)(JNIEnv* e, jobject self, jlong h, jintArray doc, jintArray word)
{
    (void)self;
    return score_common(e, h, doc, word, NULL, NULL);
}

JNIEXPORT jdoubleArray JNICALL FN(scoreFlowMin)(JNIEnv* e, jobject self, jlong h, jintArray sdoc, jintArray sword,
                                                jintArray ddoc, jintArray dword)
{
    (void)self;
    return score_common(e, h, sdoc, sword, ddoc, dword);
}

/* indices of the k most suspicious events (score <= threshold, ascending, ties by index) */
JNIEXPORT jlongArray JNICALL FN(topK)(JNIEnv* e, jobject self, jlong h, jdoubleArray score, jdouble threshold, jlong k)
{
    (void)self;
    const jsize n = (*e)->GetArrayLength(e, score);
    int64_t cap = (k < 0 || k > n) ? n : k, got = 0;
    int64_t* idx = malloc(sizeof(int64_t) * (size_t)(cap > 0 ? cap : 1));
    double* val = malloc(sizeof(double) * (size_t)(cap > 0 ? cap : 1));
    jdouble* ps = (*e)->GetPrimitiveArrayCritical(e, score, NULL);
    int rc = (idx && val && ps) ? onilda_topk((onilda_scorer*)(intptr_t)h, ps, n, threshold, k, idx, val, &got) : ONILDA_ENOMEM;
    if (ps) (*e)->ReleasePrimitiveArrayCritical(e, score, ps, JNI_ABORT);
    jlongArray out = NULL;
    if (rc == ONILDA_OK) {
        out = (*e)->NewLongArray(e, (jsize)got);
        if (out) (*e)->SetLongArrayRegion(e, out, 0, (jsize)got, (const jlong*)idx);
    } else {
        throw_last(e);
    }
    free(idx);
    free(val);
    return out;
}

JNIEXPORT void JNICALL FN(destroy)(JNIEnv* e, jobject self, jlong h)
{
    (void)e; (void)self;
    onilda_scorer_destroy((onilda_scorer*)(intptr_t)h);
}
