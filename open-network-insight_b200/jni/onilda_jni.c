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
 *
 * Arrays are reached with Get<Type>ArrayElements / Release<Type>ArrayElements, not with the "critical" variants:
 * every call here runs GPU work (copies, kernels, stream synchronisation) between get and release, and the JNI
 * specification forbids blocking inside a critical region (the collector may be held off for its whole length).
 */
#include <jni.h>
#include <stdint.h>
#include <stdlib.h>

#include "onilda.h"

#define CLS "org_opennetworkinsight_OniGpuScorer_00024"
#define FN(name) Java_org_opennetworkinsight_OniGpuScorer_00024_##name

#define PIN_I(arr) ((arr) ? (*e)->GetIntArrayElements(e, (arr), NULL) : NULL)
#define PIN_L(arr) ((arr) ? (*e)->GetLongArrayElements(e, (arr), NULL) : NULL)
#define PIN_D(arr) ((arr) ? (*e)->GetDoubleArrayElements(e, (arr), NULL) : NULL)
#define UNPIN_I(arr, p, mode) do { if (p) (*e)->ReleaseIntArrayElements(e, (arr), (p), (mode)); } while (0)
#define UNPIN_L(arr, p, mode) do { if (p) (*e)->ReleaseLongArrayElements(e, (arr), (p), (mode)); } while (0)
#define UNPIN_D(arr, p, mode) do { if (p) (*e)->ReleaseDoubleArrayElements(e, (arr), (p), (mode)); } while (0)

static void throw_last(JNIEnv* e)
{
    jclass c = (*e)->FindClass(e, "java/lang/RuntimeException");
    if (c) (*e)->ThrowNew(e, c, onilda_last_error());
}

JNIEXPORT jlong JNICALL FN(create)(JNIEnv* e, jobject self, jdoubleArray theta, jlong num_docs, jdoubleArray phi,
                                   jint num_terms, jint k, jint device)
{
    (void)self;
    jdouble* t = PIN_D(theta);
    jdouble* p = PIN_D(phi);
    onilda_scorer* s = NULL;
    int rc = (t && p) ? onilda_scorer_create(t, num_docs, p, num_terms, k, device, &s) : ONILDA_ENOMEM;
    UNPIN_D(phi, p, JNI_ABORT);
    UNPIN_D(theta, t, JNI_ABORT);
    if (rc != ONILDA_OK) { throw_last(e); return 0; }
    return (jlong)(intptr_t)s;
}

static jdoubleArray score_common(JNIEnv* e, jlong h, jintArray a, jintArray b, jintArray c, jintArray d)
{
    const jsize n = (*e)->GetArrayLength(e, a);
    jdoubleArray out = (*e)->NewDoubleArray(e, n);
    if (!out) return NULL;
    jint *pa = PIN_I(a), *pb = PIN_I(b), *pc = PIN_I(c), *pd = PIN_I(d);
    jdouble* po = PIN_D(out);
    int rc = ONILDA_ENOMEM;
    if (pa && pb && po && (!c || (pc && pd)))
        rc = c ? onilda_score_flow_min((onilda_scorer*)(intptr_t)h, (const int32_t*)pa, (const int32_t*)pb,
                                       (const int32_t*)pc, (const int32_t*)pd, n, po)
               : onilda_score((onilda_scorer*)(intptr_t)h, (const int32_t*)pa, (const int32_t*)pb, n, po);
    UNPIN_D(out, po, 0);
    UNPIN_I(d, pd, JNI_ABORT); UNPIN_I(c, pc, JNI_ABORT); UNPIN_I(b, pb, JNI_ABORT); UNPIN_I(a, pa, JNI_ABORT);
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
    jdouble* ps = PIN_D(score);
    int rc = (idx && val && ps) ? onilda_topk((onilda_scorer*)(intptr_t)h, ps, n, threshold, k, idx, val, &got) : ONILDA_ENOMEM;
    UNPIN_D(score, ps, JNI_ABORT);
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

/* ---------------------------------------------------------------------------------------------
 * The steps either side of LDA.  Scala side:
 *
 *   @native def flowWords(hour: Array[Int], minute: Array[Int], second: Array[Int], ibyt: Array[Long], ipkt: Array[Long],
 *                         sport: Array[Int], dport: Array[Int], cuts: Array[Double] (25: time, ibyt, ipkt),
 *                         srcCode: Array[Int], dstCode: Array[Int]): Unit          // FlowWordCreation.flowWordCreation
 *   @native def flowWordString(code: Int): String
 *   @native def pairsCount(doc: Array[Int], word: Array[Int], numDocs: Long, numTerms: Int,
 *                          docOut: Array[Int], wordOut: Array[Int], countOut: Array[Int]): Long   // reduceByKey(_ + _)
 *   @native def corpusFromTriples(doc: Array[Int], word: Array[Int], count: Array[Int], numDocs: Long, numTerms: Int): Long
 *   @native def corpusSaveDat(corpus: Long, path: String): Unit                   // createModel's lines, when the file is wanted
 *   @native def flowResults(h: Long, sdoc: Array[Int], sword: Array[Int], ddoc: Array[Int], dword: Array[Int],
 *                           threshold: Double, k: Long, idx: Array[Long], srcScore: Array[Double], dstScore: Array[Double]): Long
 */

JNIEXPORT void JNICALL FN(flowWords)(JNIEnv* e, jobject self, jintArray hour, jintArray minute, jintArray second, jlongArray ibyt,
                                     jlongArray ipkt, jintArray sport, jintArray dport, jdoubleArray cuts, jintArray src_code,
                                     jintArray dst_code)
{
    (void)self;
    const jsize n = (*e)->GetArrayLength(e, hour);
    jint *h = PIN_I(hour), *m = PIN_I(minute), *s = PIN_I(second), *sp = PIN_I(sport), *dp = PIN_I(dport), *sc = PIN_I(src_code), *dc = PIN_I(dst_code);
    jlong *b = PIN_L(ibyt), *p = PIN_L(ipkt);
    jdouble* c = PIN_D(cuts);
    onilda_flow_cuts fc;
    int rc = onilda_flow_words((const int32_t*)h, (const int32_t*)m, (const int32_t*)s, (const int64_t*)b, (const int64_t*)p,
                               (const int32_t*)sp, (const int32_t*)dp, n, -1, &fc, (int32_t*)sc, (int32_t*)dc);
    if (rc == ONILDA_OK && c) {
        for (int q = 0; q < 10; q++) { c[q] = fc.time[q]; c[10 + q] = fc.ibyt[q]; }
        for (int q = 0; q < 5; q++) c[20 + q] = fc.ipkt[q];
    }
    UNPIN_D(cuts, c, 0); UNPIN_I(dst_code, dc, 0); UNPIN_I(src_code, sc, 0);
    UNPIN_L(ipkt, p, JNI_ABORT); UNPIN_L(ibyt, b, JNI_ABORT); UNPIN_I(dport, dp, JNI_ABORT); UNPIN_I(sport, sp, JNI_ABORT);
    UNPIN_I(second, s, JNI_ABORT); UNPIN_I(minute, m, JNI_ABORT); UNPIN_I(hour, h, JNI_ABORT);
    if (rc != ONILDA_OK) throw_last(e);
}

JNIEXPORT jstring JNICALL FN(flowWordString)(JNIEnv* e, jobject self, jint code)
{
    (void)self;
    char buf[64];
    if (onilda_flow_word_string(code, buf, sizeof buf) < 0) { throw_last(e); return NULL; }
    return (*e)->NewStringUTF(e, buf);
}

JNIEXPORT jlong JNICALL FN(pairsCount)(JNIEnv* e, jobject self, jintArray doc, jintArray word, jlong num_docs, jint num_terms,
                                       jintArray doc_out, jintArray word_out, jintArray count_out)
{
    (void)self;
    const jsize n = (*e)->GetArrayLength(e, doc);
    jint *d = PIN_I(doc), *w = PIN_I(word), *od = PIN_I(doc_out), *ow = PIN_I(word_out), *oc = PIN_I(count_out);
    int64_t m = 0;
    int rc = onilda_pairs_count((const int32_t*)d, (const int32_t*)w, n, num_docs, num_terms, -1, (int32_t*)od, (int32_t*)ow,
                                (int32_t*)oc, &m);
    UNPIN_I(count_out, oc, 0); UNPIN_I(word_out, ow, 0); UNPIN_I(doc_out, od, 0); UNPIN_I(word, w, JNI_ABORT); UNPIN_I(doc, d, JNI_ABORT);
    if (rc != ONILDA_OK) { throw_last(e); return 0; }
    return (jlong)m;
}

JNIEXPORT jlong JNICALL FN(corpusFromTriples)(JNIEnv* e, jobject self, jintArray doc, jintArray word, jintArray count, jlong num_docs,
                                              jint num_terms)
{
    (void)self;
    const jsize n = (*e)->GetArrayLength(e, doc);
    jint *d = PIN_I(doc), *w = PIN_I(word), *c = PIN_I(count);
    onilda_corpus* corpus = NULL;
    int rc = onilda_corpus_from_triples((const int32_t*)d, (const int32_t*)w, (const int32_t*)c, n, num_docs, num_terms, -1, &corpus);
    UNPIN_I(count, c, JNI_ABORT); UNPIN_I(word, w, JNI_ABORT); UNPIN_I(doc, d, JNI_ABORT);
    if (rc != ONILDA_OK) { throw_last(e); return 0; }
    return (jlong)(intptr_t)corpus;
}

JNIEXPORT void JNICALL FN(corpusSaveDat)(JNIEnv* e, jobject self, jlong corpus, jstring path)
{
    (void)self;
    const onilda_corpus* c = (const onilda_corpus*)(intptr_t)corpus;
    int64_t docs = 0;
    const char* p = (*e)->GetStringUTFChars(e, path, NULL);
    int rc = p ? onilda_corpus_dims(c, &docs, NULL, NULL, NULL, NULL) : ONILDA_ENOMEM;
    if (rc == ONILDA_OK) rc = onilda_corpus_save_dat(c, 0, docs, p);
    if (p) (*e)->ReleaseStringUTFChars(e, path, p);
    if (rc != ONILDA_OK) throw_last(e);
}

JNIEXPORT jlong JNICALL FN(flowResults)(JNIEnv* e, jobject self, jlong h, jintArray sdoc, jintArray sword, jintArray ddoc,
                                        jintArray dword, jdouble threshold, jlong k, jlongArray idx, jdoubleArray src_score,
                                        jdoubleArray dst_score)
{
    (void)self;
    const jsize n = (*e)->GetArrayLength(e, sdoc);
    jint *a = PIN_I(sdoc), *b = PIN_I(sword), *c = PIN_I(ddoc), *d = PIN_I(dword);
    jlong* oi = PIN_L(idx);
    jdouble *os = PIN_D(src_score), *od = PIN_D(dst_score);
    int64_t kept = 0;
    int rc = onilda_flow_results((onilda_scorer*)(intptr_t)h, (const int32_t*)a, (const int32_t*)b, (const int32_t*)c,
                                 (const int32_t*)d, n, threshold, k, (int64_t*)oi, os, od, &kept);
    UNPIN_D(dst_score, od, 0); UNPIN_D(src_score, os, 0); UNPIN_L(idx, oi, 0);
    UNPIN_I(dword, d, JNI_ABORT); UNPIN_I(ddoc, c, JNI_ABORT); UNPIN_I(sword, b, JNI_ABORT); UNPIN_I(sdoc, a, JNI_ABORT);
    if (rc != ONILDA_OK) { throw_last(e); return 0; }
    return (jlong)kept;
}

/* dns / proxy rows (onilda_results_rows):
 *   @native def resultsRows(h: Long, doc: Array[Int], word: Array[Int], threshold: Double, k: Long,
 *                           idx: Array[Long], score: Array[Double]): Long */
JNIEXPORT jlong JNICALL FN(resultsRows)(JNIEnv* e, jobject self, jlong h, jintArray doc, jintArray word, jdouble threshold, jlong k,
                                        jlongArray idx, jdoubleArray score)
{
    (void)self;
    const jsize n = (*e)->GetArrayLength(e, doc);
    jint *a = PIN_I(doc), *b = PIN_I(word);
    jlong* oi = PIN_L(idx);
    jdouble* os = PIN_D(score);
    int64_t kept = 0;
    int rc = onilda_results_rows((onilda_scorer*)(intptr_t)h, (const int32_t*)a, (const int32_t*)b, n, threshold, k, (int64_t*)oi, os, &kept);
    UNPIN_D(score, os, 0); UNPIN_L(idx, oi, 0); UNPIN_I(word, b, JNI_ABORT); UNPIN_I(doc, a, JNI_ABORT);
    if (rc != ONILDA_OK) { throw_last(e); return 0; }
    return (jlong)kept;
}
