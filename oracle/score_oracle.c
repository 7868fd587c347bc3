/*
 * score_oracle.c - plain-C restatement of the post-LDA arithmetic that the
 * reference does in Scala (no JVM exists in this image, so it cannot be run).
 * TEST INFRASTRUCTURE (see oracle.h).  Doubles, sums taken left to right from
 * 0.0 exactly as Scala's Array[Double].sum does.
 *
 * Pinned by the exact doubles in
 * oni-ml/src/test/scala/org/opennetworkinsight/OniLDACWrapperTest.scala:14-150
 * (theta / phi normalisation).  score, min, filter and top-k have no reference
 * test: "parity unpinned" for those.
 */
#include "oracle.h"

#include <math.h>
#include <pthread.h>
#include <stdlib.h>

/* libgomp is not installed in this image, so the all-cores scoring baseline
 * uses a small pthread fork/join over contiguous event ranges. */
static int g_threads = 1;
void oracle_set_threads(int n) { g_threads = n < 1 ? 1 : n; }

typedef struct {
    int64_t begin, end;
    const int32_t *sdoc, *sword, *ddoc, *dword;
    int64_t D; int V, K;
    const double *theta, *phi_wk;
    double* out;
} score_job;

/* oni-ml/src/main/scala/org/opennetworkinsight/OniLDACWrapper.scala:132-144 */
void oracle_theta_from_gamma(int64_t D, int K, const double* gamma, double* theta)
{
    for (int64_t d = 0; d < D; d++) {
        double s = 0.0;
        for (int k = 0; k < K; k++) s += gamma[d * K + k];
        for (int k = 0; k < K; k++) theta[d * K + k] = (s > 0) ? gamma[d * K + k] / s : 0.0;
    }
}

/* OniLDACWrapper.scala:121-130 (per topic line: exp, sum, divide) and :178-195
 * (transpose to word-major).  log_beta is [K][V]; phi_wk is [V][K]. */
void oracle_phi_from_logbeta(int K, int V, const double* log_beta, double* phi_wk)
{
    for (int k = 0; k < K; k++) {
        double s = 0.0;
        for (int w = 0; w < V; w++) {
            double p = exp(log_beta[(size_t)k * V + w]);
            phi_wk[(size_t)w * K + k] = p;
            s += p;
        }
        for (int w = 0; w < V; w++) phi_wk[(size_t)w * K + k] /= s;
    }
}

/* SuspiciousConnectsScoreFunction.scala:11-23 : zip, multiply, sum */
static double score_one(int32_t doc, int32_t word, int64_t D, int V, int K,
                        const double* theta, const double* phi_wk)
{
    double uniform = 1.0 / K, s = 0.0;
    const double* t = (doc >= 0 && doc < D) ? theta + (int64_t)doc * K : NULL;
    const double* p = (word >= 0 && word < V) ? phi_wk + (int64_t)word * K : NULL;
    for (int k = 0; k < K; k++) s += (t ? t[k] : uniform) * (p ? p[k] : uniform);
    return s;
}

static void* score_range(void* arg)
{
    score_job* j = (score_job*)arg;
    for (int64_t i = j->begin; i < j->end; i++) {
        double a = score_one(j->sdoc[i], j->sword[i], j->D, j->V, j->K, j->theta, j->phi_wk);
        if (j->ddoc) {
            double b = score_one(j->ddoc[i], j->dword[i], j->D, j->V, j->K, j->theta, j->phi_wk);
            j->out[i] = (a <= b) ? a : b;      /* scala.math.min */
            /* netflow/FlowPostLDA.scala:94: the events are inner-joined with the document table on the ip, for the
             * source and again for the destination, so an event with an ip that has no theta row leaves the result
             * set.  Per-event output has to keep the slot: +infinity, which no threshold keeps and no top-k takes. */
            if (!(j->sdoc[i] >= 0 && j->sdoc[i] < j->D) || !(j->ddoc[i] >= 0 && j->ddoc[i] < j->D)) j->out[i] = INFINITY;
        } else {
            j->out[i] = a;
        }
    }
    return NULL;
}

static void score_all(score_job proto, int64_t n)
{
    int t = g_threads;
    if (t <= 1 || n < 4096) { proto.begin = 0; proto.end = n; score_range(&proto); return; }
    pthread_t* tid = malloc(sizeof(pthread_t) * (size_t)t);
    score_job* jobs = malloc(sizeof(score_job) * (size_t)t);
    for (int i = 0; i < t; i++) {
        jobs[i] = proto;
        jobs[i].begin = n * i / t;
        jobs[i].end = n * (i + 1) / t;
        pthread_create(&tid[i], NULL, score_range, &jobs[i]);
    }
    for (int i = 0; i < t; i++) pthread_join(tid[i], NULL);
    free(tid); free(jobs);
}

void oracle_score(int64_t n, const int32_t* doc, const int32_t* word, int64_t D, int V, int K,
                  const double* theta, const double* phi_wk, double* out)
{
    score_job j = {0, 0, doc, word, NULL, NULL, D, V, K, theta, phi_wk, out};
    score_all(j, n);
}

/* netflow/FlowPostLDA.scala:99-106 (per side) and :115-125 (min of the two) */
void oracle_score_flow_min(int64_t n, const int32_t* sdoc, const int32_t* sword, const int32_t* ddoc,
                           const int32_t* dword, int64_t D, int V, int K, const double* theta,
                           const double* phi_wk, double* out)
{
    score_job j = {0, 0, sdoc, sword, ddoc, dword, D, V, K, theta, phi_wk, out};
    score_all(j, n);
}

typedef struct { double s; int64_t i; } scored;

static int by_score_then_index(const void* x, const void* y)
{
    const scored* a = (const scored*)x;
    const scored* b = (const scored*)y;
    if (a->s < b->s) return -1;
    if (a->s > b->s) return 1;
    return (a->i > b->i) - (a->i < b->i);
}

/* netflow/FlowPostLDA.scala:79-80 ; dns/DNSPostLDA.scala:49-70 ;
 * utilities/DataFrameUtils.scala:21-38.  The reference leaves tie order
 * unspecified; here ties are ordered by event index. */
int64_t oracle_topk(int64_t n, const double* score, double threshold, int64_t k,
                    int64_t* idx_out, double* score_out)
{
    scored* kept = malloc(sizeof(scored) * (size_t)(n > 0 ? n : 1));
    int64_t m = 0;
    for (int64_t i = 0; i < n; i++)
        if (score[i] <= threshold && !isinf(score[i])) { kept[m].s = score[i]; kept[m].i = i; m++; }   /* +inf: a row the flow join dropped */
    qsort(kept, (size_t)m, sizeof(scored), by_score_then_index);
    int64_t take = (k == -1 || m < k) ? m : k;
    for (int64_t j = 0; j < take; j++) { idx_out[j] = kept[j].i; score_out[j] = kept[j].s; }
    free(kept);
    return take;
}
