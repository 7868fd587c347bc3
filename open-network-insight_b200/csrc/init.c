/*
 * Start models for `lda est ... random|seeded` (reference:
 * oni-ml/oni-lda-c/lda-model.c:104-144 driven from lda-estimate.c:147-162, with
 * the Mersenne Twister of cokus.c:52-145 behind lda-model.h:11's myrand()).
 *
 * Host code on purpose: it is one pass of K*V sequential RNG draws made once
 * per run, and reproducing the reference's exact stream (so that a given seed
 * yields the reference's own 000.beta) needs the generator's serial order.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "onilda_internal.h"

enum { MT_LEN = 624, MT_LAG = 397 };

typedef struct {
    uint32_t word[MT_LEN];
    int next;
} mt19937;

/* cokus.c:52-101: seed forced odd, then x <- 69069 * x (mod 2^32) */
static void mt_seed(mt19937* g, uint32_t seed)
{
    uint32_t x = seed | 1u;
    for (int i = 0; i < MT_LEN; i++) { g->word[i] = x; x *= 69069u; }
    g->next = MT_LEN;
}

static void mt_twist(mt19937* g)
{
    for (int i = 0; i < MT_LEN; i++) {
        uint32_t a = g->word[i], b = g->word[(i + 1) % MT_LEN];
        uint32_t y = (a & 0x80000000u) | (b & 0x7FFFFFFFu);
        g->word[i] = g->word[(i + MT_LAG) % MT_LEN] ^ (y >> 1) ^ ((b & 1u) ? 0x9908B0DFu : 0u);
    }
    g->next = 0;
}

static uint32_t mt_draw(mt19937* g)
{
    if (g->next >= MT_LEN) mt_twist(g);
    uint32_t y = g->word[g->next++];
    y ^= y >> 11;
    y ^= (y << 7) & 0x9D2C5680u;
    y ^= (y << 15) & 0xEFC60000u;
    y ^= y >> 18;
    return y;
}

/* lda-model.h:11 : randomMT() / 2^32 */
static double mt_uniform(mt19937* g) { return (double)(mt_draw(g) / 4294967296.); }

/* lda_mle without the alpha part (lda-model.c:27-44) on [K][V] statistics */
static void normalise_log(int K, int V, const double* cw, const double* ct, double* log_beta)
{
    for (int k = 0; k < K; k++) {
        const double lt = log(ct[k]);
        for (int w = 0; w < V; w++) {
            double x = cw[(size_t)k * V + w];
            log_beta[(size_t)k * V + w] = x > 0 ? log(x) - lt : -100;
        }
    }
}

int onilda_model_init_random(onilda_model* m, uint32_t seed)
{
    if (!m) return onilda_fail(ONILDA_EINVAL, "model_init_random: null model");
    const int K = m->num_topics, V = m->num_terms;
    double* cw = (double*)calloc((size_t)K * V + 1, sizeof(double));
    double* ct = (double*)calloc((size_t)K, sizeof(double));
    if (!cw || !ct) { free(cw); free(ct); return onilda_fail(ONILDA_ENOMEM, "model_init_random: out of host memory"); }
    mt19937 g;
    mt_seed(&g, seed);
    for (int k = 0; k < K; k++)                       /* k-major, n-minor draw order (lda-model.c:109-116) */
        for (int w = 0; w < V; w++) {
            cw[(size_t)k * V + w] += 1.0 / V + mt_uniform(&g);
            ct[k] += cw[(size_t)k * V + w];
        }
    normalise_log(K, V, cw, ct, m->log_beta);
    free(cw); free(ct);
    return ONILDA_OK;
}

int onilda_model_init_seeded(onilda_model* m, const onilda_corpus* c, uint32_t seed)
{
    if (!m || !c) return onilda_fail(ONILDA_EINVAL, "model_init_seeded: null argument");
    if (c->num_docs <= 0) return onilda_fail(ONILDA_EINVAL, "model_init_seeded: empty corpus");
    if (c->num_terms > m->num_terms) return onilda_fail(ONILDA_EINVAL, "model_init_seeded: corpus has more terms than the model");
    const int K = m->num_topics, V = m->num_terms;
    double* cw = (double*)calloc((size_t)K * V + 1, sizeof(double));
    double* ct = (double*)calloc((size_t)K, sizeof(double));
    if (!cw || !ct) { free(cw); free(ct); return onilda_fail(ONILDA_ENOMEM, "model_init_seeded: out of host memory"); }
    mt19937 g;
    mt_seed(&g, seed);
    for (int k = 0; k < K; k++) {
        /* NUM_INIT = 1 document per topic (lda-model.h:12, lda-model.c:128-137) */
        int64_t d = (int64_t)floor(mt_uniform(&g) * (double)(int)c->num_docs);
        for (int64_t e = c->offsets[d]; e < c->offsets[d + 1]; e++) cw[(size_t)k * V + c->words[e]] += c->counts[e];
        for (int w = 0; w < V; w++) {
            cw[(size_t)k * V + w] += 1.0;
            ct[k] = ct[k] + cw[(size_t)k * V + w];
        }
    }
    normalise_log(K, V, cw, ct, m->log_beta);
    free(cw); free(ct);
    return ONILDA_OK;
}
