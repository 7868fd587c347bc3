/*
 * lda_oracle.c - plain-C restatement of the reference oni-lda-c algorithm on
 * flat CSR / row-major arrays.  TEST INFRASTRUCTURE (see oracle.h): it is the
 * checker for the CUDA path and the "port" CPU baseline, never the product.
 *
 * Every function cites the reference lines it follows (paths relative to
 * /root/reference/).  Floating-point expressions keep the reference's operand
 * order so that, built with the same compiler and libm, results are
 * bit-identical to oracle/_ref/lda (checked by tests/test_oracle_vs_ref.py).
 */
#include "oracle.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------ utils.c */

/* oni-ml/oni-lda-c/utils.c:8-21 */
double oracle_log_sum(double log_a, double log_b)
{
    if (log_a < log_b) return log_b + log(1 + exp(log_a - log_b));
    return log_a + log(1 + exp(log_b - log_a));
}

/* oni-ml/oni-lda-c/utils.c:32-47 (A&S 6.4.11/6.4.12 with six recurrence steps) */
double oracle_trigamma(double x)
{
    double y = x + 6;
    double p = 1 / (y * y);
    p = (((((0.075757575757576 * p - 0.033333333333333) * p + 0.0238095238095238) * p
           - 0.033333333333333) * p + 0.166666666666667) * p + 1) / y + 0.5 * p;
    for (int step = 0; step < 6; step++) {
        y = y - 1;
        p = 1 / (y * y) + p;
    }
    return p;
}

/* oni-ml/oni-lda-c/utils.c:55-64 (truncated-decimal coefficients are the reference's) */
double oracle_digamma(double x)
{
    double y = x + 6;
    double p = 1 / (y * y);
    p = (((0.004166666666667 * p - 0.003968253986254) * p + 0.008333333333333) * p
         - 0.083333333333333) * p;
    p = p + log(y) - 0.5 / y - 1 / (y - 1) - 1 / (y - 2) - 1 / (y - 3) - 1 / (y - 4)
        - 1 / (y - 5) - 1 / (y - 6);
    return p;
}

/* ------------------------------------------------------------------ cokus.c */
/* MT19937 as in oni-ml/oni-lda-c/cokus.c:52-145, written with 32-bit words and
 * an index instead of the reference's pointer walk; the output stream is the
 * same (seed forced odd, state filled by x *= 69069, in-place regeneration of
 * all 624 words, standard tempering). */
#define MT_N 624
#define MT_M 397
static uint32_t mt_state[MT_N];
static int mt_pos = MT_N;   /* MT_N => regenerate before the next draw */
static int mt_seeded = 0;

void oracle_mt_seed(unsigned long seed)
{
    uint32_t x = ((uint32_t)seed) | 1u;
    for (int j = 0; j < MT_N; j++) {
        mt_state[j] = x;
        x *= 69069u;
    }
    mt_pos = MT_N;
    mt_seeded = 1;
}

static void mt_regenerate(void)
{
    for (int i = 0; i < MT_N; i++) {
        uint32_t cur = mt_state[i];
        uint32_t nxt = mt_state[(i + 1) % MT_N];
        uint32_t mixed = (cur & 0x80000000u) | (nxt & 0x7FFFFFFFu);
        mt_state[i] = mt_state[(i + MT_M) % MT_N] ^ (mixed >> 1) ^ ((nxt & 1u) ? 0x9908B0DFu : 0u);
    }
    mt_pos = 0;
}

unsigned long oracle_mt_next(void)
{
    if (!mt_seeded) oracle_mt_seed(4357u);   /* cokus.c:109-110 */
    if (mt_pos >= MT_N) mt_regenerate();
    uint32_t y = mt_state[mt_pos++];
    y ^= (y >> 11);
    y ^= (y << 7) & 0x9D2C5680u;
    y ^= (y << 15) & 0xEFC60000u;
    y ^= (y >> 18);
    return (unsigned long)y;
}

/* oni-ml/oni-lda-c/lda-model.h:11 */
double oracle_myrand(void) { return (double)(oracle_mt_next() / 4294967296.); }

/* ------------------------------------------------------------ lda-estimate.c */

/* oni-ml/oni-lda-c/lda-estimate.c:331-350 */
int oracle_read_settings(const char* path, oracle_settings* out)
{
    FILE* f = fopen(path, "r");
    char action[100];
    if (!f) return -1;
    action[0] = 0;
    fscanf(f, "var max iter %d\n", &out->var_max_iter);
    fscanf(f, "var convergence %f\n", &out->var_converged);
    fscanf(f, "em max iter %d\n", &out->em_max_iter);
    fscanf(f, "em convergence %f\n", &out->em_converged);
    fscanf(f, "alpha %s", action);
    out->estimate_alpha = (strcmp(action, "fixed") == 0) ? 0 : 1;
    fclose(f);
    return 0;
}

/* --------------------------------------------------------------- lda-data.c */

/* oni-ml/oni-lda-c/lda-data.c:22-59.  Same fscanf conversions ("%10d" then
 * "%10d:%10d" per entry, newline-agnostic, OFFSET 0, num_terms = max id + 1,
 * total = sum of counts), but filling CSR arrays instead of per-doc mallocs. */
int oracle_read_dat(const char* path, int64_t** offsets_out, int32_t** words_out, int32_t** counts_out,
                    int32_t** totals_out, int64_t* num_docs, int64_t* nnz, int32_t* num_terms)
{
    FILE* f = fopen(path, "r");
    if (!f) return -1;
    int64_t dcap = 1024, ncap = 1 << 16, nd = 0, nn = 0;
    int64_t* offsets = malloc(sizeof(int64_t) * (size_t)(dcap + 1));
    int32_t* totals = malloc(sizeof(int32_t) * (size_t)dcap);
    int32_t* words = malloc(sizeof(int32_t) * (size_t)ncap);
    int32_t* counts = malloc(sizeof(int32_t) * (size_t)ncap);
    int length, word = 0, count = 0, nw = 0;
    offsets[0] = 0;
    while (fscanf(f, "%10d", &length) != EOF) {
        if (nd == dcap) {
            dcap *= 2;
            offsets = realloc(offsets, sizeof(int64_t) * (size_t)(dcap + 1));
            totals = realloc(totals, sizeof(int32_t) * (size_t)dcap);
        }
        int total = 0;
        for (int n = 0; n < length; n++) {
            fscanf(f, "%10d:%10d", &word, &count);
            if (nn == ncap) {
                ncap *= 2;
                words = realloc(words, sizeof(int32_t) * (size_t)ncap);
                counts = realloc(counts, sizeof(int32_t) * (size_t)ncap);
            }
            words[nn] = word;
            counts[nn] = count;
            nn++;
            total += count;
            if (word >= nw) nw = word + 1;
        }
        totals[nd] = total;
        nd++;
        offsets[nd] = nn;
    }
    fclose(f);
    *offsets_out = offsets; *words_out = words; *counts_out = counts; *totals_out = totals;
    *num_docs = nd; *nnz = nn; *num_terms = nw;
    return 0;
}

void oracle_free(void* p) { free(p); }

/* -------------------------------------------------------------- lda-alpha.c */

/* oni-ml/oni-lda-c/lda-alpha.c:27-34 */
static double alpha_objective(double a, double ss, int D, int K)
{ return D * (lgamma(K * a) - K * lgamma(a)) + (a - 1) * ss; }
static double alpha_gradient(double a, double ss, int D, int K)
{ return D * (K * oracle_digamma(K * a) - K * oracle_digamma(a)) + ss; }
static double alpha_curvature(double a, int D, int K)
{ return D * (K * K * oracle_trigamma(K * a) - K * oracle_trigamma(a)); }

/* oni-ml/oni-lda-c/lda-alpha.c:42-68 : Newton on log(alpha) from 100 */
double oracle_opt_alpha(double ss, int D, int K)
{
    double start = 100, log_a = log(start), a, grad, curv;
    int iter = 0;
    do {
        iter++;
        a = exp(log_a);
        if (isnan(a)) {
            start = start * 10;
            a = start;
            log_a = log(a);
        }
        (void)alpha_objective(a, ss, D, K);
        grad = alpha_gradient(a, ss, D, K);
        curv = alpha_curvature(a, D, K);
        log_a = log_a - grad / (curv * a + grad);
    } while ((fabs(grad) > 1e-5) && (iter < 1000));   /* lda-alpha.h:11-12 */
    return exp(log_a);
}

/* -------------------------------------------------------------- lda-model.c */

/* oni-ml/oni-lda-c/lda-model.c:27-44 (alpha handled by the caller) */
void oracle_lda_mle(int K, int V, const double* class_word, const double* class_total, double* log_beta)
{
    for (int k = 0; k < K; k++)
        for (int w = 0; w < V; w++) {
            double cw = class_word[(size_t)k * V + w];
            log_beta[(size_t)k * V + w] = (cw > 0) ? log(cw) - log(class_total[k]) : -100;
        }
}

/* oni-ml/oni-lda-c/lda-model.c:104-117, then lda_mle as in lda-estimate.c:158-160 */
void oracle_random_init(int K, int V, double* log_beta)
{
    double* cw = calloc((size_t)K * V, sizeof(double));
    double* ct = calloc((size_t)K, sizeof(double));
    for (int k = 0; k < K; k++)
        for (int w = 0; w < V; w++) {
            cw[(size_t)k * V + w] += 1.0 / V + oracle_myrand();
            ct[k] += cw[(size_t)k * V + w];
        }
    oracle_lda_mle(K, V, cw, ct, log_beta);
    free(cw); free(ct);
}

/* oni-ml/oni-lda-c/lda-model.c:120-144 (NUM_INIT = 1), then lda_mle (lda-estimate.c:151-152) */
void oracle_seeded_init(int K, int V, int64_t D, const int64_t* offsets, const int32_t* words,
                        const int32_t* counts, double* log_beta)
{
    double* cw = calloc((size_t)K * V, sizeof(double));
    double* ct = calloc((size_t)K, sizeof(double));
    for (int k = 0; k < K; k++) {
        int d = (int)floor(oracle_myrand() * D);
        for (int64_t e = offsets[d]; e < offsets[d + 1]; e++)
            cw[(size_t)k * V + words[e]] += counts[e];
        for (int w = 0; w < V; w++) {
            cw[(size_t)k * V + w] += 1.0;
            ct[k] = ct[k] + cw[(size_t)k * V + w];
        }
    }
    oracle_lda_mle(K, V, cw, ct, log_beta);
    free(cw); free(ct);
}

/* oni-ml/oni-lda-c/lda-model.c:193-217 */
int oracle_save_model(const char* root, int K, int V, double alpha, const double* log_beta)
{
    char name[4096];
    snprintf(name, sizeof name, "%s.beta", root);
    FILE* f = fopen(name, "w");
    if (!f) return -1;
    for (int k = 0; k < K; k++) {
        for (int w = 0; w < V; w++) fprintf(f, " %5.10f", log_beta[(size_t)k * V + w]);
        fprintf(f, "\n");
    }
    fclose(f);
    snprintf(name, sizeof name, "%s.other", root);
    f = fopen(name, "w");
    if (!f) return -1;
    fprintf(f, "num_topics %d\n", K);
    fprintf(f, "num_terms %d\n", V);
    fprintf(f, "alpha %5.10f\n", alpha);
    fclose(f);
    return 0;
}

int oracle_load_model_dims(const char* root, int* K, int* V)
{
    char name[4096];
    snprintf(name, sizeof name, "%s.other", root);
    FILE* f = fopen(name, "r");
    if (!f) return -1;
    fscanf(f, "num_topics %d\n", K);
    fscanf(f, "num_terms %d\n", V);
    fclose(f);
    return 0;
}

/* oni-ml/oni-lda-c/lda-model.c:220-251 : beta and alpha go through FLOAT */
int oracle_load_model(const char* root, int K, int V, double* alpha, double* log_beta)
{
    char name[4096];
    int k2, v2;
    float a, x;
    snprintf(name, sizeof name, "%s.other", root);
    FILE* f = fopen(name, "r");
    if (!f) return -1;
    fscanf(f, "num_topics %d\n", &k2);
    fscanf(f, "num_terms %d\n", &v2);
    fscanf(f, "alpha %f\n", &a);
    fclose(f);
    if (k2 != K || v2 != V) return -2;
    *alpha = a;
    snprintf(name, sizeof name, "%s.beta", root);
    f = fopen(name, "r");
    if (!f) return -1;
    for (int k = 0; k < K; k++)
        for (int w = 0; w < V; w++) {
            fscanf(f, "%f", &x);
            log_beta[(size_t)k * V + w] = x;
        }
    fclose(f);
    return 0;
}

/* oni-ml/oni-lda-c/lda-estimate.c:88-104 */
int oracle_save_gamma(const char* path, int64_t D, int K, const double* gamma)
{
    FILE* f = fopen(path, "w");
    if (!f) return -1;
    for (int64_t d = 0; d < D; d++) {
        fprintf(f, "%5.10f", gamma[d * K]);
        for (int k = 1; k < K; k++) fprintf(f, " %5.10f", gamma[d * K + k]);
        fprintf(f, "\n");
    }
    fclose(f);
    return 0;
}

/* ---------------------------------------------------------- lda-inference.c */

/* oni-ml/oni-lda-c/lda-inference.c:93-127 ; k-outer / n-inner summation order */
double oracle_compute_likelihood(int len, const int32_t* words, const int32_t* counts, int K, int V,
                                 double alpha, const double* log_beta, const double* phi,
                                 const double* gamma)
{
    double lik = 0, digsum = 0, gamma_sum = 0;
    double* dig = malloc(sizeof(double) * (size_t)K);
    for (int k = 0; k < K; k++) {
        dig[k] = oracle_digamma(gamma[k]);
        gamma_sum += gamma[k];
    }
    digsum = oracle_digamma(gamma_sum);
    lik = lgamma(alpha * K) - K * lgamma(alpha) - (lgamma(gamma_sum));
    for (int k = 0; k < K; k++) {
        lik += (alpha - 1) * (dig[k] - digsum) + lgamma(gamma[k]) - (gamma[k] - 1) * (dig[k] - digsum);
        for (int n = 0; n < len; n++) {
            double p = phi[(size_t)n * K + k];
            if (p > 0)
                lik += counts[n] * (p * ((dig[k] - digsum) - log(p) + log_beta[(size_t)k * V + words[n]]));
        }
    }
    free(dig);
    return lik;
}

/* oni-ml/oni-lda-c/lda-inference.c:27-85 */
double oracle_lda_inference(int len, const int32_t* words, const int32_t* counts, int total, int K, int V,
                            double alpha, const double* log_beta, int var_max_iter, float var_converged,
                            double* gamma, double* phi, int* sweeps)
{
    double converged = 1, phisum = 0, lik = 0, lik_old = 0;
    double* oldphi = malloc(sizeof(double) * (size_t)K);
    double* dig = malloc(sizeof(double) * (size_t)K);
    int it = 0;

    for (int k = 0; k < K; k++) {
        gamma[k] = alpha + (total / ((double)K));
        dig[k] = oracle_digamma(gamma[k]);
        for (int n = 0; n < len; n++) phi[(size_t)n * K + k] = 1.0 / K;
    }
    while ((converged > var_converged) && ((it < var_max_iter) || (var_max_iter == -1))) {
        it++;
        for (int n = 0; n < len; n++) {
            double* row = phi + (size_t)n * K;
            phisum = 0;
            for (int k = 0; k < K; k++) {
                oldphi[k] = row[k];
                row[k] = dig[k] + log_beta[(size_t)k * V + words[n]];
                if (k > 0) phisum = oracle_log_sum(phisum, row[k]);
                else phisum = row[k];
            }
            for (int k = 0; k < K; k++) {
                row[k] = exp(row[k] - phisum);
                gamma[k] = gamma[k] + counts[n] * (row[k] - oldphi[k]);
                dig[k] = oracle_digamma(gamma[k]);
            }
        }
        lik = oracle_compute_likelihood(len, words, counts, K, V, alpha, log_beta, phi, gamma);
        if (isnan(lik)) { fprintf(stderr, "oracle: NaN likelihood (reference asserts here)\n"); abort(); }
        converged = (lik_old - lik) / lik_old;
        lik_old = lik;
    }
    if (sweeps) *sweeps = it;
    free(oldphi); free(dig);
    return lik;
}

/* ------------------------------------------------------------ lda-estimate.c */

/* doc_e_step (lda-estimate.c:28-60) over the docs one rank owns (:201-209) */
double oracle_e_step(int64_t D, const int64_t* offsets, const int32_t* words, const int32_t* counts,
                     const int32_t* totals, int K, int V, double alpha, const double* log_beta,
                     int var_max_iter, float var_converged, int rank, int nproc,
                     double* gamma, double* class_word, double* class_total, double* alpha_ss,
                     int* num_docs, double* doc_likelihood)
{
    int64_t max_len = 0;
    for (int64_t d = 0; d < D; d++)
        if (offsets[d + 1] - offsets[d] > max_len) max_len = offsets[d + 1] - offsets[d];
    double* phi = malloc(sizeof(double) * (size_t)(max_len > 0 ? max_len : 1) * K);
    double lik_sum = 0;
    for (int64_t d = rank; d < D; d += nproc) {
        int len = (int)(offsets[d + 1] - offsets[d]);
        const int32_t* w = words + offsets[d];
        const int32_t* c = counts + offsets[d];
        double* g = gamma + d * K;
        double lik = oracle_lda_inference(len, w, c, totals[d], K, V, alpha, log_beta,
                                          var_max_iter, var_converged, g, phi, NULL);
        double gamma_sum = 0;
        for (int k = 0; k < K; k++) {
            gamma_sum += g[k];
            *alpha_ss += oracle_digamma(g[k]);
        }
        *alpha_ss -= K * oracle_digamma(gamma_sum);
        for (int n = 0; n < len; n++)
            for (int k = 0; k < K; k++) {
                class_word[(size_t)k * V + w[n]] += c[n] * phi[(size_t)n * K + k];
                class_total[k] += c[n] * phi[(size_t)n * K + k];
            }
        *num_docs = *num_docs + 1;
        if (doc_likelihood) doc_likelihood[d] = lik;
        lik_sum += lik;
    }
    free(phi);
    return lik_sum;
}

/* run_em (lda-estimate.c:112-321) from a given start model.  With nproc > 1
 * each emulated rank fills its own suffstats and the totals are formed as
 * 0.0 + rank0 + rank1 + ... , which is what oracle/mpi_stub's Allreduce does. */
int oracle_run_em(int64_t D, const int64_t* offsets, const int32_t* words, const int32_t* counts,
                  const int32_t* totals, int K, int V, double* alpha, double* log_beta,
                  const oracle_settings* settings, int nproc, double* gamma,
                  double* lik_hist, double* conv_hist, int* n_iter)
{
    int var_max_iter = settings->var_max_iter;       /* mutated at lda-estimate.c:261 */
    size_t kv = (size_t)K * V;
    if (nproc < 1) nproc = 1;
    double* cw = malloc(sizeof(double) * kv * (size_t)nproc);
    double* ct = malloc(sizeof(double) * (size_t)K * nproc);
    double* cw_sum = nproc > 1 ? malloc(sizeof(double) * kv) : cw;
    double* ct_sum = nproc > 1 ? malloc(sizeof(double) * (size_t)K) : ct;
    int i = 0;
    double lik, lik_old = 0, converged = 1;

    while (((converged < 0) || (converged > settings->em_converged) || (i <= 2)) && (i <= settings->em_max_iter)) {
        i++;
        double alpha_ss = 0.0, lik_tot = 0.0;
        int ndocs = 0;
        memset(cw, 0, sizeof(double) * kv * (size_t)nproc);
        memset(ct, 0, sizeof(double) * (size_t)K * nproc);
        if (nproc == 1) {
            lik_tot = oracle_e_step(D, offsets, words, counts, totals, K, V, *alpha, log_beta,
                                    var_max_iter, settings->var_converged, 0, 1, gamma,
                                    cw, ct, &alpha_ss, &ndocs, NULL);
        } else {
            for (int r = 0; r < nproc; r++) {
                double a_r = 0.0;
                int n_r = 0;
                double l_r = oracle_e_step(D, offsets, words, counts, totals, K, V, *alpha, log_beta,
                                           var_max_iter, settings->var_converged, r, nproc, gamma,
                                           cw + kv * r, ct + (size_t)K * r, &a_r, &n_r, NULL);
                alpha_ss += a_r; ndocs += n_r; lik_tot += l_r;
            }
            for (size_t j = 0; j < kv; j++) {
                double acc = 0.0;
                for (int r = 0; r < nproc; r++) acc += cw[kv * r + j];
                cw_sum[j] = acc;
            }
            for (int k = 0; k < K; k++) {
                double acc = 0.0;
                for (int r = 0; r < nproc; r++) acc += ct[(size_t)K * r + k];
                ct_sum[k] = acc;
            }
        }
        lik = lik_tot;
        /* M-step: lda_mle (lda-model.c:27-54) */
        oracle_lda_mle(K, V, cw_sum, ct_sum, log_beta);
        if (settings->estimate_alpha == 1) *alpha = oracle_opt_alpha(alpha_ss, ndocs, K);
        /* lda-estimate.c:260-262 */
        converged = (lik_old - lik) / (lik_old);
        if (converged < 0) var_max_iter = var_max_iter * 2;
        lik_old = lik;
        lik_hist[i - 1] = lik;
        conv_hist[i - 1] = converged;
    }
    *n_iter = i;
    if (nproc > 1) { free(cw_sum); free(ct_sum); }
    free(cw); free(ct);
    return 0;
}

/* Sweeps (var_iter, lda-inference.c:48) each document takes with a fixed model; test aid. */
int oracle_doc_sweeps(int64_t D, const int64_t* offsets, const int32_t* words, const int32_t* counts,
                      const int32_t* totals, int K, int V, double alpha, const double* log_beta,
                      int var_max_iter, float var_converged, int32_t* sweeps_out)
{
    double* gamma = malloc(sizeof(double) * (size_t)K);
    for (int64_t d = 0; d < D; d++) {
        int len = (int)(offsets[d + 1] - offsets[d]), it = 0;
        double* phi = malloc(sizeof(double) * (size_t)(len > 0 ? len : 1) * K);
        oracle_lda_inference(len, words + offsets[d], counts + offsets[d], totals[d], K, V, alpha, log_beta,
                             var_max_iter, var_converged, gamma, phi, &it);
        sweeps_out[d] = it;
        free(phi);
    }
    free(gamma);
    return 0;
}

/* infer (lda-estimate.c:358-388) */
int oracle_infer(int64_t D, const int64_t* offsets, const int32_t* words, const int32_t* counts,
                 const int32_t* totals, int K, int V, double alpha, const double* log_beta,
                 const oracle_settings* settings, double* gamma, double* doc_likelihood)
{
    for (int64_t d = 0; d < D; d++) {
        int len = (int)(offsets[d + 1] - offsets[d]);
        double* phi = malloc(sizeof(double) * (size_t)(len > 0 ? len : 1) * K);
        doc_likelihood[d] = oracle_lda_inference(len, words + offsets[d], counts + offsets[d], totals[d],
                                                 K, V, alpha, log_beta, settings->var_max_iter,
                                                 settings->var_converged, gamma + d * K, phi, NULL);
        free(phi);
    }
    return 0;
}
