/*
 * oracle.h - CPU restatement of the reference hot path.  TEST INFRASTRUCTURE.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may
 * load liboracle.so.  The product library (libonilda.so) never links, loads or
 * calls anything declared here.
 *
 * Parity status: the lda-c part (parser, init, E-step, M-step, alpha Newton, EM
 * driver, model/gamma text I/O) is PINNED byte-for-byte against the unmodified
 * reference compiled into oracle/_ref/lda (tests/test_oracle_vs_ref.py) and
 * against fixtures that binary wrote (tests/golden/).  The scoring part is
 * PINNED against the exact doubles in the reference's own
 * OniLDACWrapperTest.scala for theta/phi normalisation; score/min/filter/top-k
 * have no reference test or fixture ("parity unpinned" for those four rows,
 * see DESIGN.md).
 *
 * All citations are relative to /root/reference/.
 */
#ifndef ONI_ORACLE_H
#define ONI_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
    int var_max_iter;      /* lda-inference.h:11  (int)   */
    float var_converged;   /* lda-inference.h:10  (float) */
    int em_max_iter;       /* lda-estimate.h:22           */
    float em_converged;    /* lda-estimate.h:21   (float) */
    int estimate_alpha;    /* lda-estimate.h:23           */
} oracle_settings;

/* ---- oni-ml/oni-lda-c/utils.c ---- */
double oracle_log_sum(double log_a, double log_b);   /* utils.c:8-21  */
double oracle_trigamma(double x);                    /* utils.c:32-47 */
double oracle_digamma(double x);                     /* utils.c:55-64 */

/* ---- oni-ml/oni-lda-c/cokus.c, lda-model.h:11 ---- */
void oracle_mt_seed(unsigned long seed);             /* cokus.c:52-101 */
unsigned long oracle_mt_next(void);                  /* cokus.c:104-145 */
double oracle_myrand(void);                          /* lda-model.h:11 */

/* ---- lda-estimate.c:331-350 ---- */
int oracle_read_settings(const char* path, oracle_settings* out);

/* ---- lda-data.c:22-59 : token-based .dat reader into CSR.  Buffers are
 * malloc'ed; release with oracle_free. */
int oracle_read_dat(const char* path, int64_t** offsets, int32_t** words, int32_t** counts,
                    int32_t** totals, int64_t* num_docs, int64_t* nnz, int32_t* num_terms);
void oracle_free(void* p);

/* ---- lda-alpha.c ---- */
double oracle_opt_alpha(double ss, int D, int K);    /* lda-alpha.c:42-68 */

/* ---- lda-model.c ---- log_beta is row-major [K][V] everywhere in the oracle. */
void oracle_lda_mle(int K, int V, const double* class_word, const double* class_total,
                    double* log_beta);                                            /* lda-model.c:27-44 */
void oracle_random_init(int K, int V, double* log_beta);                          /* lda-model.c:104-117 + lda_mle */
void oracle_seeded_init(int K, int V, int64_t D, const int64_t* offsets, const int32_t* words,
                        const int32_t* counts, double* log_beta);                 /* lda-model.c:120-144 + lda_mle */
int oracle_save_model(const char* root, int K, int V, double alpha, const double* log_beta); /* lda-model.c:193-217 */
int oracle_load_model_dims(const char* root, int* K, int* V);
int oracle_load_model(const char* root, int K, int V, double* alpha, double* log_beta);      /* lda-model.c:220-251 */
int oracle_save_gamma(const char* path, int64_t D, int K, const double* gamma);              /* lda-estimate.c:88-104 */

/* ---- lda-inference.c ---- */
double oracle_compute_likelihood(int len, const int32_t* words, const int32_t* counts, int K, int V,
                                 double alpha, const double* log_beta, const double* phi,
                                 const double* gamma);                            /* lda-inference.c:93-127 */
/* phi: len*K scratch, row n at phi[n*K].  Returns the last sweep's likelihood. */
double oracle_lda_inference(int len, const int32_t* words, const int32_t* counts, int total, int K, int V,
                            double alpha, const double* log_beta, int var_max_iter, float var_converged,
                            double* gamma, double* phi, int* sweeps);             /* lda-inference.c:27-85 */

/* One E-step pass over docs d = rank, rank+nproc, ... (lda-estimate.c:28-60, 201-209).
 * Accumulates INTO class_word [K][V], class_total [K], *alpha_ss, *num_docs and
 * returns the summed likelihood.  gamma is [D][K]; only the visited rows are written. */
double oracle_e_step(int64_t D, const int64_t* offsets, const int32_t* words, const int32_t* counts,
                     const int32_t* totals, int K, int V, double alpha, const double* log_beta,
                     int var_max_iter, float var_converged, int rank, int nproc,
                     double* gamma, double* class_word, double* class_total, double* alpha_ss,
                     int* num_docs, double* doc_likelihood /* may be NULL, [D] */);

/* run_em from a given model (lda-estimate.c:112-321).  nproc emulates the
 * reference's round-robin ranks with a rank-ordered Allreduce.  lik/conv hold
 * one entry per EM iteration (capacity em_max_iter + 1).  log_beta/alpha are
 * updated in place to the final model; gamma [D][K] is from the last E-step. */
int oracle_run_em(int64_t D, const int64_t* offsets, const int32_t* words, const int32_t* counts,
                  const int32_t* totals, int K, int V, double* alpha, double* log_beta,
                  const oracle_settings* settings, int nproc, double* gamma,
                  double* lik, double* conv, int* n_iter);

int oracle_doc_sweeps(int64_t D, const int64_t* offsets, const int32_t* words, const int32_t* counts,
                      const int32_t* totals, int K, int V, double alpha, const double* log_beta,
                      int var_max_iter, float var_converged, int32_t* sweeps_out);

/* infer (lda-estimate.c:358-388): per-document likelihood + gamma with a fixed model. */
int oracle_infer(int64_t D, const int64_t* offsets, const int32_t* words, const int32_t* counts,
                 const int32_t* totals, int K, int V, double alpha, const double* log_beta,
                 const oracle_settings* settings, double* gamma, double* doc_likelihood);

/* ---- post-LDA (Scala) restatement: score_oracle.c ---- */
/* host threads used by oracle_score / oracle_score_flow_min (default 1) */
void oracle_set_threads(int n);
/* OniLDACWrapper.scala:132-144 : theta row = gamma/sum, or zeros when sum <= 0. */
void oracle_theta_from_gamma(int64_t D, int K, const double* gamma, double* theta);
/* OniLDACWrapper.scala:121-130, 178-195 : phi[w][k] = exp(lb[k][w]) / sum_w exp(lb[k][w]). */
void oracle_phi_from_logbeta(int K, int V, const double* log_beta, double* phi_wk);
/* SuspiciousConnectsScoreFunction.scala:11-23 ; id < 0 => uniform 1/K. */
void oracle_score(int64_t n, const int32_t* doc, const int32_t* word, int64_t D, int V, int K,
                  const double* theta, const double* phi_wk, double* out);
/* FlowPostLDA.scala:99-106,115-125 : min(src, dst) per event. */
void oracle_score_flow_min(int64_t n, const int32_t* sdoc, const int32_t* sword, const int32_t* ddoc,
                           const int32_t* dword, int64_t D, int V, int K, const double* theta,
                           const double* phi_wk, double* out);
/* FlowPostLDA.scala:79-80, DataFrameUtils.scala:21-38 : keep score <= threshold,
 * ascending, first k (k == -1: all); ties broken by event index. Returns count. */
int64_t oracle_topk(int64_t n, const double* score, double threshold, int64_t k,
                    int64_t* idx_out, double* score_out);

#ifdef __cplusplus
}
#endif
#endif
