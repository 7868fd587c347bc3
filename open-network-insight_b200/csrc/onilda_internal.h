/* Internal declarations shared by the host C files and the CUDA files of libonilda. */
#ifndef ONILDA_INTERNAL_H
#define ONILDA_INTERNAL_H

#include <stddef.h>
#include <stdint.h>

#include "onilda.h"

#ifdef __cplusplus
extern "C" {
#endif

struct onilda_corpus {
    int64_t num_docs;
    int64_t nnz;
    int32_t num_terms;
    int64_t tokens;
    int32_t max_len;
    int64_t* offsets;  /* [D+1] */
    int32_t* words;    /* [N]   */
    int32_t* counts;   /* [N]   */
    int32_t* totals;   /* [D] int sum of counts (lda.h:27) */
    int pinned;        /* host arrays registered with the CUDA driver */
};

struct onilda_model {
    int num_topics;
    int num_terms;
    double alpha;
    double* log_beta;  /* [K][V] */
};

/* error.c */
int onilda_fail(int code, const char* fmt, ...);

/* hostmath.c : the reference's own series (utils.c:32-64) for the host-side alpha Newton */
double onilda_host_digamma(double x);
double onilda_host_trigamma(double x);
double onilda_opt_alpha(double ss, int D, int K);
/* tables of device_math.cuh's psi_exp_tab: rows[(octaves << mant_bits)][6], tab[n] */
void onilda_build_psi_table(double* rows, int mant_bits, int octaves);
void onilda_build_exp_table(double* tab, int n);

/* em.cu */
void onilda_corpus_unpin(onilda_corpus* c);
/* device views of a handle's results for the scorer hand-off (score.cu); waits for the handle's stream */
int onilda_em_device_views(onilda_em* em, int* device, int64_t* num_docs, int* num_topics, int* num_terms, int* kpad,
                           const double** gamma_dev, const double** log_beta_wk_dev);

/* textio.c */
/* formats v as printf("%5.10f") would; returns the number of chars written (no NUL). */
int onilda_format_fixed10(double v, char* out);

#ifdef __cplusplus
}
#endif
#endif
