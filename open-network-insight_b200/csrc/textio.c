/*
 * Host-side file boundary of the path: the settings file, the `.dat` corpus,
 * `.beta` / `.other` / `.gamma` text files.  These formats ARE the reference's
 * plug-in interface (OniLDACWrapper.scala:83-112 execs `lda` and reads the
 * files back), so parsing and printing are bit/byte exact.
 *
 * Reference: oni-ml/oni-lda-c/lda-data.c:22-59, lda-estimate.c:88-104, 331-350,
 * lda-model.c:193-251.
 */
#include <ctype.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "onilda_internal.h"

/* ------------------------------------------------------------------ settings */

/* read_settings (lda-estimate.c:331-350): literal lines in a fixed order; %f into float. */
int onilda_settings_read(const char* path, onilda_settings* out)
{
    if (!path || !out) return onilda_fail(ONILDA_EINVAL, "settings_read: null argument");
    FILE* f = fopen(path, "r");
    if (!f) return onilda_fail(ONILDA_EIO, "cannot open settings file %s", path);
    char action[100] = "";
    int got = 0;
    got += fscanf(f, "var max iter %d\n", &out->var_max_iter) == 1;
    got += fscanf(f, "var convergence %f\n", &out->var_converged) == 1;
    got += fscanf(f, "em max iter %d\n", &out->em_max_iter) == 1;
    got += fscanf(f, "em convergence %f\n", &out->em_converged) == 1;
    got += fscanf(f, "alpha %99s", action) == 1;
    fclose(f);
    if (got != 5) return onilda_fail(ONILDA_EINVAL, "settings file %s: expected the five lda-c lines", path);
    out->estimate_alpha = strcmp(action, "fixed") == 0 ? 0 : 1;
    return ONILDA_OK;
}

/* -------------------------------------------------------------------- corpus */

typedef struct {
    const unsigned char* p;
    const unsigned char* end;
} cursor;

/* One "%10d" conversion: skip white space, then an optional sign and decimal
 * digits, at most 10 characters in all.  Returns 1 (converted), 0 (matching
 * failure) or -1 (end of input before any character: scanf's EOF). */
static int scan_int10(cursor* c, int* out)
{
    while (c->p < c->end && isspace(*c->p)) c->p++;
    if (c->p >= c->end) return -1;
    const unsigned char* q = c->p;
    int width = 10, neg = 0, ndig = 0;
    int64_t v = 0;
    if (*q == '+' || *q == '-') { neg = (*q == '-'); q++; width--; }
    while (q < c->end && width > 0 && *q >= '0' && *q <= '9') { v = v * 10 + (*q - '0'); q++; width--; ndig++; }
    if (ndig == 0) return 0;
    c->p = q;
    *out = (int)(uint32_t)(neg ? -v : v);   /* int conversion wraps as scanf's does */
    return 1;
}

static int grow(void** buf, size_t elem, int64_t* cap, int64_t need)
{
    if (need <= *cap) return 1;
    int64_t ncap = *cap ? *cap : 1024;
    while (ncap < need) ncap *= 2;
    void* nb = realloc(*buf, elem * (size_t)ncap);
    if (!nb) return 0;
    *buf = nb;
    *cap = ncap;
    return 1;
}

static void corpus_finish(onilda_corpus* c)
{
    int64_t tokens = 0;
    int32_t max_len = 0;
    for (int64_t d = 0; d < c->num_docs; d++) {
        int64_t len = c->offsets[d + 1] - c->offsets[d];
        if (len > max_len) max_len = (int32_t)len;
    }
    for (int64_t e = 0; e < c->nnz; e++) tokens += c->counts[e];
    c->tokens = tokens;
    c->max_len = max_len;
}

/* read_data (lda-data.c:22-59) into CSR. */
int onilda_corpus_from_dat(const char* path, onilda_corpus** out)
{
    if (!path || !out) return onilda_fail(ONILDA_EINVAL, "corpus_from_dat: null argument");
    FILE* f = fopen(path, "rb");
    if (!f) return onilda_fail(ONILDA_EIO, "cannot open corpus %s", path);
    fseek(f, 0, SEEK_END);
    long size = ftell(f);
    fseek(f, 0, SEEK_SET);
    unsigned char* text = (unsigned char*)malloc((size_t)size + 1);
    if (!text) { fclose(f); return onilda_fail(ONILDA_ENOMEM, "corpus %s: %ld bytes do not fit in memory", path, size); }
    if (size > 0 && fread(text, 1, (size_t)size, f) != (size_t)size) {
        fclose(f); free(text);
        return onilda_fail(ONILDA_EIO, "short read on %s", path);
    }
    fclose(f);

    onilda_corpus* c = (onilda_corpus*)calloc(1, sizeof *c);
    int64_t dcap = 0, tcap = 0, ncap_w = 0, ncap_c = 0;
    cursor cur = { text, text + size };
    int length = 0, word = 0, count = 0, nterms = 0, rc = ONILDA_OK;
    if (!c || !grow((void**)&c->offsets, sizeof(int64_t), &dcap, 1)) { rc = ONILDA_ENOMEM; goto done; }
    c->offsets[0] = 0;
    for (;;) {
        int r = scan_int10(&cur, &length);
        if (r < 0) break;                               /* EOF ends the outer while (lda-data.c:35) */
        if (r == 0) { rc = onilda_fail(ONILDA_EINVAL, "corpus %s: expected a document length at byte %ld", path,
                                       (long)(cur.p - text)); goto done; }
        if (!grow((void**)&c->offsets, sizeof(int64_t), &dcap, c->num_docs + 2) ||
            !grow((void**)&c->totals, sizeof(int32_t), &tcap, c->num_docs + 1) ||
            (length > 0 && (!grow((void**)&c->words, sizeof(int32_t), &ncap_w, c->nnz + length) ||
                            !grow((void**)&c->counts, sizeof(int32_t), &ncap_c, c->nnz + length)))) {
            rc = ONILDA_ENOMEM; goto done;
        }
        int total = 0;
        for (int n = 0; n < length; n++) {
            /* "%10d:%10d": on a failed conversion the reference keeps the previous
             * values of word / count (it ignores fscanf's result, lda-data.c:44). */
            int r1 = scan_int10(&cur, &word);
            if (r1 == 1 && cur.p < cur.end && *cur.p == ':') {
                cur.p++;
                scan_int10(&cur, &count);
            } else if (r1 == 0) {
                rc = onilda_fail(ONILDA_EINVAL, "corpus %s: expected word:count at byte %ld", path, (long)(cur.p - text));
                goto done;
            }
            if (word < 0) {
                rc = onilda_fail(ONILDA_EINVAL, "corpus %s: negative word id %d in document %lld", path, word,
                                 (long long)c->num_docs);
                goto done;
            }
            c->words[c->nnz] = word;
            c->counts[c->nnz] = count;
            c->nnz++;
            total += count;
            if (word >= nterms) nterms = word + 1;
        }
        c->totals[c->num_docs] = total;
        c->num_docs++;
        c->offsets[c->num_docs] = c->nnz;
    }
    c->num_terms = nterms;
    corpus_finish(c);
done:
    free(text);
    if (rc != ONILDA_OK) {
        if (rc == ONILDA_ENOMEM) onilda_fail(rc, "corpus %s: out of host memory", path);
        onilda_corpus_free(c);
        return rc;
    }
    *out = c;
    return ONILDA_OK;
}

int onilda_corpus_from_csr(const int64_t* offsets, const int32_t* words, const int32_t* counts,
                           int64_t num_docs, int32_t num_terms, onilda_corpus** out)
{
    if (!offsets || !out || num_docs < 0) return onilda_fail(ONILDA_EINVAL, "corpus_from_csr: bad argument");
    int64_t nnz = offsets[num_docs];
    if (offsets[0] != 0 || nnz < 0 || (nnz > 0 && (!words || !counts)))
        return onilda_fail(ONILDA_EINVAL, "corpus_from_csr: offsets must start at 0 and words/counts be non-null");
    onilda_corpus* c = (onilda_corpus*)calloc(1, sizeof *c);
    if (!c) return onilda_fail(ONILDA_ENOMEM, "corpus_from_csr: out of host memory");
    c->num_docs = num_docs;
    c->nnz = nnz;
    c->offsets = (int64_t*)malloc(sizeof(int64_t) * (size_t)(num_docs + 1));
    c->totals = (int32_t*)malloc(sizeof(int32_t) * (size_t)(num_docs > 0 ? num_docs : 1));
    c->words = (int32_t*)malloc(sizeof(int32_t) * (size_t)(nnz > 0 ? nnz : 1));
    c->counts = (int32_t*)malloc(sizeof(int32_t) * (size_t)(nnz > 0 ? nnz : 1));
    if (!c->offsets || !c->totals || !c->words || !c->counts) {
        onilda_corpus_free(c);
        return onilda_fail(ONILDA_ENOMEM, "corpus_from_csr: out of host memory");
    }
    memcpy(c->offsets, offsets, sizeof(int64_t) * (size_t)(num_docs + 1));
    if (nnz > 0) {
        memcpy(c->words, words, sizeof(int32_t) * (size_t)nnz);
        memcpy(c->counts, counts, sizeof(int32_t) * (size_t)nnz);
    }
    int32_t nterms = 0;
    for (int64_t d = 0; d < num_docs; d++) {
        if (offsets[d + 1] < offsets[d]) {
            onilda_corpus_free(c);
            return onilda_fail(ONILDA_EINVAL, "corpus_from_csr: offsets decrease at document %lld", (long long)d);
        }
        int total = 0;
        for (int64_t e = offsets[d]; e < offsets[d + 1]; e++) {
            if (words[e] < 0) {
                onilda_corpus_free(c);
                return onilda_fail(ONILDA_EINVAL, "corpus_from_csr: negative word id at entry %lld", (long long)e);
            }
            total += counts[e];
            if (words[e] >= nterms) nterms = words[e] + 1;
        }
        c->totals[d] = total;
    }
    c->num_terms = num_terms > nterms ? num_terms : nterms;
    corpus_finish(c);
    *out = c;
    return ONILDA_OK;
}

int onilda_corpus_dims(const onilda_corpus* c, int64_t* num_docs, int64_t* nnz, int32_t* num_terms,
                       int64_t* tokens, int32_t* max_len)
{
    if (!c) return onilda_fail(ONILDA_EINVAL, "corpus_dims: null corpus");
    if (num_docs) *num_docs = c->num_docs;
    if (nnz) *nnz = c->nnz;
    if (num_terms) *num_terms = c->num_terms;
    if (tokens) *tokens = c->tokens;
    if (max_len) *max_len = c->max_len;
    return ONILDA_OK;
}

int onilda_corpus_csr(const onilda_corpus* c, const int64_t** offsets, const int32_t** words,
                      const int32_t** counts, const int32_t** totals)
{
    if (!c) return onilda_fail(ONILDA_EINVAL, "corpus_csr: null corpus");
    if (offsets) *offsets = c->offsets;
    if (words) *words = c->words;
    if (counts) *counts = c->counts;
    if (totals) *totals = c->totals;
    return ONILDA_OK;
}

static char* put_int(char* p, long long v)
{
    char tmp[24];
    int n = 0;
    unsigned long long u = v < 0 ? (unsigned long long)(-v) : (unsigned long long)v;
    if (v < 0) *p++ = '-';
    do { tmp[n++] = (char)('0' + u % 10); u /= 10; } while (u);
    while (n) *p++ = tmp[--n];
    return p;
}

int onilda_corpus_save_dat(const onilda_corpus* c, int64_t doc_begin, int64_t doc_end, const char* path)
{
    if (!c || !path || doc_begin < 0 || doc_end > c->num_docs || doc_begin > doc_end)
        return onilda_fail(ONILDA_EINVAL, "corpus_save_dat: bad argument");
    FILE* f = fopen(path, "w");
    if (!f) return onilda_fail(ONILDA_EIO, "cannot write %s", path);
    size_t cap = 1 << 20;
    char* buf = (char*)malloc(cap + 64);
    if (!buf) { fclose(f); return onilda_fail(ONILDA_ENOMEM, "corpus_save_dat: out of memory"); }
    char* p = buf;
    int rc = ONILDA_OK;
    for (int64_t d = doc_begin; d < doc_end && rc == ONILDA_OK; d++) {
        p = put_int(p, c->offsets[d + 1] - c->offsets[d]);
        for (int64_t e = c->offsets[d]; e < c->offsets[d + 1]; e++) {
            *p++ = ' ';
            p = put_int(p, c->words[e]);
            *p++ = ':';
            p = put_int(p, c->counts[e]);
            if ((size_t)(p - buf) > cap) {
                if (fwrite(buf, 1, (size_t)(p - buf), f) != (size_t)(p - buf)) { rc = onilda_fail(ONILDA_EIO, "write failed on %s", path); break; }
                p = buf;
            }
        }
        *p++ = '\n';
        if ((size_t)(p - buf) > cap) {
            if (fwrite(buf, 1, (size_t)(p - buf), f) != (size_t)(p - buf)) rc = onilda_fail(ONILDA_EIO, "write failed on %s", path);
            p = buf;
        }
    }
    if (rc == ONILDA_OK && p > buf && fwrite(buf, 1, (size_t)(p - buf), f) != (size_t)(p - buf)) rc = onilda_fail(ONILDA_EIO, "write failed on %s", path);
    free(buf);
    if (fclose(f) != 0 && rc == ONILDA_OK) rc = onilda_fail(ONILDA_EIO, "close failed on %s", path);
    return rc;
}

void onilda_corpus_free(onilda_corpus* c)
{
    if (!c) return;
    if (c->pinned) onilda_corpus_unpin(c);
    free(c->offsets); free(c->words); free(c->counts); free(c->totals);
    free(c);
}

/* --------------------------------------------------------------------- model */

int onilda_model_new(int num_topics, int num_terms, onilda_model** out)
{
    if (!out || num_topics <= 0 || num_terms < 0) return onilda_fail(ONILDA_EINVAL, "model_new: bad dimensions");
    onilda_model* m = (onilda_model*)calloc(1, sizeof *m);
    if (!m) return onilda_fail(ONILDA_ENOMEM, "model_new: out of host memory");
    m->num_topics = num_topics;
    m->num_terms = num_terms;
    m->alpha = 1.0;                       /* lda-model.c:159 */
    size_t n = (size_t)num_topics * (size_t)num_terms;
    m->log_beta = (double*)calloc(n ? n : 1, sizeof(double));
    if (!m->log_beta) { free(m); return onilda_fail(ONILDA_ENOMEM, "model_new: %d x %d table does not fit", num_topics, num_terms); }
    *out = m;
    return ONILDA_OK;
}

int onilda_model_dims(const onilda_model* m, int* num_topics, int* num_terms)
{
    if (!m) return onilda_fail(ONILDA_EINVAL, "model_dims: null model");
    if (num_topics) *num_topics = m->num_topics;
    if (num_terms) *num_terms = m->num_terms;
    return ONILDA_OK;
}

double onilda_model_alpha(const onilda_model* m) { return m ? m->alpha : NAN; }

int onilda_model_set_alpha(onilda_model* m, double alpha)
{
    if (!m) return onilda_fail(ONILDA_EINVAL, "model_set_alpha: null model");
    m->alpha = alpha;
    return ONILDA_OK;
}

double* onilda_model_log_beta(onilda_model* m) { return m ? m->log_beta : NULL; }

void onilda_model_free(onilda_model* m)
{
    if (!m) return;
    free(m->log_beta);
    free(m);
}

/* printf("%5.10f") without printf.  The value is split into integer and
 * fractional part (both exact in binary64); frac * 1e10 carries an error far
 * below 1e-3, so unless it lands within 1e-3 of a rounding tie the ten digits
 * are decided; the rare undecided case goes through snprintf. */
int onilda_format_fixed10(double v, char* out)
{
    double a = fabs(v);
    if (!(a < 9.0e5)) return sprintf(out, "%5.10f", v);   /* also NaN / inf */
    double ip = floor(a);
    double scaled = (a - ip) * 1e10;
    double r = floor(scaled);
    double rem = scaled - r;
    if (fabs(rem - 0.5) < 1e-3) return sprintf(out, "%5.10f", v);
    if (rem > 0.5) r += 1.0;
    uint64_t frac = (uint64_t)r;
    uint32_t whole = (uint32_t)ip;
    if (frac >= 10000000000ull) { frac -= 10000000000ull; whole++; }
    char* p = out;
    if (signbit(v)) *p++ = '-';
    char tmp[16];
    int n = 0;
    do { tmp[n++] = (char)('0' + whole % 10); whole /= 10; } while (whole);
    while (n) *p++ = tmp[--n];
    *p++ = '.';
    for (int i = 9; i >= 0; i--) { p[i] = (char)('0' + frac % 10); frac /= 10; }
    p += 10;
    return (int)(p - out);
}

/* save_lda_model (lda-model.c:193-217), byte for byte; written to a temporary
 * name and renamed so a failed run never leaves a partial final.* behind. */
int onilda_model_save(const onilda_model* m, const char* model_root)
{
    if (!m || !model_root) return onilda_fail(ONILDA_EINVAL, "model_save: null argument");
    size_t L = strlen(model_root);
    char* name = (char*)malloc(L + 16);
    char* tmpname = (char*)malloc(L + 24);
    if (!name || !tmpname) { free(name); free(tmpname); return onilda_fail(ONILDA_ENOMEM, "model_save: out of memory"); }
    int rc = ONILDA_OK;
    const int V = m->num_terms, K = m->num_topics;
    size_t cap = (size_t)V * 24 + 16;
    char* line = (char*)malloc(cap);
    if (!line) { rc = onilda_fail(ONILDA_ENOMEM, "model_save: out of memory"); goto out; }

    sprintf(name, "%s.beta", model_root);
    sprintf(tmpname, "%s.beta.tmp", model_root);
    FILE* f = fopen(tmpname, "w");
    if (!f) { rc = onilda_fail(ONILDA_EIO, "cannot write %s", tmpname); goto out; }
    for (int k = 0; k < K; k++) {
        char* p = line;
        const double* row = m->log_beta + (size_t)k * V;
        for (int w = 0; w < V; w++) {
            *p++ = ' ';
            p += onilda_format_fixed10(row[w], p);
        }
        *p++ = '\n';
        if (fwrite(line, 1, (size_t)(p - line), f) != (size_t)(p - line)) { rc = onilda_fail(ONILDA_EIO, "write failed on %s", tmpname); break; }
    }
    if (fclose(f) != 0 && rc == ONILDA_OK) rc = onilda_fail(ONILDA_EIO, "close failed on %s", tmpname);
    if (rc != ONILDA_OK) { remove(tmpname); goto out; }
    if (rename(tmpname, name) != 0) { rc = onilda_fail(ONILDA_EIO, "cannot rename %s", tmpname); goto out; }

    sprintf(name, "%s.other", model_root);
    sprintf(tmpname, "%s.other.tmp", model_root);
    f = fopen(tmpname, "w");
    if (!f) { rc = onilda_fail(ONILDA_EIO, "cannot write %s", tmpname); goto out; }
    fprintf(f, "num_topics %d\n", K);
    fprintf(f, "num_terms %d\n", V);
    fprintf(f, "alpha %5.10f\n", m->alpha);
    if (fclose(f) != 0) { rc = onilda_fail(ONILDA_EIO, "close failed on %s", tmpname); remove(tmpname); goto out; }
    if (rename(tmpname, name) != 0) rc = onilda_fail(ONILDA_EIO, "cannot rename %s", tmpname);
out:
    free(line); free(name); free(tmpname);
    return rc;
}

/* Parses one float token the way fscanf("%f") + widening does: strtof. */
static int next_float(cursor* c, float* out)
{
    while (c->p < c->end && isspace(*c->p)) c->p++;
    if (c->p >= c->end) return 0;
    char* stop = NULL;
    *out = strtof((const char*)c->p, &stop);
    if ((const unsigned char*)stop == c->p) return 0;
    c->p = (const unsigned char*)stop;
    return 1;
}

/* load_lda_model (lda-model.c:220-251): alpha and beta go through FLOAT. */
int onilda_model_load(const char* model_root, onilda_model** out)
{
    if (!model_root || !out) return onilda_fail(ONILDA_EINVAL, "model_load: null argument");
    size_t L = strlen(model_root);
    char* name = (char*)malloc(L + 16);
    if (!name) return onilda_fail(ONILDA_ENOMEM, "model_load: out of memory");
    int K = 0, V = 0, rc = ONILDA_OK;
    float alpha = 0;
    onilda_model* m = NULL;
    unsigned char* text = NULL;

    sprintf(name, "%s.other", model_root);
    FILE* f = fopen(name, "r");
    if (!f) { rc = onilda_fail(ONILDA_EIO, "cannot open %s", name); goto out; }
    int got = 0;
    got += fscanf(f, "num_topics %d\n", &K) == 1;
    got += fscanf(f, "num_terms %d\n", &V) == 1;
    got += fscanf(f, "alpha %f\n", &alpha) == 1;
    fclose(f);
    if (got != 3 || K <= 0 || V < 0) { rc = onilda_fail(ONILDA_EINVAL, "%s: not an lda-c .other file", name); goto out; }
    rc = onilda_model_new(K, V, &m);
    if (rc != ONILDA_OK) goto out;
    m->alpha = alpha;

    sprintf(name, "%s.beta", model_root);
    f = fopen(name, "rb");
    if (!f) { rc = onilda_fail(ONILDA_EIO, "cannot open %s", name); goto out; }
    fseek(f, 0, SEEK_END);
    long size = ftell(f);
    fseek(f, 0, SEEK_SET);
    text = (unsigned char*)malloc((size_t)size + 1);
    if (!text) { fclose(f); rc = onilda_fail(ONILDA_ENOMEM, "%s does not fit in memory", name); goto out; }
    if (size > 0 && fread(text, 1, (size_t)size, f) != (size_t)size) { fclose(f); rc = onilda_fail(ONILDA_EIO, "short read on %s", name); goto out; }
    fclose(f);
    text[size] = 0;
    cursor cur = { text, text + size };
    for (size_t i = 0, n = (size_t)K * (size_t)V; i < n; i++) {
        float x;
        if (!next_float(&cur, &x)) { rc = onilda_fail(ONILDA_EINVAL, "%s: fewer than %d x %d values", name, K, V); goto out; }
        m->log_beta[i] = x;
    }
out:
    free(text); free(name);
    if (rc != ONILDA_OK) { onilda_model_free(m); return rc; }
    *out = m;
    return ONILDA_OK;
}

/* save_gamma (lda-estimate.c:88-104) */
int onilda_gamma_save(const char* path, const double* gamma, int64_t num_docs, int num_topics)
{
    if (!path || (!gamma && num_docs > 0) || num_topics <= 0) return onilda_fail(ONILDA_EINVAL, "gamma_save: bad argument");
    size_t L = strlen(path);
    char* tmpname = (char*)malloc(L + 8);
    char* line = (char*)malloc((size_t)num_topics * 24 + 16);
    if (!tmpname || !line) { free(tmpname); free(line); return onilda_fail(ONILDA_ENOMEM, "gamma_save: out of memory"); }
    sprintf(tmpname, "%s.tmp", path);
    int rc = ONILDA_OK;
    FILE* f = fopen(tmpname, "w");
    if (!f) { rc = onilda_fail(ONILDA_EIO, "cannot write %s", tmpname); goto out; }
    for (int64_t d = 0; d < num_docs; d++) {
        const double* row = gamma + d * num_topics;
        char* p = line;
        for (int k = 0; k < num_topics; k++) {
            if (k) *p++ = ' ';
            p += onilda_format_fixed10(row[k], p);
        }
        *p++ = '\n';
        if (fwrite(line, 1, (size_t)(p - line), f) != (size_t)(p - line)) { rc = onilda_fail(ONILDA_EIO, "write failed on %s", tmpname); break; }
    }
    if (fclose(f) != 0 && rc == ONILDA_OK) rc = onilda_fail(ONILDA_EIO, "close failed on %s", tmpname);
    if (rc != ONILDA_OK) { remove(tmpname); goto out; }
    if (rename(tmpname, path) != 0) rc = onilda_fail(ONILDA_EIO, "cannot rename %s", tmpname);
out:
    free(tmpname); free(line);
    return rc;
}
