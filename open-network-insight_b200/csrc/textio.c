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
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

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

/* Host threads for parsing and printing the text files (ONILDA_IO_THREADS overrides; 1 = the
 * sequential code only).  The files are the reference's interface, not its hot path, but at
 * cfg 2 (1 GB of model.dat, 280 MB of final.gamma) one thread of text work would outlast EM. */
static int io_threads(void)
{
    const char* e = getenv("ONILDA_IO_THREADS");
    long n = e && *e ? atol(e) : sysconf(_SC_NPROCESSORS_ONLN);
    if (n < 1) n = 1;
    if (n > 64) n = 64;
    return (int)n;
}

/* The file into memory, by several threads where it is large (a copy out of the page cache runs at a few GB/s per
 * thread, and model.dat is 850 MB at cfg 2). */
typedef struct { int fd; unsigned char* dst; size_t begin, end; int failed; } read_job;

static void* read_worker(void* arg)
{
    read_job* j = (read_job*)arg;
    size_t pos = j->begin;
    while (pos < j->end) {
        ssize_t got = pread(j->fd, j->dst + pos, j->end - pos, (off_t)pos);
        if (got <= 0) { j->failed = 1; return NULL; }
        pos += (size_t)got;
    }
    return NULL;
}

static int read_whole(FILE* f, unsigned char* dst, size_t size)
{
    int T = io_threads();
    if (size < ((size_t)8 << 20)) T = 1;
    if (T > 16) T = 16;
    read_job job[16];
    pthread_t tid[16];
    const int fd = fileno(f);
    for (int t = 0; t < T; t++) {
        job[t].fd = fd; job[t].dst = dst; job[t].failed = 0;
        job[t].begin = size / (size_t)T * (size_t)t;
        job[t].end = t == T - 1 ? size : size / (size_t)T * (size_t)(t + 1);
    }
    for (int t = 1; t < T; t++) pthread_create(&tid[t], NULL, read_worker, &job[t]);
    read_worker(&job[0]);
    for (int t = 1; t < T; t++) pthread_join(tid[t], NULL);
    for (int t = 0; t < T; t++) if (job[t].failed) return -1;
    return 0;
}

/* ---- parallel reader for the canonical layout
 * OniLDACWrapper.createModel (OniLDACWrapper.scala:146-168) writes one document per line:
 * "<n> <w>:<c> ... <w>:<c>\n", plain decimal, single spaces.  On such a file every line can be
 * parsed on its own, and what a line yields is by construction what read_data's token stream
 * yields (numbers of at most 9 digits never meet "%10d"'s width limit or int overflow).  Any
 * deviation - a sign, a longer number, a pair missing or in excess, another separator - makes
 * the chunk give up and the whole file goes through the sequential scanf emulation below, which
 * is the definition of the format. */
typedef struct {
    const unsigned char* begin;
    const unsigned char* end;
    int64_t docs, nnz;          /* pass 1: lines with content, ':' characters */
    int64_t doc0, nnz0;         /* pass 2: where this chunk's output starts */
    onilda_corpus* c;
    int nterms;
    int ok;
} dat_chunk;

static void* dat_count(void* arg)
{
    dat_chunk* k = (dat_chunk*)arg;
    int64_t docs = 0, nnz = 0;
    int content = 0;
    for (const unsigned char* p = k->begin; p < k->end; p++) {
        unsigned char ch = *p;
        if (ch == '\n') { docs += content; content = 0; }
        else if (ch == ':') { nnz++; content = 1; }
        else if (ch != ' ' && ch != '\r' && ch != '\t') content = 1;
    }
    k->docs = docs + content;
    k->nnz = nnz;
    return NULL;
}

/* decimal digits, 1..9 of them; returns the value or -1 */
static inline int64_t canon_uint(const unsigned char** pp, const unsigned char* end)
{
    const unsigned char* p = *pp;
    int64_t v = 0;
    int nd = 0;
    while (p < end && *p >= '0' && *p <= '9' && nd < 10) { v = v * 10 + (*p - '0'); p++; nd++; }
    if (nd == 0 || nd > 9) return -1;
    *pp = p;
    return v;
}

static void* dat_parse(void* arg)
{
    dat_chunk* k = (dat_chunk*)arg;
    onilda_corpus* c = k->c;
    const unsigned char* p = k->begin;
    const unsigned char* end = k->end;
    int64_t d = k->doc0, e = k->nnz0;
    const int64_t d_end = k->doc0 + k->docs, e_end = k->nnz0 + k->nnz;
    int nterms = 0;
    k->ok = 0;
    while (p < end) {
        if (*p == '\n') { p++; continue; }                      /* empty line: white space to read_data */
        int64_t len = canon_uint(&p, end);
        if (len < 0 || d >= d_end || e + len > e_end) return NULL;
        int total = 0;
        for (int64_t n = 0; n < len; n++) {
            if (p >= end || *p != ' ') return NULL;
            p++;
            int64_t w = canon_uint(&p, end);
            if (w < 0 || p >= end || *p != ':') return NULL;
            p++;
            int64_t cnt = canon_uint(&p, end);
            if (cnt < 0) return NULL;
            c->words[e] = (int32_t)w;
            c->counts[e] = (int32_t)cnt;
            e++;
            total = (int)((unsigned int)total + (unsigned int)cnt);
            if (w >= nterms) nterms = (int)w + 1;
        }
        if (p < end && *p == '\r') p++;
        if (p < end && *p != '\n') return NULL;
        p++;
        c->totals[d] = total;
        c->offsets[d + 1] = len;                                  /* lengths now, prefix sums after the join */
        d++;
    }
    if (d != d_end || e != e_end) return NULL;
    k->nterms = nterms;
    k->ok = 1;
    return NULL;
}

/* Returns 1 with *out set, 0 when the file is not in the canonical layout (caller falls back),
 * or a negative error code. */
static int corpus_from_dat_parallel(const unsigned char* text, size_t size, int threads, onilda_corpus** out)
{
    dat_chunk ch[64];
    pthread_t tid[64];
    int T = threads;
    if ((size_t)T > size / (1u << 20) + 1) T = (int)(size / (1u << 20) + 1);
    const unsigned char* pos = text;
    const unsigned char* end = text + size;
    int used = 0;
    for (int t = 0; t < T && pos < end; t++) {
        const unsigned char* stop = t == T - 1 ? end : text + (size / (size_t)T) * (size_t)(t + 1);
        if (stop < pos) stop = pos;
        while (stop < end && stop[-1] != '\n' && stop > pos) stop++;      /* to the end of the line */
        if (stop == pos && pos < end) { const unsigned char* nl = memchr(pos, '\n', (size_t)(end - pos)); stop = nl ? nl + 1 : end; }
        ch[used].begin = pos;
        ch[used].end = stop;
        pos = stop;
        used++;
    }
    if (pos < end) { ch[used - 1].end = end; }
    T = used;
    for (int t = 1; t < T; t++) pthread_create(&tid[t], NULL, dat_count, &ch[t]);
    dat_count(&ch[0]);
    for (int t = 1; t < T; t++) pthread_join(tid[t], NULL);
    int64_t docs = 0, nnz = 0;
    for (int t = 0; t < T; t++) { ch[t].doc0 = docs; ch[t].nnz0 = nnz; docs += ch[t].docs; nnz += ch[t].nnz; }
    onilda_corpus* c = (onilda_corpus*)calloc(1, sizeof *c);
    if (!c) return ONILDA_ENOMEM;
    c->offsets = (int64_t*)malloc(sizeof(int64_t) * (size_t)(docs + 1));
    c->totals = (int32_t*)malloc(sizeof(int32_t) * (size_t)(docs > 0 ? docs : 1));
    c->words = (int32_t*)malloc(sizeof(int32_t) * (size_t)(nnz > 0 ? nnz : 1));
    c->counts = (int32_t*)malloc(sizeof(int32_t) * (size_t)(nnz > 0 ? nnz : 1));
    if (!c->offsets || !c->totals || !c->words || !c->counts) { onilda_corpus_free(c); return ONILDA_ENOMEM; }
    for (int t = 0; t < T; t++) ch[t].c = c;
    for (int t = 1; t < T; t++) pthread_create(&tid[t], NULL, dat_parse, &ch[t]);
    dat_parse(&ch[0]);
    for (int t = 1; t < T; t++) pthread_join(tid[t], NULL);
    int nterms = 0;
    for (int t = 0; t < T; t++) {
        if (!ch[t].ok) { onilda_corpus_free(c); return 0; }
        if (ch[t].nterms > nterms) nterms = ch[t].nterms;
    }
    c->offsets[0] = 0;
    for (int64_t d = 0; d < docs; d++) c->offsets[d + 1] += c->offsets[d];
    c->num_docs = docs;
    c->nnz = nnz;
    c->num_terms = nterms;
    corpus_finish(c);
    *out = c;
    return 1;
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
    if (size > 0 && read_whole(f, text, (size_t)size) != 0) {
        fclose(f); free(text);
        return onilda_fail(ONILDA_EIO, "short read on %s", path);
    }
    fclose(f);

    {
        const int threads = io_threads();
        if (threads > 1 && size > 0) {
            int pr = corpus_from_dat_parallel(text, (size_t)size, threads, out);
            if (pr != 0) {
                free(text);
                if (pr < 0) return onilda_fail(pr, "corpus %s: out of host memory", path);
                return ONILDA_OK;
            }
        }
    }

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

/* ---- parallel "%5.10f" matrix printer (save_lda_model's .beta, save_gamma)
 * The flat element range is split over the host threads; each formats its part into its own
 * buffer, the buffers are written in order.  lead != 0: a space before every number (.beta),
 * else a space between numbers (.gamma). */
typedef struct {
    const double* v;
    int64_t begin, end;
    int cols, lead;
    char* buf;
    size_t len;
    int failed;
} fmt_job;

static void* fmt_worker(void* arg)
{
    fmt_job* j = (fmt_job*)arg;
    size_t cap = (size_t)(j->end - j->begin) * 20 + 1024;
    char* buf = (char*)malloc(cap);
    char* p = buf;
    j->failed = buf == NULL;
    if (!buf) return NULL;
    int col = (int)(j->begin % j->cols);
    for (int64_t i = j->begin; i < j->end; i++) {
        if ((size_t)(p - buf) + 400 > cap) {                     /* "%5.10f" of a huge value is long */
            size_t used = (size_t)(p - buf);
            cap = cap * 2 + 1024;
            char* nb = (char*)realloc(buf, cap);
            if (!nb) { free(buf); j->failed = 1; return NULL; }
            buf = nb;
            p = buf + used;
        }
        if (j->lead || col) *p++ = ' ';
        p += onilda_format_fixed10(j->v[i], p);
        if (++col == j->cols) { *p++ = '\n'; col = 0; }
    }
    j->buf = buf;
    j->len = (size_t)(p - buf);
    return NULL;
}

static int write_matrix(FILE* f, const double* v, int64_t rows, int cols, int lead, const char* name)
{
    const int64_t n = rows * (int64_t)cols;
    if (n == 0) return ONILDA_OK;
    int T = io_threads();
    if (n < 200000) T = 1;
    fmt_job job[64];
    pthread_t tid[64];
    for (int t = 0; t < T; t++) {
        job[t].v = v; job[t].cols = cols; job[t].lead = lead; job[t].buf = NULL; job[t].len = 0; job[t].failed = 0;
        job[t].begin = n * t / T;
        job[t].end = n * (t + 1) / T;
    }
    for (int t = 1; t < T; t++) pthread_create(&tid[t], NULL, fmt_worker, &job[t]);
    fmt_worker(&job[0]);
    for (int t = 1; t < T; t++) pthread_join(tid[t], NULL);
    int rc = ONILDA_OK;
    for (int t = 0; t < T && rc == ONILDA_OK; t++) {
        if (job[t].failed) rc = onilda_fail(ONILDA_ENOMEM, "out of memory while formatting %s", name);
        else if (job[t].len && fwrite(job[t].buf, 1, job[t].len, f) != job[t].len) rc = onilda_fail(ONILDA_EIO, "write failed on %s", name);
    }
    for (int t = 0; t < T; t++) free(job[t].buf);
    return rc;
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
    rc = write_matrix(f, m->log_beta, K, V, 1, tmpname);
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
    rc = write_matrix(f, gamma, num_docs, num_topics, 0, tmpname);
    if (fclose(f) != 0 && rc == ONILDA_OK) rc = onilda_fail(ONILDA_EIO, "close failed on %s", tmpname);
    if (rc != ONILDA_OK) { remove(tmpname); goto out; }
    if (rename(tmpname, path) != 0) rc = onilda_fail(ONILDA_EIO, "cannot rename %s", tmpname);
out:
    free(tmpname); free(line);
    return rc;
}


/*
 * The results file.  The reference ends each analysis with
 *     rows.map(_.mkString(outputDelimiter)).saveAsTextFile(path)
 * (netflow/FlowPostLDA.scala:80, dns/DNSPostLDA.scala:69, proxy/ProxySuspiciousConnectsAnalysis.scala:51) and ml_ops.sh
 * merges the parts into <type>_results.csv, which oni-oa reads (oa/flow/flow_oa.py:101).  A row is the event's own
 * columns, then its word column(s), then the score column(s): SourceScore, DestinationScore for flow (the minimum is
 * dropped, :80), Score for dns and proxy.  mkString prints a Double with Java's Double.toString.
 */

/* Java's Double.toString layout with the shortest digits that read back as the same double: plain decimal for
 * 1e-3 <= |v| < 1e7 (at least one digit after the point), otherwise d.dddE[-]n.  JDK 19+ prints exactly this; older
 * JDKs (Spark 1.6 ran on 7/8) print one digit more for a small share of values - the same double when parsed, which
 * is how oni-oa reads the column.  Returns the length (no NUL); out needs 32 bytes. */
int onilda_format_java_double(double v, char* out)
{
    if (isnan(v)) { memcpy(out, "NaN", 3); return 3; }
    int n = 0;
    if (signbit(v)) { out[n++] = '-'; v = -v; }
    if (isinf(v)) { memcpy(out + n, "Infinity", 8); return n + 8; }
    if (v == 0.0) { memcpy(out + n, "0.0", 3); return n + 3; }
    char buf[40];
    int prec;
    for (prec = 2; prec <= 17; prec++) {            /* shortest digit string (two digits at least, as Java) that round-trips */
        snprintf(buf, sizeof buf, "%.*e", prec - 1, v);
        if (strtod(buf, NULL) == v) break;
    }
    /* buf = d[.ddd]e[+-]xx */
    char digits[20];
    int nd = 0;
    const char* p = buf;
    for (; *p && *p != 'e'; p++) if (*p >= '0' && *p <= '9') digits[nd++] = *p;
    int e10 = atoi(p + 1);
    while (nd > 1 && digits[nd - 1] == '0') nd--;
    if (v >= 1e-3 && v < 1e7) {
        if (e10 >= 0) {
            for (int i = 0; i <= e10; i++) out[n++] = i < nd ? digits[i] : '0';
            out[n++] = '.';
            if (nd > e10 + 1) for (int i = e10 + 1; i < nd; i++) out[n++] = digits[i];
            else out[n++] = '0';
        } else {
            out[n++] = '0';
            out[n++] = '.';
            for (int i = 0; i < -e10 - 1; i++) out[n++] = '0';
            for (int i = 0; i < nd; i++) out[n++] = digits[i];
        }
    } else {
        out[n++] = digits[0];
        out[n++] = '.';
        if (nd > 1) for (int i = 1; i < nd; i++) out[n++] = digits[i];
        else out[n++] = '0';
        out[n++] = 'E';
        n += sprintf(out + n, "%d", e10);
    }
    return n;
}

/* rows[i] + delimiter + score_a[i] (+ delimiter + score_b[i] when score_b is given), one line per kept row, in the
 * order given (the order of onilda_flow_results / onilda_results_rows).  Written to a temporary file and renamed. */
int onilda_results_write(const char* path, const char* const* rows, int64_t m, const double* score_a, const double* score_b,
                         const char* delimiter)
{
    if (!path || m < 0 || (m > 0 && (!rows || !score_a)) || !delimiter)
        return onilda_fail(ONILDA_EINVAL, "results_write: bad argument");
    size_t L = strlen(path);
    char* tmpname = (char*)malloc(L + 8);
    if (!tmpname) return onilda_fail(ONILDA_ENOMEM, "results_write: out of memory");
    sprintf(tmpname, "%s.tmp", path);
    int rc = ONILDA_OK;
    FILE* f = fopen(tmpname, "w");
    if (!f) { rc = onilda_fail(ONILDA_EIO, "cannot write %s", tmpname); free(tmpname); return rc; }
    char num[40];
    for (int64_t i = 0; i < m && rc == ONILDA_OK; i++) {
        if (!rows[i]) { rc = onilda_fail(ONILDA_EINVAL, "results_write: row %lld is null", (long long)i); break; }
        int k = onilda_format_java_double(score_a[i], num);
        if (fputs(rows[i], f) < 0 || fputs(delimiter, f) < 0 || fwrite(num, 1, (size_t)k, f) != (size_t)k) rc = onilda_fail(ONILDA_EIO, "write failed on %s", tmpname);
        if (rc == ONILDA_OK && score_b) {
            k = onilda_format_java_double(score_b[i], num);
            if (fputs(delimiter, f) < 0 || fwrite(num, 1, (size_t)k, f) != (size_t)k) rc = onilda_fail(ONILDA_EIO, "write failed on %s", tmpname);
        }
        if (rc == ONILDA_OK && fputc('\n', f) == EOF) rc = onilda_fail(ONILDA_EIO, "write failed on %s", tmpname);
    }
    if (fclose(f) != 0 && rc == ONILDA_OK) rc = onilda_fail(ONILDA_EIO, "close failed on %s", tmpname);
    if (rc != ONILDA_OK) remove(tmpname);
    else if (rename(tmpname, path) != 0) rc = onilda_fail(ONILDA_EIO, "cannot rename %s", tmpname);
    free(tmpname);
    return rc;
}
