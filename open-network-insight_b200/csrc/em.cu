// EM handle: run_em's state for one shard of documents on one device
// (reference: oni-ml/oni-lda-c/lda-estimate.c:112-321) and `infer` (:358-388).
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <vector>


#include "device_math.cuh"
#include "kernels.cuh"
#include "nccl_dyn.h"

using namespace onilda;

struct onilda_em {
    const onilda_corpus* corpus;   // borrowed (for onilda_em_reload)
    int device;
    int sm_count;
    cudaStream_t stream;
    cudaStream_t long_stream;   // long documents run beside the bulk launch
    cudaEvent_t ev0, ev1, ev2, ev3, ev_fork, ev_forked, ev_join;
    cudaEvent_t ev_long_end, ev_bulk_end;   // ONILDA_TIMING=1: when each of the two concurrent launches ended
    int timing;
    int64_t num_long;           // documents of at least long_len entries: prefix of d_order
    int64_t long_max_total;     // most tokens in one of them (decides whether the table-only long kernel may run)
    int64_t bulk_max_total;     // most tokens in a document of the bulk launch (decides which variant of the kernel runs)
    int long_len, long_group, long_smem, long_threads, long_per_sm, bulk_group, bulk_per_sm, bulk_threads, pdl, carveout_pct;
    // shard
    int64_t doc_begin, doc_end, num_docs, nnz;
    int K, V;
    TopicLayout lay;
    onilda_settings settings;
    int var_max_iter;           // mutated as at lda-estimate.c:261
    double alpha;
    // EM loop state (lda-estimate.c:172-186)
    int iteration;
    double likelihood_old;
    double converged;
    // device buffers
    int64_t* d_offsets;
    int32_t *d_words, *d_counts, *d_totals, *d_order;
    int32_t *d_order_dyn, *d_win_off;   // the order the bulk launch uses (re-bundled by the last sweep counts) and its windows
    int num_windows, rebundle, sweeps_valid;
    double *d_log_beta_wk, *d_beta_scaled, *d_lbmax;
    double* d_ss;               // [(V + kTablePadRows)*kpad statistics | SS_TAIL scalars | kpad class_total]
    double* d_partial;
    double* d_phi;              // [nnz][kpad]
    double *d_gamma, *d_doc_lik, *d_doc_ass;
    int32_t* d_doc_sweeps;
    unsigned long long* d_queue;   // [2]: bulk, long
    int* d_error;               // [0] NaN likelihood seen; [1] words redone in log space in the last E-step
    double* d_kv;               // [K][V] staging for import / export
    double* d_math_tab;         // psi rows, then the 2^(i/64) table (device_math.cuh)
    // stats
    double estep_ms, mstep_ms;
    int64_t launches;
    int64_t sweeps;
    double word_steps;          // entries x sweeps of the last E-step, summed over all shards at the M-step
    int estep_done;
    // sufficient statistics across shards (NCCL); nranks == 1: none
    ncclComm_t comm;
    int nranks, rank;
    cudaEvent_t ev_c0, ev_c1, ev_c2, ev_c3;   // around the reduce-scatter and the all-gathers
    double rs_ms, ag_ms;
};

namespace {

struct DeviceGuard {
    int prev;
    bool ok;
    explicit DeviceGuard(int dev) : prev(-1), ok(true)
    {
        if (cudaGetDevice(&prev) != cudaSuccess) { ok = false; return; }
        if (dev != prev && cudaSetDevice(dev) != cudaSuccess) ok = false;
    }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

template <typename T>
int dev_alloc(T** p, size_t n)
{
    *p = nullptr;
    cudaError_t e = cudaMalloc((void**)p, sizeof(T) * (n ? n : 1));
    if (e != cudaSuccess) {
        cudaGetLastError();
        return onilda_fail(ONILDA_ENOMEM, "cudaMalloc of %zu bytes failed: %s", sizeof(T) * n, cudaGetErrorString(e));
    }
    return ONILDA_OK;
}

// documents by decreasing length (stable counting sort): longest first for the bundle queue
void order_by_length(const int64_t* offsets, int64_t D, int32_t max_len, std::vector<int32_t>& order)
{
    std::vector<int64_t> start((size_t)max_len + 2, 0);
    for (int64_t d = 0; d < D; d++) start[(size_t)(max_len - (offsets[d + 1] - offsets[d])) + 1]++;
    for (size_t i = 1; i < start.size(); i++) start[i] += start[i - 1];
    order.resize((size_t)D);
    for (int64_t d = 0; d < D; d++) order[(size_t)start[(size_t)(max_len - (offsets[d + 1] - offsets[d]))]++] = (int32_t)d;
}

// doubles of the padded statistics / model tables; the scalar tail and class_total follow in d_ss
size_t table_doubles(const onilda_em* em) { return ((size_t)em->V + kTablePadRows) * em->lay.kpad; }

int upload_model(onilda_em* em, const double* log_beta_kv)
{
    const size_t kv = (size_t)em->K * em->V;
    if (kv == 0) return ONILDA_OK;
    ONILDA_CUDA(cudaMemcpyAsync(em->d_kv, log_beta_kv, sizeof(double) * kv, cudaMemcpyHostToDevice, em->stream));
    int l = 0;
    int rc = launch_tables_from_log_beta(em->d_kv, em->K, em->V, em->lay, em->d_log_beta_wk, em->d_beta_scaled,
                                         em->d_lbmax, em->stream, &l);
    em->launches += l;
    return rc;
}

}  // namespace

extern "C" int onilda_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

// Creates the CUDA context of the first `ndev` devices (all when ndev <= 0).  Context creation takes several hundred
// milliseconds; a caller with host work to do first (reading model.dat) runs this on another thread meanwhile.
extern "C" int onilda_device_warmup(int ndev)
{
    const int avail = onilda_device_count();
    if (avail == 0) return onilda_fail(ONILDA_ECUDA, "device_warmup: no CUDA device is usable");
    if (ndev <= 0 || ndev > avail) ndev = avail;
    int prev = 0;
    ONILDA_CUDA(cudaGetDevice(&prev));
    for (int d = 0; d < ndev; d++) {
        ONILDA_CUDA(cudaSetDevice(d));
        ONILDA_CUDA(cudaFree(0));
    }
    ONILDA_CUDA(cudaSetDevice(prev));
    return ONILDA_OK;
}

extern "C" int onilda_corpus_pin(onilda_corpus* c)
{
    if (!c) return onilda_fail(ONILDA_EINVAL, "corpus_pin: null corpus");
    if (c->pinned) return ONILDA_OK;
    if (onilda_device_count() == 0) return onilda_fail(ONILDA_ECUDA, "corpus_pin: no CUDA device is usable");
    cudaError_t e = cudaSuccess;
    if (c->nnz > 0) {
        e = cudaHostRegister(c->words, sizeof(int32_t) * (size_t)c->nnz, cudaHostRegisterDefault);
        if (e == cudaSuccess) {
            e = cudaHostRegister(c->counts, sizeof(int32_t) * (size_t)c->nnz, cudaHostRegisterDefault);
            if (e != cudaSuccess) cudaHostUnregister(c->words);
        }
    }
    if (e != cudaSuccess) { cudaGetLastError(); return onilda_fail(ONILDA_ECUDA, "corpus_pin: %s", cudaGetErrorString(e)); }
    c->pinned = 1;
    return ONILDA_OK;
}

extern "C" void onilda_corpus_unpin(onilda_corpus* c)
{
    if (!c || !c->pinned) return;
    if (c->nnz > 0) { cudaHostUnregister(c->words); cudaHostUnregister(c->counts); }
    cudaGetLastError();
    c->pinned = 0;
}

extern "C" void onilda_em_destroy(onilda_em* em)
{
    if (!em) return;
    DeviceGuard g(em->device);
    cudaFree(em->d_offsets); cudaFree(em->d_words); cudaFree(em->d_counts); cudaFree(em->d_totals); cudaFree(em->d_order); cudaFree(em->d_order_dyn); cudaFree(em->d_win_off);
    cudaFree(em->d_log_beta_wk); cudaFree(em->d_beta_scaled); cudaFree(em->d_lbmax); cudaFree(em->d_ss);
    cudaFree(em->d_partial); cudaFree(em->d_phi); cudaFree(em->d_gamma);
    cudaFree(em->d_doc_lik); cudaFree(em->d_doc_ass); cudaFree(em->d_doc_sweeps); cudaFree(em->d_queue);
    cudaFree(em->d_error); cudaFree(em->d_kv); cudaFree(em->d_math_tab);
    if (em->ev0) cudaEventDestroy(em->ev0);
    if (em->ev1) cudaEventDestroy(em->ev1);
    if (em->ev2) cudaEventDestroy(em->ev2);
    if (em->ev3) cudaEventDestroy(em->ev3);
    if (em->ev_fork) cudaEventDestroy(em->ev_fork);
    if (em->ev_join) cudaEventDestroy(em->ev_join);
    if (em->ev_forked) cudaEventDestroy(em->ev_forked);
    if (em->ev_long_end) cudaEventDestroy(em->ev_long_end);
    if (em->ev_bulk_end) cudaEventDestroy(em->ev_bulk_end);
    if (em->comm) { const NcclApi* nccl = nccl_api(); if (nccl) nccl->CommDestroy(em->comm); }
    if (em->ev_c0) cudaEventDestroy(em->ev_c0);
    if (em->ev_c1) cudaEventDestroy(em->ev_c1);
    if (em->ev_c2) cudaEventDestroy(em->ev_c2);
    if (em->ev_c3) cudaEventDestroy(em->ev_c3);
    if (em->long_stream) cudaStreamDestroy(em->long_stream);
    if (em->stream) cudaStreamDestroy(em->stream);
    free(em);
}

extern "C" int onilda_em_create(const onilda_corpus* c, int64_t doc_begin, int64_t doc_end, const onilda_model* start,
                                const onilda_settings* settings, int device, onilda_em** out)
{
    if (!c || !start || !settings || !out) return onilda_fail(ONILDA_EINVAL, "em_create: null argument");
    if (doc_begin < 0 || doc_end > c->num_docs || doc_begin > doc_end)
        return onilda_fail(ONILDA_EINVAL, "em_create: shard [%lld, %lld) outside the corpus of %lld documents",
                           (long long)doc_begin, (long long)doc_end, (long long)c->num_docs);
    if (c->num_terms > start->num_terms)
        return onilda_fail(ONILDA_EINVAL, "em_create: corpus has %d terms, model only %d", c->num_terms, start->num_terms);
    TopicLayout lay;
    if (!choose_layout(start->num_topics, &lay))
        return onilda_fail(ONILDA_EINVAL, "em_create: %d topics unsupported (1..256)", start->num_topics);
    if (onilda_device_count() == 0)
        return onilda_fail(ONILDA_ECUDA, "em_create: no CUDA device is usable and this library has no CPU path");
    if (device < 0) ONILDA_CUDA(cudaGetDevice(&device));
    DeviceGuard guard(device);
    if (!guard.ok) return onilda_fail(ONILDA_ECUDA, "em_create: cannot select device %d", device);

    onilda_em* em = (onilda_em*)calloc(1, sizeof *em);
    if (!em) return onilda_fail(ONILDA_ENOMEM, "em_create: out of host memory");
    em->corpus = c;
    em->device = device;
    em->doc_begin = doc_begin;
    em->doc_end = doc_end;
    em->num_docs = doc_end - doc_begin;
    em->nnz = c->offsets[doc_end] - c->offsets[doc_begin];
    em->K = start->num_topics;
    em->V = start->num_terms;
    em->lay = lay;
    em->settings = *settings;
    em->var_max_iter = settings->var_max_iter;
    em->alpha = start->alpha;
    em->iteration = 0;
    em->likelihood_old = 0;
    em->converged = 1;

    int rc = ONILDA_OK;
#define TRY(x) do { rc = (x); if (rc != ONILDA_OK) { onilda_em_destroy(em); return rc; } } while (0)
#define TRYCUDA(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { rc = onilda_fail(ONILDA_ECUDA, "%s: %s", #x, cudaGetErrorString(e_)); onilda_em_destroy(em); return rc; } } while (0)
    TRYCUDA(cudaDeviceGetAttribute(&em->sm_count, cudaDevAttrMultiProcessorCount, device));
    TRYCUDA(cudaStreamCreateWithFlags(&em->stream, cudaStreamNonBlocking));
    TRYCUDA(cudaEventCreate(&em->ev0));
    TRYCUDA(cudaEventCreate(&em->ev1));
    TRYCUDA(cudaEventCreate(&em->ev2));
    TRYCUDA(cudaEventCreate(&em->ev3));
    {
        // highest priority: when both launches are ready the block scheduler takes the long documents first
        int prio_lo = 0, prio_hi = 0;
        TRYCUDA(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
        TRYCUDA(cudaStreamCreateWithPriority(&em->long_stream, cudaStreamNonBlocking, prio_hi));
    }
    TRYCUDA(cudaEventCreateWithFlags(&em->ev_fork, cudaEventDisableTiming));
    TRYCUDA(cudaEventCreateWithFlags(&em->ev_join, cudaEventDisableTiming));
    TRYCUDA(cudaEventCreateWithFlags(&em->ev_forked, cudaEventDisableTiming));
    TRYCUDA(cudaEventCreate(&em->ev_long_end));
    TRYCUDA(cudaEventCreate(&em->ev_bulk_end));
    em->timing = getenv("ONILDA_TIMING") != nullptr;

    const int64_t D = em->num_docs, N = em->nnz;
    const size_t table = (size_t)em->V * lay.kpad;
    TRY(dev_alloc(&em->d_offsets, (size_t)D + 1));
    TRY(dev_alloc(&em->d_words, (size_t)N));
    TRY(dev_alloc(&em->d_counts, (size_t)N));
    TRY(dev_alloc(&em->d_totals, (size_t)D));
    TRY(dev_alloc(&em->d_order, (size_t)D));
    TRY(dev_alloc(&em->d_order_dyn, (size_t)D));
    // every table has kTablePadRows spare rows (kernels.cuh); beta_scaled and lbmax one more, the idle row
    const size_t table_pad = table + (size_t)kTablePadRows * lay.kpad;
    TRY(dev_alloc(&em->d_log_beta_wk, table_pad));
    TRY(dev_alloc(&em->d_beta_scaled, table_pad + lay.kpad));
    TRY(dev_alloc(&em->d_lbmax, (size_t)em->V + kTablePadRows + 1));
    TRY(dev_alloc(&em->d_ss, table_pad + SS_TAIL + lay.kpad));
    em->nranks = 1;
    TRYCUDA(cudaEventCreate(&em->ev_c0));
    TRYCUDA(cudaEventCreate(&em->ev_c1));
    TRYCUDA(cudaEventCreate(&em->ev_c2));
    TRYCUDA(cudaEventCreate(&em->ev_c3));
    TRYCUDA(cudaMemset(em->d_log_beta_wk, 0, sizeof(double) * table_pad));
    TRYCUDA(cudaMemset(em->d_beta_scaled, 0, sizeof(double) * table_pad));
    TRYCUDA(cudaMemset(em->d_lbmax, 0, sizeof(double) * ((size_t)em->V + kTablePadRows)));
    TRY(dev_alloc(&em->d_partial, (size_t)mstep_partial_doubles(em->V, lay, D)));
    TRY(dev_alloc(&em->d_phi, (size_t)N * lay.kpad));
    TRY(dev_alloc(&em->d_gamma, (size_t)D * em->K));
    TRY(dev_alloc(&em->d_doc_lik, (size_t)D));
    TRY(dev_alloc(&em->d_doc_ass, (size_t)D));
    TRY(dev_alloc(&em->d_doc_sweeps, (size_t)D));
    TRY(dev_alloc(&em->d_queue, 2));
    TRY(dev_alloc(&em->d_error, 2));
    TRY(dev_alloc(&em->d_kv, (size_t)em->K * em->V));
    TRY(dev_alloc(&em->d_math_tab, (size_t)6 * kPsiRows + kExpTabSize));
    {
        std::vector<double> tab((size_t)6 * kPsiRows + kExpTabSize);
        onilda_build_psi_table(tab.data(), kPsiMantBits, kPsiOctaves);
        onilda_build_exp_table(tab.data() + (size_t)6 * kPsiRows, kExpTabSize);
        TRYCUDA(cudaMemcpy(em->d_math_tab, tab.data(), sizeof(double) * tab.size(), cudaMemcpyHostToDevice));
        std::vector<double> idle((size_t)lay.kpad + 1, 0.0);     // beta' = 1 for the real topics, then lbmax = 0
        for (int k = 0; k < em->K; k++) idle[(size_t)k] = 1.0;
        TRYCUDA(cudaMemcpy(em->d_beta_scaled + table_pad, idle.data(), sizeof(double) * lay.kpad, cudaMemcpyHostToDevice));
        TRYCUDA(cudaMemcpy(em->d_lbmax + em->V + kTablePadRows, idle.data() + lay.kpad, sizeof(double), cudaMemcpyHostToDevice));
    }

    // shard-local CSR
    {
        std::vector<int64_t> off((size_t)D + 1);
        const int64_t base = c->offsets[doc_begin];
        int32_t max_len = 0;
        for (int64_t d = 0; d <= D; d++) off[(size_t)d] = c->offsets[doc_begin + d] - base;
        for (int64_t d = 0; d < D; d++) {
            int64_t len = off[(size_t)d + 1] - off[(size_t)d];
            if (len > max_len) max_len = (int32_t)len;
        }
        std::vector<int32_t> order;
        order_by_length(off.data(), D, max_len, order);
        // Long documents (a prefix of the order) get more lanes each and their own launch: the
        // word-by-word chain of one document cannot be split, so the longest document bounds the
        // E-step; fewer topics per lane makes each of its steps shorter.
        const char* env_len = getenv("ONILDA_LONG_LEN");
        const char* env_grp = getenv("ONILDA_LONG_GROUP");
        em->long_len = env_len ? atoi(env_len) : 1024;
        em->long_group = env_grp ? atoi(env_grp) : 32;     // one topic per lane where K <= 32: shortest step
        if (em->long_group < lay.group || (em->long_group & (em->long_group - 1)) || em->long_group > 32) em->long_group = lay.group;
        if ((em->K + em->long_group - 1) / em->long_group > 8) em->long_group = lay.group;
        const char* env_bulk = getenv("ONILDA_BULK_GROUP");
        em->bulk_group = env_bulk ? atoi(env_bulk) : lay.group;
        if (em->bulk_group < lay.group || (em->bulk_group & (em->bulk_group - 1)) || em->bulk_group > 32) em->bulk_group = lay.group;
        const char* env_bps = getenv("ONILDA_BULK_PER_SM");
        em->bulk_per_sm = env_bps ? atoi(env_bps) : 0;
        // The long-document launch: blocks of eight warps, one block per SM, i.e. two long documents per SM
        // sub-partition (measured at cfg 2 in round 1: with three on some sub-partitions their documents
        // finish last, 155-160 -> 145 ms).
        const char* env_smem = getenv("ONILDA_LONG_SMEM");
        const char* env_lt = getenv("ONILDA_LONG_THREADS");
        const char* env_lps = getenv("ONILDA_LONG_PER_SM");
        const char* env_bt = getenv("ONILDA_BULK_THREADS");
        em->long_smem = env_smem ? atoi(env_smem) : 0;
        em->long_threads = env_lt ? atoi(env_lt) : 256;
        em->long_per_sm = env_lps ? atoi(env_lps) : 1;
        em->bulk_threads = env_bt ? atoi(env_bt) : 256;    // eight warps share one copy of the tables
        const char* env_co = getenv("ONILDA_CARVEOUT_PCT");
        // Both launches ask for the SAME shared-memory carve-out, large enough for what is to be resident together (one
        // long-document block, 25 KB, and two bulk blocks, 62 KB each: the 132 KB setting).  Left to the driver, each
        // kernel gets the carve-out that suits it alone; an SM set up for the first one then cannot take blocks of the
        // other (the split is only changed on an idle SM) and the E-step was 110 ms instead of 90.  More than needed
        // costs L1: 106 ms at the 196 / 228 KB settings.
        em->carveout_pct = env_co ? atoi(env_co) : 58;
        const char* env_pdl = getenv("ONILDA_PDL");
        em->pdl = env_pdl ? atoi(env_pdl) : 1;
        em->num_long = 0;
        if (em->long_len > 0 && em->long_group != lay.group)
            while (em->num_long < D && off[(size_t)order[(size_t)em->num_long] + 1] - off[(size_t)order[(size_t)em->num_long]] >= em->long_len) em->num_long++;
        em->long_max_total = 0;
        for (int64_t i = 0; i < em->num_long; i++) {
            const int64_t t = c->totals[doc_begin + order[(size_t)i]];
            if (t > em->long_max_total) em->long_max_total = t;
        }
        em->bulk_max_total = 0;
        for (int64_t i = em->num_long; i < D; i++) {
            const int64_t t = c->totals[doc_begin + order[(size_t)i]];
            if (t > em->bulk_max_total) em->bulk_max_total = t;
        }
        TRYCUDA(cudaMemcpyAsync(em->d_offsets, off.data(), sizeof(int64_t) * (size_t)(D + 1), cudaMemcpyHostToDevice, em->stream));
        if (D > 0) {
            TRYCUDA(cudaMemcpyAsync(em->d_totals, c->totals + doc_begin, sizeof(int32_t) * (size_t)D, cudaMemcpyHostToDevice, em->stream));
            TRYCUDA(cudaMemcpyAsync(em->d_order, order.data(), sizeof(int32_t) * (size_t)D, cudaMemcpyHostToDevice, em->stream));
            TRYCUDA(cudaMemcpyAsync(em->d_order_dyn, order.data(), sizeof(int32_t) * (size_t)D, cudaMemcpyHostToDevice, em->stream));
        }
        // windows of the bulk part of the order for the re-bundling (mstep.cu): at most 2048 documents, the longest at
        // most 1.1 times the shortest
        std::vector<int32_t> win;
        const char* env_rb = getenv("ONILDA_REBUNDLE");
        em->rebundle = env_rb ? atoi(env_rb) : 1;
        em->sweeps_valid = 0;
        if (em->rebundle && D < ((int64_t)1 << 31)) {
            int64_t at = em->num_long;
            win.push_back((int32_t)at);
            while (at < D) {
                const int64_t first_len = off[(size_t)order[(size_t)at] + 1] - off[(size_t)order[(size_t)at]];
                int64_t end = at + 1;
                while (end < D && end - at < 2048) {
                    const int64_t len = off[(size_t)order[(size_t)end] + 1] - off[(size_t)order[(size_t)end]];
                    if (len * 11 < first_len * 10) break;
                    end++;
                }
                win.push_back((int32_t)end);
                at = end;
            }
        }
        em->num_windows = win.empty() ? 0 : (int)win.size() - 1;
        if (em->num_windows > 0) {
            TRY(dev_alloc(&em->d_win_off, win.size()));
            TRYCUDA(cudaMemcpyAsync(em->d_win_off, win.data(), sizeof(int32_t) * win.size(), cudaMemcpyHostToDevice, em->stream));
        }
        if (N > 0) {
            TRYCUDA(cudaMemcpyAsync(em->d_words, c->words + base, sizeof(int32_t) * (size_t)N, cudaMemcpyHostToDevice, em->stream));
            TRYCUDA(cudaMemcpyAsync(em->d_counts, c->counts + base, sizeof(int32_t) * (size_t)N, cudaMemcpyHostToDevice, em->stream));
        }
        TRYCUDA(cudaStreamSynchronize(em->stream));   // the host vectors go out of scope
    }
    TRYCUDA(cudaMemsetAsync(em->d_ss, 0, sizeof(double) * (table_pad + SS_TAIL + lay.kpad), em->stream));
    TRYCUDA(cudaMemsetAsync(em->d_error, 0, 2 * sizeof(int), em->stream));
    TRY(upload_model(em, start->log_beta));
    TRYCUDA(cudaStreamSynchronize(em->stream));
#undef TRY
#undef TRYCUDA
    *out = em;
    return ONILDA_OK;
}

extern "C" int onilda_em_reload_corpus(onilda_em* em)
{
    if (!em) return onilda_fail(ONILDA_EINVAL, "em_reload_corpus: null argument");
    DeviceGuard g(em->device);
    const onilda_corpus* c = em->corpus;
    const int64_t base = c->offsets[em->doc_begin];
    // In pieces of 16 MB: a copy engine works through its queue in order, and a small copy or memset of ANOTHER handle
    // (its model, its cleared statistics: the start of its E-step) that lands behind one 430 MB copy waits for all of it.
    const size_t piece = (size_t)4 << 20;      // elements
    for (size_t at = 0; at < (size_t)em->nnz; at += piece) {
        const size_t n = (size_t)em->nnz - at < piece ? (size_t)em->nnz - at : piece;
        ONILDA_CUDA(cudaMemcpyAsync(em->d_words + at, c->words + base + at, sizeof(int32_t) * n, cudaMemcpyHostToDevice, em->stream));
        ONILDA_CUDA(cudaMemcpyAsync(em->d_counts + at, c->counts + base + at, sizeof(int32_t) * n, cudaMemcpyHostToDevice, em->stream));
    }
    if (em->num_docs > 0)
        ONILDA_CUDA(cudaMemcpyAsync(em->d_totals, c->totals + em->doc_begin, sizeof(int32_t) * (size_t)em->num_docs, cudaMemcpyHostToDevice, em->stream));
    // the host arrays belong to the caller: they may be reused as soon as this returns
    ONILDA_CUDA(cudaStreamSynchronize(em->stream));
    return ONILDA_OK;
}

extern "C" int onilda_em_reload_model(onilda_em* em, const onilda_model* m)
{
    if (!em || !m) return onilda_fail(ONILDA_EINVAL, "em_reload_model: null argument");
    if (m->num_topics != em->K || m->num_terms != em->V)
        return onilda_fail(ONILDA_EINVAL, "em_reload_model: model is %d x %d, handle is %d x %d", m->num_topics, m->num_terms, em->K, em->V);
    DeviceGuard g(em->device);
    int rc = upload_model(em, m->log_beta);
    if (rc != ONILDA_OK) return rc;
    em->alpha = m->alpha;
    em->var_max_iter = em->settings.var_max_iter;
    em->iteration = 0;
    em->likelihood_old = 0;
    em->converged = 1;
    em->estep_done = 0;
    em->sweeps_valid = 0;      // the next E-step is iteration 1 of another run: its sweep counts are not these
    ONILDA_CUDA(cudaStreamSynchronize(em->stream));
    return ONILDA_OK;
}

extern "C" int onilda_em_reload(onilda_em* em, const onilda_model* m)
{
    if (!em || !m) return onilda_fail(ONILDA_EINVAL, "em_reload: null argument");
    if (m->num_topics != em->K || m->num_terms != em->V)
        return onilda_fail(ONILDA_EINVAL, "em_reload: model is %d x %d, handle is %d x %d", m->num_topics, m->num_terms, em->K, em->V);
    int rc = onilda_em_reload_corpus(em);
    if (rc != ONILDA_OK) return rc;
    return onilda_em_reload_model(em, m);
}

static int run_estep(onilda_em* em, bool accumulate)
{
    const size_t table = table_doubles(em);
    int l = 0;
    ONILDA_CUDA(cudaEventRecord(em->ev0, em->stream));
    if (accumulate) {
        ONILDA_CUDA(cudaMemsetAsync(em->d_ss, 0, sizeof(double) * (table + SS_TAIL + em->lay.kpad), em->stream));   // zero_initialize_ss
    }
    EStepArgs a;
    a.offsets = em->d_offsets; a.words = em->d_words; a.counts = em->d_counts; a.totals = em->d_totals;
    a.order = em->d_order;
    // bundles of documents that stopped together in the last EM iteration (not after a reload: another run's sweeps)
    if (em->num_windows > 0 && em->sweeps_valid) {
        int rl = 0;
        int rrc = launch_rebundle(em->d_order, em->d_win_off, em->num_windows, em->d_doc_sweeps, em->d_order_dyn, em->stream, &rl);
        if (rrc != ONILDA_OK) return rrc;
        l += rl;
    }
    a.carveout_pct = em->num_long > 0 ? em->carveout_pct : 0;
    a.psi_tab = em->d_math_tab; a.exp_tab = em->d_math_tab + (size_t)6 * kPsiRows;
    a.beta_scaled = em->d_beta_scaled; a.lbmax = em->d_lbmax; a.num_topics = em->K; a.num_terms = em->V; a.idle_row = em->V + kTablePadRows; a.kstride = em->lay.kpad;
    a.alpha = em->alpha;
    a.lik_const = lgamma(em->alpha * em->K) - em->K * lgamma(em->alpha);
    a.var_max_iter = em->var_max_iter;
    a.var_converged = (double)em->settings.var_converged;
    a.phi = em->d_phi; a.gamma = em->d_gamma; a.doc_lik = em->d_doc_lik; a.doc_alpha_ss = em->d_doc_ass;
    a.doc_sweeps = em->d_doc_sweeps;
    a.class_word = accumulate ? em->d_ss : nullptr;
    a.error_flag = em->d_error;
    ONILDA_CUDA(cudaMemsetAsync(em->d_queue, 0, 2 * sizeof(unsigned long long), em->stream));
    ONILDA_CUDA(cudaMemsetAsync(em->d_error + 1, 0, sizeof(int), em->stream));
    // The general kernel without its rare paths (estep.cu) where no gamma can leave the math tables (every gamma of a
    // document lies within [alpha, alpha + its tokens]) and no word's sum can leave the range of its reciprocal
    // (alpha >= 0.0016).  ONILDA_RARE_PATHS=1 forces the full variant.
    auto needs_rare_paths = [&](int64_t max_total) {
        return getenv("ONILDA_RARE_PATHS") != nullptr ||
               !(em->alpha >= 0.0016 && em->alpha + (double)max_total + 1.0 < (double)(1 << kPsiOctaves) - 1.0);
    };
    a.rare_paths = 1;
    int rc;
    if (em->num_long > 0 && em->pdl) {
        // Both launches in ONE stream, the bulk one as the secondary of a programmatic dependent launch: its blocks are
        // scheduled when every long-document block is resident (each executes griddepcontrol.launch_dependents when it
        // starts), not when that kernel has finished.  With two streams the order in which the block scheduler took
        // the two grids was a race: when the bulk grid, sized for the empty machine, got there first it filled every
        // SM, the long documents ran only after it, and the E-step took 150-200 ms instead of 105 (about three
        // iterations in eighteen of a bench run).
        EStepArgs b = a;
        b.order_begin = 0; b.order_count = em->num_long; b.queue = em->d_queue + 1;
        b.rare_paths = needs_rare_paths(em->long_max_total);
        const bool tables_cover = em->alpha >= 1e-8 && em->alpha + (double)em->long_max_total + 1.0 < (double)(1 << kPsiOctaves) - 1.0;
        if (em->long_group == 32 && em->K <= 64 && tables_cover && !getenv("ONILDA_LONG_GENERIC"))
            rc = launch_estep_long(b, em->sm_count, em->long_threads, em->long_per_sm, em->stream, &l);
        else
            rc = launch_estep(em->long_group, b, em->sm_count, em->long_threads, em->long_smem, em->long_per_sm, em->stream, &l);
        if (rc != ONILDA_OK) return rc;
        a.order_begin = em->num_long; a.order_count = em->num_docs - em->num_long; a.queue = em->d_queue;
        a.order = em->d_order_dyn;
        a.rare_paths = needs_rare_paths(em->bulk_max_total);
        rc = launch_estep(em->bulk_group, a, em->sm_count, em->bulk_threads, 0, em->bulk_per_sm, em->stream, &l, true);
        if (rc != ONILDA_OK) return rc;
    } else {
        if (em->num_long > 0) {
            // fork: long documents first, on their own stream, one warp per block so that the block
            // scheduler spreads them over the SMs; the bulk launch fills the rest of the machine
            ONILDA_CUDA(cudaEventRecord(em->ev_fork, em->stream));
            ONILDA_CUDA(cudaStreamWaitEvent(em->long_stream, em->ev_fork, 0));
            // ... and the main stream waits until the long stream has got past that wait: both launches then become
            // runnable at the same moment and the long one, enqueued first on the higher-priority stream, gets its
            // blocks placed first.  Without this the bulk launch can start alone, fill every SM with its persistent
            // blocks, and the long documents run only after it (measured: 157 ms instead of 105 in one iteration of ten).
            ONILDA_CUDA(cudaEventRecord(em->ev_forked, em->long_stream));
            ONILDA_CUDA(cudaStreamWaitEvent(em->stream, em->ev_forked, 0));
            EStepArgs b = a;
            b.order_begin = 0; b.order_count = em->num_long; b.queue = em->d_queue + 1;
            b.rare_paths = needs_rare_paths(em->long_max_total);
            // The staged one-document-per-warp kernel reads psi and exp(psi) from the tables only: every gamma of its
            // documents, within [alpha, alpha + total], must lie inside their range (device_math.cuh).  Otherwise the
            // general kernel, which has the closed forms, takes the long documents too.
            const bool tables_cover = em->alpha >= 1e-8 && em->alpha + (double)em->long_max_total + 1.0 < (double)(1 << kPsiOctaves) - 1.0;
            if (em->long_group == 32 && em->K <= 64 && tables_cover && !getenv("ONILDA_LONG_GENERIC"))
                rc = launch_estep_long(b, em->sm_count, em->long_threads, em->long_per_sm, em->long_stream, &l);
            else
                rc = launch_estep(em->long_group, b, em->sm_count, em->long_threads, em->long_smem, em->long_per_sm, em->long_stream, &l);
            if (rc != ONILDA_OK) return rc;
            if (em->timing) ONILDA_CUDA(cudaEventRecord(em->ev_long_end, em->long_stream));
            ONILDA_CUDA(cudaEventRecord(em->ev_join, em->long_stream));
        }
        a.order_begin = em->num_long; a.order_count = em->num_docs - em->num_long; a.queue = em->d_queue;
        a.order = em->d_order_dyn;
        a.rare_paths = needs_rare_paths(em->bulk_max_total);
        rc = launch_estep(em->bulk_group, a, em->sm_count, em->bulk_threads, 0, em->num_long > 0 ? em->bulk_per_sm : 0, em->stream, &l);
        if (rc != ONILDA_OK) return rc;
        if (em->timing) ONILDA_CUDA(cudaEventRecord(em->ev_bulk_end, em->stream));
        if (em->num_long > 0) ONILDA_CUDA(cudaStreamWaitEvent(em->stream, em->ev_join, 0));
    }
    rc = launch_reduce_doc_stats(em->d_doc_lik, em->d_doc_ass, em->d_offsets, em->d_doc_sweeps, em->num_docs, em->d_partial,
                                 em->d_ss + table, em->stream, &l);
    if (rc != ONILDA_OK) return rc;
    ONILDA_CUDA(cudaEventRecord(em->ev1, em->stream));
    em->launches = l;
    em->estep_done = 1;
    em->sweeps_valid = 1;
    return ONILDA_OK;
}

extern "C" int onilda_em_estep(onilda_em* em)
{
    if (!em) return onilda_fail(ONILDA_EINVAL, "em_estep: null handle");
    DeviceGuard g(em->device);
    return run_estep(em, true);
}

extern "C" int onilda_em_suffstats_device(onilda_em* em, double** dev_ptr, int64_t* count)
{
    if (!em || !dev_ptr || !count) return onilda_fail(ONILDA_EINVAL, "em_suffstats_device: null argument");
    *dev_ptr = em->d_ss;
    *count = (int64_t)table_doubles(em) + SS_TAIL;
    return ONILDA_OK;
}

extern "C" void* onilda_em_stream(onilda_em* em) { return em ? (void*)em->stream : nullptr; }

// The M-step over all shards of one process (n handles; one per process under torchrun).
//
// One shard: column sums -> lda_mle rows (lda-model.c:27-44) on the device.
// P shards (a communicator attached): what the reference does with 2K + 3 MPI_Allreduce calls and a replicated
// lda_mle on every rank (lda-estimate.c:211-256) becomes, stream-ordered and without a host synchronisation,
//     ncclReduceScatter of class_word      each rank receives the sums of its own ceil(V/P) rows
//     column sums of the own rows          partial class_total
//     ncclAllReduce of 4 + kpad doubles    likelihood, alpha statistic, document count, class_total
//     lda_mle on the own rows              log(cw) - log(ct) | -100, beta', lbmax
//     ncclAllGather x 3                    log_beta, beta', lbmax: every rank has the whole model again
// which moves the same bytes as an all-reduce of K x V doubles and normalises each row once instead of P times.
static int mstep_group(onilda_em** ems, int n, double* likelihood, double* converged, int* keep_going)
{
    for (int i = 0; i < n; i++) {
        if (!ems[i]) return onilda_fail(ONILDA_EINVAL, "em_mstep: null handle");
        if (!ems[i]->estep_done) return onilda_fail(ONILDA_EINVAL, "em_mstep: call onilda_em_estep first");
        if (ems[i]->nranks != ems[0]->nranks || ems[i]->V != ems[0]->V || ems[i]->K != ems[0]->K)
            return onilda_fail(ONILDA_EINVAL, "em_mstep: handles of different shapes in one group");
    }
    const int P = ems[0]->nranks;
    if (P > 1 && !ems[0]->comm) return onilda_fail(ONILDA_EINVAL, "em_mstep: no communicator attached");
    const NcclApi* nccl = P > 1 ? nccl_api() : nullptr;
    if (P > 1 && !nccl) return ONILDA_ECUDA;
    int prev = 0;
    ONILDA_CUDA(cudaGetDevice(&prev));
    struct Restore { int d; ~Restore() { cudaSetDevice(d); } } restore{prev};
    const int kpad = ems[0]->lay.kpad;
    const size_t table = table_doubles(ems[0]);
    const int64_t rows_per_rank = P > 1 ? ((int64_t)ems[0]->V + P - 1) / P : (int64_t)ems[0]->V;
    std::vector<int> launches((size_t)n, 0);
#define NCCL_TRY(call)                                                                                              \
    do {                                                                                                            \
        ncclResult_t r__ = (call);                                                                                  \
        if (r__ != ncclSuccess) return onilda_fail(ONILDA_ECUDA, "%s failed: %s", #call, nccl->GetErrorString(r__)); \
    } while (0)

    for (int i = 0; i < n; i++) {
        ONILDA_CUDA(cudaSetDevice(ems[i]->device));
        ONILDA_CUDA(cudaEventRecord(ems[i]->ev2, ems[i]->stream));
        ONILDA_CUDA(cudaEventRecord(ems[i]->ev_c0, ems[i]->stream));
    }
    if (P > 1) {
        NCCL_TRY(nccl->GroupStart());
        for (int i = 0; i < n; i++) {
            onilda_em* em = ems[i];
            const size_t cnt = (size_t)rows_per_rank * kpad;
            NCCL_TRY(nccl->ReduceScatter(em->d_ss, em->d_ss + (size_t)em->rank * cnt, cnt, ncclDouble, ncclSum, em->comm, em->stream));
        }
        NCCL_TRY(nccl->GroupEnd());
    }
    for (int i = 0; i < n; i++) {
        onilda_em* em = ems[i];
        ONILDA_CUDA(cudaSetDevice(em->device));
        ONILDA_CUDA(cudaEventRecord(em->ev_c1, em->stream));
        const size_t lo = P > 1 ? (size_t)em->rank * rows_per_rank : 0;
        int rc = launch_column_sums(em->d_ss + lo * kpad, (int)rows_per_rank, em->lay, em->sm_count, em->d_partial,
                                    em->d_ss + table + SS_TAIL, em->stream, &launches[(size_t)i]);
        if (rc != ONILDA_OK) return rc;
    }
    if (P > 1) {
        NCCL_TRY(nccl->GroupStart());
        for (int i = 0; i < n; i++)
            NCCL_TRY(nccl->AllReduce(ems[i]->d_ss + table, ems[i]->d_ss + table, (size_t)SS_TAIL + kpad, ncclDouble, ncclSum,
                                   ems[i]->comm, ems[i]->stream));
        NCCL_TRY(nccl->GroupEnd());
    }
    for (int i = 0; i < n; i++) {
        onilda_em* em = ems[i];
        ONILDA_CUDA(cudaSetDevice(em->device));
        const size_t lo = P > 1 ? (size_t)em->rank * rows_per_rank : 0;
        int rc = launch_normalise(em->d_ss + lo * kpad, em->d_ss + table + SS_TAIL, em->K, (int)rows_per_rank, em->lay,
                                  em->d_log_beta_wk + lo * kpad, em->d_beta_scaled + lo * kpad, em->d_lbmax + lo, em->stream,
                                  &launches[(size_t)i]);
        if (rc != ONILDA_OK) return rc;
        ONILDA_CUDA(cudaEventRecord(em->ev_c2, em->stream));
    }
    if (P > 1) {
        NCCL_TRY(nccl->GroupStart());
        for (int i = 0; i < n; i++) {
            onilda_em* em = ems[i];
            const size_t cnt = (size_t)rows_per_rank * kpad, lo = (size_t)em->rank * rows_per_rank;
            NCCL_TRY(nccl->AllGather(em->d_log_beta_wk + lo * kpad, em->d_log_beta_wk, cnt, ncclDouble, em->comm, em->stream));
            NCCL_TRY(nccl->AllGather(em->d_beta_scaled + lo * kpad, em->d_beta_scaled, cnt, ncclDouble, em->comm, em->stream));
            NCCL_TRY(nccl->AllGather(em->d_lbmax + lo, em->d_lbmax, (size_t)rows_per_rank, ncclDouble, em->comm, em->stream));
        }
        NCCL_TRY(nccl->GroupEnd());
    }
#undef NCCL_TRY
    std::vector<double> tails((size_t)n * SS_TAIL);
    std::vector<int> errs((size_t)n, 0);
    for (int i = 0; i < n; i++) {
        onilda_em* em = ems[i];
        ONILDA_CUDA(cudaSetDevice(em->device));
        ONILDA_CUDA(cudaEventRecord(em->ev_c3, em->stream));
        ONILDA_CUDA(cudaMemcpyAsync(&tails[(size_t)i * SS_TAIL], em->d_ss + table, sizeof(double) * SS_TAIL, cudaMemcpyDeviceToHost, em->stream));
        ONILDA_CUDA(cudaMemcpyAsync(&errs[(size_t)i], em->d_error, sizeof(int), cudaMemcpyDeviceToHost, em->stream));
        ONILDA_CUDA(cudaEventRecord(em->ev3, em->stream));
    }
    for (int i = 0; i < n; i++) {
        onilda_em* em = ems[i];
        ONILDA_CUDA(cudaSetDevice(em->device));
        ONILDA_CUDA(cudaStreamSynchronize(em->stream));
        em->launches += launches[(size_t)i];
        em->estep_done = 0;
        float ms = 0;
        if (cudaEventElapsedTime(&ms, em->ev0, em->ev1) == cudaSuccess) em->estep_ms = ms;
        if (em->timing && em->pdl && em->num_long > 0) {
            fprintf(stderr, "[onilda] E-step %.2f ms (the two launches share one stream; ONILDA_PDL=0 times them apart)\n", em->estep_ms);
        } else if (em->timing) {
            float tl = 0, tb = 0;
            if (em->num_long > 0) cudaEventElapsedTime(&tl, em->ev0, em->ev_long_end);
            cudaEventElapsedTime(&tb, em->ev0, em->ev_bulk_end);
            fprintf(stderr, "[onilda] E-step %.2f ms: long-document launch ended at %.2f ms, bulk launch at %.2f ms\n",
                    em->estep_ms, tl, tb);
        }
        if (cudaEventElapsedTime(&ms, em->ev2, em->ev3) == cudaSuccess) em->mstep_ms = ms;
        if (cudaEventElapsedTime(&ms, em->ev_c0, em->ev_c1) == cudaSuccess) em->rs_ms = ms;
        if (cudaEventElapsedTime(&ms, em->ev_c2, em->ev_c3) == cudaSuccess) em->ag_ms = ms;
        cudaGetLastError();
    }
    for (int i = 0; i < n; i++)
        if (errs[(size_t)i] || isnan(tails[(size_t)i * SS_TAIL + SS_LIK]))
            return onilda_fail(ONILDA_ENAN, "NaN likelihood in the E-step (the reference asserts at lda-inference.c:78)");

    // every shard holds the same reduced scalars; the host part of lda_mle and run_em's bookkeeping
    // (lda-model.c:45-53, lda-estimate.c:186-188, 260-262) runs once and is copied to each handle
    const double* tail = tails.data();
    const double lik = tail[SS_LIK];
    onilda_em* e0 = ems[0];
    double alpha = e0->alpha;
    if (e0->settings.estimate_alpha == 1) alpha = onilda_opt_alpha(tail[SS_ALPHA], (int)llround(tail[SS_NDOCS]), e0->K);
    const double conv = (e0->likelihood_old - lik) / (e0->likelihood_old);
    for (int i = 0; i < n; i++) {
        onilda_em* em = ems[i];
        em->alpha = alpha;
        em->iteration++;
        em->converged = conv;
        if (em->converged < 0) em->var_max_iter = em->var_max_iter * 2;
        em->likelihood_old = lik;
        em->word_steps = tail[SS_STEPS];
    }
    if (likelihood) *likelihood = lik;
    if (converged) *converged = conv;
    if (keep_going)
        *keep_going = ((e0->converged < 0) || (e0->converged > e0->settings.em_converged) || (e0->iteration <= 2)) &&
                      (e0->iteration <= e0->settings.em_max_iter);
    return ONILDA_OK;
}

extern "C" int onilda_em_mstep(onilda_em* em, double* likelihood, double* converged, int* keep_going)
{
    return mstep_group(&em, 1, likelihood, converged, keep_going);
}

extern "C" int onilda_em_mstep_group(onilda_em** ems, int n, double* likelihood, double* converged, int* keep_going)
{
    if (!ems || n <= 0) return onilda_fail(ONILDA_EINVAL, "em_mstep_group: no handles");
    return mstep_group(ems, n, likelihood, converged, keep_going);
}

extern "C" int onilda_comm_unique_id(unsigned char* id)
{
    if (!id) return onilda_fail(ONILDA_EINVAL, "comm_unique_id: null argument");
    static_assert(sizeof(ncclUniqueId) == ONILDA_COMM_ID_BYTES, "ONILDA_COMM_ID_BYTES must be NCCL's unique-id size");
    const NcclApi* nccl = nccl_api();
    if (!nccl) return ONILDA_ECUDA;
    ncclUniqueId u;
    ncclResult_t r = nccl->GetUniqueId(&u);
    if (r != ncclSuccess) return onilda_fail(ONILDA_ECUDA, "ncclGetUniqueId: %s", nccl->GetErrorString(r));
    memcpy(id, &u, sizeof u);
    return ONILDA_OK;
}

extern "C" int onilda_em_comm_attach(onilda_em* em, const unsigned char* id, int nranks, int rank)
{
    if (!em || !id) return onilda_fail(ONILDA_EINVAL, "em_comm_attach: null argument");
    if (nranks < 1 || nranks > kTablePadRows || rank < 0 || rank >= nranks)
        return onilda_fail(ONILDA_EINVAL, "em_comm_attach: rank %d of %d (1..%d ranks)", rank, nranks, (int)kTablePadRows);
    if (em->comm) return onilda_fail(ONILDA_EINVAL, "em_comm_attach: the handle already has a communicator");
    if (nranks == 1) return ONILDA_OK;
    const NcclApi* nccl = nccl_api();
    if (!nccl) return ONILDA_ECUDA;
    DeviceGuard g(em->device);
    ncclUniqueId u;
    memcpy(&u, id, sizeof u);
    ncclResult_t r = nccl->CommInitRank(&em->comm, nranks, u, rank);
    if (r != ncclSuccess) { em->comm = nullptr; return onilda_fail(ONILDA_ECUDA, "ncclCommInitRank: %s", nccl->GetErrorString(r)); }
    em->nranks = nranks;
    em->rank = rank;
    return ONILDA_OK;
}

extern "C" int onilda_em_comm_attach_local(onilda_em** ems, int n)
{
    if (!ems || n < 1 || n > kTablePadRows) return onilda_fail(ONILDA_EINVAL, "em_comm_attach_local: 1..%d handles", (int)kTablePadRows);
    if (n == 1) return ONILDA_OK;
    std::vector<int> devs((size_t)n);
    for (int i = 0; i < n; i++) {
        if (!ems[i] || ems[i]->comm) return onilda_fail(ONILDA_EINVAL, "em_comm_attach_local: null handle or communicator present");
        devs[(size_t)i] = ems[i]->device;
    }
    const NcclApi* nccl = nccl_api();
    if (!nccl) return ONILDA_ECUDA;
    std::vector<ncclComm_t> comms((size_t)n);
    ncclResult_t r = nccl->CommInitAll(comms.data(), n, devs.data());
    if (r != ncclSuccess) return onilda_fail(ONILDA_ECUDA, "ncclCommInitAll: %s", nccl->GetErrorString(r));
    for (int i = 0; i < n; i++) {
        ems[i]->comm = comms[(size_t)i];
        ems[i]->nranks = n;
        ems[i]->rank = i;
    }
    return ONILDA_OK;
}

// The exchange of one M-step by itself: `repeats` reduce-scatters of the statistics table and as many rounds of the
// three all-gathers, each set behind a barrier-like small all-reduce and timed with events on the handle's stream.
// The statistics buffer is left undefined (the next E-step clears it); the model tables are gathered onto themselves.
extern "C" int onilda_em_collective_time(onilda_em** ems, int n, int repeats, double* rs_ms, double* ag_ms, int64_t* bytes)
{
    if (!ems || n <= 0 || repeats <= 0) return onilda_fail(ONILDA_EINVAL, "em_collective_time: bad argument");
    const int P = ems[0]->nranks;
    if (P < 2 || !ems[0]->comm) return onilda_fail(ONILDA_EINVAL, "em_collective_time: no communicator attached");
    const NcclApi* nccl = nccl_api();
    if (!nccl) return ONILDA_ECUDA;
    int prev = 0;
    ONILDA_CUDA(cudaGetDevice(&prev));
    struct Restore { int d; ~Restore() { cudaSetDevice(d); } } restore{prev};
    const int kpad = ems[0]->lay.kpad;
    const size_t table = table_doubles(ems[0]);
    const size_t rows = (size_t)(((int64_t)ems[0]->V + P - 1) / P), cnt = rows * kpad;
#define NCCL_TRY(call)                                                                                              \
    do {                                                                                                            \
        ncclResult_t r__ = (call);                                                                                  \
        if (r__ != ncclSuccess) return onilda_fail(ONILDA_ECUDA, "%s failed: %s", #call, nccl->GetErrorString(r__)); \
    } while (0)
    for (int phase = 0; phase < 2; phase++) {
        NCCL_TRY(nccl->GroupStart());
        for (int i = 0; i < n; i++)
            NCCL_TRY(nccl->AllReduce(ems[i]->d_ss + table, ems[i]->d_ss + table, 1, ncclDouble, ncclSum, ems[i]->comm, ems[i]->stream));
        NCCL_TRY(nccl->GroupEnd());
        for (int i = 0; i < n; i++) { ONILDA_CUDA(cudaSetDevice(ems[i]->device)); ONILDA_CUDA(cudaEventRecord(ems[i]->ev_c0, ems[i]->stream)); }
        for (int it = 0; it < repeats; it++) {
            NCCL_TRY(nccl->GroupStart());
            for (int i = 0; i < n; i++) {
                onilda_em* em = ems[i];
                const size_t lo = (size_t)em->rank * rows;
                if (phase == 0) {
                    NCCL_TRY(nccl->ReduceScatter(em->d_ss, em->d_ss + lo * kpad, cnt, ncclDouble, ncclSum, em->comm, em->stream));
                } else {
                    NCCL_TRY(nccl->AllGather(em->d_log_beta_wk + lo * kpad, em->d_log_beta_wk, cnt, ncclDouble, em->comm, em->stream));
                    NCCL_TRY(nccl->AllGather(em->d_beta_scaled + lo * kpad, em->d_beta_scaled, cnt, ncclDouble, em->comm, em->stream));
                    NCCL_TRY(nccl->AllGather(em->d_lbmax + lo, em->d_lbmax, rows, ncclDouble, em->comm, em->stream));
                }
            }
            NCCL_TRY(nccl->GroupEnd());
        }
        double worst = 0;
        for (int i = 0; i < n; i++) {
            ONILDA_CUDA(cudaSetDevice(ems[i]->device));
            ONILDA_CUDA(cudaEventRecord(ems[i]->ev_c1, ems[i]->stream));
            ONILDA_CUDA(cudaStreamSynchronize(ems[i]->stream));
            float ms = 0;
            ONILDA_CUDA(cudaEventElapsedTime(&ms, ems[i]->ev_c0, ems[i]->ev_c1));
            if (ms > worst) worst = ms;
        }
        if (phase == 0 && rs_ms) *rs_ms = worst / repeats;
        if (phase == 1 && ag_ms) *ag_ms = worst / repeats;
    }
#undef NCCL_TRY
    if (bytes) *bytes = (int64_t)(cnt * P * sizeof(double));
    return ONILDA_OK;
}

extern "C" int onilda_em_comm_stats(onilda_em* em, int* nranks, int* rank, double* reduce_scatter_ms, double* all_gather_ms)
{
    if (!em) return onilda_fail(ONILDA_EINVAL, "em_comm_stats: null handle");
    if (nranks) *nranks = em->nranks;
    if (rank) *rank = em->rank;
    if (reduce_scatter_ms) *reduce_scatter_ms = em->rs_ms;
    if (all_gather_ms) *all_gather_ms = em->ag_ms;
    return ONILDA_OK;
}

extern "C" int onilda_em_iterate(onilda_em* em, double* likelihood, double* converged, int* keep_going)
{
    int rc = onilda_em_estep(em);
    if (rc != ONILDA_OK) return rc;
    return onilda_em_mstep(em, likelihood, converged, keep_going);
}

extern "C" int onilda_em_get_model(onilda_em* em, onilda_model* out)
{
    if (!em || !out) return onilda_fail(ONILDA_EINVAL, "em_get_model: null argument");
    if (out->num_topics != em->K || out->num_terms != em->V)
        return onilda_fail(ONILDA_EINVAL, "em_get_model: model is %d x %d, handle is %d x %d", out->num_topics,
                           out->num_terms, em->K, em->V);
    DeviceGuard g(em->device);
    int l = 0;
    int rc = launch_export_log_beta(em->d_log_beta_wk, em->K, em->V, em->lay, em->d_kv, em->stream, &l);
    if (rc != ONILDA_OK) return rc;
    if ((size_t)em->K * em->V > 0)
        ONILDA_CUDA(cudaMemcpyAsync(out->log_beta, em->d_kv, sizeof(double) * (size_t)em->K * em->V, cudaMemcpyDeviceToHost, em->stream));
    ONILDA_CUDA(cudaStreamSynchronize(em->stream));
    out->alpha = em->alpha;
    return ONILDA_OK;
}

extern "C" int onilda_em_get_gamma(onilda_em* em, double* gamma_out)
{
    if (!em || (!gamma_out && em->num_docs > 0)) return onilda_fail(ONILDA_EINVAL, "em_get_gamma: null argument");
    DeviceGuard g(em->device);
    if (em->num_docs > 0)
        ONILDA_CUDA(cudaMemcpyAsync(gamma_out, em->d_gamma, sizeof(double) * (size_t)em->num_docs * em->K, cudaMemcpyDeviceToHost, em->stream));
    ONILDA_CUDA(cudaStreamSynchronize(em->stream));
    return ONILDA_OK;
}

extern "C" int onilda_em_get_doc_likelihood(onilda_em* em, double* out)
{
    if (!em || (!out && em->num_docs > 0)) return onilda_fail(ONILDA_EINVAL, "em_get_doc_likelihood: null argument");
    DeviceGuard g(em->device);
    if (em->num_docs > 0)
        ONILDA_CUDA(cudaMemcpyAsync(out, em->d_doc_lik, sizeof(double) * (size_t)em->num_docs, cudaMemcpyDeviceToHost, em->stream));
    ONILDA_CUDA(cudaStreamSynchronize(em->stream));
    return ONILDA_OK;
}

extern "C" int onilda_em_device_views(onilda_em* em, int* device, int64_t* num_docs, int* num_topics, int* num_terms, int* kpad,
                                      const double** gamma_dev, const double** log_beta_wk_dev)
{
    if (!em) return onilda_fail(ONILDA_EINVAL, "em_device_views: null handle");
    DeviceGuard g(em->device);
    ONILDA_CUDA(cudaStreamSynchronize(em->stream));
    *device = em->device; *num_docs = em->num_docs; *num_topics = em->K; *num_terms = em->V; *kpad = em->lay.kpad;
    *gamma_dev = em->d_gamma; *log_beta_wk_dev = em->d_log_beta_wk;
    return ONILDA_OK;
}

extern "C" int onilda_em_get_doc_sweeps(onilda_em* em, int32_t* out)
{
    if (!em || (!out && em->num_docs > 0)) return onilda_fail(ONILDA_EINVAL, "em_get_doc_sweeps: null argument");
    DeviceGuard g(em->device);
    if (em->num_docs > 0)
        ONILDA_CUDA(cudaMemcpyAsync(out, em->d_doc_sweeps, sizeof(int32_t) * (size_t)em->num_docs, cudaMemcpyDeviceToHost, em->stream));
    ONILDA_CUDA(cudaStreamSynchronize(em->stream));
    return ONILDA_OK;
}

// Words of the last E-step whose sum over topics left the range of the product form and were redone in log space
// (estep.cu header; the reference works in log space throughout, lda-inference.c:52-68).  A diagnostic: the tests
// use it to know that a constructed case really took that path.
extern "C" int onilda_em_log_space_words(onilda_em* em, int64_t* count)
{
    if (!em || !count) return onilda_fail(ONILDA_EINVAL, "em_log_space_words: null argument");
    DeviceGuard g(em->device);
    int v = 0;
    ONILDA_CUDA(cudaMemcpyAsync(&v, em->d_error + 1, sizeof v, cudaMemcpyDeviceToHost, em->stream));
    ONILDA_CUDA(cudaStreamSynchronize(em->stream));
    *count = v;
    return ONILDA_OK;
}

// (document, word, sweep) steps of the last iteration's E-step - entries x variational sweeps, summed over all
// documents of all shards: the E-step's unit of work.  It grows as the topics sharpen (more documents run to
// the sweep cap), which is why an EM iteration late in a run takes longer than an early one on the same corpus.
extern "C" int onilda_em_word_steps(onilda_em* em, double* steps)
{
    if (!em || !steps) return onilda_fail(ONILDA_EINVAL, "em_word_steps: null argument");
    *steps = em->word_steps;
    return ONILDA_OK;
}

extern "C" int onilda_em_stats(onilda_em* em, double* estep_ms, double* mstep_ms, int64_t* launches, int64_t* sweeps)
{
    if (!em) return onilda_fail(ONILDA_EINVAL, "em_stats: null handle");
    if (estep_ms) *estep_ms = em->estep_ms;
    if (mstep_ms) *mstep_ms = em->mstep_ms;
    if (launches) *launches = em->launches;
    if (sweeps) {
        DeviceGuard g(em->device);
        std::vector<int32_t> sw((size_t)em->num_docs);
        if (em->num_docs > 0)
            ONILDA_CUDA(cudaMemcpy(sw.data(), em->d_doc_sweeps, sizeof(int32_t) * (size_t)em->num_docs, cudaMemcpyDeviceToHost));
        int64_t tot = 0;
        for (int32_t v : sw) tot += v;
        *sweeps = tot;
    }
    return ONILDA_OK;
}

// infer (lda-estimate.c:358-388): E-step without sufficient statistics
extern "C" int onilda_infer(const onilda_model* m, const onilda_corpus* c, const onilda_settings* settings, int device,
                            double* gamma_out, double* likelihood_out)
{
    if (!m || !c || !settings) return onilda_fail(ONILDA_EINVAL, "infer: null argument");
    onilda_em* em = nullptr;
    int rc = onilda_em_create(c, 0, c->num_docs, m, settings, device, &em);
    if (rc != ONILDA_OK) return rc;
    {
        DeviceGuard g(em->device);
        rc = run_estep(em, false);
        if (rc == ONILDA_OK) {
            int err = 0;
            cudaError_t e = cudaMemcpyAsync(&err, em->d_error, sizeof err, cudaMemcpyDeviceToHost, em->stream);
            if (e == cudaSuccess) e = cudaStreamSynchronize(em->stream);
            if (e != cudaSuccess) rc = onilda_fail(ONILDA_ECUDA, "infer: %s", cudaGetErrorString(e));
            else if (err) rc = onilda_fail(ONILDA_ENAN, "NaN likelihood in inference (the reference asserts at lda-inference.c:78)");
        }
    }
    if (rc == ONILDA_OK && gamma_out) rc = onilda_em_get_gamma(em, gamma_out);
    if (rc == ONILDA_OK && likelihood_out) rc = onilda_em_get_doc_likelihood(em, likelihood_out);
    onilda_em_destroy(em);
    return rc;
}
