// run_em / infer drivers (reference: oni-ml/oni-lda-c/lda-estimate.c:112-321, 358-388)
// and the in-process sum of sufficient statistics over NVLink peer memory that
// stands where the reference has its MPI_Allreduce calls (:211-252).
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>

#include <chrono>
#include <string>
#include <vector>

#include "kernels.cuh"

using namespace onilda;

namespace {

constexpr int kMaxPeers = 16;

struct PeerPtrs {
    double* buf[kMaxPeers];
    int n;
};

// Device `me` owns elements [lo, hi): it reads that slice from every peer's
// buffer (rank order 0..n-1: the cross-device sum has a fixed order; each device's own buffer was filled
// with fp64 atomics, so its inputs still vary in the last bits from run to run), adds, and
// stores the sum into every peer's buffer.  Loads and stores on peer pointers
// travel over NVLink; slices of different devices are disjoint.
__global__ void __launch_bounds__(256) peer_sum_kernel(PeerPtrs p, int64_t lo, int64_t hi)
{
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = lo + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += stride) {
        double acc = 0.0;
        for (int r = 0; r < p.n; r++) acc += p.buf[r][i];
        for (int r = 0; r < p.n; r++) p.buf[r][i] = acc;
    }
}

}  // namespace

extern "C" int onilda_em_allreduce_peer(onilda_em** ems, int n)
{
    if (!ems || n <= 0 || n > kMaxPeers) return onilda_fail(ONILDA_EINVAL, "em_allreduce_peer: need 1..%d handles", kMaxPeers);
    if (n == 1) return ONILDA_OK;
    PeerPtrs p;
    p.n = n;
    int64_t count = 0;
    std::vector<int> dev(n);
    std::vector<cudaStream_t> st(n);
    for (int r = 0; r < n; r++) {
        int64_t c = 0;
        int rc = onilda_em_suffstats_device(ems[r], &p.buf[r], &c);
        if (rc != ONILDA_OK) return rc;
        if (r == 0) count = c;
        else if (c != count) return onilda_fail(ONILDA_EINVAL, "em_allreduce_peer: handles disagree on the buffer size");
        st[r] = (cudaStream_t)onilda_em_stream(ems[r]);
        cudaPointerAttributes at;
        ONILDA_CUDA(cudaPointerGetAttributes(&at, p.buf[r]));
        dev[r] = at.device;
    }
    int prev = 0;
    ONILDA_CUDA(cudaGetDevice(&prev));
    // peer access, both directions (idempotent)
    for (int a = 0; a < n; a++) {
        ONILDA_CUDA(cudaSetDevice(dev[a]));
        for (int b = 0; b < n; b++) {
            if (dev[a] == dev[b]) continue;
            int can = 0;
            ONILDA_CUDA(cudaDeviceCanAccessPeer(&can, dev[a], dev[b]));
            if (!can) { cudaSetDevice(prev); return onilda_fail(ONILDA_ECUDA, "device %d cannot map device %d's memory", dev[a], dev[b]); }
            cudaError_t e = cudaDeviceEnablePeerAccess(dev[b], 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
                cudaSetDevice(prev);
                return onilda_fail(ONILDA_ECUDA, "cudaDeviceEnablePeerAccess(%d -> %d): %s", dev[a], dev[b], cudaGetErrorString(e));
            }
            cudaGetLastError();
        }
    }
    // every E-step must have finished before any slice is summed
    for (int r = 0; r < n; r++) { ONILDA_CUDA(cudaSetDevice(dev[r])); ONILDA_CUDA(cudaStreamSynchronize(st[r])); }
    const int64_t chunk = (count + n - 1) / n;
    for (int r = 0; r < n; r++) {
        int64_t lo = chunk * r, hi = lo + chunk;
        if (hi > count) hi = count;
        if (lo >= hi) continue;
        ONILDA_CUDA(cudaSetDevice(dev[r]));
        int blocks = (int)((hi - lo + 255) / 256), sms = 0;
        ONILDA_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev[r]));
        if (blocks > sms * 8) blocks = sms * 8;
        peer_sum_kernel<<<blocks, 256, 0, st[r]>>>(p, lo, hi);
        ONILDA_CUDA(cudaGetLastError());
    }
    // ... and every slice must have landed everywhere before any M-step reads it
    for (int r = 0; r < n; r++) { ONILDA_CUDA(cudaSetDevice(dev[r])); ONILDA_CUDA(cudaStreamSynchronize(st[r])); }
    ONILDA_CUDA(cudaSetDevice(prev));
    return ONILDA_OK;
}

namespace {

// contiguous document ranges balanced by nnz (+1 per document for the per-document work)
void split_by_nnz(const onilda_corpus* c, int n, std::vector<int64_t>& bounds)
{
    bounds.assign((size_t)n + 1, c->num_docs);
    bounds[0] = 0;
    const double total = (double)c->nnz + (double)c->num_docs;
    int64_t d = 0;
    for (int r = 1; r < n; r++) {
        const double want = total * r / n;
        while (d < c->num_docs && (double)c->offsets[d] + (double)d < want) d++;
        bounds[(size_t)r] = d;
    }
}

struct EmSet {
    std::vector<onilda_em*> ems;
    ~EmSet() { for (onilda_em* e : ems) onilda_em_destroy(e); }
};

}  // namespace

extern "C" int onilda_run_em(const char* start, const char* directory, const onilda_corpus* c, int num_topics,
                             double initial_alpha, const onilda_settings* settings, int ndev, uint32_t seed)
{
    if (!start || !directory || !c || !settings) return onilda_fail(ONILDA_EINVAL, "run_em: null argument");
    const int avail = onilda_device_count();
    if (avail == 0) return onilda_fail(ONILDA_ECUDA, "run_em: no CUDA device is usable and this library has no CPU path");
    if (ndev <= 0 || ndev > avail) ndev = avail;
    if (ndev > kMaxPeers) ndev = kMaxPeers;
    if ((int64_t)ndev > c->num_docs) ndev = c->num_docs > 0 ? (int)c->num_docs : 1;

    // initialise the model (lda-estimate.c:147-167)
    onilda_model* model = nullptr;
    int rc;
    if (strcmp(start, "seeded") == 0) {
        rc = onilda_model_new(num_topics, c->num_terms, &model);
        if (rc == ONILDA_OK) rc = onilda_model_init_seeded(model, c, seed);
        if (rc == ONILDA_OK) model->alpha = initial_alpha;
    } else if (strcmp(start, "random") == 0) {
        rc = onilda_model_new(num_topics, c->num_terms, &model);
        if (rc == ONILDA_OK) rc = onilda_model_init_random(model, seed);
        if (rc == ONILDA_OK) model->alpha = initial_alpha;
    } else {
        rc = onilda_model_load(start, &model);
    }
    if (rc != ONILDA_OK) { onilda_model_free(model); return rc; }
    // A loaded model brings its own topic count (the reference saves gamma with model->num_topics too,
    // lda-estimate.c:313-317); everything below is sized by it, not by the argument.
    const int K = model->num_topics;

    mkdir(directory, S_IRUSR | S_IWUSR | S_IXUSR);   // utils.c:86-89, errors ignored as there
    const std::string dir(directory);
    rc = onilda_model_save(model, (dir + "/000").c_str());
    if (rc != ONILDA_OK) { onilda_model_free(model); return rc; }

    const bool timing = getenv("ONILDA_TIMING") != nullptr;
    const auto now = [] { return std::chrono::steady_clock::now(); };
    const auto secs = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) {
        return std::chrono::duration<double>(b - a).count();
    };
    const auto t_setup = now();
    EmSet set;
    std::vector<int64_t> bounds;
    split_by_nnz(c, ndev, bounds);
    for (int r = 0; r < ndev; r++) {
        onilda_em* em = nullptr;
        rc = onilda_em_create(c, bounds[(size_t)r], bounds[(size_t)r + 1], model, settings, r, &em);
        if (rc != ONILDA_OK) { onilda_model_free(model); return rc; }
        set.ems.push_back(em);
    }

    // several devices: NCCL reduce-scatter -> lda_mle of the own rows -> all-gather (em.cu), unless
    // ONILDA_ALLREDUCE=peer asks for the fixed-order peer-memory sum
    const char* env_ar = getenv("ONILDA_ALLREDUCE");
    const bool peer_sum = env_ar && strcmp(env_ar, "peer") == 0;
    if (ndev > 1 && !peer_sum) {
        rc = onilda_em_comm_attach_local(set.ems.data(), ndev);
        if (rc != ONILDA_OK) { onilda_model_free(model); return rc; }
    }
    FILE* lf = fopen((dir + "/likelihood.dat").c_str(), "w");
    if (!lf) { onilda_model_free(model); return onilda_fail(ONILDA_EIO, "cannot write %s/likelihood.dat", directory); }

    const auto t_em = now();
    // lda-estimate.c:186: the loop is entered at least once (i = 0 <= 2 and 0 <= EM_MAX_ITER for any sane setting)
    int keep = (0 <= settings->em_max_iter);
    int iter = 0;
    while (keep) {
        iter++;
        printf("**** em iteration %d ****\n", iter);
        for (onilda_em* em : set.ems) { rc = onilda_em_estep(em); if (rc != ONILDA_OK) break; }
        double lik = 0, conv = 0;
        if (rc == ONILDA_OK && peer_sum) {
            // the fixed-order alternative: sums over peer memory in rank order, then the whole M-step on every device
            if (ndev > 1) rc = onilda_em_allreduce_peer(set.ems.data(), ndev);
            if (rc == ONILDA_OK)
                for (onilda_em* em : set.ems) { rc = onilda_em_mstep(em, &lik, &conv, &keep); if (rc != ONILDA_OK) break; }
        } else if (rc == ONILDA_OK) {
            rc = onilda_em_mstep_group(set.ems.data(), ndev, &lik, &conv, &keep);
        }
        if (rc != ONILDA_OK) break;
        fprintf(lf, "%10.10f\t%5.5e\n", lik, conv);   // lda-estimate.c:266
        fflush(lf);
    }
    fclose(lf);
    if (rc != ONILDA_OK) { onilda_model_free(model); return rc; }

    // final model + gamma (lda-estimate.c:272-317); each shard owns its gamma rows, no gather traffic
    const auto t_save = now();
    rc = onilda_em_get_model(set.ems[0], model);
    if (rc == ONILDA_OK) rc = onilda_model_save(model, (dir + "/final").c_str());
    if (rc == ONILDA_OK) {
        std::vector<double> gamma((size_t)c->num_docs * K + 1);
        for (int r = 0; r < ndev && rc == ONILDA_OK; r++)
            rc = onilda_em_get_gamma(set.ems[(size_t)r], gamma.data() + (size_t)bounds[(size_t)r] * K);
        if (rc == ONILDA_OK) rc = onilda_gamma_save((dir + "/final.gamma").c_str(), gamma.data(), c->num_docs, K);
    }
    if (timing)
        fprintf(stderr, "[timing] device setup %.3f s, %d EM iterations %.3f s, final.* %.3f s\n", secs(t_setup, t_em), iter,
                secs(t_em, t_save), secs(t_save, now()));
    onilda_model_free(model);
    return rc;
}

extern "C" int onilda_infer_files(const char* model_root, const char* name, const onilda_corpus* c,
                                  const onilda_settings* settings, int device)
{
    if (!model_root || !name || !c || !settings) return onilda_fail(ONILDA_EINVAL, "infer_files: null argument");
    onilda_model* model = nullptr;
    int rc = onilda_model_load(model_root, &model);
    if (rc != ONILDA_OK) return rc;
    std::vector<double> gamma((size_t)c->num_docs * model->num_topics + 1), lik((size_t)c->num_docs + 1);
    rc = onilda_infer(model, c, settings, device, gamma.data(), lik.data());
    if (rc == ONILDA_OK) {
        const std::string base(name);
        FILE* f = fopen((base + "-lda-lhood.dat").c_str(), "w");   // lda-estimate.c:371-383
        if (!f) rc = onilda_fail(ONILDA_EIO, "cannot write %s-lda-lhood.dat", name);
        else {
            for (int64_t d = 0; d < c->num_docs; d++) fprintf(f, "%5.5f\n", lik[(size_t)d]);
            fclose(f);
            rc = onilda_gamma_save((base + "-gamma.dat").c_str(), gamma.data(), c->num_docs, model->num_topics);
        }
    }
    onilda_model_free(model);
    return rc;
}
