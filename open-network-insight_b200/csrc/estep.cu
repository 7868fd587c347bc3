// E-step of variational-EM LDA on sm_100a: the per-document fixed point
// (reference: oni-ml/oni-lda-c/lda-inference.c:27-85), its likelihood bound
// (:93-127) and the sufficient-statistics update of doc_e_step
// (lda-estimate.c:28-60), for every document of a shard in two concurrent
// launches of one kernel template (em.cu): the documents of at least 1024
// entries with G = 32, all others with G = 4 (K <= 32).
//
// Mapping
//   * a document is owned by a group of G lanes; lane j keeps topics
//     k = t*G + j in registers (gamma, E, and psi(gamma) as two parts).  A warp
//     therefore advances 32/G documents at once.  G = 4 is the efficient form
//     (no idle lanes at K = 20, 72 instructions per document-word); G = 32, one
//     topic per lane, is the low-latency form for long documents, whose serial
//     word chain would otherwise bound the whole E-step.
//   * words of a document are visited strictly in file order and gamma /
//     digamma are refreshed after EVERY word (lda-inference.c:66-74): the
//     Gauss-Seidel order of the reference is kept; parallelism is over topics
//     and over documents only.
//   * persistent warps pull "bundles" of 32/G consecutive documents of a
//     length-sorted order from an atomic queue (longest first), so the groups
//     of a warp run documents of near-equal length in lockstep and a warp that
//     finishes early steals the next bundle.  Each document leaves the sweep
//     loop on its own convergence test (in-kernel early exit).
//
// Arithmetic (all fp64).  With E_k = exp(psi(gamma_k) - shift) and the table
// beta'[w][k] = exp(log_beta[k][w] - lbmax[w]) built once per EM iteration,
//     phi_nk = E_k beta'_kw / s_n,   s_n = sum_j E_j beta'_jw
// which is the reference's exp(psi_k + log_beta_kw - log_sum) without a
// log_sum/exp per (word, topic): one log per WORD (log s_n) remains.  Per (word,
// topic) the update costs two reciprocals, one exp and one log (device_math.cuh).  shift is
// psi(alpha + total/K), the value every topic starts from, so E starts at 1 and
// s_n can neither overflow nor vanish in practice; if it ever leaves
// [1e-290, 1e290] that word is redone in log space (the reference's own form).
// The bound needs log phi_nk = psi_k(at word n) + log_beta_kw - lse_n; inserting
// it and gamma_k = alpha + sum_n c_n phi_nk into lda-inference.c:106-125 the
// (psi_k - psi_sum) terms cancel exactly and
//     L = lgamma(alpha K) - K lgamma(alpha) - lgamma(sum gamma) + sum_k lgamma(gamma_k)
//         - sum_n sum_k c_n phi_nk psi_k(at word n) + sum_n c_n lse_n .
// tests/ check it against the oracle to ~1e-13 relative per document.
#include "device_math.cuh"
#include "kernels.cuh"

#ifndef ONILDA_PHI_AHEAD
#define ONILDA_PHI_AHEAD 8
#endif

namespace onilda {

// Sum / max over the G lanes of a group.  Always executed by the whole warp with the full
// mask (the kernel keeps its warps converged in the hot loop; partial-mask shuffles cost a
// MATCH/VOTE/branch sequence each).
template <int G>
__device__ __forceinline__ double group_sum(double v)
{
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

template <int G>
__device__ __forceinline__ double group_max(double v)
{
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

template <int TPL>
struct WordItem {
    int word;            // -1: this group idles at this word
    double count;
    double lbmax;
    double beta[TPL];
    double old_phi[TPL];
};

template <int G, int TPL>
__global__ void __launch_bounds__(128) estep_kernel(EStepArgs a, unsigned long long num_bundles)
{
    constexpr int DPW = 32 / G;      // documents per warp
    const int lane = threadIdx.x & 31;
    const int grp = lane / G;
    const int j = lane % G;
    const int K = a.num_topics;
    const int KS = a.kstride;
    const double inv_K = 1.0 / K;
    bool valid[TPL];
#pragma unroll
    for (int t = 0; t < TPL; t++) valid[t] = (t * G + j) < K;

    const double* __restrict__ beta_scaled = a.beta_scaled;
    double* __restrict__ phi = a.phi;

    // Where several warps share an SM sub-partition (G < 32) the two logarithms per word go through a
    // 128-entry table in shared memory (10 FP64 instructions instead of 19).  The one-document-per-warp
    // instantiation keeps the table-free form: alone on its sub-partition it is bound by the latency of
    // its own instruction stream, and the table's shared-memory round trip costs it more than the
    // nine instructions (measured: 155 -> 186 ms per E-step at cfg 2 with the table).
    constexpr bool kTableLog = (G < 32);
    __shared__ LogTabEntry logtab[kTableLog ? 128 : 1];
    if constexpr (kTableLog) {
        for (int i = threadIdx.x; i < 128; i += blockDim.x) logtab[i] = kLogTab[i];
        __syncthreads();
    }
    auto log_of = [&](double v) -> double {
        if constexpr (kTableLog) return fast_log_tab(v, logtab);
        else return fast_log(v);
    };

    for (;;) {
        // Warp-uniform values are produced with redux/vote (their results live in uniform
        // registers), so the compiler can prove the warp converged at every shuffle below and
        // emits them without a divergence branch (BRA.DIV), which would otherwise fence its
        // instruction scheduling around each one.
        unsigned int claimed = 0;
        if (lane == 0) claimed = (unsigned int)atomicAdd(a.queue, 1ull);
        const unsigned long long bundle = __reduce_max_sync(0xffffffffu, claimed);
        if (bundle >= num_bundles) break;

        const int64_t slot = (int64_t)bundle * DPW + grp;
        const bool has_doc = slot < a.order_count;
        const int64_t d = has_doc ? a.order[a.order_begin + slot] : 0;
        const int64_t beg = has_doc ? a.offsets[d] : 0;
        const int len = has_doc ? (int)(a.offsets[d + 1] - beg) : 0;
        const int total = has_doc ? a.totals[d] : 0;
        const int max_len = __reduce_max_sync(0xffffffffu, len);

        // posterior Dirichlet start (lda-inference.c:37-43)
        const double gamma0 = a.alpha + (total / ((double)K));
        const DigammaParts d0 = digamma_rational(gamma0);
        const double shift = d0.rational + fast_log(d0.y);
        // per topic: gamma, E = exp(psi(gamma) - shift), and psi(gamma) kept as its two parts
        // (rational part, y = gamma + 6) so that the logarithm can be taken one word later
        double gam[TPL], E[TPL], rat[TPL], yy[TPL];
#pragma unroll
        for (int t = 0; t < TPL; t++) {
            gam[t] = valid[t] ? gamma0 : 1.0;
            E[t] = 1.0;
            rat[t] = d0.rational;
            yy[t] = d0.y;
        }

        double lik = 0.0, lik_old = 0.0, converged = 1.0;
        int sweeps = 0;
        bool active = has_doc && (converged > a.var_converged) && ((0 < a.var_max_iter) || (a.var_max_iter == -1));

        while (__any_sync(0xffffffffu, active)) {
            if (active) sweeps++;
            const bool first = (sweeps == 1);   // same for every active group of the warp
            double acc_lse = 0.0, count_sum = 0.0;
            double acc_psi[TPL];
#pragma unroll
            for (int t = 0; t < TPL; t++) acc_psi[t] = 0.0;

            // Two-stage software pipeline over the words: ids (word, count) are fetched two words
            // ahead, the rows they select (beta', old phi, lbmax) one word ahead, so neither load
            // latency nor the id -> address dependency sits on the serial update chain.
            // A group that is idle at word n (document shorter than the bundle's longest, or already
            // converged) still runs the arithmetic, on count = 0 and beta' = 1: every update is then
            // multiplied by zero and nothing of it is stored.  That keeps the warp converged.
            constexpr int kPhiAhead = ONILDA_PHI_AHEAD;
            auto load_ids = [&](int n, int& w, int& c) {
                const bool on = active && n < len;
                w = on ? a.words[beg + n] : -1;
                c = on ? a.counts[beg + n] : 0;
            };
            auto load_rows = [&](int n, int w, int c, WordItem<TPL>& it) {
                const bool on = w >= 0;
                it.word = w;
                it.count = (double)c;
                it.lbmax = on ? a.lbmax[w] : 0.0;
                const double* brow = beta_scaled + (int64_t)w * KS + j;
                const double* prow = phi + (beg + n) * KS + j;
                // phi is the one stream that comes from HBM (17 GB at cfg 2, re-read every sweep); ask L2 for
                // the row kPhiAhead words ahead so that the load one word ahead finds it there
                if (kPhiAhead > 0 && !first && active && n + kPhiAhead < len)
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(prow + (int64_t)kPhiAhead * KS));
#pragma unroll
                for (int t = 0; t < TPL; t++) {
                    it.beta[t] = (on && valid[t]) ? __ldg(brow + t * G) : (valid[t] ? 1.0 : 0.0);
                    it.old_phi[t] = (first || !on || !valid[t]) ? (valid[t] ? inv_K : 0.0) : prow[t * G];
                }
            };

            // word n's log(s_n) is also taken one word later (s_prev, with its weight and lbmax)
            double s_prev = 1.0, c_prev = 0.0, lb_prev = 0.0;

            // one word of one sweep for every group of the warp
            auto step = [&](const WordItem<TPL>& it, int n) {
                const bool on = it.word >= 0;

                // The recurrence E -> s -> phi -> gamma -> E (lda-inference.c:52-73) starts here ...
                double x[TPL], part = 0.0;
#pragma unroll
                for (int t = 0; t < TPL; t++) {
                    x[t] = E[t] * it.beta[t];
                    part += x[t];
                }
                const double s = group_sum<G>(part);

                // ... and, off the recurrence, the two logarithms left over from the previous word.
                // They depend only on values carried into this iteration, so they fill the latency
                // gaps of the serial chain (the shuffles above, the reciprocals below).
                double psi[TPL];
#pragma unroll
                for (int t = 0; t < TPL; t++) psi[t] = rat[t] + log_of(yy[t]);
                acc_lse = fma(c_prev, lb_prev + log_of(s_prev), acc_lse);
                const bool bad = !(s > 1e-290 && s < 1e290);
                double inv = fast_rcp(s);
                s_prev = s;
                c_prev = it.count;
                lb_prev = it.lbmax;
                if (__any_sync(0xffffffffu, bad)) {
                    // rare: redo this word in log space, as the reference does (lda-inference.c:52-68)
                    double av[TPL], m = -INFINITY;
#pragma unroll
                    for (int t = 0; t < TPL; t++) {
                        av[t] = (valid[t] && it.beta[t] > 0.0) ? (psi[t] - shift) + log(it.beta[t]) : -INFINITY;
                        m = fmax(m, av[t]);
                    }
                    m = group_max<G>(m);
                    double y[TPL], ps = 0.0;
#pragma unroll
                    for (int t = 0; t < TPL; t++) {
                        y[t] = (av[t] > -INFINITY) ? exp(av[t] - m) : 0.0;
                        ps += y[t];
                    }
                    ps = group_sum<G>(ps);
                    if (bad) {
#pragma unroll
                        for (int t = 0; t < TPL; t++) x[t] = y[t];
                        inv = 1.0 / ps;
                        acc_lse = fma(it.count, it.lbmax + m + log(ps), acc_lse);
                        c_prev = 0.0;      // this word's log-sum is already in
                        s_prev = 1.0;
                    }
                }
                double* prow = phi + (beg + n) * KS + j;
#pragma unroll
                for (int t = 0; t < TPL; t++) {
                    const double p = x[t] * inv;
                    if (on && valid[t]) prow[t * G] = p;
                    acc_psi[t] = fma(it.count * p, psi[t], acc_psi[t]);
                    gam[t] = gam[t] + it.count * (p - it.old_phi[t]);
                    const DigammaParts dp = digamma_rational(gam[t]);
                    // E = exp(psi - shift) = y * exp(rational - shift): no logarithm on the chain.
                    // Padding topics need no mask here: their beta' is 0, so their E never counts.
                    E[t] = dp.y * fast_exp(dp.rational - shift);
                    rat[t] = dp.rational;
                    yy[t] = dp.y;
                }
                count_sum += it.count;
            };

            // The loop is unrolled by two over a pair of row buffers (A, B) so that prefetched rows
            // are never copied: a register move of a value still in flight would wait for the load.
            // Rows for word n+1 are requested before word n is worked on and first touched a whole
            // step later; ids run one word further ahead.
            // (The one-document-per-warp instantiation keeps a single-step loop: it is bound by the
            // latency of its own instruction stream, and doubling the loop body costs it more in
            // instruction fetch than the copies do.)
            int wa, ca, wb, cb;
            WordItem<TPL> A, B;
            load_ids(0, wa, ca);
            load_ids(1, wb, cb);
            load_rows(0, wa, ca, A);
            if constexpr (G < 32) {
                for (int n = 0; n < max_len; n += 2) {
                    load_rows(n + 1, wb, cb, B);
                    load_ids(n + 2, wa, ca);
                    step(A, n);
                    load_rows(n + 2, wa, ca, A);
                    load_ids(n + 3, wb, cb);
                    step(B, n + 1);
                }
            } else {
                for (int n = 0; n < max_len; n++) {   // wb, cb hold the ids of word n + 1
                    load_rows(n + 1, wb, cb, B);
                    load_ids(n + 2, wb, cb);
                    step(A, n);
                    A = B;
                }
            }
            acc_lse = fma(c_prev, lb_prev + log_of(s_prev), acc_lse);

            // likelihood bound of this sweep (lda-inference.c:93-127, see header); computed by every
            // group, used by the active ones
            double gs = 0.0, lg = 0.0, ap = 0.0;
#pragma unroll
            for (int t = 0; t < TPL; t++) {
                if (valid[t]) {
                    gs += gam[t];
                    lg += lgamma(gam[t]);
                }
                ap += acc_psi[t];
            }
            gs = group_sum<G>(gs);
            lg = group_sum<G>(lg);
            ap = group_sum<G>(ap);
            if (active) {
                lik = a.lik_const - lgamma(gs) + lg - ap + (acc_lse + shift * count_sum);
                if (isnan(lik)) *a.error_flag = 1;   // the reference asserts here (lda-inference.c:78)
                converged = (lik_old - lik) / lik_old;
                lik_old = lik;
                active = (converged > a.var_converged) && ((sweeps < a.var_max_iter) || (a.var_max_iter == -1));
            }
        }

        // doc_e_step (lda-estimate.c:40-46): alpha_suffstats += sum_k psi(gamma_k) - K psi(sum gamma)
        double gs = 0.0, ps = 0.0;
#pragma unroll
        for (int t = 0; t < TPL; t++) {
            if (valid[t]) {
                gs += gam[t];
                ps += rat[t] + fast_log(yy[t]);
                if (has_doc) a.gamma[d * K + t * G + j] = gam[t];
            }
        }
        gs = group_sum<G>(gs);
        ps = group_sum<G>(ps);
        if (has_doc && j == 0) {
            a.doc_lik[d] = lik;
            a.doc_alpha_ss[d] = ps - K * digamma(gs);
            a.doc_sweeps[d] = sweeps;
        }
        // class_word[k][w] += c_n phi_nk (lda-estimate.c:48-55); class_total is the column sum (mstep.cu)
        if (a.class_word != nullptr) {
            // four words per trip, all loads issued before the first atomic: the pass is pure
            // memory latency otherwise (phi rows come back from L2/HBM)
            constexpr int U = 4;
            for (int n0 = 0; n0 < len; n0 += U) {
                int w[U];
                double c[U], p[U][TPL];
#pragma unroll
                for (int u = 0; u < U; u++) {
                    const bool in = n0 + u < len;
                    const int64_t e = beg + n0 + u;
                    w[u] = in ? a.words[e] : -1;
                    c[u] = in ? (double)a.counts[e] : 0.0;
                    const double* prow = phi + e * KS + j;
#pragma unroll
                    for (int t = 0; t < TPL; t++)
                        p[u][t] = (in && valid[t] && sweeps != 0) ? prow[t * G] : inv_K;
                }
#pragma unroll
                for (int u = 0; u < U; u++) {
                    if (w[u] >= 0) {
                        double* crow = a.class_word + (int64_t)w[u] * KS + j;
#pragma unroll
                        for (int t = 0; t < TPL; t++)
                            if (valid[t]) atomicAdd(crow + t * G, c[u] * p[u][t]);
                    }
                }
            }
        }
    }
}

// Grid for `num_bundles` bundles of one (G, TPL) instantiation: persistent, at most the resident
// capacity of `sm_count` SMs (or `max_blocks` when the launch shares the device with another).
template <int G, int TPL>
static int launch_one(const EStepArgs& args, int sm_count, int block_threads, int reserve_smem, int max_per_sm,
                      cudaStream_t stream)
{
    constexpr int DPW = 32 / G;
    const unsigned long long num_bundles = (unsigned long long)((args.order_count + DPW - 1) / DPW);
    if (num_bundles == 0) return ONILDA_OK;
    // reserve_smem: dynamic shared memory the kernel never touches, asked for only to cap how many
    // of these blocks an SM takes (em.cu sets it for the long-document launch: two warps per sub-partition)
    if (reserve_smem > 0)
        ONILDA_CUDA(cudaFuncSetAttribute(estep_kernel<G, TPL>, cudaFuncAttributeMaxDynamicSharedMemorySize, reserve_smem));
    int per_sm = 0;
    ONILDA_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, estep_kernel<G, TPL>, block_threads, (size_t)reserve_smem));
    if (per_sm < 1) per_sm = 1;
    if (max_per_sm > 0 && per_sm > max_per_sm) per_sm = max_per_sm;
    const unsigned long long wpb = (unsigned long long)(block_threads / 32);
    unsigned long long want = (num_bundles + wpb - 1) / wpb;
    unsigned long long grid = (unsigned long long)sm_count * per_sm;
    if (grid > want) grid = want;
    estep_kernel<G, TPL><<<(unsigned)grid, block_threads, (size_t)reserve_smem, stream>>>(args, num_bundles);
    ONILDA_CUDA(cudaGetLastError());
    return ONILDA_OK;
}

template <int G>
static int launch_group(int tpl, const EStepArgs& args, int sm_count, int block_threads, int reserve_smem, int max_per_sm,
                        cudaStream_t stream)
{
    switch (tpl) {
        case 1: return launch_one<G, 1>(args, sm_count, block_threads, reserve_smem, max_per_sm, stream);
        case 2: return launch_one<G, 2>(args, sm_count, block_threads, reserve_smem, max_per_sm, stream);
        case 3: return launch_one<G, 3>(args, sm_count, block_threads, reserve_smem, max_per_sm, stream);
        case 4: return launch_one<G, 4>(args, sm_count, block_threads, reserve_smem, max_per_sm, stream);
        case 5: return launch_one<G, 5>(args, sm_count, block_threads, reserve_smem, max_per_sm, stream);
        case 6: return launch_one<G, 6>(args, sm_count, block_threads, reserve_smem, max_per_sm, stream);
        case 7: return launch_one<G, 7>(args, sm_count, block_threads, reserve_smem, max_per_sm, stream);
        case 8: return launch_one<G, 8>(args, sm_count, block_threads, reserve_smem, max_per_sm, stream);
    }
    return onilda_fail(ONILDA_EINVAL, "unsupported topics-per-lane %d", tpl);
}

// One launch over documents order[order_begin, order_begin + order_count) with `group` lanes per
// document and `block_threads` (32..128) threads per block.  args.queue must be zero (the caller
// clears it on the same stream).
int launch_estep(int group, const EStepArgs& args, int sm_count, int block_threads, int reserve_smem, int max_per_sm,
                 cudaStream_t stream, int* launches)
{
    if (args.order_count <= 0) return ONILDA_OK;
    const int tpl = (args.num_topics + group - 1) / group;
    int rc;
    switch (group) {
        case 4: rc = launch_group<4>(tpl, args, sm_count, block_threads, reserve_smem, max_per_sm, stream); break;
        case 8: rc = launch_group<8>(tpl, args, sm_count, block_threads, reserve_smem, max_per_sm, stream); break;
        case 16: rc = launch_group<16>(tpl, args, sm_count, block_threads, reserve_smem, max_per_sm, stream); break;
        case 32: rc = launch_group<32>(tpl, args, sm_count, block_threads, reserve_smem, max_per_sm, stream); break;
        default: return onilda_fail(ONILDA_EINVAL, "unsupported lane group %d", group);
    }
    if (rc == ONILDA_OK && launches) *launches += 1;
    return rc;
}

}  // namespace onilda
