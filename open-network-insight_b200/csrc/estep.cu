// E-step of variational-EM LDA on sm_100a: the per-document fixed point
// (reference: oni-ml/oni-lda-c/lda-inference.c:27-85), its likelihood bound
// (:93-127) and the sufficient-statistics update of doc_e_step
// (lda-estimate.c:28-60), for every document of a shard in two concurrent
// launches of one kernel template (em.cu): the documents of at least 1024
// entries with G = 32, all others with G = 4 (K <= 32).
//
// Mapping
//   * a document is owned by a group of G lanes; lane j keeps topics
//     k = t*G + j in registers (gamma, psi(gamma), E = exp(psi(gamma))).  A warp
//     therefore advances 32/G documents at once.  G = 4 is the efficient form
//     (no idle lanes at K = 20); G = 32, one topic per lane, is the low-latency
//     form for long documents, whose serial word chain would otherwise bound
//     the whole E-step.
//   * words of a document are visited strictly in file order and gamma /
//     digamma are refreshed after EVERY word (lda-inference.c:66-74): the
//     Gauss-Seidel order of the reference is kept; parallelism is over topics
//     and over documents only.
//   * persistent warps take "bundles" of 32/G consecutive documents of a
//     length-sorted order from an atomic queue (longest first), so the groups of a
//     warp run documents of near-equal length in lockstep and a warp that
//     finishes early steals the next bundle.  Each document leaves the sweep
//     loop on its own convergence test (in-kernel early exit).
//
// Arithmetic (all fp64).  With E_k = exp(psi(gamma_k)) and the table
// beta'[w][k] = exp(log_beta[k][w] - lbmax[w]) built once per EM iteration,
//     phi_nk = E_k beta'_kw / s_n,   s_n = sum_j E_j beta'_jw
// which is the reference's exp(psi_k + log_beta_kw - log_sum) without a
// log_sum/exp per (word, topic): one log per WORD (log s_n) remains.  psi and
// exp(psi) come from shared-memory tables (device_math.cuh: psi = R(gamma) - 1/gamma
// with R a piecewise degree-5 fit of the reference's own series, exp with a
// 64-entry table of 2^(i/64)): 28 FP64 instructions per (word, topic) where the
// closed forms took 68.  If s_n ever leaves [1e-290, 1e290] that word is redone
// in log space (the reference's own form); a gamma beyond the table (>= 2^20 - 1)
// takes the closed forms.
// The bound needs log phi_nk = psi_k(at word n) + log_beta_kw - lse_n; inserting
// it and gamma_k = alpha + sum_n c_n phi_nk into lda-inference.c:106-125 the
// (psi_k - psi_sum) terms cancel exactly and
//     L = lgamma(alpha K) - K lgamma(alpha) - lgamma(sum gamma) + sum_k lgamma(gamma_k)
//         - sum_n sum_k c_n phi_nk psi_k(at word n) + sum_n c_n lse_n .
// tests/ check it against the oracle to ~1e-13 relative per document.
#include "estep_common.cuh"

#ifndef ONILDA_PHI_AHEAD
#define ONILDA_PHI_AHEAD 8
#endif
#ifndef ONILDA_LONG_NB
#define ONILDA_LONG_NB 3
#endif
#ifndef ONILDA_PIN
#define ONILDA_PIN 1
#endif
#ifndef ONILDA_ESTEP_MAX_THREADS
#define ONILDA_ESTEP_MAX_THREADS 512
#endif

namespace onilda {

// KC: the number of topics when known at compile time (0: taken from the arguments).  With it the
// row stride, the padding tests and every row address are constants or one multiply-add.
// kRare: the kernel carries its two rare paths - the closed forms of psi / exp(psi) for a bundle whose gamma may leave
// the tables (alpha < 1e-8 or alpha + tokens >= 2^20 - 2) and the log-space redo of a word whose sum s leaves
// [2^-963, 2^963).  Both are decided on the host per launch (em.cu): s >= min_k exp(psi(gamma_k)) >= exp(psi(alpha))
// because the largest beta' of a word is exactly 1, so for alpha >= 0.0016 (psi > -626, exp > 2^-903) no sum can
// leave the range, and the documents' token counts are known.  Without them the word loop has no call, no taken
// branch and no spills (2.6 % of the E-step at cfg 2).
template <int G, int TPL, int KC, bool kRare>
__global__ void __launch_bounds__(ONILDA_ESTEP_MAX_THREADS) estep_kernel(EStepArgs a, unsigned long long num_bundles)
{
    constexpr int DPW = 32 / G;      // documents per warp
    // The one-document-per-warp form waits on its own dependency chain: shallow (Estrin) polynomials.
    // The others are bound by FP64 issue: Horner, fewest instructions.
    constexpr bool kShallow = (G == 32);
    const int lane = threadIdx.x & 31;
    const int grp = lane / G;
    const int j = lane % G;
    const int K = KC ? KC : a.num_topics;
    const uint32_t KS = KC ? (uint32_t)kpad_for(KC) : (uint32_t)a.kstride;
    const int32_t idle_word = a.idle_row;      // beta' = 1, lbmax = 0
    const double inv_K = 1.0 / K;
    bool valid[TPL];
    double start_phi[TPL];                     // phi before the first sweep: 1/K (lda-inference.c:41)
#pragma unroll
    for (int t = 0; t < TPL; t++) {
        valid[t] = (t * G + j) < K;
        start_phi[t] = valid[t] ? inv_K : 0.0;
    }

    // kernel parameters used in the word loop are pinned in registers the same way (a constant-bank read
    // has a scoreboard latency that the serial chain would wait for)
    const double* __restrict__ beta_scaled = a.beta_scaled + j;
    const double* __restrict__ lbmax = a.lbmax;
#if ONILDA_PIN
    asm volatile("" : "+l"(beta_scaled), "+l"(lbmax));
#endif

    // Tables in shared memory, one copy per block: lanes index them divergently (constant memory
    // would serialise, L1 would fetch a sector per lane).
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ LogTabEntry logtab[128];
    {
        const int4* src = reinterpret_cast<const int4*>(a.psi_tab);
        int4* dst = reinterpret_cast<int4*>(smem_raw);
        constexpr int chunks = (int)((sizeof(PsiRow) * kPsiRows + sizeof(double) * kExpTabSize) / sizeof(int4));
        for (int i = threadIdx.x; i < chunks; i += blockDim.x) dst[i] = src[i];     // psi rows, then 2^(i/64)
        for (int i = threadIdx.x; i < 128; i += blockDim.x) logtab[i] = kLogTab[i];
        __syncthreads();
    }
    // The shared-window addresses are taken once and pinned in registers (volatile asm): the compiler would
    // otherwise re-derive the window base from a special register at every use, a ~15-cycle read on the chain.
    // a launch that must share the SMs with this one may be waiting for every block of this grid to be resident
    // (programmatic dependent launch, em.cu); a no-op otherwise
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    SmemTables tb;
    tb.psi = (uint32_t)__cvta_generic_to_shared(smem_raw);
    tb.log = (uint32_t)__cvta_generic_to_shared(logtab);
#if ONILDA_PIN
    asm volatile("" : "+r"(tb.psi), "+r"(tb.log));
#endif
    tb.exp2 = tb.psi + (uint32_t)(sizeof(PsiRow) * kPsiRows);

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
        const int32_t* __restrict__ doc_words = a.words + beg;
        const int32_t* __restrict__ doc_counts = a.counts + beg;
        double* __restrict__ doc_phi = a.phi + beg * (int64_t)KS + j;

        // Every gamma of a document stays within [alpha, alpha + total] up to rounding.  Where that
        // interval leaves the tables' range the whole bundle takes the closed forms (warp-uniform).
        const bool closed_forms = kRare && __any_sync(0xffffffffu, !(a.alpha >= kPsiTableMin && a.alpha + total + 1.0 < kPsiTableMax));
        auto psi_exp = [&](double (&g)[TPL], double (&psi)[TPL], double (&E)[TPL]) {
            if (__builtin_expect(closed_forms, 0)) {
#pragma unroll
                for (int t = 0; t < TPL; t++) {
                    const PsiExp pe = psi_exp_closed(g[t]);
                    psi[t] = pe.psi;
                    E[t] = pe.e;
                }
            } else {
#pragma unroll
                for (int t = 0; t < TPL; t++) {
                    const PsiExp pe = psi_exp_tab<kShallow>(g[t], tb);
                    psi[t] = pe.psi;
                    E[t] = pe.e;
                }
            }
        };

        // posterior Dirichlet start (lda-inference.c:37-43); per topic: gamma, psi(gamma), E = exp(psi(gamma))
        const double gamma0 = a.alpha + (total / ((double)K));
        double gam[TPL], psi[TPL], E[TPL];
#pragma unroll
        // a group without a document idles on gamma = 1: with its total of 0 it would start at alpha, and for
        // alpha below ~0.0015 exp(digamma(alpha)) underflows, which would send every word of the warp down the rare path
        for (int t = 0; t < TPL; t++) gam[t] = (valid[t] && has_doc) ? gamma0 : 1.0;
        psi_exp(gam, psi, E);

        double lik = 0.0, lik_old = 0.0, converged = 1.0;
        int sweeps = 0;
        bool active = has_doc && (converged > a.var_converged) && ((0 < a.var_max_iter) || (a.var_max_iter == -1));

        while (__any_sync(0xffffffffu, active)) {
            if (active) sweeps++;
            const bool first = (sweeps == 1);   // same for every active group of the warp
            double acc_lse = 0.0;
            double acc_psi[TPL];
#pragma unroll
            for (int t = 0; t < TPL; t++) acc_psi[t] = 0.0;

            // Two-stage software pipeline over the words: ids (word, count) are fetched two words
            // ahead, the rows they select (beta', old phi, lbmax) one word ahead, so neither load
            // latency nor the id -> address dependency sits on the serial update chain.
            // A group that is idle at word n (document shorter than the bundle's longest, or already
            // converged) still runs the arithmetic, on count = 0 and the table's spare row (beta' = 1):
            // every update is then multiplied by zero and nothing of it is stored.  That keeps the
            // warp converged.
            constexpr int kPhiAhead = ONILDA_PHI_AHEAD;
            const int live_len = active ? len : 0;
            auto load_ids = [&](int n, int& w, int& c) {
                w = idle_word;
                c = 0;
                if (n < live_len) {
                    w = doc_words[n];
                    c = doc_counts[n];
                }
            };
            auto load_rows = [&](int n, int w, int c, WordItem<TPL>& it) {
                it.word = w;
                it.count = (double)c;
                it.lbmax = __ldg(lbmax + w);
                const double* brow = beta_scaled + (uint64_t)(uint32_t)w * KS;
                const double* prow = doc_phi + (uint64_t)(uint32_t)n * KS;
                const bool reread = !first && w != idle_word;
                // phi is the one stream that comes from HBM (17 GB at cfg 2, re-read every sweep); ask L2 for
                // the row kPhiAhead words ahead so that the load one word ahead finds it there
                if (kPhiAhead > 0 && !first && valid[0] && n + kPhiAhead < live_len)
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(prow + (uint64_t)kPhiAhead * KS));
#pragma unroll
                for (int t = 0; t < TPL; t++) {
                    it.beta[t] = 0.0;
                    if (valid[t]) it.beta[t] = __ldg(brow + t * G);
                    it.old_phi[t] = start_phi[t];
                    if (reread && valid[t]) it.old_phi[t] = prow[t * G];
                }
            };

            // word n's log(s_n) is taken one word later (s_prev, with its weight and lbmax): it depends only
            // on values carried into the iteration, so it fills the latency gaps of the serial chain
            double s_prev = 1.0, c_prev = 0.0, lb_prev = 0.0;

            // one word of one sweep for every group of the warp
            auto step = [&](const WordItem<TPL>& it, int n) {
                // the recurrence E -> s -> phi -> gamma -> E (lda-inference.c:52-73)
                double x[TPL], part = 0.0;
#pragma unroll
                for (int t = 0; t < TPL; t++) {
                    x[t] = E[t] * it.beta[t];
                    part += x[t];
                }
                // off the chain: gamma - c*old_phi, so that one fma remains after phi is known
                double gpre[TPL];
#pragma unroll
                for (int t = 0; t < TPL; t++) gpre[t] = fma(-it.count, it.old_phi[t], gam[t]);
                const double s = chain_sum<G>(part);
                acc_lse = fma(c_prev, lb_prev + fast_log_tab(s_prev, tb), acc_lse);
                // s outside [2^-963, 2^963) (about 1e-290 .. 1e290), zero, negative or NaN: an integer test
                // on the exponent field, off the FP64 pipe
                const bool bad = (uint32_t)(double_hi(s) - 0x03C00000) >= (uint32_t)(0x7C200000 - 0x03C00000);
                // phi = x / s as x*r + (x*r)*e with r the hardware seed (2^-20) and e = d + d*d, d = 1 - s*r:
                // the cubic step of fast_rcp folded into the product, three dependent operations after the seed
                double r0 = rcp_seed(s);
                double e = fma(-s, r0, 1.0);
                e = fma(e, e, e);
                s_prev = s;
                c_prev = it.count;
                lb_prev = it.lbmax;
                if (kRare && __builtin_expect(__any_sync(0xffffffffu, bad), 0)) {
                    // rare: redo this word in log space, as the reference does (lda-inference.c:52-68); out of
                    // line, through a scratch block that only this branch touches
                    if (bad && j == 0) atomicAdd(a.error_flag + 1, 1);
                    LogSpaceWord<TPL> ls;
#pragma unroll
                    for (int t = 0; t < TPL; t++) {
                        ls.psi[t] = psi[t];
                        ls.beta[t] = valid[t] ? it.beta[t] : 0.0;
                    }
                    redo_word_in_log_space<G, TPL>(&ls);
                    if (bad) {
#pragma unroll
                        for (int t = 0; t < TPL; t++) x[t] = ls.y[t];
                        r0 = ls.inv_sum;
                        e = 0.0;
                        acc_lse = fma(it.count, it.lbmax + ls.lse, acc_lse);
                        c_prev = 0.0;      // this word's log-sum is already in
                        s_prev = 1.0;
                    }
                }
                double* prow = doc_phi + (uint64_t)(uint32_t)n * KS;
                const bool keep = it.word != idle_word;
#pragma unroll
                for (int t = 0; t < TPL; t++) {
                    const double xr = x[t] * r0;
                    const double p = fma(xr, e, xr);
                    gam[t] = fma(it.count, p, gpre[t]);
                    if (keep && valid[t]) prow[t * G] = p;
                    acc_psi[t] = fma(it.count * p, psi[t], acc_psi[t]);
                }
                // Padding topics need no mask: their beta' is 0, so their E never counts.
                psi_exp(gam, psi, E);
            };

            // The loop is unrolled over a ring of NB row buffers so that prefetched rows are never copied (a
            // register move of a value still in flight would wait for the load).  Rows for word n + NB - 1 are
            // requested before word n is worked on; ids run one word further ahead.  NB = 2 where several
            // documents share the warp (a step is long and registers are short), 3 for the one-document-per-warp
            // form, whose step is shorter than a round trip to L2 under load.
            constexpr int NB = (G == 32) ? ONILDA_LONG_NB : 2;
            WordItem<TPL> ring[NB];
            int wn, cn;
#pragma unroll
            for (int b = 0; b < NB - 1; b++) {
                load_ids(b, wn, cn);
                load_rows(b, wn, cn, ring[b]);
            }
            load_ids(NB - 1, wn, cn);
            for (int n = 0; n < max_len; n += NB) {
#pragma unroll
                for (int b = 0; b < NB; b++) {
                    load_rows(n + b + NB - 1, wn, cn, ring[(b + NB - 1) % NB]);
                    load_ids(n + b + NB, wn, cn);
                    step(ring[b], n + b);
                }
            }
            acc_lse = fma(c_prev, lb_prev + fast_log_tab(s_prev, tb), acc_lse);

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
                lik = a.lik_const - lgamma(gs) + lg - ap + acc_lse;
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
                ps += psi[t];
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
            double* __restrict__ class_word = a.class_word + j;
            for (int n0 = 0; n0 < len; n0 += U) {
                int w[U];
                double c[U], p[U][TPL];
#pragma unroll
                for (int u = 0; u < U; u++) {
                    const bool in = n0 + u < len;
                    w[u] = in ? doc_words[n0 + u] : -1;
                    c[u] = in ? (double)doc_counts[n0 + u] : 0.0;
                    const double* prow = doc_phi + (uint64_t)(uint32_t)(n0 + u) * KS;
#pragma unroll
                    for (int t = 0; t < TPL; t++)
                        p[u][t] = (in && valid[t] && sweeps != 0) ? prow[t * G] : inv_K;
                }
#pragma unroll
                for (int u = 0; u < U; u++) {
                    if (w[u] >= 0) {
                        double* crow = class_word + (uint64_t)(uint32_t)w[u] * KS;
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
template <int G, int TPL, int KC, bool kRare = false>
static int launch_one(const EStepArgs& args, int sm_count, int block_threads, int extra_smem, int max_per_sm,
                      cudaStream_t stream, bool after_resident)
{
    constexpr int DPW = 32 / G;
    const unsigned long long num_bundles = (unsigned long long)((args.order_count + DPW - 1) / DPW);
    if (num_bundles == 0) return ONILDA_OK;
    // extra_smem: dynamic shared memory beyond the tables that the kernel never touches, asked for only to
    // cap how many of these blocks an SM takes
    if (block_threads > ONILDA_ESTEP_MAX_THREADS || block_threads < 32 || (block_threads & 31))
        return onilda_fail(ONILDA_EINVAL, "E-step block of %d threads (32..%d, whole warps)", block_threads, ONILDA_ESTEP_MAX_THREADS);
    const size_t smem = kEstepTableBytes + (size_t)extra_smem;
    ONILDA_CUDA(cudaFuncSetAttribute(estep_kernel<G, TPL, KC, kRare>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    if (args.carveout_pct > 0)
        ONILDA_CUDA(cudaFuncSetAttribute(estep_kernel<G, TPL, KC, kRare>, cudaFuncAttributePreferredSharedMemoryCarveout, args.carveout_pct));
    int per_sm = 0;
    ONILDA_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, estep_kernel<G, TPL, KC, kRare>, block_threads, smem));
    if (per_sm < 1) per_sm = 1;
    if (max_per_sm > 0 && per_sm > max_per_sm) per_sm = max_per_sm;
    const unsigned long long wpb = (unsigned long long)(block_threads / 32);
    unsigned long long want = (num_bundles + wpb - 1) / wpb;
    unsigned long long grid = (unsigned long long)sm_count * per_sm;
    if (grid > want) grid = want;
    if (after_resident) {
        // secondary of a programmatic dependent launch: its blocks are scheduled once every block of the kernel before
        // it in this stream has executed griddepcontrol.launch_dependents, i.e. is resident (it does not wait for
        // that kernel to finish: no griddepcontrol.wait here)
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)grid);
        cfg.blockDim = dim3((unsigned)block_threads);
        cfg.dynamicSmemBytes = smem;
        cfg.stream = stream;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = at;
        cfg.numAttrs = 1;
        ONILDA_CUDA(cudaLaunchKernelEx(&cfg, estep_kernel<G, TPL, KC, kRare>, args, num_bundles));
    } else {
        estep_kernel<G, TPL, KC, kRare><<<(unsigned)grid, block_threads, smem, stream>>>(args, num_bundles);
    }
    ONILDA_CUDA(cudaGetLastError());
    return ONILDA_OK;
}

template <int G>
static int launch_group(int tpl, const EStepArgs& args, int sm_count, int block_threads, int extra_smem, int max_per_sm,
                        cudaStream_t stream, bool after_resident)
{
    if (args.rare_paths) {      // no kernels with the topic count compiled in for these
        switch (tpl) {
            case 1: return launch_one<G, 1, 0, true>(args, sm_count, block_threads, extra_smem, max_per_sm, stream, after_resident);
            case 2: return launch_one<G, 2, 0, true>(args, sm_count, block_threads, extra_smem, max_per_sm, stream, after_resident);
            case 3: return launch_one<G, 3, 0, true>(args, sm_count, block_threads, extra_smem, max_per_sm, stream, after_resident);
            case 4: return launch_one<G, 4, 0, true>(args, sm_count, block_threads, extra_smem, max_per_sm, stream, after_resident);
            case 5: return launch_one<G, 5, 0, true>(args, sm_count, block_threads, extra_smem, max_per_sm, stream, after_resident);
            case 6: return launch_one<G, 6, 0, true>(args, sm_count, block_threads, extra_smem, max_per_sm, stream, after_resident);
            case 7: return launch_one<G, 7, 0, true>(args, sm_count, block_threads, extra_smem, max_per_sm, stream, after_resident);
            case 8: return launch_one<G, 8, 0, true>(args, sm_count, block_threads, extra_smem, max_per_sm, stream, after_resident);
        }
        return onilda_fail(ONILDA_EINVAL, "unsupported topics-per-lane %d", tpl);
    }
    // ONI's own topic counts (ml_ops.sh: TOPIC_COUNT=20; the proxy shape of BASELINE.json: 50) get kernels
    // with the count compiled in
    if (args.num_topics == 20 && args.kstride == kpad_for(20) && tpl == (20 + G - 1) / G)
        return launch_one<G, (20 + G - 1) / G, 20>(args, sm_count, block_threads, extra_smem, max_per_sm, stream, after_resident);
    if (args.num_topics == 50 && args.kstride == kpad_for(50) && tpl == (50 + G - 1) / G) {
        if constexpr (G >= 8)
            return launch_one<G, (50 + G - 1) / G, 50>(args, sm_count, block_threads, extra_smem, max_per_sm, stream, after_resident);
    }
    switch (tpl) {
        case 1: return launch_one<G, 1, 0>(args, sm_count, block_threads, extra_smem, max_per_sm, stream, after_resident);
        case 2: return launch_one<G, 2, 0>(args, sm_count, block_threads, extra_smem, max_per_sm, stream, after_resident);
        case 3: return launch_one<G, 3, 0>(args, sm_count, block_threads, extra_smem, max_per_sm, stream, after_resident);
        case 4: return launch_one<G, 4, 0>(args, sm_count, block_threads, extra_smem, max_per_sm, stream, after_resident);
        case 5: return launch_one<G, 5, 0>(args, sm_count, block_threads, extra_smem, max_per_sm, stream, after_resident);
        case 6: return launch_one<G, 6, 0>(args, sm_count, block_threads, extra_smem, max_per_sm, stream, after_resident);
        case 7: return launch_one<G, 7, 0>(args, sm_count, block_threads, extra_smem, max_per_sm, stream, after_resident);
        case 8: return launch_one<G, 8, 0>(args, sm_count, block_threads, extra_smem, max_per_sm, stream, after_resident);
    }
    return onilda_fail(ONILDA_EINVAL, "unsupported topics-per-lane %d", tpl);
}

// One launch over documents order[order_begin, order_begin + order_count) with `group` lanes per
// document and `block_threads` threads per block.  args.queue must be zero (the caller
// clears it on the same stream).
int launch_estep(int group, const EStepArgs& args, int sm_count, int block_threads, int extra_smem, int max_per_sm,
                 cudaStream_t stream, int* launches, bool after_resident)
{
    if (args.order_count <= 0) return ONILDA_OK;
    const int tpl = (args.num_topics + group - 1) / group;
    int rc;
    switch (group) {
        case 4: rc = launch_group<4>(tpl, args, sm_count, block_threads, extra_smem, max_per_sm, stream, after_resident); break;
        case 8: rc = launch_group<8>(tpl, args, sm_count, block_threads, extra_smem, max_per_sm, stream, after_resident); break;
        case 16: rc = launch_group<16>(tpl, args, sm_count, block_threads, extra_smem, max_per_sm, stream, after_resident); break;
        case 32: rc = launch_group<32>(tpl, args, sm_count, block_threads, extra_smem, max_per_sm, stream, after_resident); break;
        default: return onilda_fail(ONILDA_EINVAL, "unsupported lane group %d", group);
    }
    if (rc == ONILDA_OK && launches) *launches += 1;
    return rc;
}

}  // namespace onilda
