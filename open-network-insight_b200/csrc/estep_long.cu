// E-step for LONG documents (K <= 64): one document per warp, one topic per lane (two for K > 32), the rows of the
// words ahead staged through a per-warp ring in shared memory.
//
// Why a kernel of its own.  The words of a document are a serial chain (Gauss-Seidel order of
// lda-inference.c:50-75: gamma and digamma(gamma) change after every word), so the E-step cannot end
// before the longest document has walked len x sweeps dependent steps.  This kernel is built around the
// latency of ONE step (measured with ncu, profiles/ncu_r2_long_word_trace.txt: 340 cycles): nothing that can be
// known in advance may sit on the chain E -> s -> phi -> gamma -> psi -> E.
//   * rows (beta' of the word, last sweep's phi, lbmax) are fetched kDepth = 6 words ahead with cp.async into a
//     ring of kRing = 8 slots per warp and read from shared memory one word ahead: a trip to L2 under load is
//     up to ~1000 cycles, to HBM more - register prefetch two words ahead still left 340 cycles of scoreboard
//     stall per word on the first multiply.
//   * word ids come 32 at a time in one coalesced load per lane, three blocks ahead; a word's id is a
//     shuffle from the lane that holds it.
//   * s = sum over the lanes of x = E beta' through the FP64 tensor pipe: two mma.m8n8k4.f64 and one add,
//     60 cycles against 168 for five butterfly rounds (estep_common.cuh).  A word starts with them; the
//     staging work for later words is issued while they are in flight.
//   * phi = x/s is x*r + (x*r)*e with the hardware reciprocal seed r, gamma is updated with one fma after e
//     (gamma - c*phi_old + c*x*r is formed beside the chain).
//   * K <= 32: the lane keeps the psi-table row of its gamma in registers; a word whose gamma left the row's
//     cell, like a word whose s is unusable, makes the warp leave the word loop, repair (reload the row /
//     redo the word in log space) and enter again - the loop itself has no taken branch.
// The arithmetic is that of estep.cu (see its header): psi and exp(psi) from tables, the likelihood bound in
// the algebraic form, a log-space redo of a word whose sum leaves the safe range.
#include "estep_common.cuh"

#ifndef ONILDA_LONG_MAX_THREADS
#define ONILDA_LONG_MAX_THREADS 256
#endif
#ifndef ONILDA_LONG_RING
#define ONILDA_LONG_RING 8
#endif
#ifndef ONILDA_LONG_KEEP_ROWS
#define ONILDA_LONG_KEEP_ROWS 1
#endif
#ifndef ONILDA_LONG_MIN_BLOCKS
#define ONILDA_LONG_MIN_BLOCKS 2
#endif

namespace onilda {

namespace {

constexpr int kRing = ONILDA_LONG_RING;        // slots of the per-warp ring (power of two)
constexpr int kDepth = ONILDA_LONG_RING - 2;   // words ahead; the slot written at word n was last read at word n - 2

// cp.async under a predicate, without a branch
__device__ __forceinline__ void cp_async16_if(bool on, uint32_t dst, const void* src)
{
    asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.b32 q, %0, 0;\n\t@q cp.async.cg.shared.global [%1], [%2], 16;\n\t}" ::"r"((int)on), "r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async8_if(bool on, uint32_t dst, const void* src)
{
    asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.b32 q, %0, 0;\n\t@q cp.async.ca.shared.global [%1], [%2], 8;\n\t}" ::"r"((int)on), "r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ double lds64(uint32_t addr)
{
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr) : "memory");
    return v;
}

// bytes of math tables a block keeps in shared memory (see the kernel)
template <int TPL>
struct LongTables {
    static constexpr uint32_t bytes = (ONILDA_LONG_KEEP_ROWS && TPL == 1) ? (uint32_t)(8u * kExpTabSize) : (uint32_t)kEstepTableBytes;
};

template <int TPL>
struct LongWord {
    double count, lbmax, beta[TPL], old_phi[TPL];
};

}  // namespace

// TPL: topics per lane (1 for K <= 32, 2 for K <= 64); lane j holds topics j and j + 32
// kGuard: test every word's sum s for the range of its reciprocal and redo the word in log space where it fails.
// The host leaves it out when alpha >= 0.0016: no sum can then leave the range (estep.cu, kRare).
template <int KC, int TPL, bool kGuard>
__global__ void __launch_bounds__(ONILDA_LONG_MAX_THREADS, ONILDA_LONG_MIN_BLOCKS) estep_long_kernel(EStepArgs a, unsigned long long num_docs)
{
    constexpr unsigned FULL = 0xffffffffu;
    constexpr bool kKeepRows = ONILDA_LONG_KEEP_ROWS && TPL == 1;      // two rows per lane do not fit the registers
    const int lane = threadIdx.x & 31;
    const int K = KC ? KC : a.num_topics;
    const uint32_t KS = KC ? (uint32_t)kpad_for(KC) : (uint32_t)a.kstride;
    const int32_t idle_word = a.idle_row;
    bool valid[TPL];
#pragma unroll
    for (int t = 0; t < TPL; t++) valid[t] = t * 32 + lane < K;
    const double inv_K = 1.0 / K;

    const double* __restrict__ beta_scaled = a.beta_scaled;
    const double* __restrict__ lbmax = a.lbmax;
    asm volatile("" : "+l"(beta_scaled), "+l"(lbmax));

    // Shared memory: [psi rows, only where the lanes do not keep their row in registers] [2^(i/64)] [constants, rings].
    // With kept rows the psi table is read from global memory on the rare reload (it stays in L1: 61 KB, and the
    // 61 KB of shared memory not taken here are L1 for this block and the bulk blocks beside it).
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ LogTabEntry logtab[128];
    constexpr uint32_t kPsiBytes = (uint32_t)(sizeof(PsiRow) * kPsiRows);
    constexpr uint32_t kTabBytes = LongTables<TPL>::bytes;
    {
        const int4* src = reinterpret_cast<const int4*>(static_cast<const unsigned char*>(a.psi_tab) + (kPsiBytes + 8u * kExpTabSize - kTabBytes));
        int4* dst = reinterpret_cast<int4*>(smem_raw);
        constexpr int chunks = (int)(kTabBytes / sizeof(int4));
        for (int i = threadIdx.x; i < chunks; i += blockDim.x) dst[i] = src[i];
        for (int i = threadIdx.x; i < 128; i += blockDim.x) logtab[i] = kLogTab[i];
        // constants and rings start as zeros (a phantom word may read a slot that was never staged)
        double* rings = reinterpret_cast<double*>(smem_raw + kTabBytes);
        const int ring_doubles = 2 + (int)(blockDim.x >> 5) * kRing * (int)(2 * KS + 2);
        for (int i = threadIdx.x; i < ring_doubles; i += blockDim.x) rings[i] = (i == 1) ? inv_K : 0.0;
        __syncthreads();
    }
    // the bulk launch behind this one in the stream waits for every block of this grid to get here, i.e. to be
    // resident (programmatic dependent launch, em.cu): the two launches then share every SM from the start
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    SmemTables tb;
    const uint32_t smem_base = (uint32_t)__cvta_generic_to_shared(smem_raw);
    tb.exp2 = smem_base + kTabBytes - 8u * kExpTabSize;
    tb.psi = smem_base;                                  // meaningful only when the rows are in shared memory
    tb.log = (uint32_t)__cvta_generic_to_shared(logtab);
    asm volatile("" : "+r"(tb.exp2), "+r"(tb.log));
    const PsiRow* __restrict__ psi_rows_global = static_cast<const PsiRow*>(a.psi_tab);
    // slot: beta'[KS] | phi_old[KS] | lbmax | pad
    const uint32_t slot_bytes = 2u * KS * 8u + 16u;
    const uint32_t ring_base = smem_base + kTabBytes;
    const uint32_t ring = ring_base + 16u + (uint32_t)(threadIdx.x >> 5) * (uint32_t)kRing * slot_bytes;
    // two constants in front of the rings, read by address where a lane has nothing of its own to read:
    // 0.0 (beta' of a lane beyond K) and phi before the first sweep (1/K, lda-inference.c:41; 0 beyond K)
    const uint32_t zero_word = ring_base;
    uint32_t start_word[TPL];
#pragma unroll
    for (int t = 0; t < TPL; t++) start_word[t] = valid[t] ? ring_base + 8u : ring_base;
    // this lane's part in staging a word.  One topic per lane (rows of at most 32 doubles): lanes 0..15 move 16-byte pieces
    // of the beta' row, lanes 16..31 of the phi row, one copy instruction.  Two topics per lane (rows of up to 64 doubles):
    // every lane moves a piece of the beta' row and, with a second instruction, the same piece of the phi row.
    const int piece = TPL == 1 ? (lane & 15) : lane;
    const bool stages_beta = TPL == 1 ? lane < 16 : true;
    const bool stages = (uint32_t)piece < KS / 2;
    const uint32_t stage_dst = (stages_beta ? 0u : KS * 8u) + 16u * (uint32_t)piece;

    // Documents are claimed from an atomic counter, longest first - but strided: claim q takes the document of rank
    // (q mod 8) * ceil(n/8) + q / 8 of the length-sorted list, so the eight warps of a block, which claim together,
    // get one document from each eighth of the list (one very long one per SM, not eight), and a block that the
    // scheduler places late (it may have to wait for a bulk block to leave) simply takes shorter documents instead of
    // holding back long ones that were promised to it.
    const unsigned long long stratum = (num_docs + 7) / 8;

    for (;;) {
        unsigned int claimed = 0;
        if (lane == 0) claimed = (unsigned int)atomicAdd(a.queue, 1ull);
        const unsigned long long q = __reduce_max_sync(FULL, claimed);
        if (q >= 8 * stratum) break;
        const unsigned long long slot_in_order = (q & 7) * stratum + (q >> 3);
        if (slot_in_order >= num_docs) continue;

        const int64_t d = a.order[a.order_begin + (int64_t)slot_in_order];
        const int64_t beg = a.offsets[d];
        const int len = (int)(a.offsets[d + 1] - beg);
        const int total = a.totals[d];
        const int32_t* __restrict__ doc_words = a.words + beg;
        const int32_t* __restrict__ doc_counts = a.counts + beg;
        double* __restrict__ doc_phi = a.phi + beg * (int64_t)KS;
        const double* stage_src_phi = doc_phi + 2 * piece;

        // every gamma of a document stays within [alpha, alpha + total]: the caller (em.cu) sends only documents for
        // which that interval lies inside the tables' range
        auto psi_exp = [&](const double (&g)[TPL], double (&psi)[TPL], double (&E)[TPL]) {
#pragma unroll
            for (int t = 0; t < TPL; t++) {
                const PsiExp pe = psi_exp_tab<true>(g[t], tb);
                psi[t] = pe.psi;
                E[t] = pe.e;
            }
        };

        // The same with the lane's table row kept in registers.  A long document's gamma_k seldom leaves its cell of the
        // table (1/64 octave of gamma + 1) from one word to the next, and the row's three 16-byte loads - 20 lanes, 20
        // different rows, bank conflicts - are 70 cycles of the chain.  psi_exp_kept evaluates with the row it holds and
        // reports a lane whose cell has changed; the caller then leaves the word loop, reloads (psi_exp_reload) and
        // comes back.
        double row[TPL][6], row_at[TPL];      // the row and (its cell's centre) - 1: u = t - centre = gamma - row_at, one operation from gamma
        int row_key[TPL];
        auto psi_exp_reload = [&](const double (&g)[TPL], double (&psi)[TPL], double (&E)[TPL]) {
#pragma unroll
            for (int t = 0; t < TPL; t++) {
                const double tt = g[t] + 1.0;
                row_key[t] = psi_row_of(tt);
                const double2* src = reinterpret_cast<const double2*>(psi_rows_global + row_key[t]);
#pragma unroll
                for (int i = 0; i < 3; i++) {
                    const double2 v = __ldg(src + i);
                    row[t][2 * i] = v.x;
                    row[t][2 * i + 1] = v.y;
                }
                row_at[t] = (tt - psi_cell_offset(tt)) - 1.0;      // exact: the centre has seven significant bits
                psi[t] = psi_poly<true>(row[t], g[t] - row_at[t]) - fast_rcp(g[t]);
                E[t] = fast_exp_tab<true>(psi[t], tb);
            }
        };
        auto psi_exp_kept = [&](const double (&g)[TPL], double (&psi)[TPL], double (&E)[TPL]) -> bool {
            bool moved = false;
#pragma unroll
            for (int t = 0; t < TPL; t++) {
                moved |= psi_row_of(g[t] + 1.0) != row_key[t];      // beside the chain
                psi[t] = psi_poly<true>(row[t], g[t] - row_at[t]) - fast_rcp(g[t]);
                E[t] = fast_exp_tab<true>(psi[t], tb);
            }
            return moved;
        };

        const double gamma0 = a.alpha + (total / ((double)K));
        double gam[TPL], psi[TPL], E[TPL];
#pragma unroll
        for (int t = 0; t < TPL; t++) gam[t] = valid[t] ? gamma0 : 1.0;
        if (kKeepRows) psi_exp_reload(gam, psi, E);
        else psi_exp(gam, psi, E);

        double lik = 0.0, lik_old = 0.0, converged = 1.0;
        int sweeps = 0;
        bool active = (converged > a.var_converged) && ((0 < a.var_max_iter) || (a.var_max_iter == -1));

        while (active) {
            sweeps++;
            const bool first = (sweeps == 1);
            double acc_lse = 0.0, acc_psi = 0.0;

            // ids, 32 words per block: the current block, the next two, and one in flight
            int blk = 0;
            int w0, c0, w1, c1, w2, c2, w3, c3;
            auto load_block = [&](int b, int& w, int& c) {
                const int m = b * 32 + lane;
                w = idle_word;
                c = 0;
                if (m < len) {
                    w = doc_words[m];
                    c = doc_counts[m];
                }
            };
            load_block(0, w0, c0);
            load_block(1, w1, c1);
            load_block(2, w2, c2);
            load_block(3, w3, c3);

            // Everything below up to the rare-path test at the end of a word is ONE basic block (predicated
            // copies, no branches), so that the compiler can place the staging work of later words into the
            // latency gaps of the chain of the current one.
            // rows of word m -> slot m % kRing (m is warp-uniform); always one commit group per call
            auto stage = [&](int m) {
                const int w = __shfl_sync(FULL, ((m >> 5) == blk) ? w0 : w1, m & 31);
                const uint32_t slot = ring + (uint32_t)(m & (kRing - 1)) * slot_bytes;
                const double* src = stages_beta ? beta_scaled + (uint64_t)(uint32_t)w * KS + 2 * piece
                                                : stage_src_phi + (uint64_t)(uint32_t)m * KS;
                // past the end of the document the id is the table's idle row (beta' = 1, lbmax = 0): a phantom word
                // then finds a harmless slot; phi rows past the end belong to other documents and are not touched
                cp_async16_if(stages && (stages_beta || (!first && m < len)), slot + stage_dst, src);
                if (TPL > 1) cp_async16_if(stages && !first && m < len, slot + KS * 8u + 16u * (uint32_t)piece, stage_src_phi + (uint64_t)(uint32_t)m * KS);
                cp_async8_if(lane == 31, slot + 2u * KS * 8u, lbmax + w);
                cp_async_commit();
            };
            // the staged rows of word m into registers (m's group has landed and the warp has synchronised)
            auto fetch = [&](int m, LongWord<TPL>& it) {
                const uint32_t slot = ring + (uint32_t)(m & (kRing - 1)) * slot_bytes;
                it.count = (double)__shfl_sync(FULL, ((m >> 5) == blk) ? c0 : c1, m & 31);
                it.lbmax = lds64(slot + 2u * KS * 8u);
                // lanes beyond K read the slot's zero word: beta' = 0 keeps them inert
#pragma unroll
                for (int t = 0; t < TPL; t++) {
                    const uint32_t k8 = 8u * (uint32_t)(t * 32 + lane);
                    it.beta[t] = lds64(valid[t] ? slot + k8 : zero_word);
                    it.old_phi[t] = lds64((valid[t] && !first) ? slot + KS * 8u + k8 : start_word[t]);
                }
            };

            // s = sum over the lanes of x, the same bits in every lane: two FP64 tensor-core instructions and one add
            // (estep_common.cuh), 108 cycles shorter than five butterfly rounds.  (An all-gather through shared
            // memory - every lane stores x, reads all K back with broadcast 16-byte loads and adds them in one fixed
            // tree - was measured slower than the butterfly: 85.0 ms against 81.3 ms for 592 hub documents alone.)
            auto sum_over_topics = [&](double x) -> double { return chain_sum<32>(x); };

            // The likelihood needs sum_n c_n (lbmax_n + log s_n).  s_n is the same in every lane, so lane n mod 32 keeps word
            // n's (s, c, lbmax) and every 32 words all lanes take ONE logarithm each: 1/32 of the logarithms per lane,
            // and the word step loses ~20 instructions.  acc_lse is then a per-lane partial, summed with the other
            // reductions at the end of the sweep.
            double s_keep = 1.0, c_keep = 0.0, lb_keep = 0.0;
            auto flush_logs = [&]() {
                acc_lse = fma(c_keep, lb_keep + fast_log_tab(s_keep, tb), acc_lse);
                c_keep = 0.0;
                s_keep = 1.0;
            };
            // One word of the chain E -> s -> phi -> gamma -> psi -> E, in two parts so that the staging work of a later word
            // (advance, below) runs while the sum is in the tensor pipe.  head: x = E beta', s = sum over the lanes.
            auto head = [&](const LongWord<TPL>& it, double (&x)[TPL]) -> double {
#pragma unroll
                for (int t = 0; t < TPL; t++) x[t] = E[t] * it.beta[t];
                double part = x[0];
#pragma unroll
                for (int t = 1; t < TPL; t++) part += x[t];
                return sum_over_topics(part);
            };
            // the state before the word, for the log-space redo
            double gam_in[TPL], psi_in[TPL], acc_psi_in = 0.0, s_in = 1.0;
            // tail: phi = x / s, gamma, psi, E.  True when s is outside [2^-963, 2^963), zero, negative or NaN (an integer
            // test on the exponent field; s is the same in every lane, the vote only tells the compiler so): the word
            // must then be done again in log space (redo, below) before the next one.
            auto tail = [&](const LongWord<TPL>& it, int n, const double (&x)[TPL], double s) -> bool {
                double gpre[TPL];
                acc_psi_in = acc_psi;
#pragma unroll
                for (int t = 0; t < TPL; t++) {
                    gam_in[t] = gam[t];
                    psi_in[t] = psi[t];
                    gpre[t] = fma(-it.count, it.old_phi[t], gam[t]);      // beside the chain
                }
                const double r0 = rcp_seed(s);
                double e = fma(-s, r0, 1.0);
                e = fma(e, e, e);
                const bool keeper = lane == (n & 31);
                s_keep = keeper ? s : s_keep;
                c_keep = keeper ? it.count : c_keep;
                lb_keep = keeper ? it.lbmax : lb_keep;
                double* const prow = doc_phi + (uint64_t)(uint32_t)n * KS + lane;
                const bool in_doc = n < len;
#pragma unroll
                for (int t = 0; t < TPL; t++) {
                    const double xr = x[t] * r0;
                    const double p = fma(xr, e, xr);
                    // gamma - c phi_old + c phi with phi = xr + xr e: the part without e is formed while e is computed
                    const double cx = it.count * xr;
                    gam[t] = fma(cx, e, cx + gpre[t]);
                    if (valid[t] && in_doc) prow[t * 32] = p;
                    acc_psi = fma(it.count * p, psi[t], acc_psi);
                }
                if (kGuard) s_in = s;
                bool leave = kGuard && (uint32_t)(double_hi(s) - 0x03C00000) >= (uint32_t)(0x7C200000 - 0x03C00000);
                if (kKeepRows) leave |= psi_exp_kept(gam, psi, E);
                else psi_exp(gam, psi, E);
                if (!kGuard && !kKeepRows) return false;
                return __any_sync(FULL, leave);
            };
            // the word again in log space, as the reference does it (lda-inference.c:52-68), from the saved state.  Rare;
            // reached only from behind the word loop, so that the loop itself has no taken branch.
            auto redo = [&](const LongWord<TPL>& it, int n) {
                if (!kGuard || (uint32_t)(double_hi(s_in) - 0x03C00000) < (uint32_t)(0x7C200000 - 0x03C00000)) {
                    psi_exp_reload(gam, psi, E);      // the sum was fine: a lane's gamma has moved to another cell of the table
                    return;
                }
                if (lane == 0) atomicAdd(a.error_flag + 1, 1);
                LogSpaceWord<TPL> ls;
#pragma unroll
                for (int t = 0; t < TPL; t++) {
                    ls.psi[t] = psi_in[t];
                    ls.beta[t] = it.beta[t];
                }
                redo_word_in_log_space<32, TPL>(&ls);
                acc_psi = acc_psi_in;
                double* const prow = doc_phi + (uint64_t)(uint32_t)n * KS + lane;
#pragma unroll
                for (int t = 0; t < TPL; t++) {
                    const double p2 = ls.y[t] * ls.inv_sum;
                    gam[t] = fma(it.count, p2, fma(-it.count, it.old_phi[t], gam_in[t]));
                    if (valid[t] && n < len) prow[t * 32] = p2;
                    acc_psi = fma(it.count * p2, psi_in[t], acc_psi);
                }
                if (lane == (n & 31)) {      // this word's log-sum goes in directly
                    acc_lse = fma(it.count, it.lbmax + ls.lse, acc_lse);
                    c_keep = 0.0;
                    s_keep = 1.0;
                }
                if (kKeepRows) psi_exp_reload(gam, psi, E);
                else psi_exp(gam, psi, E);
            };

            for (int m = 0; m < kDepth; m++) stage(m);
            cp_async_wait<kDepth - 1>();
            __syncwarp();
            LongWord<TPL> A, B;
            fetch(0, A);
            // word n: its sum first, then - while the sum is on its way - the copies for word n + kDepth are issued and the
            // rows of word n + 1 (landed by now) are read, then the rest of word n
            auto advance = [&](int n, const LongWord<TPL>& now, LongWord<TPL>& next) -> bool {
                double x[TPL];
                const double s = head(now, x);
                stage(n + kDepth);
                cp_async_wait<kDepth - 1>();
                __syncwarp();
                fetch(n + 1, next);
                return tail(now, n, x, s);
            };
            // Words four at a time (a document may end on up to three phantom words: the idle row, count 0, nothing
            // stored), 32 to a block of ids.  The block loaded at a rotation is first copied a whole block later: a
            // register move of a value still in flight would wait for its load.  The loop is left at a word that needs
            // the redo and entered again behind it.
            auto next_block = [&]() {
                flush_logs();
                blk++;
                w0 = w1; c0 = c1;
                w1 = w2; c1 = c2;
                w2 = w3; c2 = c3;
                load_block(blk + 3, w3, c3);
            };
            int n = 0;
            for (;;) {
                int sub = -1;
                while (n < len) {
                    const int n1 = ((n | 31) + 1 < len) ? (n | 31) + 1 : len;
                    do {
                        if (__builtin_expect(advance(n, A, B), 0)) { sub = 0; break; }
                        if (__builtin_expect(advance(n + 1, B, A), 0)) { sub = 1; break; }
                        if (__builtin_expect(advance(n + 2, A, B), 0)) { sub = 2; break; }
                        if (__builtin_expect(advance(n + 3, B, A), 0)) { sub = 3; break; }
                        n += 4;
                    } while (n < n1);
                    if (sub >= 0) break;
                    next_block();
                }
                if (sub < 0) break;
                redo((sub & 1) ? B : A, n + sub);
                for (int j = sub + 1; j < 4; j++) {
                    const bool again = (j & 1) ? advance(n + j, B, A) : advance(n + j, A, B);
                    if (again) redo((j & 1) ? B : A, n + j);
                }
                n += 4;
                if ((n & 31) == 0 || n >= len) next_block();
            }
            cp_async_wait<0>();
            __syncwarp();

            // likelihood bound of this sweep (lda-inference.c:93-127; estep.cu header)
            double gs = 0.0, lg = 0.0, ap = acc_psi, ls_all = acc_lse;
#pragma unroll
            for (int t = 0; t < TPL; t++) {
                if (valid[t]) {
                    gs += gam[t];
                    lg += lgamma(gam[t]);
                }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                gs += __shfl_xor_sync(FULL, gs, o);
                lg += __shfl_xor_sync(FULL, lg, o);
                ap += __shfl_xor_sync(FULL, ap, o);
                ls_all += __shfl_xor_sync(FULL, ls_all, o);
            }
            lik = a.lik_const - lgamma(gs) + lg - ap + ls_all;
            if (isnan(lik)) *a.error_flag = 1;   // the reference asserts here (lda-inference.c:78)
            converged = (lik_old - lik) / lik_old;
            lik_old = lik;
            active = (converged > a.var_converged) && ((sweeps < a.var_max_iter) || (a.var_max_iter == -1));
        }

        // doc_e_step (lda-estimate.c:40-46)
        double gs = 0.0, ps = 0.0;
#pragma unroll
        for (int t = 0; t < TPL; t++) {
            if (valid[t]) {
                gs += gam[t];
                ps += psi[t];
                a.gamma[d * K + t * 32 + lane] = gam[t];
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            gs += __shfl_xor_sync(FULL, gs, o);
            ps += __shfl_xor_sync(FULL, ps, o);
        }
        if (lane == 0) {
            a.doc_lik[d] = lik;
            a.doc_alpha_ss[d] = ps - K * digamma(gs);
            a.doc_sweeps[d] = sweeps;
        }
        // class_word[k][w] += c_n phi_nk (lda-estimate.c:48-55)
        if (a.class_word != nullptr) {
            constexpr int U = TPL == 1 ? 8 : 4;
            double* __restrict__ class_word = a.class_word + lane;
            for (int n0 = 0; n0 < len; n0 += U) {
                int w[U];
                double c[U], p[U][TPL];
#pragma unroll
                for (int u = 0; u < U; u++) {
                    const bool in = n0 + u < len;
                    w[u] = in ? doc_words[n0 + u] : -1;
                    c[u] = in ? (double)doc_counts[n0 + u] : 0.0;
#pragma unroll
                    for (int t = 0; t < TPL; t++)
                        p[u][t] = (in && valid[t] && sweeps != 0) ? doc_phi[(uint64_t)(uint32_t)(n0 + u) * KS + t * 32 + lane] : inv_K;
                }
#pragma unroll
                for (int u = 0; u < U; u++)
#pragma unroll
                    for (int t = 0; t < TPL; t++)
                        if (w[u] >= 0 && valid[t]) atomicAdd(class_word + (uint64_t)(uint32_t)w[u] * KS + t * 32, c[u] * p[u][t]);
            }
        }
    }
}

template <int KC, int TPL, bool kGuard = false>
static int launch_long_one(const EStepArgs& args, int sm_count, int block_threads, int max_per_sm, cudaStream_t stream)
{
    if (block_threads > ONILDA_LONG_MAX_THREADS || block_threads < 32 || (block_threads & 31))
        return onilda_fail(ONILDA_EINVAL, "long-document E-step block of %d threads (32..256, whole warps)", block_threads);
    const size_t slot_bytes = 2 * (size_t)args.kstride * 8 + 16;
    const size_t smem = LongTables<TPL>::bytes + 16 + (size_t)(block_threads / 32) * kRing * slot_bytes;
    ONILDA_CUDA(cudaFuncSetAttribute(estep_long_kernel<KC, TPL, kGuard>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    if (args.carveout_pct > 0)
        ONILDA_CUDA(cudaFuncSetAttribute(estep_long_kernel<KC, TPL, kGuard>, cudaFuncAttributePreferredSharedMemoryCarveout, args.carveout_pct));
    int per_sm = 0;
    ONILDA_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, estep_long_kernel<KC, TPL, kGuard>, block_threads, smem));
    if (per_sm < 1) per_sm = 1;
    if (max_per_sm > 0 && per_sm > max_per_sm) per_sm = max_per_sm;
    const unsigned long long wpb = (unsigned long long)(block_threads / 32);
    const unsigned long long docs = (unsigned long long)args.order_count;
    unsigned long long grid = (unsigned long long)sm_count * per_sm;
    if (grid > (docs + wpb - 1) / wpb) grid = (docs + wpb - 1) / wpb;
    estep_long_kernel<KC, TPL, kGuard><<<(unsigned)grid, block_threads, smem, stream>>>(args, docs);
    ONILDA_CUDA(cudaGetLastError());
    return ONILDA_OK;
}

// The documents order[order_begin, order_begin + order_count), one per warp.  K <= 64 and an even row stride <= 64.
int launch_estep_long(const EStepArgs& args, int sm_count, int block_threads, int max_per_sm, cudaStream_t stream, int* launches)
{
    if (args.order_count <= 0) return ONILDA_OK;
    if (args.num_topics > 64 || args.kstride > 64 || (args.kstride & 1))
        return onilda_fail(ONILDA_EINVAL, "long-document E-step: %d topics with row stride %d unsupported", args.num_topics, args.kstride);
    int rc;
    if (args.rare_paths) {      // the guarded variant, topic count not compiled in
        if (args.num_topics <= 32) rc = launch_long_one<0, 1, true>(args, sm_count, block_threads, max_per_sm, stream);
        else rc = launch_long_one<0, 2, true>(args, sm_count, block_threads, max_per_sm, stream);
    } else if (args.num_topics == 20 && args.kstride == kpad_for(20)) rc = launch_long_one<20, 1>(args, sm_count, block_threads, max_per_sm, stream);
    else if (args.num_topics == 50 && args.kstride == kpad_for(50)) rc = launch_long_one<50, 2>(args, sm_count, block_threads, max_per_sm, stream);
    else if (args.num_topics <= 32) rc = launch_long_one<0, 1>(args, sm_count, block_threads, max_per_sm, stream);
    else rc = launch_long_one<0, 2>(args, sm_count, block_threads, max_per_sm, stream);
    if (rc == ONILDA_OK && launches) *launches += 1;
    return rc;
}

}  // namespace onilda
