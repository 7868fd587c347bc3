// Pieces shared by the two E-step kernels (estep.cu: several documents per warp; estep_long.cu: one long
// document per warp with its rows staged through shared memory).
#pragma once

#include "device_math.cuh"
#include "kernels.cuh"

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

// The math tables in shared memory, addressed through the 32-bit shared window (a generic pointer
// costs a window-base computation per access).
struct SmemTables {
    uint32_t psi, exp2, log;
    __device__ __forceinline__ static void lds2(uint32_t addr, double& x, double& y)
    {
        asm("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(x), "=d"(y) : "r"(addr));
    }
    __device__ __forceinline__ void log_entry(int i, double& inv_c, double& log_c) const { lds2(log + (uint32_t)i * 16u, inv_c, log_c); }
    __device__ __forceinline__ double exp2_frac(int i) const
    {
        double v;
        asm("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(exp2 + (uint32_t)i * 8u));
        return v;
    }
    __device__ __forceinline__ void psi_row(int row, double (&a)[6]) const
    {
        const uint32_t p = psi + (uint32_t)row * (uint32_t)sizeof(PsiRow);
        lds2(p, a[0], a[1]);
        lds2(p + 16, a[2], a[3]);
        lds2(p + 32, a[4], a[5]);
    }
};

template <int TPL>
struct WordItem {
    int word;            // the sentinel row (num_terms) where this group idles at this word
    double count;
    double lbmax;
    double beta[TPL];
    double old_phi[TPL];
};

// The rare paths are kept out of line so that the word loop stays compact in the instruction cache.
static __device__ __noinline__ PsiExp psi_exp_closed(double g)
{
    const DigammaParts dp = digamma_rational(g);
    PsiExp out;
    out.psi = dp.rational + fast_log(dp.y);
    out.e = fast_exp(out.psi);
    return out;
}

// One word of one document group in log space: phi_k = exp(psi_k + log beta'_k - lse).  Called by the whole
// warp (the shuffles inside need it converged).
template <int TPL>
struct LogSpaceWord {
    double psi[TPL], beta[TPL];       // in: beta' = 0 marks a padding topic or a true zero
    double y[TPL], inv_sum, lse;      // out: phi_k = y_k * inv_sum; lse is relative to lbmax
};

template <int G, int TPL>
static __device__ __noinline__ void redo_word_in_log_space(LogSpaceWord<TPL>* w)
{
    double av[TPL], m = -INFINITY;
#pragma unroll
    for (int t = 0; t < TPL; t++) {
        av[t] = (w->beta[t] > 0.0) ? w->psi[t] + log(w->beta[t]) : -INFINITY;
        m = fmax(m, av[t]);
    }
    m = group_max<G>(m);
    double ps = 0.0;
#pragma unroll
    for (int t = 0; t < TPL; t++) {
        w->y[t] = (av[t] > -INFINITY) ? exp(av[t] - m) : 0.0;
        ps += w->y[t];
    }
    ps = group_sum<G>(ps);
    w->inv_sum = 1.0 / ps;
    w->lse = m + log(ps);
}

// largest gamma the psi table serves (t = gamma + 1 < 2^kPsiOctaves), smallest the table exp takes (psi > -1e9)
__device__ constexpr double kPsiTableMax = (double)(1 << kPsiOctaves) - 1.0;
__device__ constexpr double kPsiTableMin = 1e-8;


// dynamic shared memory of the tables: the psi rows and the 2^(i/64) table
static constexpr size_t kEstepTableBytes = sizeof(PsiRow) * kPsiRows + sizeof(double) * kExpTabSize;

}  // namespace onilda
