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

// The same sums through the FP64 tensor-core instruction, used for its LATENCY (measured on B200,
// tools/warp_sum_latency.cu: a warp-wide sum costs 60 cycles this way against 168 for five butterfly rounds, and the
// word-by-word chain of a document waits for exactly this sum).  mma.m8n8k4.f64 computes D[m][n] = sum_k A[m][k] B[k][n];
// lane l holds A[l/4][l%4], B[l%4][l/4] and D[l/4][2(l%4)], D[l/4][2(l%4)+1].  The products are by 1.0, the additions
// run in the fixed order k = 0..3 in fp64: a deterministic sum like the butterfly's, in another order.
#ifndef ONILDA_MMA_SUM
#define ONILDA_MMA_SUM 1      // bit 0: one document per warp; bit 1: groups of four lanes (measured: no gain where several documents share a warp, 153 against 150 ms)
#endif
__device__ __forceinline__ void dmma884(double a, double b, double& d0, double& d1)
{
    const double z = 0.0;
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%4,%5};"
                 : "=d"(d0), "=d"(d1) : "d"(a), "d"(b), "d"(z), "d"(z));
}

// every lane receives the sum of v over the warp
__device__ __forceinline__ double warp_sum_mma(double v)
{
    double c0, c1, t0, t1;
    dmma884(1.0, v, c0, c1);          // with A = 1: the sums of lanes 8q..8q+3 and 8q+4..8q+7, q = lane % 4
    dmma884(1.0, c0 + c1, t0, t1);    // the four partials p[q] sit in B[q][*]: their sum, in every lane
    return t0;
}

// every lane receives the sum of v over its aligned group of four lanes
__device__ __forceinline__ double quad_sum_mma(double v)
{
    double d0, d1;
    dmma884(v, 1.0, d0, d1);          // with B = 1: row m = lane / 4 of D is the sum of A's row m, the lane's own four
    return d0;
}

// the per-word sum on the serial chain
template <int G>
__device__ __forceinline__ double chain_sum(double v)
{
    if constexpr (G == 32 && (ONILDA_MMA_SUM & 1)) return warp_sum_mma(v);
    if constexpr (G == 4 && (ONILDA_MMA_SUM & 2)) return quad_sum_mma(v);
    return group_sum<G>(v);
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
