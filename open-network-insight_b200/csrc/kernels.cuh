// Shared declarations between the CUDA translation units of libonilda.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "onilda_internal.h"

#define ONILDA_CUDA(call)                                                                       \
    do {                                                                                        \
        cudaError_t e__ = (call);                                                               \
        if (e__ != cudaSuccess)                                                                 \
            return onilda_fail(ONILDA_ECUDA, "%s failed at %s:%d: %s", #call, __FILE__, __LINE__, \
                               cudaGetErrorString(e__));                                        \
    } while (0)

namespace onilda {

// Topic-to-lane layout.  A document is worked on by a group of G lanes of a
// warp; lane j of the group owns topics k = t*G + j, t = 0..TPL-1, so that for
// a fixed t the G lanes read G consecutive doubles of a word's topic row.
// kpad = G*TPL >= K; padded topics carry beta' = 0 and stay inert.
struct TopicLayout {
    int group;   // G   in {4, 8, 16, 32}
    int tpl;     // TPL in 1..8
    int kpad;    // G*TPL
};

// K <= 32 -> 4 lanes, K <= 64 -> 8, K <= 128 -> 16, K <= 256 -> 32.
__host__ __device__ constexpr int group_for(int K) { return K <= 32 ? 4 : K <= 64 ? 8 : K <= 128 ? 16 : 32; }
__host__ __device__ constexpr int kpad_for(int K) { return group_for(K) * ((K + group_for(K) - 1) / group_for(K)); }

inline bool choose_layout(int K, TopicLayout* out)
{
    if (K <= 0 || K > 256) return false;
    int g = group_for(K);
    out->group = g;
    out->tpl = (K + g - 1) / g;
    out->kpad = out->group * out->tpl;
    return true;
}

// Tail of the packed sufficient-statistics buffer, after the V*kpad table.
enum { SS_LIK = 0, SS_ALPHA = 1, SS_NDOCS = 2, SS_STEPS = 3, SS_TAIL = 4 };   // SS_STEPS: sum over documents of entries x sweeps
// The tables carry kTablePadRows spare rows after row V: with P <= 16 ranks each owning ceil(V/P) rows, the
// reduce-scatter / all-gather of the distributed M-step covers at most V + 15 rows; the idle row sits behind them.
enum { kTablePadRows = 16 };

struct EStepArgs {
    // corpus shard, CSR, offsets local to the shard (offsets[0] == 0)
    const int64_t* offsets;
    const int32_t* words;
    const int32_t* counts;
    const int32_t* totals;
    const int32_t* order;       // documents by decreasing length
    int64_t order_begin;        // this launch works on order[order_begin, order_begin + order_count)
    int64_t order_count;
    // model tables, word-major [V][kpad]
    const double* beta_scaled;  // exp(log_beta[w][k] - lbmax[w]); 0 in padding
    const double* lbmax;        // [V] max_k log_beta[w][k]
    int rare_paths;             // host side only: the launch needs the kernel variant with the closed forms and the log-space redo (estep.cu)
    int carveout_pct;           // host side only: preferred shared-memory carve-out of the launch (percent of the maximum; 0 = the driver's choice)
    const void* psi_tab;        // device_math.cuh PsiRow[kPsiRows] (hostmath.c builds it)
    const double* exp_tab;      // 2^(i/64), i < 64
    int num_topics;
    int num_terms;              // V
    int idle_row;               // row of beta_scaled / lbmax with beta' = 1 for k < K and lbmax = 0: what an idle group reads
    int kstride;                // row stride of the tables and of phi, in doubles (TopicLayout::kpad)
    double alpha;
    double lik_const;           // lgamma(alpha*K) - K*lgamma(alpha)
    int var_max_iter;
    double var_converged;       // the reference's float threshold, widened
    // scratch and outputs
    double* phi;                // [nnz][kstride]
    double* gamma;              // [D][K]
    double* doc_lik;            // [D]
    double* doc_alpha_ss;       // [D]
    int32_t* doc_sweeps;        // [D]
    double* class_word;         // [V][kpad], accumulated with atomics; NULL: inference only
    unsigned long long* queue;  // bundle counter
    int* error_flag;            // set to 1 on a NaN likelihood
};

int launch_estep(int group, const EStepArgs& args, int sm_count, int block_threads, int extra_smem, int max_per_sm,
                 cudaStream_t stream, int* launches, bool after_resident = false);

// estep_long.cu: documents one per warp with staged rows (K <= 32)
int launch_estep_long(const EStepArgs& args, int sm_count, int block_threads, int max_per_sm, cudaStream_t stream, int* launches);

// M-step and table kernels (mstep.cu)
int launch_tables_from_log_beta(const double* log_beta_kv, int K, int V, const TopicLayout& lay, double* log_beta_wk,
                                double* beta_scaled, double* lbmax, cudaStream_t stream, int* launches);
int launch_rebundle(const int32_t* order_static, const int32_t* win_off, int num_windows, const int32_t* doc_sweeps,
                    int32_t* order_out, cudaStream_t stream, int* launches);
int launch_reduce_doc_stats(const double* doc_lik, const double* doc_alpha_ss, const int64_t* offsets, const int32_t* doc_sweeps,
                            int64_t num_docs, double* partial,
                            double* ss_tail, cudaStream_t stream, int* launches);
int launch_column_sums(const double* table, int rows, const TopicLayout& lay, int sm_count, double* partial, double* total,
                       cudaStream_t stream, int* launches);
int launch_normalise(const double* class_word, const double* class_total, int K, int rows, const TopicLayout& lay,
                     double* log_beta_wk, double* beta_scaled, double* lbmax, cudaStream_t stream, int* launches);
int launch_export_log_beta(const double* log_beta_wk, int K, int V, const TopicLayout& lay, double* log_beta_kv,
                           cudaStream_t stream, int* launches);
int mstep_partial_doubles(int V, const TopicLayout& lay, int64_t num_docs);

}  // namespace onilda
