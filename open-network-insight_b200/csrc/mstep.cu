// M-step and model-table kernels (reference: lda_mle, oni-ml/oni-lda-c/lda-model.c:27-44;
// the scalar alpha Newton stays on the host, lda-alpha.c:42-68).
//
// Device model layout: word-major [V][kpad] so one word's topic row is one
// contiguous, sector-aligned vector for the E-step gather:
//     log_beta_wk[w][k]   log p(w | k)                        (saved / exported)
//     lbmax[w]            max_k log_beta_wk[w][k]
//     beta_scaled[w][k]   exp(log_beta_wk[w][k] - lbmax[w])   (what the E-step reads)
// Sufficient statistics class_word are kept in the same [V][kpad] layout and
// class_total[k] is formed here as its column sum (fixed-order two-stage
// reduction: this step adds no order dependence of its own; class_word itself was
// accumulated with fp64 atomics in the E-step and is reproducible only to ~1e-14 relative).
#include "kernels.cuh"

namespace onilda {

static constexpr int kMaxColBlocks = 1024;   // partial[] is sized for this many; launches use 4 x the SM count
static constexpr int kColThreads = 256;

// partial[b][k] = sum over this block's words of table[w][k]
__global__ void __launch_bounds__(kColThreads) column_partial_kernel(const double* __restrict__ table, int V, int kpad,
                                                                      double* __restrict__ partial)
{
    extern __shared__ double sh[];   // [rows_per_block][kpad]
    const int rows = kColThreads / kpad;          // thread rows that each walk a strided set of words
    const int k = threadIdx.x % kpad;
    const int r = threadIdx.x / kpad;
    double acc = 0.0;
    if (r < rows) {
        const int64_t chunk = ((int64_t)V + gridDim.x - 1) / gridDim.x;
        const int64_t lo = chunk * blockIdx.x;
        int64_t hi = lo + chunk;
        if (hi > V) hi = V;
        for (int64_t w = lo + r; w < hi; w += rows) acc += table[w * kpad + k];
        sh[r * kpad + k] = acc;
    }
    __syncthreads();
    if (r == 0) {
        double tot = sh[k];
        for (int i = 1; i < rows; i++) tot += sh[i * kpad + k];
        partial[(int64_t)blockIdx.x * kpad + k] = tot;
    }
}

__global__ void column_final_kernel(const double* __restrict__ partial, int nblocks, int kpad, double* __restrict__ total)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= kpad) return;
    double acc = 0.0;
    for (int b = 0; b < nblocks; b++) acc += partial[(int64_t)b * kpad + k];
    total[k] = acc;
}

// one thread per word: log-normalise its K statistics and build the E-step row
__global__ void __launch_bounds__(256) normalise_kernel(const double* __restrict__ class_word,
                                                        const double* __restrict__ class_total, int K, int V, int kpad,
                                                        double* __restrict__ log_beta_wk, double* __restrict__ beta_scaled,
                                                        double* __restrict__ lbmax)
{
    const int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= V) return;
    const double* cw = class_word + w * kpad;
    double* lb = log_beta_wk + w * kpad;
    double* bs = beta_scaled + w * kpad;
    double m = -INFINITY;
    for (int k = 0; k < K; k++) {
        const double x = cw[k];
        // lda-model.c:35-43: log(cw) - log(ct) when cw > 0, else the -100 sentinel
        const double v = x > 0 ? log(x) - log(class_total[k]) : -100.0;
        lb[k] = v;
        m = fmax(m, v);
    }
    lbmax[w] = m;
    for (int k = 0; k < K; k++) bs[k] = exp(lb[k] - m);
    for (int k = K; k < kpad; k++) { lb[k] = 0.0; bs[k] = 0.0; }
}

// host layout [K][V] -> device tables; one thread per word (coalesced over w for each k)
__global__ void __launch_bounds__(256) tables_from_kv_kernel(const double* __restrict__ log_beta_kv, int K, int V, int kpad,
                                                             double* __restrict__ log_beta_wk,
                                                             double* __restrict__ beta_scaled, double* __restrict__ lbmax)
{
    const int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= V) return;
    double* lb = log_beta_wk + w * kpad;
    double* bs = beta_scaled + w * kpad;
    double m = -INFINITY;
    for (int k = 0; k < K; k++) {
        const double v = log_beta_kv[(int64_t)k * V + w];
        lb[k] = v;
        m = fmax(m, v);
    }
    lbmax[w] = m;
    for (int k = 0; k < K; k++) bs[k] = exp(lb[k] - m);
    for (int k = K; k < kpad; k++) { lb[k] = 0.0; bs[k] = 0.0; }
}

__global__ void __launch_bounds__(256) export_kv_kernel(const double* __restrict__ log_beta_wk, int K, int V, int kpad,
                                                        double* __restrict__ log_beta_kv)
{
    const int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= V) return;
    for (int k = 0; k < K; k++) log_beta_kv[(int64_t)k * V + w] = log_beta_wk[w * kpad + k];
}

// fixed-order sum of the per-document likelihood and alpha statistics
static constexpr int kDocBlocks = 296;
static constexpr int kDocThreads = 256;

__global__ void __launch_bounds__(kDocThreads) doc_partial_kernel(const double* __restrict__ doc_lik,
                                                                  const double* __restrict__ doc_ass,
                                                                  const int64_t* __restrict__ offsets,
                                                                  const int32_t* __restrict__ doc_sweeps, int64_t D,
                                                                  double* __restrict__ partial)
{
    __shared__ double sl[kDocThreads], sa[kDocThreads], sw[kDocThreads];
    const int64_t chunk = (D + gridDim.x - 1) / gridDim.x;
    const int64_t lo = chunk * blockIdx.x;
    int64_t hi = lo + chunk;
    if (hi > D) hi = D;
    double l = 0.0, s = 0.0, w = 0.0;      // w: (document, word, sweep) steps = entries x sweeps, the E-step's unit of work
    for (int64_t d = lo + threadIdx.x; d < hi; d += kDocThreads) {
        l += doc_lik[d];
        s += doc_ass[d];
        w += (double)(offsets[d + 1] - offsets[d]) * (double)doc_sweeps[d];
    }
    sl[threadIdx.x] = l;
    sa[threadIdx.x] = s;
    sw[threadIdx.x] = w;
    __syncthreads();
    for (int o = kDocThreads / 2; o > 0; o >>= 1) {
        if (threadIdx.x < o) { sl[threadIdx.x] += sl[threadIdx.x + o]; sa[threadIdx.x] += sa[threadIdx.x + o]; sw[threadIdx.x] += sw[threadIdx.x + o]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) { partial[3 * blockIdx.x] = sl[0]; partial[3 * blockIdx.x + 1] = sa[0]; partial[3 * blockIdx.x + 2] = sw[0]; }
}

__global__ void doc_final_kernel(const double* __restrict__ partial, int nblocks, int64_t D, double* __restrict__ tail)
{
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    double l = 0.0, s = 0.0, w = 0.0;
    for (int b = 0; b < nblocks; b++) { l += partial[3 * b]; s += partial[3 * b + 1]; w += partial[3 * b + 2]; }
    tail[SS_LIK] = l;
    tail[SS_ALPHA] = s;
    tail[SS_NDOCS] = (double)D;
    tail[SS_STEPS] = w;
}

// Re-bundling of the bulk launch's documents.  A warp of the general kernel sweeps its documents together until the
// slowest has converged; the sweeps a document took in the last EM iteration predict the next ones well (correlation
// 0.7-0.95 at cfg 2).  Inside windows of the length-sorted order (at most 2048 documents of nearly the same length,
// boundaries from the host) the documents are put in the order of their last sweep counts, most first, stably: bundles
// then hold documents that stop together.  One warp per window; a window is walked 32 positions at a time.
__global__ void __launch_bounds__(32) rebundle_kernel(const int32_t* __restrict__ order_static, const int32_t* __restrict__ win_off,
                                                      const int32_t* __restrict__ doc_sweeps, int32_t* __restrict__ order_out)
{
    constexpr int kKeys = 64;
    __shared__ unsigned int cursor[kKeys];
    const int lane = threadIdx.x;
    const int b = win_off[blockIdx.x], e = win_off[blockIdx.x + 1];
    for (int k = lane; k < kKeys; k += 32) cursor[k] = 0;
    __syncwarp();
    // key 0 = most sweeps
    auto key_of = [&](int pos) {
        const int s = doc_sweeps[order_static[pos]];
        return kKeys - 1 - (s < 0 ? 0 : (s > kKeys - 1 ? kKeys - 1 : s));
    };
    for (int p = b + lane; p < e; p += 32) atomicAdd(&cursor[key_of(p)], 1u);
    __syncwarp();
    // exclusive scan of the 64 counts: two per lane
    unsigned int c0 = cursor[2 * lane], c1 = cursor[2 * lane + 1], inc = c0 + c1;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned int t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    __syncwarp();
    cursor[2 * lane] = inc - c0 - c1;
    cursor[2 * lane + 1] = inc - c1;
    __syncwarp();
    const unsigned int lt = (1u << lane) - 1u;
    for (int p0 = b; p0 < e; p0 += 32) {
        const int p = p0 + lane;
        const bool live = p < e;
        const int doc = live ? order_static[p] : 0;
        const int k = live ? key_of(p) : kKeys;          // dead lanes match only each other
        const unsigned int peers = __match_any_sync(0xffffffffu, k);
        unsigned int base = 0;
        if (live) base = cursor[k];
        __syncwarp();
        if (live) {
            order_out[b + base + __popc(peers & lt)] = doc;
            if ((peers & lt) == 0) cursor[k] = base + __popc(peers);
        }
        __syncwarp();
    }
}

int launch_rebundle(const int32_t* order_static, const int32_t* win_off, int num_windows, const int32_t* doc_sweeps,
                    int32_t* order_out, cudaStream_t stream, int* launches)
{
    if (num_windows <= 0) return ONILDA_OK;
    rebundle_kernel<<<(unsigned)num_windows, 32, 0, stream>>>(order_static, win_off, doc_sweeps, order_out);
    ONILDA_CUDA(cudaGetLastError());
    if (launches) *launches += 1;
    return ONILDA_OK;
}

int mstep_partial_doubles(int V, const TopicLayout& lay, int64_t num_docs)
{
    (void)V; (void)num_docs;
    int a = kMaxColBlocks * lay.kpad, b = 3 * kDocBlocks;
    return a > b ? a : b;
}

static int blocks_for(int64_t n, int threads) { return (int)((n + threads - 1) / threads); }

int launch_tables_from_log_beta(const double* log_beta_kv, int K, int V, const TopicLayout& lay, double* log_beta_wk,
                                double* beta_scaled, double* lbmax, cudaStream_t stream, int* launches)
{
    if (V == 0) return ONILDA_OK;
    tables_from_kv_kernel<<<blocks_for(V, 256), 256, 0, stream>>>(log_beta_kv, K, V, lay.kpad, log_beta_wk, beta_scaled, lbmax);
    ONILDA_CUDA(cudaGetLastError());
    if (launches) *launches += 1;
    return ONILDA_OK;
}

int launch_export_log_beta(const double* log_beta_wk, int K, int V, const TopicLayout& lay, double* log_beta_kv,
                           cudaStream_t stream, int* launches)
{
    if (V == 0) return ONILDA_OK;
    export_kv_kernel<<<blocks_for(V, 256), 256, 0, stream>>>(log_beta_wk, K, V, lay.kpad, log_beta_kv);
    ONILDA_CUDA(cudaGetLastError());
    if (launches) *launches += 1;
    return ONILDA_OK;
}

int launch_reduce_doc_stats(const double* doc_lik, const double* doc_alpha_ss, const int64_t* offsets, const int32_t* doc_sweeps,
                            int64_t num_docs, double* partial,
                            double* ss_tail, cudaStream_t stream, int* launches)
{
    doc_partial_kernel<<<kDocBlocks, kDocThreads, 0, stream>>>(doc_lik, doc_alpha_ss, offsets, doc_sweeps, num_docs, partial);
    ONILDA_CUDA(cudaGetLastError());
    doc_final_kernel<<<1, 32, 0, stream>>>(partial, kDocBlocks, num_docs, ss_tail);
    ONILDA_CUDA(cudaGetLastError());
    if (launches) *launches += 2;
    return ONILDA_OK;
}

static int col_blocks(int sm_count)
{
    int b = 4 * (sm_count > 0 ? sm_count : 1);
    return b > kMaxColBlocks ? kMaxColBlocks : b;
}

// class_total[k] = sum over `rows` table rows of column k, in a fixed order (two stages)
int launch_column_sums(const double* table, int rows, const TopicLayout& lay, int sm_count, double* partial, double* total,
                       cudaStream_t stream, int* launches)
{
    const int blocks = col_blocks(sm_count);
    const size_t shmem = sizeof(double) * (size_t)(kColThreads / lay.kpad) * lay.kpad;
    column_partial_kernel<<<blocks, kColThreads, shmem, stream>>>(table, rows, lay.kpad, partial);
    ONILDA_CUDA(cudaGetLastError());
    column_final_kernel<<<blocks_for(lay.kpad, 64), 64, 0, stream>>>(partial, blocks, lay.kpad, total);
    ONILDA_CUDA(cudaGetLastError());
    if (launches) *launches += 2;
    return ONILDA_OK;
}

// lda_mle over `rows` rows (all pointers already at the first of them)
int launch_normalise(const double* class_word, const double* class_total, int K, int rows, const TopicLayout& lay,
                     double* log_beta_wk, double* beta_scaled, double* lbmax, cudaStream_t stream, int* launches)
{
    if (rows <= 0) return ONILDA_OK;
    normalise_kernel<<<blocks_for(rows, 256), 256, 0, stream>>>(class_word, class_total, K, rows, lay.kpad, log_beta_wk,
                                                                beta_scaled, lbmax);
    ONILDA_CUDA(cudaGetLastError());
    if (launches) *launches += 1;
    return ONILDA_OK;
}

}  // namespace onilda
