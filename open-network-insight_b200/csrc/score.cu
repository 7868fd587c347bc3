// Post-LDA suspicious-connects scoring on sm_100a.
//
// Reference (Scala, Spark UDFs over broadcast maps):
//   theta/phi normalisation  OniLDACWrapper.scala:121-144, 170-195
//   score(ip, word)          SuspiciousConnectsScoreFunction.scala:11-23
//   flow: two sides + min    netflow/FlowPostLDA.scala:94-125
//   filter + ascending top-k netflow/FlowPostLDA.scala:79-80, dns/DNSPostLDA.scala:49-70,
//                            utilities/DataFrameUtils.scala:21-38
//
// The scoring kernel is a streaming gather: per event side it reads two int32
// ids, gathers one theta row and one phi row (K doubles each) and writes one
// double.  A warp takes 32 events; its lanes first fetch the 32+32 rows
// cooperatively (consecutive lanes read consecutive doubles of a row, so every
// 32-byte sector that is fetched is used in full) into a per-warp shared-memory
// tile, then each lane forms its own event's dot product from the tile with the
// products rounded and added left to right - the order of Scala's
// zip/map/sum - so the result is bit-identical to a sequential CPU evaluation.
#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "kernels.cuh"

struct onilda_scorer {
    int device;
    int sm_count;
    int K;
    int64_t D;
    int32_t V;
    double* d_theta;   // [(D+1)][K], row D = uniform 1/K (unknown ip)
    double* d_phi;     // [(V+1)][K], row V = uniform 1/K (unknown word)
    cudaStream_t stream[2];
    cudaEvent_t ev0, ev1;
    // staging for host-buffer calls (two chunks in flight)
    int64_t chunk;
    int32_t* d_ids[2];     // 4 * chunk
    double* d_out[2];      // chunk
    // selection scratch
    unsigned long long* d_hist;   // [2048 + 8]
    int64_t* d_sel_idx;
    double* d_sel_score;
    int64_t sel_cap;
    int64_t* d_cand_idx;          // candidates still in question after two radix digits
    double* d_cand_score;
    int64_t cand_cap;
    double kernel_ms;
    int64_t launches;
};

namespace {

using namespace onilda;

struct DeviceGuard {
    int prev;
    explicit DeviceGuard(int dev) : prev(-1) { cudaGetDevice(&prev); if (dev != prev) cudaSetDevice(dev); }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

constexpr int kScoreThreads = 256;            // 8 warps
constexpr int kScoreWarps = kScoreThreads / 32;

__device__ __forceinline__ int tile_stride(int K) { return K | 1; }   // odd stride: conflict-free 64-bit rows

// Fetch 32 rows (one per lane's `row` index) of `table` into tile[32][stride].
__device__ __forceinline__ void stage_rows(const double* __restrict__ table, int64_t row, int K, int stride,
                                           double* tile, int lane)
{
    int r = 0, k = lane;
    while (k >= K) { k -= K; r++; }
    for (; r < 32;) {
        const int64_t src_row = __shfl_sync(0xffffffffu, row, r & 31);
        tile[r * stride + k] = __ldg(table + src_row * K + k);
        k += 32;
        while (k >= K) { k -= K; r++; }
    }
}

__device__ __forceinline__ double dot_in_order(const double* a, const double* b, int K)
{
    double s = 0.0;
    for (int k = 0; k < K; k++) s = __dadd_rn(s, __dmul_rn(a[k], b[k]));   // no fma: products are rounded first
    return s;
}

template <bool FLOW>
__global__ void __launch_bounds__(kScoreThreads) score_kernel(const double* __restrict__ theta, int64_t D,
                                                              const double* __restrict__ phi, int32_t V, int K,
                                                              const int32_t* __restrict__ sdoc,
                                                              const int32_t* __restrict__ sword,
                                                              const int32_t* __restrict__ ddoc,
                                                              const int32_t* __restrict__ dword, int64_t n,
                                                              double* __restrict__ out)
{
    extern __shared__ double smem[];
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int stride = tile_stride(K);
    double* tile_t = smem + (size_t)warp * 2 * 32 * stride;
    double* tile_p = tile_t + 32 * stride;
    const int64_t tiles = (n + 31) / 32;
    const int64_t warps_total = (int64_t)gridDim.x * kScoreWarps;
    for (int64_t tile = (int64_t)blockIdx.x * kScoreWarps + warp; tile < tiles; tile += warps_total) {
        const int64_t i = tile * 32 + lane;
        const bool live = i < n;
        int32_t d0 = live ? sdoc[i] : -1, w0 = live ? sword[i] : -1;
        // ids outside the tables (e.g. -1) take the uniform row (SuspiciousConnectsScoreFunction.scala:13-18).
        // Flow is the exception for the ip: FlowPostLDA joins the events with the document table (an inner join,
        // FlowPostLDA.scala:94), so an event whose source or destination ip has no theta row is gone from its
        // results.  Here such an event keeps its slot and gets +infinity: no threshold keeps it, no top-k takes it.
        bool dropped = FLOW && !(d0 >= 0 && d0 < D);
        int64_t drow = (d0 >= 0 && d0 < D) ? d0 : D;
        int64_t wrow = (w0 >= 0 && w0 < V) ? w0 : V;
        stage_rows(theta, drow, K, stride, tile_t, lane);
        stage_rows(phi, wrow, K, stride, tile_p, lane);
        __syncwarp();
        double s = dot_in_order(tile_t + lane * stride, tile_p + lane * stride, K);
        __syncwarp();
        if (FLOW) {
            int32_t d1 = live ? ddoc[i] : -1, w1 = live ? dword[i] : -1;
            dropped = dropped || !(d1 >= 0 && d1 < D);
            drow = (d1 >= 0 && d1 < D) ? d1 : D;
            wrow = (w1 >= 0 && w1 < V) ? w1 : V;
            stage_rows(theta, drow, K, stride, tile_t, lane);
            stage_rows(phi, wrow, K, stride, tile_p, lane);
            __syncwarp();
            const double s2 = dot_in_order(tile_t + lane * stride, tile_p + lane * stride, K);
            __syncwarp();
            s = (s <= s2) ? s : s2;   // scala.math.min (FlowPostLDA.scala:117-119)
        }
        if (live) out[i] = dropped ? INFINITY : s;
    }
}

// Even K (the usual case: ONI uses K = 20): rows are whole numbers of 16-byte pairs, and only the
// PRODUCTS need staging.  Lanes fetch pair c*32+lane of the 32 theta rows and the matching pair of
// the 32 phi rows (two coalesced 128-bit gathers), multiply (each product is rounded on its own, so
// doing them in parallel changes nothing), store the pair of products, and the owning lane then
// adds its row's K products left to right - the order of Scala's sum.  Half the shared-memory
// traffic of staging both tables, no shuffles, 128-bit accesses throughout.
__device__ __forceinline__ double2 ldg_pair(const double* p)
{
    return __ldg(reinterpret_cast<const double2*>(p));
}

// CC: pairs per row when known at compile time (10 for ONI's K = 20, 25 for K = 50; 0: taken from K).  With it the
// gather loop unrolls into 2*CC independent 16-byte loads per lane and side - the kernel waits on the memory system
// (ncu: long scoreboard 16.8 per issued instruction with the rolled loop), so loads in flight are what it needs.
template <bool FLOW, int CC>
__global__ void __launch_bounds__(kScoreThreads) score_pairs_kernel(const double* __restrict__ theta, int64_t D,
                                                                    const double* __restrict__ phi, int32_t V, int K,
                                                                    const int32_t* __restrict__ sdoc,
                                                                    const int32_t* __restrict__ sword,
                                                                    const int32_t* __restrict__ ddoc,
                                                                    const int32_t* __restrict__ dword, int64_t n,
                                                                    double* __restrict__ out)
{
    extern __shared__ double smem[];
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int C = CC ? CC : (K >> 1);     // pairs per row
    const int stride = 2 * C + 2;         // C + 1 pairs: odd pair stride -> conflict-free 128-bit rows
    double* tile = smem + (size_t)warp * (32 * stride + 64);
    int* rows_t = reinterpret_cast<int*>(tile + 32 * stride);
    int* rows_p = rows_t + 32;
    // (row, pair) walked by this lane: element e = c*32 + lane, row = e / C, pair = e % C
    const int r0 = lane / C, q0 = lane % C, dr = 32 / C, dq = 32 % C;
    const int64_t tiles = (n + 31) / 32;
    const int64_t warps_total = (int64_t)gridDim.x * kScoreWarps;
    for (int64_t tile_id = (int64_t)blockIdx.x * kScoreWarps + warp; tile_id < tiles; tile_id += warps_total) {
        const int64_t i = tile_id * 32 + lane;
        const bool live = i < n;
        double result = 0.0;
        bool dropped = false;     // flow only: an ip without a theta row (the inner join of FlowPostLDA.scala:94; see score_kernel)
#pragma unroll
        for (int side = 0; side < (FLOW ? 2 : 1); side++) {
            const int32_t d0 = live ? (side ? ddoc[i] : sdoc[i]) : -1;
            const int32_t w0 = live ? (side ? dword[i] : sword[i]) : -1;
            // ids outside the tables (e.g. -1) take the uniform row (SuspiciousConnectsScoreFunction.scala:13-18)
            if (FLOW) dropped = dropped || !(d0 >= 0 && d0 < D);
            rows_t[lane] = (d0 >= 0 && d0 < D) ? d0 : (int)D;
            rows_p[lane] = (w0 >= 0 && w0 < V) ? w0 : (int)V;
            __syncwarp();
            int r = r0, q = q0;
            if (CC) {
                // all loads first, then the products: 2*CC requests in flight per lane
                double2 t2[CC ? CC : 1], p2[CC ? CC : 1];
                int rr[CC ? CC : 1], qq[CC ? CC : 1];
#pragma unroll
                for (int c = 0; c < CC; c++) {
                    rr[c] = r;
                    qq[c] = q;
                    t2[c] = ldg_pair(theta + (int64_t)rows_t[r] * (2 * CC) + 2 * q);
                    p2[c] = ldg_pair(phi + (int64_t)rows_p[r] * (2 * CC) + 2 * q);
                    r += dr;
                    q += dq;
                    if (q >= C) { q -= C; r++; }
                }
#pragma unroll
                for (int c = 0; c < CC; c++) {
                    double2 pr;
                    pr.x = __dmul_rn(t2[c].x, p2[c].x);
                    pr.y = __dmul_rn(t2[c].y, p2[c].y);
                    *reinterpret_cast<double2*>(tile + rr[c] * stride + 2 * qq[c]) = pr;
                }
            } else {
                for (int c = 0; c < C; c++) {
                    const double2 t2 = ldg_pair(theta + (int64_t)rows_t[r] * K + 2 * q);
                    const double2 p2 = ldg_pair(phi + (int64_t)rows_p[r] * K + 2 * q);
                    double2 pr;
                    pr.x = __dmul_rn(t2.x, p2.x);
                    pr.y = __dmul_rn(t2.y, p2.y);
                    *reinterpret_cast<double2*>(tile + r * stride + 2 * q) = pr;
                    r += dr;
                    q += dq;
                    if (q >= C) { q -= C; r++; }
                }
            }
            __syncwarp();
            double s = 0.0;
            const double* mine = tile + lane * stride;
#pragma unroll
            for (int c = 0; c < C; c++) {
                const double2 v = *reinterpret_cast<const double2*>(mine + 2 * c);
                s = __dadd_rn(s, v.x);      // no fma, fixed order: bit-identical to the sequential sum
                s = __dadd_rn(s, v.y);
            }
            __syncwarp();
            result = (side == 0) ? s : ((result <= s) ? result : s);   // scala.math.min (FlowPostLDA.scala:117-119)
        }
        if (live) out[i] = dropped ? INFINITY : result;
    }
}

// theta row = gamma / sum (left-to-right sum), zeros when sum <= 0 (OniLDACWrapper.scala:132-144)
__global__ void theta_kernel(const double* __restrict__ gamma, int64_t D, int K, double* __restrict__ theta)
{
    const int64_t d = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (d >= D) return;
    const double* g = gamma + d * K;
    double s = 0.0;
    for (int k = 0; k < K; k++) s = __dadd_rn(s, g[k]);
    for (int k = 0; k < K; k++) theta[d * K + k] = (s > 0) ? g[k] / s : 0.0;
}

// phi: per topic exp, sum over the vocabulary, divide, transpose (OniLDACWrapper.scala:121-130, 178-195)
__global__ void __launch_bounds__(256) phi_sum_kernel(const double* __restrict__ log_beta_kv, int V, double* __restrict__ sums)
{
    __shared__ double sh[256];
    const int k = blockIdx.x;
    double acc = 0.0;
    for (int w = threadIdx.x; w < V; w += 256) acc += exp(log_beta_kv[(int64_t)k * V + w]);
    sh[threadIdx.x] = acc;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) sums[k] = sh[0];
}

__global__ void phi_norm_kernel(const double* __restrict__ log_beta_kv, const double* __restrict__ sums, int K, int V,
                                double* __restrict__ phi_wk)
{
    const int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= V) return;
    for (int k = 0; k < K; k++) phi_wk[w * K + k] = exp(log_beta_kv[(int64_t)k * V + w]) / sums[k];
}

// phi straight from the E-step's word-major table log_beta_wk [V][kpad] (no host round trip):
// per-block partial sums of exp over a slice of the vocabulary, then a fixed-order final sum
constexpr int kPhiBlocks = 296;
__global__ void __launch_bounds__(256) phi_partial_wk_kernel(const double* __restrict__ lb_wk, int V, int K, int kpad,
                                                             double* __restrict__ partial)
{
    extern __shared__ double sh[];                       // [rows][K]
    const int rows = 256 / K, k = threadIdx.x % K, r = threadIdx.x / K;
    const int64_t chunk = ((int64_t)V + gridDim.x - 1) / gridDim.x, lo = chunk * blockIdx.x;
    int64_t hi = lo + chunk;
    if (hi > V) hi = V;
    if (r < rows) {
        double acc = 0.0;
        for (int64_t w = lo + r; w < hi; w += rows) acc += exp(lb_wk[w * kpad + k]);
        sh[r * K + k] = acc;
    }
    __syncthreads();
    if (r == 0) {
        double tot = sh[k];
        for (int i = 1; i < rows; i++) tot += sh[i * K + k];
        partial[(int64_t)blockIdx.x * K + k] = tot;
    }
}

__global__ void phi_final_wk_kernel(const double* __restrict__ partial, int nblocks, int K, double* __restrict__ sums)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= K) return;
    double acc = 0.0;
    for (int b = 0; b < nblocks; b++) acc += partial[(int64_t)b * K + k];
    sums[k] = acc;
}

__global__ void phi_norm_wk_kernel(const double* __restrict__ lb_wk, const double* __restrict__ sums, int V, int K, int kpad,
                                   double* __restrict__ phi_wk)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (int64_t)V * K) return;
    const int64_t w = i / K;
    const int k = (int)(i % K);
    phi_wk[i] = exp(lb_wk[w * kpad + k]) / sums[k];
}

__global__ void fill_kernel(double* p, int n, double v)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}

// ---- selection: k smallest (score, index) pairs with score <= threshold ----

__device__ __forceinline__ unsigned long long order_key(double x)
{
    const unsigned long long b = (unsigned long long)__double_as_longlong(x);
    return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}

constexpr int kDigitBits = 11;
constexpr int kBins = 1 << kDigitBits;

// Histogram of one digit.  Items considered: key <= max_key, (key & key_mask) == key_prefix and,
// when on_index, key == key_prefix exactly and (idx & idx_mask) == idx_prefix.
// idx_map: event index of element i when the array is a compacted candidate list (nullptr: i itself).
__global__ void __launch_bounds__(256) select_hist_kernel(const double* __restrict__ score,
                                                          const int64_t* __restrict__ idx_map, int64_t n,
                                                          unsigned long long max_key, unsigned long long key_prefix,
                                                          unsigned long long key_mask, int on_index,
                                                          unsigned long long idx_prefix, unsigned long long idx_mask,
                                                          int shift, unsigned int digit_mask,
                                                          unsigned long long* __restrict__ hist)
{
    __shared__ unsigned int sh[kBins];
    for (int b = threadIdx.x; b < kBins; b += blockDim.x) sh[b] = 0;
    __syncthreads();
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const unsigned long long key = order_key(score[i]);
        if (key > max_key || (key & key_mask) != key_prefix) continue;
        unsigned long long digit_src = key;
        if (on_index) {
            const unsigned long long ev = idx_map ? (unsigned long long)idx_map[i] : (unsigned long long)i;
            if ((ev & idx_mask) != idx_prefix) continue;
            digit_src = ev;
        }
        atomicAdd(&sh[(unsigned int)(digit_src >> shift) & digit_mask], 1u);
    }
    __syncthreads();
    for (int b = threadIdx.x; b < kBins; b += blockDim.x)
        if (sh[b]) atomicAdd(&hist[b], (unsigned long long)sh[b]);
}

// Collect every item with (key, idx) <= (cut_key, cut_idx) and key <= max_key (unordered).
__global__ void __launch_bounds__(256) select_collect_kernel(const double* __restrict__ score,
                                                             const int64_t* __restrict__ idx_map, int64_t n,
                                                             unsigned long long max_key, unsigned long long cut_key,
                                                             unsigned long long cut_idx, int64_t cap,
                                                             unsigned long long* __restrict__ counter,
                                                             int64_t* __restrict__ idx_out, double* __restrict__ score_out)
{
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const double s = score[i];
        const unsigned long long key = order_key(s);
        if (key > max_key) continue;
        const int64_t ev = idx_map ? idx_map[i] : i;
        if (key < cut_key || (key == cut_key && (unsigned long long)ev <= cut_idx)) {
            const unsigned long long slot = atomicAdd(counter, 1ull);
            if ((int64_t)slot < cap) { idx_out[slot] = ev; score_out[slot] = s; }
        }
    }
}

// After two digits of the radix select: events whose key prefix is BELOW the cut prefix are selected
// for certain (written straight to the output, counter[0]); those EQUAL to it are the only ones still
// in question and are compacted (score + event index, counter[1]) so that the remaining passes read
// that short list instead of all n scores.
__global__ void __launch_bounds__(256) select_split_kernel(const double* __restrict__ score, int64_t n,
                                                           unsigned long long max_key, unsigned long long prefix,
                                                           unsigned long long mask, int64_t sel_cap, int64_t cand_cap,
                                                           unsigned long long* __restrict__ counter,
                                                           int64_t* __restrict__ sel_idx, double* __restrict__ sel_score,
                                                           int64_t* __restrict__ cand_idx, double* __restrict__ cand_score)
{
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const double s = score[i];
        const unsigned long long key = order_key(s);
        if (key > max_key) continue;
        const unsigned long long pre = key & mask;
        if (pre < prefix) {
            const unsigned long long slot = atomicAdd(&counter[0], 1ull);
            if ((int64_t)slot < sel_cap) { sel_idx[slot] = i; sel_score[slot] = s; }
        } else if (pre == prefix) {
            const unsigned long long slot = atomicAdd(&counter[1], 1ull);
            if ((int64_t)slot < cand_cap) { cand_idx[slot] = i; cand_score[slot] = s; }
        }
    }
}

unsigned long long host_order_key(double x)
{
    unsigned long long b;
    memcpy(&b, &x, sizeof b);
    return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}

int ensure_sel(onilda_scorer* s, int64_t cap)
{
    if (cap <= s->sel_cap) return ONILDA_OK;
    cudaFree(s->d_sel_idx); cudaFree(s->d_sel_score);
    s->d_sel_idx = nullptr; s->d_sel_score = nullptr; s->sel_cap = 0;
    if (cudaMalloc((void**)&s->d_sel_idx, sizeof(int64_t) * (size_t)cap) != cudaSuccess ||
        cudaMalloc((void**)&s->d_sel_score, sizeof(double) * (size_t)cap) != cudaSuccess) {
        cudaGetLastError();
        return onilda_fail(ONILDA_ENOMEM, "top-k: cannot allocate room for %lld candidates on the device", (long long)cap);
    }
    s->sel_cap = cap;
    return ONILDA_OK;
}

int ensure_cand(onilda_scorer* s, int64_t cap)
{
    if (cap <= s->cand_cap) return ONILDA_OK;
    cudaFree(s->d_cand_idx); cudaFree(s->d_cand_score);
    s->d_cand_idx = nullptr; s->d_cand_score = nullptr; s->cand_cap = 0;
    if (cudaMalloc((void**)&s->d_cand_idx, sizeof(int64_t) * (size_t)cap) != cudaSuccess ||
        cudaMalloc((void**)&s->d_cand_score, sizeof(double) * (size_t)cap) != cudaSuccess) {
        cudaGetLastError();
        return onilda_fail(ONILDA_ENOMEM, "top-k: cannot allocate room for %lld candidates on the device", (long long)cap);
    }
    s->cand_cap = cap;
    return ONILDA_OK;
}

// k smallest of score_dev[0..n) (device) under `threshold`; ties by index.  Host outputs.
// Radix select, most significant digit first, on the composite (order_key(score), index): each
// pass is one histogram kernel plus a 16 KB read-back.  After two digits the events still in
// question are normally a sliver of n; they are compacted once (select_split_kernel) and the
// remaining eight passes read that list, so all n scores are read three times, not eleven.
int select_topk_device(onilda_scorer* s, const double* score_dev, int64_t n, double threshold, int64_t k,
                       int64_t* idx_out, double* score_out, int64_t* n_out)
{
    cudaStream_t st = s->stream[0];
    // +infinity marks an event the flow join dropped (score_kernel): no threshold, however large, keeps it
    const unsigned long long max_key = isnan(threshold) ? 0ull : host_order_key(threshold > DBL_MAX ? DBL_MAX : threshold);
    const int blocks = s->sm_count * 4;
    std::vector<unsigned long long> hist(kBins);
    // the array the digit passes read: all n scores at first, the compacted candidates later
    const double* cur_score = score_dev;
    const int64_t* cur_map = nullptr;
    int64_t cur_n = n;
    auto run_hist = [&](unsigned long long kp, unsigned long long km, int on_idx, unsigned long long ip,
                        unsigned long long im, int shift, int bits) -> int {
        ONILDA_CUDA(cudaMemsetAsync(s->d_hist, 0, sizeof(unsigned long long) * kBins, st));
        select_hist_kernel<<<blocks, 256, 0, st>>>(cur_score, cur_map, cur_n, max_key, kp, km, on_idx, ip, im, shift,
                                                   (1u << bits) - 1u, s->d_hist);
        ONILDA_CUDA(cudaGetLastError());
        ONILDA_CUDA(cudaMemcpyAsync(hist.data(), s->d_hist, sizeof(unsigned long long) * kBins, cudaMemcpyDeviceToHost, st));
        ONILDA_CUDA(cudaStreamSynchronize(st));
        s->launches++;
        return ONILDA_OK;
    };
    // returns the bin holding the `remaining`-th item and subtracts what lies below it
    auto pick = [&](unsigned long long& remaining) -> unsigned long long {
        unsigned long long acc = 0;
        int b = 0;
        for (; b < kBins - 1; b++) { if (acc + hist[(size_t)b] >= remaining) break; acc += hist[(size_t)b]; }
        remaining -= acc;
        return (unsigned long long)b;
    };
    static const int key_pass[6][2] = { {53, 11}, {42, 11}, {31, 11}, {20, 11}, {9, 11}, {0, 9} };
    static const int idx_pass[4][2] = { {33, 11}, {22, 11}, {11, 11}, {0, 11} };   // indices < 2^44
    constexpr int64_t kMaxCandidates = (int64_t)1 << 26;

    unsigned long long cut_key = max_key, cut_idx = ~0ull;
    int64_t take = 0, already = 0;       // already: rows written to d_sel by the split pass
    int rc;
    if (n > 0) {
        // the first key pass also yields the number of events under the threshold
        rc = run_hist(0, 0, 0, 0, 0, key_pass[0][0], key_pass[0][1]);
        if (rc != ONILDA_OK) return rc;
        unsigned long long under = 0;
        for (unsigned long long h : hist) under += h;
        take = (k == -1 || (unsigned long long)k > under) ? (int64_t)under : k;
        if (take > 0) { rc = ensure_sel(s, take); if (rc != ONILDA_OK) return rc; }
        if (take > 0 && (unsigned long long)take < under) {
            unsigned long long remaining = (unsigned long long)take, kp = 0, km = 0;
            for (int p = 0; p < 6; p++) {
                const int shift = key_pass[p][0], bits = key_pass[p][1];
                if (p > 0) { rc = run_hist(kp, km, 0, 0, 0, shift, bits); if (rc != ONILDA_OK) return rc; }
                const unsigned long long b = pick(remaining);
                const unsigned long long in_bin = hist[(size_t)b];
                kp |= b << shift;
                km |= ((1ull << bits) - 1ull) << shift;
                // two digits fixed and the undecided set is small next to n: split once, then work on the short list
                if (p == 1 && cur_map == nullptr && (int64_t)in_bin <= kMaxCandidates && (int64_t)in_bin * 4 <= n) {
                    rc = ensure_cand(s, (int64_t)in_bin);
                    if (rc != ONILDA_OK) return rc;
                    ONILDA_CUDA(cudaMemsetAsync(s->d_hist, 0, 2 * sizeof(unsigned long long), st));
                    select_split_kernel<<<blocks, 256, 0, st>>>(score_dev, n, max_key, kp, km, take, (int64_t)in_bin, s->d_hist,
                                                                s->d_sel_idx, s->d_sel_score, s->d_cand_idx, s->d_cand_score);
                    ONILDA_CUDA(cudaGetLastError());
                    s->launches++;
                    unsigned long long cnt[2] = {0, 0};
                    ONILDA_CUDA(cudaMemcpyAsync(cnt, s->d_hist, sizeof cnt, cudaMemcpyDeviceToHost, st));
                    ONILDA_CUDA(cudaStreamSynchronize(st));
                    if (cnt[1] != in_bin || cnt[0] != (unsigned long long)take - remaining)
                        return onilda_fail(ONILDA_ECUDA, "top-k split pass: %llu below / %llu candidates, expected %llu / %llu",
                                           cnt[0], cnt[1], (unsigned long long)take - remaining, in_bin);
                    already = (int64_t)cnt[0];
                    cur_score = s->d_cand_score;
                    cur_map = s->d_cand_idx;
                    cur_n = (int64_t)in_bin;
                }
            }
            cut_key = kp;
            // among events whose score equals the cut, keep the `remaining` smallest indices
            unsigned long long ip = 0, im = 0;
            for (int p = 0; p < 4; p++) {
                const int shift = idx_pass[p][0], bits = idx_pass[p][1];
                rc = run_hist(cut_key, ~0ull, 1, ip, im, shift, bits);
                if (rc != ONILDA_OK) return rc;
                const unsigned long long b = pick(remaining);
                ip |= b << shift;
                im |= ((1ull << bits) - 1ull) << shift;
            }
            cut_idx = ip;
        }
    }
    *n_out = take;
    if (take == 0) return ONILDA_OK;
    // collect: from the candidate list (appending to what the split pass already wrote) or from all scores
    unsigned long long start = (unsigned long long)already;
    ONILDA_CUDA(cudaMemcpyAsync(s->d_hist, &start, sizeof start, cudaMemcpyHostToDevice, st));
    select_collect_kernel<<<blocks, 256, 0, st>>>(cur_score, cur_map, cur_n, max_key, cut_key, cut_idx, take, s->d_hist,
                                                  s->d_sel_idx, s->d_sel_score);
    ONILDA_CUDA(cudaGetLastError());
    s->launches++;
    std::vector<int64_t> idx((size_t)take);
    std::vector<double> sc((size_t)take);
    unsigned long long got = 0;
    ONILDA_CUDA(cudaMemcpyAsync(&got, s->d_hist, sizeof got, cudaMemcpyDeviceToHost, st));
    ONILDA_CUDA(cudaMemcpyAsync(idx.data(), s->d_sel_idx, sizeof(int64_t) * (size_t)take, cudaMemcpyDeviceToHost, st));
    ONILDA_CUDA(cudaMemcpyAsync(sc.data(), s->d_sel_score, sizeof(double) * (size_t)take, cudaMemcpyDeviceToHost, st));
    ONILDA_CUDA(cudaStreamSynchronize(st));
    if ((int64_t)got != take)
        return onilda_fail(ONILDA_ECUDA, "top-k selection collected %llu rows, expected %lld", got, (long long)take);
    // ascending by (score, index): the at most `take` survivors are ordered on the host
    std::vector<int64_t> perm((size_t)take);
    for (int64_t i = 0; i < take; i++) perm[(size_t)i] = i;
    std::sort(perm.begin(), perm.end(), [&](int64_t a, int64_t b) {
        if (sc[(size_t)a] != sc[(size_t)b]) return sc[(size_t)a] < sc[(size_t)b];
        return idx[(size_t)a] < idx[(size_t)b];
    });
    for (int64_t i = 0; i < take; i++) { idx_out[i] = idx[(size_t)perm[(size_t)i]]; score_out[i] = sc[(size_t)perm[(size_t)i]]; }
    return ONILDA_OK;
}

size_t score_smem_bytes(int K) { return sizeof(double) * (size_t)kScoreWarps * 2 * 32 * (size_t)(K | 1); }
size_t score_pairs_smem_bytes(int K) { return sizeof(double) * (size_t)kScoreWarps * (32 * (size_t)(K + 2) + 64); }
bool use_pairs_kernel(const onilda_scorer* s) { return (s->K % 2) == 0 && s->K >= 2 && s->K <= 64 && s->D < 0x7fffffff; }

int launch_score(onilda_scorer* s, cudaStream_t st, const int32_t* sdoc, const int32_t* sword, const int32_t* ddoc,
                 const int32_t* dword, int64_t n, double* out)
{
    if (n <= 0) return ONILDA_OK;
    int64_t tiles = (n + 31) / 32;
    int64_t blocks = (tiles + kScoreWarps - 1) / kScoreWarps;
    const int64_t cap = (int64_t)s->sm_count * 8;
    if (blocks > cap) blocks = cap;
    if (use_pairs_kernel(s)) {
        const size_t sh = score_pairs_smem_bytes(s->K);
#define ONILDA_PAIRS_LAUNCH(FLOWV, CCV) \
        score_pairs_kernel<FLOWV, CCV><<<(unsigned)blocks, kScoreThreads, sh, st>>>(s->d_theta, s->D, s->d_phi, s->V, s->K, sdoc, sword, ddoc, dword, n, out)
        if (s->K == 20) { if (ddoc) ONILDA_PAIRS_LAUNCH(true, 10); else ONILDA_PAIRS_LAUNCH(false, 10); }
        else if (s->K == 50) { if (ddoc) ONILDA_PAIRS_LAUNCH(true, 25); else ONILDA_PAIRS_LAUNCH(false, 25); }
        else { if (ddoc) ONILDA_PAIRS_LAUNCH(true, 0); else ONILDA_PAIRS_LAUNCH(false, 0); }
#undef ONILDA_PAIRS_LAUNCH
        ONILDA_CUDA(cudaGetLastError());
        s->launches++;
        return ONILDA_OK;
    }
    const size_t shmem = score_smem_bytes(s->K);
    if (ddoc) score_kernel<true><<<(unsigned)blocks, kScoreThreads, shmem, st>>>(s->d_theta, s->D, s->d_phi, s->V, s->K, sdoc, sword, ddoc, dword, n, out);
    else score_kernel<false><<<(unsigned)blocks, kScoreThreads, shmem, st>>>(s->d_theta, s->D, s->d_phi, s->V, s->K, sdoc, sword, nullptr, nullptr, n, out);
    ONILDA_CUDA(cudaGetLastError());
    s->launches++;
    return ONILDA_OK;
}

}  // namespace

extern "C" void onilda_scorer_destroy(onilda_scorer* s)
{
    if (!s) return;
    DeviceGuard g(s->device);
    cudaFree(s->d_theta); cudaFree(s->d_phi);
    for (int i = 0; i < 2; i++) { cudaFree(s->d_ids[i]); cudaFree(s->d_out[i]); if (s->stream[i]) cudaStreamDestroy(s->stream[i]); }
    cudaFree(s->d_hist); cudaFree(s->d_sel_idx); cudaFree(s->d_sel_score); cudaFree(s->d_cand_idx); cudaFree(s->d_cand_score);
    if (s->ev0) cudaEventDestroy(s->ev0);
    if (s->ev1) cudaEventDestroy(s->ev1);
    free(s);
}

extern "C" int onilda_scorer_create(const double* theta, int64_t num_docs, const double* phi, int32_t num_terms,
                                    int num_topics, int device, onilda_scorer** out)
{
    if (!out || num_docs < 0 || num_terms < 0 || num_topics <= 0 || (num_docs > 0 && !theta) || (num_terms > 0 && !phi))
        return onilda_fail(ONILDA_EINVAL, "scorer_create: bad argument");
    if (onilda_device_count() == 0)
        return onilda_fail(ONILDA_ECUDA, "scorer_create: no CUDA device is usable and this library has no CPU path");
    if (device < 0) ONILDA_CUDA(cudaGetDevice(&device));
    DeviceGuard g(device);
    onilda_scorer* s = (onilda_scorer*)calloc(1, sizeof *s);
    if (!s) return onilda_fail(ONILDA_ENOMEM, "scorer_create: out of host memory");
    s->device = device; s->K = num_topics; s->D = num_docs; s->V = num_terms;
    s->chunk = 1 << 24;
    const int K = num_topics;
    cudaError_t e = cudaSuccess;
#define STEP(x) if (e == cudaSuccess) e = (x)
    STEP(cudaDeviceGetAttribute(&s->sm_count, cudaDevAttrMultiProcessorCount, device));
    STEP(cudaStreamCreateWithFlags(&s->stream[0], cudaStreamNonBlocking));
    STEP(cudaStreamCreateWithFlags(&s->stream[1], cudaStreamNonBlocking));
    STEP(cudaEventCreate(&s->ev0));
    STEP(cudaEventCreate(&s->ev1));
    STEP(cudaMalloc((void**)&s->d_theta, sizeof(double) * (size_t)(num_docs + 1) * K));
    STEP(cudaMalloc((void**)&s->d_phi, sizeof(double) * (size_t)(num_terms + 1) * K));
    STEP(cudaMalloc((void**)&s->d_hist, sizeof(unsigned long long) * (kBins + 8)));
    if (num_docs > 0) STEP(cudaMemcpyAsync(s->d_theta, theta, sizeof(double) * (size_t)num_docs * K, cudaMemcpyHostToDevice, s->stream[0]));
    if (num_terms > 0) STEP(cudaMemcpyAsync(s->d_phi, phi, sizeof(double) * (size_t)num_terms * K, cudaMemcpyHostToDevice, s->stream[0]));
    if (e == cudaSuccess) {
        fill_kernel<<<(K + 63) / 64, 64, 0, s->stream[0]>>>(s->d_theta + (size_t)num_docs * K, K, 1.0 / K);
        fill_kernel<<<(K + 63) / 64, 64, 0, s->stream[0]>>>(s->d_phi + (size_t)num_terms * K, K, 1.0 / K);
        e = cudaGetLastError();
    }
    // phi (16 MB at V = 100k, K = 20) is read by every event; ask L2 to keep it (a persisting access-policy window on
    // the scorer's streams) so that the theta rows of rarely seen documents, streaming through, do not evict it.
    // ONILDA_SCORE_L2_PERSIST=0 turns it off (for measurements).
    {
        const char* env = getenv("ONILDA_SCORE_L2_PERSIST");
        int max_persist = 0, max_window = 0;
        cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, device);
        cudaDeviceGetAttribute(&max_window, cudaDevAttrMaxAccessPolicyWindowSize, device);
        size_t bytes = sizeof(double) * (size_t)(num_terms + 1) * K;
        if (!(env && atoi(env) == 0) && e == cudaSuccess && max_persist > 0 && max_window > 0 && bytes > 0) {
            if (bytes > (size_t)max_window) bytes = (size_t)max_window;
            size_t carve = bytes < (size_t)max_persist ? bytes : (size_t)max_persist;
            if (cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, carve) == cudaSuccess) {
                cudaStreamAttrValue av;
                memset(&av, 0, sizeof av);
                av.accessPolicyWindow.base_ptr = s->d_phi;
                av.accessPolicyWindow.num_bytes = bytes;
                av.accessPolicyWindow.hitRatio = carve >= bytes ? 1.0f : (float)carve / (float)bytes;
                av.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
                av.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
                for (int i = 0; i < 2; i++) cudaStreamSetAttribute(s->stream[i], cudaStreamAttributeAccessPolicyWindow, &av);
            }
            cudaGetLastError();      // best effort: a refusal leaves the default policy
        }
    }
    const size_t shmem = score_smem_bytes(K);
    STEP(cudaFuncSetAttribute(score_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)shmem));
    STEP(cudaFuncSetAttribute(score_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)shmem));
    {
        const int sh = (int)score_pairs_smem_bytes(K);
        STEP(cudaFuncSetAttribute(score_pairs_kernel<true, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, sh));
        STEP(cudaFuncSetAttribute(score_pairs_kernel<false, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, sh));
        STEP(cudaFuncSetAttribute(score_pairs_kernel<true, 10>, cudaFuncAttributeMaxDynamicSharedMemorySize, sh));
        STEP(cudaFuncSetAttribute(score_pairs_kernel<false, 10>, cudaFuncAttributeMaxDynamicSharedMemorySize, sh));
        STEP(cudaFuncSetAttribute(score_pairs_kernel<true, 25>, cudaFuncAttributeMaxDynamicSharedMemorySize, sh));
        STEP(cudaFuncSetAttribute(score_pairs_kernel<false, 25>, cudaFuncAttributeMaxDynamicSharedMemorySize, sh));
    }
    STEP(cudaStreamSynchronize(s->stream[0]));
#undef STEP
    if (e != cudaSuccess) {
        int rc = onilda_fail(e == cudaErrorMemoryAllocation ? ONILDA_ENOMEM : ONILDA_ECUDA, "scorer_create: %s", cudaGetErrorString(e));
        cudaGetLastError();
        onilda_scorer_destroy(s);
        return rc;
    }
    *out = s;
    return ONILDA_OK;
}

// theta / phi taken from an EM handle on the device: replaces writing final.gamma / final.beta, reading them
// back and normalising on the host (OniLDACWrapper.scala:88-144, 170-195) when EM and scoring share a process
extern "C" int onilda_scorer_create_from_em(onilda_em* em, onilda_scorer** out)
{
    if (!em || !out) return onilda_fail(ONILDA_EINVAL, "scorer_create_from_em: null argument");
    int device = 0, K = 0, V = 0, kpad = 0;
    int64_t D = 0;
    const double *gamma_dev = nullptr, *lb_dev = nullptr;
    int rc = onilda_em_device_views(em, &device, &D, &K, &V, &kpad, &gamma_dev, &lb_dev);
    if (rc != ONILDA_OK) return rc;
    if (K > 256) return onilda_fail(ONILDA_EINVAL, "scorer_create_from_em: %d topics unsupported", K);
    // an empty scorer of the right shape (uniform rows, streams, scratch), then the tables are filled on the device
    rc = onilda_scorer_create(nullptr, 0, nullptr, 0, K, device, out);
    if (rc != ONILDA_OK) return rc;
    onilda_scorer* s = *out;
    DeviceGuard g(device);
    cudaFree(s->d_theta); cudaFree(s->d_phi);
    s->d_theta = nullptr; s->d_phi = nullptr;
    s->D = D; s->V = V;
    double *partial = nullptr, *sums = nullptr;
    cudaError_t e = cudaMalloc((void**)&s->d_theta, sizeof(double) * (size_t)(D + 1) * K);
    if (e == cudaSuccess) e = cudaMalloc((void**)&s->d_phi, sizeof(double) * (size_t)(V + 1) * K);
    if (e == cudaSuccess) e = cudaMalloc((void**)&partial, sizeof(double) * (size_t)kPhiBlocks * K);
    if (e == cudaSuccess) e = cudaMalloc((void**)&sums, sizeof(double) * (size_t)K);
    if (e == cudaSuccess) {
        cudaStream_t st = s->stream[0];
        if (D > 0) theta_kernel<<<(unsigned)((D + 255) / 256), 256, 0, st>>>(gamma_dev, D, K, s->d_theta);
        if (V > 0) {
            phi_partial_wk_kernel<<<kPhiBlocks, 256, sizeof(double) * (size_t)(256 / K) * K, st>>>(lb_dev, V, K, kpad, partial);
            phi_final_wk_kernel<<<(K + 63) / 64, 64, 0, st>>>(partial, kPhiBlocks, K, sums);
            phi_norm_wk_kernel<<<(unsigned)(((int64_t)V * K + 255) / 256), 256, 0, st>>>(lb_dev, sums, V, K, kpad, s->d_phi);
        }
        fill_kernel<<<(K + 63) / 64, 64, 0, st>>>(s->d_theta + (size_t)D * K, K, 1.0 / K);
        fill_kernel<<<(K + 63) / 64, 64, 0, st>>>(s->d_phi + (size_t)V * K, K, 1.0 / K);
        e = cudaGetLastError();
        if (e == cudaSuccess) e = cudaStreamSynchronize(st);
        s->launches += 6;
    }
    cudaFree(partial); cudaFree(sums);
    if (e != cudaSuccess) {
        rc = onilda_fail(e == cudaErrorMemoryAllocation ? ONILDA_ENOMEM : ONILDA_ECUDA, "scorer_create_from_em: %s", cudaGetErrorString(e));
        cudaGetLastError();
        onilda_scorer_destroy(s);
        *out = nullptr;
        return rc;
    }
    return ONILDA_OK;
}

static int ensure_staging(onilda_scorer* s)
{
    if (s->d_ids[0]) return ONILDA_OK;
    for (int i = 0; i < 2; i++) {
        ONILDA_CUDA(cudaMalloc((void**)&s->d_ids[i], sizeof(int32_t) * 4 * (size_t)s->chunk));
        ONILDA_CUDA(cudaMalloc((void**)&s->d_out[i], sizeof(double) * (size_t)s->chunk));
    }
    return ONILDA_OK;
}

// host buffers: chunks alternate between two streams so copies overlap the kernel
static int score_host(onilda_scorer* s, const int32_t* sdoc, const int32_t* sword, const int32_t* ddoc,
                      const int32_t* dword, int64_t n, double* out)
{
    if (!s || n < 0 || (n > 0 && (!sdoc || !sword || !out))) return onilda_fail(ONILDA_EINVAL, "score: bad argument");
    DeviceGuard g(s->device);
    int rc = ensure_staging(s);
    if (rc != ONILDA_OK) return rc;
    const int64_t C = s->chunk;
    int which = 0;
    ONILDA_CUDA(cudaEventRecord(s->ev0, s->stream[0]));
    for (int64_t lo = 0; lo < n; lo += C, which ^= 1) {
        const int64_t m = (n - lo < C) ? n - lo : C;
        cudaStream_t st = s->stream[which];
        int32_t* ids = s->d_ids[which];
        ONILDA_CUDA(cudaMemcpyAsync(ids, sdoc + lo, sizeof(int32_t) * (size_t)m, cudaMemcpyHostToDevice, st));
        ONILDA_CUDA(cudaMemcpyAsync(ids + C, sword + lo, sizeof(int32_t) * (size_t)m, cudaMemcpyHostToDevice, st));
        if (ddoc) {
            ONILDA_CUDA(cudaMemcpyAsync(ids + 2 * C, ddoc + lo, sizeof(int32_t) * (size_t)m, cudaMemcpyHostToDevice, st));
            ONILDA_CUDA(cudaMemcpyAsync(ids + 3 * C, dword + lo, sizeof(int32_t) * (size_t)m, cudaMemcpyHostToDevice, st));
        }
        rc = launch_score(s, st, ids, ids + C, ddoc ? ids + 2 * C : nullptr, ddoc ? ids + 3 * C : nullptr, m, s->d_out[which]);
        if (rc != ONILDA_OK) return rc;
        ONILDA_CUDA(cudaMemcpyAsync(out + lo, s->d_out[which], sizeof(double) * (size_t)m, cudaMemcpyDeviceToHost, st));
    }
    ONILDA_CUDA(cudaStreamSynchronize(s->stream[1]));
    ONILDA_CUDA(cudaEventRecord(s->ev1, s->stream[0]));
    ONILDA_CUDA(cudaStreamSynchronize(s->stream[0]));
    float ms = 0;
    if (cudaEventElapsedTime(&ms, s->ev0, s->ev1) == cudaSuccess) s->kernel_ms = ms;
    return ONILDA_OK;
}

extern "C" int onilda_score(onilda_scorer* s, const int32_t* doc, const int32_t* word, int64_t n, double* out)
{
    return score_host(s, doc, word, nullptr, nullptr, n, out);
}

extern "C" int onilda_score_flow_min(onilda_scorer* s, const int32_t* sdoc, const int32_t* sword, const int32_t* ddoc,
                                     const int32_t* dword, int64_t n, double* out)
{
    if (n > 0 && (!ddoc || !dword)) return onilda_fail(ONILDA_EINVAL, "score_flow_min: null destination ids");
    return score_host(s, sdoc, sword, ddoc, dword, n, out);
}

extern "C" int onilda_topk(onilda_scorer* s, const double* score, int64_t n, double threshold, int64_t k,
                           int64_t* idx_out, double* score_out, int64_t* n_out)
{
    if (!s || n < 0 || (n > 0 && !score) || !n_out || k < -1) return onilda_fail(ONILDA_EINVAL, "topk: bad argument");
    DeviceGuard g(s->device);
    double* d_score = nullptr;
    if (n > 0) {
        ONILDA_CUDA(cudaMalloc((void**)&d_score, sizeof(double) * (size_t)n));
        cudaError_t e = cudaMemcpyAsync(d_score, score, sizeof(double) * (size_t)n, cudaMemcpyHostToDevice, s->stream[0]);
        if (e != cudaSuccess) { cudaFree(d_score); return onilda_fail(ONILDA_ECUDA, "topk: %s", cudaGetErrorString(e)); }
    }
    int rc = select_topk_device(s, d_score, n, threshold, k, idx_out, score_out, n_out);
    cudaFree(d_score);
    return rc;
}

extern "C" int onilda_score_flow_topk_device(onilda_scorer* s, const int32_t* sdoc_dev, const int32_t* sword_dev,
                                             const int32_t* ddoc_dev, const int32_t* dword_dev, int64_t n,
                                             double threshold, int64_t k, int64_t* idx_out, double* score_out,
                                             int64_t* n_out, double* score_dev_out)
{
    if (!s || n < 0 || !n_out || k < -1 || (n > 0 && (!sdoc_dev || !sword_dev)))
        return onilda_fail(ONILDA_EINVAL, "score_flow_topk_device: bad argument");
    DeviceGuard g(s->device);
    double* d_score = score_dev_out;
    if (!d_score && n > 0) ONILDA_CUDA(cudaMalloc((void**)&d_score, sizeof(double) * (size_t)n));
    cudaStream_t st = s->stream[0];
    cudaEventRecord(s->ev0, st);
    int rc = launch_score(s, st, sdoc_dev, sword_dev, ddoc_dev, dword_dev, n, d_score);
    cudaEventRecord(s->ev1, st);
    if (rc == ONILDA_OK) rc = select_topk_device(s, d_score, n, threshold, k, idx_out, score_out, n_out);
    if (rc == ONILDA_OK) {
        cudaStreamSynchronize(st);
        float ms = 0;
        if (cudaEventElapsedTime(&ms, s->ev0, s->ev1) == cudaSuccess) s->kernel_ms = ms;
    }
    if (!score_dev_out) cudaFree(d_score);
    return rc;
}

// FlowPostLDA.scala:75-80 in one call: both role scores, their minimum, filter(min <= threshold), orderBy(min),
// limit(k); what comes back is what the results file needs per kept row (the row index to join the event's own
// columns, SourceScore and DestinationScore - the minimum itself is dropped from the file, :80).
extern "C" int onilda_flow_results(onilda_scorer* s, const int32_t* sdoc, const int32_t* sword, const int32_t* ddoc,
                                   const int32_t* dword, int64_t n, double threshold, int64_t k, int64_t* idx_out,
                                   double* src_score_out, double* dst_score_out, int64_t* n_out)
{
    if (!s || n < 0 || !n_out || k < -1 || (n > 0 && (!sdoc || !sword || !ddoc || !dword)))
        return onilda_fail(ONILDA_EINVAL, "flow_results: bad argument");
    *n_out = 0;
    if (n == 0 || k == 0) return ONILDA_OK;
    if (!idx_out || !src_score_out || !dst_score_out) return onilda_fail(ONILDA_EINVAL, "flow_results: null output");
    DeviceGuard g(s->device);
    int32_t* d_ids = nullptr;
    double* d_score = nullptr;
    const size_t b = (sizeof(int32_t) * (size_t)n + 255) & ~(size_t)255;
    if (cudaMalloc((void**)&d_ids, 4 * b) != cudaSuccess || cudaMalloc((void**)&d_score, sizeof(double) * (size_t)n) != cudaSuccess) {
        cudaGetLastError();
        cudaFree(d_ids);
        return onilda_fail(ONILDA_ENOMEM, "flow_results: device memory for %lld events not available", (long long)n);
    }
    struct Free { void *a, *b; ~Free() { cudaFree(a); cudaFree(b); } } fr{d_ids, d_score};
    const int32_t* host[4] = {sdoc, sword, ddoc, dword};
    int32_t* dev[4];
    for (int c = 0; c < 4; c++) {
        dev[c] = (int32_t*)((char*)d_ids + (size_t)c * b);
        ONILDA_CUDA(cudaMemcpyAsync(dev[c], host[c], sizeof(int32_t) * (size_t)n, cudaMemcpyHostToDevice, s->stream[0]));
    }
    const int64_t cap = k == -1 ? n : (k < n ? k : n);
    std::vector<double> min_score((size_t)cap);
    int rc = onilda_score_flow_topk_device(s, dev[0], dev[1], dev[2], dev[3], n, threshold, k, idx_out, min_score.data(), n_out, d_score);
    if (rc != ONILDA_OK) return rc;
    // the two role scores of the kept rows (at most k of them): the same kernel on the gathered ids
    const int64_t m = *n_out;
    std::vector<int32_t> ids((size_t)(4 * m > 0 ? 4 * m : 1));
    for (int64_t r = 0; r < m; r++)
        for (int c = 0; c < 4; c++) ids[(size_t)(c * m + r)] = host[c][idx_out[r]];
    if (m > 0) {
        rc = onilda_score(s, ids.data(), ids.data() + m, m, src_score_out);
        if (rc == ONILDA_OK) rc = onilda_score(s, ids.data() + 2 * m, ids.data() + 3 * m, m, dst_score_out);
    }
    return rc;
}

// dns / proxy: one score per event (DNSPostLDA.scala:44, 75-92; ProxySuspiciousConnectsModel.scala:41-71), then
// filter(score <= threshold), takeOrdered(k) with k == -1 meaning all, ascending (DNSPostLDA.scala:49-70,
// ProxySuspiciousConnectsAnalysis.scala:45-51, DataFrameUtils.scala:21-38) in one call.  Ties, which Spark leaves in
// no particular order, come out by event index.
extern "C" int onilda_results_rows(onilda_scorer* s, const int32_t* doc, const int32_t* word, int64_t n, double threshold,
                                   int64_t k, int64_t* idx_out, double* score_out, int64_t* n_out)
{
    if (!s || n < 0 || !n_out || k < -1 || (n > 0 && (!doc || !word)))
        return onilda_fail(ONILDA_EINVAL, "results_rows: bad argument");
    *n_out = 0;
    if (n == 0 || k == 0) return ONILDA_OK;
    if (!idx_out || !score_out) return onilda_fail(ONILDA_EINVAL, "results_rows: null output");
    DeviceGuard g(s->device);
    int32_t* d_ids = nullptr;
    double* d_score = nullptr;
    const size_t b = (sizeof(int32_t) * (size_t)n + 255) & ~(size_t)255;
    if (cudaMalloc((void**)&d_ids, 2 * b) != cudaSuccess || cudaMalloc((void**)&d_score, sizeof(double) * (size_t)n) != cudaSuccess) {
        cudaGetLastError();
        cudaFree(d_ids);
        return onilda_fail(ONILDA_ENOMEM, "results_rows: device memory for %lld events not available", (long long)n);
    }
    struct Free { void *a, *b; ~Free() { cudaFree(a); cudaFree(b); } } fr{d_ids, d_score};
    int32_t* d_doc = d_ids;
    int32_t* d_word = (int32_t*)((char*)d_ids + b);
    ONILDA_CUDA(cudaMemcpyAsync(d_doc, doc, sizeof(int32_t) * (size_t)n, cudaMemcpyHostToDevice, s->stream[0]));
    ONILDA_CUDA(cudaMemcpyAsync(d_word, word, sizeof(int32_t) * (size_t)n, cudaMemcpyHostToDevice, s->stream[0]));
    return onilda_score_flow_topk_device(s, d_doc, d_word, nullptr, nullptr, n, threshold, k, idx_out, score_out, n_out, d_score);
}

extern "C" int onilda_scorer_stats(onilda_scorer* s, double* kernel_ms, int64_t* launches)
{
    if (!s) return onilda_fail(ONILDA_EINVAL, "scorer_stats: null handle");
    if (kernel_ms) *kernel_ms = s->kernel_ms;
    if (launches) *launches = s->launches;
    return ONILDA_OK;
}

extern "C" int onilda_theta_from_gamma(const double* gamma, int64_t num_docs, int num_topics, int device, double* theta_out)
{
    if (num_docs < 0 || num_topics <= 0 || (num_docs > 0 && (!gamma || !theta_out)))
        return onilda_fail(ONILDA_EINVAL, "theta_from_gamma: bad argument");
    if (num_docs == 0) return ONILDA_OK;
    if (onilda_device_count() == 0) return onilda_fail(ONILDA_ECUDA, "theta_from_gamma: no CUDA device is usable");
    if (device < 0) ONILDA_CUDA(cudaGetDevice(&device));
    DeviceGuard g(device);
    const size_t bytes = sizeof(double) * (size_t)num_docs * num_topics;
    double *d_in = nullptr, *d_out = nullptr;
    ONILDA_CUDA(cudaMalloc((void**)&d_in, bytes));
    if (cudaMalloc((void**)&d_out, bytes) != cudaSuccess) { cudaFree(d_in); cudaGetLastError(); return onilda_fail(ONILDA_ENOMEM, "theta_from_gamma: device memory"); }
    cudaError_t e = cudaMemcpy(d_in, gamma, bytes, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
        theta_kernel<<<(unsigned)((num_docs + 255) / 256), 256>>>(d_in, num_docs, num_topics, d_out);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpy(theta_out, d_out, bytes, cudaMemcpyDeviceToHost);
    cudaFree(d_in); cudaFree(d_out);
    if (e != cudaSuccess) return onilda_fail(ONILDA_ECUDA, "theta_from_gamma: %s", cudaGetErrorString(e));
    return ONILDA_OK;
}

extern "C" int onilda_phi_from_log_beta(const double* log_beta, int num_topics, int num_terms, int device, double* phi_out)
{
    if (num_topics <= 0 || num_terms < 0 || (num_terms > 0 && (!log_beta || !phi_out)))
        return onilda_fail(ONILDA_EINVAL, "phi_from_log_beta: bad argument");
    if (num_terms == 0) return ONILDA_OK;
    if (onilda_device_count() == 0) return onilda_fail(ONILDA_ECUDA, "phi_from_log_beta: no CUDA device is usable");
    if (device < 0) ONILDA_CUDA(cudaGetDevice(&device));
    DeviceGuard g(device);
    const size_t bytes = sizeof(double) * (size_t)num_topics * num_terms;
    double *d_in = nullptr, *d_out = nullptr, *d_sum = nullptr;
    ONILDA_CUDA(cudaMalloc((void**)&d_in, bytes));
    if (cudaMalloc((void**)&d_out, bytes) != cudaSuccess || cudaMalloc((void**)&d_sum, sizeof(double) * num_topics) != cudaSuccess) {
        cudaFree(d_in); cudaFree(d_out); cudaGetLastError();
        return onilda_fail(ONILDA_ENOMEM, "phi_from_log_beta: device memory");
    }
    cudaError_t e = cudaMemcpy(d_in, log_beta, bytes, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
        phi_sum_kernel<<<num_topics, 256>>>(d_in, num_terms, d_sum);
        phi_norm_kernel<<<(num_terms + 255) / 256, 256>>>(d_in, d_sum, num_topics, num_terms, d_out);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpy(phi_out, d_out, bytes, cudaMemcpyDeviceToHost);
    cudaFree(d_in); cudaFree(d_out); cudaFree(d_sum);
    if (e != cudaSuccess) return onilda_fail(ONILDA_ECUDA, "phi_from_log_beta: %s", cudaGetErrorString(e));
    return ONILDA_OK;
}

// ---------------------------------------------------------------------------------------------
// Top-k without any allocation.
//
// Some callers of the scorer may not allocate device memory at the moment they need a result:
//   * a Spark executor (the reference's post-LDA scoring runs inside one) whose GPU memory has
//     been handed in full to a pooling allocator at start-up - a cudaMalloc from this library
//     then fails with out-of-memory although the pool still has room;
//   * a streaming scorer that ranks every micro-batch and must not stall the device -
//     cudaMalloc and cudaFree synchronise the whole GPU, and they are illegal while a stream
//     is being captured.
// onilda_topk / onilda_score_flow_topk_device allocate (a device copy of the scores, candidate
// buffers that grow on demand, a scorer handle with its own streams).  The entry point below
// allocates nothing, on the device or on the host heap: the radix-select histogram, the
// candidate counter and the at most ONILDA_NOALLOC_MAX_K selected rows live in the file-scope
// array g_topk_scratch (82 KB, held by every GPU that loads the library until the process
// exits); host temporaries are on the stack; all work goes to the legacy default stream.
//
// In host code the name g_topk_scratch is only a stand-in, not a device address: the address the
// kernels need is asked for once per device with cudaGetSymbolAddress (with that device
// current) and handed to them as an ordinary pointer argument.
namespace {

constexpr int64_t kNoAllocMaxK = 4096;     // ONILDA_NOALLOC_MAX_K in onilda.h

struct TopkScratch {
    unsigned long long hist[kBins];        // one radix digit's histogram
    unsigned long long counter[32];        // collect-pass row counter (padded to 256 B)
    int64_t idx[kNoAllocMaxK];             // selected rows, unordered
    double score[kNoAllocMaxK];
};

static __device__ __align__(256) unsigned char g_topk_scratch[sizeof(TopkScratch)];

constexpr int kMaxDevices = 64;
TopkScratch* g_topk_scratch_addr[kMaxDevices];   // per device; filled on the first call on that device

struct ScoredRow {
    double s;
    int64_t i;
};

}  // namespace

extern "C" int onilda_topk_device_noalloc(const double* score_dev, int64_t n, double threshold, int64_t k,
                                          int64_t* idx_out, double* score_out, int64_t* n_out)
{
    if (n < 0 || (n > 0 && !score_dev) || !n_out || (k > 0 && (!idx_out || !score_out)))
        return onilda_fail(ONILDA_EINVAL, "topk_device_noalloc: bad argument");
    if (k < 0 || k > kNoAllocMaxK)
        return onilda_fail(ONILDA_EINVAL,
                           "topk_device_noalloc: k = %lld is outside 0..%lld, the most rows the fixed scratch holds "
                           "(use onilda_topk for more, or for k = -1)", (long long)k, (long long)kNoAllocMaxK);
    if (n >= ((int64_t)1 << 44))
        return onilda_fail(ONILDA_EINVAL, "topk_device_noalloc: more than 2^44 events");
    int dev = 0, sms = 0;
    ONILDA_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) return onilda_fail(ONILDA_EINVAL, "topk_device_noalloc: device index %d unsupported", dev);
    if (!g_topk_scratch_addr[dev]) {
        // first call on this device: resolve the device address of the scratch symbol
        void* p = nullptr;
        ONILDA_CUDA(cudaGetSymbolAddress(&p, g_topk_scratch));
        g_topk_scratch_addr[dev] = static_cast<TopkScratch*>(p);
    }
    TopkScratch* sc = g_topk_scratch_addr[dev];
    ONILDA_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const int blocks = sms * 4;
    // +infinity marks an event the flow join dropped (score_kernel): no threshold, however large, keeps it
    const unsigned long long max_key = isnan(threshold) ? 0ull : host_order_key(threshold > DBL_MAX ? DBL_MAX : threshold);
    unsigned long long hist[kBins];          // 16 KB on the stack
    *n_out = 0;
    if (n == 0 || k == 0) return ONILDA_OK;

    // one histogram pass: the scratch keeps the previous call's numbers, so it is cleared first
    auto run_hist = [&](unsigned long long kp, unsigned long long km, int on_idx, unsigned long long ip,
                        unsigned long long im, int shift, int bits) -> int {
        ONILDA_CUDA(cudaMemsetAsync(sc->hist, 0, sizeof sc->hist, 0));
        select_hist_kernel<<<blocks, 256, 0, 0>>>(score_dev, nullptr, n, max_key, kp, km, on_idx, ip, im, shift,
                                                  (1u << bits) - 1u, sc->hist);
        ONILDA_CUDA(cudaGetLastError());
        ONILDA_CUDA(cudaMemcpy(hist, sc->hist, sizeof hist, cudaMemcpyDeviceToHost));
        return ONILDA_OK;
    };
    auto pick = [&](unsigned long long& remaining) -> unsigned long long {
        unsigned long long acc = 0;
        int b = 0;
        for (; b < kBins - 1; b++) { if (acc + hist[b] >= remaining) break; acc += hist[b]; }
        remaining -= acc;
        return (unsigned long long)b;
    };
    static const int key_pass[6][2] = { {53, 11}, {42, 11}, {31, 11}, {20, 11}, {9, 11}, {0, 9} };
    static const int idx_pass[4][2] = { {33, 11}, {22, 11}, {11, 11}, {0, 11} };

    int rc = run_hist(0, 0, 0, 0, 0, key_pass[0][0], key_pass[0][1]);
    if (rc != ONILDA_OK) return rc;
    unsigned long long under = 0;
    for (int b = 0; b < kBins; b++) under += hist[b];
    const int64_t take = (unsigned long long)k > under ? (int64_t)under : k;
    unsigned long long cut_key = max_key, cut_idx = ~0ull;
    if (take > 0 && (unsigned long long)take < under) {
        unsigned long long remaining = (unsigned long long)take, kp = 0, km = 0;
        for (int p = 0; p < 6; p++) {
            const int shift = key_pass[p][0], bits = key_pass[p][1];
            if (p > 0) { rc = run_hist(kp, km, 0, 0, 0, shift, bits); if (rc != ONILDA_OK) return rc; }
            kp |= pick(remaining) << shift;
            km |= ((1ull << bits) - 1ull) << shift;
        }
        cut_key = kp;
        unsigned long long ip = 0, im = 0;
        for (int p = 0; p < 4; p++) {
            const int shift = idx_pass[p][0], bits = idx_pass[p][1];
            rc = run_hist(cut_key, ~0ull, 1, ip, im, shift, bits);
            if (rc != ONILDA_OK) return rc;
            ip |= pick(remaining) << shift;
            im |= ((1ull << bits) - 1ull) << shift;
        }
        cut_idx = ip;
    }
    if (take == 0) return ONILDA_OK;
    ONILDA_CUDA(cudaMemsetAsync(sc->counter, 0, sizeof sc->counter, 0));
    select_collect_kernel<<<blocks, 256, 0, 0>>>(score_dev, nullptr, n, max_key, cut_key, cut_idx, take, sc->counter, sc->idx, sc->score);
    ONILDA_CUDA(cudaGetLastError());
    unsigned long long got = 0;
    ONILDA_CUDA(cudaMemcpy(&got, sc->counter, sizeof got, cudaMemcpyDeviceToHost));
    if ((int64_t)got != take)
        return onilda_fail(ONILDA_ECUDA, "topk_device_noalloc: collected %llu rows, expected %lld", got, (long long)take);
    ONILDA_CUDA(cudaMemcpy(idx_out, sc->idx, sizeof(int64_t) * (size_t)take, cudaMemcpyDeviceToHost));
    ONILDA_CUDA(cudaMemcpy(score_out, sc->score, sizeof(double) * (size_t)take, cudaMemcpyDeviceToHost));
    // ascending by (score, index), in place, no heap: at most 4096 rows (64 KB) on the stack
    ScoredRow rows[kNoAllocMaxK];
    for (int64_t i = 0; i < take; i++) { rows[i].s = score_out[i]; rows[i].i = idx_out[i]; }
    std::sort(rows, rows + take, [](const ScoredRow& a, const ScoredRow& b) { return a.s != b.s ? a.s < b.s : a.i < b.i; });
    for (int64_t i = 0; i < take; i++) { score_out[i] = rows[i].s; idx_out[i] = rows[i].i; }
    *n_out = take;
    return ONILDA_OK;
}
