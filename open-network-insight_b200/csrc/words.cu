// Flow word creation on integer columns: quantile cuts, binning, port logic -> integer word codes.
//
// SURVEY.md section 8(f), row N4 (the per-event work upstream of the path).  Reference:
//   utilities/Quantiles.scala:21-99      computeEcdf / computeQuantiles / bin
//   netflow/FlowWordCreation.scala:14-84 time deciles, ibyt deciles, ipkt quintiles, the three bins
//   netflow/FlowWordCreation.scala:92-146 adjustPort -> word port, direction marker
//
// Quantiles: "the least x in the data with Pr(X <= x) >= t" = the k-th smallest value, k the least count with
// (double)k / (double)n >= t (the reference compares countToLeft / totalCount, doubles, with t).  No sort: the
// k-th smallest of an int64 column comes from a most-significant-digit-first radix select that follows all
// thresholds at once (one 2048-bin histogram per distinct prefix per level, every level one read of the column).
// The time of day h + m/60 + s/3600 is strictly increasing in the second of the day, so its deciles come from
// one 86400-bin histogram and the cut is evaluated by the reference's own expression.
#include <limits.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "kernels.cuh"

namespace {

constexpr int kSelBits = 11;
constexpr int kSelBins = 1 << kSelBits;
constexpr int kMaxQ = 16;

struct MinMax {
    long long lo, hi;
};

__global__ void __launch_bounds__(256) minmax_kernel(const int64_t* __restrict__ v, int64_t n, MinMax* __restrict__ out)
{
    long long lo = LLONG_MAX, hi = LLONG_MIN;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const long long x = v[i];
        lo = min(lo, x);
        hi = max(hi, x);
    }
    for (int o = 16; o > 0; o >>= 1) {
        lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o));
        hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o));
    }
    if ((threadIdx.x & 31) == 0) {
        atomicMin(&out->lo, lo);
        atomicMax(&out->hi, hi);
    }
}

struct SelLevel {
    int nprefix;
    int shift;                               // digit = (key >> shift) & (kSelBins - 1)
    unsigned long long prefix[kMaxQ];        // key >> (shift + kSelBits) must equal this
};

// hist[p][digit] += 1 for every key whose bits above the digit equal prefix p.  key = value - lo (unsigned).
// Up to 16 x 2048 shared counters (128 KB) allow one or two blocks per SM, so the blocks are 1024 threads wide and
// every thread keeps four loads in flight.
constexpr int kSelThreads = 1024;
constexpr int kSelUnroll = 4;

__device__ __forceinline__ void select_count(unsigned long long key, const SelLevel& lv, int up, unsigned int* sh, bool live)
{
    const unsigned long long hi = up >= 64 ? 0ull : key >> up;
    const unsigned int d = (unsigned int)(key >> lv.shift) & (kSelBins - 1);
    int slot = -1;
    if (live)
        for (int p = 0; p < lv.nprefix; p++)
            if (hi == lv.prefix[p]) slot = p * kSelBins + (int)d;
    // byte counts are heavy-tailed: most of a warp often lands in one bin, so equal slots add once.  The match is
    // on the whole slot: whichever lanes happen to arrive together, a group shares prefix and digit.
    const unsigned int peers = __match_any_sync(0xffffffffu, slot);
    if (slot >= 0 && (peers & ((1u << (threadIdx.x & 31)) - 1u)) == 0) atomicAdd(&sh[slot], (unsigned int)__popc(peers));
}

__global__ void __launch_bounds__(kSelThreads) multi_select_hist_kernel(const int64_t* __restrict__ v, int64_t n, long long lo, SelLevel lv,
                                                                        unsigned long long* __restrict__ hist)
{
    extern __shared__ unsigned int sh[];     // [nprefix][kSelBins]
    for (int b = threadIdx.x; b < lv.nprefix * kSelBins; b += blockDim.x) sh[b] = 0;
    __syncthreads();
    const int up = lv.shift + kSelBits;
    const int64_t stride = (int64_t)gridDim.x * kSelThreads;
    // whole warps stay in the loop together (the bound is rounded up to the stride), so the match is over full warps
    for (int64_t base = (int64_t)blockIdx.x * kSelThreads + threadIdx.x; base - threadIdx.x % 32 < n; base += stride * kSelUnroll) {
        long long x[kSelUnroll];
#pragma unroll
        for (int u = 0; u < kSelUnroll; u++) {
            const int64_t i = base + u * stride;
            x[u] = i < n ? v[i] : 0;
        }
#pragma unroll
        for (int u = 0; u < kSelUnroll; u++)
            select_count((unsigned long long)x[u] - (unsigned long long)lo, lv, up, sh, base + u * stride < n);
    }
    __syncthreads();
    for (int b = threadIdx.x; b < lv.nprefix * kSelBins; b += blockDim.x)
        if (sh[b]) atomicAdd(&hist[b], (unsigned long long)sh[b]);
}

__global__ void __launch_bounds__(256) time_hist_kernel(const int32_t* __restrict__ hour, const int32_t* __restrict__ minute,
                                                        const int32_t* __restrict__ second, int64_t n,
                                                        unsigned long long* __restrict__ hist /* [86400] */, int* __restrict__ bad)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const int h = hour[i], m = minute[i], s = second[i];
        if ((unsigned)h >= 24u || (unsigned)m >= 60u || (unsigned)s >= 60u) {
            *bad = 1;
            continue;
        }
        atomicAdd(&hist[h * 3600 + m * 60 + s], 1ull);
    }
}

struct FlowCutsDev {
    int time_key[10];            // second of the day of each time cut
    long long ibyt[10];
    long long ipkt[5];
};

// FlowWordCreation.scala:92-146.  code = (((port_index * 10 + timeBin) * 10 + ibytBin) * 5 + ipktBin) * 2 + marker,
// port_index = the port, 65536 for 111111, 65537 for 333333; marker 1 = the "-1_" prefix.
__device__ __forceinline__ void adjust_port(int dport, int sport, int tb, int bb, int pb, int32_t* src_code, int32_t* dst_code)
{
    int word_port, port_case;
    const int lo = min(dport, sport);
    if ((dport <= 1024 || sport <= 1024) && (dport > 1024 || sport > 1024) && lo != 0) {
        port_case = 2;
        word_port = lo;
    } else if (dport > 1024 && sport > 1024) {
        port_case = 3;
        word_port = 65537;                    // 333333
    } else if (dport == 0 && sport != 0) {
        port_case = 4;
        word_port = sport;
    } else if (sport == 0 && dport != 0) {
        port_case = 4;
        word_port = dport;
    } else {
        port_case = 1;
        word_port = lo == 0 ? max(dport, sport) : 65536;    // 111111
    }
    const int base = (((word_port * 10 + tb) * 10 + bb) * 5 + pb) * 2;
    int s = 0, d = 0;
    if (port_case == 2 && dport < sport) d = 1;
    else if (port_case == 2 && sport < dport) s = 1;
    else if (port_case == 4 && dport == 0) s = 1;
    else if (port_case == 4 && sport == 0) d = 1;
    *src_code = base + s;
    *dst_code = base + d;
}

__global__ void __launch_bounds__(256) flow_words_kernel(const int32_t* __restrict__ hour, const int32_t* __restrict__ minute,
                                                         const int32_t* __restrict__ second, const int64_t* __restrict__ ibyt,
                                                         const int64_t* __restrict__ ipkt, const int32_t* __restrict__ sport,
                                                         const int32_t* __restrict__ dport, int64_t n, FlowCutsDev cuts,
                                                         int32_t* __restrict__ src_code, int32_t* __restrict__ dst_code,
                                                         int* __restrict__ bad)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int tkey = hour[i] * 3600 + minute[i] * 60 + second[i];
    const long long b = ibyt[i], p = ipkt[i];
    // Quantiles.bin: the first cut with value <= cut; the last cut is the maximum of the column, so one always exists
    int tb = 0, bb = 0, pb = 0;
#pragma unroll
    for (int q = 0; q < 9; q++) tb += tkey > cuts.time_key[q];
#pragma unroll
    for (int q = 0; q < 9; q++) bb += b > cuts.ibyt[q];
#pragma unroll
    for (int q = 0; q < 4; q++) pb += p > cuts.ipkt[q];
    const int dp = dport[i], sp = sport[i];
    if ((unsigned)dp > 65535u || (unsigned)sp > 65535u) {
        *bad = 1;
        return;
    }
    int32_t sc, dc;
    adjust_port(dp, sp, tb, bb, pb, &sc, &dc);
    src_code[i] = sc;
    dst_code[i] = dc;
}

// least k in 1..n with (double)k / (double)n >= t (Quantiles.scala:31-38, 62-66)
long long rank_of_threshold(double t, long long n)
{
    const double dn = (double)n;
    long long k = (long long)ceil(t * dn);
    if (k < 1) k = 1;
    if (k > n) k = n;
    while (k > 1 && (double)(k - 1) / dn >= t) k--;
    while (k < n && !((double)k / dn >= t)) k++;
    return k;
}

// The k-th smallest (1-based ranks, ascending) of a device int64 column, all ranks at once.
int select_ranks_device(const int64_t* d_v, int64_t n, const long long* ranks, int nq, long long* out, cudaStream_t st)
{
    if (nq > kMaxQ) return onilda_fail(ONILDA_EINVAL, "quantiles: at most %d thresholds", kMaxQ);
    MinMax mm{LLONG_MAX, LLONG_MIN};
    MinMax* d_mm = nullptr;
    unsigned long long* d_hist = nullptr;
    ONILDA_CUDA(cudaMalloc(&d_mm, sizeof mm));
    if (cudaMalloc(&d_hist, sizeof(unsigned long long) * kMaxQ * kSelBins) != cudaSuccess) {
        cudaFree(d_mm);
        cudaGetLastError();
        return onilda_fail(ONILDA_ENOMEM, "quantiles: device scratch not available");
    }
    struct Guard { void *a, *b; ~Guard() { cudaFree(a); cudaFree(b); } } guard{d_mm, d_hist};
    ONILDA_CUDA(cudaMemcpyAsync(d_mm, &mm, sizeof mm, cudaMemcpyHostToDevice, st));
    minmax_kernel<<<148 * 8, 256, 0, st>>>(d_v, n, d_mm);
    ONILDA_CUDA(cudaMemcpyAsync(&mm, d_mm, sizeof mm, cudaMemcpyDeviceToHost, st));
    ONILDA_CUDA(cudaStreamSynchronize(st));
    const unsigned long long span = (unsigned long long)mm.hi - (unsigned long long)mm.lo;
    int bits = 0;
    while (bits < 64 && (span >> bits) != 0) bits++;
    const int levels = bits == 0 ? 0 : (bits + kSelBits - 1) / kSelBits;
    std::vector<unsigned long long> prefix((size_t)nq, 0ull), hist((size_t)kMaxQ * kSelBins);
    std::vector<long long> remaining(ranks, ranks + nq);
    // per device, so on every call (the current device may differ from the last call's)
    ONILDA_CUDA(cudaFuncSetAttribute(multi_select_hist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)(sizeof(unsigned int) * kMaxQ * kSelBins)));
    for (int level = levels - 1; level >= 0; level--) {
        SelLevel lv;
        lv.shift = level * kSelBits;
        std::vector<unsigned long long> distinct(prefix);
        std::sort(distinct.begin(), distinct.end());
        distinct.erase(std::unique(distinct.begin(), distinct.end()), distinct.end());
        lv.nprefix = (int)distinct.size();
        for (int p = 0; p < lv.nprefix; p++) lv.prefix[p] = distinct[(size_t)p];
        ONILDA_CUDA(cudaMemsetAsync(d_hist, 0, sizeof(unsigned long long) * (size_t)lv.nprefix * kSelBins, st));
        multi_select_hist_kernel<<<148 * (lv.nprefix <= 8 ? 2 : 1), kSelThreads, sizeof(unsigned int) * (size_t)lv.nprefix * kSelBins, st>>>(d_v, n, mm.lo, lv, d_hist);
        ONILDA_CUDA(cudaGetLastError());
        ONILDA_CUDA(cudaMemcpyAsync(hist.data(), d_hist, sizeof(unsigned long long) * (size_t)lv.nprefix * kSelBins,
                                    cudaMemcpyDeviceToHost, st));
        ONILDA_CUDA(cudaStreamSynchronize(st));
        for (int q = 0; q < nq; q++) {
            const int p = (int)(std::lower_bound(distinct.begin(), distinct.end(), prefix[(size_t)q]) - distinct.begin());
            const unsigned long long* h = hist.data() + (size_t)p * kSelBins;
            long long r = remaining[(size_t)q];
            int d = 0;
            while (d < kSelBins - 1 && (long long)h[d] < r) { r -= (long long)h[d]; d++; }
            remaining[(size_t)q] = r;
            prefix[(size_t)q] = (prefix[(size_t)q] << kSelBits) | (unsigned long long)d;
        }
    }
    for (int q = 0; q < nq; q++) out[q] = (long long)((unsigned long long)mm.lo + prefix[(size_t)q]);
    return ONILDA_OK;
}

struct DeviceScope {
    int prev = 0;
    bool set = false;
    int enter(int device)
    {
        ONILDA_CUDA(cudaGetDevice(&prev));
        if (device < 0) device = prev;
        ONILDA_CUDA(cudaSetDevice(device));
        set = true;
        return ONILDA_OK;
    }
    ~DeviceScope() { if (set) cudaSetDevice(prev); }
};

struct DevMem {
    void* p = nullptr;
    ~DevMem() { cudaFree(p); }
    int alloc(size_t bytes, const char* what)
    {
        if (cudaMalloc(&p, bytes ? bytes : 1) != cudaSuccess) {
            cudaGetLastError();
            return onilda_fail(ONILDA_ENOMEM, "%s: %zu bytes of device memory not available", what, bytes);
        }
        return ONILDA_OK;
    }
};

}  // namespace

// Quantiles.computeQuantiles (Quantiles.scala:53-75) of an int64 column in device memory.
extern "C" int onilda_quantiles_i64_device(const int64_t* values, int64_t n, const double* thresholds, int nq, double* cuts_out,
                                           void* stream)
{
    if (n < 0 || nq < 0 || (nq > 0 && (!thresholds || !cuts_out)) || (n > 0 && !values))
        return onilda_fail(ONILDA_EINVAL, "quantiles: bad argument");
    if (nq > kMaxQ) return onilda_fail(ONILDA_EINVAL, "quantiles: at most %d thresholds", kMaxQ);
    if (n == 0) {                                                    // Quantiles.scala:72: the trivial cutoffs
        for (int q = 0; q < nq; q++) cuts_out[q] = INFINITY;
        return ONILDA_OK;
    }
    long long ranks[kMaxQ] = {0}, vals[kMaxQ] = {0};
    for (int q = 0; q < nq; q++) {
        // a threshold above 1 is never reached: the cutoff stays +infinity
        ranks[q] = thresholds[q] > 1.0 || thresholds[q] != thresholds[q] ? -1 : rank_of_threshold(thresholds[q], n);
    }
    long long use[kMaxQ] = {0};
    for (int q = 0; q < nq; q++) use[q] = ranks[q] < 0 ? 1 : ranks[q];
    int rc = select_ranks_device(values, n, use, nq, vals, (cudaStream_t)stream);
    if (rc != ONILDA_OK) return rc;
    for (int q = 0; q < nq; q++) cuts_out[q] = ranks[q] < 0 ? INFINITY : (double)vals[q];
    return ONILDA_OK;
}

extern "C" int onilda_quantiles_i64(const int64_t* values, int64_t n, const double* thresholds, int nq, int device, double* cuts_out)
{
    if (n < 0 || (n > 0 && !values)) return onilda_fail(ONILDA_EINVAL, "quantiles: bad argument");
    if (onilda_device_count() == 0)
        return onilda_fail(ONILDA_ECUDA, "quantiles: no CUDA device is usable and this library has no CPU path");
    DeviceScope scope;
    int rc = scope.enter(device);
    if (rc != ONILDA_OK) return rc;
    DevMem d;
    if (n > 0) {
        if ((rc = d.alloc(sizeof(int64_t) * (size_t)n, "quantiles")) != ONILDA_OK) return rc;
        ONILDA_CUDA(cudaMemcpy(d.p, values, sizeof(int64_t) * (size_t)n, cudaMemcpyHostToDevice));
    }
    return onilda_quantiles_i64_device((const int64_t*)d.p, n, thresholds, nq, cuts_out, nullptr);
}

// FlowWordCreation.flowWordCreation on columns in device memory.  cuts_out (host): time deciles as the reference's
// doubles, ibyt deciles, ipkt quintiles.  src_code / dst_code: device, n entries each.
extern "C" int onilda_flow_words_device(const int32_t* hour, const int32_t* minute, const int32_t* second, const int64_t* ibyt,
                                        const int64_t* ipkt, const int32_t* sport, const int32_t* dport, int64_t n,
                                        onilda_flow_cuts* cuts_out, int32_t* src_code, int32_t* dst_code, void* stream)
{
    if (n < 0 || !cuts_out || (n > 0 && (!hour || !minute || !second || !ibyt || !ipkt || !sport || !dport || !src_code || !dst_code)))
        return onilda_fail(ONILDA_EINVAL, "flow_words: bad argument");
    for (int q = 0; q < 10; q++) cuts_out->time[q] = cuts_out->ibyt[q] = INFINITY;
    for (int q = 0; q < 5; q++) cuts_out->ipkt[q] = INFINITY;
    if (n == 0) return ONILDA_OK;
    cudaStream_t st = (cudaStream_t)stream;
    static const double deciles[10] = {0.1, 0.2, 0.3, 0.4, 0.5, 0.6, 0.7, 0.8, 0.9, 1.0};      // Quantiles.scala:86
    static const double quintiles[5] = {0.2, 0.4, 0.6, 0.8, 1.0};                               // Quantiles.scala:94
    FlowCutsDev cuts;
    // time of day: histogram over the second of the day
    DevMem d_hist, d_bad;
    int rc;
    if ((rc = d_hist.alloc(sizeof(unsigned long long) * 86400, "flow_words")) != ONILDA_OK) return rc;
    if ((rc = d_bad.alloc(sizeof(int), "flow_words")) != ONILDA_OK) return rc;
    ONILDA_CUDA(cudaMemsetAsync(d_hist.p, 0, sizeof(unsigned long long) * 86400, st));
    ONILDA_CUDA(cudaMemsetAsync(d_bad.p, 0, sizeof(int), st));
    time_hist_kernel<<<148 * 8, 256, 0, st>>>(hour, minute, second, n, (unsigned long long*)d_hist.p, (int*)d_bad.p);
    ONILDA_CUDA(cudaGetLastError());
    std::vector<unsigned long long> th(86400);
    int bad = 0;
    ONILDA_CUDA(cudaMemcpyAsync(th.data(), d_hist.p, sizeof(unsigned long long) * 86400, cudaMemcpyDeviceToHost, st));
    ONILDA_CUDA(cudaMemcpyAsync(&bad, d_bad.p, sizeof bad, cudaMemcpyDeviceToHost, st));
    ONILDA_CUDA(cudaStreamSynchronize(st));
    if (bad) return onilda_fail(ONILDA_EINVAL, "flow_words: hour/minute/second outside 0-23 / 0-59 / 0-59");
    {
        long long rank[10];
        for (int q = 0; q < 10; q++) rank[q] = rank_of_threshold(deciles[q], n);
        long long cum = 0;
        int q = 0;
        for (int key = 0; key < 86400 && q < 10; key++) {
            cum += (long long)th[(size_t)key];
            while (q < 10 && cum >= rank[q]) {
                cuts.time_key[q] = key;
                const int h = key / 3600, m = key / 60 % 60, s = key % 60;
                cuts_out->time[q] = (double)h + (double)m / 60 + (double)s / 3600;           // addTime, FlowWordCreation.scala:88-90
                q++;
            }
        }
    }
    double c10[10], c5[5];
    if ((rc = onilda_quantiles_i64_device(ibyt, n, deciles, 10, c10, stream)) != ONILDA_OK) return rc;
    if ((rc = onilda_quantiles_i64_device(ipkt, n, quintiles, 5, c5, stream)) != ONILDA_OK) return rc;
    for (int q = 0; q < 10; q++) { cuts_out->ibyt[q] = c10[q]; cuts.ibyt[q] = (long long)c10[q]; }
    for (int q = 0; q < 5; q++) { cuts_out->ipkt[q] = c5[q]; cuts.ipkt[q] = (long long)c5[q]; }
    flow_words_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(hour, minute, second, ibyt, ipkt, sport, dport, n, cuts, src_code,
                                                                    dst_code, (int*)d_bad.p);
    ONILDA_CUDA(cudaGetLastError());
    ONILDA_CUDA(cudaMemcpyAsync(&bad, d_bad.p, sizeof bad, cudaMemcpyDeviceToHost, st));
    ONILDA_CUDA(cudaStreamSynchronize(st));
    if (bad) return onilda_fail(ONILDA_EINVAL, "flow_words: port outside 0-65535");
    return ONILDA_OK;
}

extern "C" int onilda_flow_words(const int32_t* hour, const int32_t* minute, const int32_t* second, const int64_t* ibyt,
                                 const int64_t* ipkt, const int32_t* sport, const int32_t* dport, int64_t n, int device,
                                 onilda_flow_cuts* cuts_out, int32_t* src_code, int32_t* dst_code)
{
    if (n < 0 || !cuts_out || (n > 0 && (!hour || !minute || !second || !ibyt || !ipkt || !sport || !dport || !src_code || !dst_code)))
        return onilda_fail(ONILDA_EINVAL, "flow_words: bad argument");
    if (onilda_device_count() == 0)
        return onilda_fail(ONILDA_ECUDA, "flow_words: no CUDA device is usable and this library has no CPU path");
    DeviceScope scope;
    int rc = scope.enter(device);
    if (rc != ONILDA_OK) return rc;
    if (n == 0) return onilda_flow_words_device(nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, 0, cuts_out, nullptr, nullptr, nullptr);
    DevMem d;
    const size_t b4 = (sizeof(int32_t) * (size_t)n + 255) & ~(size_t)255, b8 = (sizeof(int64_t) * (size_t)n + 255) & ~(size_t)255;
    if ((rc = d.alloc(7 * b4 + 2 * b8, "flow_words")) != ONILDA_OK) return rc;
    char* p = (char*)d.p;
    int64_t* d_ibyt = (int64_t*)p; p += b8;
    int64_t* d_ipkt = (int64_t*)p; p += b8;
    int32_t* col[7];
    for (int k = 0; k < 7; k++) { col[k] = (int32_t*)p; p += b4; }
    const int32_t* src[5] = {hour, minute, second, sport, dport};
    for (int k = 0; k < 5; k++) ONILDA_CUDA(cudaMemcpyAsync(col[k], src[k], sizeof(int32_t) * (size_t)n, cudaMemcpyHostToDevice, 0));
    ONILDA_CUDA(cudaMemcpyAsync(d_ibyt, ibyt, sizeof(int64_t) * (size_t)n, cudaMemcpyHostToDevice, 0));
    ONILDA_CUDA(cudaMemcpyAsync(d_ipkt, ipkt, sizeof(int64_t) * (size_t)n, cudaMemcpyHostToDevice, 0));
    rc = onilda_flow_words_device(col[0], col[1], col[2], d_ibyt, d_ipkt, col[3], col[4], n, cuts_out, col[5], col[6], nullptr);
    if (rc != ONILDA_OK) return rc;
    ONILDA_CUDA(cudaMemcpyAsync(src_code, col[5], sizeof(int32_t) * (size_t)n, cudaMemcpyDeviceToHost, 0));
    ONILDA_CUDA(cudaMemcpyAsync(dst_code, col[6], sizeof(int32_t) * (size_t)n, cudaMemcpyDeviceToHost, 0));
    ONILDA_CUDA(cudaStreamSynchronize(0));
    return ONILDA_OK;
}

// The reference's word string of a code: ["-1_"]<port>.0_<timeBin>.0_<ibytBin>.0_<ipktBin>.0 (FlowWordCreation.scala:131-145;
// the four parts are Scala Doubles printed with toString).  Returns the length, or a negative error code.
extern "C" int onilda_flow_word_string(int32_t code, char* buf, size_t cap)
{
    if (code < 0 || !buf) return onilda_fail(ONILDA_EINVAL, "flow_word_string: bad argument");
    const int marker = code & 1;
    int rest = code >> 1;
    const int pb = rest % 5; rest /= 5;
    const int bb = rest % 10; rest /= 10;
    const int tb = rest % 10; rest /= 10;
    if (rest > 65537) return onilda_fail(ONILDA_EINVAL, "flow_word_string: %d is not a word code", code);
    const int port = rest == 65536 ? 111111 : rest == 65537 ? 333333 : rest;
    const int len = snprintf(buf, cap, "%s%d.0_%d.0_%d.0_%d.0", marker ? "-1_" : "", port, tb, bb, pb);
    if (len < 0 || (size_t)len >= cap) return onilda_fail(ONILDA_EINVAL, "flow_word_string: buffer of %zu bytes is too small", cap);
    return len;
}

// =============================================================================== DNS word creation
// dns/DNSWordCreation.scala:32-236 and utilities/Entropy.scala:16-23 on a byte buffer of query names.
namespace {

// ISO country-code suffixes the reference tests the last label against (DNSWordCreation.scala:17-30; the set
// also holds "krd" and the empty string, which a last label can never equal after String.split drops trailing
// empty parts).  Sorted, two letters each; "krd" is tested apart.
static const char kCountry2[] =
    "acadaeafagaialamanaoaqarasatauawaxazbabbbdbebfbgbhbibjbmbnbobqbrbsbtbvbwbybzcacccdcfcgchcickclcmcncocrcucvcwcxcycz"
    "dedjdkdmdodzeceeegeheresetfifjfkfmfofrgagbgdgegfggghgiglgmgngpgqgrgsgtgugwgyhkhmhnhrhthuidieiliminioiqirisit"
    "jejmjojpkekgkhkikmknkpkrkwkykzlalblclilklrlsltlulvlymamcmdmemgmhmkmlmmmnmompmqmrmsmtmumvmwmxmymznancnenfngninlnonpnrnunz"
    "ompapepfpgphpkplpmpnprpsptpwpyqarerorsrurwsasbscsdsesgshsisjskslsmsnsosrssstsusvsxsysztctdtftgthtjtktltmtntotptrtttvtwtz"
    "uaugukusuyuzvavcvevgvivnvuwfwsyeytzazmzweu";
constexpr int kCountry2Count = (int)(sizeof(kCountry2) - 1) / 2;

// 26 x 26 bits, one per pair of lower-case letters; filled from kCountry2 on each call (per device)
__constant__ unsigned int kCountryBits[22];

__device__ __forceinline__ bool is_country(const unsigned char* p, int len)
{
    if (len == 3) return p[0] == 'k' && p[1] == 'r' && p[2] == 'd';
    if (len != 2) return false;
    const unsigned int a = p[0] - 'a', b = p[1] - 'a';
    if (a >= 26u || b >= 26u) return false;
    const unsigned int bit = a * 26 + b;
    return (kCountryBits[bit >> 5] >> (bit & 31)) & 1u;
}

__device__ __forceinline__ bool bytes_equal(const unsigned char* p, int len, const char* lit, int lit_len)
{
    if (len != lit_len) return false;
    for (int i = 0; i < len; i++)
        if (p[i] != (unsigned char)lit[i]) return false;
    return true;
}

__host__ __device__ __forceinline__ unsigned long long fnv1a(const unsigned char* p, int len)
{
    unsigned long long h = 1469598103934665603ull;
    for (int i = 0; i < len; i++) {
        h ^= p[i];
        h *= 1099511628211ull;
    }
    return h;
}

struct TopTable {
    const int32_t* slot;        // index into the top-domain list or -1; capacity a power of two
    unsigned long long mask;
    const unsigned char* bytes;
    const int64_t* offsets;
};

// Entropy.stringEntropy: sum over the distinct characters of -p log10(p) / log10(2), p = occurrences / length.
// The terms come from a host-made table term[len][count] (the host's log10, so the device agrees bit for bit with a
// host evaluation) and are added largest count first, which makes the value a function of the multiset of counts
// (the reference adds in the iteration order of a Scala Map, which is not specified).
__device__ double string_entropy(const unsigned char* p, int len, const double* __restrict__ term /* [256][256] */)
{
    if (len <= 0) return 0.0;
    unsigned char counts[256];            // counts of the distinct characters, in order of first occurrence
    int nd = 0;
    for (int i = 0; i < len; i++) {
        const unsigned char c = p[i];
        bool seen = false;
        for (int j = 0; j < i; j++)
            if (p[j] == c) { seen = true; break; }
        if (seen) continue;
        int k = 1;
        for (int j = i + 1; j < len; j++) k += (p[j] == c);
        counts[nd++] = (unsigned char)(k > 255 ? 255 : k);
    }
    double e = 0.0;
    for (int want = len > 255 ? 255 : len; want >= 1; want--)        // largest count first
        for (int d = 0; d < nd; d++)
            if (counts[d] == want) e += len <= 255 ? term[len * 256 + want] : -((double)want / len) * log10((double)want / len) / log10(2.0);
    return e;
}

// DNSWordCreation.extractSubdomain (:155-196) + stringEntropy of the subdomain column + getTopDomain (:147-153)
// One name: s[0, len) (generic pointer: shared-memory stage or global), results to index i.
__device__ __forceinline__ void dns_features_one(const unsigned char* s, int len, int64_t i, const TopTable& top,
                                                 const double* __restrict__ term, int64_t* __restrict__ sub_len,
                                                 int64_t* __restrict__ num_parts, int64_t* __restrict__ entropy_bits,
                                                 int32_t* __restrict__ top_out, int32_t* __restrict__ spans)
{
    // String.split("[.]"): trailing empty parts are dropped; a string without a dot is one part
    int e = len, dots = 0;
    for (int k = 0; k < len; k++) dots += (s[k] == '.');
    if (dots > 0)
        while (e > 0 && s[e - 1] == '.') e--;
    int parts = 0, d1 = -1, d2 = -1, d3 = -1;        // the last three dots of s[0, e)
    if (dots == 0) parts = 1;
    else if (e > 0) {
        parts = 1;
        for (int k = 0; k < e; k++)
            if (s[k] == '.') { parts++; d3 = d2; d2 = d1; d1 = k; }
    }
    int dom_b = -1, dom_e = -1, sub_e = -1;          // domain = s[dom_b, dom_e), subdomain = s[0, sub_e); -1 = "None"
    if (parts > 2) {
        const unsigned char* last = s + d1 + 1;
        const int last_len = e - d1 - 1;
        const bool ip = bytes_equal(last, last_len, "arpa", 4) && bytes_equal(s + d2 + 1, d1 - d2 - 1, "in-addr", 7);
        if (!ip) {
            if (is_country(last, last_len)) {
                dom_b = d3 + 1; dom_e = d2;
                if (parts - 3 >= 1) sub_e = d3;
            } else {
                dom_b = d2 + 1; dom_e = d1;
                sub_e = d2;
            }
        }
    }
    // the reference keeps "None" as a string: a subdomain that reads None counts as absent, and the entropy column is
    // the entropy of whatever string is there - 2.0 for "None"
    const bool sub_none = sub_e < 0 || bytes_equal(s, sub_e, "None", 4);
    sub_len[i] = sub_none ? 0 : sub_e;
    num_parts[i] = parts;
    const unsigned char none[4] = {'N', 'o', 'n', 'e'};
    const double ent = sub_e < 0 ? string_entropy(none, 4, term) : string_entropy(s, sub_e, term);
    entropy_bits[i] = ent > 0.0 ? __double_as_longlong(ent) : 0;
    // a missing domain is the string "None" in the reference and is looked up like any other
    const unsigned char* dp = dom_b >= 0 ? s + dom_b : none;
    const int dl = dom_b >= 0 ? dom_e - dom_b : 4;
    int t = 0;
    if (bytes_equal(dp, dl, "intel", 5)) t = 2;
    else if (top.slot) {
        unsigned long long h = fnv1a(dp, dl) & top.mask;
        for (;;) {
            const int32_t idx = top.slot[h];
            if (idx < 0) break;
            const int64_t b = top.offsets[idx];
            if (bytes_equal(dp, dl, (const char*)(top.bytes + b), (int)(top.offsets[idx + 1] - b))) { t = 1; break; }
            h = (h + 1) & top.mask;
        }
    }
    top_out[i] = t;
    if (spans) {
        spans[3 * i] = dom_b;
        spans[3 * i + 1] = dom_e;
        spans[3 * i + 2] = sub_e;
    }
}

// DNSWordCreation.extractSubdomain (:155-196) + stringEntropy of the subdomain column + getTopDomain (:147-153).
// One thread per name.  The names of a warp are contiguous in the buffer: the warp copies them into shared memory
// with 16-byte loads first (a thread walking its own name byte by byte in global memory costs a cache-line lookup
// per lane per byte), and falls back to global memory when its 32 names do not fit the stage.
constexpr int kDnsStage = 2048;

__global__ void __launch_bounds__(128) dns_features_kernel(const unsigned char* __restrict__ names, const int64_t* __restrict__ off,
                                                           int64_t n, TopTable top, const double* __restrict__ term,
                                                           int64_t* __restrict__ sub_len, int64_t* __restrict__ num_parts,
                                                           int64_t* __restrict__ entropy_bits, int32_t* __restrict__ top_out,
                                                           int* __restrict__ bad, int32_t* __restrict__ spans /* [n][3] or null */)
{
    __shared__ __align__(16) unsigned char stage[4][kDnsStage + 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const bool in_range = i < n;
    const int64_t b = in_range ? off[i] : 0, e = in_range ? off[i + 1] : 0;
    const int64_t len64 = e - b;
    bool live = in_range;
    if (in_range && (b < 0 || len64 < 0 || len64 > 255)) {          // a DNS name has at most 253 characters
        *bad = 1;
        live = false;
    }
    // the byte range of the warp's names (lane 0 holds the lowest index of the warp)
    const int64_t first = __shfl_sync(0xffffffffu, b, 0);
    long long rel_end = live ? e - first : 0;
    if (rel_end < 0) rel_end = 0;
    if (rel_end > (1ll << 30)) rel_end = 1ll << 30;
    const int span = __reduce_max_sync(0xffffffffu, (int)rel_end);
    const bool ordered = __all_sync(0xffffffffu, !live || b >= first);
    const unsigned char* s = names + b;
    const uintptr_t base_addr = (uintptr_t)(names + first);
    const int lead = (int)(base_addr & 15);                   // the copy starts at the 16-byte boundary below the first byte
    if (ordered && span > 0 && span + lead <= kDnsStage) {
        const uint4* src = reinterpret_cast<const uint4*>(base_addr - lead);
        uint4* dst = reinterpret_cast<uint4*>(stage[warp]);
        const int chunks = (span + lead + 15) / 16;
        for (int c = lane; c < chunks; c += 32) dst[c] = src[c];
        __syncwarp();
        s = stage[warp] + lead + (b - first);
    }
    if (live) dns_features_one(s, (int)len64, i, top, term, sub_len, num_parts, entropy_bits, top_out, spans);
}

struct DnsCutsDev {
    long long time[10], frame[10], sub_len[5], entropy[5], periods[5];
};

__global__ void __launch_bounds__(256) dns_words_kernel(const int32_t* __restrict__ frame_len, const int64_t* __restrict__ tstamp,
                                                        const int64_t* __restrict__ sub_len, const int64_t* __restrict__ entropy_bits,
                                                        const int64_t* __restrict__ num_parts, const int32_t* __restrict__ top,
                                                        const int32_t* __restrict__ qtype, const int32_t* __restrict__ rcode, int64_t n,
                                                        DnsCutsDev cuts, int64_t* __restrict__ code, int* __restrict__ bad)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const long long f = frame_len[i], t = tstamp[i], sl = sub_len[i], eb = entropy_bits[i], np = num_parts[i];
    int fb = 0, tb = 0, sb = 0, ebn = 0, pb = 0;
#pragma unroll
    for (int q = 0; q < 9; q++) { fb += f > cuts.frame[q]; tb += t > cuts.time[q]; }
#pragma unroll
    for (int q = 0; q < 4; q++) { sb += sl > cuts.sub_len[q]; ebn += eb > cuts.entropy[q]; pb += np > cuts.periods[q]; }
    const int qt = qtype[i], rc = rcode[i];
    if ((unsigned)qt > 65535u || (unsigned)rc > 65535u) {
        *bad = 1;
        return;
    }
    long long c = top[i];
    c = c * 10 + fb; c = c * 10 + tb; c = c * 5 + sb; c = c * 5 + ebn; c = c * 5 + pb;
    code[i] = (c << 32) | ((long long)qt << 16) | rc;
}

__global__ void __launch_bounds__(256) widen_kernel(const int32_t* __restrict__ in, int64_t n, int64_t* __restrict__ out)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = in[i];
}

__global__ void __launch_bounds__(256) count_positive_kernel(const int64_t* __restrict__ v, int64_t n, unsigned long long* __restrict__ out)
{
    unsigned int c = 0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) c += v[i] > 0;
    c = __reduce_add_sync(0xffffffffu, c);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(out, (unsigned long long)c);
}

// Quantiles.computeQuantiles over the entries > 0 of a non-negative column (the reference filters "> 0" first,
// DNSWordCreation.scala:96-121): the zeros sort first, so the ranks shift by their number.
int quantiles_positive(const int64_t* d_v, int64_t n, const double* thresholds, int nq, long long* out /* LLONG_MAX = +inf */,
                       cudaStream_t st)
{
    unsigned long long* d_cnt = nullptr;
    ONILDA_CUDA(cudaMalloc(&d_cnt, sizeof *d_cnt));
    struct G { void* p; ~G() { cudaFree(p); } } g{d_cnt};
    ONILDA_CUDA(cudaMemsetAsync(d_cnt, 0, sizeof *d_cnt, st));
    count_positive_kernel<<<148 * 8, 256, 0, st>>>(d_v, n, d_cnt);
    unsigned long long valid = 0;
    ONILDA_CUDA(cudaMemcpyAsync(&valid, d_cnt, sizeof valid, cudaMemcpyDeviceToHost, st));
    ONILDA_CUDA(cudaStreamSynchronize(st));
    if (valid == 0) {
        for (int q = 0; q < nq; q++) out[q] = LLONG_MAX;
        return ONILDA_OK;
    }
    long long ranks[kMaxQ];
    for (int q = 0; q < nq; q++) ranks[q] = rank_of_threshold(thresholds[q], (long long)valid) + (n - (long long)valid);
    return select_ranks_device(d_v, n, ranks, nq, out, st);
}

}  // namespace

// features_out (device, may be null): [0] subdomain length, [1] parts, [2] entropy bits, each int64[n]; top flag int32[n];
// spans int32[n][3] = domain begin, domain end, subdomain end (-1 = "None")
struct DnsFeatureOut {
    int64_t *sub_len, *parts, *entropy_bits;
    int32_t *top, *spans;
};

static int dns_words_core(const char* names, const int64_t* name_offsets, const int32_t* frame_len, const int64_t* unix_tstamp,
                          const int32_t* qry_type, const int32_t* qry_rcode, int64_t n, const char* top_bytes,
                          const int64_t* top_offsets, const int32_t* top_slots, int64_t top_capacity, onilda_dns_cuts* cuts_out,
                          int64_t* word_code, void* stream, const DnsFeatureOut* fo)
{
    int32_t* features_spans = fo ? fo->spans : nullptr;
    if (n < 0 || !cuts_out || (n > 0 && (!names || !name_offsets || !frame_len || !unix_tstamp || !qry_type || !qry_rcode || !word_code)))
        return onilda_fail(ONILDA_EINVAL, "dns_words: bad argument");
    if (top_capacity < 0 || (top_capacity & (top_capacity - 1)) || (top_capacity > 0 && (!top_bytes || !top_offsets || !top_slots)))
        return onilda_fail(ONILDA_EINVAL, "dns_words: the top-domain table needs a power-of-two capacity");
    for (int q = 0; q < 10; q++) cuts_out->time[q] = cuts_out->frame_length[q] = INFINITY;
    for (int q = 0; q < 5; q++) cuts_out->subdomain_length[q] = cuts_out->entropy[q] = cuts_out->num_periods[q] = INFINITY;
    if (n == 0) return ONILDA_OK;
    cudaStream_t st = (cudaStream_t)stream;
    static const double deciles[10] = {0.1, 0.2, 0.3, 0.4, 0.5, 0.6, 0.7, 0.8, 0.9, 1.0};
    static const double quintiles[5] = {0.2, 0.4, 0.6, 0.8, 1.0};
    // the table of entropy terms, from the host's log10
    std::vector<double> term(256 * 256, 0.0);
    for (int len = 1; len < 256; len++)
        for (int k = 1; k <= len; k++) {
            const double p = (double)k / (double)len;
            term[(size_t)len * 256 + k] = -p * log10(p) / log10(2.0);
        }
    DevMem d_term, d_cols, d_top, d_bad;
    int rc;
    const size_t b8 = (sizeof(int64_t) * (size_t)n + 255) & ~(size_t)255, b4 = (sizeof(int32_t) * (size_t)n + 255) & ~(size_t)255;
    if ((rc = d_term.alloc(sizeof(double) * term.size(), "dns_words")) != ONILDA_OK) return rc;
    if ((rc = d_cols.alloc(4 * b8 + b4, "dns_words")) != ONILDA_OK) return rc;
    if ((rc = d_bad.alloc(sizeof(int), "dns_words")) != ONILDA_OK) return rc;
    ONILDA_CUDA(cudaMemcpyAsync(d_term.p, term.data(), sizeof(double) * term.size(), cudaMemcpyHostToDevice, st));
    ONILDA_CUDA(cudaMemsetAsync(d_bad.p, 0, sizeof(int), st));
    char* p = (char*)d_cols.p;
    int64_t* d_sub = (int64_t*)p; p += b8;
    int64_t* d_parts = (int64_t*)p; p += b8;
    int64_t* d_ent = (int64_t*)p; p += b8;
    int64_t* d_frame = (int64_t*)p; p += b8;
    int32_t* d_topflag = (int32_t*)p;
    {
        unsigned int bits[22] = {0};
        for (int i = 0; i < kCountry2Count; i++) {
            const unsigned int bit = (unsigned int)(kCountry2[2 * i] - 'a') * 26 + (unsigned int)(kCountry2[2 * i + 1] - 'a');
            bits[bit >> 5] |= 1u << (bit & 31);
        }
        ONILDA_CUDA(cudaMemcpyToSymbolAsync(kCountryBits, bits, sizeof bits, 0, cudaMemcpyHostToDevice, st));
    }
    TopTable top{top_capacity > 0 ? top_slots : nullptr, (unsigned long long)(top_capacity > 0 ? top_capacity - 1 : 0),
                 (const unsigned char*)top_bytes, top_offsets};
    dns_features_kernel<<<(unsigned)((n + 127) / 128), 128, 0, st>>>((const unsigned char*)names, name_offsets, n, top,
                                                                      (const double*)d_term.p, d_sub, d_parts, d_ent, d_topflag,
                                                                      (int*)d_bad.p, features_spans);
    widen_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(frame_len, n, d_frame);
    ONILDA_CUDA(cudaGetLastError());
    {
        int bad0 = 0;
        ONILDA_CUDA(cudaMemcpyAsync(&bad0, d_bad.p, sizeof bad0, cudaMemcpyDeviceToHost, st));
        ONILDA_CUDA(cudaStreamSynchronize(st));
        if (bad0) return onilda_fail(ONILDA_EINVAL, "dns_words: a query name is longer than 255 bytes (or the offsets decrease)");
    }
    DnsCutsDev cuts;
    double c10[10];
    if ((rc = onilda_quantiles_i64_device(unix_tstamp, n, deciles, 10, c10, stream)) != ONILDA_OK) return rc;
    for (int q = 0; q < 10; q++) { cuts_out->time[q] = c10[q]; cuts.time[q] = (long long)c10[q]; }
    if ((rc = onilda_quantiles_i64_device(d_frame, n, deciles, 10, c10, stream)) != ONILDA_OK) return rc;
    for (int q = 0; q < 10; q++) { cuts_out->frame_length[q] = c10[q]; cuts.frame[q] = (long long)c10[q]; }
    if ((rc = quantiles_positive(d_sub, n, quintiles, 5, cuts.sub_len, st)) != ONILDA_OK) return rc;
    if ((rc = quantiles_positive(d_ent, n, quintiles, 5, cuts.entropy, st)) != ONILDA_OK) return rc;
    if ((rc = quantiles_positive(d_parts, n, quintiles, 5, cuts.periods, st)) != ONILDA_OK) return rc;
    for (int q = 0; q < 5; q++) {
        cuts_out->subdomain_length[q] = cuts.sub_len[q] == LLONG_MAX ? INFINITY : (double)cuts.sub_len[q];
        cuts_out->num_periods[q] = cuts.periods[q] == LLONG_MAX ? INFINITY : (double)cuts.periods[q];
        if (cuts.entropy[q] == LLONG_MAX) cuts_out->entropy[q] = INFINITY;
        else memcpy(&cuts_out->entropy[q], &cuts.entropy[q], sizeof(double));
    }
    dns_words_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(frame_len, unix_tstamp, d_sub, d_ent, d_parts, d_topflag, qry_type,
                                                                   qry_rcode, n, cuts, word_code, (int*)d_bad.p);
    ONILDA_CUDA(cudaGetLastError());
    int bad = 0;
    ONILDA_CUDA(cudaMemcpyAsync(&bad, d_bad.p, sizeof bad, cudaMemcpyDeviceToHost, st));
    ONILDA_CUDA(cudaStreamSynchronize(st));
    if (bad) return onilda_fail(ONILDA_EINVAL, "dns_words: query type / response code outside 0-65535");
    if (fo) {
        ONILDA_CUDA(cudaMemcpyAsync(fo->sub_len, d_sub, sizeof(int64_t) * (size_t)n, cudaMemcpyDeviceToDevice, st));
        ONILDA_CUDA(cudaMemcpyAsync(fo->parts, d_parts, sizeof(int64_t) * (size_t)n, cudaMemcpyDeviceToDevice, st));
        ONILDA_CUDA(cudaMemcpyAsync(fo->entropy_bits, d_ent, sizeof(int64_t) * (size_t)n, cudaMemcpyDeviceToDevice, st));
        ONILDA_CUDA(cudaMemcpyAsync(fo->top, d_topflag, sizeof(int32_t) * (size_t)n, cudaMemcpyDeviceToDevice, st));
        ONILDA_CUDA(cudaStreamSynchronize(st));
    }
    return ONILDA_OK;
}

extern "C" int onilda_dns_words_device(const char* names, const int64_t* name_offsets, const int32_t* frame_len,
                                       const int64_t* unix_tstamp, const int32_t* qry_type, const int32_t* qry_rcode, int64_t n,
                                       const char* top_bytes, const int64_t* top_offsets, const int32_t* top_slots,
                                       int64_t top_capacity, onilda_dns_cuts* cuts_out, int64_t* word_code, void* stream)
{
    return dns_words_core(names, name_offsets, frame_len, unix_tstamp, qry_type, qry_rcode, n, top_bytes, top_offsets, top_slots,
                          top_capacity, cuts_out, word_code, stream, nullptr);
}

extern "C" int onilda_dns_words(const char* names, const int64_t* name_offsets, const int32_t* frame_len, const int64_t* unix_tstamp,
                                const int32_t* qry_type, const int32_t* qry_rcode, int64_t n, const char* top_domains,
                                const int64_t* top_offsets, int64_t n_top, int device, onilda_dns_cuts* cuts_out, int64_t* word_code)
{
    return onilda_dns_words_features(names, name_offsets, frame_len, unix_tstamp, qry_type, qry_rcode, n, top_domains, top_offsets,
                                     n_top, device, cuts_out, word_code, nullptr, nullptr, nullptr, nullptr, nullptr);
}

// The same with the per-event columns DNSPostLDA keeps beside the word (subdomain length, number of parts, subdomain
// entropy, top-domain flag) and the byte spans of domain and subdomain inside each name ({-1,-1} / -1 = "None").
extern "C" int onilda_dns_words_features(const char* names, const int64_t* name_offsets, const int32_t* frame_len,
                                         const int64_t* unix_tstamp, const int32_t* qry_type, const int32_t* qry_rcode, int64_t n,
                                         const char* top_domains, const int64_t* top_offsets, int64_t n_top, int device,
                                         onilda_dns_cuts* cuts_out, int64_t* word_code, double* subdomain_length_out,
                                         double* num_parts_out, double* entropy_out, int32_t* top_out, int32_t* spans_out)
{
    if (n < 0 || n_top < 0 || !cuts_out || (n > 0 && (!names || !name_offsets || !frame_len || !unix_tstamp || !qry_type || !qry_rcode || !word_code)) ||
        (n_top > 0 && (!top_domains || !top_offsets)))
        return onilda_fail(ONILDA_EINVAL, "dns_words: bad argument");
    if (onilda_device_count() == 0)
        return onilda_fail(ONILDA_ECUDA, "dns_words: no CUDA device is usable and this library has no CPU path");
    DeviceScope scope;
    int rc = scope.enter(device);
    if (rc != ONILDA_OK) return rc;
    if (n == 0) return onilda_dns_words_device(nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, 0, nullptr, nullptr, nullptr, 0, cuts_out, nullptr, nullptr);
    const bool want_features = subdomain_length_out || num_parts_out || entropy_out || top_out || spans_out;
    // hash set of the top domains: open addressing over FNV-1a, exact (the device compares the bytes)
    int64_t cap = 0;
    std::vector<int32_t> slots;
    if (n_top > 0) {
        if (n_top >= ((int64_t)1 << 30)) return onilda_fail(ONILDA_EINVAL, "dns_words: too many top domains");
        cap = 1;
        while (cap < 2 * n_top) cap <<= 1;
        slots.assign((size_t)cap, -1);
        for (int64_t t = 0; t < n_top; t++) {
            const unsigned char* b = (const unsigned char*)top_domains + top_offsets[t];
            const int len = (int)(top_offsets[t + 1] - top_offsets[t]);
            unsigned long long h = fnv1a(b, len) & (unsigned long long)(cap - 1);
            bool dup = false;
            while (slots[(size_t)h] >= 0) {
                const int32_t o = slots[(size_t)h];
                const int olen = (int)(top_offsets[o + 1] - top_offsets[o]);
                if (olen == len && memcmp(top_domains + top_offsets[o], b, (size_t)len) == 0) { dup = true; break; }
                h = (h + 1) & (unsigned long long)(cap - 1);
            }
            if (!dup) slots[(size_t)h] = (int32_t)t;
        }
    }
    const size_t name_bytes = (size_t)name_offsets[n], top_bytes = n_top > 0 ? (size_t)top_offsets[n_top] : 0;
    DevMem d;
    const size_t a = 256;
    auto up = [&](size_t x) { return (x + a - 1) & ~(a - 1); };
    const size_t total = up(name_bytes) + up(sizeof(int64_t) * (size_t)(n + 1)) + 3 * up(sizeof(int32_t) * (size_t)n) +
                         2 * up(sizeof(int64_t) * (size_t)n) + up(top_bytes) + up(sizeof(int64_t) * (size_t)(n_top + 1)) +
                         up(sizeof(int32_t) * (size_t)cap);
    if ((rc = d.alloc(total, "dns_words")) != ONILDA_OK) return rc;
    char* p = (char*)d.p;
    auto put = [&](const void* src, size_t bytes) -> void* {
        void* dst = p;
        if (bytes) cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, 0);
        p += up(bytes);
        return dst;
    };
    const char* d_names = (const char*)put(names, name_bytes);
    const int64_t* d_off = (const int64_t*)put(name_offsets, sizeof(int64_t) * (size_t)(n + 1));
    const int32_t* d_frame = (const int32_t*)put(frame_len, sizeof(int32_t) * (size_t)n);
    const int32_t* d_qt = (const int32_t*)put(qry_type, sizeof(int32_t) * (size_t)n);
    const int32_t* d_rc = (const int32_t*)put(qry_rcode, sizeof(int32_t) * (size_t)n);
    const int64_t* d_ts = (const int64_t*)put(unix_tstamp, sizeof(int64_t) * (size_t)n);
    int64_t* d_code = (int64_t*)p; p += up(sizeof(int64_t) * (size_t)n);
    const char* d_top = (const char*)put(top_domains, top_bytes);
    const int64_t* d_topoff = (const int64_t*)put(top_offsets, n_top > 0 ? sizeof(int64_t) * (size_t)(n_top + 1) : 0);
    const int32_t* d_slots = (const int32_t*)put(slots.data(), sizeof(int32_t) * (size_t)cap);
    ONILDA_CUDA(cudaGetLastError());
    DevMem d_feat;
    DnsFeatureOut fo{nullptr, nullptr, nullptr, nullptr, nullptr};
    if (want_features) {
        const size_t f8 = up(sizeof(int64_t) * (size_t)n), f4 = up(sizeof(int32_t) * (size_t)n);
        if ((rc = d_feat.alloc(3 * f8 + f4 + 3 * f4, "dns_words")) != ONILDA_OK) return rc;
        char* q = (char*)d_feat.p;
        fo.sub_len = (int64_t*)q; q += f8;
        fo.parts = (int64_t*)q; q += f8;
        fo.entropy_bits = (int64_t*)q; q += f8;
        fo.top = (int32_t*)q; q += f4;
        fo.spans = (int32_t*)q;
    }
    rc = dns_words_core(d_names, d_off, d_frame, d_ts, d_qt, d_rc, n, d_top, d_topoff, d_slots, cap, cuts_out, d_code, nullptr,
                        want_features ? &fo : nullptr);
    if (rc != ONILDA_OK) return rc;
    ONILDA_CUDA(cudaMemcpy(word_code, d_code, sizeof(int64_t) * (size_t)n, cudaMemcpyDeviceToHost));
    if (want_features) {
        std::vector<int64_t> tmp((size_t)n);
        if (subdomain_length_out) {
            ONILDA_CUDA(cudaMemcpy(tmp.data(), fo.sub_len, sizeof(int64_t) * (size_t)n, cudaMemcpyDeviceToHost));
            for (int64_t i = 0; i < n; i++) subdomain_length_out[i] = (double)tmp[(size_t)i];
        }
        if (num_parts_out) {
            ONILDA_CUDA(cudaMemcpy(tmp.data(), fo.parts, sizeof(int64_t) * (size_t)n, cudaMemcpyDeviceToHost));
            for (int64_t i = 0; i < n; i++) num_parts_out[i] = (double)tmp[(size_t)i];
        }
        if (entropy_out) ONILDA_CUDA(cudaMemcpy(entropy_out, fo.entropy_bits, sizeof(int64_t) * (size_t)n, cudaMemcpyDeviceToHost));
        if (top_out) ONILDA_CUDA(cudaMemcpy(top_out, fo.top, sizeof(int32_t) * (size_t)n, cudaMemcpyDeviceToHost));
        if (spans_out) ONILDA_CUDA(cudaMemcpy(spans_out, fo.spans, 3 * sizeof(int32_t) * (size_t)n, cudaMemcpyDeviceToHost));
    }
    return ONILDA_OK;
}

// The reference's word of a code: top_frameBin_timeBin_subdomainLengthBin_entropyBin_periodsBin_queryType_responseCode
// (DNSWordCreation.scala:226-235).  Returns the length or a negative error code.
extern "C" int onilda_dns_word_string(int64_t code, char* buf, size_t cap)
{
    if (code < 0 || !buf) return onilda_fail(ONILDA_EINVAL, "dns_word_string: bad argument");
    const int rc = (int)(code & 0xffff), qt = (int)((code >> 16) & 0xffff);
    long long c = code >> 32;
    const int pb = (int)(c % 5); c /= 5;
    const int eb = (int)(c % 5); c /= 5;
    const int sb = (int)(c % 5); c /= 5;
    const int tb = (int)(c % 10); c /= 10;
    const int fb = (int)(c % 10); c /= 10;
    if (c > 2) return onilda_fail(ONILDA_EINVAL, "dns_word_string: not a word code");
    const int len = snprintf(buf, cap, "%d_%d_%d_%d_%d_%d_%d_%d", (int)c, fb, tb, sb, eb, pb, qt, rc);
    if (len < 0 || (size_t)len >= cap) return onilda_fail(ONILDA_EINVAL, "dns_word_string: buffer of %zu bytes is too small", cap);
    return len;
}
