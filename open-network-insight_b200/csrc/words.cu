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
    static bool attr_set = false;
    if (!attr_set) {
        ONILDA_CUDA(cudaFuncSetAttribute(multi_select_hist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)(sizeof(unsigned int) * kMaxQ * kSelBins)));
        attr_set = true;
    }
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
