// (document, word, count) triples -> CSR corpus on the device.
//
// the step immediately before the path.  In the reference,
// OniLDACWrapper.runLDA turns the RDD of OniLDACInput(doc, word, count) into one text line per
// document ("<#entries> <word>:<count> ...", createModel, OniLDACWrapper.scala:146-168), writes
// model.dat, and lda-c parses it back (lda-data.c:22-59).  Here the same grouping is a stable
// sort of the triples by document id on the GPU: documents come out in id order, the entries of
// a document keep their input order (Spark's groupByKey order is unspecified; input order is the
// deterministic choice), duplicates (word repeated within a document) are kept as separate
// entries exactly as read_data keeps them.  String -> id dictionaries stay with the caller
// (OniLDACWrapper.scala:40-60 already owns them).
//
// Algorithm: least-significant-digit radix sort on the document id, 8 bits per pass, carrying the
// original position as payload (so the sort is stable by construction), then one gather of
// (word, count) and a histogram + exclusive scan for the CSR offsets.  Per pass: per-block digit
// histograms, one exclusive scan over the bin-major (digit, block) table, stable scatter.
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "kernels.cuh"

namespace {

constexpr int kMaxRadixBits = 8;
constexpr int kMaxBins = 1 << kMaxRadixBits;
constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
// items per thread: ITEMS (template parameter); a warp owns 32 * ITEMS consecutive triples, a block kThreads * ITEMS

// block b, digit d -> hist[d * nblocks + b]
template <int ITEMS>
__global__ void __launch_bounds__(kThreads) radix_hist_kernel(const int32_t* __restrict__ key, int64_t n, int shift, unsigned int mask,
                                                              int nblocks, unsigned int* __restrict__ hist)
{
    constexpr int kItems = ITEMS, kTile = kThreads * ITEMS;
    __shared__ unsigned int sh[kMaxBins];
    sh[threadIdx.x] = 0;
    __syncthreads();
    const int64_t base = (int64_t)blockIdx.x * kTile;
    int32_t k[kItems];
#pragma unroll
    for (int r = 0; r < kItems; r++) {
        const int64_t i = base + r * kThreads + threadIdx.x;
        k[r] = i < n ? key[i] : -1;
    }
    const bool full = base + kTile <= n;
#pragma unroll
    for (int r = 0; r < kItems; r++) {
        const unsigned int d = ((unsigned int)k[r] >> shift) & mask;
        // a corpus dominated by one document gives whole warps one digit: add those once instead of 32 ways
        const unsigned int d0 = __shfl_sync(0xffffffffu, d, 0);
        if (full && __all_sync(0xffffffffu, d == d0)) {
            if ((threadIdx.x & 31) == 0) atomicAdd(&sh[d], 32u);
        } else if (base + r * kThreads + threadIdx.x < n) {
            atomicAdd(&sh[d], 1u);
        }
    }
    __syncthreads();
    if (threadIdx.x <= mask) hist[(int64_t)threadIdx.x * nblocks + blockIdx.x] = sh[threadIdx.x];
}

// ---- exclusive scan of an unsigned array (sums below 2^31): block sums, scan of the sums by one block, final pass
constexpr int kScanTile = 2048;                  // per block: 256 threads x 8

__global__ void __launch_bounds__(kThreads) scan_sums_kernel(const unsigned int* __restrict__ in, int64_t n, unsigned int* __restrict__ sums)
{
    __shared__ unsigned int sh[kWarps];
    const int64_t base = (int64_t)blockIdx.x * kScanTile;
    unsigned int acc = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) {
        const int64_t i = base + k * kThreads + threadIdx.x;
        if (i < n) acc += in[i];
    }
    acc = __reduce_add_sync(0xffffffffu, acc);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned int t = 0;
        for (int w = 0; w < kWarps; w++) t += sh[w];
        sums[blockIdx.x] = t;
    }
}

// inclusive scan across a block of 1024 threads; returns this thread's inclusive prefix
__device__ __forceinline__ unsigned int block_inclusive_scan_1024(unsigned int v, unsigned int* warp_sums /* [32] */)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned int t = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += t;
    }
    if (lane == 31) warp_sums[warp] = v;
    __syncthreads();
    if (warp == 0) {
        unsigned int w = warp_sums[lane];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned int t = __shfl_up_sync(0xffffffffu, w, o);
            if (lane >= o) w += t;
        }
        warp_sums[lane] = w;
    }
    __syncthreads();
    if (warp > 0) v += warp_sums[warp - 1];
    return v;
}

__global__ void __launch_bounds__(1024) scan_of_sums_kernel(unsigned int* __restrict__ sums, int64_t m)
{
    // one block walks the block sums in chunks of 1024 (m is a 2048th of the scanned length)
    __shared__ unsigned int warp_sums[32];
    __shared__ unsigned int carry_s;
    unsigned int carry = 0;
    for (int64_t base = 0; base < m; base += 1024) {
        const int64_t i = base + threadIdx.x;
        const unsigned int v = i < m ? sums[i] : 0;
        const unsigned int inc = block_inclusive_scan_1024(v, warp_sums);
        if (i < m) sums[i] = carry + inc - v;   // exclusive
        if (threadIdx.x == 1023) carry_s = carry + inc;
        __syncthreads();
        carry = carry_s;
    }
}

// out[i] = exclusive prefix of in[0..i)
__global__ void __launch_bounds__(kThreads) scan_final_kernel(const unsigned int* __restrict__ in, int64_t n,
                                                              const unsigned int* __restrict__ sums, unsigned int* __restrict__ out)
{
    __shared__ unsigned int sh[kWarps];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t base = (int64_t)blockIdx.x * kScanTile + threadIdx.x * 8;
    unsigned int v[8];
    unsigned int acc = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) {
        v[k] = base + k < n ? in[base + k] : 0;
        acc += v[k];
    }
    unsigned int inc = acc;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned int t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) sh[warp] = inc;
    __syncthreads();
    unsigned int run = sums[blockIdx.x] + inc - acc;
    for (int w = 0; w < warp; w++) run += sh[w];
#pragma unroll
    for (int k = 0; k < 8; k++) {
        if (base + k < n) out[base + k] = run;
        run += v[k];
    }
}

// Stable scatter of one pass, carrying (word, count) with the document id.  A warp owns kWarpTile consecutive
// items and takes them 32 at a time, so the input order is (block, warp, round, lane).  An item's place inside
// the block's sorted tile = items of lower digits + items of its digit in lower warps + in earlier rounds of its
// warp + in lower lanes of its round.  The tile is put in that order in shared memory, one array at a time, and
// written out from there: a digit's items leave as one contiguous run (kTile / bins items on average) instead
// of one 4-byte store per sector.
template <bool TWO, int ITEMS>   // TWO: carries (word, count); otherwise one payload (count_in / count_out unused)
__global__ void __launch_bounds__(kThreads, ITEMS >= 16 ? 3 : 5) radix_scatter_kernel(const int32_t* __restrict__ key_in, const int32_t* __restrict__ word_in,
                                                                 const int32_t* __restrict__ count_in, int64_t n, int shift,
                                                                 unsigned int mask, int nblocks, const unsigned int* __restrict__ offs,
                                                                 int32_t* __restrict__ key_out, int32_t* __restrict__ word_out,
                                                                 int32_t* __restrict__ count_out)
{
    constexpr int kItems = ITEMS, kWarpTile = 32 * ITEMS, kTile = kThreads * ITEMS;
    __shared__ unsigned int cnt[kWarps][kMaxBins + 1];     // [.][kMaxBins]: the dead items of a partial tile
    __shared__ int32_t tile[kTile];
    __shared__ unsigned int gdelta[kMaxBins];              // global start of the digit's run - its start in the tile
    __shared__ unsigned int wsum[kWarps];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned int lt = (1u << lane) - 1u;
    for (int b = threadIdx.x; b < kWarps * (kMaxBins + 1); b += kThreads) (&cnt[0][0])[b] = 0;
    const int64_t block_base = (int64_t)blockIdx.x * kTile;
    const int64_t base = block_base + warp * kWarpTile + lane;
    const int live_count = (int)(n - block_base < kTile ? n - block_base : kTile);
    int32_t k[kItems], w[kItems], c[kItems];
#pragma unroll
    for (int r = 0; r < kItems; r++) {
        const int64_t i = base + r * 32;
        const bool live = i < n;
        k[r] = live ? key_in[i] : 0;
        w[r] = live ? word_in[i] : 0;
        c[r] = (TWO && live) ? count_in[i] : 0;
    }
    __syncthreads();
    unsigned int lp[kItems];
    unsigned int* mine = cnt[warp];
#pragma unroll
    for (int r = 0; r < kItems; r++) {
        const bool live = base + r * 32 < n;
        const unsigned int d = live ? (((unsigned int)k[r] >> shift) & mask) : kMaxBins;
        const unsigned int peers = __match_any_sync(0xffffffffu, d);
        const unsigned int before = mine[d];
        lp[r] = before + __popc(peers & lt);
        __syncwarp();
        if ((peers & lt) == 0) mine[d] = before + __popc(peers);
        __syncwarp();
    }
    __syncthreads();
    // per digit: counts of the warps -> exclusive prefix over warps; then an exclusive scan over the digits
    unsigned int tot = 0;
    if (threadIdx.x <= mask) {
#pragma unroll
        for (int q = 0; q < kWarps; q++) {
            const unsigned int t = cnt[q][threadIdx.x];
            cnt[q][threadIdx.x] = tot;
            tot += t;
        }
    }
    unsigned int inc = tot;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned int t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) wsum[warp] = inc;
    __syncthreads();
    unsigned int lstart = inc - tot;
    for (int q = 0; q < warp; q++) lstart += wsum[q];
    if (threadIdx.x <= mask) {
#pragma unroll
        for (int q = 0; q < kWarps; q++) cnt[q][threadIdx.x] += lstart;
        gdelta[threadIdx.x] = offs[(int64_t)threadIdx.x * nblocks + blockIdx.x] - lstart;
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < kItems; r++) {
        if (base + r * 32 < n) {
            lp[r] += mine[((unsigned int)k[r] >> shift) & mask];
            tile[lp[r]] = k[r];
        }
    }
    __syncthreads();
    unsigned int dst[kItems];
#pragma unroll
    for (int r = 0; r < kItems; r++) {
        const int j = r * kThreads + threadIdx.x;
        if (j < live_count) {
            const int32_t kk = tile[j];
            dst[r] = (unsigned int)j + gdelta[((unsigned int)kk >> shift) & mask];
            key_out[dst[r]] = kk;
        }
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < kItems; r++)
        if (base + r * 32 < n) tile[lp[r]] = w[r];
    __syncthreads();
#pragma unroll
    for (int r = 0; r < kItems; r++) {
        const int j = r * kThreads + threadIdx.x;
        if (j < live_count) word_out[dst[r]] = tile[j];
    }
    if (!TWO) return;
    __syncthreads();
#pragma unroll
    for (int r = 0; r < kItems; r++)
        if (base + r * 32 < n) tile[lp[r]] = c[r];
    __syncthreads();
#pragma unroll
    for (int r = 0; r < kItems; r++) {
        const int j = r * kThreads + threadIdx.x;
        if (j < live_count) count_out[dst[r]] = tile[j];
    }
}

// ---- (document, word) occurrences -> counts: run heads of the sorted pairs, their ranks, run lengths
__global__ void __launch_bounds__(256) pair_heads_kernel(const int32_t* __restrict__ doc, const int32_t* __restrict__ word, int64_t n,
                                                         unsigned int* __restrict__ head)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) head[i] = (i == 0 || doc[i] != doc[i - 1] || word[i] != word[i - 1]) ? 1u : 0u;
}

// rank[i] = heads before i.  A head writes its ids and where its run starts; the last element writes the number of runs.
__global__ void __launch_bounds__(256) pair_emit_kernel(const int32_t* __restrict__ doc, const int32_t* __restrict__ word, int64_t n,
                                                        const unsigned int* __restrict__ rank, int32_t* __restrict__ doc_out,
                                                        int32_t* __restrict__ word_out, unsigned int* __restrict__ start,
                                                        long long* __restrict__ n_out)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const unsigned int r = rank[i];
    const bool is_head = i == 0 || doc[i] != doc[i - 1] || word[i] != word[i - 1];
    if (is_head) {
        doc_out[r] = doc[i];
        word_out[r] = word[i];
        start[r] = (unsigned int)i;
    }
    if (i == n - 1) {
        const unsigned int m = r + (is_head ? 1u : 0u);
        start[m] = (unsigned int)n;
        *n_out = (long long)m;
    }
}

__global__ void __launch_bounds__(256) pair_counts_kernel(const unsigned int* __restrict__ start, const long long* __restrict__ n_out,
                                                          int32_t* __restrict__ count_out)
{
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j < *n_out) count_out[j] = (int32_t)(start[j + 1] - start[j]);
}

// Per-document int count sum (document.total, lda.h:27; wraps as lda-c's int does).  Equal documents are
// adjacent: a warp adds each of its runs once, and the block's first document (the only one that can cover
// whole blocks) is summed in shared memory first.
__global__ void __launch_bounds__(256) doc_totals_kernel(const int32_t* __restrict__ key, const int32_t* __restrict__ count, int64_t n,
                                                         int32_t* __restrict__ totals)
{
    __shared__ int first_sum;
    const int64_t block_base = (int64_t)blockIdx.x * blockDim.x;
    const int64_t i = block_base + threadIdx.x;
    const bool live = i < n;
    const int32_t first = key[block_base];
    if (threadIdx.x == 0) first_sum = 0;
    __syncthreads();
    const int32_t d = live ? key[i] : -1;
    const int32_t c = live ? count[i] : 0;
    const unsigned int peers = __match_any_sync(0xffffffffu, d);
    const int sum = __reduce_add_sync(peers, c);
    if (live && (peers & ((1u << (threadIdx.x & 31)) - 1u)) == 0) {
        if (d == first) atomicAdd(&first_sum, sum);
        else atomicAdd(&totals[d], sum);
    }
    __syncthreads();
    if (threadIdx.x == 0) atomicAdd(&totals[first], first_sum);
}

// offsets[d] = first position whose document id is >= d (the ids are sorted); longest document
__global__ void __launch_bounds__(256) doc_offsets_kernel(const int32_t* __restrict__ key, int64_t n, int64_t num_docs,
                                                          int64_t* __restrict__ offsets, int32_t* __restrict__ max_len)
{
    const int64_t d = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int len = 0;
    if (d <= num_docs) {
        auto lower = [&](int64_t v) {
            int64_t lo = 0, hi = n;
            while (lo < hi) {
                const int64_t mid = (lo + hi) >> 1;
                if (key[mid] < v) lo = mid + 1; else hi = mid;
            }
            return lo;
        };
        const int64_t a = lower(d);
        offsets[d] = a;
        if (max_len && d < num_docs) len = (int)(lower(d + 1) - a);
    }
    if (max_len) {
        len = __reduce_max_sync(0xffffffffu, len);
        if ((threadIdx.x & 31) == 0 && len > 0) atomicMax(max_len, len);
    }
}

// range[0] = max document id, [1] = max word id, [2] = 1 if any id is negative, [3] = sum of counts
__global__ void __launch_bounds__(256) triple_range_kernel(const int32_t* __restrict__ doc, const int32_t* __restrict__ word,
                                                           const int32_t* __restrict__ count, int64_t n, long long* __restrict__ range)
{
    int md = -1, mw = -1, neg = 0;
    long long tok = 0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const int d = doc[i], w = word[i];
        neg |= (d < 0) | (w < 0);
        md = max(md, d);
        mw = max(mw, w);
        tok += count[i];
    }
    md = __reduce_max_sync(0xffffffffu, md);
    mw = __reduce_max_sync(0xffffffffu, mw);
    neg = __reduce_or_sync(0xffffffffu, (unsigned)neg);
    for (int o = 16; o > 0; o >>= 1) tok += __shfl_down_sync(0xffffffffu, tok, o);
    if ((threadIdx.x & 31) == 0) {
        atomicMax(&range[0], (long long)md);
        atomicMax(&range[1], (long long)mw);
        if (neg) atomicMax(&range[2], 1ll);
        atomicAdd((unsigned long long*)&range[3], (unsigned long long)tok);
    }
}

// exclusive scan on `st`; `sums` holds (n + kScanTile - 1) / kScanTile entries
void exclusive_scan(const unsigned int* in, int64_t n, unsigned int* sums, unsigned int* out, cudaStream_t st)
{
    const int64_t m = (n + kScanTile - 1) / kScanTile;
    scan_sums_kernel<<<(unsigned)m, kThreads, 0, st>>>(in, n, sums);
    scan_of_sums_kernel<<<1, 1024, 0, st>>>(sums, m);
    scan_final_kernel<<<(unsigned)m, kThreads, 0, st>>>(in, n, sums, out);
}

// items per thread of the sort passes (8: five blocks per SM, 16: three; ONILDA_SORT_ITEMS overrides)
int sort_items()
{
    static int items = 0;
    if (!items) {
        const char* e = getenv("ONILDA_SORT_ITEMS");
        items = e && atoi(e) == 16 ? 16 : 8;     // measured at 110 M triples: 4.3 ms with 8, 4.8 ms with 16
    }
    return items;
}

template <bool TWO>
void sort_pass(int items, const int32_t* key_in, const int32_t* a_in, const int32_t* b_in, int64_t n, int shift, unsigned int mask,
               int nblocks, unsigned int* hist, unsigned int* sums, unsigned int* offs, int32_t* key_out, int32_t* a_out, int32_t* b_out,
               cudaStream_t st)
{
    if (items == 16) radix_hist_kernel<16><<<nblocks, kThreads, 0, st>>>(key_in, n, shift, mask, nblocks, hist);
    else radix_hist_kernel<8><<<nblocks, kThreads, 0, st>>>(key_in, n, shift, mask, nblocks, hist);
    exclusive_scan(hist, ((int64_t)mask + 1) * nblocks, sums, offs, st);
    if (items == 16)
        radix_scatter_kernel<TWO, 16><<<nblocks, kThreads, 0, st>>>(key_in, a_in, b_in, n, shift, mask, nblocks, offs, key_out, a_out, b_out);
    else
        radix_scatter_kernel<TWO, 8><<<nblocks, kThreads, 0, st>>>(key_in, a_in, b_in, n, shift, mask, nblocks, offs, key_out, a_out, b_out);
}

size_t align256(size_t b) { return (b + 255) & ~(size_t)255; }

}  // namespace

// Device-resident form: all pointers are device memory, everything is ordered on `stream`, nothing
// synchronises.  Scratch (about 16 bytes per triple) comes from the stream-ordered pool.
// offsets_out[num_docs + 1], words_out[n], counts_out[n]; totals_out[num_docs] and max_len_out[1] may be null.
extern "C" int onilda_triples_group_device(const int32_t* doc, const int32_t* word, const int32_t* count, int64_t n,
                                           int64_t num_docs, int64_t* offsets_out, int32_t* words_out, int32_t* counts_out,
                                           int32_t* totals_out, int32_t* max_len_out, void* stream)
{
    if (n < 0 || num_docs < 0 || !offsets_out || (n > 0 && (!doc || !word || !count || !words_out || !counts_out)))
        return onilda_fail(ONILDA_EINVAL, "triples_group_device: bad argument");
    if (n >= ((int64_t)1 << 31)) return onilda_fail(ONILDA_EINVAL, "triples_group_device: at most 2^31 - 1 triples per call");
    if (n > 0 && num_docs == 0) return onilda_fail(ONILDA_EINVAL, "triples_group_device: triples but no documents");
    cudaStream_t st = (cudaStream_t)stream;
    if (totals_out && num_docs > 0) cudaMemsetAsync(totals_out, 0, sizeof(int32_t) * (size_t)num_docs, st);
    if (max_len_out) cudaMemsetAsync(max_len_out, 0, sizeof(int32_t), st);
    char* arena = nullptr;
    const int32_t* sorted_key = doc;
    if (n > 0) {
        // stable LSD radix sort of (document id, original position), digits sized so that the passes are even
        int bits = 1;
        while (((int64_t)1 << bits) < num_docs) bits++;
        const int passes = (bits + kMaxRadixBits - 1) / kMaxRadixBits;
        const int digit = (bits + passes - 1) / passes;
        const unsigned int mask = (1u << digit) - 1u;
        const int items = sort_items(), tile = kThreads * items;
        const int nblocks = (int)((n + tile - 1) / tile);
        const int64_t scan_len = ((int64_t)mask + 1) * nblocks;
        const size_t b_key = align256(sizeof(int32_t) * (size_t)n), b_hist = align256(sizeof(unsigned int) * (size_t)scan_len),
                     b_sums = align256(sizeof(unsigned int) * (size_t)((scan_len + kScanTile - 1) / kScanTile + 1));
        const size_t total = 4 * b_key + 2 * b_hist + b_sums;
        if (cudaMallocAsync((void**)&arena, total, st) != cudaSuccess) {
            cudaGetLastError();
            return onilda_fail(ONILDA_ENOMEM, "triples_group_device: %zu bytes of device scratch not available", total);
        }
        char* p = arena;
        int32_t* key[2];
        key[0] = (int32_t*)p; p += b_key; key[1] = (int32_t*)p; p += b_key;
        int32_t* wtmp = (int32_t*)p; p += b_key;
        int32_t* ctmp = (int32_t*)p; p += b_key;
        unsigned int* hist = (unsigned int*)p; p += b_hist;
        unsigned int* offs = (unsigned int*)p; p += b_hist;
        unsigned int* sums = (unsigned int*)p;
        const int32_t *key_in = doc, *word_in = word, *count_in = count;
        for (int pass = 0; pass < passes; pass++) {
            const int shift = pass * digit;
            // the last pass writes the caller's arrays; the passes before it alternate so that it never reads them
            const bool to_out = ((passes - 1 - pass) & 1) == 0;
            int32_t* w_out = to_out ? words_out : wtmp;
            int32_t* c_out = to_out ? counts_out : ctmp;
            sort_pass<true>(items, key_in, word_in, count_in, n, shift, mask, nblocks, hist, sums, offs, key[pass & 1], w_out, c_out, st);
            key_in = key[pass & 1];
            word_in = w_out;
            count_in = c_out;
        }
        if (totals_out) doc_totals_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(key_in, counts_out, n, totals_out);
        sorted_key = key_in;
    }
    doc_offsets_kernel<<<(unsigned)((num_docs + 1 + 255) / 256), 256, 0, st>>>(sorted_key, n, num_docs, offsets_out, max_len_out);
    const cudaError_t err = cudaGetLastError();
    if (arena) cudaFreeAsync(arena, st);
    if (err != cudaSuccess) return onilda_fail(ONILDA_ECUDA, "triples_group_device: launch failed: %s", cudaGetErrorString(err));
    return ONILDA_OK;
}

// reduceByKey(_ + _) over ((document, word), 1): FlowPreLDA.scala:131-137, DNSPreLDA.scala:114-120.
// Sort by word, then (stably) by document, one payload; run heads -> ranks -> run lengths.  Output sorted by
// (document, word).  doc_out / word_out / count_out have room for n entries; *n_out (device) = distinct pairs.
extern "C" int onilda_pairs_count_device(const int32_t* doc, const int32_t* word, int64_t n, int64_t num_docs, int32_t num_terms,
                                         int32_t* doc_out, int32_t* word_out, int32_t* count_out, long long* n_out, void* stream)
{
    if (n < 0 || num_docs < 0 || num_terms < 0 || !n_out || (n > 0 && (!doc || !word || !doc_out || !word_out || !count_out)))
        return onilda_fail(ONILDA_EINVAL, "pairs_count_device: bad argument");
    if (n >= ((int64_t)1 << 31) - 1) return onilda_fail(ONILDA_EINVAL, "pairs_count_device: at most 2^31 - 2 pairs per call");
    if (n > 0 && (num_docs == 0 || num_terms == 0)) return onilda_fail(ONILDA_EINVAL, "pairs_count_device: pairs but no ids");
    cudaStream_t st = (cudaStream_t)stream;
    if (n == 0) {
        cudaMemsetAsync(n_out, 0, sizeof(long long), st);
        return ONILDA_OK;
    }
    const int items = sort_items(), tile = kThreads * items;
    const int nblocks = (int)((n + tile - 1) / tile);
    const int64_t scan_len = (int64_t)kMaxBins * nblocks > n + 1 ? (int64_t)kMaxBins * nblocks : n + 1;
    const size_t b_key = align256(sizeof(int32_t) * (size_t)(n + 1)), b_hist = align256(sizeof(unsigned int) * (size_t)kMaxBins * nblocks),
                 b_sums = align256(sizeof(unsigned int) * (size_t)((scan_len + kScanTile - 1) / kScanTile + 1));
    char* arena = nullptr;
    const size_t total = 4 * b_key + 2 * b_hist + b_sums;
    if (cudaMallocAsync((void**)&arena, total, st) != cudaSuccess) {
        cudaGetLastError();
        return onilda_fail(ONILDA_ENOMEM, "pairs_count_device: %zu bytes of device scratch not available", total);
    }
    char* p = arena;
    int32_t* buf[2][2];      // [pair][0 = the array that was the key of the pass, 1 = its payload]
    buf[0][0] = (int32_t*)p; p += b_key; buf[0][1] = (int32_t*)p; p += b_key;
    buf[1][0] = (int32_t*)p; p += b_key; buf[1][1] = (int32_t*)p; p += b_key;
    unsigned int* hist = (unsigned int*)p; p += b_hist;
    unsigned int* offs = (unsigned int*)p; p += b_hist;
    unsigned int* sums = (unsigned int*)p;
    const int32_t *key_in = word, *pay_in = doc;
    int flip = 0;
    for (int phase = 0; phase < 2; phase++) {
        const int64_t range = phase == 0 ? (int64_t)num_terms : num_docs;
        int bits = 1;
        while (((int64_t)1 << bits) < range) bits++;
        const int passes = (bits + kMaxRadixBits - 1) / kMaxRadixBits;
        const int digit = (bits + passes - 1) / passes;
        const unsigned int mask = (1u << digit) - 1u;
        for (int pass = 0; pass < passes; pass++, flip ^= 1) {
            sort_pass<false>(items, key_in, pay_in, nullptr, n, pass * digit, mask, nblocks, hist, sums, offs, buf[flip][0], buf[flip][1],
                             nullptr, st);
            key_in = buf[flip][0];
            pay_in = buf[flip][1];
        }
        if (phase == 0) {            // now by document, carrying the word
            const int32_t* t = key_in; key_in = pay_in; pay_in = t;
        }
    }
    // key_in = documents, pay_in = words, sorted by (document, word); the other pair of buffers is free
    unsigned int* head = (unsigned int*)buf[flip][0];
    unsigned int* rank = (unsigned int*)buf[flip][1];
    const unsigned int grid = (unsigned int)((n + 255) / 256);
    pair_heads_kernel<<<grid, 256, 0, st>>>(key_in, pay_in, n, head);
    exclusive_scan(head, n, sums, rank, st);
    pair_emit_kernel<<<grid, 256, 0, st>>>(key_in, pay_in, n, rank, doc_out, word_out, head /* reused: run starts */, n_out);
    pair_counts_kernel<<<grid, 256, 0, st>>>(head, n_out, count_out);
    const cudaError_t err = cudaGetLastError();
    cudaFreeAsync(arena, st);
    if (err != cudaSuccess) return onilda_fail(ONILDA_ECUDA, "pairs_count_device: launch failed: %s", cudaGetErrorString(err));
    return ONILDA_OK;
}

// Host arrays in, host arrays out (room for n entries each); *n_out = distinct pairs.
extern "C" int onilda_pairs_count(const int32_t* doc, const int32_t* word, int64_t n, int64_t num_docs, int32_t num_terms, int device,
                                  int32_t* doc_out, int32_t* word_out, int32_t* count_out, int64_t* n_out)
{
    if (n < 0 || !n_out || (n > 0 && (!doc || !word || !doc_out || !word_out || !count_out)))
        return onilda_fail(ONILDA_EINVAL, "pairs_count: bad argument");
    if (onilda_device_count() == 0)
        return onilda_fail(ONILDA_ECUDA, "pairs_count: no CUDA device is usable and this library has no CPU path");
    *n_out = 0;
    if (n == 0) return ONILDA_OK;
    int prev = 0;
    ONILDA_CUDA(cudaGetDevice(&prev));
    if (device < 0) device = prev;
    ONILDA_CUDA(cudaSetDevice(device));
    struct Restore { int d; ~Restore() { cudaSetDevice(d); } } restore{prev};
    cudaStream_t st = 0;
    struct DevBuf { void* p = nullptr; ~DevBuf() { cudaFree(p); } } d;
    const size_t b_n = align256(sizeof(int32_t) * (size_t)n);
    if (cudaMalloc(&d.p, 5 * b_n + 256) != cudaSuccess) {
        cudaGetLastError();
        return onilda_fail(ONILDA_ENOMEM, "pairs_count: %zu bytes of device memory not available", 5 * b_n);
    }
    char* base = (char*)d.p;
    int32_t* d_doc = (int32_t*)base; int32_t* d_word = (int32_t*)(base + b_n);
    int32_t* o_doc = (int32_t*)(base + 2 * b_n); int32_t* o_word = (int32_t*)(base + 3 * b_n); int32_t* o_cnt = (int32_t*)(base + 4 * b_n);
    long long* d_range = (long long*)(base + 5 * b_n);          // [0..3]: id range check, [4]: number of runs
    long long range[5] = {-1, -1, 0, 0, 0};
    ONILDA_CUDA(cudaMemcpyAsync(d_doc, doc, sizeof(int32_t) * (size_t)n, cudaMemcpyHostToDevice, st));
    ONILDA_CUDA(cudaMemcpyAsync(d_word, word, sizeof(int32_t) * (size_t)n, cudaMemcpyHostToDevice, st));
    ONILDA_CUDA(cudaMemcpyAsync(d_range, range, sizeof range, cudaMemcpyHostToDevice, st));
    triple_range_kernel<<<148 * 8, 256, 0, st>>>(d_doc, d_word, d_word, n, d_range);
    ONILDA_CUDA(cudaGetLastError());
    ONILDA_CUDA(cudaMemcpyAsync(range, d_range, sizeof range, cudaMemcpyDeviceToHost, st));
    ONILDA_CUDA(cudaStreamSynchronize(st));
    if (range[2]) return onilda_fail(ONILDA_EINVAL, "pairs_count: negative document or word id");
    if (num_docs == 0) num_docs = range[0] + 1;
    if (num_terms == 0) num_terms = (int32_t)(range[1] + 1);
    if (range[0] >= num_docs || range[1] >= num_terms)
        return onilda_fail(ONILDA_EINVAL, "pairs_count: id out of range (document %lld of %lld, word %lld of %d)", range[0],
                           (long long)num_docs, range[1], num_terms);
    int rc = onilda_pairs_count_device(d_doc, d_word, n, num_docs, num_terms, o_doc, o_word, o_cnt, d_range + 4, st);
    if (rc != ONILDA_OK) return rc;
    long long m = 0;
    ONILDA_CUDA(cudaMemcpyAsync(&m, d_range + 4, sizeof m, cudaMemcpyDeviceToHost, st));
    ONILDA_CUDA(cudaStreamSynchronize(st));
    ONILDA_CUDA(cudaMemcpyAsync(doc_out, o_doc, sizeof(int32_t) * (size_t)m, cudaMemcpyDeviceToHost, st));
    ONILDA_CUDA(cudaMemcpyAsync(word_out, o_word, sizeof(int32_t) * (size_t)m, cudaMemcpyDeviceToHost, st));
    ONILDA_CUDA(cudaMemcpyAsync(count_out, o_cnt, sizeof(int32_t) * (size_t)m, cudaMemcpyDeviceToHost, st));
    ONILDA_CUDA(cudaStreamSynchronize(st));
    *n_out = m;
    return ONILDA_OK;
}

// Device triples (ids in range) -> grouped on the device -> host corpus.
static int corpus_from_device_triples(const int32_t* d_doc, const int32_t* d_word, const int32_t* d_count, int64_t n, int64_t num_docs,
                                      int32_t num_terms, long long tokens, cudaStream_t st, onilda_corpus** out)
{
    struct DevBuf {
        void* p = nullptr;
        ~DevBuf() { cudaFree(p); }
    } d_out;
    const size_t b_n = align256(sizeof(int32_t) * (size_t)n);
    onilda_corpus* c = (onilda_corpus*)calloc(1, sizeof *c);
    if (!c) return onilda_fail(ONILDA_ENOMEM, "corpus_from_triples: out of host memory");
    c->num_docs = num_docs;
    c->nnz = n;
    c->tokens = tokens;
    c->num_terms = num_terms;
    c->offsets = (int64_t*)malloc(sizeof(int64_t) * (size_t)(num_docs + 1));
    c->totals = (int32_t*)malloc(sizeof(int32_t) * (size_t)(num_docs > 0 ? num_docs : 1));
    c->words = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
    c->counts = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
    if (!c->offsets || !c->totals || !c->words || !c->counts) {
        onilda_corpus_free(c);
        return onilda_fail(ONILDA_ENOMEM, "corpus_from_triples: out of host memory");
    }
    const size_t b_off = align256(sizeof(int64_t) * (size_t)(num_docs + 1)), b_tot = align256(sizeof(int32_t) * (size_t)(num_docs + 1));
    if (cudaMalloc(&d_out.p, b_off + b_tot + 2 * b_n + 256) != cudaSuccess) {
        cudaGetLastError();
        onilda_corpus_free(c);
        return onilda_fail(ONILDA_ENOMEM, "corpus_from_triples: device memory for the grouped corpus not available");
    }
    char* o = (char*)d_out.p;
    int64_t* d_off = (int64_t*)o;
    int32_t* d_tot = (int32_t*)(o + b_off);
    int32_t* d_w = (int32_t*)(o + b_off + b_tot);
    int32_t* d_c = (int32_t*)(o + b_off + b_tot + b_n);
    int32_t* d_maxlen = (int32_t*)(o + b_off + b_tot + 2 * b_n);
    int rc = onilda_triples_group_device(d_doc, d_word, d_count, n, num_docs, d_off, d_w, d_c, d_tot, d_maxlen, st);
    cudaError_t err = cudaSuccess;
    if (rc == ONILDA_OK) {
        err = cudaMemcpyAsync(c->offsets, d_off, sizeof(int64_t) * (size_t)(num_docs + 1), cudaMemcpyDeviceToHost, st);
        if (err == cudaSuccess && num_docs > 0) err = cudaMemcpyAsync(c->totals, d_tot, sizeof(int32_t) * (size_t)num_docs, cudaMemcpyDeviceToHost, st);
        if (err == cudaSuccess && n > 0) err = cudaMemcpyAsync(c->words, d_w, sizeof(int32_t) * (size_t)n, cudaMemcpyDeviceToHost, st);
        if (err == cudaSuccess && n > 0) err = cudaMemcpyAsync(c->counts, d_c, sizeof(int32_t) * (size_t)n, cudaMemcpyDeviceToHost, st);
        if (err == cudaSuccess) err = cudaMemcpyAsync(&c->max_len, d_maxlen, sizeof(int32_t), cudaMemcpyDeviceToHost, st);
        if (err == cudaSuccess) err = cudaStreamSynchronize(st);
        if (err != cudaSuccess) rc = onilda_fail(ONILDA_ECUDA, "corpus_from_triples: %s", cudaGetErrorString(err));
    }
    if (rc != ONILDA_OK) {
        onilda_corpus_free(c);
        return rc;
    }
    *out = c;
    return ONILDA_OK;
}

extern "C" int onilda_corpus_from_triples(const int32_t* doc, const int32_t* word, const int32_t* count, int64_t n,
                                          int64_t num_docs, int32_t num_terms, int device, onilda_corpus** out)
{
    if (!out || n < 0 || (n > 0 && (!doc || !word || !count)) || num_docs < 0)
        return onilda_fail(ONILDA_EINVAL, "corpus_from_triples: bad argument");
    if (n >= ((int64_t)1 << 31)) return onilda_fail(ONILDA_EINVAL, "corpus_from_triples: at most 2^31 - 1 triples per call");
    if (onilda_device_count() == 0)
        return onilda_fail(ONILDA_ECUDA, "corpus_from_triples: no CUDA device is usable and this library has no CPU path");
    int prev = 0;
    ONILDA_CUDA(cudaGetDevice(&prev));
    if (device < 0) device = prev;
    ONILDA_CUDA(cudaSetDevice(device));
    struct Restore { int d; ~Restore() { cudaSetDevice(d); } } restore{prev};
    cudaStream_t st = 0;

    struct DevBuf {
        void* p = nullptr;
        ~DevBuf() { cudaFree(p); }
    } d_in, d_range;
    const size_t b_n = align256(sizeof(int32_t) * (size_t)n);
    long long range[4] = {-1, -1, 0, 0};
    if (n > 0) {
        if (cudaMalloc(&d_in.p, 3 * b_n) != cudaSuccess || cudaMalloc(&d_range.p, sizeof range) != cudaSuccess) {
            cudaGetLastError();
            return onilda_fail(ONILDA_ENOMEM, "corpus_from_triples: %zu bytes of device memory not available", 3 * b_n);
        }
        char* in = (char*)d_in.p;
        ONILDA_CUDA(cudaMemcpyAsync(in, doc, sizeof(int32_t) * (size_t)n, cudaMemcpyHostToDevice, st));
        ONILDA_CUDA(cudaMemcpyAsync(in + b_n, word, sizeof(int32_t) * (size_t)n, cudaMemcpyHostToDevice, st));
        ONILDA_CUDA(cudaMemcpyAsync(in + 2 * b_n, count, sizeof(int32_t) * (size_t)n, cudaMemcpyHostToDevice, st));
        ONILDA_CUDA(cudaMemcpyAsync(d_range.p, range, sizeof range, cudaMemcpyHostToDevice, st));
        triple_range_kernel<<<148 * 8, 256, 0, st>>>((const int32_t*)in, (const int32_t*)(in + b_n), (const int32_t*)(in + 2 * b_n), n,
                                                     (long long*)d_range.p);
        ONILDA_CUDA(cudaGetLastError());
        ONILDA_CUDA(cudaMemcpyAsync(range, d_range.p, sizeof range, cudaMemcpyDeviceToHost, st));
        ONILDA_CUDA(cudaStreamSynchronize(st));
        if (range[2]) return onilda_fail(ONILDA_EINVAL, "corpus_from_triples: negative document or word id");
        if (num_docs == 0) num_docs = range[0] + 1;
        if (range[0] >= num_docs)
            return onilda_fail(ONILDA_EINVAL, "corpus_from_triples: document id %lld >= num_docs %lld", range[0], (long long)num_docs);
    }
    const char* in = (const char*)d_in.p;
    return corpus_from_device_triples((const int32_t*)in, (const int32_t*)(in + b_n), (const int32_t*)(in + 2 * b_n), n, num_docs,
                                      num_terms > (int32_t)(range[1] + 1) ? num_terms : (int32_t)(range[1] + 1), range[3], st, out);
}

// ---- FlowPreLDA in one call: event columns -> words -> dictionary -> word counts per role -> corpus
namespace {

__global__ void __launch_bounds__(256) mark_codes_kernel(const int32_t* __restrict__ a, const int32_t* __restrict__ b, int64_t n,
                                                         unsigned int* __restrict__ bitmap)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const unsigned int ca = (unsigned int)a[i], cb = (unsigned int)b[i];
    // most events of a warp share a handful of words: test before the atomic
    if (!(bitmap[ca >> 5] >> (ca & 31) & 1u)) atomicOr(&bitmap[ca >> 5], 1u << (ca & 31));
    if (!(bitmap[cb >> 5] >> (cb & 31) & 1u)) atomicOr(&bitmap[cb >> 5], 1u << (cb & 31));
}

__global__ void __launch_bounds__(256) popc_kernel(const unsigned int* __restrict__ bitmap, int64_t words, unsigned int* __restrict__ cnt)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < words) cnt[i] = (unsigned int)__popc(bitmap[i]);
}

// word id = number of distinct codes below this one (the dictionary is the sorted list of distinct codes)
__global__ void __launch_bounds__(256) code_ranks_kernel(const int32_t* __restrict__ code, int64_t n, const unsigned int* __restrict__ bitmap,
                                                         const unsigned int* __restrict__ before, int32_t* __restrict__ word)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const unsigned int c = (unsigned int)code[i];
    word[i] = (int32_t)(before[c >> 5] + (unsigned int)__popc(bitmap[c >> 5] & ((1u << (c & 31)) - 1u)));
}

__global__ void __launch_bounds__(256) vocab_kernel(const unsigned int* __restrict__ bitmap, int64_t words,
                                                    const unsigned int* __restrict__ before, int32_t* __restrict__ vocab)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= words) return;
    unsigned int m = bitmap[i], r = before[i];
    while (m) {
        const int b = __ffs(m) - 1;
        vocab[r++] = (int32_t)(i * 32 + b);
        m &= m - 1;
    }
}

}  // namespace

extern "C" int onilda_flow_prelda(const int32_t* hour, const int32_t* minute, const int32_t* second, const int64_t* ibyt,
                                  const int64_t* ipkt, const int32_t* sport, const int32_t* dport, const int32_t* sip,
                                  const int32_t* dip, int64_t n, int64_t num_docs, int device, onilda_flow_cuts* cuts_out,
                                  onilda_corpus** corpus_out, int32_t* vocab_codes_out, int32_t vocab_cap, int32_t* num_terms_out,
                                  int32_t* src_word_out, int32_t* dst_word_out)
{
    if (n < 0 || num_docs < 0 || !cuts_out || !corpus_out || !num_terms_out || vocab_cap < 0 || (vocab_cap > 0 && !vocab_codes_out) ||
        (n > 0 && (!hour || !minute || !second || !ibyt || !ipkt || !sport || !dport || !sip || !dip)))
        return onilda_fail(ONILDA_EINVAL, "flow_prelda: bad argument");
    if (2 * n >= ((int64_t)1 << 31) - 1) return onilda_fail(ONILDA_EINVAL, "flow_prelda: at most 2^30 - 1 events per call");
    if (onilda_device_count() == 0)
        return onilda_fail(ONILDA_ECUDA, "flow_prelda: no CUDA device is usable and this library has no CPU path");
    int prev = 0;
    ONILDA_CUDA(cudaGetDevice(&prev));
    if (device < 0) device = prev;
    ONILDA_CUDA(cudaSetDevice(device));
    struct Restore { int d; ~Restore() { cudaSetDevice(d); } } restore{prev};
    cudaStream_t st = 0;
    *num_terms_out = 0;
    if (n == 0) {
        int rc = onilda_flow_words_device(nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, 0, cuts_out, nullptr, nullptr, nullptr);
        if (rc != ONILDA_OK) return rc;
        std::vector<int64_t> off((size_t)num_docs + 1, 0);
        int32_t dummy = 0;
        return onilda_corpus_from_csr(off.data(), &dummy, &dummy, num_docs, 0, corpus_out);
    }
    struct DevBuf { void* p = nullptr; ~DevBuf() { cudaFree(p); } } d_cols, d_work;
    const size_t b4 = align256(sizeof(int32_t) * (size_t)n), b8 = align256(sizeof(int64_t) * (size_t)n);
    const int64_t bm_words = (ONILDA_FLOW_CODE_SPACE + 31) / 32;
    const size_t b_bm = align256(sizeof(unsigned int) * (size_t)bm_words);
    if (cudaMalloc(&d_cols.p, 7 * b4 + 2 * b8 + 4 * b4) != cudaSuccess ||
        cudaMalloc(&d_work.p, 3 * b_bm + align256(sizeof(unsigned int) * (size_t)(bm_words / 2048 + 2)) + 6 * b4 + 512) != cudaSuccess) {
        cudaGetLastError();
        return onilda_fail(ONILDA_ENOMEM, "flow_prelda: device memory for %lld events not available", (long long)n);
    }
    char* p = (char*)d_cols.p;
    int64_t* d_ibyt = (int64_t*)p; p += b8;
    int64_t* d_ipkt = (int64_t*)p; p += b8;
    int32_t* col[11];
    for (int k = 0; k < 11; k++) { col[k] = (int32_t*)p; p += b4; }
    const int32_t* src[7] = {hour, minute, second, sport, dport, sip, dip};
    for (int k = 0; k < 7; k++) ONILDA_CUDA(cudaMemcpyAsync(col[k], src[k], sizeof(int32_t) * (size_t)n, cudaMemcpyHostToDevice, st));
    ONILDA_CUDA(cudaMemcpyAsync(d_ibyt, ibyt, sizeof(int64_t) * (size_t)n, cudaMemcpyHostToDevice, st));
    ONILDA_CUDA(cudaMemcpyAsync(d_ipkt, ipkt, sizeof(int64_t) * (size_t)n, cudaMemcpyHostToDevice, st));
    int32_t *d_sip = col[5], *d_dip = col[6], *d_scode = col[7], *d_dcode = col[8], *d_sw = col[9], *d_dw = col[10];
    char* q = (char*)d_work.p;
    unsigned int* bitmap = (unsigned int*)q; q += b_bm;
    unsigned int* cnt = (unsigned int*)q; q += b_bm;
    unsigned int* before = (unsigned int*)q; q += b_bm;
    unsigned int* sums = (unsigned int*)q; q += align256(sizeof(unsigned int) * (size_t)(bm_words / 2048 + 2));
    int32_t* o_doc = (int32_t*)q; q += 2 * b4;
    int32_t* o_word = (int32_t*)q; q += 2 * b4;
    int32_t* o_cnt = (int32_t*)q; q += 2 * b4;
    long long* d_scalars = (long long*)q;      // [0..3] id range, [4] m1, [5] m2
    // address ids
    long long range[6] = {-1, -1, 0, 0, 0, 0};
    ONILDA_CUDA(cudaMemcpyAsync(d_scalars, range, sizeof range, cudaMemcpyHostToDevice, st));
    triple_range_kernel<<<148 * 8, 256, 0, st>>>(d_sip, d_dip, d_dip, n, d_scalars);
    ONILDA_CUDA(cudaGetLastError());
    ONILDA_CUDA(cudaMemcpyAsync(range, d_scalars, sizeof range, cudaMemcpyDeviceToHost, st));
    ONILDA_CUDA(cudaStreamSynchronize(st));
    if (range[2]) return onilda_fail(ONILDA_EINVAL, "flow_prelda: negative address id");
    const long long max_ip = range[0] > range[1] ? range[0] : range[1];
    if (num_docs == 0) num_docs = max_ip + 1;
    if (max_ip >= num_docs) return onilda_fail(ONILDA_EINVAL, "flow_prelda: address id %lld >= num_docs %lld", max_ip, (long long)num_docs);
    // words (FlowWordCreation), then the dictionary: the sorted distinct codes
    int rc = onilda_flow_words_device(col[0], col[1], col[2], d_ibyt, d_ipkt, col[3], col[4], n, cuts_out, d_scode, d_dcode, nullptr);
    if (rc != ONILDA_OK) return rc;
    const unsigned int grid_n = (unsigned int)((n + 255) / 256), grid_w = (unsigned int)((bm_words + 255) / 256);
    ONILDA_CUDA(cudaMemsetAsync(bitmap, 0, sizeof(unsigned int) * (size_t)bm_words, st));
    mark_codes_kernel<<<grid_n, 256, 0, st>>>(d_scode, d_dcode, n, bitmap);
    popc_kernel<<<grid_w, 256, 0, st>>>(bitmap, bm_words, cnt);
    exclusive_scan(cnt, bm_words, sums, before, st);
    unsigned int last_before = 0, last_cnt = 0;
    ONILDA_CUDA(cudaMemcpyAsync(&last_before, before + bm_words - 1, sizeof(unsigned int), cudaMemcpyDeviceToHost, st));
    ONILDA_CUDA(cudaMemcpyAsync(&last_cnt, cnt + bm_words - 1, sizeof(unsigned int), cudaMemcpyDeviceToHost, st));
    ONILDA_CUDA(cudaStreamSynchronize(st));
    const int32_t V = (int32_t)(last_before + last_cnt);
    *num_terms_out = V;
    if ((int64_t)V > bm_words) return onilda_fail(ONILDA_EINVAL, "flow_prelda: %d distinct words is more than this call handles", V);
    if (vocab_codes_out && V > vocab_cap)
        return onilda_fail(ONILDA_EINVAL, "flow_prelda: %d distinct words, room for %d (the number is in *num_terms_out)", V, vocab_cap);
    code_ranks_kernel<<<grid_n, 256, 0, st>>>(d_scode, n, bitmap, before, d_sw);
    code_ranks_kernel<<<grid_n, 256, 0, st>>>(d_dcode, n, bitmap, before, d_dw);
    if (vocab_codes_out) {
        vocab_kernel<<<grid_w, 256, 0, st>>>(bitmap, bm_words, before, (int32_t*)cnt /* reused: V <= bm_words */);
        ONILDA_CUDA(cudaMemcpyAsync(vocab_codes_out, cnt, sizeof(int32_t) * (size_t)V, cudaMemcpyDeviceToHost, st));
    }
    ONILDA_CUDA(cudaGetLastError());
    // word counts per role (FlowPreLDA.scala:131-137), the union, the corpus
    rc = onilda_pairs_count_device(d_sip, d_sw, n, num_docs, V, o_doc, o_word, o_cnt, d_scalars + 4, st);
    if (rc != ONILDA_OK) return rc;
    long long m1 = 0, m2 = 0;
    ONILDA_CUDA(cudaMemcpyAsync(&m1, d_scalars + 4, sizeof m1, cudaMemcpyDeviceToHost, st));
    ONILDA_CUDA(cudaStreamSynchronize(st));
    rc = onilda_pairs_count_device(d_dip, d_dw, n, num_docs, V, o_doc + m1, o_word + m1, o_cnt + m1, d_scalars + 5, st);
    if (rc != ONILDA_OK) return rc;
    ONILDA_CUDA(cudaMemcpyAsync(&m2, d_scalars + 5, sizeof m2, cudaMemcpyDeviceToHost, st));
    ONILDA_CUDA(cudaStreamSynchronize(st));
    if (src_word_out) ONILDA_CUDA(cudaMemcpyAsync(src_word_out, d_sw, sizeof(int32_t) * (size_t)n, cudaMemcpyDeviceToHost, st));
    if (dst_word_out) ONILDA_CUDA(cudaMemcpyAsync(dst_word_out, d_dw, sizeof(int32_t) * (size_t)n, cudaMemcpyDeviceToHost, st));
    return corpus_from_device_triples(o_doc, o_word, o_cnt, m1 + m2, num_docs, V, 2 * (long long)n, st, corpus_out);
}
