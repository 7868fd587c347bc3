// What can the memory system deliver to a gather of whole table rows?  The scorer's reads are rows of K doubles
// (160 B at K = 20) picked by event ids from a phi table of 16 MB and a theta table of 160 MB; both are served
// mostly by the 126 MB L2 (ncu: 84 % hits).  This program measures the ceiling for exactly that access shape, so that
// the scoring kernel's roofline has a measured L2 denominator instead of the HBM figure its traffic never meets:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/l2_gather tools/l2_gather_bench.cu && /tmp/l2_gather
// For each table size it gathers `rows` random rows with 16-byte loads, 32 lanes covering consecutive 16-byte pieces
// of 3.2 rows per instruction (the scorer's pattern), many independent loads in flight, and reports GB/s of rows
// delivered.  Sizes: 16 MB (phi: L2 resident), 64 MB, 160 MB (theta: beyond the 126 MB L2 for uniform ids),
// 1 GB (HBM gather).  Output is kept in profiles/.
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

constexpr int K = 20, C = K / 2;          // 10 sixteen-byte pieces per row

__global__ void __launch_bounds__(256) gather(const double* __restrict__ table, const int32_t* __restrict__ ids, int64_t n,
                                              double* __restrict__ out)
{
    const int lane = threadIdx.x & 31;
    const int64_t warp = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int64_t warps = (int64_t)gridDim.x * (blockDim.x >> 5);
    double acc = 0.0;
    for (int64_t tile = warp; tile * 32 < n; tile += warps) {
        const int32_t my = ids[tile * 32 + lane];
        int r = lane / C, q = lane % C;
        double2 v[C];
#pragma unroll
        for (int c = 0; c < C; c++) {          // piece c*32 + lane of the tile's 32 rows
            const int32_t row = __shfl_sync(0xffffffffu, my, r);
            v[c] = __ldg(reinterpret_cast<const double2*>(table + (int64_t)row * K + 2 * q));
            r += 32 / C;
            q += 32 % C;
            if (q >= C) { q -= C; r++; }
        }
#pragma unroll
        for (int c = 0; c < C; c++) acc += v[c].x + v[c].y;
    }
    if (acc == 123.456) out[0] = acc;          // keep the loads
}

int main()
{
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    const int64_t n = 1ll << 28;               // rows gathered per launch: 43 GB of rows
    int32_t* ids;
    double* out;
    CK(cudaMalloc(&ids, n * 4));
    CK(cudaMalloc(&out, 8));
    int32_t* h = (int32_t*)malloc(n * 4);
    const double sizes_mb[] = {16, 64, 160, 1024};
    for (double mb : sizes_mb) {
        const int64_t rows = (int64_t)(mb * 1e6 / (K * 8));
        double* table;
        CK(cudaMalloc(&table, rows * K * 8));
        CK(cudaMemset(table, 0, rows * K * 8));
        uint64_t x = 88172645463325252ull;
        for (int64_t i = 0; i < n; i++) { x ^= x << 13; x ^= x >> 7; x ^= x << 17; h[i] = (int32_t)(x % (uint64_t)rows); }
        CK(cudaMemcpy(ids, h, n * 4, cudaMemcpyHostToDevice));
        cudaEvent_t a, b;
        CK(cudaEventCreate(&a));
        CK(cudaEventCreate(&b));
        for (int occ : {4, 8}) {
            for (int it = 0; it < 2; it++) gather<<<sms * occ, 256>>>(table, ids, n, out);
            CK(cudaEventRecord(a));
            for (int it = 0; it < 3; it++) gather<<<sms * occ, 256>>>(table, ids, n, out);
            CK(cudaEventRecord(b));
            CK(cudaDeviceSynchronize());
            float ms;
            CK(cudaEventElapsedTime(&ms, a, b));
            ms /= 3;
            printf("table %6.0f MB (%9lld rows of %d B), %d blocks/SM: %7.2f ms per 2^28 rows -> %7.1f GB/s of rows, %5.2f G rows/s\n",
                   mb, (long long)rows, K * 8, occ, ms, n * (double)(K * 8) / ms / 1e6, n / ms / 1e6);
        }
        CK(cudaFree(table));
    }
    return 0;
}
