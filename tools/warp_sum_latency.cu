// Latency of a warp-wide FP64 all-reduce (every lane ends with the sum of 32 values), as a dependent chain:
//   (a) five butterfly rounds (2 SHFL.BFLY + DADD each), what group_sum<32> does;
//   (b) two mma.sync.m8n8k4.f64 with A = ones (column sums over the 4 lanes of a quad-row, then over the 8
//       partials) and one DADD between them.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o warp_sum_latency tools/warp_sum_latency.cu
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ double bfly_sum(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__device__ __forceinline__ void dmma_ones(double b, double& d0, double& d1)
{
    const double one = 1.0, z = 0.0;
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%4,%5};"
                 : "=d"(d0), "=d"(d1) : "d"(one), "d"(b), "d"(z), "d"(z));
}

__device__ __forceinline__ double mma_sum(double v)
{
    double c0, c1, t0, t1;
    dmma_ones(v, c0, c1);          // c0, c1: sums over lanes 8q..8q+3 and 8q+4..8q+7, q = lane % 4
    dmma_ones(c0 + c1, t0, t1);    // sum over q of the four partials
    return t0;
}

template <int MODE>
__global__ void chain(double* out, long long* cycles, int iters, double seed)
{
    double v = seed + threadIdx.x * 1e-3;
    long long t0 = clock64();
    for (int i = 0; i < iters; i++) {
        double s = MODE == 0 ? bfly_sum(v) : mma_sum(v);
        v = fma(s, 1e-3, v * 0.5);       // next input depends on the sum (2 more FP64 ops on the chain)
    }
    long long t1 = clock64();
    out[threadIdx.x] = v;
    if (threadIdx.x == 0) *cycles = t1 - t0;
}

__global__ void check(double* out)
{
    double v = 1.0 + threadIdx.x * 0.25 + (threadIdx.x % 3) * 1e-9;
    out[threadIdx.x] = bfly_sum(v);
    out[32 + threadIdx.x] = mma_sum(v);
}

int main()
{
    double* d; long long* c;
    cudaMalloc(&d, 64 * sizeof(double)); cudaMalloc(&c, sizeof(long long));
    double h[64];
    check<<<1, 32>>>(d);
    cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
    double worst = 0;
    for (int i = 0; i < 32; i++) { double e = (h[32 + i] - h[i]) / h[i]; if (e < 0) e = -e; if (e > worst) worst = e; }
    printf("check: butterfly %.17g mma %.17g worst rel diff %.3g (all lanes)\n", h[0], h[32], worst);
    const int iters = 100000;
    for (int rep = 0; rep < 2; rep++) {
        long long cy;
        chain<0><<<1, 32>>>(d, c, iters, 1.0); cudaMemcpy(&cy, c, 8, cudaMemcpyDeviceToHost);
        printf("butterfly x5 + 2 fp64: %.1f cycles per reduction\n", (double)cy / iters);
        chain<1><<<1, 32>>>(d, c, iters, 1.0); cudaMemcpy(&cy, c, 8, cudaMemcpyDeviceToHost);
        printf("2 x dmma + dadd + 2 fp64: %.1f cycles per reduction\n", (double)cy / iters);
    }
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
