// Micro-measurements that shaped device_math.cuh / estep.cu (run on the B200 box):
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/microbench tools/microbench.cu && /tmp/microbench
// Prints dependent-issue latencies (cycles) of the FP64 ops on the E-step's serial chain, the
// accuracy of the MUFU.RCP64H seed, and the FP64 issue rate of one warp / one SM.
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

template <int OP>
__global__ void chain(double* out, long long* cycles, int iters, double a, double b)
{
    double x = a + threadIdx.x * 1e-9;
    long long t0 = clock64();
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int u = 0; u < 16; u++) {
            if (OP == 0) x = fma(x, b, a);
            if (OP == 1) x = x * b;
            if (OP == 2) x = x + b;
            if (OP == 3) { double r; asm volatile("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x)); x = r; }
            if (OP == 4) x = __shfl_xor_sync(0xffffffffu, x, 1);
            if (OP == 5) x = __shfl_xor_sync(0xffffffffu, x, 1) + x;
            if (OP == 6) { float f = (float)x; f = __log2f(f); x = (double)f; }
        }
    }
    long long t1 = clock64();
    if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
    out[blockIdx.x * blockDim.x + threadIdx.x] = x;
}

// ILP independent chains per thread: FP64 issue rate
template <int ILP>
__global__ void tput(double* out, long long* cycles, int iters, double b)
{
    double x[ILP];
#pragma unroll
    for (int i = 0; i < ILP; i++) x[i] = 1.0 + threadIdx.x * 1e-9 + i;
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int u = 0; u < 8; u++)
#pragma unroll
            for (int i = 0; i < ILP; i++) x[i] = fma(x[i], b, 1e-9);
    }
    long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int i = 0; i < ILP; i++) s += x[i];
    if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void rcp_seed_err(const double* x, double* err, int n)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x[i]));
    err[i] = fabs(r * x[i] - 1.0);
}

int main()
{
    double* out; long long* cyc;
    CK(cudaMalloc(&out, sizeof(double) * 1 << 20));
    CK(cudaMallocManaged(&cyc, sizeof(long long)));
    const int iters = 2000;
    const char* names[] = {"DFMA", "DMUL", "DADD", "MUFU.RCP64H", "SHFL.64", "SHFL.64+DADD", "F2F+MUFU.LG2+F2F"};
#define RUN(OP) chain<OP><<<1, 32>>>(out, cyc, iters, 1.0000001, 0.9999999); CK(cudaDeviceSynchronize()); \
    printf("dependent latency %-18s %.1f cycles\n", names[OP], (double)*cyc / (iters * 16.0));
    RUN(0) RUN(1) RUN(2) RUN(3) RUN(4) RUN(5) RUN(6)
    // issue rate, one warp
#define TP(ILP, W) tput<ILP><<<1, 32 * W>>>(out, cyc, iters, 0.9999999); CK(cudaDeviceSynchronize()); \
    printf("DFMA issue: %d warp(s) x ILP %d : %.2f cycles per warp-instruction (block view)\n", W, ILP, (double)*cyc / (iters * 8.0 * ILP));
    TP(1, 1) TP(2, 1) TP(4, 1) TP(8, 1) TP(8, 4) TP(8, 8) TP(8, 16) TP(4, 16)
    // seed accuracy
    const int n = 1 << 20;
    double *hx = (double*)malloc(sizeof(double) * n), *dx, *derr, *herr = (double*)malloc(sizeof(double) * n);
    srand(1);
    for (int i = 0; i < n; i++) hx[i] = exp(((double)rand() / RAND_MAX - 0.5) * 1200.0) * (1.0 + (double)rand() / RAND_MAX);
    CK(cudaMalloc(&dx, sizeof(double) * n)); CK(cudaMalloc(&derr, sizeof(double) * n));
    CK(cudaMemcpy(dx, hx, sizeof(double) * n, cudaMemcpyHostToDevice));
    rcp_seed_err<<<n / 256, 256>>>(dx, derr, n);
    CK(cudaMemcpy(herr, derr, sizeof(double) * n, cudaMemcpyDeviceToHost));
    double mx = 0;
    for (int i = 0; i < n; i++) if (herr[i] > mx) mx = herr[i];
    printf("rcp.approx.ftz.f64 seed: max |r*x - 1| = %.3e  (2^%.1f)\n", mx, log2(mx));
    return 0;
}
