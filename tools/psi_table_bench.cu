// Decision data for the round-2 plan in DESIGN.md 4.1: psi_ref(x) = h(y) - 1/(y - 6), y = x + 6, with h from a
// shared-memory table of piecewise polynomials in u = y - 5, against the present digamma_rational + table logarithm.
// Prints the table's worst error (against a long-double evaluation of the reference's own formula, utils.c:56-64)
// and the throughput of both forms on one B200 at the bulk kernel's shape (4 warps per sub-partition, 5 chains per lane).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -Iopen-network-insight_b200/csrc -Iinclude \
//        -o /tmp/psi_bench tools/psi_table_bench.cu && /tmp/psi_bench
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include <cuda_runtime.h>

#include "device_math.cuh"

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

constexpr int kSub = 5;            // 2^5 intervals per octave of u
constexpr int kOct = 20;           // u in [1, 2^20)
#ifndef DEG
#define DEG 5
#endif
constexpr int kRow = DEG + 1 <= 6 ? 6 : 8;     // doubles per interval

static long double h_ref(long double y)
{
    long double p = 1.0L / (y * y);
    p = (((0.004166666666667L * p - 0.003968253986254L) * p + 0.008333333333333L) * p - 0.083333333333333L) * p;
    return p + logl(y) - 0.5L / y - 1.0L / (y - 1) - 1.0L / (y - 2) - 1.0L / (y - 3) - 1.0L / (y - 4) - 1.0L / (y - 5);
}

static void fit_interval(long double u0, long double w, double* row)
{
    const int n = DEG + 1;
    long double s[8], f[8], a[8][9];
    for (int i = 0; i < n; i++) {
        s[i] = cosl(3.14159265358979323846264338327950288L * (2 * i + 1) / (2 * n));       // Chebyshev nodes on [-1, 1]
        f[i] = h_ref(u0 + (s[i] + 1) * w / 2 + 5);
    }
    for (int i = 0; i < n; i++) {
        long double v = 1;
        for (int j = 0; j < n; j++) { a[i][j] = v; v *= s[i]; }
        a[i][n] = f[i];
    }
    for (int c = 0; c < n; c++) {                                                          // Gauss-Jordan with pivoting
        int piv = c;
        for (int r = c + 1; r < n; r++) if (fabsl(a[r][c]) > fabsl(a[piv][c])) piv = r;
        for (int j = 0; j <= n; j++) { long double t = a[c][j]; a[c][j] = a[piv][j]; a[piv][j] = t; }
        for (int r = 0; r < n; r++) {
            if (r == c) continue;
            const long double m = a[r][c] / a[c][c];
            for (int j = c; j <= n; j++) a[r][j] -= m * a[c][j];
        }
    }
    long double cs[8], b[8] = {0};
    for (int j = 0; j < n; j++) cs[j] = a[j][n] / a[j][j];                                  // p(s) = sum cs_j s^j
    // s = 2t/w - 1  ->  coefficients in t
    for (int j = 0; j < n; j++) {
        long double binom = 1;                                                              // C(j, i)
        for (int i = 0; i <= j; i++) {
            const long double term = cs[j] * binom * powl(2 / w, i) * (((j - i) & 1) ? -1 : 1);
            b[i] += term;
            binom = binom * (j - i) / (i + 1);
        }
    }
    for (int j = 0; j < kRow; j++) row[j] = j < n ? (double)b[j] : 0.0;
}

__device__ __forceinline__ double psi_table(double x, const double* __restrict__ tab, double* rcp_out)
{
    const double y = x + 6.0;
    const double xr = y - 6.0;
    const double u = y - 5.0;
    const long long bits = __double_as_longlong(u);
    const int idx = (int)((bits >> (52 - kSub)) - (1023ll << kSub));
    const double u0 = __longlong_as_double(bits & ~((1ll << (52 - kSub)) - 1));
    const double t = u - u0;
    const double* row = tab + idx * kRow;
    double h;
    if (DEG == 5) {
        const double2 c01 = *reinterpret_cast<const double2*>(row), c23 = *reinterpret_cast<const double2*>(row + 2),
                      c45 = *reinterpret_cast<const double2*>(row + 4);
        const double t2 = t * t;
        h = fma(fma(fma(c45.y, t, c45.x), t2, fma(c23.y, t, c23.x)), t2, fma(c01.y, t, c01.x));
    } else {
        const double2 c01 = *reinterpret_cast<const double2*>(row), c23 = *reinterpret_cast<const double2*>(row + 2),
                      c45 = *reinterpret_cast<const double2*>(row + 4), c67 = *reinterpret_cast<const double2*>(row + 6);
        const double t2 = t * t, t4 = t2 * t2;
        h = fma(fma(fma(c67.y, t, c67.x), t2, fma(c45.y, t, c45.x)), t4, fma(fma(c23.y, t, c23.x), t2, fma(c01.y, t, c01.x)));
    }
    const double r = onilda::fast_rcp(xr);
    *rcp_out = r;
    return h - r;
}

template <int MODE>     // 0: present form (rational + table log + exp), 1: table psi + exp
__global__ void __launch_bounds__(128) bench_kernel(const double* __restrict__ x, double* __restrict__ out, int n, int iters,
                                                    const double* __restrict__ gtab, int tab_doubles, double shift)
{
    extern __shared__ __align__(16) double sh[];
    onilda::LogTabEntry* logtab = reinterpret_cast<onilda::LogTabEntry*>(sh);
    double* tab = sh + 256;
    for (int i = threadIdx.x; i < 128; i += blockDim.x) logtab[i] = onilda::kLogTab[i];
    if (MODE == 1)
        for (int i = threadIdx.x; i < tab_doubles; i += blockDim.x) tab[i] = gtab[i];
    __syncthreads();
    const int gid = blockIdx.x * blockDim.x + threadIdx.x;
    double g[5], acc = 0.0;
    for (int t = 0; t < 5; t++) g[t] = x[(gid * 5 + t) % n];
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int t = 0; t < 5; t++) {
            double psi, E;
            if (MODE == 0) {
                const onilda::DigammaParts dp = onilda::digamma_rational(g[t]);
                psi = dp.rational + onilda::fast_log_tab(dp.y, logtab);
                E = dp.y * onilda::fast_exp(dp.rational - shift);
            } else {
                double r;
                psi = psi_table(g[t], tab, &r);
                E = onilda::fast_exp(psi - shift);
            }
            acc += psi;
            g[t] = fma(E, 1e-9, g[t]);            // keeps the chain alive, barely moves the argument
        }
    }
    out[gid] = acc + g[0];
}

template <int MODE>
__global__ void eval_kernel(const double* __restrict__ x, double* __restrict__ out, int n, const double* __restrict__ gtab)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (MODE == 0) out[i] = onilda::digamma(x[i]);
    else { double r; out[i] = psi_table(x[i], gtab, &r); }
}

int main()
{
    const int rows = kOct << kSub;
    std::vector<double> tab((size_t)rows * kRow);
    for (int i = 0; i < rows; i++) {
        const int e = i >> kSub, k = i & ((1 << kSub) - 1);
        const long double w = ldexpl(1.0L, e - kSub), u0 = ldexpl(1.0L, e) + k * w;
        fit_interval(u0, w, &tab[(size_t)i * kRow]);
    }
    printf("table: degree %d, %d intervals, %zu bytes\n", DEG, rows, tab.size() * sizeof(double));
    // accuracy on a grid of arguments, against the long-double formula
    const int n = 1 << 20;
    std::vector<double> hx(n);
    for (int i = 0; i < n; i++) {
        const double f = (double)rand() / RAND_MAX;
        hx[i] = i % 3 == 0 ? 0.001 + 2.0 * f : exp(log(1e-3) + f * log(5e5 / 1e-3));
    }
    double *dx, *dout, *dtab;
    CK(cudaMalloc(&dx, n * sizeof(double))); CK(cudaMalloc(&dout, n * sizeof(double))); CK(cudaMalloc(&dtab, tab.size() * sizeof(double)));
    CK(cudaMemcpy(dx, hx.data(), n * sizeof(double), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dtab, tab.data(), tab.size() * sizeof(double), cudaMemcpyHostToDevice));
    std::vector<double> o0(n), o1(n);
    eval_kernel<0><<<n / 256, 256>>>(dx, dout, n, dtab); CK(cudaMemcpy(o0.data(), dout, n * sizeof(double), cudaMemcpyDeviceToHost));
    eval_kernel<1><<<n / 256, 256>>>(dx, dout, n, dtab); CK(cudaMemcpy(o1.data(), dout, n * sizeof(double), cudaMemcpyDeviceToHost));
    double e0 = 0, e1 = 0;
    for (int i = 0; i < n; i++) {
        const long double y = (long double)(hx[i] + 6.0);
        const long double ref = h_ref(y) - 1.0L / (long double)((hx[i] + 6.0) - 6.0);
        const double scale = fmax(1.0, fabs((double)ref));
        e0 = fmax(e0, fabs((double)(o0[i] - ref)) / scale);
        e1 = fmax(e1, fabs((double)(o1[i] - ref)) / scale);
    }
    printf("max |psi - reference formula| / max(1, |psi|): present form %.3e, table form %.3e\n", e0, e1);
    // throughput at the bulk kernel's shape
    cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, 0));
    const int blocks = prop.multiProcessorCount * 4, threads = 128, iters = 2000;
    const size_t smem = 2048 + tab.size() * sizeof(double);
    CK(cudaFuncSetAttribute(bench_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    CK(cudaFuncSetAttribute(bench_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    double* dbig; CK(cudaMalloc(&dbig, (size_t)blocks * threads * sizeof(double)));
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    for (int mode = 0; mode < 2; mode++) {
        for (int rep = 0; rep < 2; rep++) {
            cudaEventRecord(a);
            if (mode == 0) bench_kernel<0><<<blocks, threads, smem>>>(dx, dbig, n, iters, dtab, (int)tab.size(), 1.5);
            else bench_kernel<1><<<blocks, threads, smem>>>(dx, dbig, n, iters, dtab, (int)tab.size(), 1.5);
            cudaEventRecord(b);
            CK(cudaDeviceSynchronize());
            float ms; cudaEventElapsedTime(&ms, a, b);
            if (rep == 1)
                printf("%s: %.3f ms for %d x %d lanes x %d iterations x 5 -> %.1f G psi+exp per second\n",
                       mode == 0 ? "present form (rational + table log + exp)" : "table form (h table + reciprocal + exp) ", ms, blocks, threads,
                       iters, (double)blocks * threads * iters * 5 / ms / 1e6);
        }
    }
    return 0;
}
