// Host build of open-network-insight_b200/csrc/device_math.cuh (the same source the kernels
// compile; the MUFU.RCP64H seed is replaced by a 20-bit truncation) for tests/test_device_math.py.
#include "device_math.cuh"

extern "C" {
void dm_rcp(const double* x, double* y, long n) { for (long i = 0; i < n; i++) y[i] = onilda::fast_rcp(x[i]); }
void dm_log(const double* x, double* y, long n) { for (long i = 0; i < n; i++) y[i] = onilda::fast_log(x[i]); }
void dm_exp(const double* x, double* y, long n) { for (long i = 0; i < n; i++) y[i] = onilda::fast_exp(x[i]); }
void dm_log_tab(const double* x, double* y, long n) { for (long i = 0; i < n; i++) y[i] = onilda::fast_log_tab(x[i], onilda::kLogTab); }
void dm_digamma(const double* x, double* y, long n) { for (long i = 0; i < n; i++) y[i] = onilda::digamma(x[i]); }
}
