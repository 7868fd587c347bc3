// Host build of open-network-insight_b200/csrc/device_math.cuh (the same source the kernels
// compile; the MUFU.RCP64H seed is replaced by a 20-bit truncation) for tests/test_device_math.py.
#include "device_math.cuh"
#include <vector>
extern "C" {
void onilda_build_psi_table(double* rows, int mant_bits, int octaves);
void onilda_build_exp_table(double* tab, int n);
}
static std::vector<onilda::PsiRow> g_psi;
static double g_exp[onilda::kExpTabSize];
static onilda::MathTablesPtr g_tb;
static void tables()
{
    if (!g_psi.empty()) return;
    g_psi.resize(onilda::kPsiRows);
    onilda_build_psi_table(&g_psi[0].a[0], onilda::kPsiMantBits, onilda::kPsiOctaves);
    onilda_build_exp_table(g_exp, onilda::kExpTabSize);
    g_tb.psi = g_psi.data();
    g_tb.exp2 = g_exp;
    g_tb.log = onilda::kLogTab;
}

extern "C" {
void dm_rcp(const double* x, double* y, long n) { for (long i = 0; i < n; i++) y[i] = onilda::fast_rcp(x[i]); }
void dm_log(const double* x, double* y, long n) { for (long i = 0; i < n; i++) y[i] = onilda::fast_log(x[i]); }
void dm_exp(const double* x, double* y, long n) { for (long i = 0; i < n; i++) y[i] = onilda::fast_exp(x[i]); }
void dm_log_tab(const double* x, double* y, long n) { for (long i = 0; i < n; i++) y[i] = onilda::fast_log_tab(x[i], onilda::kLogTab); }
void dm_digamma(const double* x, double* y, long n) { for (long i = 0; i < n; i++) y[i] = onilda::digamma(x[i]); }
void dm_exp_tab(const double* x, double* y, long n) { tables(); for (long i = 0; i < n; i++) y[i] = onilda::fast_exp_tab<false>(x[i], g_tb); }
void dm_exp_tab_shallow(const double* x, double* y, long n) { tables(); for (long i = 0; i < n; i++) y[i] = onilda::fast_exp_tab<true>(x[i], g_tb); }
void dm_psi_tab(const double* x, double* y, long n) { tables(); for (long i = 0; i < n; i++) y[i] = onilda::psi_exp_tab<false>(x[i], g_tb).psi; }
void dm_psi_tab_shallow(const double* x, double* y, long n) { tables(); for (long i = 0; i < n; i++) y[i] = onilda::psi_exp_tab<true>(x[i], g_tb).psi; }
void dm_epsi_tab(const double* x, double* y, long n) { tables(); for (long i = 0; i < n; i++) y[i] = onilda::psi_exp_tab<false>(x[i], g_tb).e; }
}
