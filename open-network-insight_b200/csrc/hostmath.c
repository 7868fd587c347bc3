/*
 * Host-side scalar math for the one step north_star leaves on the CPU: the
 * Newton iteration for alpha (reference: oni-ml/oni-lda-c/lda-alpha.c:27-68),
 * with the reference's own digamma / trigamma series (utils.c:32-64) - their
 * truncated-decimal coefficients are part of the result, so they are kept.
 */
#include <math.h>

#include "onilda_internal.h"

/* psi(x): shift by 6, asymptotic series in 1/y^2, then the six recurrence terms (utils.c:55-64) */
double onilda_host_digamma(double x)
{
    const double y = x + 6;
    double q = 1 / (y * y);
    q = (((0.004166666666667 * q - 0.003968253986254) * q + 0.008333333333333) * q - 0.083333333333333) * q;
    return q + log(y) - 0.5 / y - 1 / (y - 1) - 1 / (y - 2) - 1 / (y - 3) - 1 / (y - 4) - 1 / (y - 5) - 1 / (y - 6);
}

/* psi'(x) (utils.c:32-47) */
double onilda_host_trigamma(double x)
{
    double y = x + 6;
    double q = 1 / (y * y);
    q = (((((0.075757575757576 * q - 0.033333333333333) * q + 0.0238095238095238) * q - 0.033333333333333) * q
          + 0.166666666666667) * q + 1) / y + 0.5 * q;
    for (int j = 0; j < 6; j++) {
        y = y - 1;
        q = 1 / (y * y) + q;
    }
    return q;
}

/* opt_alpha (lda-alpha.c:42-68): Newton on log(alpha) from alpha = 100; D and K are ints there. */
double onilda_opt_alpha(double ss, int D, int K)
{
    double guess = 100, log_a = log(guess), a, df, d2f;
    int it = 0;
    do {
        it++;
        a = exp(log_a);
        if (isnan(a)) {
            guess = guess * 10;
            a = guess;
            log_a = log(a);
        }
        df = D * (K * onilda_host_digamma(K * a) - K * onilda_host_digamma(a)) + ss;
        d2f = D * (K * K * onilda_host_trigamma(K * a) - K * onilda_host_trigamma(a));
        log_a = log_a - df / (d2f * a + df);
    } while (fabs(df) > 1e-5 && it < 1000);
    return exp(log_a);
}

/*
 * Tables of the E-step's per-(word, topic) math (device_math.cuh: psi_exp_tab).
 *
 * psi rows: the reference's digamma without its pole, R(x) = digamma(x) + 1/x (utils.c:55-64 with
 * the 1/(y-6) term left out), as a function of t = x + 1 on cells of 1/64 octave, t in [1, 2^20).
 * Each row holds the degree-5 polynomial in u = t - centre that interpolates R at the six Chebyshev
 * nodes of its cell, worked out in long double.  The fit is to the REFERENCE's series (its
 * truncated-decimal coefficients are part of the result), not to the true digamma.
 */
static long double ref_digamma_no_pole(long double x)
{
    const long double y = x + 6.0L;
    long double q = 1.0L / (y * y);
    q = (((0.004166666666667L * q - 0.003968253986254L) * q + 0.008333333333333L) * q - 0.083333333333333L) * q;
    return q + logl(y) - 0.5L / y - 1.0L / (x + 5.0L) - 1.0L / (x + 4.0L) - 1.0L / (x + 3.0L) - 1.0L / (x + 2.0L)
           - 1.0L / (x + 1.0L);
}

void onilda_build_psi_table(double* rows, int mant_bits, int octaves)
{
    const long double pi = 3.14159265358979323846264338327950288L;
    const int per_oct = 1 << mant_bits;
    for (int e = 0; e < octaves; e++) {
        const long double h = ldexpl(1.0L, e - mant_bits - 1);           /* half a cell */
        for (int i = 0; i < per_oct; i++) {
            const long double centre = ldexpl(1.0L, e) + (2 * i + 1) * h;
            long double f[6], c[6];
            for (int m = 0; m < 6; m++) f[m] = ref_digamma_no_pole(centre + h * cosl((2 * m + 1) * pi / 12.0L) - 1.0L);
            for (int k = 0; k < 6; k++) {                                 /* Chebyshev coefficients */
                long double s = 0;
                for (int m = 0; m < 6; m++) s += f[m] * cosl(k * (2 * m + 1) * pi / 12.0L);
                c[k] = s * (k == 0 ? 1.0L : 2.0L) / 6.0L;
            }
            /* T0..T5 to powers of z = u/h */
            const long double b[6] = {c[0] - c[2] + c[4], c[1] - 3 * c[3] + 5 * c[5], 2 * c[2] - 8 * c[4],
                                      4 * c[3] - 20 * c[5], 8 * c[4], 16 * c[5]};
            double* row = rows + 6 * ((size_t)e * per_oct + i);
            long double scale = 1.0L;
            for (int j = 0; j < 6; j++) {
                row[j] = (double)(b[j] * scale);
                scale /= h;
            }
        }
    }
}

/* 2^(i/n), i = 0..n-1 */
void onilda_build_exp_table(double* tab, int n)
{
    for (int i = 0; i < n; i++) tab[i] = (double)exp2l((long double)i / n);
}
