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
