// Device-side scalar math for the E-step (fp64).
//
// digamma follows the reference's own series (oni-ml/oni-lda-c/utils.c:55-64):
//     y = x + 6,  p = 1/y^2
//     psi(x) = (((c4 p - c3) p + c2) p - c1) p + log(y) - 0.5/y - sum_{j=0..5} 1/(x + j)
// with the reference's truncated-decimal coefficients (they are part of the
// result the oracle produces, so they are kept verbatim).  The reference spends
// eight divisions on it; here the six recurrence terms are folded into one
// rational P(x)/Q(x), Q = x(x+1)...(x+5), P = Q', and 1/y and 1/Q come out of a
// single reciprocal of y*Q.  All coefficients of P and Q are positive and x > 0,
// so Horner evaluation is well conditioned (a few ulp).
#pragma once

#include <cuda_runtime.h>

namespace onilda {

struct DigammaParts {
    double rational;   // psi(x) - log(x + 6)
    double y;          // x + 6
};

// psi(x) - log(x+6): everything except the logarithm (no transcendental).
__device__ __forceinline__ DigammaParts digamma_rational(double x)
{
    const double y = x + 6.0;
    // Q(x)/x = x^5 + 15 x^4 + 85 x^3 + 225 x^2 + 274 x + 120
    double q = x + 15.0;
    q = fma(q, x, 85.0);
    q = fma(q, x, 225.0);
    q = fma(q, x, 274.0);
    q = fma(q, x, 120.0);
    q *= x;
    // P(x) = Q'(x) = 6 x^5 + 75 x^4 + 340 x^3 + 675 x^2 + 548 x + 120
    double p = fma(6.0, x, 75.0);
    p = fma(p, x, 340.0);
    p = fma(p, x, 675.0);
    p = fma(p, x, 548.0);
    p = fma(p, x, 120.0);
    const double inv = 1.0 / (y * q);
    const double inv_y = q * inv;
    const double inv_q = y * inv;
    const double z = inv_y * inv_y;
    double series = fma(0.004166666666667, z, -0.003968253986254);
    series = fma(series, z, 0.008333333333333);
    series = fma(series, z, -0.083333333333333);
    series *= z;
    DigammaParts out;
    out.rational = series - 0.5 * inv_y - p * inv_q;
    out.y = y;
    return out;
}

__device__ __forceinline__ double digamma(double x)
{
    const DigammaParts d = digamma_rational(x);
    return d.rational + log(d.y);
}

}  // namespace onilda
