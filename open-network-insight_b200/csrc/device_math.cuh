// Scalar fp64 math of the E-step, written for the B200 FP64 pipe.
//
// Why not libdevice's log/exp/division: profiled on B200 (profiles/), the
// E-step's inner loop was 1170 SASS instructions per word of which only 37 %
// were FP64 math - the rest was 64-bit constant materialisation (UMOV pairs),
// special-case branches (denormals, inf/nan, division slow paths) and moves.
// Every argument here is a finite, normal, positive quantity with a known
// range, so the routines below drop the special cases, keep their coefficients
// in __constant__ memory (a DFMA can read c[bank][offset] directly) and are
// evaluated in Estrin form: the Gauss-Seidel order of lda_inference
// (oni-ml/oni-lda-c/lda-inference.c:50-75) makes one word's update a serial
// dependency chain, so depth matters as much as instruction count.
//
// Accuracy (measured on the host against libm/mpmath by tests/test_device_math.py,
// which compiles this header for the CPU): rcp <= 1 ulp, log and exp <= 2 ulp,
// digamma within 4e-15 (of max(|psi|, 1)) of the reference's series.
//
// digamma follows the reference's own formula (utils.c:55-64):
//     y = x + 6,  p = 1/y^2
//     psi(x) = (((c4 p - c3) p + c2) p - c1) p + log(y) - 0.5/y - sum_{j=0..5} 1/(x + j)
// with its truncated-decimal coefficients kept verbatim.  The six recurrence
// terms are paired, 1/(x+j) + 1/(x+5-j) = (2x+5)/(u + j(5-j)) with u = x(x+5):
//     sum_j 1/(x+j) = (2x+5) (3u^2 + 20u + 24) / (u (u+4) (u+6))
// so that one reciprocal (of y * u(u+4)(u+6)) serves both 1/y and the sum.
#pragma once

#include <math.h>
#include <stdint.h>
#include <string.h>

#ifdef __CUDACC__
#include <cuda_runtime.h>
#define ONILDA_HD __host__ __device__ __forceinline__
#else
#define ONILDA_HD inline
#endif

namespace onilda {

#if defined(__CUDA_ARCH__)
#define ONILDA_CONST __constant__
#else
#define ONILDA_CONST static const
#endif

// log(m) = 2s + 2 s z P(z), s = (m-1)/(m+1), z = s^2, m in [sqrt(1/2), sqrt(2)): near-minimax, 4.5e-18
ONILDA_CONST double kLogP[7] = {0.3333333333333335, 0.19999999999949722, 0.14285714313001635, 0.11111105565129939,
                                0.09091444727837474, 0.07665855309105386, 0.07308292712444311};
// exp(r), |r| <= ln2/2: degree 11, 4.2e-18
ONILDA_CONST double kExpP[12] = {1.0, 1.0, 0.5000000000000019, 0.1666666666666668, 0.04166666666648795,
                                 0.008333333333319589, 0.0013888888952352863, 0.00019841269890076403,
                                 2.4801485441561313e-05, 2.755724088722987e-06, 2.763265472252779e-07,
                                 2.5110049204818658e-08};
// ln2 split: HI has 21 trailing zero bits, so k*HI is exact for |k| < 2^20
ONILDA_CONST double kLn2[3] = {0.6931471803691238, 1.9082149292705877e-10, 1.4426950408889634};
// the reference's digamma series coefficients (utils.c:60-61), signs applied
ONILDA_CONST double kPsiSeries[4] = {-0.083333333333333, 0.008333333333333, -0.003968253986254, 0.004166666666667};

ONILDA_HD double hilo_to_double(int hi, int lo)
{
#if defined(__CUDA_ARCH__)
    return __hiloint2double(hi, lo);
#else
    uint64_t b = ((uint64_t)(uint32_t)hi << 32) | (uint32_t)lo;
    double d;
    memcpy(&d, &b, sizeof d);
    return d;
#endif
}

ONILDA_HD int double_hi(double x)
{
#if defined(__CUDA_ARCH__)
    return __double2hiint(x);
#else
    uint64_t b;
    memcpy(&b, &x, sizeof b);
    return (int)(b >> 32);
#endif
}

ONILDA_HD int double_lo(double x)
{
#if defined(__CUDA_ARCH__)
    return __double2loint(x);
#else
    uint64_t b;
    memcpy(&b, &x, sizeof b);
    return (int)(uint32_t)b;
#endif
}

// 1/x for finite normal x.  The hardware seed (MUFU.RCP64H) is good to 2^-20 (measured on B200,
// tools/microbench.cu), so one cubic step reaches 2^-60; no special cases.
ONILDA_HD double fast_rcp(double x)
{
    double r;
#if defined(__CUDA_ARCH__)
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
#else
    r = hilo_to_double(double_hi(1.0 / x) & ~0x1, 0);     // host stand-in for the seed: 19 good bits
#endif
    double e = fma(-x, r, 1.0);
    e = fma(e, e, e);
    return fma(r, e, r);
}

// log(x) for finite normal x > 0.
ONILDA_HD double fast_log(double x)
{
    int hi = double_hi(x);
    const int lo = double_lo(x);
    int e = (hi >> 20) - 1023;
    hi = (hi & 0x000FFFFF) | 0x3FF00000;          // mantissa in [1, 2)
    if (hi >= 0x3FF6A09F) { hi -= 0x00100000; e += 1; }   // -> [sqrt(1/2), sqrt(2))
    const double m = hilo_to_double(hi, lo);
    const double f = m - 1.0;
    const double s = f * fast_rcp(m + 1.0);
    const double z = s * s;
    // Estrin on P(z), degree 6
    const double z2 = z * z;
    const double p01 = fma(kLogP[1], z, kLogP[0]);
    const double p23 = fma(kLogP[3], z, kLogP[2]);
    const double p45 = fma(kLogP[5], z, kLogP[4]);
    const double z4 = z2 * z2;
    const double p0123 = fma(p23, z2, p01);
    const double p456 = fma(kLogP[6], z2, p45);
    const double p = fma(p456, z4, p0123);
    const double ed = (double)e;
    const double s2 = s + s;
    // e*ln2 + 2s + 2s*z*P
    double r = fma(s2 * z, p, fma(ed, kLn2[1], s2));
    return fma(ed, kLn2[0], r);
}

// exp(x) for x <= ~700.  Very negative x gives a tiny positive value (about 2^-1021), never NaN:
// the binary exponent is clamped as an integer, which keeps the clamp off the FP64 dependency chain.
ONILDA_HD double fast_exp(double x)
{
    const double magic = 6755399441055744.0;            // 1.5 * 2^52: round-to-nearest-integer trick
    const double t = fma(x, kLn2[2], magic);
    int k = double_lo(t);
    const double kd = t - magic;
    double r = fma(kd, -kLn2[0], x);
    r = fma(kd, -kLn2[1], r);
    // Estrin, degree 11
    const double r2 = r * r;
    const double q01 = fma(kExpP[1], r, kExpP[0]);
    const double q23 = fma(kExpP[3], r, kExpP[2]);
    const double q45 = fma(kExpP[5], r, kExpP[4]);
    const double q67 = fma(kExpP[7], r, kExpP[6]);
    const double q89 = fma(kExpP[9], r, kExpP[8]);
    const double qab = fma(kExpP[11], r, kExpP[10]);
    const double r4 = r2 * r2;
    const double q0123 = fma(q23, r2, q01);
    const double q4567 = fma(q67, r2, q45);
    const double q89ab = fma(qab, r2, q89);
    const double r8 = r4 * r4;
    double q = fma(q89ab, r8, fma(q4567, r4, q0123));
    // x below about -1e15 would make r meaningless; such x only ever mean "zero"
    q = (x > -1.0e9) ? q : 1.0;
    k = k > -1021 ? k : -1021;
    // scale by 2^k through the exponent field (q in [0.7, 1.42], k in [-1021, 1011])
    return hilo_to_double(double_hi(q) + (k << 20), double_lo(q));
}

struct DigammaParts {
    double rational;   // psi(x) - log(x + 6)
    double y;          // x + 6
    double inv_y;      // 1 / (x + 6)
};

// psi(x) - log(x+6): everything except the logarithm.  x > 0.
// Two independent reciprocals (of y and of the recurrence denominator) instead of one shared
// one: four more instructions, but the asymptotic series then runs beside the denominator chain
// and only "multiply, subtract" follow the second reciprocal (the update of one word is a
// serial chain; see estep.cu).
ONILDA_HD DigammaParts digamma_rational(double x)
{
    const double y = x + 6.0;
    // The reference forms its recurrence terms from the ROUNDED shifted argument, 1/(y-1) ... 1/(y-6)
    // (utils.c:62): for tiny x that is 1/fl(fl(x+6)-6), not 1/x.  xr reproduces it.
    const double xr = y - 6.0;
    const double u = xr * (xr + 5.0);
    const double den = (u * (u + 4.0)) * (u + 6.0);          // xr(xr+1)...(xr+5)
    const double num = fma(fma(3.0, u, 20.0), u, 24.0);
    const double tn = fma(2.0, xr, 5.0) * num;
    const double inv_y = fast_rcp(y);
    const double z = inv_y * inv_y;
    const double z2 = z * z;
    const double s01 = fma(kPsiSeries[1], z, kPsiSeries[0]);
    const double s23 = fma(kPsiSeries[3], z, kPsiSeries[2]);
    const double head = fma(-0.5, inv_y, fma(s23, z2, s01) * z);
    const double inv_den = fast_rcp(den);
    DigammaParts out;
    out.rational = fma(-tn, inv_den, head);                  // head - sum_{j=0..5} 1/(xr+j)
    out.y = y;
    out.inv_y = inv_y;
    return out;
}

ONILDA_HD double digamma(double x)
{
    const DigammaParts d = digamma_rational(x);
    return d.rational + fast_log(d.y);
}

}  // namespace onilda
