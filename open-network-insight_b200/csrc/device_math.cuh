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

// log by table: x = 2^e * m, m in [1,2); the top 7 mantissa bits pick c_i = 1 + (i + 1/2)/128 with
// 1/c_i and log(c_i) tabulated (log of the ROUNDED reciprocal, so the pair is consistent), and
// r = m/c_i - 1, |r| < 2^-8, goes through six terms of log1p.  10 FP64 instructions instead of the
// 19 of fast_log (no reciprocal); the 2 KB table is copied to shared memory by each block because
// lanes index it divergently (constant memory would serialise).  <= 2 ulp like fast_log.
struct alignas(16) LogTabEntry {
    double inv_c;
    double log_c;
};
#if defined(__CUDA_ARCH__)
#define ONILDA_TABLE __device__ const
#else
#define ONILDA_TABLE static const
#endif
ONILDA_TABLE LogTabEntry kLogTab[128] = {
    {0.9961089494163424, 0.003898640415657309},
    {0.9884169884169884, 0.01165061721997525},
    {0.9808429118773946, 0.019342962843130987},
    {0.973384030418251, 0.026976587698202083},
    {0.9660377358490566, 0.03455238150665973},
    {0.9588014981273408, 0.042071213920687044},
    {0.9516728624535316, 0.049533935122276676},
    {0.9446494464944649, 0.05694137640013845},
    {0.9377289377289377, 0.06429435070539725},
    {0.9309090909090909, 0.07159365318700882},
    {0.924187725631769, 0.078840061707776},
    {0.9175627240143369, 0.08603433734180316},
    {0.9110320284697508, 0.09317722485418334},
    {0.9045936395759717, 0.10026945316367517},
    {0.8982456140350877, 0.10731173578908804},
    {0.89198606271777, 0.11430477128005863},
    {0.8858131487889274, 0.12124924363286965},
    {0.8797250859106529, 0.12814582269193006},
    {0.8737201365187713, 0.13499516453750482},
    {0.8677966101694915, 0.1417979118602574},
    {0.8619528619528619, 0.1485546943231372},
    {0.8561872909698997, 0.15526612891112396},
    {0.8504983388704319, 0.16193282026931324},
    {0.8448844884488449, 0.16855536102980664},
    {0.839344262295082, 0.17513433212784915},
    {0.8338762214983714, 0.18167030310763463},
    {0.8284789644012945, 0.18816383241818294},
    {0.8231511254019293, 0.19461546769967167},
    {0.8178913738019169, 0.2010257460605908},
    {0.8126984126984127, 0.2073951943460706},
    {0.807570977917981, 0.21372432939771818},
    {0.8025078369905956, 0.22001365830528213},
    {0.7975077881619937, 0.2262636786504534},
    {0.7925696594427245, 0.232474878743094},
    {0.7876923076923077, 0.238647737850175},
    {0.7828746177370031, 0.24478272641769092},
    {0.7781155015197568, 0.25088030628580943},
    {0.7734138972809668, 0.2569409308975004},
    {0.7687687687687688, 0.26296504550088134},
    {0.764179104477612, 0.26895308734550394},
    {0.7596439169139466, 0.2749054858727992},
    {0.7551622418879056, 0.2808226629008878},
    {0.750733137829912, 0.2867050328039543},
    {0.7463556851311953, 0.29255300268637746},
    {0.7420289855072464, 0.2983669725517973},
    {0.7377521613832853, 0.3041473354672968},
    {0.7335243553008596, 0.3098944777228647},
    {0.7293447293447294, 0.3156087789863033},
    {0.7252124645892352, 0.32129061245373425},
    {0.7211267605633803, 0.3269403449958533},
    {0.7170868347338936, 0.3325583373000766},
    {0.713091922005571, 0.3381449440087164},
    {0.7091412742382271, 0.34370051385331846},
    {0.7052341597796143, 0.3492253897852883},
    {0.7013698630136986, 0.354719909102929},
    {0.6975476839237057, 0.3601844035750078},
    {0.6937669376693767, 0.3656191995609647},
    {0.6900269541778976, 0.37102461812787263},
    {0.6863270777479893, 0.376400975164253},
    {0.6826666666666666, 0.3817485814908484},
    {0.6790450928381963, 0.3870677429684483},
    {0.6754617414248021, 0.3923587606028639},
    {0.6719160104986877, 0.3976219306471385},
    {0.6684073107049608, 0.4028575447010835},
    {0.6649350649350649, 0.4080658898082217},
    {0.661498708010336, 0.41324724855021927},
    {0.6580976863753213, 0.41840189913888387},
    {0.6547314578005116, 0.4235301155058032},
    {0.6513994910941476, 0.42863216738969867},
    {0.6481012658227848, 0.4337083204215594},
    {0.6448362720403022, 0.43875883620762796},
    {0.6416040100250626, 0.44378397241030104},
    {0.6384039900249376, 0.4487839828270067},
    {0.6352357320099256, 0.4537591174671205},
    {0.6320987654320988, 0.4587096226269767},
    {0.628992628992629, 0.46363574096303256},
    {0.6259168704156479, 0.46853771156323926},
    {0.6228710462287105, 0.4734157700166721},
    {0.6198547215496368, 0.47827014848147026},
    {0.6168674698795181, 0.48310107575113576},
    {0.6139088729016786, 0.48790877731923904},
    {0.6109785202863962, 0.4926934754425752},
    {0.6080760095011877, 0.4974553892028189},
    {0.6052009456264775, 0.5021947345667155},
    {0.6023529411764705, 0.5069117244448544},
    {0.5995316159250585, 0.5116065687490621},
    {0.5967365967365967, 0.5162794744484545},
    {0.5939675174013921, 0.5209306456241853},
    {0.5912240184757506, 0.5255602835229274},
    {0.5885057471264368, 0.5301685866091216},
    {0.585812356979405, 0.5347557506160276},
    {0.5831435079726651, 0.5393219685956089},
    {0.5804988662131519, 0.5438674309672835},
    {0.5778781038374717, 0.5483923255655733},
    {0.5752808988764045, 0.5528968376866776},
    {0.5727069351230425, 0.5573811501340064},
    {0.5701559020044543, 0.5618454432626918},
    {0.5676274944567627, 0.5662898950231159},
    {0.565121412803532, 0.5707146810034716},
    {0.5626373626373626, 0.575119974471388},
    {0.5601750547045952, 0.5795059464146423},
    {0.5577342047930284, 0.5838727655809826},
    {0.5553145336225597, 0.588220598517086},
    {0.5529157667386609, 0.5925496096066716},
    {0.5505376344086022, 0.5968599611077938},
    {0.5481798715203426, 0.6011518131893347},
    {0.5458422174840085, 0.6054253239667169},
    {0.5435244161358811, 0.6096806495368553},
    {0.5412262156448203, 0.6139179440123704},
    {0.5389473684210526, 0.6181373595550788},
    {0.5366876310272537, 0.6223390464087787},
    {0.534446764091858, 0.6265231529313529},
    {0.5322245322245323, 0.6306898256261987},
    {0.5300207039337475, 0.6348392091730102},
    {0.5278350515463918, 0.6389714464579207},
    {0.5256673511293635, 0.6430866786030273},
    {0.523517382413088, 0.6471850449953095},
    {0.5213849287169042, 0.6512666833149582},
    {0.5192697768762677, 0.6553317295631277},
    {0.5171717171717172, 0.6593803180891278},
    {0.5150905432595574, 0.6634125816170662},
    {0.5130260521042084, 0.6674286512719563},
    {0.5109780439121756, 0.6714286566053024},
    {0.5089463220675944, 0.6754127256201768},
    {0.5069306930693069, 0.6793809847957973},
    {0.504930966469428, 0.6833335591116206},
    {0.5029469548133595, 0.6872705720709603},
    {0.5009784735812133, 0.691192145724142}
};

// Table access is a template parameter of the routines below: the kernels read the tables from shared
// memory through 32-bit shared-window addresses (estep.cu), the host tests through plain pointers.
struct MathTablesPtr {
    const struct PsiRow* psi;
    const double* exp2;
    const LogTabEntry* log;
    ONILDA_HD void log_entry(int i, double& inv_c, double& log_c) const { inv_c = log[i].inv_c; log_c = log[i].log_c; }
    ONILDA_HD double exp2_frac(int i) const { return exp2[i]; }
    ONILDA_HD void psi_row(int row, double (&a)[6]) const;
};

template <class Tables>
ONILDA_HD double fast_log_tab(double x, const Tables& tb)
{
    const int hi = double_hi(x);
    const int e = (hi >> 20) - 1023;
    double inv_c, log_c;
    tb.log_entry((hi >> 13) & 127, inv_c, log_c);
    const double m = hilo_to_double((hi & 0x000FFFFF) | 0x3FF00000, double_lo(x));
    const double r = fma(m, inv_c, -1.0);
    // log1p(r) = r - r^2/2 + r^3/3 - r^4/4 + r^5/5 - r^6/6
    const double r2 = r * r;
    const double a = fma(r, 0.3333333333333333, -0.5);
    const double b = fma(r, 0.2, -0.25);
    const double c = fma(r2, -0.16666666666666666, b);      // -1/4 + r/5 - r^2/6
    const double p = fma(r2, c, a);                           // -1/2 + r/3 - r^2/4 + r^3/5 - r^4/6
    const double ed = (double)e;
    const double lo = fma(ed, kLn2[1], fma(r2, p, r));
    return fma(ed, kLn2[0], log_c) + lo;
}

ONILDA_HD double fast_log_tab(double x, const LogTabEntry* tab)
{
    MathTablesPtr tb;
    tb.psi = nullptr;
    tb.exp2 = nullptr;
    tb.log = tab;
    return fast_log_tab(x, tb);
}

// 1/x for finite normal x.  The hardware seed (MUFU.RCP64H) is good to 2^-20 (measured on B200,
// tools/microbench.cu), so one cubic step reaches 2^-60; no special cases.
ONILDA_HD double rcp_seed(double x)
{
    double r;
#if defined(__CUDA_ARCH__)
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
#else
    r = hilo_to_double(double_hi(1.0 / x) & ~0x1, 0);     // host stand-in for the seed: 19 good bits
#endif
    return r;
}

ONILDA_HD double fast_rcp(double x)
{
    const double r = rcp_seed(x);
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

// ---------------------------------------------------------------------------------------------
// psi and exp by table: the per-(word, topic) forms of the E-step.
//
// The reference's digamma (utils.c:55-64) splits exactly as psi(x) = R(x) - 1/x, with
//     R(x) = series(1/(x+6)^2) + log(x+6) - 0.5/(x+6) - sum_{j=1..5} 1/(x+j)
// analytic for x > -1.  R is tabulated as a function of t = x + 1 >= 1: the exponent and the top
// kPsiMantBits mantissa bits of t pick a cell (64 per octave, kPsiOctaves octaves, t < 2^20), and a
// degree-5 polynomial in u = t - (cell centre) interpolates the REFERENCE's series (truncated
// coefficients included) at six Chebyshev nodes of the cell; the host builds the rows in long
// double (hostmath.c).  Interpolation error <= 8e-15 absolute on [1, 2), less above.  One
// reciprocal supplies the pole.  7 + 4 FP64 instructions against 37 for the rational form plus its
// logarithm.  48-byte rows: three 16-byte shared-memory loads whose 16-byte chunk index 3i + j
// spreads over all eight bank groups (a 64-byte row would put every lane in two of them).
//
// exp: x = (64 k + i) ln2/64 + r, |r| <= ln2/128; 2^(i/64) from a 64-entry table, degree-5 Taylor
// in r (3.5e-17), the scale 2^k through the exponent field.  8 FP64 instructions against 14.
constexpr int kPsiMantBits = 6;
constexpr int kPsiOctaves = 20;
constexpr int kPsiRows = kPsiOctaves << kPsiMantBits;     // 1280 rows, 61440 bytes
constexpr int kExpTabSize = 64;
struct alignas(16) PsiRow {
    double a[6];
};

// row of t (t >= 1); rows past the table (t >= 2^20) are the caller's business
ONILDA_HD int psi_row_of(double t) { return (double_hi(t) >> (20 - kPsiMantBits)) - (1023 << kPsiMantBits); }

// u = t - centre of t's cell (exact)
ONILDA_HD double psi_cell_offset(double t)
{
    const int keep = ~((1 << (20 - kPsiMantBits)) - 1);
    const int hc = (double_hi(t) & keep) | (1 << (19 - kPsiMantBits));
    return t - hilo_to_double(hc, 0);
}

ONILDA_HD void MathTablesPtr::psi_row(int row, double (&a)[6]) const
{
    for (int i = 0; i < 6; i++) a[i] = psi[row].a[i];
}

// R by Horner (five dependent DFMA: the throughput form) or Estrin (seven, depth three: the form
// for the one-document-per-warp kernel, which waits on its own dependency chain)
template <bool kShallow>
ONILDA_HD double psi_poly(const double (&a)[6], double u)
{
    if (kShallow) {
        const double u2 = u * u;
        const double p01 = fma(a[1], u, a[0]);
        const double p23 = fma(a[3], u, a[2]);
        const double p45 = fma(a[5], u, a[4]);
        const double u4 = u2 * u2;
        return fma(p45, u4, fma(p23, u2, p01));
    } else {
        double p = fma(a[5], u, a[4]);
        p = fma(p, u, a[3]);
        p = fma(p, u, a[2]);
        p = fma(p, u, a[1]);
        return fma(p, u, a[0]);
    }
}

// 64/ln2 and ln2/64.  The reduction r = x - kd*(ln2/64) is ONE fma with the rounded constant: its error,
// |kd| * 1.2e-18, is 3e-15 for |x| <= 30 (where exp(psi) matters) and 8e-14 at x = -700.
ONILDA_CONST double kExpTabC[2] = {92.33248261689366, 0.010830424696249145};
// 1/2, 1/6, 1/24, 1/120
ONILDA_CONST double kExpTabP[4] = {0.5, 0.16666666666666666, 0.041666666666666664, 0.008333333333333333};

// exp(x) for -1e9 < x <= ~700 with the table of 2^(i/64); x below about -708 gives a tiny positive value.
// (Callers guarantee x > -1e9: psi(gamma) with gamma >= 1e-8, see estep.cu.)
template <bool kShallow, class Tables>
ONILDA_HD double fast_exp_tab(double x, const Tables& tb)
{
    // the constants as literals: the instructions then take them from the compiler's own constant bank as operands (an
    // element of a __constant__ array costs a load instruction per use when registers are short)
    const double magic = 6755399441055744.0;
    const double c_hi = 92.33248261689366, c_lo = 0.010830424696249145;      // kExpTabC
    const double p2 = 0.5, p3 = 0.16666666666666666, p4 = 0.041666666666666664, p5 = 0.008333333333333333;      // kExpTabP
    const double t = fma(x, c_hi, magic);
    const int n = double_lo(t);
    const double kd = t - magic;
    const double r = fma(kd, -c_lo, x);
    const double T = tb.exp2_frac(n & (kExpTabSize - 1));
    int k = n >> 6;
    double q;                                              // exp(r) - 1 = r + r^2/2 + ... + r^5/120
    if (kShallow) {
        const double r2 = r * r;
        const double qa = fma(p2, r, 1.0);
        const double q34 = fma(p4, r, p3);
        const double r3 = r2 * r;
        const double qb = fma(p5, r2, q34);
        q = fma(r3, qb, r * qa);
    } else {
        q = fma(p5, r, p4);
        q = fma(q, r, p3);
        q = fma(q, r, p2);
        q = fma(q, r, 1.0);
        q = q * r;
    }
    const double v = fma(T, q, T);                         // in [0.99, 2.02)
    k = k > -1021 ? k : -1021;
    return hilo_to_double(double_hi(v) + (k << 20), double_lo(v));
}

struct PsiExp {
    double psi;   // the reference's digamma(x)
    double e;     // exp(psi)
};

// psi(x) and exp(psi(x)) for 1e-8 <= x, x + 1 < 2^kPsiOctaves (the caller keeps other arguments away:
// the row index is not clamped)
template <bool kShallow, class Tables>
ONILDA_HD PsiExp psi_exp_tab(double x, const Tables& tb)
{
    const double t = x + 1.0;
    double a[6];
    tb.psi_row(psi_row_of(t), a);
    const double u = psi_cell_offset(t);
    const double inv = fast_rcp(x);
    PsiExp out;
    out.psi = psi_poly<kShallow>(a, u) - inv;
    out.e = fast_exp_tab<kShallow>(out.psi, tb);
    return out;
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
