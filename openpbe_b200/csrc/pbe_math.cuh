// pbe_math.cuh -- FP64 device math for the PBE source kernels (sm_100a).
//
// The MOC kernel spends >70 % of its FP64 issue slots in exp(): a branch-free version with a
// known instruction count (15 FP64-pipe instructions) is what the roofline model in DESIGN.md
// counts.  The functions are __host__ __device__ so that tests/test_device_math_host.py can
// compile this header with g++ and check the arithmetic against libm without a GPU; the
// product only ever runs the device instantiation.
#pragma once

#include <math.h>
#include <stdint.h>
#include <string.h>

#if defined(__CUDACC__)
#define PBE_HD __host__ __device__ __forceinline__
#else
#define PBE_HD inline
#endif

namespace pbe {

PBE_HD double fma_(double a, double b, double c) {
#if defined(__CUDA_ARCH__)
    return __fma_rn(a, b, c);
#else
    return fma(a, b, c);
#endif
}

PBE_HD int hi_word(double x) {
#if defined(__CUDA_ARCH__)
    return __double2hiint(x);
#else
    uint64_t u;
    memcpy(&u, &x, 8);
    return (int)(u >> 32);
#endif
}
PBE_HD int lo_word(double x) {
#if defined(__CUDA_ARCH__)
    return __double2loint(x);
#else
    uint64_t u;
    memcpy(&u, &x, 8);
    return (int)(u & 0xffffffffu);
#endif
}
PBE_HD double make_double(int hi, int lo) {
#if defined(__CUDA_ARCH__)
    return __hiloint2double(hi, lo);
#else
    uint64_t u = ((uint64_t)(uint32_t)hi << 32) | (uint32_t)lo;
    double x;
    memcpy(&x, &u, 8);
    return x;
#endif
}

// exp(x) for x <= 0 (the only sign the Coulaloglou-Tavlarides efficiencies produce:
// coalescenceKernels/CoulaloglouTavlarides.C:84-88, breakupKernels/CoulaloglouTavlarides.C:86-90).
// Cody-Waite reduction x = k ln2 + r, |r| <= ln2/2, degree-11 polynomial from
// tools/gen_exp_coeffs.py (0.15 ulp in exact arithmetic; < 1 ulp measured after rounding),
// 2^k applied by an integer add on the exponent field.  15 FP64-pipe instructions; the range
// handling is done on the integer pipe so it costs no FP64 issue slot:
//   * x <= -800 (including -inf): 0;  results below 2^-1021 are flushed to 0 (libm returns a
//     denormal there; the difference is < 2.3e-308 absolute);
//   * NaN propagates (k = 0 for the quiet NaN, the polynomial carries the NaN);
//   * 0 < x <= 709 also works (not on the measured path); larger x is outside the contract.
PBE_HD double exp_neg(double x) {
    const double L2E = 1.4426950408889634074;         // 1/ln2
    const double SHIFT = 6755399441055744.0;          // 1.5 * 2^52: round-to-nearest-int trick
    const double LN2_HI = 6.93147180369123816490e-01; // high 33 bits of ln2
    const double LN2_LO = 1.90821492927058770002e-10;
    double t = fma_(x, L2E, SHIFT);
    int k = lo_word(t);
    double kd = t - SHIFT;
    double r = fma_(kd, -LN2_HI, x);
    r = fma_(kd, -LN2_LO, r);
    double p = 2.510134692835034e-08;
    p = fma_(p, r, 2.7621338669142925e-07);
    p = fma_(p, r, 2.75572664186123e-06);
    p = fma_(p, r, 2.4801518641395993e-05);
    p = fma_(p, r, 0.00019841269864380513);
    p = fma_(p, r, 0.0013888888918939168);
    p = fma_(p, r, 0.008333333333329793);
    p = fma_(p, r, 0.04166666666662064);
    p = fma_(p, r, 0.16666666666666669);
    p = fma_(p, r, 0.5000000000000001);
    p = fma_(p, r, 1.0);
    p = fma_(p, r, 1.0);
    // p in [0.70, 1.42): its exponent field is 1022 or 1023; adding k keeps it normal for
    // k in [-1021, 1023].
    int hi = hi_word(p) + (int)((unsigned)k << 20);
    double res = make_double(hi, lo_word(p));
    // hi word of x in [hi(-800.0), hi(-inf)]  <=>  x <= -800
    const unsigned ux = (unsigned)hi_word(x);
    const bool under = (ux - 0xC0890000u) <= (0xFFF00000u - 0xC0890000u);
    res = (under || k < -1021) ? 0.0 : res;
    return res;
}

// 2^y for y <= 0, |y| < 2^30 (the caller clamps; NaN propagates).  The pair loop of the TMEM
// MOC kernel calls this once per class pair with y = (K_c log2 e) H_ij, so the argument
// multiply replaces Cody-Waite: t = y + 1.5*2^52 rounds y to the nearest integer k, f = y - k
// is exact, 2^f = 1 + f s(f) with s of degree 9 (tools/gen_exp2_coeffs.py: 5.7 ulp in exact
// arithmetic; the rounding of y itself costs |y| ulp, which dominates).  13 FP64-pipe
// instructions and 2 integer ones.
PBE_HD double exp2_neg_fast(double y) {
    const double SHIFT = 6755399441055744.0; // 1.5 * 2^52
    const double t = y + SHIFT;
    const int k = lo_word(t);
    const double f = y - (t - SHIFT);
    double p = 7.0713011514345125e-09;
    p = fma_(p, f, 1.020646529769823e-07);
    p = fma_(p, f, 1.3215450204669348e-06);
    p = fma_(p, f, 1.5252670450809624e-05);
    p = fma_(p, f, 0.00015403530426708997);
    p = fma_(p, f, 0.0013333558204141327);
    p = fma_(p, f, 0.009618129107617852);
    p = fma_(p, f, 0.055504108664637584);
    p = fma_(p, f, 0.24022650695910078);
    p = fma_(p, f, 0.6931471805599463);
    p = fma_(p, f, 1.0);
    // 2^k by an integer add on the exponent field (p is in [0.70, 1.42): field 1022 or 1023).
    // k is clamped at -1022 instead of testing for underflow: results below 2^-1022 come out
    // as a denormal-range value < 2.3e-308 where exp() itself is denormal.
    const int kc = (k > -1022) ? k : -1022;
    const int hi = hi_word(p) + (int)((unsigned)kc << 20);
    const int lo = lo_word(p);
    return make_double(hi, lo);
}

#if defined(__CUDACC__)
// exp_neg with its polynomial in constant memory: FP64 immediates cost two UMOVs each wherever
// the call is inlined into straight-line code (the QMOM kernel has nine calls and no loop to
// hoist them out of; ncu r01f: 218 UMOV per cell); c[bank][offset] operands cost nothing.
static __constant__ double PBE_EXP_C[16] = {2.510134692835034e-08,  2.7621338669142925e-07, 2.75572664186123e-06,
                                     2.4801518641395993e-05, 0.00019841269864380513, 0.0013888888918939168,
                                     0.008333333333329793,   0.04166666666662064,    0.16666666666666669,
                                     0.5000000000000001,     1.4426950408889634074,  6755399441055744.0,
                                     -6.93147180369123816490e-01, -1.90821492927058770002e-10, 0.0, 0.0};
__device__ __forceinline__ double exp_neg_c(double x) {
    const double SHIFT = PBE_EXP_C[11];
    double t = __fma_rn(x, PBE_EXP_C[10], SHIFT);
    int k = __double2loint(t);
    double kd = t - SHIFT;
    double r = __fma_rn(kd, PBE_EXP_C[12], x);
    r = __fma_rn(kd, PBE_EXP_C[13], r);
    double p = PBE_EXP_C[0];
#pragma unroll
    for (int i = 1; i < 10; ++i) p = __fma_rn(p, r, PBE_EXP_C[i]);
    p = __fma_rn(p, r, 1.0);
    p = __fma_rn(p, r, 1.0);
    int hi = __double2hiint(p) + (int)((unsigned)k << 20);
    double res = __hiloint2double(hi, __double2loint(p));
    const unsigned ux = (unsigned)__double2hiint(x);
    const bool under = (ux - 0xC0890000u) <= (0xFFF00000u - 0xC0890000u);
    return (under || k < -1021) ? 0.0 : res;
}

// 1/x for normal x by MUFU.RCP64H + two Newton steps: no denormal / special-case slow path
// (the IEEE division emits one behind a branch).  <= 1 ulp; 0, inf, NaN behave like 1/x.
__device__ __forceinline__ double rcp_fast(double x) {
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
    r = __fma_rn(__fma_rn(-x, r, 1.0), r, r);
    r = __fma_rn(__fma_rn(-x, r, 1.0), r, r);
    return r;
}

// x^(1/3) for positive normal x without libm's cbrt() (~30 FP64 instructions and a call): the
// exponent is split off (x = xs 8^k, xs in [1/4, 8)), an FP32 seed u0 = 2^(-log2(xs)/3) (MUFU.LG2 /
// MUFU.EX2, |e| < 1e-6) is corrected ONCE with the second-order series of (1 - e)^(-1/3),
// e = 1 - xs u0^3, and x^(1/3) = xs u^2: 10 FP64-pipe instructions, <= 3 ulp.  Everything else
// (<= 0, denormal, inf, NaN) takes the libm call behind one rarely taken branch, with pow(x, 1/3)'s
// sign convention of pow_third() below.
__device__ __forceinline__ double cbrt_pos_fast(double x) {
    const int hi = __double2hiint(x);
    const unsigned ex_field = (unsigned)(hi >> 20) - 1u; // sign bit set or field 0 / 0x7ff -> out of range
    if (ex_field >= 0x7feu) return (x < 0.0) ? NAN : cbrt(x);
    const int ex = (int)ex_field - 1022;
    const int k = (ex * 21846 + 32768) >> 16; // ~ ex / 3, rounded
    const double xs = __hiloint2double(hi - ((3 * k) << 20), __double2loint(x));
    const float lf = __log2f((float)xs);
    const double u0 = (double)exp2f(lf * (-1.0f / 3.0f));
    const double u2 = u0 * u0;
    const double e = __fma_rn(-(xs * u0), u2, 1.0);
    const double c = __fma_rn(2.0 / 9.0, e, 1.0 / 3.0);
    const double us = __fma_rn(u0 * e, c, u0);
    const double rs = xs * (us * us);
    return __hiloint2double(__double2hiint(rs) + (k << 20), __double2loint(rs));
}

// sqrt(x) for positive normal x: MUFU.RSQ64H seed, one Newton step on 1/sqrt, one correction of
// the root (8 FP64-pipe instructions, <= 1 ulp, no slow-path call); 0, inf, negative, NaN as sqrt().
__device__ __forceinline__ double sqrt_pos_fast(double x) {
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    const double h = 0.5 * x;
    y = y * __fma_rn(-(h * y), y, 1.5);
    double s = x * y;
    s = __fma_rn(__fma_rn(-s, s, x), 0.5 * y, s);
    // x = 0 -> y = inf, s = NaN; x = inf -> y = 0, s = NaN; x < 0 -> NaN already
    return (x == 0.0 || x == __longlong_as_double(0x7ff0000000000000LL)) ? x : s;
}

// ninth roots of a node volume: r = x^(1/9), u = x^(-1/9), from which the Coulaloglou-
// Tavlarides kernels take x^(1/3) = r^3, x^(2/9) = r^2, x^(-2/9) = u^2, x^(-5/9) = u^5
// (coalescenceKernels/CoulaloglouTavlarides.C:75-88, breakupKernels/CoulaloglouTavlarides.C:
// 83-92).  x is scaled by 2^(-9k) into [2^-5, 2^5]; an FP32 seed 2^(-log2(x)/9) (MUFU.LG2 /
// MUFU.EX2, |e| < 1e-5) is corrected ONCE with the third-order series of (1 - e)^(-1/9),
// e = 1 - x u^9, division-free: 13 FP64 instructions for both roots against ~60 for two cbrt()
// and a division.  pow() semantics at the edges: x < 0 or NaN -> NaN, 0 -> (0, inf),
// inf -> (inf, 0); denormal x is outside the contract.
__device__ __forceinline__ void ninth_roots(double x, double &r, double &u) {
    const int hi = __double2hiint(x), lo = __double2loint(x);
    const int ex = ((hi >> 20) & 0x7ff) - 1023;
    const int k = (ex * 7282 + 32768) >> 16; // ~ ex / 9
    const double xs = __hiloint2double(hi - ((9 * k) << 20), lo);
    const float lf = __log2f((float)xs);
    const double u0 = (double)exp2f(lf * (-1.0f / 9.0f));
    const double u2 = u0 * u0, u4 = u2 * u2, u8 = u4 * u4;
    const double e = __fma_rn(-(xs * u0), u8, 1.0);
    const double c = __fma_rn(__fma_rn(95.0 / 2187.0, e, 5.0 / 81.0), e, 1.0 / 9.0);
    const double us = __fma_rn(u0 * e, c, u0);
    const double v2 = us * us, v4 = v2 * v2;
    const double rs = xs * (v4 * v4);
    // undo the scaling: u = us 2^-k, r = rs 2^k
    u = __hiloint2double(__double2hiint(us) - (k << 20), __double2loint(us));
    r = __hiloint2double(__double2hiint(rs) + (k << 20), __double2loint(rs));
    const bool pos_normal = (x > 0.0) && (x < __longlong_as_double(0x7ff0000000000000LL));
    if (!pos_normal) {
        const double bad = (x == 0.0) ? 0.0 : ((x > 0.0) ? x : NAN); // 0, +inf, NaN (x < 0 or NaN)
        r = bad;
        u = (x == 0.0) ? __longlong_as_double(0x7ff0000000000000LL) : ((x > 0.0) ? 0.0 : NAN);
    }
}
#endif

// 2^y for y <= ~0, |y| < 2^27 (the caller clamps; NaN propagates), table-driven:
//   2^LOGT y = 2^LOGT k + j + 2^LOGT f,  j in 0..2^LOGT-1,  |f| <= 2^-(LOGT+1):
//   2^y = 2^k T[j] (1 + f s(f)),   T[j] = 2^(j / 2^LOGT) read from shared memory.
// t = y + 1.5*2^(52-LOGT) rounds y to a multiple of 2^-LOGT whose integer 2^LOGT k + j sits in
// the low word of t.  Three sizes (tools/gen_exp2_table_coeffs.py):
//   LOGT 4:  16 entries (128 B = the 32 banks once: any mix of j across a warp is conflict-free),
//            s of degree 5, 0.11 ulp in exact arithmetic, 10 FP64-pipe instructions
//   LOGT 5:  32 entries, s of degree 4 (1.2e-16),  9 FP64-pipe instructions, <= 2-way conflicts
//   LOGT 7: 128 entries, s of degree 3 (1.3e-16),  8 FP64-pipe instructions, <= 8-way conflicts
// against 13 for exp2_neg_fast; the table read and the integer work run on other pipes.
#define PBE_EXP2_TAB128_INIT                                                                       \
    {1.0, 1.0054299011128027, 1.0108892860517005, 1.016378314910953, \
     1.0218971486541166, 1.0274459491187637, 1.0330248790212284, 1.0386341019613787, \
     1.0442737824274138, 1.0499440858006872, 1.0556451783605572, 1.061377227289262, \
     1.0671404006768237, 1.0729348675259756, 1.0787607977571199, 1.0846183622133092, \
     1.0905077326652577, 1.0964290818163769, 1.102382583307841, 1.1083684117236787, \
     1.1143867425958924, 1.1204377524096067, 1.1265216186082418, 1.1326385195987192, \
     1.1387886347566916, 1.1449721444318042, 1.1511892299529827, 1.1574400736337511, \
     1.1637248587775775, 1.1700437696832502, 1.1763969916502812, 1.182784710984341, \
     1.189207115002721, 1.1956643920398273, 1.202156731452703, 1.2086843236265816, \
     1.215247359980469, 1.2218460329727576, 1.22848053610687, 1.2351510639369334, \
     1.241857812073484, 1.2486009771892048, 1.255380757024691, 1.2621973503942507, \
     1.2690509571917332, 1.275941778396392, 1.2828700160787783, 1.2898358734066657, \
     1.2968395546510096, 1.3038812651919358, 1.3109612115247644, 1.318079601266064, \
     1.3252366431597413, 1.3324325470831615, 1.339667524053303, 1.3469417862329458, \
     1.3542555469368927, 1.3616090206382248, 1.3690024229745905, 1.3764359707545302, \
     1.383909881963832, 1.3914243757719262, 1.3989796725383112, 1.4065759938190154, \
     1.4142135623730951, 1.4218926021691656, 1.42961333839197, 1.4373759974489824, \
     1.4451808069770467, 1.4530279958490526, 1.460917794180647, 1.4688504333369818, \
     1.4768261459394993, 1.4848451658727524, 1.4929077282912648, 1.5010140696264256, \
     1.5091644275934228, 1.5173590411982147, 1.5255981507445384, 1.533881997840956, \
     1.5422108254079407, 1.550584877685, 1.559004400237837, 1.567469639965553, \
     1.5759808451078865, 1.5845382652524937, 1.593142151342267, 1.6017927556826934, \
     1.6104903319492543, 1.6192351351948637, 1.6280274218573478, 1.6368674497669644, \
     1.645755478153965, 1.6546917676561943, 1.6636765803267364, 1.6727101796415966, \
     1.681792830507429, 1.6909247992693053, 1.7001063537185235, 1.709337763100463, \
     1.718619298122478, 1.7279512309618377, 1.7373338352737062, 1.746767386199169, \
     1.7562521603732995, 1.7657884359332727, 1.7753764925265212, 1.785016611318935, \
     1.7947090750031072, 1.804454167806624, 1.8142521755003989, 1.8241033854070534, \
     1.8340080864093424, 1.843966568958626, 1.8539791250833855, 1.864046048397789, \
     1.8741676341103, 1.8843441790323345, 1.8945759815869656, 1.9048633418176741, \
     1.9152065613971474, 1.925605943636125, 1.9360617934922943, 1.9465744175792332, \
     1.9571441241754002, 1.9677712232331759, 1.978456026387951, 1.9891988469672663}
#if defined(__CUDACC__)
static __constant__ double PBE_EXP2_TAB128[128] = PBE_EXP2_TAB128_INIT; // staged into shared memory by the kernels
#endif
template <int LOGT>
PBE_HD double exp2_neg_tab(double y, const double *T) {
    static_assert(LOGT == 4 || LOGT == 5 || LOGT == 7, "table sizes with a fitted polynomial");
    const double SHIFT = 6755399441055744.0 / (double)(1 << LOGT); // 1.5 * 2^(52 - LOGT)
    const double t = y + SHIFT;
    const int m = lo_word(t);
    const double f = y - (t - SHIFT);
    double s;
    if (LOGT == 4) {
        s = 0.00015403727871878627;
        s = fma_(s, f, 0.0013333786067768005);
        s = fma_(s, f, 0.009618129106890753);
        s = fma_(s, f, 0.05550410865630706);
        s = fma_(s, f, 0.24022650695910075);
        s = fma_(s, f, 0.6931471805599457);
    } else if (LOGT == 5) {
        s = 0.001333360562976466;
        s = fma_(s, f, 0.009618177060432826);
        s = fma_(s, f, 0.05550410866452594);
        s = fma_(s, f, 0.24022650695611508);
        s = fma_(s, f, 0.6931471805599453);
    } else {
        s = 0.009618131505263852;
        s = fma_(s, f, 0.05550412941916238);
        s = fma_(s, f, 0.24022650695909606);
        s = fma_(s, f, 0.6931471805599049);
    }
    // Integer side kept on the ALU pipe (rotate + mask instead of shifts, which ptxas lowers to
    // IMAD on the FMA pipe: tools/ubench_fp64_issue.cu shows FMA-pipe instructions taking FP64
    // dispatch slots on sm_100a).  Table byte offset = (m & (2^LOGT - 1)) * 8 = rotl(m, 3) & mask.
    constexpr int NT = 1 << LOGT;
#if defined(__CUDA_ARCH__)
    const double Tj = *reinterpret_cast<const double *>(reinterpret_cast<const char *>(T) +
                                                        (__funnelshift_l(m, m, 3) & ((NT - 1) << 3)));
#else
    const double Tj = T[m & (NT - 1)];
#endif
    const double p = fma_(Tj * f, s, Tj); // in [1, 2): exponent field 1023
    // 2^k by an integer add on the exponent field, k = m >> LOGT clamped at -1022 (denormal-range
    // results come out below 4.5e-308 instead of being rounded, like exp2_neg_fast).  -1022 is the
    // lowest safe clamp: p can be just below 1 (f < 0, T[0] = 1: exponent field 1022), and field
    // 1022 - 1023 would wrap into the sign bit (measured: NaN sources in the extreme-cell test).
    // k << 20 = rotl(m, 20 - LOGT) & 0xFFF00000
    const int mc = (m > -1022 * NT) ? m : -1022 * NT;
#if defined(__CUDA_ARCH__)
    const int kbits = (int)(__funnelshift_l(mc, mc, 20 - LOGT) & 0xFFF00000u);
#else
    const int kbits = (int)((unsigned)(mc >> LOGT) << 20);
#endif
    return make_double(hi_word(p) + kbits, lo_word(p));
}

#if defined(__CUDACC__)
// exp2_neg_tab with the 2^k scaling folded into the table value: TB[j] holds the bit pattern of
// 2^(j/2^LOGT) MINUS j << (52 - LOGT), so that adding m << (52 - LOGT) (m = k 2^LOGT + j, one LEA on
// the high word) gives 2^k 2^(j/2^LOGT) directly -- no mask, no separate exponent add.  The
// polynomial is linear in the table value, and scaling by 2^k commutes with every rounding while the
// result stays normal, so the values are bit-identical to exp2_neg_tab there.  3 integer
// instructions per call (index mask, clamp, LEA) instead of 6.
__device__ __forceinline__ double exp2_tab_bias(double t, int j, int logt) {
    return make_double(hi_word(t) - (j << (20 - logt)), lo_word(t));
}
template <int LOGT>
__device__ __forceinline__ double exp2_neg_tabb(double y, const double *TB) {
    static_assert(LOGT == 7, "coefficients fitted for the 128-entry table");
    const double SHIFT = 6755399441055744.0 / (double)(1 << LOGT);
    const double t = y + SHIFT;
    const int m = lo_word(t);
    const double f = y - (t - SHIFT);
    double s = 0.009618131505263852;
    s = fma_(s, f, 0.05550412941916238);
    s = fma_(s, f, 0.24022650695909606);
    s = fma_(s, f, 0.6931471805599049);
    constexpr int NT = 1 << LOGT;
    const double Tb = TB[m & (NT - 1)];
    const int mc = (m > -1022 * NT) ? m : -1022 * NT;
    const double Ts = make_double(hi_word(Tb) + (int)((unsigned)mc << (20 - LOGT)), lo_word(Tb));
    return fma_(Ts * f, s, Ts);
}
#endif

#if defined(__CUDACC__)
// exp(x) = 2^(x log2 e) for x <= 0 given y = x log2(e), through the 128-entry table: 8 FP64-pipe
// instructions; results below 2^-1100 (incl. y = -inf) are 0, NaN propagates (its hi word is
// outside [hi(-1100.0), hi(-inf)]).
__device__ __forceinline__ double exp2_neg_tab_safe(double y, const double *T128) {
    const double r = exp2_neg_tab<7>(y, T128);
    const unsigned uy = (unsigned)__double2hiint(y);
    const bool under = (uy - 0xC0913000u) <= (0xFFF00000u - 0xC0913000u);
    return under ? 0.0 : r;
}
#endif

// pow(x, 1/3) with C pow() semantics for the arguments the reference feeds it
// (coalescenceKernels/CoulaloglouTavlarides.C:75-76): negative finite x -> NaN, +-0 -> 0.
PBE_HD double pow_third(double x) { return x < 0.0 ? NAN : cbrt(x); }

// x^k for integer k >= 0 by repeated multiplication; pow(x, 0) = 1 for every x including 0 and
// NaN, as C pow() (QMOM.C:248,255,260 call pow(double, label)).
PBE_HD double pow_int(double x, int k) {
    double r = 1.0;
    for (int i = 0; i < k; ++i) r *= x;
    return r;
}

} // namespace pbe
