// FP64 elementary functions tuned for the evaluation kernels.
//
// Why not libdevice exp()/1.0/x: the objective kernel is bound by the FP64 pipe (one warp
// instruction every two cycles per SM sub-partition) and, almost as tightly, by instruction issue.
// libdevice's exp costs 15 FP64 instructions plus ~25 issue slots of 64-bit immediates (UMOV pairs)
// and slow-path branches; its division adds a denormal/overflow fix-up call.  The versions below
//   kst_exp : 64-entry 2^(j/64) table in shared memory + degree-5 polynomial on
//             |r| <= ln2/128 (truncation error 4.4e-18)                    -> 10 FP64 instructions
//   kst_log : 64-entry (c_i ~ 1/m_i, -log c_i) table + degree-7 log1p polynomial on
//             |r| <= 2^-7 (truncation error 1.7e-18)                       -> 12 FP64 instructions
//   kst_rcp : MUFU.RCP64H seed + one cubic step (e + e^2)                  ->  3 FP64 instructions
// keep double accuracy (exp <= 1 ulp, rcp <= 2 ulp, log <= 2e-16 absolute on [1,2); checked against mpmath
// through kst_math_selftest) and take their constants from the constant bank as instruction operands.
//
// Measured issue model on B200 (tools/fp64_probe.cu): a warp-wide FP64 instruction holds the sub-partition's
// dispatch port for 2 cycles and every other instruction for 1, so the kernels pay
// 2 * (#FP64) + (#other) cycles per warp: integer glue is trimmed as carefully as the FP64 count.
#pragma once
#include <math.h>

__constant__ double kst_exp2_tab_c[64] = {  // 2^(j/64), correctly rounded
    0x1.0000000000000p+0, 0x1.02c9a3e778061p+0, 0x1.059b0d3158574p+0, 0x1.0874518759bc8p+0,
    0x1.0b5586cf9890fp+0, 0x1.0e3ec32d3d1a2p+0, 0x1.11301d0125b51p+0, 0x1.1429aaea92de0p+0,
    0x1.172b83c7d517bp+0, 0x1.1a35beb6fcb75p+0, 0x1.1d4873168b9aap+0, 0x1.2063b88628cd6p+0,
    0x1.2387a6e756238p+0, 0x1.26b4565e27cddp+0, 0x1.29e9df51fdee1p+0, 0x1.2d285a6e4030bp+0,
    0x1.306fe0a31b715p+0, 0x1.33c08b26416ffp+0, 0x1.371a7373aa9cbp+0, 0x1.3a7db34e59ff7p+0,
    0x1.3dea64c123422p+0, 0x1.4160a21f72e2ap+0, 0x1.44e086061892dp+0, 0x1.486a2b5c13cd0p+0,
    0x1.4bfdad5362a27p+0, 0x1.4f9b2769d2ca7p+0, 0x1.5342b569d4f82p+0, 0x1.56f4736b527dap+0,
    0x1.5ab07dd485429p+0, 0x1.5e76f15ad2148p+0, 0x1.6247eb03a5585p+0, 0x1.6623882552225p+0,
    0x1.6a09e667f3bcdp+0, 0x1.6dfb23c651a2fp+0, 0x1.71f75e8ec5f74p+0, 0x1.75feb564267c9p+0,
    0x1.7a11473eb0187p+0, 0x1.7e2f336cf4e62p+0, 0x1.82589994cce13p+0, 0x1.868d99b4492edp+0,
    0x1.8ace5422aa0dbp+0, 0x1.8f1ae99157736p+0, 0x1.93737b0cdc5e5p+0, 0x1.97d829fde4e50p+0,
    0x1.9c49182a3f090p+0, 0x1.a0c667b5de565p+0, 0x1.a5503b23e255dp+0, 0x1.a9e6b5579fdbfp+0,
    0x1.ae89f995ad3adp+0, 0x1.b33a2b84f15fbp+0, 0x1.b7f76f2fb5e47p+0, 0x1.bcc1e904bc1d2p+0,
    0x1.c199bdd85529cp+0, 0x1.c67f12e57d14bp+0, 0x1.cb720dcef9069p+0, 0x1.d072d4a07897cp+0,
    0x1.d5818dcfba487p+0, 0x1.da9e603db3285p+0, 0x1.dfc97337b9b5fp+0, 0x1.e502ee78b3ff6p+0,
    0x1.ea4afa2a490dap+0, 0x1.efa1bee615a27p+0, 0x1.f50765b6e4540p+0, 0x1.fa7c1819e90d8p+0};

// 0: 64/ln2   1: 1.5*2^52   2: -ln2/64 (high 38 bits)   3: -ln2/64 (low)
// 4..7: c5, c4, c3, c2 of exp(r) ~ 1 + r + r^2 (c2 + c3 r + c4 r^2 + c5 r^3) on |r| <= ln2/128
// (interpolation at Chebyshev nodes; truncation error 4.4e-18 relative)
// 8: -ln2/64 rounded to double (one-step reduction)
__constant__ double kst_mc[9] = {0x1.71547652b82fep+6, 6755399441055744.0, -0x1.62e42fefa0000p-7,
                                 -0x1.cf79abc9e3b3ap-46, 0x1.11111d8fbe766p-7, 0x1.55556b3304ec0p-5,
                                 0x1.5555555555255p-3, 0x1.ffffffffff57fp-2, -0x1.62e42fefa39efp-7};

#define KST_EXP_TAB_N 64
#define KST_EXP_TAB_SHIFT 6
#define KST_LOG_TAB_N 64

// (c_i, -log c_i), c_i = 1/(1 + (i + 0.5)/64) rounded to 13 bits, so that r = m c_i - 1 has |r| <= 2^-7
__constant__ double kst_log_tab_c[2 * KST_LOG_TAB_N] = {
    0x1.fc00000000000p-1, 0x1.010157588de71p-7, 0x1.f440000000000p-1, 0x1.7c61b1cf5dee0p-6,
    0x1.ecc0000000000p-1, 0x1.39f07ba0ebd62p-5, 0x1.e570000000000p-1, 0x1.b46bd74da7698p-5,
    0x1.de60000000000p-1, 0x1.163d6ef957a03p-4, 0x1.d780000000000p-1, 0x1.518874226130ap-4,
    0x1.d0d0000000000p-1, 0x1.8c0b5d97a1585p-4, 0x1.ca50000000000p-1, 0x1.c5ba492f840cbp-4,
    0x1.c400000000000p-1, 0x1.fe89139dbd566p-4, 0x1.bdd0000000000p-1, 0x1.1b7f2d5cba843p-3,
    0x1.b7d0000000000p-1, 0x1.373f423fed9a1p-3, 0x1.b200000000000p-1, 0x1.527e5e4a1b58dp-3,
    0x1.ac50000000000p-1, 0x1.6d827eb7c1e5ap-3, 0x1.a6d0000000000p-1, 0x1.87fa865210911p-3,
    0x1.a170000000000p-1, 0x1.a22e42098f496p-3, 0x1.9c30000000000p-1, 0x1.bc19e74ffcbd8p-3,
    0x1.9710000000000p-1, 0x1.d5b996b97fb5ep-3, 0x1.9210000000000p-1, 0x1.ef095cbde9936p-3,
    0x1.8d30000000000p-1, 0x1.0402994b4f041p-2, 0x1.8870000000000p-1, 0x1.10547f9d26abcp-2,
    0x1.83d0000000000p-1, 0x1.1c784c3bcb34cp-2, 0x1.7f40000000000p-1, 0x1.2896a13e086a4p-2,
    0x1.7ad0000000000p-1, 0x1.348399ada9d94p-2, 0x1.7680000000000p-1, 0x1.403d086cea79cp-2,
    0x1.7240000000000p-1, 0x1.4becf95d97913p-2, 0x1.6e20000000000p-1, 0x1.5765f1749da6bp-2,
    0x1.6a10000000000p-1, 0x1.62d2ef3a0e933p-2, 0x1.6620000000000p-1, 0x1.6e056aa4421d5p-2,
    0x1.6240000000000p-1, 0x1.792955fdf47a2p-2, 0x1.5e70000000000p-1, 0x1.843dddc4fb136p-2,
    0x1.5ac0000000000p-1, 0x1.8f12e873862c8p-2, 0x1.5720000000000p-1, 0x1.99d5d8130607dp-2,
    0x1.5390000000000p-1, 0x1.a485c90ebdb0cp-2, 0x1.5010000000000p-1, 0x1.af21d34189d09p-2,
    0x1.4cb0000000000p-1, 0x1.b977c980d2d99p-2, 0x1.4950000000000p-1, 0x1.c3e8ba8cacf28p-2,
    0x1.4610000000000p-1, 0x1.ce10b0c581285p-2, 0x1.42d0000000000p-1, 0x1.d852b28be5a0ep-2,
    0x1.3fb0000000000p-1, 0x1.e248c1a7c8c26p-2, 0x1.3ca0000000000p-1, 0x1.ec241d5e2ffd0p-2,
    0x1.3990000000000p-1, 0x1.f617ffab248ecp-2, 0x1.36a0000000000p-1, 0x1.ffbb60ca863b3p-2,
    0x1.33b0000000000p-1, 0x1.04bb19dca36b3p-1, 0x1.30d0000000000p-1, 0x1.098994f72c539p-1,
    0x1.2e00000000000p-1, 0x1.0e4898611cce1p-1, 0x1.2b40000000000p-1, 0x1.12f799594efbcp-1,
    0x1.2890000000000p-1, 0x1.17960bd022649p-1, 0x1.25e0000000000p-1, 0x1.1c3f41fa97c6bp-1,
    0x1.2340000000000p-1, 0x1.20d74d2fbafe5p-1, 0x1.20b0000000000p-1, 0x1.255d9cf910a56p-1,
    0x1.1e30000000000p-1, 0x1.29d19fed0c082p-1, 0x1.1bb0000000000p-1, 0x1.2e4fa37fc9578p-1,
    0x1.1940000000000p-1, 0x1.32bab3a7b21e8p-1, 0x1.16e0000000000p-1, 0x1.37123b54987b6p-1,
    0x1.1480000000000p-1, 0x1.3b7344be40311p-1, 0x1.1230000000000p-1, 0x1.3fc0186800513p-1,
    0x1.0ff0000000000p-1, 0x1.43f81e307de5fp-1, 0x1.0db0000000000p-1, 0x1.48391d226992dp-1,
    0x1.0b80000000000p-1, 0x1.4c649aff0ee16p-1, 0x1.0950000000000p-1, 0x1.5098dcce9fabap-1,
    0x1.0730000000000p-1, 0x1.54b6e67ef251cp-1, 0x1.0520000000000p-1, 0x1.58be1b857aebep-1,
    0x1.0310000000000p-1, 0x1.5ccd7df529a38p-1, 0x1.0100000000000p-1, 0x1.60e52f45788e3p-1};

// 0: 2^52 + 2^31 + 1023   1: ln2 (high 42 bits)   2: ln2 (low)   3..8: 1/7, -1/6, 1/5, -1/4, 1/3, -1/2
__constant__ double kst_lc[9] = {4503601774855167.0, 0x1.62e42fefa3800p-1, 0x1.ef35793c76730p-45,
                                 0x1.2492492492492p-3, -0x1.5555555555555p-3, 0x1.999999999999ap-3,
                                 -0.25, 0x1.5555555555555p-2, -0.5};

// shared-memory footprint of both tables, in doubles
#define KST_MATH_SMEM (KST_EXP_TAB_N + 2 * KST_LOG_TAB_N)

// The tables live in shared memory and are addressed through a 32-bit shared-space offset (KstTab):
// a generic pointer would make the compiler rebuild the shared window base (S2UR SR_CgaCtaId + ULEA,
// both variable latency) in front of every lookup.  Layout: exp table, then the log table.
typedef unsigned KstTab;

// copy the tables to shared memory (s_tab: KST_MATH_SMEM doubles, 16-byte aligned); the caller
// synchronises the CTA afterwards
__device__ __forceinline__ KstTab kst_math_init(double *s_tab)
{
    for (int i = threadIdx.x; i < KST_MATH_SMEM; i += blockDim.x)
        s_tab[i] = i < KST_EXP_TAB_N ? kst_exp2_tab_c[i] : kst_log_tab_c[i - KST_EXP_TAB_N];
    return (KstTab)__cvta_generic_to_shared(s_tab);
}

// 1/x for normal, finite, non-zero x (pivots, 1 + e): seed good to ~2^-20 or better, one cubic step
// (error e^3), no fix-up paths.  Result within 2 ulp.
__device__ __forceinline__ double kst_rcp(double x)
{
    double y;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    double e = fma(-x, y, 1.0);
    e = fma(e, e, e);
    return fma(y, e, y);
}

// exp(x).  Exact contract: full accuracy for |x| <= 708; beyond that the result saturates near the
// largest / smallest normal double instead of overflowing to Inf or flushing to 0 (integer clamp of the
// binary exponent, no FP64 compare); |x| above ~4e7 and non-finite x give NaN.
// RED = 2: two-step Cody-Waite reduction, result within 1 ulp of exp(x).  RED = 1: one step with ln2/64
// rounded to double, which adds at most |x| 2^-53 relative error -- no more than the rounding an argument
// that is itself a product (h log x, a t, ...) already carries; the model functors use this form.
template <int RED = 2>
__device__ __forceinline__ double kst_exp(double x, KstTab tab)
{
    const double t = fma(x, kst_mc[0], kst_mc[1]);
    const int n = __double2loint(t);
    const double nd = t - kst_mc[1];
    double r;
    if (RED == 2) {
        r = fma(nd, kst_mc[2], x);
        r = fma(nd, kst_mc[3], r);
    } else {
        r = fma(nd, kst_mc[8], x);
    }
    double p = fma(kst_mc[4], r, kst_mc[5]);
    p = fma(p, r, kst_mc[6]);
    p = fma(p, r, kst_mc[7]);
    const double r2 = r * r;
    p = fma(r2, p, r);
    double T;
    asm("ld.shared.f64 %0, [%1];" : "=d"(T) : "r"(tab + (((unsigned)n & (KST_EXP_TAB_N - 1u)) << 3)));
    const double y = fma(T, p, T);
    const int k = min(max(n >> KST_EXP_TAB_SHIFT, -1021), 1022);
    return __hiloint2double(__double2hiint(y) + (k << 20), __double2loint(y));
}

// log(x) for positive, normal, finite x.
__device__ __forceinline__ double kst_log(double x, KstTab tab)
{
    const int hi = __double2hiint(x);
    const double m = __hiloint2double((hi & 0x000fffff) | 0x3ff00000, __double2loint(x));
    double c, l;
    asm("ld.shared.v2.f64 {%0, %1}, [%2];"
        : "=d"(c), "=d"(l)
        : "r"(tab + (unsigned)(KST_EXP_TAB_N * 8) + (unsigned)(((hi >> 14) & (KST_LOG_TAB_N - 1)) << 4)));
    const double r = fma(m, c, -1.0);
    // binary exponent as a double without a conversion instruction
    const double ed = __hiloint2double(0x43300000, (hi >> 20) ^ 0x80000000) - kst_lc[0];
    double q = fma(kst_lc[3], r, kst_lc[4]);
    q = fma(q, r, kst_lc[5]);
    q = fma(q, r, kst_lc[6]);
    q = fma(q, r, kst_lc[7]);
    q = fma(q, r, kst_lc[8]);
    const double r2 = r * r;
    const double p = fma(r2, q, r);
    return fma(ed, kst_lc[1], l) + fma(ed, kst_lc[2], p);
}
