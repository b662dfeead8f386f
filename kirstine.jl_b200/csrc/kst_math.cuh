// FP64 elementary functions tuned for the evaluation kernels.
//
// Why not libdevice exp()/1.0/x: the objective kernel is bound by the FP64 pipe (one warp
// instruction every two cycles per SM sub-partition) and, almost as tightly, by instruction issue.
// libdevice's exp costs 15 FP64 instructions plus ~25 issue slots of 64-bit immediates (UMOV pairs)
// and slow-path branches; its division adds a denormal/overflow fix-up call.  The versions below
//   kst_exp : 32-entry 2^(j/32) table in shared memory + degree-6 Taylor polynomial on
//             |r| <= ln2/64 (truncation error 3.5e-18)                     -> 11 FP64 instructions
//   kst_rcp : MUFU.RCP64H seed (2^-20) + two Newton steps (2^-80)          ->  4 FP64 instructions
// keep full double accuracy (<= 1 ulp, checked against mpmath through kst_math_selftest) and take
// their constants from the constant bank as instruction operands.
#pragma once
#include <math.h>

__constant__ double kst_exp2_tab_c[32] = {
    0x1.0000000000000p+0, 0x1.059b0d3158574p+0, 0x1.0b5586cf9890fp+0, 0x1.11301d0125b51p+0,
    0x1.172b83c7d517bp+0, 0x1.1d4873168b9aap+0, 0x1.2387a6e756238p+0, 0x1.29e9df51fdee1p+0,
    0x1.306fe0a31b715p+0, 0x1.371a7373aa9cbp+0, 0x1.3dea64c123422p+0, 0x1.44e086061892dp+0,
    0x1.4bfdad5362a27p+0, 0x1.5342b569d4f82p+0, 0x1.5ab07dd485429p+0, 0x1.6247eb03a5585p+0,
    0x1.6a09e667f3bcdp+0, 0x1.71f75e8ec5f74p+0, 0x1.7a11473eb0187p+0, 0x1.82589994cce13p+0,
    0x1.8ace5422aa0dbp+0, 0x1.93737b0cdc5e5p+0, 0x1.9c49182a3f090p+0, 0x1.a5503b23e255dp+0,
    0x1.ae89f995ad3adp+0, 0x1.b7f76f2fb5e47p+0, 0x1.c199bdd85529cp+0, 0x1.cb720dcef9069p+0,
    0x1.d5818dcfba487p+0, 0x1.dfc97337b9b5fp+0, 0x1.ea4afa2a490dap+0, 0x1.f50765b6e4540p+0};

// 0: 32/ln2   1: 1.5*2^52   2: -ln2/32 (high 38 bits)   3: -ln2/32 (low)   4..8: 1/6!, 1/5!, 1/4!, 1/3!, 1/2!
__constant__ double kst_mc[9] = {0x1.71547652b82fep+5, 6755399441055744.0, -0x1.62e42fefa0000p-6,
                                 -0x1.cf79abc9e3b3ap-45, 0x1.6c16c16c16c17p-10, 0x1.1111111111111p-7,
                                 0x1.5555555555555p-5, 0x1.5555555555555p-3, 0.5};

#define KST_EXP_TAB_N 32

// copy the table to shared memory; the caller synchronises the CTA afterwards
__device__ __forceinline__ void kst_math_init(double *s_tab)
{
    if (threadIdx.x < KST_EXP_TAB_N) s_tab[threadIdx.x] = kst_exp2_tab_c[threadIdx.x];
}

// 1/x for positive, normal, finite x (pivots, 1 + e): no fix-up paths.
__device__ __forceinline__ double kst_rcp(double x)
{
    double y;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    double e = fma(-x, y, 1.0);
    y = fma(y, e, y);
    e = fma(-x, y, 1.0);
    y = fma(y, e, y);
    return y;
}

// exp(x) for |x| < 708 (no range check)
__device__ __forceinline__ double kst_exp_core(double x, const double *__restrict__ tab)
{
    const double t = fma(x, kst_mc[0], kst_mc[1]);
    const int n = __double2loint(t);
    const double nd = t - kst_mc[1];
    double r = fma(nd, kst_mc[2], x);
    r = fma(nd, kst_mc[3], r);
    double p = fma(kst_mc[4], r, kst_mc[5]);
    p = fma(p, r, kst_mc[6]);
    p = fma(p, r, kst_mc[7]);
    p = fma(p, r, kst_mc[8]);
    const double r2 = r * r;
    p = fma(r2, p, r);
    const double T = tab[n & (KST_EXP_TAB_N - 1)];
    const double y = fma(T, p, T);
    return __hiloint2double(__double2hiint(y) + ((n >> 5) << 20), __double2loint(y));
}

// exp(x) for any x: the table path inside (-708, 708), libdevice outside (rare: overflow, underflow, NaN)
__device__ __forceinline__ double kst_exp(double x, const double *__restrict__ tab)
{
    if ((__double2hiint(x) & 0x7fffffff) >= 0x40862000) return exp(x);
    return kst_exp_core(x, tab);
}

// exp(x) with x first clamped to [-700, 700] (NaN passes through as NaN)
__device__ __forceinline__ double kst_exp_clamped(double x, const double *__restrict__ tab)
{
    if ((__double2hiint(x) & 0x7fffffff) >= 0x4085e000) {
        if (x != x) return x;
        x = x > 0.0 ? 700.0 : -700.0;
    }
    return kst_exp_core(x, tab);
}
