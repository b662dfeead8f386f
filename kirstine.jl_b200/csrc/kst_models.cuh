// Device functors for the registered model families.
//
// A functor replaces the user callbacks of the reference (src/user.jl:18-98: unit_length,
// jacobianmatrix!, map_to_covariate!, update_model_vcov!, dimension) for one family.
//
// Design (not a translation of jacobianmatrix!): every Jacobian row is factored as
//     g(c, theta) = diag(cs(theta)) * gt(pc(c), dc(theta))
// with
//   cs  per-draw column scales  (depend on theta only, computed once per problem),
//   dc  per-draw constants      (theta-only transcendentals hoisted out of the hot loop),
//   pc  per-row constants       (covariate-only transcendentals, once per particle),
//   gt  the reduced row, the only thing evaluated per (row, draw).
// The normalised information matrix is then M = D Mt D with D = diag(cs) and
// Mt = sum_rows omega * gt gt'; log det, traces of inverses and DeltaMethod products are taken
// on Mt with D folded into per-draw constants (kst_linalg.cuh).  1/sigma of the unit covariance
// is folded into cs as well, so omega = w_k * varfac with varfac = 1 for constant variance.
//
// "Trivial" rows have gt = e_CONST_COL (or 0 when the family has no constant column): they
// contribute nothing to the per-draw accumulators and are compacted away per particle.
//
// row() returns the WEIGHTED reduced row  g = sqrt(omega) * gt  (sw = sqrt(omega) comes from the
// per-particle prologue), so that Mt += g g' needs no separate weight multiplications; each family
// folds sw into its cheapest product chain.  post_row() lets a family precompute sw-dependent
// per-row constants in the prologue.
#pragma once
#include <math.h>

#ifndef KST_SU_FMA
#define KST_SU_FMA 1
#endif
#ifndef KST_EXP_RED
#define KST_EXP_RED 1
#endif

#include "kst_math.cuh"
#include "../../include/kirstine_b200.h"

struct KstModelParams {
    double v[4];    // model parameters (sigma or sigma_squared)
    double cpv[16]; // parameters of the covariate parameterization (fixed dose, fixed times, log base)
    int m;   // unit length
    int cp;  // covariate parameterization id
    int d;   // design-region dimension
    int pad;
};

// Emax: test/example-emax.jl:41-47.  row = [1, x/(x+ec50), -emax x/(x+ec50)^2]
struct KstEmax {
    static constexpr int R = 3, NDC = 1, NPC = 2, CONST_COL = 0;
    static constexpr int ROWS = 4;  // rows interleaved per loop iteration in the objective kernel
    __device__ static void prep_draw(const double *th, const KstModelParams &mp, double *dc, double *cs)
    {
        const double isig = 1.0 / sqrt(mp.v[0]);  // v[0] = sigma_squared
        dc[0] = th[2];
        cs[0] = isig;
        cs[1] = isig;
        cs[2] = -th[1] * isig;
    }
    // returns true when the row is trivial
    __device__ static bool prep_row(const double *z, int, const KstModelParams &, double *pc, double &varfac)
    {
        pc[0] = z[0];
        pc[1] = 0.0;
        varfac = 1.0;
        return z[0] == 0.0;
    }
    __device__ static void post_row(double *pc, double sw) { pc[1] = sw * pc[0]; }
    static constexpr bool FAST_EXP = false;
    __device__ static double arg_bound(const double *, const double *) { return 0.0; }
    template <bool CLAMP = true>
    __device__ __forceinline__ static void row(const double *pc, const double *dc, double sw, double *g, KstTab)
    {
        const double rho = kst_rcp(pc[0] + dc[0]);
        const double sa = pc[1] * rho;  // sw x / (x + ec50)
        g[0] = sw;
        g[1] = sa;
        g[2] = sa * rho;
    }
};

// sigmoid Emax: docs/src/tutorial.md:127-142.
//   A = x^h/(x^h+ed50^h) = 1/(1+e),  e = exp(q),  q = h (log ed50 - log x),  A B = emax e/(1+e)^2
//   row = [1, A, -A B h/ed50, A B log(x/ed50)] = diag(1, 1, -emax h/ed50, -emax/h) [1, A, u, u q]
//   dose == 0 -> [1, 0, 0, 0] (trivial row)
// |q| beyond 708 saturates inside kst_exp, so e stays finite and u = e/(1+e)^2 never becomes Inf*0;
// a NaN parameter still reaches the row through q and makes the draw singular like in the reference.
struct KstSigEmax {
    static constexpr int R = 4, NDC = 2, NPC = 1, CONST_COL = 0;
    static constexpr int ROWS = 4;
    __device__ static void prep_draw(const double *th, const KstModelParams &mp, double *dc, double *cs)
    {
        const double isig = 1.0 / fabs(mp.v[0]);  // v[0] = sigma
        dc[0] = th[3];
        dc[1] = th[3] * log(th[2]);
        cs[0] = isig;
        cs[1] = isig;
        cs[2] = -th[1] * th[3] / th[2] * isig;
        cs[3] = -th[1] / th[3] * isig;
    }
    __device__ static bool prep_row(const double *z, int, const KstModelParams &, double *pc, double &varfac)
    {
        pc[0] = log(z[0]);
        varfac = 1.0;
        return z[0] == 0.0;
    }
    __device__ static void post_row(double *, double) {}
    // upper bound of |q| over all draws for a row, from the per-problem maxima of |dc| (see KstObjParams::dcmax)
    static constexpr bool FAST_EXP = true;
    __device__ static double arg_bound(const double *pc, const double *dcmax) { return fma(dcmax[0], fabs(pc[0]), dcmax[1]); }
    template <bool CLAMP = true>
    __device__ __forceinline__ static void row(const double *pc, const double *dc, double sw, double *g, KstTab tab)
    {
        const double q = fma(-dc[0], pc[0], dc[1]);
        const double e = kst_exp<KST_EXP_RED, CLAMP>(q, tab);
        const double a = kst_rcp(1.0 + e);
        const double sa = sw * a;
#if KST_SU_FMA
        const double su = fma(-sa, a, sa);  // sw a (1 - a) = sw e / (1+e)^2, one instruction
#else
        const double su = sa * (e * a);  // sw e / (1+e)^2
#endif
        g[0] = sw;
        g[1] = sa;
        g[2] = su;
        g[3] = su * q;
    }
};

// one-compartment model (Atkinson et al. 1993): test/example-compartment.jl:14-21
//   row = [A s t, -E s t, E - A],  A = exp(-a t), E = exp(-e t)
// VUMod (test/example-vector.jl:9-50) is the same row at times (i-1)*z with unit covariance
// diag(sigma_squared (1 + t_i)).
struct KstTpc {
    static constexpr int R = 3, NDC = 2, NPC = 2, CONST_COL = -1;
    static constexpr int ROWS = 2;  // two exponentials per row: pairs already give four chains
    __device__ static void prep_draw(const double *th, const KstModelParams &mp, double *dc, double *cs)
    {
        const double isig = 1.0 / fabs(mp.v[0]);  // v[0] = sigma
        dc[0] = -th[0];
        dc[1] = -th[1];
        cs[0] = th[2] * isig;
        cs[1] = -th[2] * isig;
        cs[2] = isig;
    }
    __device__ static bool prep_row(const double *z, int, const KstModelParams &, double *pc, double &varfac)
    {
        pc[0] = z[0];
        pc[1] = 0.0;
        varfac = 1.0;
        return z[0] == 0.0;
    }
    __device__ static void post_row(double *pc, double sw) { pc[1] = sw * pc[0]; }
    static constexpr bool FAST_EXP = true;
    __device__ static double arg_bound(const double *pc, const double *dcmax) { return fmax(dcmax[0], dcmax[1]) * fabs(pc[0]); }
    template <bool CLAMP = true>
    __device__ __forceinline__ static void row(const double *pc, const double *dc, double sw, double *g, KstTab tab)
    {
        const double A = kst_exp<KST_EXP_RED, CLAMP>(dc[0] * pc[0], tab);
        const double E = kst_exp<KST_EXP_RED, CLAMP>(dc[1] * pc[0], tab);
        g[0] = pc[1] * A;  // sw t A
        g[1] = pc[1] * E;
        g[2] = sw * (E - A);
    }
};

struct KstVu : KstTpc {
    __device__ static void prep_draw(const double *th, const KstModelParams &mp, double *dc, double *cs)
    {
        const double isig = 1.0 / sqrt(mp.v[0]);  // v[0] = sigma_squared
        dc[0] = -th[0];
        dc[1] = -th[1];
        cs[0] = th[2] * isig;
        cs[1] = -th[2] * isig;
        cs[2] = isig;
    }
    __device__ static bool prep_row(const double *z, int i, const KstModelParams &, double *pc, double &varfac)
    {
        const double t = (double)i * z[0];
        pc[0] = t;
        pc[1] = 0.0;
        varfac = 1.0 / (1.0 + t);
        return t == 0.0;
    }
};

// Trait: the unit covariance Sigma(c) of the family is a dense m x m matrix, i.e. its update_model_vcov!
// (src/user.jl:85) fills more than the diagonal.  Such a family provides
//     static void vcov(const double *z, const KstModelParams &mp, double *S)     S: m x m, column-major, full
// and the kernels do what update_model_workspace! / average_fishermatrix! do with it
// (src/model-nonlinear-regression.jl:39-49,60-86): Cholesky factor Sigma = U'U once per design point, every unit's
// Jacobian whitened as U^-T J before the rank-m update.  Families with a diagonal Sigma (all of the reference's own
// examples) keep the cheaper form: one scalar `varfac` per row folded into the row weight.
template <class MF>
struct KstDense {
    static constexpr bool value = false;
};

// Vector units with correlated errors: the mean function and covariates of VUMod (test/example-vector.jl:9-50: m
// measurements at times (i-1) z), unit covariance AR(1)-correlated with the time-dependent variances of VUMod,
//     Sigma_ij = sigma_squared rho^|i-j| sqrt((1 + t_i)(1 + t_j)),     |rho| < 1.
// rho = 0 is VUMod itself.  model_params = {sigma_squared, rho}.  Not a model of the reference's tests or docs: it is
// the registered example of a NonlinearRegression whose Sigma(c) is not diagonal, which the reference supports
// for every user model through potrf! + trsm!.
struct KstVuAr : KstTpc {
    __device__ static void prep_draw(const double *th, const KstModelParams &, double *dc, double *cs)
    {
        dc[0] = -th[0];
        dc[1] = -th[1];
        cs[0] = th[2];
        cs[1] = -th[2];
        cs[2] = 1.0;  // sigma_squared lives in Sigma
    }
    __device__ static bool prep_row(const double *z, int i, const KstModelParams &, double *pc, double &varfac)
    {
        pc[0] = (double)i * z[0];
        pc[1] = 0.0;
        varfac = 1.0;
        return false;  // no row of a correlated unit can be dropped on its own
    }
    __device__ static void vcov(const double *z, const KstModelParams &mp, double *S)
    {
        const int m = mp.m;
        for (int j = 0; j < m; j++)
            for (int i = 0; i < m; i++) {
                const double ti = (double)i * z[0], tj = (double)j * z[0];
                S[i + j * m] = mp.v[0] * pow(mp.v[1], (double)(i > j ? i - j : j - i)) * sqrt((1.0 + ti) * (1.0 + tj));
            }
    }
};
template <>
struct KstDense<KstVuAr> {
    static constexpr bool value = true;
};

// dose-time-response model: test/example-dtr.jl:27-47, docs/src/dtr.md:87-129.  theta = (a, e, e0, emax, ec50)
//   A = exp(-a t), B = exp(-e t), C = B - A, rd = e - a, den = dose C a - rd ec50, w = 1/den
//   row = [-dose ec50 emax (A t a rd + e C) w^2, dose a ec50 emax (B t rd + C) w^2, 1, dose a C w, dose a emax C rd w^2]
//       = diag(-ec50 emax, a ec50 emax, 1, a, a emax rd) [dose (A t a rd + e C) w^2, dose (B t rd + C) w^2, 1, dose C w, dose C w^2]
// dose == 0 or t == 0 -> [0, 0, 1, 0, 0] (trivial row).
// Covariate parameterizations (docs/src/dtr.md): CopyBoth :197-203, FixedDose :331-339, FixedTimes :401-409,
// EquidistantTimes :484-495, LogEquidistantTimes :663-681.
struct KstDtr {
    static constexpr int R = 5, NDC = 7, NPC = 3, CONST_COL = 2;
    static constexpr int ROWS = 2;
    __device__ static void prep_draw(const double *th, const KstModelParams &mp, double *dc, double *cs)
    {
        const double isig = 1.0 / fabs(mp.v[0]);  // v[0] = sigma
        const double a = th[0], e = th[1], emax = th[3], ec50 = th[4], rd = e - a;
        dc[0] = -a;
        dc[1] = -e;
        dc[2] = a;
        dc[3] = rd;
        dc[4] = a * rd;
        dc[5] = e;
        dc[6] = -rd * ec50;
        cs[0] = -ec50 * emax * isig;
        cs[1] = a * ec50 * emax * isig;
        cs[2] = isig;
        cs[3] = a * isig;
        cs[4] = a * emax * rd * isig;
    }
    __device__ static void covariate(const double *z, int i, const KstModelParams &mp, double &dose, double &t)
    {
        switch (mp.cp) {
        case KST_CP_DTR_FIXEDDOSE: dose = mp.cpv[0]; t = i == 0 ? z[0] : 0.0; break;
        case KST_CP_DTR_FIXEDTIMES: dose = z[0]; t = mp.cpv[i]; break;
        case KST_CP_DTR_EQUIDISTANT: dose = z[0]; t = (double)i * z[1]; break;
        case KST_CP_DTR_LOGEQUIDISTANT: {
            dose = z[0];
            t = i == 0 ? 0.0 : z[1];
            for (int j = 2; j <= i; j++) t = mp.cpv[0] * t;  // time[j] = b * time[j-1], same rounding as the reference
            break;
        }
        default: dose = z[0]; t = i == 0 ? z[1] : 0.0;  // CopyBoth
        }
    }
    __device__ static bool prep_row(const double *z, int i, const KstModelParams &mp, double *pc, double &varfac)
    {
        double dose, t;
        covariate(z, i, mp, dose, t);
        pc[0] = t;
        pc[1] = dose;
        pc[2] = 0.0;
        varfac = 1.0;
        return dose == 0.0 || t == 0.0;
    }
    __device__ static void post_row(double *pc, double sw) { pc[2] = sw * pc[1]; }
    static constexpr bool FAST_EXP = true;
    __device__ static double arg_bound(const double *pc, const double *dcmax) { return fmax(dcmax[0], dcmax[1]) * fabs(pc[0]); }
    template <bool CLAMP = true>
    __device__ __forceinline__ static void row(const double *pc, const double *dc, double sw, double *g, KstTab tab)
    {
        const double t = pc[0];
        const double A = kst_exp<KST_EXP_RED, CLAMP>(dc[0] * t, tab);
        const double B = kst_exp<KST_EXP_RED, CLAMP>(dc[1] * t, tab);
        const double C = B - A;
        const double w = kst_rcp(fma(pc[1] * C, dc[2], dc[6]));  // 1 / (dose C a - rd ec50)
        const double sdw = pc[2] * w;                             // sw dose w
        const double sdw2 = sdw * w;
        g[0] = sdw2 * fma(A * t, dc[4], dc[5] * C);
        g[1] = sdw2 * fma(B * t, dc[3], C);
        g[2] = sw;
        g[3] = sdw * C;
        g[4] = sdw2 * C;
    }
};

// ------------------------------------------------------------------------------------------
// logistic regression: the user-defined Model supertype of docs/src/extend-model.md:59-106.
//   M(zeta, theta) = sum_k w_k p_k (1 - p_k) J_k' J_k,   p_k = expit(eta(c_k, theta)),   J_k = d eta / d theta  (1 x r)
// As a weighted reduced row: g = sqrt(w_k) sqrt(p (1 - p)) J,  and with h = exp(-eta / 2)
//   sqrt(p (1 - p)) = h / (1 + h^2)
// (no square root, no cancellation in 1 - p).  |eta| beyond 1416 saturates inside kst_exp: h stays finite, h^2
// overflows to Inf, its reciprocal is 0 and the row vanishes, as p (1 - p) does in the reference.
// There is no unit covariance and no trivial row.
// ------------------------------------------------------------------------------------------

// linear predictor eta = beta_0 + sum_j beta_j x_j on a C-dimensional covariate copied from the design point
// (LinLogReg + CopyVector, docs/src/extend-model.md:123-179).  theta = (beta_0, ..., beta_C), J = [1, x_1, ..., x_C].
template <int C>
struct KstLinLogReg {
    static constexpr int R = C + 1, NDC = C + 1, NPC = C, CONST_COL = -1;
    static constexpr int ROWS = 4;
    __device__ static void prep_draw(const double *th, const KstModelParams &, double *dc, double *cs)
    {
#pragma unroll
        for (int i = 0; i < R; i++) {
            dc[i] = -0.5 * th[i];
            cs[i] = 1.0;
        }
    }
    __device__ static bool prep_row(const double *z, int, const KstModelParams &, double *pc, double &varfac)
    {
#pragma unroll
        for (int i = 0; i < C; i++) pc[i] = z[i];
        varfac = 1.0;
        return false;
    }
    __device__ static void post_row(double *, double) {}
    static constexpr bool FAST_EXP = false;
    __device__ static double arg_bound(const double *, const double *) { return 0.0; }
    template <bool CLAMP = true>
    __device__ __forceinline__ static void row(const double *pc, const double *dc, double sw, double *g, KstTab tab)
    {
        double q = dc[0];  // -eta / 2
#pragma unroll
        for (int i = 0; i < C; i++) q = fma(dc[1 + i], pc[i], q);
        const double h = kst_exp<2>(q, tab);
        const double s = sw * h * kst_rcp(fma(h, h, 1.0));
        g[0] = s;
#pragma unroll
        for (int i = 0; i < C; i++) g[1 + i] = s * pc[i];
    }
};

// eta = a sin(2 pi b (t - c)), theta = (a, b, c)  (SinLogReg, docs/src/extend-model.md:330-355)
//   J = [sin phi, a cos phi 2 pi (t - c), -a cos phi 2 pi b] = diag(1, 2 pi a, -2 pi a b) [sin phi, cos phi (t - c), cos phi]
struct KstSinLogReg {
    static constexpr int R = 3, NDC = 3, NPC = 1, CONST_COL = -1;
    static constexpr int ROWS = 2;
    __device__ static void prep_draw(const double *th, const KstModelParams &, double *dc, double *cs)
    {
        const double twopi = 2 * 3.14159265358979323846;
        dc[0] = twopi * th[1];
        dc[1] = th[2];
        dc[2] = -0.5 * th[0];
        cs[0] = 1.0;
        cs[1] = twopi * th[0];
        cs[2] = -twopi * th[0] * th[1];
    }
    __device__ static bool prep_row(const double *z, int, const KstModelParams &, double *pc, double &varfac)
    {
        pc[0] = z[0];
        varfac = 1.0;
        return false;
    }
    __device__ static void post_row(double *, double) {}
    static constexpr bool FAST_EXP = false;
    __device__ static double arg_bound(const double *, const double *) { return 0.0; }
    template <bool CLAMP = true>
    __device__ __forceinline__ static void row(const double *pc, const double *dc, double sw, double *g, KstTab tab)
    {
        const double u = pc[0] - dc[1];
        double sn, cn;
        sincos(dc[0] * u, &sn, &cn);
        const double h = kst_exp<2>(dc[2] * sn, tab);
        const double s = sw * h * kst_rcp(fma(h, h, 1.0));
        const double sc = s * cn;
        g[0] = s * sn;
        g[1] = sc * u;
        g[2] = sc;
    }
};

// Trait: draws evaluated per thread and pass over a particle's rows in the objective kernel (1 or 2).  Two pay off
// where the per-draw bookkeeping weighs as much as the rows themselves (few, cheap rows).
#ifndef KST_NDRAW_EMAX
#define KST_NDRAW_EMAX 2
#endif
#ifndef KST_NDRAW_SIGEMAX
#define KST_NDRAW_SIGEMAX 2
#endif
#ifndef KST_NDRAW_TPC
#define KST_NDRAW_TPC 2
#endif
template <class MF>
struct KstDraws {
    static constexpr int value = 1;
};
template <>
struct KstDraws<KstEmax> {
    static constexpr int value = KST_NDRAW_EMAX;
};
template <>
struct KstDraws<KstSigEmax> {
    static constexpr int value = KST_NDRAW_SIGEMAX;
};
template <>
struct KstDraws<KstTpc> {
    static constexpr int value = KST_NDRAW_TPC;
};
// (the dose-time-response family, r = 5, stays at one draw: two sets of 15 accumulators spill; measured -11 %)
template <>
struct KstDraws<KstVu> {
    static constexpr int value = 2;
};
template <int C>
struct KstDraws<KstLinLogReg<C>> {
    static constexpr int value = 2;
};
template <>
struct KstDraws<KstSinLogReg> {
    static constexpr int value = 2;
};
