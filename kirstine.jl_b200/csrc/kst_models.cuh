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

#include "kst_math.cuh"

struct KstModelParams {
    double v[4];
    int m;   // unit length
    int cp;  // covariate parameterization id
    int d;   // design-region dimension
    int pad;
};

// Emax: test/example-emax.jl:41-47.  row = [1, x/(x+ec50), -emax x/(x+ec50)^2]
struct KstEmax {
    static constexpr int R = 3, NDC = 1, NPC = 2, CONST_COL = 0;
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
    __device__ __forceinline__ static void row(const double *pc, const double *dc, double sw, double *g, KstTab tab)
    {
        const double q = fma(-dc[0], pc[0], dc[1]);
        const double e = kst_exp(q, tab);
        const double a = kst_rcp(1.0 + e);
        const double sa = sw * a;
        const double su = sa * (e * a);  // sw e / (1+e)^2
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
    __device__ __forceinline__ static void row(const double *pc, const double *dc, double sw, double *g, KstTab tab)
    {
        const double A = kst_exp(dc[0] * pc[0], tab);
        const double E = kst_exp(dc[1] * pc[0], tab);
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
