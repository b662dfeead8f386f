// In-register FP64 small-matrix code for the evaluation kernels.
//
// Replaces the BLAS/LAPACK calls of the reference hot path: syrk! ->
// kst_accum, potrf!/log_det! -> kst_ldlt + one log of the pivot product, posv! + mul! ->
// forward substitution on the unit-lower factor, potri! + tr -> columns of L^-1.
// All loops are compile-time unrolled so every matrix entry lives in a register.
//
// Factorisation: square-root-free Cholesky  Mt = L diag(dv) L'.  The pivots dv[j] are the
// squares of the diagonal of the reference's upper Cholesky factor, so "pivot <= 0 or NaN"
// is exactly the reference's "potrf! failed / U[i,i] <= 0" condition.
#pragma once
#include <math.h>

#include "kst_math.cuh"

#define KST_INF (__longlong_as_double(0x7ff0000000000000LL))

template <int R>
__device__ __forceinline__ constexpr int kst_pidx(int i, int j)  // i <= j, packed upper, row-major
{
    return i * R - (i * (i - 1)) / 2 + (j - i);
}
template <int R>
struct KstPacked {
    static constexpr int N = R * (R + 1) / 2;
};

// Mt += g g'  (upper triangle) for a weighted row g = sqrt(omega) gt.  When CC >= 0 the (CC, CC)
// entry is not accumulated at all: it is the per-particle sum of row weights (gt[CC] == 1).
template <int R, int CC>
__device__ __forceinline__ void kst_accum(double *M, const double *g)
{
#pragma unroll
    for (int i = 0; i < R; i++) {
#pragma unroll
        for (int j = i; j < R; j++) {
            if (i == CC && j == CC) continue;
            M[kst_pidx<R>(i, j)] = fma(g[i], g[j], M[kst_pidx<R>(i, j)]);
        }
    }
}

// Mt = g g' for the first row of a particle (saves zeroing the accumulators).
template <int R, int CC>
__device__ __forceinline__ void kst_accum_init(double *M, const double *g)
{
#pragma unroll
    for (int i = 0; i < R; i++) {
#pragma unroll
        for (int j = i; j < R; j++) {
            if (i == CC && j == CC) continue;
            M[kst_pidx<R>(i, j)] = g[i] * g[j];
        }
    }
}

// Mt (packed upper) -> unit lower L (strictly lower part used), pivots dv, 1/dv.
// Returns false at the first pivot that is not > 0 (also catches NaN).  With INV0 the reciprocal of
// the first pivot is supplied by the caller (per-particle constant when the family has a constant column).
template <int R, bool INV0 = false, bool CHECK = true>
__device__ __forceinline__ bool kst_ldlt(const double *M, double (&L)[R][R], double (&dv)[R],
                                         double (&dinv)[R], bool need_last_inv, double inv0 = 0.0)
{
    double W[R][R];  // W[i][j] = L[i][j] * dv[j]
    bool ok = true;
#pragma unroll
    for (int j = 0; j < R; j++) {
        double dj = M[kst_pidx<R>(j, j)];
#pragma unroll
        for (int k = 0; k < j; k++) dj = fma(-L[j][k], W[j][k], dj);
        dv[j] = dj;
        if (CHECK) ok = ok && (dj > 0.0);
        if (j < R - 1 || need_last_inv)
            dinv[j] = (INV0 && j == 0) ? inv0 : kst_rcp(dj);  // garbage when dj <= 0, but then ok == false
#pragma unroll
        for (int i = j + 1; i < R; i++) {
            double v = M[kst_pidx<R>(j, i)];
#pragma unroll
            for (int k = 0; k < j; k++) v = fma(-L[i][k], W[j][k], v);
            W[i][j] = v;
            L[i][j] = v * dinv[j];
        }
    }
    return ok;
}

// Determinant of Mt as ONE product plus the OR of the high words of its leading minors (or pivots), without
// any reciprocal where a closed form is cheap:
//   r = 4, first pivot a per-particle constant: Schur complement of M00 (3 x 3), then its determinant by
//          cofactors along the last row, which yields m1 = S00, m2 = S00 S11 - S01^2, m3 = det S (19 FP64);
//   r = 3: the same one level down, or the plain cofactor expansion when there is no constant column;
//   otherwise: LDL' pivots.
// Positive definite <=> all leading minors (equivalently all pivots) > 0 <=> sign bit of `signs` clear and
// prod > 0; this is the reference's "potrf! succeeded and every U[i,i] > 0" (src/design-common.jl:213-220).
// CLOSED = false forces the pivot form: it is invariant under diagonal scaling of the matrix, which the cofactor
// expansions are not -- needed for DT M^-1 DT' of a DeltaMethod transformation, whose rows carry the scales of DT
// (the column-scaled Mt of the Identity path is well scaled by construction).
template <int R, bool INV0, bool CLOSED = true>
__device__ __forceinline__ void kst_det_signs(const double *M, double inv0, double &prod, int &signs)
{
    if (CLOSED && R == 4 && INV0) {
        const double t1 = M[kst_pidx<R>(0, 1)] * inv0, t2 = M[kst_pidx<R>(0, 2)] * inv0, t3 = M[kst_pidx<R>(0, 3)] * inv0;
        const double s00 = fma(-t1, M[kst_pidx<R>(0, 1)], M[kst_pidx<R>(1, 1)]);
        const double s01 = fma(-t1, M[kst_pidx<R>(0, 2)], M[kst_pidx<R>(1, 2)]);
        const double s02 = fma(-t1, M[kst_pidx<R>(0, 3)], M[kst_pidx<R>(1, 3)]);
        const double s11 = fma(-t2, M[kst_pidx<R>(0, 2)], M[kst_pidx<R>(2, 2)]);
        const double s12 = fma(-t2, M[kst_pidx<R>(0, 3)], M[kst_pidx<R>(2, 3)]);
        const double s22 = fma(-t3, M[kst_pidx<R>(0, 3)], M[kst_pidx<R>(3, 3)]);
        const double m2 = fma(s00, s11, -(s01 * s01));
        const double c12 = fma(s00, s12, -(s01 * s02));
        const double c02 = fma(s01, s12, -(s11 * s02));
        const double m3 = fma(s22, m2, fma(s02, c02, -(s12 * c12)));
        prod = M[kst_pidx<R>(0, 0)] * m3;
        signs = __double2hiint(s00) | __double2hiint(m2) | __double2hiint(m3);
    } else if (CLOSED && R == 3 && INV0) {
        const double t1 = M[kst_pidx<R>(0, 1)] * inv0, t2 = M[kst_pidx<R>(0, 2)] * inv0;
        const double s00 = fma(-t1, M[kst_pidx<R>(0, 1)], M[kst_pidx<R>(1, 1)]);
        const double s01 = fma(-t1, M[kst_pidx<R>(0, 2)], M[kst_pidx<R>(1, 2)]);
        const double s11 = fma(-t2, M[kst_pidx<R>(0, 2)], M[kst_pidx<R>(2, 2)]);
        const double m2 = fma(s00, s11, -(s01 * s01));
        prod = M[kst_pidx<R>(0, 0)] * m2;
        signs = __double2hiint(s00) | __double2hiint(m2);
    } else if (CLOSED && R == 3) {
        const double a = M[kst_pidx<R>(0, 0)], b = M[kst_pidx<R>(0, 1)], c = M[kst_pidx<R>(0, 2)];
        const double d = M[kst_pidx<R>(1, 1)], e = M[kst_pidx<R>(1, 2)], f = M[kst_pidx<R>(2, 2)];
        const double m2 = fma(a, d, -(b * b));
        const double c12 = fma(a, e, -(b * c));
        const double c02 = fma(b, e, -(d * c));
        prod = fma(f, m2, fma(c, c02, -(e * c12)));
        signs = __double2hiint(a) | __double2hiint(m2) | __double2hiint(prod);
    } else {
        double L[R][R], dv[R], dinv[R];
        kst_ldlt<R, INV0, false>(M, L, dv, dinv, false, inv0);
        prod = dv[0];
        signs = INV0 ? 0 : __double2hiint(dv[0]);
#pragma unroll
        for (int j = 1; j < R; j++) {
            prod *= dv[j];
            signs |= __double2hiint(dv[j]);
        }
    }
}

// log det from (prod, signs) with one log call and NO floating-point compares: the sign bit of `signs` is
// OR-ed into the high word of the product and a single unsigned test on the exponent field accepts
// "positive definite and 2^-900 < det < 2^900".  Everything else takes the slow path: -Inf unless the
// matrix is positive definite (log_det!, src/design-common.jl:213-220).
__device__ __forceinline__ double kst_log_checked(double prod, int signs, KstTab tab)
{
    const unsigned key = ((unsigned)signs & 0x80000000u) | (unsigned)__double2hiint(prod);
    if (((key >> 20) - 123u) < 1800u) return kst_log(prod, tab);
    return (signs >= 0 && prod > 0.0) ? log(prod) : -KST_INF;
}

// Per-(particle, draw) criterion value from the accumulated Mt.
//
// CRIT 0 = D, 1 = A, 2 = Dexp (det^(1/t), docs/src/extend-criterion.md:114-127).  T = 0: Identity;
// T >= 1: DeltaMethod with codomain dimension T.
// aux points at this draw's auxiliary constants with stride `ld` between entries:
//   A / Identity : R doubles  s_i = 1/cs_i^2                       (tr M^-1 = sum s_i (Mt^-1)_ii)
//   Dexp / Identity : 1 double  prod_i cs_i^(2/R)                   (det M^(1/R) = det Mt^(1/R) * that)
//   DeltaMethod  : T*R doubles DTt[a][i] = DT[a][i]/cs_i, a-major   (DT M^-1 DT' = DTt Mt^-1 DTt')
// exc is set where the reference throws (PosDefException from posv!, src/design-common.jl:197;
// SingularException from potri!, src/criterion-a.jl:39): objective! then returns -Inf no matter
// the prior weight (src/designproblem.jl:217-223).  Silent failures (log_det!,
// src/design-common.jl:213-220) just return -Inf, which the caller multiplies by the weight.
template <int R, int CRIT, int T, bool INV0>
__device__ __forceinline__ double kst_criterion(const double *M, const double *aux, size_t ld, bool &exc,
                                                double inv0, KstTab tab)
{
    double L[R][R], dv[R], dinv[R];
    exc = false;
    if (CRIT != 1 && T == 0) {
        double prod;
        int signs;
        kst_det_signs<R, INV0>(M, inv0, prod, signs);
        const double ld = kst_log_checked(prod, signs, tab);
        if (CRIT == 0) return ld;
        // a draw that is not positive definite contributes exp(-Inf) = 0
        return ld == -KST_INF ? 0.0 : kst_exp(ld * (1.0 / R), tab) * aux[0];
    }
    const bool ok = kst_ldlt<R, INV0>(M, L, dv, dinv, true, inv0);
    if (!ok) {
        exc = true;
        return -KST_INF;
    }
    if (T == 0) {
        // A / Identity: -tr(M^-1) = -sum_a s_a * sum_{k>=a} (L^-1)[k][a]^2 / dv[k]
        double tr = 0.0;
#pragma unroll
        for (int a = 0; a < R; a++) {
            double y[R];
            y[a] = 1.0;
            double ta = dinv[a];
#pragma unroll
            for (int k = a + 1; k < R; k++) {
                double v = -L[k][a];
#pragma unroll
                for (int l = a + 1; l < k; l++) v = fma(-L[k][l], y[l], v);
                y[k] = v;
                ta = fma(v * v, dinv[k], ta);
            }
            tr = fma(aux[(size_t)a * ld], ta, tr);
        }
        return -tr;
    }
    // DeltaMethod: Y = L^-1 DTt' (R x T), MTinv = Y' diag(1/dv) Y  (T x T)
    constexpr int TT = T > 0 ? T : 1;
    double Y[TT][R];
#pragma unroll
    for (int a = 0; a < TT; a++) {
#pragma unroll
        for (int k = 0; k < R; k++) {
            double v = aux[(size_t)(a * R + k) * ld];
#pragma unroll
            for (int l = 0; l < k; l++) v = fma(-L[k][l], Y[a][l], v);
            Y[a][k] = v;
        }
    }
    if (CRIT == 1) {
        double tr = 0.0;
#pragma unroll
        for (int a = 0; a < TT; a++)
#pragma unroll
            for (int k = 0; k < R; k++) tr = fma(Y[a][k] * Y[a][k], dinv[k], tr);
        return -tr;
    }
    double Mi[KstPacked<TT>::N];
#pragma unroll
    for (int a = 0; a < TT; a++)
#pragma unroll
        for (int b = a; b < TT; b++) {
            double s = 0.0;
#pragma unroll
            for (int k = 0; k < R; k++) s = fma(Y[a][k] * dinv[k], Y[b][k], s);
            Mi[kst_pidx<TT>(a, b)] = s;
        }
    double prod2;
    int signs2;
    kst_det_signs<TT, false, false>(Mi, 0.0, prod2, signs2);
    // log det of the *inverse*: the sign flips unless it is -Inf (src/criterion-d.jl:35-37)
    const double ldi = kst_log_checked(prod2, signs2, tab);
    if (CRIT == 2) return ldi == -KST_INF ? 0.0 : kst_exp(-ldi * (1.0 / TT), tab);
    return ldi == -KST_INF ? ldi : -ldi;
}
