/*
 * kirstine_oracle.c -- CPU restatement of Kirstine.jl's design-objective hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  This file is the parity oracle: a plain-C, scalar
 * restatement of the reference's algorithm in the reference's own operation
 * order (upper-triangular storage, unblocked Cholesky, trsm/syrk loop nests as
 * in reference BLAS).  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load it.  The product library
 * (libkirstine_b200.so) never links or calls anything in here.
 *
 * Parity status: PINNED.  tests/test_oracle_golden.py checks this file against
 * the known-answer tests of the reference's own suite (test/design-common.jl,
 * test/criterion-d.jl, test/criterion-a.jl, test/model-nonlinear-regression.jl,
 * test/designmeasure-abstractpoint.jl, test/designregion.jl, test/pso.jl).
 * The reference itself (Julia) cannot run in this container.
 *
 * Third-party arithmetic restated here: Julia stdlib LinearAlgebra -> OpenBLAS
 * (version unpinned, no Manifest.toml in the reference): potrf!/trsm!/syrk!/
 * posv!/potri!/symm!.  We restate the reference-BLAS/LAPACK loop nests
 * (dpotf2, dtrsm LUTN, dsyrk UT, dpotrs, dtrtri+dlauum).
 *
 * All matrices are column-major like Julia's Matrix{Float64}.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define KO_MAXR 8
#define KO_MAXM 16
#define KO_MAXC (KO_MAXM + 2)
#define KO_MAXD 4

enum { KO_EMAX = 1, KO_SIGEMAX = 2, KO_TPC = 3, KO_VU = 4, KO_DTR = 5, KO_LINLOGREG = 6, KO_SINLOGREG = 7,
       KO_VUAR = 8 /* VUMod with AR(1)-correlated errors within a unit: a dense Sigma(c), mp = {sigma_squared, rho} */ };
enum { KO_CP_COPY = 0, KO_CP_EQUITIME = 1, KO_CP_DTR_COPYBOTH = 2, KO_CP_DTR_FIXEDDOSE = 3,
       KO_CP_DTR_FIXEDTIMES = 4, KO_CP_DTR_EQUIDISTANT = 5, KO_CP_DTR_LOGEQUIDISTANT = 6 };
enum { KO_CRIT_D = 0, KO_CRIT_A = 1, KO_CRIT_DEXP = 2 };
enum { KO_TRAFO_IDENTITY = 0, KO_TRAFO_DELTA = 1 };

typedef struct {
    int model_id, cp_id, crit, trafo;
    int r, t, m, d;
    long N;
    double mp[4];
    double cpp[KO_MAXM]; /* parameters of the covariate parameterization */
    double *theta; /* N x r, AoS (one parameter struct after the other) */
    double *pw;    /* N prior weights */
    double *dt;    /* N x (t x r col-major), NULL for identity */
    int has_reg;   /* normal approximation: 0 FisherMatrix, 1 RegularizedFisherMatrix */
    double reg[KO_MAXR * KO_MAXR]; /* R (r x r col-major) of docs/src/extend-approximation.md:69-79 */
} ko_problem;

#define IDX(i, j, ld) ((i) + (size_t)(j) * (ld))

/* ------------------------------------------------------------------------- */
/* user-model callbacks (reference: src/user.jl:18-98 stubs; concrete models  */
/* are test/doc code, cited per function)                                     */
/* ------------------------------------------------------------------------- */

/* map_to_covariate!: CopyDose test/example-emax.jl:31-39, CopyTime
 * test/example-compartment.jl:23-31, JustCopy src/user.jl:185-216,
 * EquiTime test/example-vector.jl:45-50, CopyBoth test/example-dtr.jl:51-55,
 * EquidistantTimes docs/src/dtr.md:484-495 */
static void ko_covariate(const ko_problem *p, const double *z, double *c)
{
    switch (p->model_id) {
    case KO_VU:
    case KO_VUAR:
        for (int j = 0; j < p->m; j++) c[j] = (double)j * z[0];
        break;
    case KO_DTR: /* docs/src/dtr.md: CopyBoth :197-203, FixedDose :331-339, FixedTimes :401-409,
                    EquidistantTimes :484-495, LogEquidistantTimes :663-681 */
        switch (p->cp_id) {
        case KO_CP_DTR_FIXEDDOSE:
            c[0] = p->cpp[0];
            for (int j = 0; j < p->m; j++) c[1 + j] = 0.0;
            c[1] = z[0];
            break;
        case KO_CP_DTR_FIXEDTIMES:
            c[0] = z[0];
            for (int j = 0; j < p->m; j++) c[1 + j] = p->cpp[j];
            break;
        case KO_CP_DTR_EQUIDISTANT:
            c[0] = z[0];
            for (int j = 0; j < p->m; j++) c[1 + j] = (double)j * z[1];
            break;
        case KO_CP_DTR_LOGEQUIDISTANT:
            c[0] = z[0];
            c[1] = 0.0;
            if (p->m > 1) c[2] = z[1];
            for (int j = 2; j < p->m; j++) c[1 + j] = p->cpp[0] * c[j];
            break;
        default: /* CopyBoth */
            c[0] = z[0];
            for (int j = 0; j < p->m; j++) c[1 + j] = 0.0;
            c[1] = z[1];
        }
        break;
    case KO_LINLOGREG: /* CopyVector, docs/src/extend-model.md:147-155 */
        for (int j = 0; j < p->d; j++) c[j] = z[j];
        break;
    default:
        c[0] = z[0];
    }
}

/* update_model_vcov!: test/example-emax.jl:25-28 (sigma_squared),
 * @simple_model src/user.jl:320-327 (sigma^2), test/example-vector.jl:20-26
 * (sigma_squared*(1+time_j)), test/example-dtr.jl:15-20 (sigma^2 I). */
static void ko_vcov(const ko_problem *p, const double *c, double *S)
{
    int m = p->m;
    for (int i = 0; i < m * m; i++) S[i] = 0.0;
    switch (p->model_id) {
    case KO_EMAX: S[0] = p->mp[0]; break;
    case KO_SIGEMAX:
    case KO_TPC: S[0] = p->mp[0] * p->mp[0]; break;
    case KO_VU:
        for (int j = 0; j < m; j++) S[IDX(j, j, m)] = p->mp[0] * (1.0 + c[j]);
        break;
    case KO_DTR:
        for (int j = 0; j < m; j++) S[IDX(j, j, m)] = p->mp[0] * p->mp[0];
        break;
    case KO_VUAR: /* not a reference model: VUMod's variances with AR(1) correlation, to exercise the general
                   * potrf! + trsm! path of src/model-nonlinear-regression.jl:39-49,79-83 */
        for (int j = 0; j < m; j++)
            for (int i = 0; i < m; i++)
                S[IDX(i, j, m)] = p->mp[0] * pow(p->mp[1], (double)(i > j ? i - j : j - i)) *
                                  sqrt((1.0 + c[i]) * (1.0 + c[j]));
        break;
    default: S[0] = 1.0; /* logistic regression has no unit covariance */
    }
}

/* ---- logistic regression, the user-defined Model supertype of docs/src/extend-model.md:59-106 ---- */
static int ko_is_logistic(const ko_problem *p)
{
    return p->model_id == KO_LINLOGREG || p->model_id == KO_SINLOGREG;
}

/* nonlinear_predictor: LinLogReg docs/src/extend-model.md:157-167, SinLogReg :344-346 */
static double ko_predictor(const ko_problem *p, const double *c, const double *th)
{
    if (p->model_id == KO_LINLOGREG) {
        double np = th[0];
        for (int j = 0; j < p->d; j++) np += th[1 + j] * c[j];
        return np;
    }
    return th[0] * sin(2 * M_PI * th[1] * (c[0] - th[2]));
}

/* predictor_jacobianmatrix!: LinLogReg docs/src/extend-model.md:169-179, SinLogReg :348-355 */
static void ko_predictor_jacobian(const ko_problem *p, const double *c, const double *th, double *jm)
{
    if (p->model_id == KO_LINLOGREG) {
        jm[0] = 1;
        for (int j = 0; j < p->d; j++) jm[1 + j] = c[j];
        return;
    }
    double sin_term = sin(2 * M_PI * th[1] * (c[0] - th[2]));
    double cos_term = cos(2 * M_PI * th[1] * (c[0] - th[2]));
    jm[0] = sin_term;
    jm[1] = th[0] * cos_term * 2 * M_PI * (c[0] - th[2]);
    jm[2] = -th[0] * cos_term * 2 * M_PI * th[1];
}

static double ko_expit(double x) { return 1 / (1 + exp(-x)); } /* docs/src/extend-model.md:86 */

/* jacobianmatrix!: Emax test/example-emax.jl:41-47; sigmoid Emax
 * docs/src/tutorial.md:127-142; TPC test/example-compartment.jl:14-21;
 * VUMod test/example-vector.jl:32-41; DTR test/example-dtr.jl:27-47.
 * jm is m x r column-major. */
void ko_jacobian(const ko_problem *p, const double *c, const double *th, double *jm)
{
    int m = p->m;
    switch (p->model_id) {
    case KO_EMAX: {
        double x = c[0], e0 = th[0], emax = th[1], ec50 = th[2];
        (void)e0;
        jm[0] = 1.0;
        jm[1] = x / (x + ec50);
        jm[2] = -emax * x / ((x + ec50) * (x + ec50));
        break;
    }
    case KO_SIGEMAX: {
        double dose = c[0], emax = th[1], ed50 = th[2], h = th[3];
        double dose_pow_h = pow(dose, h);
        double ed50_pow_h = pow(ed50, h);
        double A = dose_pow_h / (dose_pow_h + ed50_pow_h);
        double B = ed50_pow_h * emax / (dose_pow_h + ed50_pow_h);
        jm[0] = 1.0;
        jm[1] = A;
        jm[2] = -A * B * h / ed50;
        jm[3] = dose == 0 ? 0.0 : A * B * log(dose / ed50);
        break;
    }
    case KO_TPC:
    case KO_VU:
    case KO_VUAR: {
        double a = th[0], e = th[1], s = th[2];
        for (int j = 0; j < m; j++) {
            double t = c[j];
            double A = exp(-a * t);
            double E = exp(-e * t);
            jm[IDX(j, 0, m)] = A * s * t;
            jm[IDX(j, 1, m)] = -E * s * t;
            jm[IDX(j, 2, m)] = E - A;
        }
        break;
    }
    case KO_DTR: {
        double a = th[0], e = th[1], emax = th[3], ec50 = th[4];
        double dose = c[0];
        for (int j = 0; j < m; j++) {
            double t = c[1 + j];
            double A = exp(-a * t);
            double B = exp(-e * t);
            double C = B - A;
            double rd = e - a;
            double den = dose * C * a - rd * ec50;
            double den2 = den * den;
            jm[IDX(j, 0, m)] = -dose * ec50 * emax * (A * t * a * rd + e * C) / den2;
            jm[IDX(j, 1, m)] = dose * a * ec50 * emax * (B * t * rd + C) / den2;
            jm[IDX(j, 2, m)] = 1.0;
            jm[IDX(j, 3, m)] = dose * a * C / den;
            jm[IDX(j, 4, m)] = dose * a * emax * C * rd / den2;
        }
        break;
    }
    }
}

/* ------------------------------------------------------------------------- */
/* LAPACK/BLAS loop nests the reference reaches through LinearAlgebra          */
/* ------------------------------------------------------------------------- */

/* dpotf2('U'): returns 0 or the 1-based index of the failing pivot.  On
 * failure A[j,j] holds the non-positive pivot, like LAPACK. */
int ko_potrf_upper(int n, double *A)
{
    for (int j = 0; j < n; j++) {
        double ajj = A[IDX(j, j, n)];
        for (int k = 0; k < j; k++) ajj -= A[IDX(k, j, n)] * A[IDX(k, j, n)];
        if (!(ajj > 0.0)) {
            A[IDX(j, j, n)] = ajj;
            return j + 1;
        }
        ajj = sqrt(ajj);
        A[IDX(j, j, n)] = ajj;
        for (int i = j + 1; i < n; i++) {
            double s = A[IDX(j, i, n)];
            for (int k = 0; k < j; k++) s -= A[IDX(k, j, n)] * A[IDX(k, i, n)];
            A[IDX(j, i, n)] = s / ajj;
        }
    }
    return 0;
}

/* dtrsm('L','U','T','N', 1.0, U, B): B (m x nrhs) <- U^-T B */
static void ko_trsm_lutn(int m, int nrhs, const double *U, double *B)
{
    for (int j = 0; j < nrhs; j++)
        for (int i = 0; i < m; i++) {
            double temp = B[IDX(i, j, m)];
            for (int k = 0; k < i; k++) temp -= U[IDX(k, i, m)] * B[IDX(k, j, m)];
            B[IDX(i, j, m)] = temp / U[IDX(i, i, m)];
        }
}

/* dtrsm('L','U','N','N', 1.0, U, B): B <- U^-1 B */
static void ko_trsm_lunn(int m, int nrhs, const double *U, double *B)
{
    for (int j = 0; j < nrhs; j++)
        for (int k = m - 1; k >= 0; k--) {
            if (B[IDX(k, j, m)] != 0.0) {
                B[IDX(k, j, m)] /= U[IDX(k, k, m)];
                for (int i = 0; i < k; i++)
                    B[IDX(i, j, m)] -= B[IDX(k, j, m)] * U[IDX(i, k, m)];
            }
        }
}

/* dsyrk('U','T', alpha, A (k x n), 1.0, C (n x n)): C_upper += alpha A'A */
static void ko_syrk_ut(int n, int k, double alpha, const double *A, double *C)
{
    for (int j = 0; j < n; j++)
        for (int i = 0; i <= j; i++) {
            double temp = 0.0;
            for (int l = 0; l < k; l++) temp += A[IDX(l, i, k)] * A[IDX(l, j, k)];
            C[IDX(i, j, n)] = alpha * temp + C[IDX(i, j, n)];
        }
}

/* potri!('U'): inverse from the Cholesky factor, upper triangle of the result.
 * Returns nonzero (SingularException in Julia) when a diagonal entry is exactly 0. */
static int ko_potri_upper(int n, double *A)
{
    /* dtrti2 upper non-unit */
    for (int j = 0; j < n; j++)
        if (A[IDX(j, j, n)] == 0.0) return j + 1;
    for (int j = 0; j < n; j++) {
        A[IDX(j, j, n)] = 1.0 / A[IDX(j, j, n)];
        double ajj = -A[IDX(j, j, n)];
        /* x <- T(0:j,0:j) * A(0:j, j)   (dtrmv upper, no-trans, non-unit) */
        for (int k = 0; k < j; k++) {
            double temp = A[IDX(k, j, n)];
            if (temp != 0.0) {
                for (int i = 0; i < k; i++) A[IDX(i, j, n)] += temp * A[IDX(i, k, n)];
                A[IDX(k, j, n)] = temp * A[IDX(k, k, n)];
            }
        }
        for (int i = 0; i < j; i++) A[IDX(i, j, n)] *= ajj;
    }
    /* dlauu2 upper: A <- U U' */
    for (int i = 0; i < n; i++) {
        double aii = A[IDX(i, i, n)];
        if (i < n - 1) {
            double dot = 0.0;
            for (int k = i; k < n; k++) dot += A[IDX(i, k, n)] * A[IDX(i, k, n)];
            /* gemv: A(0:i, i) = aii*A(0:i,i) + A(0:i,i+1:n) * A(i,i+1:n)' */
            for (int l = 0; l < i; l++) {
                double s = 0.0;
                for (int k = i + 1; k < n; k++) s += A[IDX(l, k, n)] * A[IDX(i, k, n)];
                A[IDX(l, i, n)] = aii * A[IDX(l, i, n)] + s;
            }
            A[IDX(i, i, n)] = dot;
        } else {
            for (int l = 0; l <= i; l++) A[IDX(l, i, n)] *= aii;
        }
    }
    return 0;
}

/* inv(Symmetric(M)) as used in gateaux_constants (src/criterion-d.jl:52,69;
 * src/criterion-a.jl:56,72).  Julia uses Bunch-Kaufman; M is symmetric positive
 * definite on every non-singular design, so we invert through Cholesky and fall
 * back to reporting "singular" when that fails (Julia: SingularException).
 * In: upper triangle of M (n x n).  Out: full symmetric inverse. */
static int ko_inv_sym(int n, const double *M, double *out)
{
    double U[KO_MAXR * KO_MAXR];
    memcpy(U, M, sizeof(double) * n * n);
    if (ko_potrf_upper(n, U)) return 1;
    if (ko_potri_upper(n, U)) return 1;
    for (int j = 0; j < n; j++)
        for (int i = 0; i <= j; i++) {
            out[IDX(i, j, n)] = U[IDX(i, j, n)];
            out[IDX(j, i, n)] = U[IDX(i, j, n)];
        }
    return 0;
}

/* log_det! (src/design-common.jl:213-220) */
double ko_log_det(int n, double *A)
{
    ko_potrf_upper(n, A); /* info ignored, like potrf! in Julia */
    double acc = 0;
    for (int i = 0; i < n; i++) acc += A[IDX(i, i, n)] > 0 ? log(A[IDX(i, i, n)]) : -INFINITY;
    return 2 * acc;
}

/* tr_prod(A, B, :U) (src/design-common.jl:224-248) */
double ko_tr_prod_upper(int k, const double *A, const double *B)
{
    double acc = 0.0;
    for (int i = 0; i < k; i++) {
        acc += A[IDX(i, i, k)] * B[IDX(i, i, k)];
        for (int j = i + 1; j < k; j++) acc += 2 * A[IDX(i, j, k)] * B[IDX(i, j, k)];
    }
    return acc;
}

/* criterion_functional! D: src/criterion-d.jl:31-39; A: src/criterion-a.jl:31-45.
 * *exc is set when Julia would throw SingularException (A, not inverted).
 * Documented deviation: when potrf! fails with a strictly negative pivot the
 * reference hands the half-factored matrix to potri! and returns a finite
 * garbage value; the true M is PSD, so this only happens by rounding on a
 * singular design.  Oracle and engine both treat any non-positive pivot as
 * singular (-Inf), the same rule log_det! applies for the D criterion. */
double ko_criterion_functional(int n, double *tnim, int is_inv, int crit, int *exc)
{
    *exc = 0;
    if (crit == KO_CRIT_D || crit == KO_CRIT_DEXP) {
        double sgn = is_inv ? -1 : 1;
        double ld = ko_log_det(n, tnim);
        double cv = ld != -INFINITY ? sgn * ld : ld;
        /* DexpCriterion, det^(1/t): docs/src/extend-criterion.md:114-127 (a singular draw contributes 0) */
        return crit == KO_CRIT_D ? cv : exp(cv / n);
    }
    if (!is_inv) {
        if (ko_potrf_upper(n, tnim)) { *exc = 1; return -INFINITY; }
        if (ko_potri_upper(n, tnim)) { *exc = 1; return -INFINITY; }
    }
    double tr = 0.0;
    for (int i = 0; i < n; i++) tr += tnim[IDX(i, i, n)];
    return -tr;
}

/* apply_transformation! (src/design-common.jl:157-205).
 * rxr: upper triangle of M (or its inverse when r_is_inv); overwritten.
 * jm: t x r; out txt; returns 1 for PosDefException. *t_is_inv reports the flag. */
int ko_apply_transformation(int r, int t, int trafo, double *rxr, int r_is_inv,
                            const double *jm, double *txt, int *t_is_inv)
{
    if (trafo == KO_TRAFO_IDENTITY) {
        memcpy(txt, rxr, sizeof(double) * r * r);
        *t_is_inv = r_is_inv;
        return 0;
    }
    if (r_is_inv) {
        /* symm!('R','U'): t_x_r = jm * Symmetric(rxr); then t_x_t = t_x_r * jm' */
        double txr[KO_MAXR * KO_MAXR];
        for (int j = 0; j < r; j++)
            for (int i = 0; i < t; i++) {
                double s = 0.0;
                for (int k = 0; k < r; k++) {
                    double a = k <= j ? rxr[IDX(k, j, r)] : rxr[IDX(j, k, r)];
                    s += jm[IDX(i, k, t)] * a;
                }
                txr[IDX(i, j, t)] = s;
            }
        for (int j = 0; j < t; j++)
            for (int i = 0; i < t; i++) {
                double s = 0.0;
                for (int k = 0; k < r; k++) s += txr[IDX(i, k, t)] * jm[IDX(j, k, t)];
                txt[IDX(i, j, t)] = s;
            }
    } else {
        double rxt[KO_MAXR * KO_MAXR];
        for (int j = 0; j < t; j++)
            for (int i = 0; i < r; i++) rxt[IDX(i, j, r)] = jm[IDX(j, i, t)];
        /* posv!('U', rxr, rxt) */
        if (ko_potrf_upper(r, rxr)) return 1;
        ko_trsm_lutn(r, t, rxr, rxt);
        ko_trsm_lunn(r, t, rxr, rxt);
        for (int j = 0; j < t; j++)
            for (int i = 0; i < t; i++) {
                double s = 0.0;
                for (int k = 0; k < r; k++) s += jm[IDX(i, k, t)] * rxt[IDX(k, j, r)];
                txt[IDX(i, j, t)] = s;
            }
    }
    *t_is_inv = 1;
    return 0;
}

/* ------------------------------------------------------------------------- */
/* evaluation core                                                            */
/* ------------------------------------------------------------------------- */

typedef struct {
    double c[16 * KO_MAXC];               /* covariates of up to 16 design points */
    double chol[16 * KO_MAXM * KO_MAXM];  /* chol(Sigma_k) */
    double mxr[KO_MAXM * KO_MAXR];
    double rxr[KO_MAXR * KO_MAXR];
    double txt[KO_MAXR * KO_MAXR];
} ko_ws;
#define KO_MAXK 16

/* update_covariates! (src/design-common.jl:51-61) +
 * update_model_workspace! (src/model-nonlinear-regression.jl:39-49) */
static void ko_update_workspace(const ko_problem *p, int K, const double *pts, ko_ws *w)
{
    for (int k = 0; k < K; k++) {
        ko_covariate(p, pts + (size_t)k * p->d, w->c + k * KO_MAXC);
        double *S = w->chol + k * KO_MAXM * KO_MAXM;
        ko_vcov(p, w->c + k * KO_MAXC, S);
        ko_potrf_upper(p->m, S);
    }
}

/* average_fishermatrix! (src/model-nonlinear-regression.jl:60-86), followed by
 * informationmatrix!(afm, na) (src/design-common.jl:74-76 for FisherMatrix: identity;
 * docs/src/extend-approximation.md:69-79 for RegularizedFisherMatrix: afm .+= R).
 * Every caller in the reference applies the two back to back (src/design-common.jl:109-110,
 * src/designproblem.jl:206-207,253-254). */
static void ko_average_fishermatrix(const ko_problem *p, int K, const double *wts,
                                    const double *th, ko_ws *w)
{
    int r = p->r, m = p->m;
    for (int i = 0; i < r * r; i++) w->rxr[i] = 0.0;
    if (ko_is_logistic(p)) { /* average_fishermatrix! of docs/src/extend-model.md:88-105 */
        for (int k = 0; k < K; k++) {
            ko_predictor_jacobian(p, w->c + k * KO_MAXC, th, w->mxr);
            double prob = ko_expit(ko_predictor(p, w->c + k * KO_MAXC, th));
            double p1mp = prob * (1 - prob);
            ko_syrk_ut(r, 1, p1mp * wts[k], w->mxr, w->rxr);
        }
    } else {
        for (int k = 0; k < K; k++) {
            ko_jacobian(p, w->c + k * KO_MAXC, th, w->mxr);
            ko_trsm_lutn(m, r, w->chol + k * KO_MAXM * KO_MAXM, w->mxr);
            ko_syrk_ut(r, m, wts[k], w->mxr, w->rxr);
        }
    }
    if (p->has_reg)
        for (int i = 0; i < r * r; i++) w->rxr[i] += p->reg[i];
}

/* informationmatrix (src/design-common.jl:96-112): full Symmetric copy */
int ko_informationmatrix(const ko_problem *p, int K, const double *pts, const double *wts,
                         long draw, double *out)
{
    if (K > KO_MAXK) return -1;
    ko_ws w;
    ko_update_workspace(p, K, pts, &w);
    ko_average_fishermatrix(p, K, wts, p->theta + (size_t)draw * p->r, &w);
    int r = p->r;
    for (int j = 0; j < r; j++)
        for (int i = 0; i <= j; i++) {
            out[IDX(i, j, r)] = w.rxr[IDX(i, j, r)];
            out[IDX(j, i, r)] = w.rxr[IDX(i, j, r)];
        }
    return 0;
}

/* objective! (src/designproblem.jl:192-224) over draws [j0, j1) */
static double ko_objective_ws(const ko_problem *p, int K, const double *pts,
                              const double *wts, ko_ws *w, long j0, long j1)
{
    ko_update_workspace(p, K, pts, w);
    double acc = 0;
    for (long j = j0; j < j1; j++) {
        ko_average_fishermatrix(p, K, wts, p->theta + (size_t)j * p->r, w);
        int t_is_inv = 0, exc = 0;
        const double *jm = p->dt ? p->dt + (size_t)j * p->t * p->r : NULL;
        if (ko_apply_transformation(p->r, p->t, p->trafo, w->rxr, 0, jm, w->txt, &t_is_inv))
            return -INFINITY; /* PosDefException */
        double cv = ko_criterion_functional(p->t, w->txt, t_is_inv, p->crit, &exc);
        if (exc) return -INFINITY; /* SingularException */
        acc += p->pw[j] * cv;
    }
    return acc;
}

double ko_objective(const ko_problem *p, int K, const double *pts, const double *wts)
{
    if (K > KO_MAXK) return NAN;
    ko_ws w;
    return ko_objective_ws(p, K, pts, wts, &w, 0, p->N);
}

/* partial objective over a shard of the prior sample (multi-GPU test oracle) */
double ko_objective_range(const ko_problem *p, int K, const double *pts, const double *wts,
                          long j0, long j1)
{
    if (K > KO_MAXK) return NAN;
    ko_ws w;
    return ko_objective_ws(p, K, pts, wts, &w, j0, j1);
}

/* pso_evaluate_objective! (src/pso.jl:173-179) for a whole swarm; OpenMP over
 * particles with one workspace per thread (the reference is serial; this is
 * the "multithreaded CPU" stand-in) */
void ko_objective_batch(const ko_problem *p, int S, int K, const double *pts,
                        const double *wts, double *fx, int nthreads)
{
#ifdef _OPENMP
    if (nthreads < 1) nthreads = 1;
#pragma omp parallel for schedule(dynamic, 1) num_threads(nthreads)
#endif
    for (int s = 0; s < S; s++) {
        ko_ws w;
        fx[s] = K > KO_MAXK ? NAN
                            : ko_objective_ws(p, K, pts + (size_t)s * K * p->d,
                                              wts + (size_t)s * K, &w, 0, p->N);
    }
}

/* gateaux_constants, four variants (src/criterion-d.jl:43-87,
 * src/criterion-a.jl:47-77).  A: N x (r x r) dense, trB: N.  Returns 1 when
 * Julia would throw SingularException. */
int ko_gateaux_constants(const ko_problem *p, int K, const double *pts, const double *wts,
                         double *A, double *trB)
{
    int r = p->r, t = p->t;
    if (K > KO_MAXK) return -1;
    ko_ws w;
    ko_update_workspace(p, K, pts, &w);
    for (long j = 0; j < p->N; j++) {
        ko_average_fishermatrix(p, K, wts, p->theta + (size_t)j * r, &w);
        double iM[KO_MAXR * KO_MAXR];
        if (ko_inv_sym(r, w.rxr, iM)) return 1;
        double *Aj = A + (size_t)j * r * r;
        const double *DT = p->dt ? p->dt + (size_t)j * t * r : NULL;
        if (p->trafo == KO_TRAFO_IDENTITY && p->crit == KO_CRIT_DEXP) {
            /* docs/src/extend-criterion.md:129-150: trB = exp(-log(det(iM))/r), A = iM trB / r */
            double tmp[KO_MAXR * KO_MAXR];
            memcpy(tmp, iM, sizeof(double) * r * r);
            trB[j] = exp(-ko_log_det(r, tmp) / r);
            for (int i = 0; i < r * r; i++) Aj[i] = iM[i] * trB[j] / r;
        } else if (p->trafo == KO_TRAFO_IDENTITY && p->crit == KO_CRIT_D) {
            memcpy(Aj, iM, sizeof(double) * r * r);
            trB[j] = r;
        } else if (p->trafo == KO_TRAFO_IDENTITY && p->crit == KO_CRIT_A) {
            double tr = 0.0;
            for (int i = 0; i < r; i++) tr += iM[IDX(i, i, r)];
            trB[j] = tr;
            for (int c = 0; c < r; c++)
                for (int i = 0; i < r; i++) {
                    double s = 0.0;
                    for (int k = 0; k < r; k++) s += iM[IDX(i, k, r)] * iM[IDX(k, c, r)];
                    Aj[IDX(i, c, r)] = s;
                }
        } else if (p->crit == KO_CRIT_D || p->crit == KO_CRIT_DEXP) {
            /* A = iM * DT' * inv(DT iM DT') * DT * iM ; trB = t.
             * DexpCriterion (docs/src/extend-criterion.md:40-52,152-182): the same A times trB/t with
             * trB = det(M_T)^(1/t).  The listing there takes the determinant of the r x r inverse instead of
             * the t x t one, against its own formula; we follow the formula (the finite-difference test in
             * tests/test_oracle_golden.py shows it is the derivative of the objective). */
            double MTinv[KO_MAXR * KO_MAXR], tmp[KO_MAXR * KO_MAXR];
            int flag;
            memcpy(tmp, iM, sizeof(double) * r * r);
            ko_apply_transformation(r, t, KO_TRAFO_DELTA, tmp, 1, DT, MTinv, &flag);
            double MT[KO_MAXR * KO_MAXR];
            if (ko_inv_sym(t, MTinv, MT)) return 1;
            /* C = DT' * MT * DT (r x r) */
            double MTDT[KO_MAXR * KO_MAXR], C[KO_MAXR * KO_MAXR], iMC[KO_MAXR * KO_MAXR];
            for (int c = 0; c < r; c++)
                for (int i = 0; i < t; i++) {
                    double s = 0.0;
                    for (int k = 0; k < t; k++) s += MT[IDX(i, k, t)] * DT[IDX(k, c, t)];
                    MTDT[IDX(i, c, t)] = s;
                }
            for (int c = 0; c < r; c++)
                for (int i = 0; i < r; i++) {
                    double s = 0.0;
                    for (int k = 0; k < t; k++) s += DT[IDX(k, i, t)] * MTDT[IDX(k, c, t)];
                    C[IDX(i, c, r)] = s;
                }
            for (int c = 0; c < r; c++)
                for (int i = 0; i < r; i++) {
                    double s = 0.0;
                    for (int k = 0; k < r; k++) s += iM[IDX(i, k, r)] * C[IDX(k, c, r)];
                    iMC[IDX(i, c, r)] = s;
                }
            for (int c = 0; c < r; c++)
                for (int i = 0; i < r; i++) {
                    double s = 0.0;
                    for (int k = 0; k < r; k++) s += iMC[IDX(i, k, r)] * iM[IDX(k, c, r)];
                    Aj[IDX(i, c, r)] = s;
                }
            trB[j] = t;
            if (p->crit == KO_CRIT_DEXP) {
                double tmp[KO_MAXR * KO_MAXR];
                memcpy(tmp, MTinv, sizeof(double) * t * t);
                trB[j] = exp(-ko_log_det(t, tmp) / t);
                for (int i = 0; i < r * r; i++) Aj[i] *= trB[j] / t;
            }
        } else {
            /* A criterion, DeltaMethod: JpJ = DT'DT; trB = tr(JpJ iM); A = iM JpJ iM */
            double JpJ[KO_MAXR * KO_MAXR], T1[KO_MAXR * KO_MAXR];
            for (int c = 0; c < r; c++)
                for (int i = 0; i < r; i++) {
                    double s = 0.0;
                    for (int k = 0; k < t; k++) s += DT[IDX(k, i, t)] * DT[IDX(k, c, t)];
                    JpJ[IDX(i, c, r)] = s;
                }
            double tr = 0.0;
            for (int i = 0; i < r; i++)
                for (int k = 0; k < r; k++) tr += JpJ[IDX(i, k, r)] * iM[IDX(k, i, r)];
            trB[j] = tr;
            for (int c = 0; c < r; c++)
                for (int i = 0; i < r; i++) {
                    double s = 0.0;
                    for (int k = 0; k < r; k++) s += iM[IDX(i, k, r)] * JpJ[IDX(k, c, r)];
                    T1[IDX(i, c, r)] = s;
                }
            for (int c = 0; c < r; c++)
                for (int i = 0; i < r; i++) {
                    double s = 0.0;
                    for (int k = 0; k < r; k++) s += T1[IDX(i, k, r)] * iM[IDX(k, c, r)];
                    Aj[IDX(i, c, r)] = s;
                }
        }
    }
    return 0;
}

/* gateauxderivative! (src/designproblem.jl:239-258) for one one-point direction,
 * over draws [j0, j1) */
static double ko_gateaux_ws(const ko_problem *p, const double *z, const double *A,
                            const double *trB, ko_ws *w, long j0, long j1)
{
    double one = 1.0;
    ko_update_workspace(p, 1, z, w);
    double acc = 0;
    int r = p->r;
    for (long j = j0; j < j1; j++) {
        ko_average_fishermatrix(p, 1, &one, p->theta + (size_t)j * r, w);
        acc += p->pw[j] * (ko_tr_prod_upper(r, A + (size_t)j * r * r, w->rxr) - trB[j]);
    }
    return acc;
}

void ko_gateaux_batch(const ko_problem *p, long G, const double *dirs, const double *A,
                      const double *trB, double *out, int nthreads)
{
#ifdef _OPENMP
    if (nthreads < 1) nthreads = 1;
#pragma omp parallel for schedule(static) num_threads(nthreads)
#endif
    for (long g = 0; g < G; g++) {
        ko_ws w;
        out[g] = ko_gateaux_ws(p, dirs + (size_t)g * p->d, A, trB, &w, 0, p->N);
    }
}

/* ------------------------------------------------------------------------- */
/* problem handle                                                             */
/* ------------------------------------------------------------------------- */

ko_problem *ko_create(int model_id, const double *mp, int nmp, int cp_id, const double *cpp, int ncpp,
                      int crit, int trafo, int r, int t, int m, int d, long N, const double *theta,
                      const double *pw, const double *dt)
{
    if (r > KO_MAXR || t > KO_MAXR || m > KO_MAXM || d > KO_MAXD) return NULL;
    ko_problem *p = (ko_problem *)calloc(1, sizeof(ko_problem));
    p->model_id = model_id; p->cp_id = cp_id; p->crit = crit; p->trafo = trafo;
    p->r = r; p->t = t; p->m = m; p->d = d; p->N = N;
    for (int i = 0; i < nmp && i < 4; i++) p->mp[i] = mp[i];
    for (int i = 0; i < ncpp && i < KO_MAXM; i++) p->cpp[i] = cpp[i];
    p->theta = (double *)malloc(sizeof(double) * N * r);
    memcpy(p->theta, theta, sizeof(double) * N * r);
    p->pw = (double *)malloc(sizeof(double) * N);
    memcpy(p->pw, pw, sizeof(double) * N);
    if (trafo == KO_TRAFO_DELTA && dt) {
        p->dt = (double *)malloc(sizeof(double) * N * t * r);
        memcpy(p->dt, dt, sizeof(double) * N * t * r);
    }
    return p;
}

/* normal_approximation = RegularizedFisherMatrix(R); R is r x r column-major, NULL switches back to FisherMatrix */
void ko_set_regularization(ko_problem *p, const double *R)
{
    p->has_reg = R != NULL;
    if (R) memcpy(p->reg, R, sizeof(double) * p->r * p->r);
}

void ko_destroy(ko_problem *p)
{
    if (!p) return;
    free(p->theta); free(p->pw); free(p->dt); free(p);
}

/* ------------------------------------------------------------------------- */
/* design-region / abstract-point geometry used by the swarm move             */
/* ------------------------------------------------------------------------- */

/* move_designpoint_how_far (src/designregion.jl:114-126) */
double ko_move_designpoint_how_far(int d, const double *dp, const double *v,
                                   const double *lb, const double *ub)
{
    double t = 1.0;
    for (int j = 0; j < d; j++) {
        double sl = dp[j] + t * v[j] > ub[j] ? (ub[j] - dp[j]) / v[j] : t;
        double sr = dp[j] + t * v[j] < lb[j] ? (lb[j] - dp[j]) / v[j] : t;
        t = fmin(sl, sr);
    }
    return t;
}

/* move_handle_fixed! (src/designmeasure-abstractpoint.jl:184-223) */
void ko_move_handle_fixed(int K, int d, double *v_atoms, double *v_w, const int *fixw,
                          const int *fixp)
{
    double sum_vw_free = 0.0;
    int n_fixw = 0;
    for (int k = 0; k < K; k++) n_fixw += fixw[k] != 0;
    for (int k = 0; k < K - 1; k++) {
        if (fixw[k]) v_w[k] = 0.0;
        else sum_vw_free += v_w[k];
    }
    if (n_fixw != K && fixw[K - 1]) {
        double mean_free = sum_vw_free / (K - n_fixw);
        for (int k = 0; k < K - 1; k++)
            if (!fixw[k]) v_w[k] -= mean_free;
    }
    for (int k = 0; k < K; k++)
        if (fixp[k])
            for (int j = 0; j < d; j++) v_atoms[k * d + j] = 0.0;
}

/* move_how_far (src/designmeasure-abstractpoint.jl:225-248) */
double ko_move_how_far(int K, int d, const double *pts, const double *w, const double *v_atoms,
                       const double *v_w, const double *lb, const double *ub)
{
    double t = 1.0;
    for (int k = 0; k < K; k++)
        t = fmin(t, ko_move_designpoint_how_far(d, pts + k * d, v_atoms + k * d, lb, ub));
    for (int k = 0; k < K - 1; k++) {
        double s = w[k] + t * v_w[k] < 0 ? -w[k] / v_w[k] : t;
        t = fmin(s, t);
    }
    double sum_x = 1.0 - w[K - 1];
    double sum_v = 0.0;
    for (int k = 0; k < K - 1; k++) sum_v += v_w[k];
    double s = sum_x + t * sum_v > 1 ? (1 - sum_x) / sum_v : t;
    t = fmin(s, t);
    return t;
}

/* move_add_v! (src/designmeasure-abstractpoint.jl:250-285) with
 * move_designpoint! (src/designregion.jl:129-135) */
void ko_move_add_v(int K, int d, double *pts, double *w, double t, const double *v_atoms,
                   const double *v_w, const double *lb, const double *ub, const int *fixw)
{
    for (int k = 0; k < K; k++)
        for (int j = 0; j < d; j++) {
            double x = pts[k * d + j] + t * v_atoms[k * d + j];
            pts[k * d + j] = fmin(fmax(x, lb[j]), ub[j]);
        }
    for (int k = 0; k < K; k++) w[k] += t * v_w[k];
    double weight_K = 1.0;
    for (int k = 0; k < K - 1; k++) {
        w[k] = fmin(fmax(w[k], 0.0), 1.0);
        weight_K -= w[k];
    }
    if (!fixw[K - 1]) w[K - 1] = weight_K;
    w[K - 1] = fmax(w[K - 1], 0.0);
}

/* ap_add!(p, v, c) (src/designmeasure-abstractpoint.jl:164-182); returns t */
double ko_ap_add(int K, int d, double *pts, double *w, double *v_atoms, double *v_w,
                 const double *lb, const double *ub, const int *fixw, const int *fixp)
{
    ko_move_handle_fixed(K, d, v_atoms, v_w, fixw, fixp);
    double t = ko_move_how_far(K, d, pts, w, v_atoms, v_w, lb, ub);
    if (t < 1e-8) t = 0.0;
    ko_move_add_v(K, d, pts, w, t, v_atoms, v_w, lb, ub, fixw);
    if (t != 1.0) {
        for (int i = 0; i < K * d; i++) v_atoms[i] *= 0;
        for (int k = 0; k < K; k++) v_w[k] *= 0;
    }
    return t;
}

/* ------------------------------------------------------------------------- */
/* counter-based RNG shared with the engine: Philox4x32-10 (Salmon et al.,     */
/* SC'11).  The reference uses Julia's task-local Xoshiro256++, whose stream   */
/* cannot be reproduced outside Julia; PSO parity with the reference is        */
/* therefore statistical, while oracle<->engine parity is exact in */
/* the random inputs because both draw from this generator.                    */
/* ------------------------------------------------------------------------- */
void ko_philox4x32(uint64_t seed, uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                   uint32_t out[4])
{
    uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
    for (int round = 0; round < 10; round++) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c0;
        uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

/* two uniforms in [0,1) with 53 random bits each */
static void ko_uniform2(uint64_t seed, uint32_t iter, uint32_t particle, uint32_t comp,
                        uint32_t stream, double *u0, double *u1)
{
    uint32_t o[4];
    ko_philox4x32(seed, iter, particle, comp, stream, o);
    uint64_t a = ((uint64_t)o[1] << 32) | o[0];
    uint64_t b = ((uint64_t)o[3] << 32) | o[2];
    *u0 = (double)(a >> 11) * 0x1.0p-53;
    *u1 = (double)(b >> 11) * 0x1.0p-53;
}

void ko_uniform2_export(uint64_t seed, uint32_t iter, uint32_t particle, uint32_t comp,
                        uint32_t stream, double *out2)
{
    ko_uniform2(seed, iter, particle, comp, stream, out2, out2 + 1);
}

/* ------------------------------------------------------------------------- */
/* PSO (src/pso.jl, src/optimize.jl:42-70)                                     */
/* ------------------------------------------------------------------------- */

typedef struct {
    int S, K, d, mode; /* mode 0: objective, 1: gateaux derivative (K must be 1) */
    double c1, c2;
    uint64_t seed;
    int iter;          /* iterations done so far (initial state = 1) */
    long n_eval;
    double *x_pts, *x_w, *p_pts, *p_w, *v_pts, *v_w, *g_pts, *g_w;
    double *fx, *fp, fg;
    double lb[KO_MAXD], ub[KO_MAXD];
    int fixw[KO_MAXK], fixp[KO_MAXK];
    const ko_problem *prob;
    const double *gA, *gtrB; /* gateaux constants for mode 1 */
    int nthreads;
} ko_pso;

static void ko_pso_evaluate(ko_pso *s)
{
    if (s->mode == 0)
        ko_objective_batch(s->prob, s->S, s->K, s->x_pts, s->x_w, s->fx, s->nthreads);
    else
        ko_gateaux_batch(s->prob, s->S, s->x_pts, s->gA, s->gtrB, s->fx, s->nthreads);
    s->n_eval += s->S;
}

/* pso_update_best! (src/pso.jl:181-193) */
static void ko_pso_update_best(ko_pso *s)
{
    int K = s->K, d = s->d;
    for (int i = 0; i < s->S; i++) {
        if (s->fx[i] > s->fp[i]) {
            memcpy(s->p_pts + (size_t)i * K * d, s->x_pts + (size_t)i * K * d, sizeof(double) * K * d);
            memcpy(s->p_w + (size_t)i * K, s->x_w + (size_t)i * K, sizeof(double) * K);
            s->fp[i] = s->fx[i];
            if (s->fx[i] > s->fg) {
                memcpy(s->g_pts, s->x_pts + (size_t)i * K * d, sizeof(double) * K * d);
                memcpy(s->g_w, s->x_w + (size_t)i * K, sizeof(double) * K);
                s->fg = s->fx[i];
            }
        }
    }
}

/* optimizer_state (src/pso.jl:83-125) with ap_rand!(d, constraints)
 * (src/designmeasure-abstractpoint.jl:73-107) and random_designpoint!
 * (src/designregion.jl:104-110).  Random numbers: component c of particle i
 * draws Philox(seed; iter=0, i, c, stream=1+attempt); points use u0 of
 * component K + k*d + j, weights use -log(1-u0) of component k.
 * Returns 0 ok, 1 "infinite objective value at best initial point",
 * 2 "objective evaluates to NaN on all initial points". */
int ko_pso_init(ko_pso *s, const double *proto_pts, const double *proto_w, int P)
{
    int S = s->S, K = s->K, d = s->d;
    for (int i = 0; i < S; i++) {
        const double *sp = proto_pts + (size_t)(i < P ? i : 0) * K * d;
        const double *sw = proto_w + (size_t)(i < P ? i : 0) * K;
        double *xp = s->x_pts + (size_t)i * K * d, *xw = s->x_w + (size_t)i * K;
        memcpy(xp, sp, sizeof(double) * K * d);
        memcpy(xw, sw, sizeof(double) * K);
        if (i >= P) {
            for (int k = 0; k < K; k++)
                if (!s->fixp[k])
                    for (int j = 0; j < d; j++) {
                        double u0, u1;
                        ko_uniform2(s->seed, 0, i, K + k * d + j, 1, &u0, &u1);
                        double v = u0;
                        v *= s->ub[j] - s->lb[j];
                        v += s->lb[j];
                        xp[k * d + j] = v;
                    }
            int all_fixed = 1;
            for (int k = 0; k < K; k++) all_fixed &= s->fixw[k] != 0;
            if (!all_fixed) {
                double fix = 0.0;
                for (int k = 0; k < K; k++)
                    if (s->fixw[k]) fix += xw[k];
                double cum_sum_fix = fmin(1.0, fix);
                double cum_sum_rand = 0.0;
                int attempt = 0;
                while (cum_sum_rand < 0x1.0p-52) {
                    for (int k = 0; k < K; k++)
                        if (!s->fixw[k]) {
                            double u0, u1;
                            ko_uniform2(s->seed, 0, i, k, 1 + attempt, &u0, &u1);
                            xw[k] = -log(1.0 - u0);
                            cum_sum_rand += xw[k];
                        }
                    attempt++;
                }
                double norm_const = (1 - cum_sum_fix) / cum_sum_rand;
                for (int k = 0; k < K; k++)
                    if (!s->fixw[k]) xw[k] *= norm_const;
            }
        }
    }
    memset(s->v_pts, 0, sizeof(double) * S * K * d);
    memset(s->v_w, 0, sizeof(double) * S * K);
    memcpy(s->p_pts, s->x_pts, sizeof(double) * S * K * d);
    memcpy(s->p_w, s->x_w, sizeof(double) * S * K);
    memcpy(s->g_pts, s->x_pts, sizeof(double) * K * d);
    memcpy(s->g_w, s->x_w, sizeof(double) * K);
    s->n_eval = 0;
    ko_pso_evaluate(s);
    for (int i = 0; i < S; i++) s->fp[i] = -INFINITY;
    s->fg = -INFINITY;
    ko_pso_update_best(s);
    s->iter = 1;
    if (s->fg == INFINITY) return 1;
    int all_nan = 1;
    for (int i = 0; i < S; i++) all_nan &= isnan(s->fx[i]) != 0;
    if (all_nan) return 2;
    return 0;
}

/* tick! (src/pso.jl:127-141): velocity (:143-164), position (:166-171),
 * evaluate (:173-179), best (:181-193).  Component order of the random
 * vectors follows ap_rand!(v, lb, ub): weights 0..K-1, then atoms. */
void ko_pso_tick(ko_pso *s)
{
    int S = s->S, K = s->K, d = s->d;
    double phi = s->c1 + s->c2;
    double chi = 2 / fabs(2 - phi - sqrt(phi * phi - 4 * phi));
    for (int i = 0; i < S; i++) {
        double *xp = s->x_pts + (size_t)i * K * d, *xw = s->x_w + (size_t)i * K;
        double *pp = s->p_pts + (size_t)i * K * d, *pw = s->p_w + (size_t)i * K;
        double *vp = s->v_pts + (size_t)i * K * d, *vw = s->v_w + (size_t)i * K;
        for (int c = 0; c < K * (d + 1); c++) {
            double r1, r2;
            ko_uniform2(s->seed, (uint32_t)s->iter, i, c, 0, &r1, &r2);
            double x = c < K ? xw[c] : xp[c - K];
            double pb = c < K ? pw[c] : pp[c - K];
            double gb = c < K ? s->g_w[c] : s->g_pts[c - K];
            double *v = c < K ? &vw[c] : &vp[c - K];
            double diffp = pb - x;
            double diffg = gb - x;
            diffp *= r1;
            diffg *= r2;
            diffp *= s->c1;
            diffg *= s->c2;
            double vv = *v;
            vv += diffp;
            vv += diffg;
            vv *= chi;
            *v = vv;
        }
        ko_ap_add(K, d, xp, xw, vp, vw, s->lb, s->ub, s->fixw, s->fixp);
    }
    ko_pso_evaluate(s);
    ko_pso_update_best(s);
    s->iter++;
}

ko_pso *ko_pso_create(const ko_problem *prob, int S, int K, double c1, double c2,
                      const double *lb, const double *ub, const int *fixw, const int *fixp,
                      uint64_t seed, int mode, const double *gA, const double *gtrB,
                      int nthreads)
{
    if (K > KO_MAXK) return NULL;
    ko_pso *s = (ko_pso *)calloc(1, sizeof(ko_pso));
    int d = prob->d;
    s->S = S; s->K = K; s->d = d; s->mode = mode; s->c1 = c1; s->c2 = c2; s->seed = seed;
    s->prob = prob; s->gA = gA; s->gtrB = gtrB; s->nthreads = nthreads;
    for (int j = 0; j < d; j++) { s->lb[j] = lb[j]; s->ub[j] = ub[j]; }
    for (int k = 0; k < K; k++) { s->fixw[k] = fixw[k]; s->fixp[k] = fixp[k]; }
    size_t np = (size_t)S * K * d, nw = (size_t)S * K;
    s->x_pts = (double *)calloc(np, 8); s->p_pts = (double *)calloc(np, 8);
    s->v_pts = (double *)calloc(np, 8);
    s->x_w = (double *)calloc(nw, 8); s->p_w = (double *)calloc(nw, 8);
    s->v_w = (double *)calloc(nw, 8);
    s->g_pts = (double *)calloc((size_t)K * d, 8); s->g_w = (double *)calloc(K, 8);
    s->fx = (double *)calloc(S, 8); s->fp = (double *)calloc(S, 8);
    return s;
}

void ko_pso_destroy(ko_pso *s)
{
    if (!s) return;
    free(s->x_pts); free(s->p_pts); free(s->v_pts); free(s->x_w); free(s->p_w); free(s->v_w);
    free(s->g_pts); free(s->g_w); free(s->fx); free(s->fp); free(s);
}

/* copy-out used by tests (mirrors what trace_state = true deep-copies) */
void ko_pso_get_state(const ko_pso *s, double *x_pts, double *x_w, double *p_pts, double *p_w,
                      double *g_pts, double *g_w, double *v_pts, double *v_w, double *fx,
                      double *fp, double *fg, long *n_eval)
{
    size_t np = (size_t)s->S * s->K * s->d, nw = (size_t)s->S * s->K;
    if (x_pts) memcpy(x_pts, s->x_pts, np * 8);
    if (x_w) memcpy(x_w, s->x_w, nw * 8);
    if (p_pts) memcpy(p_pts, s->p_pts, np * 8);
    if (p_w) memcpy(p_w, s->p_w, nw * 8);
    if (g_pts) memcpy(g_pts, s->g_pts, (size_t)s->K * s->d * 8);
    if (g_w) memcpy(g_w, s->g_w, (size_t)s->K * 8);
    if (v_pts) memcpy(v_pts, s->v_pts, np * 8);
    if (v_w) memcpy(v_w, s->v_w, nw * 8);
    if (fx) memcpy(fx, s->fx, (size_t)s->S * 8);
    if (fp) memcpy(fp, s->fp, (size_t)s->S * 8);
    if (fg) *fg = s->fg;
    if (n_eval) *n_eval = s->n_eval;
}
