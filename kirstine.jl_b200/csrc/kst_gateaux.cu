// Gateaux derivative of the design objective.
//
//   gateaux_constants  src/criterion-d.jl:43-87, src/criterion-a.jl:47-77  -> k_gateaux_prepare
//   gateauxderivative! src/designproblem.jl:239-258 (+ gateaux_integrand / tr_prod,
//                      src/design-common.jl:66-68,224-248)                  -> k_gateaux_eval
//
// Everything is done in the column-scaled space of kst_models.cuh: with M = D Mt D and
// DTt = DT D^-1 the constants become
//   At = Mt^-1 Smat Mt^-1,   Smat = DTt' W DTt,
//   W  = I (A criterion)  or  (DTt Mt^-1 DTt')^-1 (D criterion);  Identity: DTt = D^-1
// and tr[A M(delta)] = sum_rows omega * gt' At gt for a one-point direction delta.
// A_j of the reference is recovered as D^-1 At D^-1 on copy-out (kst_gateaux_constants).
//
// Evaluation mapping: one thread per direction (no cross-thread reduction at all), draws are
// staged chunk-wise in shared memory and read as broadcasts, so the per-draw constants
// (8*(NDC + r(r+1)/2) bytes) are loaded once per CTA of KST_GD_NT directions instead of once
// per (direction, draw) pair; the next chunk is fetched into registers while the current one is
// evaluated.  The draw range is split over gridDim.y so that one wave of CTAs fills the GPU (split
// taken from the kernel's occupancy); partial sums are combined in fixed order.
#include "kst_objective_kernel.cuh"

#ifndef KST_GD_UNROLL
#define KST_GD_UNROLL 4
#endif
// draws evaluated side by side per thread in k_gateaux_eval (independent FP64 chains for the scheduler)
constexpr int kGdUnroll = KST_GD_UNROLL;

// doubles per row in shared memory: sqrt(omega), pc...
template <class MF>
struct KstRowW {
    static constexpr int RW = 1 + MF::NPC;
};

// rows of one design measure (single thread): map_to_covariate! + update_model_vcov! + compaction
template <class MF>
__device__ __forceinline__ int kst_build_rows(double *rows, double &sw, int &rank, const double *pts, const double *w,
                                              int K, const KstModelParams &mp)
{
    constexpr int RW = KstRowW<MF>::RW;
    int na = 0, ntriv = 0;
    sw = 0.0;
    for (int k = 0; k < K; k++) {
        const double *z = pts + (size_t)k * mp.d;
        for (int i = 0; i < mp.m; i++) {
            double pc[MF::NPC], varfac;
            const bool trivial = MF::prep_row(z, i, mp, pc, varfac);
            const double om = w[k] * varfac;
            sw += om;
            if (om == 0.0) continue;  // zero weight: contributes nothing, not even rank
            if (!trivial) {
                const double sq = sqrt(om);
                MF::post_row(pc, sq);
                rows[na * RW] = sq;
                for (int c = 0; c < MF::NPC; c++) rows[na * RW + 1 + c] = pc[c];
                na++;
            } else {
                ntriv++;
            }
        }
    }
    rank = na + ((MF::CONST_COL >= 0 && ntriv > 0) ? 1 : 0);
    return na;
}

// One thread per draw.  Writes At (full symmetric, [R*R][ld]) and trB; sets *flag on failure.
template <class MF>
__global__ void k_gateaux_prepare(const double *__restrict__ dc, const double *__restrict__ cs,
                                  const double *__restrict__ aux, const double *__restrict__ pw, size_t ld,
                                  const double *__restrict__ reg, int64_t N, int crit, int trafo, int t,
                                  const double *__restrict__ pts,
                                  const double *__restrict__ w, int K, KstModelParams mp, double *__restrict__ gA,
                                  double *__restrict__ gAp, double *__restrict__ gtrB, double *__restrict__ wtrB, int *flag)
{
    constexpr int R = MF::R, RW = KstRowW<MF>::RW;
    extern __shared__ __align__(16) double smem[];  // rows[K*m][RW], sw, (na, rank); dense covariance: a KstTile
    __shared__ __align__(16) double s_mathtab[KST_MATH_SMEM];
    const KstTab tab = kst_math_init(s_mathtab);
    double *rows = smem;
    double *s_sw = smem + (size_t)K * mp.m * RW;
    int *s_na = (int *)(s_sw + 1);
    int design_rank;
    if constexpr (KstDense<MF>::value) {
        // the design as a one-particle tile of the objective kernel: unit factors and all (ends with __syncthreads())
        kst_build_tile<MF>(smem, pts, w, 0, 1, K, mp, 1);
        design_rank = KstTile<MF>::rank(smem, K * mp.m)[0];
    } else {
        if (threadIdx.x == 0) {
            double sw;
            int rank;
            s_na[0] = kst_build_rows<MF>(rows, sw, rank, pts, w, K, mp);
            s_na[1] = rank;
            *s_sw = sw;
        }
        __syncthreads();
        design_rank = s_na[1];
    }
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= N) return;
    double dcv[MF::NDC], M[KstPacked<R>::N];
#pragma unroll
    for (int c = 0; c < MF::NDC; c++) dcv[c] = dc[(size_t)c * ld + j];
    if constexpr (KstDense<MF>::value) {
        const int rows_max = K * mp.m;
        kst_fisher<MF>(M, (unsigned)__cvta_generic_to_shared(KstTile<MF>::rows(smem)), KstTile<MF>::nact(smem, rows_max)[0],
                       KstTile<MF>::sumw(smem, rows_max)[0], dcv, tab,
                       (unsigned)__cvta_generic_to_shared(KstTile<MF>::lmat(smem, rows_max)), mp.m);
    } else {
#pragma unroll
        for (int i = 0; i < KstPacked<R>::N; i++) M[i] = 0.0;
        const int na = *s_na;
        for (int i = 0; i < na; i++) {
            double g[R];
            MF::row(rows + i * RW + 1, dcv, rows[i * RW], g, tab);
            kst_accum<R, MF::CONST_COL>(M, g);
        }
        constexpr int cc = MF::CONST_COL >= 0 ? MF::CONST_COL : 0;
        if (MF::CONST_COL >= 0) M[kst_pidx<R>(cc, cc)] = *s_sw;
    }
    // regularised normal approximation: informationmatrix!(afm, na) adds R (docs/src/extend-approximation.md:69-79)
    double Rt[KstPacked<R>::N];
#pragma unroll
    for (int i = 0; i < KstPacked<R>::N; i++) {
        Rt[i] = reg ? reg[(size_t)i * ld + j] : 0.0;
        M[i] += Rt[i];
    }

    double L[R][R], dv[R], dinv[R];
    bool ok = kst_ldlt<R, false>(M, L, dv, dinv, true);
    ok = ok && (reg || design_rank >= R);  // fewer rows than parameters: singular by construction
    // Mt^-1 = Li' diag(dinv) Li with Li = L^-1 (unit lower)
    double Li[R][R];
#pragma unroll
    for (int a = 0; a < R; a++) {
#pragma unroll
        for (int k = 0; k < R; k++) {
            double v = (k == a) ? 1.0 : 0.0;
            if (k > a) {
#pragma unroll
                for (int l = 0; l < k; l++)
                    if (l >= a) v = fma(-L[k][l], Li[l][a], v);
            }
            Li[k][a] = v;
        }
    }
    double Mi[R][R];
#pragma unroll
    for (int a = 0; a < R; a++)
#pragma unroll
        for (int b = a; b < R; b++) {
            double s = 0.0;
#pragma unroll
            for (int k = 0; k < R; k++)
                if (k >= b) s = fma(Li[k][a] * dinv[k], Li[k][b], s);
            Mi[a][b] = s;
            Mi[b][a] = s;
        }
    // DTt (tt x R)
    double DTt[R][R];
    const int tt = trafo == KST_TRAFO_IDENTITY ? R : t;
    if (trafo == KST_TRAFO_IDENTITY) {
#pragma unroll
        for (int a = 0; a < R; a++)
#pragma unroll
            for (int i = 0; i < R; i++) DTt[a][i] = (a == i) ? 1.0 / cs[(size_t)i * ld + j] : 0.0;
    } else {
#pragma unroll
        for (int a = 0; a < R; a++)
#pragma unroll
            for (int i = 0; i < R; i++) DTt[a][i] = a < tt ? aux[(size_t)(a * R + i) * ld + j] : 0.0;
    }
    double At[R][R];
    double trB;
    // Dexp shares the D constants up to a per-draw factor: psi_j / t with psi_j = det(M_T)^(1/t) the draw's
    // criterion value, and trB = psi_j (docs/src/extend-criterion.md:40-67)
    double logdet_mt = 0.0;  // log det M_T
    if (crit != KST_CRIT_A && trafo == KST_TRAFO_IDENTITY) {
#pragma unroll
        for (int a = 0; a < R; a++)
#pragma unroll
            for (int b = 0; b < R; b++) At[a][b] = Mi[a][b];
        trB = (double)R;
#pragma unroll
        for (int k = 0; k < R; k++) {
            const double c = cs[(size_t)k * ld + j];
            logdet_mt += log(dv[k]) + log(c * c);
        }
    } else {
        // P = DTt Mt^-1 (tt x R), C = P DTt' (tt x tt)
        double P[R][R], C[R][R], W[R][R];
#pragma unroll
        for (int a = 0; a < R; a++)
#pragma unroll
            for (int i = 0; i < R; i++) {
                double s = 0.0;
#pragma unroll
                for (int k = 0; k < R; k++) s = fma(DTt[a][k], Mi[k][i], s);
                P[a][i] = s;
            }
#pragma unroll
        for (int a = 0; a < R; a++)
#pragma unroll
            for (int b = 0; b < R; b++) {
                double s = 0.0;
#pragma unroll
                for (int k = 0; k < R; k++) s = fma(P[a][k], DTt[b][k], s);
                C[a][b] = s;
            }
        if (crit == KST_CRIT_A) {
            trB = 0.0;
#pragma unroll
            for (int a = 0; a < R; a++) {
                if (a < tt) trB += C[a][a];
#pragma unroll
                for (int b = 0; b < R; b++) W[a][b] = (a == b && a < tt) ? 1.0 : 0.0;
            }
        } else {
            // W = C^-1 (tt x tt, symmetric positive definite) by Gauss-Jordan without pivoting
            trB = (double)tt;
#pragma unroll
            for (int a = 0; a < R; a++)
#pragma unroll
                for (int b = 0; b < R; b++) W[a][b] = (a == b) ? 1.0 : 0.0;
#pragma unroll
            for (int c = 0; c < R; c++) {
                if (c < tt) {
                    const double piv = C[c][c];
                    ok = ok && (piv > 0.0);
                    logdet_mt -= log(piv);  // C = M_T^-1
                    const double ip = 1.0 / piv;
#pragma unroll
                    for (int b = 0; b < R; b++) { C[c][b] *= ip; W[c][b] *= ip; }
#pragma unroll
                    for (int a = 0; a < R; a++) {
                        if (a != c && a < tt) {
                            const double f = C[a][c];
#pragma unroll
                            for (int b = 0; b < R; b++) {
                                C[a][b] = fma(-f, C[c][b], C[a][b]);
                                W[a][b] = fma(-f, W[c][b], W[a][b]);
                            }
                        }
                    }
                }
            }
#pragma unroll
            for (int a = 0; a < R; a++)
#pragma unroll
                for (int b = 0; b < R; b++)
                    if (a >= tt || b >= tt) W[a][b] = 0.0;
        }
        // At = P' W P
        double WP[R][R];
#pragma unroll
        for (int a = 0; a < R; a++)
#pragma unroll
            for (int i = 0; i < R; i++) {
                double s = 0.0;
#pragma unroll
                for (int b = 0; b < R; b++) s = fma(W[a][b], P[b][i], s);
                WP[a][i] = s;
            }
#pragma unroll
        for (int i = 0; i < R; i++)
#pragma unroll
            for (int k = 0; k < R; k++) {
                double s = 0.0;
#pragma unroll
                for (int a = 0; a < R; a++) s = fma(P[a][i], WP[a][k], s);
                At[i][k] = s;
            }
    }
    if (crit == KST_CRIT_DEXP) {
        const double psi = exp(logdet_mt / tt);
        const double f = psi / tt;
#pragma unroll
        for (int a = 0; a < R; a++)
#pragma unroll
            for (int b = 0; b < R; b++) At[a][b] *= f;
        trB = psi;
    }
    if (!ok) atomicExch(flag, 1);
#pragma unroll
    for (int a = 0; a < R; a++)
#pragma unroll
        for (int b = 0; b < R; b++) gA[(size_t)(a * R + b) * ld + j] = At[a][b];
    // the form k_gateaux_eval streams: pw_j At_j, upper triangle, off-diagonals doubled
#pragma unroll
    for (int a = 0; a < R; a++)
#pragma unroll
        for (int b = a; b < R; b++) gAp[(size_t)kst_pidx<R>(a, b) * ld + j] = pw[j] * At[a][b] * (a == b ? 1.0 : 2.0);
    gtrB[j] = trB;
    // the direction's matrix is regularised too (src/designproblem.jl:253-254): tr_prod(A, M(delta) + R, :U) adds
    // the direction-independent tr_U[A R], folded into the constant here
    double trAR = 0.0;
#pragma unroll
    for (int a = 0; a < R; a++)
#pragma unroll
        for (int b = a; b < R; b++) trAR = fma(At[a][b] * (a == b ? 1.0 : 2.0), Rt[kst_pidx<R>(a, b)], trAR);
    wtrB[j] = pw[j] * (trB - trAR);
}

// Draw loop of k_gateaux_eval for one direction (thread): sum over the CTA's draws [j0, j1) of
// pw_j * sum_rows omega gt' At_j gt.  NL >= 0: exactly NL rows, constants in registers (lsq, lpc), no branch inside
// the loop.  NL = -1: m rows with constants in shared memory (s_rows), rows outside live_mask skipped.
// CL = false: the CTA has proven that no exp argument leaves +-700 (MF::arg_bound), rows skip the exponent clamp.
template <class MF, int NL, bool CL = true>
__device__ __forceinline__ double kst_gd_draws(const double *__restrict__ dc, const double *__restrict__ gAp,
                                               size_t ld, int64_t j0, int64_t j1,
                                               double *s_draw, const double *s_rows, int m, unsigned live_mask,
                                               const double *lsq, const double (*lpc)[MF::NPC], double sw_triv,
                                               KstTab tab)
{
    constexpr int R = MF::R, RW = KstRowW<MF>::RW, NP = KstPacked<R>::N, NS = MF::NDC + NP, NSP = (NS + 1) / 2 * 2;
    static_assert(KST_GD_CHUNK == KST_GD_NT, "one staged draw per thread");
    double acc = 0.0;
    // Software pipeline: the next chunk's constants travel from L2 into registers while the current chunk is being
    // evaluated out of shared memory, so the only exposed memory latency is the first chunk's.
    double nxt[NS];
    {
        const int64_t j = j0 + threadIdx.x;
#pragma unroll
        for (int c = 0; c < NS; c++)
            nxt[c] = j < j1 ? (c < MF::NDC ? dc[(size_t)c * ld + j] : gAp[(size_t)(c - MF::NDC) * ld + j]) : 0.0;
    }
    for (int64_t jb = j0; jb < j1; jb += KST_GD_CHUNK) {
        const int nj = (int)min((int64_t)KST_GD_CHUNK, j1 - jb);
        __syncthreads();  // the previous chunk has been consumed (first pass: the math tables are in place)
#pragma unroll
        for (int c = 0; c < NSP; c++) s_draw[threadIdx.x * NSP + c] = c < NS ? nxt[c < NS ? c : 0] : 0.0;
        __syncthreads();
        {
            const int64_t j = jb + KST_GD_CHUNK + threadIdx.x;
            if (j < j1) {
#pragma unroll
                for (int c = 0; c < NS; c++)
                    nxt[c] = c < MF::NDC ? dc[(size_t)c * ld + j] : gAp[(size_t)(c - MF::NDC) * ld + j];
            }
        }
#pragma unroll kGdUnroll
        for (int jj = 0; jj < nj; jj++) {
            double dv[NSP];
#pragma unroll
            for (int c = 0; c < NSP; c += 2) {
                const double2 v = *reinterpret_cast<const double2 *>(s_draw + jj * NSP + c);
                dv[c] = v.x;
                dv[c + 1] = v.y;
            }
            const double *dcv = dv, *Ap = dv + MF::NDC;
            double qsum = 0.0;
            if (NL >= 0) {
#pragma unroll
                for (int i = 0; i < (NL > 0 ? NL : 0); i++) {
                    double gv[R];
                    MF::template row<CL>(lpc[i], dcv, lsq[i], gv, tab);  // weighted row: gv' A gv = omega * gt' A gt
                    double q = 0.0;
#pragma unroll
                    for (int a = 0; a < R; a++) {
                        double s = 0.0;
#pragma unroll
                        for (int b = a; b < R; b++) s = fma(Ap[kst_pidx<R>(a, b)], gv[b], s);
                        q = fma(s, gv[a], q);
                    }
                    qsum += q;
                }
            } else {
                for (int i = 0; i < m; i++) {
                    if (!((live_mask >> i) & 1u)) continue;  // uniform over the CTA
                    double pc[MF::NPC], gv[R];
                    const double sq = s_rows[(size_t)(i * RW) * KST_GD_NT + threadIdx.x];
#pragma unroll
                    for (int c = 0; c < MF::NPC; c++) pc[c] = s_rows[(size_t)(i * RW + 1 + c) * KST_GD_NT + threadIdx.x];
                    MF::row(pc, dcv, sq, gv, tab);
                    double q = 0.0;
#pragma unroll
                    for (int a = 0; a < R; a++) {
                        double s = 0.0;
#pragma unroll
                        for (int b = a; b < R; b++) s = fma(Ap[kst_pidx<R>(a, b)], gv[b], s);
                        q = fma(s, gv[a], q);
                    }
                    qsum += q;
                }
            }
            constexpr int cc = MF::CONST_COL >= 0 ? MF::CONST_COL : 0;
            if (MF::CONST_COL >= 0) qsum = fma(sw_triv, Ap[kst_pidx<R>(cc, cc)], qsum);
            acc += qsum;
        }
    }
    return acc;
}

struct KstDcMax {
    double v[8];  // max_j |dc[c][j]| (kst_problem::dcmax)
};

// One thread per direction; draws staged in shared memory.
// MR > 0: the unit has exactly MR rows and their constants stay in registers; MR = 0: any unit length, row
// constants in shared memory.  A draw's constants (NDC hoisted constants + packed pw_j At_j) are stored
// contiguously and read as 16-byte broadcasts.
template <class MF, int MR>
__global__ void __launch_bounds__(KST_GD_NT) k_gateaux_eval(const double *__restrict__ dc,
                                                            const double *__restrict__ gAp, size_t ld, int64_t N,
                                                            const double *__restrict__ dirs, int64_t G,
                                                            KstModelParams mp, KstDcMax dcmax,
                                                            double *__restrict__ partials)
{
    constexpr int R = MF::R, RW = KstRowW<MF>::RW, NP = KstPacked<R>::N, NS = MF::NDC + NP, NSP = (NS + 1) / 2 * 2;
    constexpr int MRR = MR > 0 ? MR : 1;
    extern __shared__ __align__(16) double smem[];
    __shared__ __align__(16) double s_mathtab[KST_MATH_SMEM];
#if KST_EXP_BIG && KST_TMA_TABLES
    __shared__ __align__(8) unsigned long long s_mbar;
    const KstTab tab = kst_math_init_async(s_mathtab, &s_mbar);  // waited for after the row set-up below
#else
    const KstTab tab = kst_math_init(s_mathtab);  // visible after the first __syncthreads() of the chunk loop
#endif
    double *s_draw = smem;                                  // [KST_GD_CHUNK][NSP]
    double *s_rows = smem + (size_t)NSP * KST_GD_CHUNK;     // [m][RW][KST_GD_NT]   (MR == 0 only)
    const int m = MR > 0 ? MR : mp.m;
    const int64_t g = (int64_t)blockIdx.x * KST_GD_NT + threadIdx.x;
    const bool live = g < G;

    // rows of this thread's direction (weight 1); trivial rows get sqrt(omega) = 0 and benign pc.
    // live_mask: bit i is set when row i is live for at least one direction of the CTA; rows that are
    // trivial for the whole CTA (time 0 of every vector unit) are skipped outright in the draw loop.
    double sw_triv = 0.0, bound = 0.0;
    unsigned live_mask = 0;
    double rsq[MRR], rpc[MRR][MF::NPC];
#pragma unroll
    for (int i = 0; i < (MR > 0 ? MR : 1); i++) {
        rsq[i] = 0.0;
#pragma unroll
        for (int c = 0; c < MF::NPC; c++) rpc[i][c] = 1.0;
    }
    for (int i = 0; i < m; i++) {
        double pc[MF::NPC], varfac = 0.0;
        bool trivial = true;
        if (live) trivial = MF::prep_row(dirs + g * mp.d, i, mp, pc, varfac);
        double sq = sqrt(varfac);
        if (trivial) {
            sw_triv += varfac;
            sq = 0.0;
            for (int c = 0; c < MF::NPC; c++) pc[c] = 1.0;
        }
        if (MF::FAST_EXP) bound = fmax(bound, MF::arg_bound(pc, dcmax.v));
        MF::post_row(pc, sq);
        if (MR > 0) {
#pragma unroll
            for (int ii = 0; ii < MRR; ii++)
                if (ii == i) {
                    rsq[ii] = sq;
#pragma unroll
                    for (int c = 0; c < MF::NPC; c++) rpc[ii][c] = pc[c];
                }
        } else {
            s_rows[(size_t)(i * RW) * KST_GD_NT + threadIdx.x] = sq;
            for (int c = 0; c < MF::NPC; c++) s_rows[(size_t)(i * RW + 1 + c) * KST_GD_NT + threadIdx.x] = pc[c];
        }
        if (__syncthreads_or(sq != 0.0)) live_mask |= 1u << i;
    }

#if KST_EXP_BIG && KST_TMA_TABLES
    kst_math_wait(&s_mbar);  // (m >= 1: at least one CTA-wide barrier lies behind us)
#endif
    // this CTA's draw range
    const int64_t per = (N + gridDim.y - 1) / gridDim.y;
    const int64_t j0 = per * blockIdx.y, j1 = min(N, j0 + per);
    double acc = 0.0;
    if (MR > 0) {
        // Compact the rows that are live for at least one direction of the CTA to the front (uniform over the CTA),
        // then run the draw loop specialised on their number: with no per-row branch inside it, the compiler
        // interleaves the rows of the kGdUnroll draws evaluated side by side.
        double lsq[MRR], lpc[MRR][MF::NPC];
        int nl = 0;
#pragma unroll
        for (int i = 0; i < MRR; i++) {
            lsq[i] = 0.0;
#pragma unroll
            for (int c = 0; c < MF::NPC; c++) lpc[i][c] = 1.0;
        }
#pragma unroll
        for (int i = 0; i < MRR; i++) {
            if ((live_mask >> i) & 1u) {
#pragma unroll
                for (int o = 0; o < MRR; o++)
                    if (o == nl) {
                        lsq[o] = rsq[i];
#pragma unroll
                        for (int c = 0; c < MF::NPC; c++) lpc[o][c] = rpc[i][c];
                    }
                nl++;
            }
        }
        // every exp argument of every direction of the CTA within +-700 for every draw: rows without the exponent
        // clamp (register-row kernels only; a NaN bound compares false)
        const bool fast = MF::FAST_EXP && __syncthreads_and(bound <= 700.0);
#define KST_GD_CASE(NLV)                                                                                          \
    acc = fast ? kst_gd_draws<MF, NLV, !MF::FAST_EXP>(dc, gAp, ld, j0, j1, s_draw, nullptr, 0, 0u, lsq, lpc, sw_triv, tab) \
               : kst_gd_draws<MF, NLV, true>(dc, gAp, ld, j0, j1, s_draw, nullptr, 0, 0u, lsq, lpc, sw_triv, tab)
        switch (nl) {
        case 0: KST_GD_CASE(0); break;
        case 1: KST_GD_CASE(1); break;
        case 2: KST_GD_CASE((MR >= 2 ? 2 : 1)); break;
        case 3: KST_GD_CASE((MR >= 3 ? 3 : 1)); break;
        default: KST_GD_CASE((MR >= 4 ? 4 : 1)); break;
        }
#undef KST_GD_CASE
    } else {
        acc = kst_gd_draws<MF, -1>(dc, gAp, ld, j0, j1, s_draw, s_rows, m, live_mask, rsq, rpc, sw_triv, tab);
    }
    if (live) partials[g * gridDim.y + blockIdx.y] = acc;
}

__global__ void k_fill(double *x, double v, int64_t n)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) x[i] = v;
}

__global__ void k_gd_finish(const double *__restrict__ partials, int n_split, double offset, int singular,
                            double *__restrict__ out, int64_t G)
{
    const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= G) return;
    if (singular) {
        out[g] = __longlong_as_double(0x7ff8000000000000LL);
        return;
    }
    double s = 0.0;
    for (int b = 0; b < n_split; b++) s += partials[g * n_split + b];
    out[g] = s - offset;
}


int kst_launch_gateaux_prepare(kst_problem *h, int K, const double *d_pts, const double *d_w)
{
    if (K < 1 || K * h->m > KST_MAX_ROWS) {
        kst_set_error("gateaux_prepare: need K >= 1 and K*m <= %d", KST_MAX_ROWS);
        return KST_EINVAL;
    }
    if (!h->gA) {
        KST_CUDA(cudaMalloc(&h->gA, sizeof(double) * h->r * h->r * h->ld));
        KST_CUDA(cudaMalloc(&h->gAp, sizeof(double) * (h->r * (h->r + 1) / 2) * h->ld));
        KST_CUDA(cudaMalloc(&h->gtrB, sizeof(double) * h->ld));
        KST_CUDA(cudaMalloc(&h->g_wtrB, sizeof(double) * h->ld));
    }
    double *wtrB = h->g_wtrB;
    KST_CUDA(cudaMemsetAsync(h->d_flag, 0, sizeof(int), h->stream));
    const int nt = 128;
    const unsigned nb = (unsigned)((h->N + nt - 1) / nt);
    KST_MODEL_SWITCH(h->model_id, {
        const size_t smem = KstDense<MF>::value ? KstTile<MF>::bytes(K * h->m)
                                                : sizeof(double) * ((size_t)K * h->m * KstRowW<MF>::RW + 2);
        if (smem + sizeof(double) * KST_MATH_SMEM > 48 * 1024)
            KST_CUDA(cudaFuncSetAttribute(k_gateaux_prepare<MF>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k_gateaux_prepare<MF><<<nb, nt, smem, h->stream>>>(h->dc, h->cs, h->aux, h->pw, h->ld, h->reg, h->N, h->crit, h->trafo,
                                                           h->t, d_pts, d_w, K, h->mp, h->gA, h->gAp, h->gtrB, wtrB, h->d_flag);
    });
    KST_CUDA(cudaGetLastError());
    // one round trip for both results: the singularity flag (offset 0 of d_flag) and the summed constant (offset 8)
    int rc = kst_sum_fixed_enqueue(h, wtrB, h->N);
    if (rc) return rc;
    struct { int flag, pad; double sum; } back;
    KST_CUDA(cudaMemcpyAsync(&back, h->d_flag, 16, cudaMemcpyDeviceToHost, h->stream));
    KST_CUDA(cudaStreamSynchronize(h->stream));
    h->g_offset = back.sum;
    h->g_singular = back.flag;
    h->g_ready = 1;
    return KST_OK;
}

// Draw split of the evaluation grid: gblocks CTAs of KST_GD_NT directions each, times n_split draw ranges.  All CTAs
// take the same time, so the aim is one wave that fills every SM's resident-CTA slots as evenly as possible (or a
// whole number of waves when there are more direction blocks than slots).
static int64_t kst_gd_split(const kst_problem *h, int64_t gblocks, int resident_per_sm)
{
    const int64_t max_split = (h->N + KST_GD_CHUNK - 1) / KST_GD_CHUNK;
    const int64_t cap = (int64_t)h->sm_count * (resident_per_sm > 0 ? resident_per_sm : 1);
    int64_t best = 1;
    double best_waste = 1e30;
    const int64_t hi = (cap + gblocks - 1) / gblocks;  // smallest split that reaches a full wave
    for (int64_t ns = 1; ns <= hi && ns <= max_split; ns++) {
        const int64_t tot = ns * gblocks, waves = (tot + cap - 1) / cap;
        // idle slots in the last wave, plus a mild preference for more CTAs (latency hiding) at equal balance
        const double waste = (double)(waves * cap - tot) / (double)(waves * cap);
        if (waste < best_waste - 1e-9) { best_waste = waste; best = ns; }
    }
    if (best > 65535) best = 65535;
    return best;
}

template <class MF, int MR>
static int launch_gd(kst_problem *h, int64_t G, const double *d_dirs, size_t smem, int *n_split_out)
{
    // vector units with many rows: opt in to a large dynamic shared-memory window (the static math tables count
    // towards the 48 KB a kernel gets without asking)
    if (smem + sizeof(double) * KST_MATH_SMEM > 48 * 1024)
        KST_CUDA(cudaFuncSetAttribute(k_gateaux_eval<MF, MR>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int resident = 0;
    KST_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&resident, k_gateaux_eval<MF, MR>, KST_GD_NT, smem));
    const int64_t gblocks = (G + KST_GD_NT - 1) / KST_GD_NT;
    const int64_t n_split = kst_gd_split(h, gblocks, resident);
    const size_t need = (size_t)G * n_split;
    if (need > h->gd_partials_cap) {
        if (h->gd_partials) KST_CUDA(cudaFree(h->gd_partials));
        h->gd_partials = nullptr;
        h->gd_partials_cap = 0;
        KST_CUDA(cudaMalloc(&h->gd_partials, sizeof(double) * need));
        h->gd_partials_cap = need;
    }
    *n_split_out = (int)n_split;
    if (h->g_singular) return KST_OK;  // the finishing kernel fills NaN
    const dim3 grid((unsigned)gblocks, (unsigned)n_split);
    KstDcMax dm;
    for (int c = 0; c < 8; c++) dm.v[c] = h->dcmax[c];
    {  // kst_profile_enable: CUDA events around the evaluation kernel, read by kst_profile_read
        const int rcp = kst_prof_begin(h, KST_PROF_EVAL);
        if (rcp) return rcp;
    }
    k_gateaux_eval<MF, MR><<<grid, KST_GD_NT, smem, h->stream>>>(h->dc, h->gAp, h->ld, h->N, d_dirs, G, h->mp, dm,
                                                                 h->gd_partials);
    KST_CUDA(cudaGetLastError());
    return kst_prof_end(h, KST_PROF_EVAL);
}

template <class MF>
static int launch_gd_model(kst_problem *h, int64_t G, const double *d_dirs, int *n_split)
{
    constexpr int NSP = (MF::NDC + KstPacked<MF::R>::N + 1) / 2 * 2;
    const bool in_regs = h->m <= 4;
    const size_t smem =
        sizeof(double) * ((size_t)NSP * KST_GD_CHUNK + (in_regs ? 0 : (size_t)h->m * KstRowW<MF>::RW * KST_GD_NT));
    switch (in_regs ? h->m : 0) {
    case 1: return launch_gd<MF, 1>(h, G, d_dirs, smem, n_split);
    case 2: return launch_gd<MF, 2>(h, G, d_dirs, smem, n_split);
    case 3: return launch_gd<MF, 3>(h, G, d_dirs, smem, n_split);
    case 4: return launch_gd<MF, 4>(h, G, d_dirs, smem, n_split);
    default: return launch_gd<MF, 0>(h, G, d_dirs, smem, n_split);
    }
}

int kst_launch_gateaux_eval(kst_problem *h, int64_t G, const double *d_dirs, double *d_out)
{
    if (!h->g_ready) {
        kst_set_error("gateaux_eval before gateaux_prepare");
        return KST_ESTATE;
    }
    if (G < 1) {
        kst_set_error("gateaux_eval: need G >= 1");
        return KST_EINVAL;
    }
    bool dense = false;
    KST_MODEL_SWITCH(h->model_id, dense = KstDense<MF>::value);
    if (dense) {
        // a dense unit covariance couples the rows of a unit: the directions go through the objective kernel with the
        // Gateaux integrand as the per-draw value (kst_objective_kernel.cuh), which whitens units; weights are 1
        if (G > 65535 * KST_TILE_P) { kst_set_error("gateaux_eval: at most %d directions per call for this model", 65535 * KST_TILE_P); return KST_ELIMIT; }
        if ((size_t)G > h->gd_ones_cap) {
            if (h->gd_ones) KST_CUDA(cudaFree(h->gd_ones));
            h->gd_ones = nullptr;
            h->gd_ones_cap = 0;
            KST_CUDA(cudaMalloc(&h->gd_ones, sizeof(double) * G));
            h->gd_ones_cap = (size_t)G;
            k_fill<<<(unsigned)((G + 255) / 256), 256, 0, h->stream>>>(h->gd_ones, 1.0, G);
            KST_CUDA(cudaGetLastError());
        }
        return kst_launch_objective(h, (int)G, 1, d_dirs, h->gd_ones, d_out, nullptr, 1);
    }
    int n_split = 1;
    int rc = KST_OK;
    KST_MODEL_SWITCH(h->model_id, rc = launch_gd_model<MF>(h, G, d_dirs, &n_split));
    if (rc) return rc;
    k_gd_finish<<<(unsigned)((G + 127) / 128), 128, 0, h->stream>>>(h->gd_partials, n_split, h->g_offset,
                                                                   h->g_singular, d_out, G);
    KST_CUDA(cudaGetLastError());
    return KST_OK;
}
