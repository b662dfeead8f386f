// Evaluation kernel of the objective and its per-family launch code (see kst_objective.cu for the mapping).
// Included by kst_objective.cu (host entry points, small kernels) and by the kst_obj_inst_*.cu files, each of
// which instantiates kst_obj_launch for some model families so that the families compile in parallel.
#pragma once
#ifndef KST_DRAW_OUTER
#define KST_DRAW_OUTER 1     // 1: a thread takes a pair of draws through ALL particles of the tile before moving on
                             // (per-particle sums in shared memory); 0: particle by particle, draws inside
#endif
#ifndef KST_ROWS_OVERRIDE
#define KST_ROWS_OVERRIDE 0  // tuning: force the row interleave factor (1, 2, 4, 8) for every family
#endif



#include "kst_common.cuh"
#include "kst_linalg.cuh"

struct KstObjParams {
    const double *pw, *dc, *aux;
    const double *reg;      // REG kernels: packed upper Rt_j, [r(r+1)/2][ld]
    size_t ld;
    int64_t N;
    const double *pts, *w;  // [S][K][d], [S][K]
    int S, K;
    double *partials;       // [n_blk][S]
    KstModelParams mp;
    double w_uniform;       // the common prior weight when all weights are equal (UNI kernels), else unused
    int tile_p;             // particles per CTA, 1..KST_TILE_P (small problems use small tiles to fill the SMs)
    double dcmax[8];        // max_j |dc[c][j]| per hoisted constant (Inf / NaN when a draw is not finite): lets a
                            // particle prove that its exp arguments stay inside the exponent range (MF::arg_bound)
    int gd;                 // 1: evaluate the Gateaux derivative integrand instead of the criterion (CRIT = 3 kernels;
                            // aux = packed pw_j At_j of the prepared design, [r(r+1)/2][ld])
};

// criterion ids inside the kernels: the public KST_CRIT_* values plus the Gateaux integrand
#define KST_KCRIT_GD 3

// ------------------------------------------------------------------------------------------
// shared-memory particle tile
// ------------------------------------------------------------------------------------------
template <class MF>
struct KstTile {
    // doubles per row: sqrt(omega), pc...; padded to an even count so rows can be read as 16-byte vectors
    static constexpr int RW = (1 + MF::NPC + 1) / 2 * 2;
    // layout (doubles): rows[P][rows_max][RW] | sumw[P] | isumw[P] | wsum[P][NT/32] ; then ints nact[P], rank[P]
    __device__ static double *rows(double *s) { return s; }
    __device__ static double *sumw(double *s, int rows_max) { return s + (size_t)KST_TILE_P * rows_max * RW; }
    __device__ static double *isumw(double *s, int rows_max) { return sumw(s, rows_max) + KST_TILE_P; }
    __device__ static double *wsum(double *s, int rows_max) { return isumw(s, rows_max) + KST_TILE_P; }
    __device__ static int *nact(double *s, int rows_max) { return (int *)(wsum(s, rows_max) + KST_TILE_P * (KST_NT / 32)); }
    // upper bound on rank(Mt): below R the design is singular whatever the rounding does
    __device__ static int *rank(double *s, int rows_max) { return nact(s, rows_max) + KST_TILE_P; }
    // 1 when every exp argument of the particle is provably within +-700 for every draw (no exponent clamp needed)
    __device__ static int *safe(double *s, int rows_max) { return rank(s, rows_max) + KST_TILE_P; }
    // Families with a dense unit covariance (KstDense): per particle the packed lower factors L_k = sqrt(w_k) U_k^-T of
    // its units, m (m + 1) / 2 doubles each, K m (m + 1) / 2 <= rows_max * LW in all
    static constexpr int LW = (KST_MAX_UNIT + 2) / 2;
    __host__ __device__ static constexpr size_t base_doubles(int rows_max)
    {
        return (size_t)KST_TILE_P * rows_max * RW + 2 * KST_TILE_P + KST_TILE_P * (KST_NT / 32) +
               (sizeof(int) * 3 * KST_TILE_P + 15) / 16 * 2;
    }
    __device__ static double *lmat(double *s, int rows_max) { return s + base_doubles(rows_max); }
    // per-thread accumulators of the draw-outer evaluation loop: sum and running determinant product of every particle
    // of the tile, [2][KST_TILE_P][KST_NT] doubles behind everything else
    __host__ __device__ static size_t acc_offset(int rows_max)
    {
        return base_doubles(rows_max) + (KstDense<MF>::value ? (size_t)KST_TILE_P * rows_max * LW : 0);
    }
    __host__ __device__ static size_t bytes(int rows_max)
    {
        return sizeof(double) * (acc_offset(rows_max) + (KST_DRAW_OUTER ? 2 * (size_t)KST_TILE_P * KST_NT : 0));
    }
};

// update_model_workspace! for one design point of a dense-covariance family (src/model-nonlinear-regression.jl:39-49):
// Sigma = U'U (upper Cholesky, the loop order of LAPACK's unblocked potrf), then Lw = sw U^-T packed by rows
// (Lw[i (i + 1) / 2 + l], l <= i), so that the whitened, weighted Jacobian of the unit is Lw J -- what trsm!('L','U','T','N')
// followed by the weight of syrk! computes (:79-83).  Returns false when Sigma is not positive definite.
template <class MF>
__device__ inline bool kst_unit_factor(const double *z, const KstModelParams &mp, double sw, double *Lw)
{
    const int m = mp.m;
    double U[KST_MAX_UNIT * KST_MAX_UNIT];
    MF::vcov(z, mp, U);
    bool ok = true;
    for (int j = 0; j < m; j++) {
        double ajj = U[j + j * m];
        for (int k = 0; k < j; k++) ajj -= U[k + j * m] * U[k + j * m];
        ok = ok && (ajj > 0.0);
        ajj = sqrt(ajj);
        U[j + j * m] = ajj;
        for (int i = j + 1; i < m; i++) {
            double v = U[j + i * m];
            for (int k = 0; k < j; k++) v -= U[k + j * m] * U[k + i * m];
            U[j + i * m] = v / ajj;
        }
    }
    // column c of L = U^-T by forward substitution with U' (lower): L[i][c] = (delta_ic - sum_{l=c}^{i-1} U[l][i] L[l][c]) / U[i][i]
    for (int c = 0; c < m; c++)
        for (int i = c; i < m; i++) {
            double v = i == c ? 1.0 : 0.0;
            for (int l = c; l < i; l++) v -= U[l + i * m] * Lw[l * (l + 1) / 2 + c];
            Lw[i * (i + 1) / 2 + c] = v / U[i + i * m];
        }
    for (int i = 0; i < m * (m + 1) / 2; i++) Lw[i] *= sw;
    return ok;
}

// One thread per particle of the tile: map design points to rows (map_to_covariate! +
// update_model_vcov!, src/design-common.jl:51-61, src/model-nonlinear-regression.jl:39-49),
// hoist covariate-only transcendentals, drop trivial and zero-weight rows.
template <class MF>
__device__ __forceinline__ void kst_build_tile(double *smem, const double *__restrict__ pts,
                                               const double *__restrict__ w, int p0, int S, int K,
                                               const KstModelParams &mp, int tile_p = KST_TILE_P,
                                               const double *dcmax = nullptr, int pidx0 = 0)
{
    // pidx0: pts / w hold the particles pidx0, pidx0 + 1, ... (the persistent swarm kernel passes its shared-memory
    // copy of the tile's particles; everyone else the whole swarm, pidx0 = 0)
    const int rows_max = K * mp.m;
    constexpr int RW = KstTile<MF>::RW;
    if ((int)threadIdx.x < tile_p) {
        const int p = p0 + threadIdx.x;
        double *rows = KstTile<MF>::rows(smem) + (size_t)threadIdx.x * rows_max * RW;
        int na = 0, ntriv = 0;
        double sw = 0.0, bound = 0.0;
        bool dense_ok = true;
        if (KstDense<MF>::value && p < S) {
            // dense unit covariance: whole units are kept or (at weight zero) dropped; every row record carries
            // weight 1, the unit's factor sqrt(w_k) U_k^-T goes to the L area
            double *lm = KstTile<MF>::lmat(smem, rows_max) + (size_t)threadIdx.x * rows_max * KstTile<MF>::LW;
            const int m = mp.m;
            for (int k = 0; k < K; k++) {
                const double *z = pts + ((size_t)(p - pidx0) * K + k) * mp.d;
                const double wk = w[(size_t)(p - pidx0) * K + k];
                sw += wk;
                if (wk == 0.0) continue;
                if constexpr (KstDense<MF>::value)
                    dense_ok = kst_unit_factor<MF>(z, mp, sqrt(wk), lm + (size_t)(na / m) * (m * (m + 1) / 2)) && dense_ok;
                for (int i = 0; i < m; i++) {
                    double pc[MF::NPC], varfac;
                    MF::prep_row(z, i, mp, pc, varfac);
                    if (MF::FAST_EXP && dcmax) bound = fmax(bound, MF::arg_bound(pc, dcmax));
                    MF::post_row(pc, 1.0);
                    rows[na * RW] = 1.0;
#pragma unroll
                    for (int c = 0; c < MF::NPC; c++) rows[na * RW + 1 + c] = pc[c];
                    na++;
                }
            }
        } else if (p < S) {
            for (int k = 0; k < K; k++) {
                const double *z = pts + ((size_t)(p - pidx0) * K + k) * mp.d;
                const double wk = w[(size_t)(p - pidx0) * K + k];
                for (int i = 0; i < mp.m; i++) {
                    double pc[MF::NPC], varfac;
                    const bool trivial = MF::prep_row(z, i, mp, pc, varfac);
                    const double om = wk * varfac;
                    sw += om;
                    if (om == 0.0) continue;  // zero weight: contributes nothing, not even rank
                    if (!trivial) {
                        const double sq = sqrt(om);
                        if (MF::FAST_EXP && dcmax) bound = fmax(bound, MF::arg_bound(pc, dcmax));
                        MF::post_row(pc, sq);
                        rows[na * RW] = sq;
#pragma unroll
                        for (int c = 0; c < MF::NPC; c++) rows[na * RW + 1 + c] = pc[c];
                        na++;
                    } else {
                        ntriv++;
                    }
                }
            }
        }
        int rank = na + ((MF::CONST_COL >= 0 && ntriv > 0) ? 1 : 0);
        // the hoisted first pivot must be a positive, finite number; otherwise the design is singular
        // (or its weights are NaN) and every draw takes the rank-deficient path
        if (MF::CONST_COL >= 0 && !(sw > 0.0 && sw < KST_INF)) rank = 0;
        if (KstDense<MF>::value && !dense_ok) rank = 0;  // a unit covariance that is not positive definite
        if (na == 0 && KstDense<MF>::value) {
            // no unit left: one weightless unit (L = 0) keeps the loop shape
            double *lm = KstTile<MF>::lmat(smem, rows_max) + (size_t)threadIdx.x * rows_max * KstTile<MF>::LW;
            for (int i = 0; i < mp.m * (mp.m + 1) / 2; i++) lm[i] = 0.0;
            const double one = 1.0;
            for (int i = 0; i < mp.m; i++) {
                double pc[MF::NPC], vf;
                MF::prep_row(&one, i, mp, pc, vf);
                MF::post_row(pc, 1.0);
                rows[i * RW] = 1.0;
#pragma unroll
                for (int c = 0; c < MF::NPC; c++) rows[i * RW + 1 + c] = pc[c];
            }
            na = mp.m;
            rank = 0;
        }
        if (na == 0) {
            // the evaluation loop always runs the first row: give it a weightless, harmless one
            double pc[MF::NPC], vf;
            const double one = 1.0;
            MF::prep_row(&one, 1, mp, pc, vf);
            MF::post_row(pc, 0.0);
            rows[0] = 0.0;
#pragma unroll
            for (int c = 0; c < MF::NPC; c++) rows[1 + c] = pc[c];
            na = 1;
        }
        KstTile<MF>::sumw(smem, rows_max)[threadIdx.x] = sw;
        KstTile<MF>::isumw(smem, rows_max)[threadIdx.x] = 1.0 / sw;
        KstTile<MF>::nact(smem, rows_max)[threadIdx.x] = na;
        KstTile<MF>::rank(smem, rows_max)[threadIdx.x] = rank;
        // (a NaN bound compares false: not safe)
        KstTile<MF>::safe(smem, rows_max)[threadIdx.x] = (MF::FAST_EXP && dcmax && bound <= 700.0) ? 1 : 0;
    }
    __syncthreads();
}

// kst_build_tile with the per-row work -- map_to_covariate!, the covariate-only
// transcendentals, the square root of the row weight -- spread over one thread per (particle, row); one thread per
// particle then compacts the rows in order and sums the weights in the reference's order, so the tile is the one
// kst_build_tile builds, bit for bit.  pts / w: the tile's particles ([np][K][d], [np][K]); s_om / s_bd: scratch
// of tile_p * K * m doubles each.  In the serial version the K * m logarithms of a sigmoid Emax particle are one
// thread's dependent chain, which costs more than the evaluation itself on the reference's problem sizes.
template <class MF>
__device__ __forceinline__ void kst_build_tile_par(double *smem, const double *pts, const double *w, int np, int K,
                                                   const KstModelParams &mp, int tile_p, const double *dcmax, double *s_om,
                                                   double *s_bd)
{
    const int m = mp.m, rows_max = K * m;
    constexpr int RW = KstTile<MF>::RW;
    double *rows_all = KstTile<MF>::rows(smem);
    for (int e = threadIdx.x; e < np * rows_max; e += KST_NT) {
        const int i = e / rows_max, q = e % rows_max, k = q / m, u = q % m;
        double pc[MF::NPC], varfac;
        const bool trivial = MF::prep_row(pts + ((size_t)i * K + k) * mp.d, u, mp, pc, varfac);
        const double om = w[(size_t)i * K + k] * varfac;
        double sq = 0.0, bound = 0.0;
        if (om != 0.0 && !trivial) {
            sq = sqrt(om);
            if (MF::FAST_EXP) bound = MF::arg_bound(pc, dcmax);
            MF::post_row(pc, sq);
        }
        double *r = rows_all + ((size_t)i * rows_max + q) * RW;
        r[0] = trivial ? -1.0 : sq;  // sq >= 0 for a live row; zero-weight rows are recognised by om == 0
#pragma unroll
        for (int c = 0; c < MF::NPC; c++) r[1 + c] = pc[c];
        s_om[e] = om;
        s_bd[e] = bound;
    }
    __syncthreads();
    if ((int)threadIdx.x < tile_p) {
        const int i = threadIdx.x;
        double *rows = rows_all + (size_t)i * rows_max * RW;
        int na = 0, ntriv = 0;
        double sw = 0.0, bound = 0.0;
        if (i < np) {
            for (int q = 0; q < rows_max; q++) {
                const double om = s_om[i * rows_max + q];
                sw += om;
                if (om == 0.0) continue;
                if (!(rows[q * RW] < 0.0)) {  // live row (a NaN weight stays a live row, as in kst_build_tile)
                    bound = fmax(bound, s_bd[i * rows_max + q]);
                    if (na != q) {
#pragma unroll
                        for (int c = 0; c < RW; c++) rows[na * RW + c] = rows[q * RW + c];
                    }
                    na++;
                } else {
                    ntriv++;
                }
            }
        }
        int rank = na + ((MF::CONST_COL >= 0 && ntriv > 0) ? 1 : 0);
        if (MF::CONST_COL >= 0 && !(sw > 0.0 && sw < KST_INF)) rank = 0;
        if (na == 0) {
            double pc[MF::NPC], vf;
            const double one = 1.0;
            MF::prep_row(&one, 1, mp, pc, vf);
            MF::post_row(pc, 0.0);
            rows[0] = 0.0;
#pragma unroll
            for (int c = 0; c < MF::NPC; c++) rows[1 + c] = pc[c];
            na = 1;
        }
        KstTile<MF>::sumw(smem, rows_max)[i] = sw;
        KstTile<MF>::isumw(smem, rows_max)[i] = 1.0 / sw;
        KstTile<MF>::nact(smem, rows_max)[i] = na;
        KstTile<MF>::rank(smem, rows_max)[i] = rank;
        KstTile<MF>::safe(smem, rows_max)[i] = (MF::FAST_EXP && dcmax && bound <= 700.0) ? 1 : 0;
    }
    __syncthreads();
}

// one row record (RW doubles at a 32-bit shared-space address) -> weighted row g
template <class MF, bool CLAMP = true>
__device__ __forceinline__ void kst_row_at(unsigned addr, const double *dcv, double *g, KstTab tab)
{
    constexpr int RW = KstTile<MF>::RW;
    double rw[RW];
#pragma unroll
    for (int c = 0; c < RW; c += 2)
        asm("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(rw[c]), "=d"(rw[c + 1]) : "r"(addr + 8u * c));
    MF::template row<CLAMP>(rw + 1, dcv, rw[0], g, tab);
}

// NR consecutive rows: all reduced rows first, then all rank-one updates, so that the NR exp / reciprocal
// chains are independent instruction streams the scheduler can interleave
template <class MF, int NR, bool CLAMP = true>
__device__ __forceinline__ void kst_rows(double *M, unsigned addr, const double *dcv, KstTab tab)
{
    constexpr int R = MF::R, RW = KstTile<MF>::RW;
    double g[NR][R];
#pragma unroll
    for (int k = 0; k < NR; k++) kst_row_at<MF, CLAMP>(addr + (unsigned)(k * RW * 8), dcv, g[k], tab);
#pragma unroll
    for (int k = 0; k < NR; k++) kst_accum<R, MF::CONST_COL>(M, g[k]);
}

// Two draws per thread (KstDraws<MF>::value == 2): every row record is loaded once and evaluated for both draws, so
// the shared-memory load, the loop bookkeeping and the per-draw set-up are paid once per pair.  NR rows x 2 draws
// give the independent chains.
#ifndef KST_F2_INIT
#define KST_F2_INIT 0   // 1: the first row block of a pair of draws initialises the accumulators instead of adding to zeros
#endif
#ifndef KST_F2_ROWS
#define KST_F2_ROWS 0   // > 0: rows per block in the two-draw loop (default: half the family's ROWS)
#endif
template <class MF, int NR, bool CLAMP = true, bool INIT = false>
__device__ __forceinline__ void kst_rows2(double *MA, double *MB, unsigned addr, const double *dcA, const double *dcB,
                                          KstTab tab)
{
    constexpr int R = MF::R, RW = KstTile<MF>::RW;
    double gA[NR][R], gB[NR][R];
#pragma unroll
    for (int k = 0; k < NR; k++) {
        double rw[RW];
#pragma unroll
        for (int c = 0; c < RW; c += 2)
            asm("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(rw[c]), "=d"(rw[c + 1]) : "r"(addr + (unsigned)(k * RW * 8) + 8u * c));
        MF::template row<CLAMP>(rw + 1, dcA, rw[0], gA[k], tab);
        MF::template row<CLAMP>(rw + 1, dcB, rw[0], gB[k], tab);
    }
#pragma unroll
    for (int k = 0; k < NR; k++) {
        if (INIT && k == 0) {
            kst_accum_init<R, MF::CONST_COL>(MA, gA[k]);
            kst_accum_init<R, MF::CONST_COL>(MB, gB[k]);
        } else {
            kst_accum<R, MF::CONST_COL>(MA, gA[k]);
            kst_accum<R, MF::CONST_COL>(MB, gB[k]);
        }
    }
}

template <class MF, bool CLAMP = true>
__device__ __forceinline__ void kst_fisher2(double *MA, double *MB, unsigned rows_addr, int na, double sw,
                                            const double *dcA, const double *dcB, KstTab tab)
{
    constexpr int R = MF::R, RW = KstTile<MF>::RW, UN = KST_F2_ROWS > 0 ? KST_F2_ROWS : (MF::ROWS >= 4 ? MF::ROWS / 2 : MF::ROWS);  // two rows x two draws
    static_assert(UN == 1 || UN == 2 || UN == 4, "rows per block");
    unsigned addr = rows_addr;
    int left = na;
#if KST_F2_INIT
    // the first block (the odd row when there is one, else a full block; na >= 1) sets the accumulators
    if (UN > 1 && (left & 1)) { kst_rows2<MF, 1, CLAMP, true>(MA, MB, addr, dcA, dcB, tab); addr += RW * 8; left -= 1; }
    else if (UN > 2 && (left & 2)) { kst_rows2<MF, 2, CLAMP, true>(MA, MB, addr, dcA, dcB, tab); addr += 2 * RW * 8; left -= 2; }
    else { kst_rows2<MF, UN, CLAMP, true>(MA, MB, addr, dcA, dcB, tab); addr += UN * RW * 8; left -= UN; }
    if (UN > 2 && (left & 2)) { kst_rows2<MF, 2, CLAMP>(MA, MB, addr, dcA, dcB, tab); addr += 2 * RW * 8; left -= 2; }
#else
#pragma unroll
    for (int i = 0; i < KstPacked<R>::N; i++) MA[i] = MB[i] = 0.0;
    if (UN > 1 && (left & 1)) { kst_rows2<MF, 1, CLAMP>(MA, MB, addr, dcA, dcB, tab); addr += RW * 8; left -= 1; }
    if (UN > 2 && (left & 2)) { kst_rows2<MF, 2, CLAMP>(MA, MB, addr, dcA, dcB, tab); addr += 2 * RW * 8; left -= 2; }
#endif
#pragma unroll 1
    for (; left > 0; left -= UN, addr += UN * RW * 8) kst_rows2<MF, UN, CLAMP>(MA, MB, addr, dcA, dcB, tab);
    constexpr int cc = MF::CONST_COL >= 0 ? MF::CONST_COL : 0;
    if (MF::CONST_COL >= 0) MA[kst_pidx<R>(cc, cc)] = MB[kst_pidx<R>(cc, cc)] = sw;
}

// Mt = sum over the particle's live rows (at least one) of g g'; rows_addr is a shared-space address.
// Remainders first (1, 2, 4 rows by the bits of na), then groups of MF::ROWS rows: the FP64 pipe only
// saturates with about four independent dependency chains per warp (tools/fp64_probe.cu), which the
// interleaved rows provide; there is no remainder code after the loop.
// Dense unit covariance (KstDense): `na` rows in units of m; per unit the m Jacobian rows are evaluated, whitened with
// the unit's packed factor Lw (shared-space address laddr, m (m + 1) / 2 doubles per unit) and accumulated:
// Mt += (Lw G)'(Lw G) -- average_fishermatrix!'s trsm! + syrk! (src/model-nonlinear-regression.jl:79-83).
template <class MF, bool CLAMP = true>
__device__ __forceinline__ void kst_fisher(double *M, unsigned rows_addr, int na, double sw, const double *dcv,
                                           KstTab tab, unsigned laddr = 0, int m = 1)
{
    constexpr int R = MF::R, RW = KstTile<MF>::RW, UN = KST_ROWS_OVERRIDE > 0 ? KST_ROWS_OVERRIDE : MF::ROWS;
#pragma unroll
    for (int i = 0; i < KstPacked<R>::N; i++) M[i] = 0.0;
    if constexpr (KstDense<MF>::value) {
        unsigned ra = rows_addr, la = laddr;
        for (int u = 0; u < na; u += m) {
            double x[KST_MAX_UNIT][R];
            for (int i = 0; i < m; i++) kst_row_at<MF, CLAMP>(ra + (unsigned)(i * RW * 8), dcv, x[i], tab);
            // last row first: rows l < i are still the raw Jacobian rows when row i needs them
            for (int i = m - 1; i >= 0; i--) {
                const unsigned li = la + (unsigned)(i * (i + 1) / 2 * 8);
                double l;
                asm("ld.shared.f64 %0, [%1];" : "=d"(l) : "r"(li + (unsigned)(i * 8)));
                double acc[R];
#pragma unroll
                for (int c = 0; c < R; c++) acc[c] = l * x[i][c];
                for (int k = 0; k < i; k++) {
                    asm("ld.shared.f64 %0, [%1];" : "=d"(l) : "r"(li + (unsigned)(k * 8)));
#pragma unroll
                    for (int c = 0; c < R; c++) acc[c] = fma(l, x[k][c], acc[c]);
                }
#pragma unroll
                for (int c = 0; c < R; c++) x[i][c] = acc[c];
            }
            for (int i = 0; i < m; i++) kst_accum<R, -1>(M, x[i]);
            ra += (unsigned)(m * RW * 8);
            la += (unsigned)(m * (m + 1) / 2 * 8);
        }
        return;
    }
    unsigned addr = rows_addr;
    int left = na;
    if (UN > 1 && (left & 1)) { kst_rows<MF, 1, CLAMP>(M, addr, dcv, tab); addr += RW * 8; left -= 1; }
    if (UN > 2 && (left & 2)) { kst_rows<MF, 2, CLAMP>(M, addr, dcv, tab); addr += 2 * RW * 8; left -= 2; }
    if (UN > 4 && (left & 4)) { kst_rows<MF, 4, CLAMP>(M, addr, dcv, tab); addr += 4 * RW * 8; left -= 4; }
#pragma unroll 1
    for (; left > 0; left -= UN, addr += UN * RW * 8) kst_rows<MF, UN, CLAMP>(M, addr, dcv, tab);
    constexpr int cc = MF::CONST_COL >= 0 ? MF::CONST_COL : 0;
    if (MF::CONST_COL >= 0) M[kst_pidx<R>(cc, cc)] = sw;
}

template <int V>
struct KstInt { static constexpr int value = V; };
struct KstTrue { static constexpr bool value = true; };
struct KstFalse { static constexpr bool value = false; };

// ------------------------------------------------------------------------------------------
// the hot kernel
// ------------------------------------------------------------------------------------------
// UNI (D criterion, Identity, all prior weights equal): sum_j w log det_j = w log prod_j det_j, so the thread
// multiplies determinants and takes one log per flush instead of one per draw.  The running product is kept
// inside (2^-600, 2^600); a draw whose determinant would leave that range or is not positive definite flushes
// the product and is evaluated on its own by the regular path.
// REG (regularised normal approximation, docs/src/extend-approximation.md:69-79): the per-draw constant matrix Rt_j
// is added to Mt before the criterion; a design with fewer rows than parameters is then no longer singular and
// the first pivot is no longer a per-particle constant.
// CRIT = KST_KCRIT_GD (Gateaux integrand, gateauxderivative! src/designproblem.jl:239-258 for a swarm of directions):
// the "criterion" of a draw is tr_U[A_j M(delta, theta_j)] with the packed, prior-weighted constants of the
// prepared design in prm.aux; the caller subtracts the summed tr B_j.  A one-point direction is rank deficient by
// construction, so the structural singularity shortcut does not apply.
//
// The body of the evaluation is a device function so that the persistent swarm kernel (kst_pso_fused.cuh) runs the
// very same code on a tile it keeps in shared memory: the particle tile of CTA (bx, p0 / tile_p) has been built in
// `smem` (kst_build_tile), `tab` are the staged math tables; the CTA's partial sums for its particles go to
// partials[bx * S + p0 + i].
#if KST_DRAW_OUTER
// Loop order: DRAWS OUTSIDE, PARTICLES INSIDE.  A thread takes its draws in pairs (j, j + stride) -- one draw at a time
// for the families without the two-draw row pass, and for an odd last draw -- and carries each pair through every
// particle of the tile before it moves to the next pair: the draw's constants are fetched once per pair instead of
// once per (particle, pair), the prefetch, loop and pipeline start-up of a draw loop are paid once per tile instead
// of once per particle, and there is no reduction (with its dependent shuffles and logarithm) between particles.
// The price: the per-particle running sum and determinant product live in per-thread shared-memory slots
// (sacc / srun [particle][thread], conflict free) instead of registers: one load and one store per (particle, pair).
// Every thread still adds its draws of a particle in ascending order and the cross-thread reduction is unchanged, so
// the results are bit for bit those of the particle-outer order.
template <class MF, int CRIT, int T, bool UNI, bool REG>
__device__ __forceinline__ void kst_objective_cta(const KstObjParams &prm, double *smem, KstTab tab, const int p0,
                                                  const int bx, const int nbx, double *__restrict__ partials)
{
    constexpr int R = MF::R, RW = KstTile<MF>::RW;
    const int rows_max = prm.K * prm.mp.m;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int np = min(prm.tile_p, prm.S - p0);
    double *wsum = KstTile<MF>::wsum(smem, rows_max);
    const int n = (int)prm.N, stride = nbx * KST_NT, j0 = bx * KST_NT + (int)threadIdx.x;
    unsigned rows_base = (unsigned)__cvta_generic_to_shared(KstTile<MF>::rows(smem));
    unsigned lbase = KstDense<MF>::value ? (unsigned)__cvta_generic_to_shared(KstTile<MF>::lmat(smem, rows_max)) : 0u;
    // shared-space bases opaque to the compiler (see kst_math.cuh: otherwise rebuilt from SR_CgaCtaId before every use)
    asm volatile("" : "+r"(tab), "+r"(rows_base));
    const double *__restrict__ pw = prm.pw;
    const double *__restrict__ dc = prm.dc;
    const unsigned ld = (unsigned)prm.ld;
    constexpr bool INV0 = MF::CONST_COL == 0 && !REG;  // 1 / first pivot hoisted per particle
    constexpr bool DUAL = MF::FAST_EXP && UNI;          // unclamped copy of the row code only where it pays
    constexpr bool PAIRS = KstDraws<MF>::value == 2 && !KstDense<MF>::value;
    const int *s_na = KstTile<MF>::nact(smem, rows_max), *s_rank = KstTile<MF>::rank(smem, rows_max);
    const int *s_safe = KstTile<MF>::safe(smem, rows_max);
    const double *s_sw = KstTile<MF>::sumw(smem, rows_max), *s_isw = KstTile<MF>::isumw(smem, rows_max);
    double *sacc = smem + KstTile<MF>::acc_offset(rows_max) + threadIdx.x;   // [p * KST_NT]
    double *srun = sacc + KST_TILE_P * KST_NT;
    for (int p = 0; p < np; p++) {
        sacc[p * KST_NT] = 0.0;
        if (UNI) srun[p * KST_NT] = 1.0;
    }

    // one particle, one draw (ND = 1) or one pair of draws (ND = 2): Mt, criterion, the thread's running sums
    auto eval = [&](auto clamp_tag, auto nd_tag, const int p, const int jA, const int jB, const double *dcA,
                    const double *dcB, const double pwA, const double pwB) {
        constexpr bool CL = decltype(clamp_tag)::value;
        constexpr int ND = decltype(nd_tag)::value;
        const unsigned rows_addr = rows_base + (unsigned)(p * rows_max * RW * 8);
        const int na = s_na[p];
        const double sw = s_sw[p], inv0 = s_isw[p];
        double MA[KstPacked<R>::N], MB[KstPacked<R>::N];
        if constexpr (ND == 2) {
            kst_fisher2<MF, CL>(MA, MB, rows_addr, na, sw, dcA, dcB, tab);
        } else {
            kst_fisher<MF, CL>(MA, rows_addr, na, sw, dcA, tab, lbase + (unsigned)(p * rows_max * KstTile<MF>::LW * 8), prm.mp.m);
        }
        if (REG) {
#pragma unroll
            for (int i = 0; i < KstPacked<R>::N; i++) {
                MA[i] += __ldg(prm.reg + ((size_t)i * ld + (size_t)jA));
                if (ND == 2) MB[i] += __ldg(prm.reg + ((size_t)i * ld + (size_t)jB));
            }
        }
        double acc = sacc[p * KST_NT];
        if (CRIT == KST_KCRIT_GD) {
            double qA = 0.0, qB = 0.0;
#pragma unroll
            for (int i = 0; i < KstPacked<R>::N; i++) {
                qA = fma(__ldg(prm.aux + ((size_t)i * ld + (size_t)jA)), MA[i], qA);
                if (ND == 2) qB = fma(__ldg(prm.aux + ((size_t)i * ld + (size_t)jB)), MB[i], qB);
            }
            acc += qA;
            if (ND == 2) acc += qB;
            sacc[p * KST_NT] = acc;
        } else if (UNI) {
            double detA, detB = 1.0;
            int signsA, signsB = 0;
            kst_det_signs<R, INV0>(MA, inv0, detA, signsA);
            if (ND == 2) kst_det_signs<R, INV0>(MB, inv0, detB, signsB);
            const double run = srun[p * KST_NT];
            const double pn = ND == 2 ? run * detA * detB : run * detA;
            const unsigned key = ((unsigned)(signsA | signsB) & 0x80000000u) | (unsigned)__double2hiint(pn);
            if (((key >> 20) - 423u) < 1200u) {
                srun[p * KST_NT] = pn;
            } else {
                acc += kst_log(run, tab) + kst_log_checked(detA, signsA, tab);
                if (ND == 2) acc += kst_log_checked(detB, signsB, tab);
                sacc[p * KST_NT] = acc;
                srun[p * KST_NT] = 1.0;
            }
        } else {
            bool excA, excB = false;
            const double vA = kst_criterion<R, (CRIT == KST_KCRIT_GD ? 0 : CRIT), T, INV0>(MA, prm.aux + jA, prm.ld, excA, inv0, tab);
            acc += excA ? -KST_INF : pwA * vA;
            if (ND == 2) {
                const double vB = kst_criterion<R, (CRIT == KST_KCRIT_GD ? 0 : CRIT), T, INV0>(MB, prm.aux + jB, prm.ld, excB, inv0, tab);
                acc += excB ? -KST_INF : pwB * vB;
            }
            sacc[p * KST_NT] = acc;
        }
    };
    // all particles of the tile for one draw / one pair (uniform over the warp: all lanes work on the same particle)
    auto tile = [&](auto nd_tag, const int jA, const int jB, const double *dcA, const double *dcB, const double pwA,
                    const double pwB) {
        constexpr int ND = decltype(nd_tag)::value;
#pragma unroll 1
        for (int p = 0; p < np; p++) {
            if (!REG && CRIT != KST_KCRIT_GD && s_rank[p] < R) {
                // fewer live rows than parameters: singular by construction (one-point designs etc.).  The reference
                // reaches the same -Inf through an exactly zero pivot: weighted by the prior weight for log_det!
                // (0 * -Inf = NaN, as there), unweighted where it throws.  (Dexp: exp(-Inf / t) = 0 for such a draw.)
                double acc = sacc[p * KST_NT];
                acc += T != 0 ? -KST_INF : (CRIT == 0 ? pwA * -KST_INF : (CRIT == 2 ? pwA * 0.0 : -KST_INF));
                if (ND == 2) acc += T != 0 ? -KST_INF : (CRIT == 0 ? pwB * -KST_INF : (CRIT == 2 ? pwB * 0.0 : -KST_INF));
                sacc[p * KST_NT] = acc;
                continue;
            }
            if (DUAL && s_safe[p]) eval(KstFalse{}, nd_tag, p, jA, jB, dcA, dcB, pwA, pwB);
            else eval(KstTrue{}, nd_tag, p, jA, jB, dcA, dcB, pwA, pwB);
        }
    };

    // software prefetch: the next draws' constants are requested before the current ones go through the tile
    double dcn[MF::NDC], pwn = 0.0, dcm[MF::NDC], pwm = 0.0;
    if (j0 < n) {
#pragma unroll
        for (int c = 0; c < MF::NDC; c++) dcn[c] = __ldg(dc + (size_t)(c * ld + (unsigned)j0));
        pwn = UNI ? prm.w_uniform : __ldg(pw + j0);  // (uniform weights: the array is never read)
    }
    int j = j0;
    if constexpr (PAIRS) {
        if (j0 + stride < n) {
#pragma unroll
            for (int c = 0; c < MF::NDC; c++) dcm[c] = __ldg(dc + (size_t)(c * ld + (unsigned)(j0 + stride)));
            pwm = UNI ? prm.w_uniform : __ldg(pw + j0 + stride);
        }
#pragma unroll 1
        for (; j + stride < n; j += 2 * stride) {
            double dcA[MF::NDC], dcB[MF::NDC];
#pragma unroll
            for (int c = 0; c < MF::NDC; c++) { dcA[c] = dcn[c]; dcB[c] = dcm[c]; }
            const double pwA = pwn, pwB = pwm;
            const int jn = j + 2 * stride;
            if (jn < n) {
#pragma unroll
                for (int c = 0; c < MF::NDC; c++) dcn[c] = __ldg(dc + (size_t)(c * ld + (unsigned)jn));
                pwn = UNI ? prm.w_uniform : __ldg(pw + jn);
                if (jn + stride < n) {
#pragma unroll
                    for (int c = 0; c < MF::NDC; c++) dcm[c] = __ldg(dc + (size_t)(c * ld + (unsigned)(jn + stride)));
                    pwm = UNI ? prm.w_uniform : __ldg(pw + jn + stride);
                }
            }
            tile(KstInt<2>{}, j, j + stride, dcA, dcB, pwA, pwB);
        }
    }
    // one draw at a time: everything for the families without pairs, else at most the odd last draw
#pragma unroll 1
    for (; j < n; j += stride) {
        double dcA[MF::NDC];
#pragma unroll
        for (int c = 0; c < MF::NDC; c++) dcA[c] = dcn[c];
        const double pwA = pwn;
        const int jn = j + stride;
        if (jn < n) {
#pragma unroll
            for (int c = 0; c < MF::NDC; c++) dcn[c] = __ldg(dc + (size_t)(c * ld + (unsigned)jn));
            pwn = UNI ? prm.w_uniform : __ldg(pw + jn);
        }
        tile(KstInt<1>{}, j, j, dcA, dcA, pwA, pwA);
    }

    // per particle: the thread's sum, then the fixed-order reduction over the CTA's draws (unrolled over the tile's
    // particles: their log + shuffle chains are independent and overlap)
#pragma unroll
    for (int p = 0; p < KST_TILE_P; p++) {
        if (p < np) {
            double acc = sacc[p * KST_NT];
            if (UNI && (REG || CRIT == KST_KCRIT_GD || s_rank[p] >= R)) acc = prm.w_uniform * (acc + kst_log(srun[p * KST_NT], tab));
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
            if (lane == 0) wsum[p * (KST_NT / 32) + warp] = acc;
        }
    }
    __syncthreads();
    if (threadIdx.x < np) {
        double s = 0.0;
#pragma unroll
        for (int wi = 0; wi < KST_NT / 32; wi++) s += wsum[threadIdx.x * (KST_NT / 32) + wi];
        partials[(size_t)bx * prm.S + (p0 + threadIdx.x)] = s;  // [nbx][S]: whoever finishes the sums reads coalesced
    }
}
#else
template <class MF, int CRIT, int T, bool UNI, bool REG>
__device__ __forceinline__ void kst_objective_cta(const KstObjParams &prm, double *smem, KstTab tab, const int p0,
                                                  const int bx, const int nbx, double *__restrict__ partials)
{
    constexpr int R = MF::R;
    const int rows_max = prm.K * prm.mp.m;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int np = min(prm.tile_p, prm.S - p0);
    double *wsum = KstTile<MF>::wsum(smem, rows_max);
    // 32-bit draw indices (the launcher guarantees N + stride < 2^31) and 32-bit shared-space addresses:
    // the loop bookkeeping is a handful of integer instructions
    const int n = (int)prm.N, stride = nbx * KST_NT, j0 = bx * KST_NT + (int)threadIdx.x;
    unsigned rows_base = (unsigned)__cvta_generic_to_shared(KstTile<MF>::rows(smem));
    // make both shared-space bases opaque: otherwise the compiler rebuilds them from SR_CgaCtaId (a
    // variable-latency S2R/S2UR + LEA) in front of every use instead of keeping them in a register
    asm volatile("" : "+r"(tab), "+r"(rows_base));
    const double *__restrict__ pw = prm.pw;
    const double *__restrict__ dc = prm.dc;
    const unsigned ld = (unsigned)prm.ld;
    constexpr bool PAIRS = KstDraws<MF>::value == 2 && !KstDense<MF>::value;

#pragma unroll 1
    for (int p = 0; p < np; p++) {
        const unsigned rows_addr = rows_base + (unsigned)(p * rows_max * KstTile<MF>::RW * 8);
        const unsigned laddr = KstDense<MF>::value
                                   ? (unsigned)__cvta_generic_to_shared(KstTile<MF>::lmat(smem, rows_max)) +
                                         (unsigned)(p * rows_max * KstTile<MF>::LW * 8)
                                   : 0u;
        const int na = KstTile<MF>::nact(smem, rows_max)[p];
        const double sw = KstTile<MF>::sumw(smem, rows_max)[p];
        // 1 / first pivot, hoisted per particle when the first pivot is the constant-column entry
        constexpr bool INV0 = MF::CONST_COL == 0 && !REG;
        constexpr bool DUAL = MF::FAST_EXP && UNI;  // the unclamped copy of the draw loop only where it pays
        const double inv0 = KstTile<MF>::isumw(smem, rows_max)[p];
        double acc = 0.0;
        if (!REG && CRIT != KST_KCRIT_GD && KstTile<MF>::rank(smem, rows_max)[p] < R) {
            // fewer live rows than parameters: singular by construction (one-point designs etc.).  The
            // reference reaches the same -Inf through an exactly zero pivot: weighted by the prior weight
            // for log_det! (0 * -Inf = NaN, as there), unweighted where it throws.
            // (Dexp: exp(-Inf / t) = 0 for such a draw.)
            for (int j = j0; j < n; j += stride)
                acc += T != 0 ? -KST_INF : (CRIT == 0 ? pw[j] * -KST_INF : (CRIT == 2 ? pw[j] * 0.0 : -KST_INF));
        } else {
            double run = 1.0;  // UNI: running product of determinants
            // software prefetch: the next draw's constants are requested before this draw's rows are evaluated
            // (asking for the thread's first draws again at the end of a particle, so that the next particle starts
            // without a load, measured no gain: +-0.2 %)
            double dcn[MF::NDC], pwn = 0.0;   // next draw of the thread
            double dcm[MF::NDC], pwm = 0.0;   // PAIRS: the pair's second draw
            if (j0 < n) {
#pragma unroll
                for (int c = 0; c < MF::NDC; c++) dcn[c] = __ldg(dc + (size_t)(c * ld + (unsigned)j0));
                pwn = __ldg(pw + j0);
            }
            if (PAIRS && j0 + stride < n) {
#pragma unroll
                for (int c = 0; c < MF::NDC; c++) dcm[c] = __ldg(dc + (size_t)(c * ld + (unsigned)(j0 + stride)));
                pwm = __ldg(pw + j0 + stride);
            }
            // the draw loop, compiled twice for the kernels worth it: with and without the exponent clamp inside exp
            auto draws = [&](auto clamp_tag, const int jstart) {
            constexpr bool CL = decltype(clamp_tag)::value;
            for (int j = jstart; j < n; j += stride) {
                double dcv[MF::NDC];
#pragma unroll
                for (int c = 0; c < MF::NDC; c++) dcv[c] = dcn[c];
                const double pwj = pwn;
                const int jn = j + stride;
                if (jn < n) {
#pragma unroll
                    for (int c = 0; c < MF::NDC; c++) dcn[c] = __ldg(dc + (size_t)(c * ld + (unsigned)jn));
                    pwn = __ldg(pw + jn);
                }
                double M[KstPacked<R>::N];
                kst_fisher<MF, CL>(M, rows_addr, na, sw, dcv, tab, laddr, prm.mp.m);
                if (REG) {
#pragma unroll
                    for (int i = 0; i < KstPacked<R>::N; i++) M[i] += __ldg(prm.reg + ((size_t)i * ld + (size_t)j));
                }
                if (CRIT == KST_KCRIT_GD) {
                    double q = 0.0;
#pragma unroll
                    for (int i = 0; i < KstPacked<R>::N; i++)
                        q = fma(__ldg(prm.aux + ((size_t)i * ld + (size_t)j)), M[i], q);
                    acc += q;
                } else if (UNI) {
                    double det;
                    int signs;
                    kst_det_signs<R, INV0>(M, inv0, det, signs);
                    const double pn = run * det;
                    const unsigned key = ((unsigned)signs & 0x80000000u) | (unsigned)__double2hiint(pn);
                    if (((key >> 20) - 423u) < 1200u) {
                        run = pn;
                    } else {
                        acc += kst_log(run, tab) + kst_log_checked(det, signs, tab);
                        run = 1.0;
                    }
                } else {
                    bool exc;
                    const double val = kst_criterion<R, (CRIT == KST_KCRIT_GD ? 0 : CRIT), T, INV0>(M, prm.aux + j, prm.ld, exc, inv0, tab);
                    acc += exc ? -KST_INF : pwj * val;
                }
            }
            };
            // The same loop over PAIRS of draws (j, j + stride): one pass over the particle's rows serves both.  Returns
            // the first draw it has not done: a thread with an odd number of draws leaves its last one to `draws`.
            auto draws2 = [&](auto clamp_tag) -> int {
            constexpr bool CL = decltype(clamp_tag)::value;
            constexpr bool liveB = true;
            int j = j0;
            for (; j + stride < n; j += 2 * stride) {
                double dcA[MF::NDC], dcB[MF::NDC];
#pragma unroll
                for (int c = 0; c < MF::NDC; c++) { dcA[c] = dcn[c]; dcB[c] = dcm[c]; }
                const double pwA = pwn, pwB = pwm;
                const int jB = j + stride;
                const int jn = j + 2 * stride;
                if (jn < n) {
#pragma unroll
                    for (int c = 0; c < MF::NDC; c++) dcn[c] = __ldg(dc + (size_t)(c * ld + (unsigned)jn));
                    pwn = __ldg(pw + jn);
                    if (jn + stride < n) {
#pragma unroll
                        for (int c = 0; c < MF::NDC; c++) dcm[c] = __ldg(dc + (size_t)(c * ld + (unsigned)(jn + stride)));
                        pwm = __ldg(pw + jn + stride);
                    }
                }
                double MA[KstPacked<R>::N], MB[KstPacked<R>::N];
                kst_fisher2<MF, CL>(MA, MB, rows_addr, na, sw, dcA, dcB, tab);
                const int jBl = liveB ? jB : j;
                if (REG) {
#pragma unroll
                    for (int i = 0; i < KstPacked<R>::N; i++) {
                        MA[i] += __ldg(prm.reg + ((size_t)i * ld + (size_t)j));
                        MB[i] += __ldg(prm.reg + ((size_t)i * ld + (size_t)jBl));
                    }
                }
                if (CRIT == KST_KCRIT_GD) {
                    double qA = 0.0, qB = 0.0;
#pragma unroll
                    for (int i = 0; i < KstPacked<R>::N; i++) {
                        qA = fma(__ldg(prm.aux + ((size_t)i * ld + (size_t)j)), MA[i], qA);
                        qB = fma(__ldg(prm.aux + ((size_t)i * ld + (size_t)jBl)), MB[i], qB);
                    }
                    acc += qA;
                    if (liveB) acc += qB;
                } else if (UNI) {
                    double detA, detB;
                    int signsA, signsB;
                    kst_det_signs<R, INV0>(MA, inv0, detA, signsA);
                    kst_det_signs<R, INV0>(MB, inv0, detB, signsB);
                        const double pn = run * detA * detB;
                    const unsigned key = ((unsigned)(signsA | signsB) & 0x80000000u) | (unsigned)__double2hiint(pn);
                    if (((key >> 20) - 423u) < 1200u) {
                        run = pn;
                    } else {
                        acc += kst_log(run, tab) + kst_log_checked(detA, signsA, tab);
                        if (liveB) acc += kst_log_checked(detB, signsB, tab);
                        run = 1.0;
                    }
                } else {
                    bool excA, excB;
                    const double vA = kst_criterion<R, (CRIT == KST_KCRIT_GD ? 0 : CRIT), T, INV0>(MA, prm.aux + j, prm.ld, excA, inv0, tab);
                    const double vB = kst_criterion<R, (CRIT == KST_KCRIT_GD ? 0 : CRIT), T, INV0>(MB, prm.aux + jBl, prm.ld, excB, inv0, tab);
                    acc += excA ? -KST_INF : pwA * vA;
                    if (liveB) acc += excB ? -KST_INF : pwB * vB;
                }
            }
            return j;
            };
            // uniform over the warp: all lanes work on the same particle
            if constexpr (PAIRS) {
                if constexpr (DUAL) {
                    if (KstTile<MF>::safe(smem, rows_max)[p]) draws(KstFalse{}, draws2(KstFalse{}));
                    else draws(KstTrue{}, draws2(KstTrue{}));
                } else {
                    draws(KstTrue{}, draws2(KstTrue{}));
                }
            } else if constexpr (DUAL) {
                if (KstTile<MF>::safe(smem, rows_max)[p]) draws(KstFalse{}, j0);
                else draws(KstTrue{}, j0);
            } else {
                draws(KstTrue{}, j0);
            }
            if (UNI) acc = prm.w_uniform * (acc + kst_log(run, tab));
        }
        // fixed-order reduction over the CTA's draws
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
        if (lane == 0) wsum[p * (KST_NT / 32) + warp] = acc;
    }
    __syncthreads();
    if (threadIdx.x < np) {
        double s = 0.0;
#pragma unroll
        for (int wi = 0; wi < KST_NT / 32; wi++) s += wsum[threadIdx.x * (KST_NT / 32) + wi];
        partials[(size_t)bx * prm.S + (p0 + threadIdx.x)] = s;  // [nbx][S]: whoever finishes the sums reads coalesced
    }
}

#endif

template <class MF, int CRIT, int T, bool UNI, bool REG = false>
__global__ void __launch_bounds__(KST_NT, KST_MINB) k_objective(const KstObjParams prm)
{
    extern __shared__ __align__(16) double smem[];
    __shared__ __align__(16) double s_mathtab[KST_MATH_SMEM];
    const int p0 = blockIdx.y * prm.tile_p;
#if KST_EXP_BIG && KST_TMA_TABLES
    // the math tables arrive by one asynchronous bulk copy while the CTA builds its particle tile
    __shared__ __align__(8) unsigned long long s_mbar;
    KstTab tab = kst_math_init_async(s_mathtab, &s_mbar);
    // (one thread per particle: the per-row builder of the persistent swarm kernel, kst_build_tile_par, was measured
    // here too: -0.6 % at 125 000 draws, +-0 at 10^6, +2 % on the mid-size configs)
    kst_build_tile<MF>(smem, prm.pts, prm.w, p0, prm.S, prm.K, prm.mp, prm.tile_p, prm.dcmax);  // ends with __syncthreads()
    kst_math_wait(&s_mbar);
#else
    KstTab tab = kst_math_init(s_mathtab);
    kst_build_tile<MF>(smem, prm.pts, prm.w, p0, prm.S, prm.K, prm.mp, prm.tile_p, prm.dcmax);  // ends with __syncthreads()
#endif
    kst_objective_cta<MF, CRIT, T, UNI, REG>(prm, smem, tab, p0, (int)blockIdx.x, (int)gridDim.x, prm.partials);
}

// Units with many rows (K * m up to KST_MAX_ROWS): the particle tile plus the static math tables can exceed the
// 48 KB of shared memory a kernel gets without asking
#define KST_OPT_IN(kernel)                                                                              \
    do {                                                                                                \
        if (smem + sizeof(double) * KST_MATH_SMEM > 48 * 1024)                                          \
            KST_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
    } while (0)

template <class MF, int CRIT>
static int launch_obj_t(kst_problem *h, const KstObjParams &prm, dim3 grid, size_t smem)
{
    const int T = h->trafo == KST_TRAFO_IDENTITY ? 0 : h->t;
#define KST_LAUNCH_T(TV)                                                                        \
    case TV:                                                                                    \
        if (TV <= MF::R) {                                                                      \
            if (h->reg) {                                                                       \
                KST_OPT_IN((k_objective<MF, CRIT, (TV <= MF::R ? TV : 0), false, true>));       \
                k_objective<MF, CRIT, (TV <= MF::R ? TV : 0), false, true><<<grid, KST_NT, smem, h->stream>>>(prm); \
            } else {                                                                            \
                KST_OPT_IN((k_objective<MF, CRIT, (TV <= MF::R ? TV : 0), false>));             \
                k_objective<MF, CRIT, (TV <= MF::R ? TV : 0), false><<<grid, KST_NT, smem, h->stream>>>(prm); \
            }                                                                                   \
            return KST_OK;                                                                      \
        }                                                                                       \
        break;
    if (prm.gd) {  // Gateaux integrand of a swarm of one-point directions (CRIT is ignored)
        KST_OPT_IN((k_objective<MF, KST_KCRIT_GD, 0, false>));
        k_objective<MF, KST_KCRIT_GD, 0, false><<<grid, KST_NT, smem, h->stream>>>(prm);
        return KST_OK;
    }
    if (CRIT == 0 && T == 0 && h->uniform_w && !h->reg) {
        KST_OPT_IN((k_objective<MF, 0, 0, true>));
        k_objective<MF, 0, 0, true><<<grid, KST_NT, smem, h->stream>>>(prm);
        return KST_OK;
    }
    switch (T) {
        KST_LAUNCH_T(0)
        KST_LAUNCH_T(1)
        KST_LAUNCH_T(2)
        KST_LAUNCH_T(3)
        KST_LAUNCH_T(4)
        KST_LAUNCH_T(5)
    }
#undef KST_LAUNCH_T
    kst_set_error("unsupported transformation dimension t = %d for r = %d", T, MF::R);
    return KST_EINVAL;
}

// criterion dispatch for one model family; instantiated explicitly in kst_obj_inst_*.cu
template <class MF>
int kst_obj_launch(kst_problem *h, const KstObjParams &prm, dim3 grid, int K)
{
    const size_t smem = KstTile<MF>::bytes(K * h->m);
    return h->crit == KST_CRIT_D   ? launch_obj_t<MF, 0>(h, prm, grid, smem)
           : h->crit == KST_CRIT_A ? launch_obj_t<MF, 1>(h, prm, grid, smem)
                                   : launch_obj_t<MF, 2>(h, prm, grid, smem);
}
#define KST_OBJ_FAMILIES(X) \
    X(KstEmax) X(KstSigEmax) X(KstTpc) X(KstVu) X(KstDtr) X(KstLinLogReg<1>) X(KstLinLogReg<2>) X(KstLinLogReg<3>) X(KstSinLogReg) \
    X(KstVuAr)

