// Persistent swarm kernel: a whole run of PSO iterations in ONE cooperative launch, for problems whose evaluation
// grid fits one wave of resident CTAs (the reference's own test and vignette sizes: tens to hundreds of particles
// against 10^3 .. 10^4 draws, and both optimizers of every Exchange step).
//
// Why: with one launch each for move, evaluation and best update, such problems are bound by launch latency and by
// the evaluation kernel's fixed prologue (math tables into shared memory), not by arithmetic: 22 us per iteration
// for a 12 us kernel on BASELINE config 1.  Here the tables are staged once, and an iteration is
//     move (tile in shared memory) -> evaluate -> ONE grid barrier -> best update
// with no state crossing a barrier other than the per-CTA partial sums and the moved positions:
//   * CTA (bx, ty) evaluates particle tile ty on draw slice bx, exactly like k_objective (same device function, same
//     grid, same summation order, hence the same bits as the three-launch path);
//   * every CTA of a tile keeps that tile's x, v and personal bests in shared memory and moves them redundantly
//     (counter-based random numbers: identical in every CTA); the CTA with bx == 0 publishes the moved positions
//     in a double-buffered global array before the barrier;
//   * after the barrier every CTA finishes ALL particles' sums from the (double-buffered) partial sums, keeps its
//     own copy of every personal-best value, and finds the same arg-max (pso_update_best!: strict >, lowest index
//     wins ties, NaN never wins); the new global best is read from the published positions.
// Double buffering by iteration parity makes one barrier per iteration enough: a CTA can only overwrite buffer
// `par` two iterations later, after every CTA has passed the barrier in between.
// The barrier is a monotonic counter in global memory; co-residency of the grid is guaranteed by the cooperative
// launch (the launcher checks the occupancy first and reports "does not fit" so that the caller can fall back to
// the three-launch path).
#pragma once
#include "kst_objective_kernel.cuh"
#include "kst_pso_dev.cuh"

#define KST_FUSED_MAX_S 2048      // every CTA keeps all personal-best values in shared memory
#define KST_FUSED_NOFIT 1         // launcher: the grid is not co-resident (or the state does not fit shared memory)

struct KstFusedParams {
    KstObjParams obj;             // pts / w / partials unused: the tile comes from shared memory
    KstPsoDev pso;                // canonical swarm state: read at the start, written at the end
    double *x2_pts, *x2_w;        // published positions, [2][S][K*d] and [2][S][K]
    double *partials2;            // [2][n_blk][S]
    unsigned long long *bar;      // grid barrier counter (monotonic over launches)
    unsigned long long bar_base;  // its value when this launch starts
    int n_steps;                  // swarm evaluations in this launch
    int first_is_init;            // the first evaluation is optimizer_state's (no move before it, fp = fg = -Inf)
    uint32_t iter0;               // iteration counter of the first move (key of the random numbers)
    double c1, c2, chi;
    double offset;                // added to every finished sum ...
    int nan_fill;                 // ... or NaN instead (Gateaux mode at a singular design)
    double *tr_fx, *tr_pts, *tr_w;  // trace of (fg, g) after every evaluation, or nullptr
    int tr_slot0;
    int *status;                  // set after the initial evaluation: 0 ok, 1 fg == +Inf, 2 all NaN
    double *snap;                 // optional packed copy of the state after the initial evaluation:
                                  // x_pts, x_w, p_pts, p_w, v_pts, v_w [S..], g_pts, g_w, fx[S], fp[S], fg
};

// ctr[0] counts arrivals; ctr[1] is an error word: a CTA that waits longer than 5 s gives up (the results are then
// garbage, the host reports the error) instead of hanging the GPU should the grid ever not be co-resident.
__device__ __forceinline__ void kst_grid_barrier(unsigned long long *ctr, unsigned long long target)
{
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(ctr, 1ull);
        unsigned long long v, t0, t1;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        unsigned spins = 0;
        do {
            asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(ctr) : "memory");
            if (v < target && (++spins & 1023u) == 0) {
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
                if (t1 - t0 > 5000000000ull) {
                    atomicExch(ctr + 1, 1ull);
                    break;
                }
            }
        } while (v < target);
        __threadfence();
    }
    __syncthreads();
}

// doubles of swarm state a CTA keeps in shared memory behind the particle tile
__host__ __device__ inline size_t kst_fused_state_doubles(int S, int K, int d, int m, int tile_p)
{
    const size_t KC = (size_t)K * (d + 1);
    return 3 * tile_p * KC      // x, v, p of the tile's particles
           + KC                 // g
           + (size_t)S          // personal-best values of the whole swarm
           + tile_p             // "improved" flags of the tile (as doubles: keeps everything 8-byte aligned)
           + 2 * (size_t)tile_p * K * m;  // per-row scratch of the parallel tile build
}
template <class MF>
__host__ __device__ inline size_t kst_fused_tile_doubles(int rows_max)
{
    return (KstTile<MF>::bytes(rows_max) + 15) / 16 * 2;
}

// number of doubles of the packed state snapshot (KstFusedParams::snap)
__host__ __device__ inline size_t kst_snapshot_doubles(int S, int K, int d)
{
    return (size_t)3 * S * K * (d + 1) + (size_t)K * (d + 1) + 2 * (size_t)S + 1;
}

template <class MF, int CRIT, int T, bool UNI>
__global__ void __launch_bounds__(KST_NT, KST_MINB) k_pso_fused(const KstFusedParams fp)
{
    extern __shared__ __align__(16) double smem[];
    __shared__ __align__(16) double s_mathtab[KST_MATH_SMEM];
    __shared__ double s_rval[KST_NT / 32];
    __shared__ int s_ridx[KST_NT / 32];
    __shared__ double s_fg;
    __shared__ int s_update, s_gi;
    const KstObjParams &prm = fp.obj;
    const KstPsoDev &ps = fp.pso;
    const int K = prm.K, d = prm.mp.d, S = prm.S, tp = prm.tile_p, KD = K * d, KC = K * (d + 1);
    const int rows_max = K * prm.mp.m;
    const int bx = blockIdx.x, nbx = gridDim.x;
    const int p0 = blockIdx.y * tp, np = min(tp, S - p0);
    const bool writer = bx == 0, leader = bx == 0 && blockIdx.y == 0;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const double NaN = __longlong_as_double(0x7ff8000000000000LL);
    KstTab tab = kst_math_init(s_mathtab);

    double *st = smem + kst_fused_tile_doubles<MF>(rows_max);
    double *sx_w = st;            st += tp * K;
    double *sx_p = st;            st += tp * KD;
    double *sv_w = st;            st += tp * K;
    double *sv_p = st;            st += tp * KD;
    double *sp_w = st;            st += tp * K;
    double *sp_p = st;            st += tp * KD;
    double *sg_w = st;            st += K;
    double *sg_p = st;            st += KD;
    double *s_fp = st;            st += S;
    double *s_impr = st;          st += tp;
    double *s_om = st;            st += tp * rows_max;
    double *s_bd = st;

    // the tile's particles, the global best and every personal-best value
    for (int e = tid; e < np * K; e += KST_NT) {
        const size_t gi = (size_t)p0 * K + e;
        sx_w[e] = ps.x_w[gi]; sv_w[e] = ps.v_w[gi]; sp_w[e] = ps.p_w[gi];
    }
    for (int e = tid; e < np * KD; e += KST_NT) {
        const size_t gi = (size_t)p0 * KD + e;
        sx_p[e] = ps.x_pts[gi]; sv_p[e] = ps.v_pts[gi]; sp_p[e] = ps.p_pts[gi];
    }
    for (int e = tid; e < K; e += KST_NT) sg_w[e] = ps.g_w[e];
    for (int e = tid; e < KD; e += KST_NT) sg_p[e] = ps.g_pts[e];
    for (int i = tid; i < S; i += KST_NT) s_fp[i] = fp.first_is_init ? -KST_DINF : ps.fp[i];
    if (tid == 0) s_fg = fp.first_is_init ? -KST_DINF : *ps.fg;
    __syncthreads();

    for (int step = 0; step < fp.n_steps; step++) {
        const bool is_init = fp.first_is_init && step == 0;
        const bool last = step == fp.n_steps - 1;
        const int par = step & 1;
        if (!is_init) {
            // tick!: velocities and constrained move of the tile's particles (src/pso.jl:127-171)
            const uint32_t iter = fp.iter0 + (uint32_t)(step - (fp.first_is_init ? 1 : 0));
            // one thread per (particle, component) for the random numbers and the velocity, then one per particle for
            // the constrained move (its sums run in the reference's order)
            for (int e = tid; e < np * KC; e += KST_NT) {
                const int i = e / KC, c = e % KC;
                double r1, r2;
                kst_uniform2(ps.seed, iter, (uint32_t)(p0 + i), (uint32_t)c, 0, r1, r2);
                const bool isw = c < K;
                const int o = isw ? i * K + c : i * KD + (c - K);
                double *v = isw ? sv_w + o : sv_p + o;
                *v = kst_pso_velocity(*v, isw ? sx_w[o] : sx_p[o], isw ? sp_w[o] : sp_p[o], isw ? sg_w[c] : sg_p[c - K], r1,
                                      r2, fp.c1, fp.c2, fp.chi);
            }
            __syncthreads();
            if (tid < np)
                kst_pso_position(K, d, sx_w + tid * K, sx_p + tid * KD, sv_w + tid * K, sv_p + tid * KD, ps.fixw, ps.fixp,
                                 ps.lb, ps.ub);
            __syncthreads();
        }
        if (writer) {  // the moved positions, for whoever needs the new global best after the barrier
            double *xw = fp.x2_w + ((size_t)par * S + p0) * K, *xp = fp.x2_pts + ((size_t)par * S + p0) * KD;
            for (int e = tid; e < np * K; e += KST_NT) xw[e] = sx_w[e];
            for (int e = tid; e < np * KD; e += KST_NT) xp[e] = sx_p[e];
        }
        // pso_evaluate_objective!: this CTA's draw slice for its tile
        if constexpr (KstDense<MF>::value)  // unit factors: one thread per particle, like everywhere else
            kst_build_tile<MF>(smem, sx_p, sx_w, p0, S, K, prm.mp, tp, prm.dcmax, p0);
        else
            kst_build_tile_par<MF>(smem, sx_p, sx_w, np, K, prm.mp, tp, prm.dcmax, s_om, s_bd);  // ends with __syncthreads()
        double *pt = fp.partials2 + (size_t)par * S * nbx;
        kst_objective_cta<MF, CRIT, T, UNI, false>(prm, smem, tab, p0, bx, nbx, pt);
        kst_grid_barrier(fp.bar, fp.bar_base + (unsigned long long)(step + 1) * (unsigned long long)(nbx * gridDim.y));

        // pso_update_best! (src/pso.jl:181-193), the same in every CTA
        double best = -KST_DINF;
        int bidx = -1;
        bool notnan = false;
        for (int i = tid; i < S; i += KST_NT) {
            double acc = 0.0;
            for (int b = 0; b < nbx; b++) acc += __ldcg(pt + (size_t)b * S + i);
            const double f = fp.nan_fill ? NaN : acc + fp.offset;
            if (leader && (last || (is_init && fp.snap))) ps.fx[i] = f;
            notnan = notnan || (f == f);
            const bool improved = f > s_fp[i];
            if (improved) {
                s_fp[i] = f;
                if (bidx < 0 || f > best) { best = f; bidx = i; }
            }
            if (i >= p0 && i < p0 + np) s_impr[i - p0] = improved ? 1.0 : 0.0;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const double vb = __shfl_down_sync(0xffffffffu, best, o);
            const int ib = __shfl_down_sync(0xffffffffu, bidx, o);
            if (ib >= 0 && (bidx < 0 || vb > best || (vb == best && ib < bidx))) { best = vb; bidx = ib; }
        }
        if (lane == 0) { s_rval[warp] = best; s_ridx[warp] = bidx; }
        const int any_notnan = __syncthreads_or(notnan ? 1 : 0);
        if (tid == 0) {
            double gv = s_rval[0];
            int gi = s_ridx[0];
            for (int wi = 1; wi < KST_NT / 32; wi++) {
                const double vb = s_rval[wi];
                const int ib = s_ridx[wi];
                if (ib >= 0 && (gi < 0 || vb > gv || (vb == gv && ib < gi))) { gv = vb; gi = ib; }
            }
            const double fg_old = s_fg;
            const int upd = (gi >= 0 && gv > fg_old) ? 1 : 0;
            const double fg_new = upd ? gv : fg_old;
            s_fg = fg_new;
            s_update = upd;
            s_gi = gi;
            if (leader) {
                if (fp.tr_fx) fp.tr_fx[fp.tr_slot0 + step] = fg_new;
                if (is_init && fp.status) *fp.status = fg_new == KST_DINF ? 1 : (any_notnan ? 0 : 2);
            }
        }
        __syncthreads();
        if (s_update) {
            const double *xw = fp.x2_w + ((size_t)par * S + s_gi) * K, *xp = fp.x2_pts + ((size_t)par * S + s_gi) * KD;
            for (int e = tid; e < K; e += KST_NT) sg_w[e] = __ldcg(xw + e);
            for (int e = tid; e < KD; e += KST_NT) sg_p[e] = __ldcg(xp + e);
        }
        for (int e = tid; e < np * K; e += KST_NT)
            if (s_impr[e / K] != 0.0) sp_w[e] = sx_w[e];
        for (int e = tid; e < np * KD; e += KST_NT)
            if (s_impr[e / KD] != 0.0) sp_p[e] = sx_p[e];
        __syncthreads();
        if (leader && fp.tr_pts)
            for (int e = tid; e < KD; e += KST_NT) fp.tr_pts[(size_t)(fp.tr_slot0 + step) * KD + e] = sg_p[e];
        if (leader && fp.tr_w)
            for (int e = tid; e < K; e += KST_NT) fp.tr_w[(size_t)(fp.tr_slot0 + step) * K + e] = sg_w[e];
        if (is_init && fp.snap) {
            // the state optimize() keeps as trace_state[1] (src/optimize.jl:51-53), packed
            double *q = fp.snap;
            double *q_xp = q, *q_xw = q_xp + (size_t)S * KD, *q_pp = q_xw + (size_t)S * K, *q_pw = q_pp + (size_t)S * KD;
            double *q_vp = q_pw + (size_t)S * K, *q_vw = q_vp + (size_t)S * KD, *q_gp = q_vw + (size_t)S * K;
            double *q_gw = q_gp + KD, *q_fx = q_gw + K, *q_fp = q_fx + S, *q_fg = q_fp + S;
            if (writer) {
                for (int e = tid; e < np * K; e += KST_NT) {
                    q_xw[(size_t)p0 * K + e] = sx_w[e]; q_pw[(size_t)p0 * K + e] = sp_w[e]; q_vw[(size_t)p0 * K + e] = sv_w[e];
                }
                for (int e = tid; e < np * KD; e += KST_NT) {
                    q_xp[(size_t)p0 * KD + e] = sx_p[e]; q_pp[(size_t)p0 * KD + e] = sp_p[e]; q_vp[(size_t)p0 * KD + e] = sv_p[e];
                }
            }
            if (leader) {
                for (int e = tid; e < KD; e += KST_NT) q_gp[e] = sg_p[e];
                for (int e = tid; e < K; e += KST_NT) q_gw[e] = sg_w[e];
                // fx (NaN and -Inf included) is what this very thread wrote to ps.fx a moment ago
                for (int i = tid; i < S; i += KST_NT) { q_fp[i] = s_fp[i]; q_fx[i] = ps.fx[i]; }
                if (tid == 0) *q_fg = s_fg;
            }
        }
    }

    // back to the canonical state arrays (kst_pso_get_state, further runs)
    if (writer) {
        for (int e = tid; e < np * K; e += KST_NT) {
            const size_t gi = (size_t)p0 * K + e;
            ps.x_w[gi] = sx_w[e]; ps.v_w[gi] = sv_w[e]; ps.p_w[gi] = sp_w[e];
        }
        for (int e = tid; e < np * KD; e += KST_NT) {
            const size_t gi = (size_t)p0 * KD + e;
            ps.x_pts[gi] = sx_p[e]; ps.v_pts[gi] = sv_p[e]; ps.p_pts[gi] = sp_p[e];
        }
    }
    if (leader) {
        for (int e = tid; e < K; e += KST_NT) ps.g_w[e] = sg_w[e];
        for (int e = tid; e < KD; e += KST_NT) ps.g_pts[e] = sg_p[e];
        for (int i = tid; i < S; i += KST_NT) ps.fp[i] = s_fp[i];
        if (tid == 0) *ps.fg = s_fg;
    }
}

template <class MF, int CRIT, int T, bool UNI>
static int kst_fused_go(kst_problem *h, KstFusedParams &fp, dim3 grid, size_t smem)
{
    auto kern = k_pso_fused<MF, CRIT, T, UNI>;
    if (smem + sizeof(double) * KST_MATH_SMEM + 256 > 200 * 1024) return KST_FUSED_NOFIT;
    if (smem + sizeof(double) * KST_MATH_SMEM + 256 > 48 * 1024)
        KST_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int resident = 0;
    KST_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&resident, kern, KST_NT, smem));
    if ((long long)resident * h->sm_count < (long long)grid.x * grid.y) return KST_FUSED_NOFIT;
    void *args[] = {(void *)&fp};
    KST_CUDA(cudaLaunchCooperativeKernel((void *)kern, grid, dim3(KST_NT), args, smem, h->stream));
    return KST_OK;
}

template <class MF, int CRIT>
static int kst_fused_t(kst_problem *h, KstFusedParams &fp, dim3 grid, size_t smem)
{
    const int T = h->trafo == KST_TRAFO_IDENTITY ? 0 : h->t;
    if (CRIT == 0 && T == 0 && h->uniform_w) return kst_fused_go<MF, 0, 0, true>(h, fp, grid, smem);
    switch (T) {
    case 0: return kst_fused_go<MF, CRIT, 0, false>(h, fp, grid, smem);
    case 1: return kst_fused_go<MF, CRIT, 1, false>(h, fp, grid, smem);
    case 2: if (MF::R >= 2) return kst_fused_go<MF, CRIT, (MF::R >= 2 ? 2 : 0), false>(h, fp, grid, smem); break;
    case 3: if (MF::R >= 3) return kst_fused_go<MF, CRIT, (MF::R >= 3 ? 3 : 0), false>(h, fp, grid, smem); break;
    case 4: if (MF::R >= 4) return kst_fused_go<MF, CRIT, (MF::R >= 4 ? 4 : 0), false>(h, fp, grid, smem); break;
    case 5: if (MF::R >= 5) return kst_fused_go<MF, CRIT, (MF::R >= 5 ? 5 : 0), false>(h, fp, grid, smem); break;
    }
    kst_set_error("unsupported transformation dimension t = %d for r = %d", T, MF::R);
    return KST_EINVAL;
}

// One cooperative launch running fp.n_steps swarm evaluations.  KST_FUSED_NOFIT (> 0): this problem needs the
// three-launch path (grid larger than one wave, or the swarm state does not fit beside the particle tile).
template <class MF>
int kst_fused_launch(kst_problem *h, KstFusedParams &fp, dim3 grid)
{
    const int K = fp.obj.K;
    const size_t smem = sizeof(double) * (kst_fused_tile_doubles<MF>(K * h->m) +
                                          kst_fused_state_doubles(fp.obj.S, K, h->d, h->m, fp.obj.tile_p));
    if (fp.obj.gd) return kst_fused_go<MF, KST_KCRIT_GD, 0, false>(h, fp, grid, smem);
    return h->crit == KST_CRIT_D   ? kst_fused_t<MF, 0>(h, fp, grid, smem)
           : h->crit == KST_CRIT_A ? kst_fused_t<MF, 1>(h, fp, grid, smem)
                                   : kst_fused_t<MF, 2>(h, fp, grid, smem);
}
