// Device-side pieces of the particle swarm shared by the stand-alone kernels (kst_pso.cu) and the persistent swarm
// kernel (kst_pso_fused.cuh) and the multi-GPU tail (kst_comm.cu): the counter-based random numbers, the device view
// of the swarm state, the velocity update of one component, the constrained move of one particle, and the pieces of
// pso_update_best!.
//
// The move must round exactly like the scalar reference arithmetic (src/pso.jl:143-171,
// src/designmeasure-abstractpoint.jl:164-285, src/designregion.jl:114-135) whatever -fmad setting the including
// translation unit is compiled with: every operation goes through the __d*_rn intrinsics, which the compiler
// never contracts into an FMA.
#pragma once
#include <stdint.h>

#include "../../include/kirstine_b200.h"

#define KST_DINF (__longlong_as_double(0x7ff0000000000000LL))

__device__ __forceinline__ void kst_philox(uint64_t seed, uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                           uint32_t (&o)[4])
{
    uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
    for (int round = 0; round < 10; round++) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        const uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
        c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    o[0] = c0; o[1] = c1; o[2] = c2; o[3] = c3;
}

__device__ __forceinline__ void kst_uniform2(uint64_t seed, uint32_t iter, uint32_t particle, uint32_t comp,
                                             uint32_t stream, double &u0, double &u1)
{
    uint32_t o[4];
    kst_philox(seed, iter, particle, comp, stream, o);
    const uint64_t a = ((uint64_t)o[1] << 32) | o[0], b = ((uint64_t)o[3] << 32) | o[2];
    u0 = (double)(a >> 11) * 0x1.0p-53;
    u1 = (double)(b >> 11) * 0x1.0p-53;
}

struct KstPsoDev {
    int S, K, d;
    double *x_pts, *x_w, *p_pts, *p_w, *v_pts, *v_w, *g_pts, *g_w, *fx, *fp, *fg;
    const int *fixw, *fixp;
    double lb[KST_MAX_D], ub[KST_MAX_D];
    uint64_t seed;
};

// pso_update_velocity! (src/pso.jl:143-164) for ONE component: v = ((v + (p-x) r1 c1) + (g-x) r2 c2) chi.  Components
// are independent of each other, so a caller may spread them over threads.
__device__ __forceinline__ double kst_pso_velocity(double v, double x, double pb, double gb, double r1, double r2,
                                                   double c1, double c2, double chi)
{
    double diffp = __dsub_rn(pb, x), diffg = __dsub_rn(gb, x);
    diffp = __dmul_rn(diffp, r1);
    diffg = __dmul_rn(diffg, r2);
    diffp = __dmul_rn(diffp, c1);
    diffg = __dmul_rn(diffg, c2);
    v = __dadd_rn(v, diffp);
    v = __dadd_rn(v, diffg);
    return __dmul_rn(v, chi);
}

// pso_update_position! (src/pso.jl:166-171) = ap_add!(x, v, constraints) for one particle whose velocity has been
// updated: fixed components, largest feasible step, move, clamp, stop at the boundary.
__device__ __forceinline__ void kst_pso_position(const int K, const int d, double *xw, double *xp, double *vw, double *vp,
                                                 const int *fixw, const int *fixp, const double *lb, const double *ub)
{
    // move_handle_fixed! (src/designmeasure-abstractpoint.jl:184-223)
    double sum_vw_free = 0.0;
    int n_fixw = 0;
    for (int k = 0; k < K; k++) n_fixw += fixw[k] != 0;
    for (int k = 0; k < K - 1; k++) {
        if (fixw[k]) vw[k] = 0.0;
        else sum_vw_free = __dadd_rn(sum_vw_free, vw[k]);
    }
    if (n_fixw != K && fixw[K - 1]) {
        const double mean_free = __ddiv_rn(sum_vw_free, (double)(K - n_fixw));
        for (int k = 0; k < K - 1; k++)
            if (!fixw[k]) vw[k] = __dsub_rn(vw[k], mean_free);
    }
    for (int k = 0; k < K; k++)
        if (fixp[k])
            for (int j = 0; j < d; j++) vp[k * d + j] = 0.0;

    // move_how_far (:225-248) with move_designpoint_how_far (src/designregion.jl:114-126)
    double t = 1.0;
    for (int k = 0; k < K; k++) {
        double tk = 1.0;
        for (int j = 0; j < d; j++) {
            const double x = xp[k * d + j], v = vp[k * d + j];
            const double probe = __dadd_rn(x, __dmul_rn(tk, v));
            const double sl = probe > ub[j] ? __ddiv_rn(__dsub_rn(ub[j], x), v) : tk;
            const double sr = probe < lb[j] ? __ddiv_rn(__dsub_rn(lb[j], x), v) : tk;
            tk = fmin(sl, sr);
        }
        t = fmin(t, tk);
    }
    for (int k = 0; k < K - 1; k++) {
        const double sk = __dadd_rn(xw[k], __dmul_rn(t, vw[k])) < 0 ? __ddiv_rn(-xw[k], vw[k]) : t;
        t = fmin(sk, t);
    }
    {
        const double sum_x = __dsub_rn(1.0, xw[K - 1]);
        double sum_v = 0.0;
        for (int k = 0; k < K - 1; k++) sum_v = __dadd_rn(sum_v, vw[k]);
        const double sk = __dadd_rn(sum_x, __dmul_rn(t, sum_v)) > 1 ? __ddiv_rn(__dsub_rn(1.0, sum_x), sum_v) : t;
        t = fmin(sk, t);
    }
    if (t < 1e-8) t = 0.0;

    // move_add_v! (:250-285) with move_designpoint! (src/designregion.jl:129-135)
    for (int k = 0; k < K; k++)
        for (int j = 0; j < d; j++) {
            const double x = __dadd_rn(xp[k * d + j], __dmul_rn(t, vp[k * d + j]));
            xp[k * d + j] = fmin(fmax(x, lb[j]), ub[j]);
        }
    for (int k = 0; k < K; k++) xw[k] = __dadd_rn(xw[k], __dmul_rn(t, vw[k]));
    double weight_K = 1.0;
    for (int k = 0; k < K - 1; k++) {
        xw[k] = fmin(fmax(xw[k], 0.0), 1.0);
        weight_K = __dsub_rn(weight_K, xw[k]);
    }
    if (!fixw[K - 1]) xw[K - 1] = weight_K;
    xw[K - 1] = fmax(xw[K - 1], 0.0);

    // stop the particle when the boundary was hit
    if (t != 1.0) {
        for (int c = 0; c < K * d; c++) vp[c] = 0.0;
        for (int k = 0; k < K; k++) vw[k] = 0.0;
    }
}

// dst[0..n) = src[0..n) by one thread, eight loads in flight: written as load batch, then store batch, because a warp
// issues in order and a plain `dst[c] = src[c]` loop is one dependent load -> store round trip per element (the
// compiler may not move a load above an earlier store that could alias it): 13 of k_pso_best's 20 us with cold caches.
__device__ __forceinline__ void kst_copy8(double *dst, const double *src, int n)
{
    for (int c0 = 0; c0 < n; c0 += 8) {
        double t[8];
#pragma unroll
        for (int u = 0; u < 8; u++)
            if (c0 + u < n) t[u] = src[c0 + u];
#pragma unroll
        for (int u = 0; u < 8; u++)
            if (c0 + u < n) dst[c0 + u] = t[u];
    }
}

// pso_update_best!: strict '>', sequential semantics => the lowest index attaining the maximum
// among improved particles becomes the global best iff it beats fg; NaN never wins.
// status (init only): 1 = fg == +Inf, 2 = all fx NaN.
// partials != nullptr: the evaluation kernel's per-CTA partial sums are still unfinished; fx[i] is their fixed-order
// sum plus `offset` (what k_obj_finish would have written), computed here to save a launch per iteration.
// The pieces below are shared by k_pso_best (kst_pso.cu) and the multi-GPU tail k_pso_tail (kst_comm.cu), which run as
// CTAs of 128 particles; the persistent swarm kernel (kst_pso_fused.cuh), where every CTA needs the whole result, has
// its own copy of the update.

// arg-max merge: the larger value, the lower index among equals; index -1 = no candidate
__device__ __forceinline__ void kst_argmax_merge(double &va, int &ia, const double vb, const int ib)
{
    if (ib >= 0 && (ia < 0 || vb > va || (vb == va && ib < ia))) { va = vb; ia = ib; }
}
__device__ __forceinline__ void kst_argmax_warp(double &best, int &bi)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const double vb = __shfl_down_sync(0xffffffffu, best, o);
        const int ib = __shfl_down_sync(0xffffffffu, bi, o);
        kst_argmax_merge(best, bi, vb, ib);
    }
}
// arg-max over a CTA of any multiple of 32 threads (<= 1024); result valid in thread 0.  s_val / s_idx: 32 entries.
__device__ __forceinline__ void kst_argmax_cta(double &best, int &bi, double *s_val, int *s_idx)
{
    const int tid = threadIdx.x;
    kst_argmax_warp(best, bi);
    if ((tid & 31) == 0) { s_val[tid >> 5] = best; s_idx[tid >> 5] = bi; }
    __syncthreads();
    if (tid < 32) {
        const int nw = (blockDim.x + 31) >> 5;
        best = tid < nw ? s_val[tid] : -KST_DINF;
        bi = tid < nw ? s_idx[tid] : -1;
        kst_argmax_warp(best, bi);
    }
}

// Requested before a particle's value is finished, needed after: its personal best value, and its position in
// L2 / L1 (cold when a bench flushes L2 between iterations).
__device__ __forceinline__ double kst_pso_best_prefetch(const KstPsoDev &s, const int i, const int init)
{
    asm volatile("prefetch.global.L1 [%0];" ::"l"(s.x_pts + (size_t)i * s.K * s.d));
    asm volatile("prefetch.global.L1 [%0];" ::"l"(s.x_w + (size_t)i * s.K));
    return init ? (-KST_DINF) : s.fp[i];
}
// this rank's part of a particle's objective value: the fixed-order sum of the evaluation kernel's per-CTA partial sums
// plus `offset` (exactly what k_obj_finish computes)
__device__ __forceinline__ double kst_obj_finish_value(const double *__restrict__ partials, const int n_blk, const int S,
                                                       const int i, const double offset, const int nan_fill)
{
    double acc = 0.0;
#pragma unroll 8
    for (int b = 0; b < n_blk; b++) acc += partials[(size_t)b * S + i];  // eight loads in flight, added in order
    return nan_fill ? __longlong_as_double(0x7ff8000000000000LL) : acc + offset;
}
__device__ __forceinline__ void kst_pso_best_update(const KstPsoDev &s, const int i, const int init, const double f,
                                                    const double fpi, bool &improved);

// One particle: finish its objective value (when partials), store it, update the personal best.  Returns the value;
// `improved` says whether it is a candidate for the global best.
__device__ __forceinline__ double kst_pso_best_particle(const KstPsoDev &s, const int i, const int init,
                                                        const double *__restrict__ partials, const int n_blk,
                                                        const double offset, const int nan_fill, bool &improved)
{
    const double fpi = kst_pso_best_prefetch(s, i, init);
    double f;
    if (partials) {
        f = kst_obj_finish_value(partials, n_blk, s.S, i, offset, nan_fill);
        s.fx[i] = f;
    } else {
        f = s.fx[i];
    }
    kst_pso_best_update(s, i, init, f, fpi, improved);
    return f;
}
// the personal-best half of pso_update_best! for one particle whose value f is known (fpi = its personal best value)
__device__ __forceinline__ void kst_pso_best_update(const KstPsoDev &s, const int i, const int init, const double f,
                                                    const double fpi, bool &improved)
{
    const int K = s.K, d = s.d;
    improved = f > fpi;
    if (improved) {
        kst_copy8(s.p_pts + (size_t)i * K * d, s.x_pts + (size_t)i * K * d, K * d);
        kst_copy8(s.p_w + (size_t)i * K, s.x_w + (size_t)i * K, K);
        s.fp[i] = f;
    } else if (init) {
        s.fp[i] = fpi;
    }
}

// The swarm-wide part, by one CTA that knows the arg-max (gv, gi; valid in thread 0), the best value so far (fg_old,
// thread 0) and whether any value was not NaN: global best, trace entry, status.  Ends with the CTA synchronised.
__device__ __forceinline__ void kst_pso_best_finish(const KstPsoDev &s, const double gv, const int gi_in, const double fg_old,
                                                    const int any_notnan, double *tr_fx, double *tr_pts, double *tr_w,
                                                    const int slot, int *status)
{
    __shared__ int s_update, s_gi;
    const int K = s.K, d = s.d, tid = threadIdx.x;
    if (tid == 0) {
        s_gi = gi_in;
        s_update = (gi_in >= 0 && gv > fg_old) ? 1 : 0;
        const double fg_new = s_update ? gv : fg_old;
        *s.fg = fg_new;
        if (tr_fx) tr_fx[slot] = fg_new;
        if (status) *status = fg_new == KST_DINF ? 1 : (any_notnan ? 0 : 2);
    }
    __syncthreads();
    // the global best moves to the winner's position (or stays); the trace entry is written from the same registers
    const int upd = s_update, gi = s_gi;
    if (upd || tr_pts) {
        const double *xp = s.x_pts + (size_t)(upd ? gi : 0) * K * d;
        for (int c = tid; c < K * d; c += blockDim.x) {
            const double v = upd ? xp[c] : s.g_pts[c];
            if (upd) s.g_pts[c] = v;
            if (tr_pts) tr_pts[(size_t)slot * K * d + c] = v;
        }
    }
    if (upd || tr_w) {
        const double *xw = s.x_w + (size_t)(upd ? gi : 0) * K;
        for (int k = tid; k < K; k += blockDim.x) {
            const double v = upd ? xw[k] : s.g_w[k];
            if (upd) s.g_w[k] = v;
            if (tr_w) tr_w[(size_t)slot * K + k] = v;
        }
    }
    __syncthreads();
}

// Scratch of the multi-CTA best update: one candidate per CTA and the arrival counter (zero between launches).
struct KstBestScratch {
    double *val;
    int *idx;       // [n_cta] candidate index, then [n_cta] "some value was not NaN"
    unsigned int *counter;
};

// The swarm-wide half of pso_update_best! for a grid of CTAs that have each updated the personal bests of their own
// particles: (best, bi) = the thread's candidate (its improved particle or -1), notnan = its value was not NaN.
// Every CTA leaves its arg-max candidate in global memory; the CTA that arrives last (one atomic counter) merges the
// candidates -- the merge rule is order independent: largest value, lowest index among equals -- and does the
// swarm-wide part (global best, trace entry, status).  fg_old: the best value so far, read by thread 0 at kernel start.
__device__ __forceinline__ void kst_pso_best_grid(const KstPsoDev &s, double best, int bi, const bool notnan,
                                                  const double fg_old, const KstBestScratch &sc, double *tr_fx,
                                                  double *tr_pts, double *tr_w, const int slot, int *status)
{
    __shared__ double s_val[32];
    __shared__ int s_idx[32];
    __shared__ int s_notnan, s_last;
    const int tid = threadIdx.x, n_cta = gridDim.x;
    if (tid == 0) s_notnan = 0;
    __syncthreads();
    if (__any_sync(0xffffffffu, notnan) && (tid & 31) == 0) atomicOr(&s_notnan, 1);
    kst_argmax_cta(best, bi, s_val, s_idx);
    __syncthreads();
    if (tid == 0) {
        sc.val[blockIdx.x] = best;
        sc.idx[blockIdx.x] = bi;
        sc.idx[n_cta + blockIdx.x] = s_notnan;
        __threadfence();
        s_last = atomicAdd(sc.counter, 1u) == (unsigned)(n_cta - 1);
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    // the last CTA: merge the candidates (written by other SMs: volatile reads)
    best = -KST_DINF;
    bi = -1;
    int nn = 0;
    for (int c = tid; c < n_cta; c += blockDim.x) {
        kst_argmax_merge(best, bi, ((volatile double *)sc.val)[c], ((volatile int *)sc.idx)[c]);
        nn |= ((volatile int *)sc.idx)[n_cta + c];
    }
    nn = __any_sync(0xffffffffu, nn != 0);
    __syncthreads();   // s_notnan, s_val, s_idx are reused
    if (tid == 0) { s_notnan = 0; *sc.counter = 0u; }   // the counter is zero again for the next launch
    __syncthreads();
    if (nn && (tid & 31) == 0) atomicOr(&s_notnan, 1);
    kst_argmax_cta(best, bi, s_val, s_idx);
    __syncthreads();
    kst_pso_best_finish(s, best, bi, fg_old, s_notnan, tr_fx, tr_pts, tr_w, slot, status);
}
