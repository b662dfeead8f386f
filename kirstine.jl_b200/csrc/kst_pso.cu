// Device-resident particle swarm: the swarm state never leaves the GPU between iterations.
//
//   optimizer_state      src/pso.jl:83-125       -> k_pso_init + evaluation + k_pso_best(init)
//   pso_update_velocity! src/pso.jl:143-164  \
//   pso_update_position! src/pso.jl:166-171   >  -> k_pso_move (one thread per particle)
//   ap_add!(p, v, c)     src/designmeasure-abstractpoint.jl:164-285 /
//   pso_update_best!     src/pso.jl:181-193      -> k_pso_best (one CTA, fixed-order arg-max)
//
// Random numbers come from Philox4x32-10 keyed on (seed; iteration, particle, component, stream),
// so every replica of the swarm (one per GPU) advances identically without any broadcast, and the
// CPU oracle can draw the very same numbers.  The reference uses Julia's Xoshiro256++ stream,
// which cannot be reproduced outside Julia: parity of *trajectories* with the reference is
// statistical, parity with the oracle is exact in the random inputs.
#include "kst_common.cuh"

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

// ap_rand!(d, constraints): src/designmeasure-abstractpoint.jl:73-107, random_designpoint!
// src/designregion.jl:104-110.  Points use u0 of component K + k*d + j, weights -log(1-u0) of
// component k (stream = 1 + attempt).
__global__ void k_pso_init(KstPsoDev s, const double *__restrict__ proto_pts, const double *__restrict__ proto_w, int P)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= s.S) return;
    const int K = s.K, d = s.d;
    const double *sp = proto_pts + (size_t)(i < P ? i : 0) * K * d;
    const double *sw = proto_w + (size_t)(i < P ? i : 0) * K;
    double *xp = s.x_pts + (size_t)i * K * d, *xw = s.x_w + (size_t)i * K;
    for (int c = 0; c < K * d; c++) xp[c] = sp[c];
    for (int k = 0; k < K; k++) xw[k] = sw[k];
    if (i >= P) {
        for (int k = 0; k < K; k++)
            if (!s.fixp[k])
                for (int j = 0; j < d; j++) {
                    double u0, u1;
                    kst_uniform2(s.seed, 0, i, K + k * d + j, 1, u0, u1);
                    double v = u0;
                    v *= s.ub[j] - s.lb[j];
                    v += s.lb[j];
                    xp[k * d + j] = v;
                }
        bool all_fixed = true;
        for (int k = 0; k < K; k++) all_fixed = all_fixed && s.fixw[k];
        if (!all_fixed) {
            double fix = 0.0;
            for (int k = 0; k < K; k++)
                if (s.fixw[k]) fix += xw[k];
            const double cum_sum_fix = fmin(1.0, fix);
            double cum_sum_rand = 0.0;
            int attempt = 0;
            while (cum_sum_rand < 0x1.0p-52) {
                for (int k = 0; k < K; k++)
                    if (!s.fixw[k]) {
                        double u0, u1;
                        kst_uniform2(s.seed, 0, i, k, 1 + attempt, u0, u1);
                        xw[k] = -log(1.0 - u0);
                        cum_sum_rand += xw[k];
                    }
                attempt++;
            }
            const double norm_const = (1 - cum_sum_fix) / cum_sum_rand;
            for (int k = 0; k < K; k++)
                if (!s.fixw[k]) xw[k] *= norm_const;
        }
    }
    double *vp = s.v_pts + (size_t)i * K * d, *vw = s.v_w + (size_t)i * K;
    double *pp = s.p_pts + (size_t)i * K * d, *pw = s.p_w + (size_t)i * K;
    for (int c = 0; c < K * d; c++) { vp[c] = 0.0; pp[c] = xp[c]; }
    for (int k = 0; k < K; k++) { vw[k] = 0.0; pw[k] = xw[k]; }
    if (i == 0) {
        for (int c = 0; c < K * d; c++) s.g_pts[c] = xp[c];
        for (int k = 0; k < K; k++) s.g_w[k] = xw[k];
    }
}

// velocity + constrained move of one particle, operation order exactly as the reference
__global__ void k_pso_move(KstPsoDev s, uint32_t iter, double c1, double c2, double chi)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= s.S) return;
    const int K = s.K, d = s.d;
    double *xp = s.x_pts + (size_t)i * K * d, *xw = s.x_w + (size_t)i * K;
    const double *pp = s.p_pts + (size_t)i * K * d, *pw = s.p_w + (size_t)i * K;
    double *vp = s.v_pts + (size_t)i * K * d, *vw = s.v_w + (size_t)i * K;

    // pso_update_velocity!: v = ((v + (p-x) r1 c1) + (g-x) r2 c2) chi, fresh r1, r2 per component
    // (weights first, then atoms: ap_rand!(v, lb, ub), src/designmeasure-abstractpoint.jl:136-144)
    for (int c = 0; c < K * (d + 1); c++) {
        double r1, r2;
        kst_uniform2(s.seed, iter, i, c, 0, r1, r2);
        const double x = c < K ? xw[c] : xp[c - K];
        const double pb = c < K ? pw[c] : pp[c - K];
        const double gb = c < K ? s.g_w[c] : s.g_pts[c - K];
        double *v = c < K ? &vw[c] : &vp[c - K];
        double diffp = pb - x, diffg = gb - x;
        diffp = __dmul_rn(diffp, r1);
        diffg = __dmul_rn(diffg, r2);
        diffp = __dmul_rn(diffp, c1);
        diffg = __dmul_rn(diffg, c2);
        double vv = *v;
        vv = __dadd_rn(vv, diffp);
        vv = __dadd_rn(vv, diffg);
        *v = __dmul_rn(vv, chi);
    }

    // move_handle_fixed! (src/designmeasure-abstractpoint.jl:184-223)
    double sum_vw_free = 0.0;
    int n_fixw = 0;
    for (int k = 0; k < K; k++) n_fixw += s.fixw[k] != 0;
    for (int k = 0; k < K - 1; k++) {
        if (s.fixw[k]) vw[k] = 0.0;
        else sum_vw_free += vw[k];
    }
    if (n_fixw != K && s.fixw[K - 1]) {
        const double mean_free = sum_vw_free / (K - n_fixw);
        for (int k = 0; k < K - 1; k++)
            if (!s.fixw[k]) vw[k] -= mean_free;
    }
    for (int k = 0; k < K; k++)
        if (s.fixp[k])
            for (int j = 0; j < d; j++) vp[k * d + j] = 0.0;

    // move_how_far (:225-248) with move_designpoint_how_far (src/designregion.jl:114-126)
    double t = 1.0;
    for (int k = 0; k < K; k++) {
        double tk = 1.0;
        for (int j = 0; j < d; j++) {
            const double x = xp[k * d + j], v = vp[k * d + j];
            const double probe = __dadd_rn(x, __dmul_rn(tk, v));
            const double sl = probe > s.ub[j] ? (s.ub[j] - x) / v : tk;
            const double sr = probe < s.lb[j] ? (s.lb[j] - x) / v : tk;
            tk = fmin(sl, sr);
        }
        t = fmin(t, tk);
    }
    for (int k = 0; k < K - 1; k++) {
        const double sk = __dadd_rn(xw[k], __dmul_rn(t, vw[k])) < 0 ? -xw[k] / vw[k] : t;
        t = fmin(sk, t);
    }
    {
        const double sum_x = 1.0 - xw[K - 1];
        double sum_v = 0.0;
        for (int k = 0; k < K - 1; k++) sum_v += vw[k];
        const double sk = __dadd_rn(sum_x, __dmul_rn(t, sum_v)) > 1 ? (1 - sum_x) / sum_v : t;
        t = fmin(sk, t);
    }
    if (t < 1e-8) t = 0.0;

    // move_add_v! (:250-285) with move_designpoint! (src/designregion.jl:129-135)
    for (int k = 0; k < K; k++)
        for (int j = 0; j < d; j++) {
            const double x = __dadd_rn(xp[k * d + j], __dmul_rn(t, vp[k * d + j]));
            xp[k * d + j] = fmin(fmax(x, s.lb[j]), s.ub[j]);
        }
    for (int k = 0; k < K; k++) xw[k] = __dadd_rn(xw[k], __dmul_rn(t, vw[k]));
    double weight_K = 1.0;
    for (int k = 0; k < K - 1; k++) {
        xw[k] = fmin(fmax(xw[k], 0.0), 1.0);
        weight_K -= xw[k];
    }
    if (!s.fixw[K - 1]) xw[K - 1] = weight_K;
    xw[K - 1] = fmax(xw[K - 1], 0.0);

    // stop the particle when the boundary was hit
    if (t != 1.0) {
        for (int c = 0; c < K * d; c++) vp[c] = 0.0;
        for (int k = 0; k < K; k++) vw[k] = 0.0;
    }
}

// pso_update_best!: strict '>', sequential semantics => the lowest index attaining the maximum
// among improved particles becomes the global best iff it beats fg; NaN never wins.
// status (init only): 1 = fg == +Inf, 2 = all fx NaN.
// partials != nullptr: the evaluation kernel's per-CTA partial sums are still unfinished; fx[i] is their fixed-order
// sum plus `offset` (what k_obj_finish would have written), computed here to save a launch per iteration.
__global__ void __launch_bounds__(1024) k_pso_best(KstPsoDev s, int init, double *tr_fx, double *tr_pts, double *tr_w,
                                                   int slot, int *status, const double *__restrict__ partials,
                                                   int n_blk, double offset)
{
    __shared__ double s_val[1024];
    __shared__ int s_idx[1024];
    __shared__ int s_notnan;
    const int K = s.K, d = s.d, tid = threadIdx.x;
    if (tid == 0) s_notnan = 0;
    __syncthreads();
    double best = -KST_DINF;
    int bidx = -1;
    bool have = false;
    bool notnan = false;
    for (int i = tid; i < s.S; i += blockDim.x) {
        double f;
        if (partials) {
            double acc = 0.0;
            for (int b = 0; b < n_blk; b++) acc += partials[(size_t)i * n_blk + b];
            f = acc + offset;
            s.fx[i] = f;
        } else {
            f = s.fx[i];
        }
        const double fpi = init ? (-KST_DINF) : s.fp[i];
        notnan = notnan || (f == f);
        if (f > fpi) {
            const double *xp = s.x_pts + (size_t)i * K * d, *xw = s.x_w + (size_t)i * K;
            double *pp = s.p_pts + (size_t)i * K * d, *pw = s.p_w + (size_t)i * K;
            for (int c = 0; c < K * d; c++) pp[c] = xp[c];
            for (int k = 0; k < K; k++) pw[k] = xw[k];
            s.fp[i] = f;
            if (!have || f > best) { best = f; bidx = i; have = true; }
        } else if (init) {
            s.fp[i] = fpi;
        }
    }
    if (notnan) atomicOr(&s_notnan, 1);
    s_val[tid] = best;
    s_idx[tid] = have ? bidx : -1;
    __syncthreads();
    for (int o = blockDim.x / 2; o > 0; o >>= 1) {
        if (tid < o) {
            const int ia = s_idx[tid], ib = s_idx[tid + o];
            const double va = s_val[tid], vb = s_val[tid + o];
            const bool take_b = ib >= 0 && (ia < 0 || vb > va || (vb == va && ib < ia));
            if (take_b) { s_val[tid] = vb; s_idx[tid] = ib; }
        }
        __syncthreads();
    }
    const int gi = s_idx[0];
    const double gv = s_val[0];
    __shared__ int s_update;
    if (tid == 0) {
        const double fg_old = init ? (-KST_DINF) : *s.fg;
        s_update = (gi >= 0 && gv > fg_old) ? 1 : 0;
        const double fg_new = s_update ? gv : fg_old;
        *s.fg = fg_new;
        if (tr_fx) tr_fx[slot] = fg_new;
        if (status) {
            int st = 0;
            if (fg_new == KST_DINF) st = 1;
            *status = st;
        }
    }
    __syncthreads();
    if (s_update) {
        const double *xp = s.x_pts + (size_t)gi * K * d, *xw = s.x_w + (size_t)gi * K;
        for (int c = tid; c < K * d; c += blockDim.x) s.g_pts[c] = xp[c];
        for (int k = tid; k < K; k += blockDim.x) s.g_w[k] = xw[k];
    }
    __syncthreads();
    if (tr_pts)
        for (int c = tid; c < K * d; c += blockDim.x) tr_pts[(size_t)slot * K * d + c] = s.g_pts[c];
    if (tr_w)
        for (int k = tid; k < K; k += blockDim.x) tr_w[(size_t)slot * K + k] = s.g_w[k];
    if (status && tid == 0 && *status == 0 && !s_notnan) *status = 2;
}

static KstPsoDev make_dev(const kst_problem *h)
{
    KstPsoDev s;
    const KstPso &p = h->pso;
    s.S = p.S; s.K = p.K; s.d = h->d;
    s.x_pts = p.x_pts; s.x_w = p.x_w; s.p_pts = p.p_pts; s.p_w = p.p_w; s.v_pts = p.v_pts; s.v_w = p.v_w;
    s.g_pts = p.g_pts; s.g_w = p.g_w; s.fx = p.fx; s.fp = p.fp; s.fg = p.fg;
    s.fixw = p.fixw; s.fixp = p.fixp;
    for (int j = 0; j < KST_MAX_D; j++) { s.lb[j] = h->lb[j]; s.ub[j] = h->ub[j]; }
    s.seed = p.seed;
    return s;
}

void kst_pso_free(kst_problem *h)
{
    KstPso &p = h->pso;
    double **ptrs[] = {&p.x_pts, &p.x_w, &p.p_pts, &p.p_w, &p.v_pts, &p.v_w, &p.g_pts, &p.g_w,
                       &p.fx, &p.fp, &p.fg, &p.tr_fx, &p.tr_pts, &p.tr_w};
    for (auto q : ptrs) {
        if (*q) cudaFree(*q);
        *q = nullptr;
    }
    if (p.fixw) cudaFree(p.fixw);
    if (p.fixp) cudaFree(p.fixp);
    p.fixw = p.fixp = nullptr;
    p.tr_cap = 0;
    p.S = p.K = 0;
}

int kst_pso_alloc(kst_problem *h, int S, int K)
{
    kst_pso_free(h);
    KstPso &p = h->pso;
    const size_t np = (size_t)S * K * h->d, nw = (size_t)S * K;
    KST_CUDA(cudaMalloc(&p.x_pts, 8 * np)); KST_CUDA(cudaMalloc(&p.p_pts, 8 * np)); KST_CUDA(cudaMalloc(&p.v_pts, 8 * np));
    KST_CUDA(cudaMalloc(&p.x_w, 8 * nw)); KST_CUDA(cudaMalloc(&p.p_w, 8 * nw)); KST_CUDA(cudaMalloc(&p.v_w, 8 * nw));
    KST_CUDA(cudaMalloc(&p.g_pts, 8 * (size_t)K * h->d)); KST_CUDA(cudaMalloc(&p.g_w, 8 * (size_t)K));
    KST_CUDA(cudaMalloc(&p.fx, 8 * (size_t)S)); KST_CUDA(cudaMalloc(&p.fp, 8 * (size_t)S)); KST_CUDA(cudaMalloc(&p.fg, 8));
    KST_CUDA(cudaMalloc(&p.fixw, sizeof(int) * K)); KST_CUDA(cudaMalloc(&p.fixp, sizeof(int) * K));
    p.S = S; p.K = K;
    return KST_OK;
}

int kst_launch_pso_init(kst_problem *h, const double *d_proto_pts, const double *d_proto_w, int P)
{
    const KstPsoDev s = make_dev(h);
    k_pso_init<<<(s.S + 127) / 128, 128, 0, h->stream>>>(s, d_proto_pts, d_proto_w, P);
    KST_CUDA(cudaGetLastError());
    return KST_OK;
}

int kst_launch_pso_move(kst_problem *h)
{
    const KstPsoDev s = make_dev(h);
    const double phi = h->pso.c1 + h->pso.c2;
    const double chi = 2 / fabs(2 - phi - sqrt(phi * phi - 4 * phi));  // src/pso.jl:134-135
    k_pso_move<<<(s.S + 127) / 128, 128, 0, h->stream>>>(s, (uint32_t)h->pso.iter, h->pso.c1, h->pso.c2, chi);
    KST_CUDA(cudaGetLastError());
    return KST_OK;
}

int kst_launch_pso_best(kst_problem *h, int init, int trace_slot)
{
    const KstPsoDev s = make_dev(h);
    KstPso &p = h->pso;
    k_pso_best<<<1, 1024, 0, h->stream>>>(s, init, trace_slot >= 0 ? p.tr_fx : nullptr,
                                           trace_slot >= 0 ? p.tr_pts : nullptr, trace_slot >= 0 ? p.tr_w : nullptr,
                                           trace_slot >= 0 ? trace_slot : 0, init ? h->d_flag : nullptr,
                                           p.pending_n_blk ? h->partials : nullptr, p.pending_n_blk, h->offset);
    p.pending_n_blk = 0;
    KST_CUDA(cudaGetLastError());
    return KST_OK;
}
