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
#include "kst_pso_dev.cuh"

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

// velocity + constrained move, one thread per particle (kst_pso_move_particle, kst_pso_dev.cuh)
__global__ void k_pso_move(KstPsoDev s, uint32_t iter, double c1, double c2, double chi)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= s.S) return;
    const int K = s.K, d = s.d;
    kst_pso_move_particle(K, d, s.x_w + (size_t)i * K, s.x_pts + (size_t)i * K * d, s.p_w + (size_t)i * K,
                          s.p_pts + (size_t)i * K * d, s.v_w + (size_t)i * K, s.v_pts + (size_t)i * K * d, s.g_w, s.g_pts,
                          s.fixw, s.fixp, s.lb, s.ub, c1, c2, chi, KstRndPhilox{s.seed, iter, (uint32_t)i});
}

// pso_update_best!: strict '>', sequential semantics => the lowest index attaining the maximum
// among improved particles becomes the global best iff it beats fg; NaN never wins.
// status (init only): 1 = fg == +Inf, 2 = all fx NaN.
// partials != nullptr: the evaluation kernel's per-CTA partial sums are still unfinished; fx[i] is their fixed-order
// sum plus `offset` (what k_obj_finish would have written), computed here to save a launch per iteration.
__global__ void __launch_bounds__(1024) k_pso_best(KstPsoDev s, int init, double *tr_fx, double *tr_pts, double *tr_w,
                                                   int slot, int *status, const double *__restrict__ partials,
                                                   int n_blk, double offset, int nan_fill)
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
            f = nan_fill ? __longlong_as_double(0x7ff8000000000000LL) : acc + offset;
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

void kst_pso_make_dev(const kst_problem *h, KstPsoDev *out)
{
    KstPsoDev &s = *out;
    const KstPso &p = h->pso;
    s.S = p.S; s.K = p.K; s.d = h->d;
    s.x_pts = p.x_pts; s.x_w = p.x_w; s.p_pts = p.p_pts; s.p_w = p.p_w; s.v_pts = p.v_pts; s.v_w = p.v_w;
    s.g_pts = p.g_pts; s.g_w = p.g_w; s.fx = p.fx; s.fp = p.fp; s.fg = p.fg;
    s.fixw = p.fixw; s.fixp = p.fixp;
    for (int j = 0; j < KST_MAX_D; j++) { s.lb[j] = h->lb[j]; s.ub[j] = h->ub[j]; }
    s.seed = p.seed;
}

static KstPsoDev make_dev(const kst_problem *h)
{
    KstPsoDev s;
    kst_pso_make_dev(h, &s);
    return s;
}

void kst_pso_free(kst_problem *h)
{
    KstPso &p = h->pso;
    double **ptrs[] = {&p.x_pts, &p.x_w, &p.p_pts, &p.p_w, &p.v_pts, &p.v_w, &p.g_pts, &p.g_w,
                       &p.fx, &p.fp, &p.fg, &p.tr_fx, &p.tr_pts, &p.tr_w, &p.x2_pts, &p.x2_w, &p.snap};
    for (auto q : ptrs) {
        if (*q) cudaFree(*q);
        *q = nullptr;
    }
    if (p.fixw) cudaFree(p.fixw);
    if (p.fixp) cudaFree(p.fixp);
    p.fixw = p.fixp = nullptr;
    if (p.bar) cudaFree(p.bar);
    p.bar = nullptr;
    p.bar_count = 0;
    p.fused_off = 0;
    p.tr_cap = 0;
    p.S = p.K = 0;
}

// (Re)allocates the swarm for S particles of K points.  A swarm of the same shape as the previous one is reused as
// it is: Exchange runs two optimizers per step, and cudaMalloc / cudaFree cost more than their iterations.
int kst_pso_alloc(kst_problem *h, int S, int K)
{
    KstPso &p = h->pso;
    if (p.S == S && p.K == K && p.x_pts) {
        p.fused_off = 0;
        return KST_OK;
    }
    // the trace buffers do not depend on the swarm's shape: keep them
    double *tfx = p.tr_fx, *tpt = p.tr_pts, *tw = p.tr_w;
    const int tcap = p.tr_cap;
    p.tr_fx = p.tr_pts = p.tr_w = nullptr;
    kst_pso_free(h);
    p.tr_fx = tfx; p.tr_pts = tpt; p.tr_w = tw; p.tr_cap = tcap;
    const size_t np = (size_t)S * K * h->d, nw = (size_t)S * K;
    KST_CUDA(cudaMalloc(&p.x_pts, 8 * np)); KST_CUDA(cudaMalloc(&p.p_pts, 8 * np)); KST_CUDA(cudaMalloc(&p.v_pts, 8 * np));
    KST_CUDA(cudaMalloc(&p.x_w, 8 * nw)); KST_CUDA(cudaMalloc(&p.p_w, 8 * nw)); KST_CUDA(cudaMalloc(&p.v_w, 8 * nw));
    KST_CUDA(cudaMalloc(&p.g_pts, 8 * (size_t)K * h->d)); KST_CUDA(cudaMalloc(&p.g_w, 8 * (size_t)K));
    KST_CUDA(cudaMalloc(&p.fx, 8 * (size_t)S)); KST_CUDA(cudaMalloc(&p.fp, 8 * (size_t)S)); KST_CUDA(cudaMalloc(&p.fg, 8));
    KST_CUDA(cudaMalloc(&p.fixw, sizeof(int) * K)); KST_CUDA(cudaMalloc(&p.fixp, sizeof(int) * K));
    KST_CUDA(cudaMalloc(&p.x2_pts, 16 * np)); KST_CUDA(cudaMalloc(&p.x2_w, 16 * nw));
    KST_CUDA(cudaMalloc(&p.snap, 8 * (3 * (np + nw) + (size_t)K * (h->d + 1) + 2 * (size_t)S + 1)));
    KST_CUDA(cudaMalloc(&p.bar, 2 * sizeof(unsigned long long)));  // arrival counter, error word
    KST_CUDA(cudaMemsetAsync(p.bar, 0, 2 * sizeof(unsigned long long), h->stream));
    p.bar_count = 0;
    p.S = S; p.K = K;
    return KST_OK;
}

// trace buffers for n entries, sized for the largest design the handle can hold (swarms of any shape reuse them)
int kst_pso_ensure_trace(kst_problem *h, int n)
{
    KstPso &p = h->pso;
    if (n <= p.tr_cap) return KST_OK;
    KST_CUDA(cudaStreamSynchronize(h->stream));
    if (p.tr_fx) { KST_CUDA(cudaFree(p.tr_fx)); KST_CUDA(cudaFree(p.tr_pts)); KST_CUDA(cudaFree(p.tr_w)); }
    p.tr_fx = p.tr_pts = p.tr_w = nullptr;
    p.tr_cap = 0;
    KST_CUDA(cudaMalloc(&p.tr_fx, sizeof(double) * n));
    KST_CUDA(cudaMalloc(&p.tr_pts, sizeof(double) * n * KST_MAX_ROWS * h->d));
    KST_CUDA(cudaMalloc(&p.tr_w, sizeof(double) * n * KST_MAX_ROWS));
    p.tr_cap = n;
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
                                           p.pending_n_blk ? h->partials : nullptr, p.pending_n_blk, p.pending_offset,
                                           p.pending_nan);
    p.pending_n_blk = 0;
    KST_CUDA(cudaGetLastError());
    return KST_OK;
}
