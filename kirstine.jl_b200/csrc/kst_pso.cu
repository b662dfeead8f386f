// Device-resident particle swarm: the swarm state never leaves the GPU between iterations.
//
//   optimizer_state      src/pso.jl:83-125       -> k_pso_init + evaluation + k_pso_best(init)
//   pso_update_velocity! src/pso.jl:143-164  \
//   pso_update_position! src/pso.jl:166-171   >  -> k_pso_move (a thread per component, then one per particle)
//   ap_add!(p, v, c)     src/designmeasure-abstractpoint.jl:164-285 /
//   pso_update_best!     src/pso.jl:181-193      -> k_pso_best (CTAs of 128 particles, order-independent arg-max merge)
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

// velocity + constrained move.  A CTA takes `ppb` particles: one thread per (particle, component) draws the two random
// numbers and updates the velocity (kst_pso_velocity; components are independent) and leaves position and new velocity
// in a shared-memory slice of the particle; then one thread per particle does the constrained move on its slice, whose
// sums run in the reference's order (kst_pso_position); then all threads write x and v back -- the same split as in
// the persistent swarm kernel.
// Why the slice: the move rewrites every component several times in loops of the form x[k] = f(x[k], v[k]).  A warp
// issues in order, so in global memory each iteration is a dependent round trip (the store drops the line from L1
// and the next iteration's load has to wait for it to issue): one thread per particle working in place took 17.5 us
// for 1024 particles x 16 components with cold caches, velocities spread over the components 11.3 us.  (A first
// staged version copied x, v, p into shared memory with plain per-element loops -- load -> store chains of the same
// kind -- and took 24.5 us.)
#define KST_MOVE_NT 128
__global__ void __launch_bounds__(KST_MOVE_NT) k_pso_move(KstPsoDev s, uint32_t iter, double c1, double c2, double chi, int ppb)
{
    extern __shared__ double s_mv[];   // [ppb][2 KC + 1]: x_w | x_pts | v_w | v_pts of a particle, odd stride
    const int K = s.K, d = s.d, KD = K * d, KC = K + KD, stride = 2 * KC + 1;
    const int i0 = blockIdx.x * ppb, np = min(ppb, s.S - i0);
    for (int e = threadIdx.x; e < np * KC; e += KST_MOVE_NT) {
        const int il = e / KC, c = e % KC, i = i0 + il;
        double r1, r2;
        kst_uniform2(s.seed, iter, (uint32_t)i, (uint32_t)c, 0, r1, r2);
        const bool isw = c < K;
        const size_t o = isw ? (size_t)i * K + c : (size_t)i * KD + (c - K);
        const double x = isw ? s.x_w[o] : s.x_pts[o], v = isw ? s.v_w[o] : s.v_pts[o];
        const double pb = isw ? s.p_w[o] : s.p_pts[o], gb = isw ? s.g_w[c] : s.g_pts[c - K];
        double *sl = s_mv + (size_t)il * stride;
        sl[c] = x;
        sl[KC + c] = kst_pso_velocity(v, x, pb, gb, r1, r2, c1, c2, chi);
    }
    __syncthreads();
    if ((int)threadIdx.x < np) {
        double *sl = s_mv + (size_t)threadIdx.x * stride;
        kst_pso_position(K, d, sl, sl + K, sl + KC, sl + KC + K, s.fixw, s.fixp, s.lb, s.ub);
    }
    __syncthreads();
    for (int e = threadIdx.x; e < np * KC; e += KST_MOVE_NT) {
        const int il = e / KC, c = e % KC, i = i0 + il;
        const bool isw = c < K;
        const size_t o = isw ? (size_t)i * K + c : (size_t)i * KD + (c - K);
        const double *sl = s_mv + (size_t)il * stride;
        if (isw) { s.x_w[o] = sl[c]; s.v_w[o] = sl[KC + c]; }
        else { s.x_pts[o] = sl[c]; s.v_pts[o] = sl[KC + c]; }
    }
}

// pso_update_best! as a kernel of its own, spread over CTAs of KST_BEST_NT particles: every CTA finishes the objective
// values of its particles and updates their personal bests; the swarm-wide part is done by the CTA that arrives last
// (kst_pso_best_grid, kst_pso_dev.cuh).
// One CTA of 1024 threads did all of it before: 20 us for 1024 particles, because a single SM had to move the partial
// sums, x and p (about 0.6 MB) on its own.
#define KST_BEST_NT 128
__global__ void __launch_bounds__(KST_BEST_NT) k_pso_best(KstPsoDev s, int init, double *tr_fx, double *tr_pts, double *tr_w,
                                                          int slot, int *status, const double *__restrict__ partials,
                                                          int n_blk, double offset, int nan_fill, KstBestScratch sc)
{
    const int tid = threadIdx.x;
    double fg_old = -KST_DINF;
    if (tid == 0 && !init) fg_old = *s.fg;   // requested first, needed last (by the last CTA only)
    const int i = blockIdx.x * KST_BEST_NT + tid;
    double best = -KST_DINF;
    int bi = -1;
    bool notnan = false;
    if (i < s.S) {
        bool improved;
        const double f = kst_pso_best_particle(s, i, init, partials, n_blk, offset, nan_fill, improved);
        notnan = f == f;
        if (improved) { best = f; bi = i; }
    }
    kst_pso_best_grid(s, best, bi, notnan, fg_old, sc, tr_fx, tr_pts, tr_w, slot, status);
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
    if (p.arena) cudaFree(p.arena);
    if (p.out) cudaFree(p.out);
    if (p.h_out) cudaFreeHost(p.h_out);
    p = KstPso();
}

// Carves the swarm for S particles of K points out of ONE device allocation that only ever grows: Exchange alternates
// between two swarm shapes every step, and a cudaMalloc / cudaFree pair costs more than a short optimizer run.
int kst_pso_alloc(kst_problem *h, int S, int K)
{
    KstPso &p = h->pso;
    const size_t np = (size_t)S * K * h->d, nw = (size_t)S * K, kd = (size_t)K * h->d;
    // doubles: barrier words (2), x p v (3 np + 3 nw), published positions (2 np + 2 nw), g, fx, fp, fg, fixw / fixp,
    // best-update scratch (per CTA one candidate value, two ints; one counter word)
    const size_t n_cta = ((size_t)S + 127) / 128;
    const size_t need = 2 + 5 * (np + nw) + kd + K + 2 * (size_t)S + 1 + (size_t)K + 2 + 2 * n_cta + 1;
    if (need > p.arena_cap) {
        KST_CUDA(cudaStreamSynchronize(h->stream));
        if (p.arena) KST_CUDA(cudaFree(p.arena));
        p.arena = nullptr;
        p.arena_cap = 0;
        const size_t cap = need + need / 2;
        KST_CUDA(cudaMalloc(&p.arena, sizeof(double) * cap));
        p.arena_cap = cap;
        KST_CUDA(cudaMemsetAsync(p.arena, 0, 2 * sizeof(double), h->stream));  // arrival counter, error word
        p.bar_count = 0;
    }
    double *q = p.arena;
    p.bar = (unsigned long long *)q; q += 2;
    p.x_pts = q; q += np; p.p_pts = q; q += np; p.v_pts = q; q += np; p.x2_pts = q; q += 2 * np;
    p.x_w = q; q += nw; p.p_w = q; q += nw; p.v_w = q; q += nw; p.x2_w = q; q += 2 * nw;
    p.g_pts = q; q += kd; p.g_w = q; q += K;
    p.fx = q; q += S; p.fp = q; q += S; p.fg = q; q += 1;
    p.fixw = (int *)q; p.fixp = p.fixw + K; q += K + 2;
    p.best_val = q; q += n_cta; p.best_idx = (int *)q; q += n_cta; p.best_counter = (unsigned int *)q;
    KST_CUDA(cudaMemsetAsync(p.best_counter, 0, sizeof(double), h->stream));  // zero between launches (the last CTA resets it)
    p.fused_off = 0;
    p.S = S; p.K = K;
    p.tr_n = 0;  // the output block has to be carved again for this shape
    return KST_OK;
}

// doubles of the packed PsoState snapshot (layout: KstFusedParams::snap)
static size_t snapshot_doubles(int S, int K, int d)
{
    return 3 * (size_t)S * K * (d + 1) + (size_t)K * (d + 1) + 2 * (size_t)S + 1;
}

// Carves what a run of n trace entries sends back to the host out of one device block, mirrored by a pinned host
// block, so that everything returns in ONE copy: [status word | trace_fx n | trace_points n K d | trace_weights n K |
// snapshot of the initial state].
int kst_pso_ensure_trace(kst_problem *h, int n)
{
    KstPso &p = h->pso;
    const size_t kd = (size_t)p.K * h->d;
    const size_t need = 1 + (size_t)n * (1 + kd + p.K) + snapshot_doubles(p.S, p.K, h->d);
    if (need > p.out_cap) {
        KST_CUDA(cudaStreamSynchronize(h->stream));
        if (p.out) KST_CUDA(cudaFree(p.out));
        if (p.h_out) KST_CUDA(cudaFreeHost(p.h_out));
        p.out = p.h_out = nullptr;
        p.out_cap = 0;
        const size_t cap = need + need / 2;
        KST_CUDA(cudaMalloc(&p.out, sizeof(double) * cap));
        KST_CUDA(cudaHostAlloc(&p.h_out, sizeof(double) * cap, cudaHostAllocDefault));
        p.out_cap = cap;
    }
    p.d_status = (int *)p.out;
    p.tr_fx = p.out + 1;
    p.tr_pts = p.tr_fx + n;
    p.tr_w = p.tr_pts + (size_t)n * kd;
    p.snap = p.tr_w + (size_t)n * p.K;
    p.tr_n = n;
    return KST_OK;
}

// Brings the output block of the last run back in one copy and waits for it.  The pointers (into the pinned host
// mirror, valid until the next run) follow the device layout; snap only when with_snap.
int kst_pso_fetch(kst_problem *h, int with_snap, int *status, const double **tr_fx, const double **tr_pts,
                  const double **tr_w, const double **snap)
{
    KstPso &p = h->pso;
    const size_t kd = (size_t)p.K * h->d;
    const size_t n = 1 + (size_t)p.tr_n * (1 + kd + p.K) + (with_snap ? snapshot_doubles(p.S, p.K, h->d) : 0);
    KST_CUDA(cudaMemcpyAsync(p.h_out, p.out, sizeof(double) * n, cudaMemcpyDeviceToHost, h->stream));
    KST_CUDA(cudaStreamSynchronize(h->stream));
    if (status) *status = *(const int *)p.h_out;
    if (tr_fx) *tr_fx = p.h_out + 1;
    if (tr_pts) *tr_pts = p.h_out + 1 + p.tr_n;
    if (tr_w) *tr_w = p.h_out + 1 + p.tr_n + (size_t)p.tr_n * kd;
    if (snap) *snap = p.h_out + 1 + (size_t)p.tr_n * (1 + kd + p.K);
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
    const int kc = s.K * (s.d + 1);
    const int ppb = kc >= KST_MOVE_NT ? 1 : KST_MOVE_NT / kc;  // particles per CTA: one thread per component where they fit
    int rc = kst_prof_begin(h, KST_PROF_MOVE);
    if (rc) return rc;
    const size_t smem = sizeof(double) * ppb * (2 * (size_t)kc + 1);
    if (smem > 48 * 1024) { kst_set_error("pso: %d components per particle exceed the move kernel's shared memory", kc); return KST_ELIMIT; }
    k_pso_move<<<(s.S + ppb - 1) / ppb, KST_MOVE_NT, smem, h->stream>>>(s, (uint32_t)h->pso.iter, h->pso.c1, h->pso.c2, chi, ppb);
    KST_CUDA(cudaGetLastError());
    return kst_prof_end(h, KST_PROF_MOVE);
}

int kst_launch_pso_best(kst_problem *h, int init, int trace_slot)
{
    const KstPsoDev s = make_dev(h);
    KstPso &p = h->pso;
    int rc = kst_prof_begin(h, KST_PROF_BEST);
    if (rc) return rc;
    KstBestScratch sc;
    sc.val = p.best_val; sc.idx = p.best_idx; sc.counter = p.best_counter;
    k_pso_best<<<(s.S + KST_BEST_NT - 1) / KST_BEST_NT, KST_BEST_NT, 0, h->stream>>>(s, init, trace_slot >= 0 ? p.tr_fx : nullptr,
                                           trace_slot >= 0 ? p.tr_pts : nullptr, trace_slot >= 0 ? p.tr_w : nullptr,
                                           trace_slot >= 0 ? trace_slot : 0, init ? p.d_status : nullptr,
                                           p.pending_n_blk ? h->partials : nullptr, p.pending_n_blk, p.pending_offset,
                                           p.pending_nan, sc);
    p.pending_n_blk = 0;
    KST_CUDA(cudaGetLastError());
    return kst_prof_end(h, KST_PROF_BEST);
}
