// Objective evaluation: for every particle (design measure) of a swarm and every prior draw,
// accumulate the weighted Fisher information, apply the transformation, evaluate the D / A
// criterion, and prior-average.  Replaces the three nested serial loops of the reference:
//   pso_evaluate_objective!  src/pso.jl:173-179            (particles)
//   objective!               src/designproblem.jl:192-224  (prior draws)
//   average_fishermatrix!    src/model-nonlinear-regression.jl:60-86 (design points)
//
// Mapping: CTA = (tile of KST_TILE_P particles) x (strided set of draw chunks).  Lanes run
// along draws (coalesced SoA loads of the per-draw constants); all lanes of a warp share the
// particle, so its compacted row list is read from shared memory as a broadcast.  The r x r
// matrix, its LDL' factor and everything after live in FP64 registers.  Per particle one
// fixed-order block reduction; per-CTA partial sums go to scratch [n_blk][S] and are summed in
// fixed order by k_obj_finish, so results are bit-reproducible run to run.
#include <stdlib.h>
#include <string.h>

#include <mutex>

#include "kst_pso_fused.cuh"

// the evaluation kernels of every family are compiled in their own translation units
#define KST_X(MF)                                                                          \
    extern template int kst_obj_launch<MF>(kst_problem *, const KstObjParams &, dim3, int); \
    extern template int kst_fused_launch<MF>(kst_problem *, KstFusedParams &, dim3);
KST_OBJ_FAMILIES(KST_X)
#undef KST_X

// ------------------------------------------------------------------------------------------
// per-draw preparation (once per problem)
// ------------------------------------------------------------------------------------------
template <class MF>
__global__ void k_prep_draws(const double *__restrict__ theta_aos, const double *__restrict__ pw,
                             const double *__restrict__ dt, int64_t N, size_t ld, int crit, int trafo,
                             int t, KstModelParams mp, double *__restrict__ theta, double *__restrict__ dc,
                             double *__restrict__ cs, double *__restrict__ aux, double *__restrict__ woff)
{
    constexpr int R = MF::R;
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= N) return;
    double th[R], dcv[MF::NDC], csv[R];
#pragma unroll
    for (int i = 0; i < R; i++) {
        th[i] = theta_aos[j * R + i];
        theta[(size_t)i * ld + j] = th[i];
    }
    MF::prep_draw(th, mp, dcv, csv);
#pragma unroll
    for (int i = 0; i < MF::NDC; i++) dc[(size_t)i * ld + j] = dcv[i];
#pragma unroll
    for (int i = 0; i < R; i++) cs[(size_t)i * ld + j] = csv[i];
    double off = 0.0;
    if (trafo == KST_TRAFO_IDENTITY) {
        if (crit == KST_CRIT_D) {
#pragma unroll
            for (int i = 0; i < R; i++) off += log(csv[i] * csv[i]);
            off *= pw[j];
        } else if (crit == KST_CRIT_DEXP) {
            double lg = 0.0;
#pragma unroll
            for (int i = 0; i < R; i++) lg += log(csv[i] * csv[i]);
            aux[j] = exp(lg / R);
        } else {
#pragma unroll
            for (int i = 0; i < R; i++) aux[(size_t)i * ld + j] = 1.0 / (csv[i] * csv[i]);
        }
    } else {
        for (int a = 0; a < t; a++)
#pragma unroll
            for (int i = 0; i < R; i++)
                aux[(size_t)(a * R + i) * ld + j] = dt[j * (int64_t)(t * R) + a + i * t] / csv[i];
    }
    woff[j] = off;
}


// nan_fill: Gateaux derivative at a design whose constants could not be computed (gateauxderivative fills NaN,
// src/designproblem.jl:281-284)
__global__ void k_obj_finish(const double *__restrict__ partials, int n_blk, double offset,
                             double *__restrict__ fx, int S, int nan_fill)
{
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= S) return;
    double s = 0.0;
#pragma unroll 8
    for (int b = 0; b < n_blk; b++) s += partials[(size_t)b * S + p];
    fx[p] = nan_fill ? __longlong_as_double(0x7ff8000000000000LL) : s + offset;
}

// informationmatrix for one draw (debugging aid of the reference, src/design-common.jl:96-112)
template <class MF>
__global__ void k_infomatrix(const double *dc, const double *cs, const double *reg, size_t ld, int64_t j,
                             const double *pts, const double *w, int K, KstModelParams mp, double *Mout)
{
    constexpr int R = MF::R;
    extern __shared__ __align__(16) double smem[];
    __shared__ __align__(16) double s_mathtab[KST_MATH_SMEM];
    const KstTab tab = kst_math_init(s_mathtab);
    kst_build_tile<MF>(smem, pts, w, 0, 1, K, mp);
    if (threadIdx.x != 0) return;
    const int rows_max = K * mp.m;
    double dcv[MF::NDC], M[KstPacked<R>::N];
    for (int c = 0; c < MF::NDC; c++) dcv[c] = dc[(size_t)c * ld + j];
    kst_fisher<MF>(M, (unsigned)__cvta_generic_to_shared(KstTile<MF>::rows(smem)), KstTile<MF>::nact(smem, rows_max)[0],
                   KstTile<MF>::sumw(smem, rows_max)[0], dcv, tab,
                   (unsigned)__cvta_generic_to_shared(KstTile<MF>::lmat(smem, rows_max)), mp.m);
    for (int a = 0; a < R; a++)
        for (int b = a; b < R; b++) {
            const double mt = M[kst_pidx<R>(a, b)] + (reg ? reg[(size_t)kst_pidx<R>(a, b) * ld + j] : 0.0);
            const double v = mt * cs[(size_t)a * ld + j] * cs[(size_t)b * ld + j];
            Mout[a + b * R] = v;
            Mout[b + a * R] = v;
        }
}

// max_j |x[c][j]| for the rows c of an [rows][ld] array, as the bit pattern of the absolute value (monotone for
// non-negative doubles; a NaN sorts above Inf and so poisons the result, which is what the caller wants)
__global__ void k_absmax(const double *__restrict__ x, int rows, size_t ld, int64_t n, unsigned long long *out)
{
    for (int c = 0; c < rows; c++) {
        unsigned long long m = 0;
        for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += (int64_t)gridDim.x * blockDim.x) {
            const unsigned long long b = (unsigned long long)__double_as_longlong(x[(size_t)c * ld + j]) & 0x7fffffffffffffffull;
            m = b > m ? b : m;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const unsigned long long t = __shfl_down_sync(0xffffffffu, m, o);
            m = t > m ? t : m;
        }
        if ((threadIdx.x & 31) == 0 && m) atomicMax(out + c, m);
    }
}

// Deterministic sum of n doubles: 1024 threads take strided slices, then a fixed tree.
__global__ void k_sum_fixed(const double *__restrict__ x, int64_t n, double *__restrict__ out)
{
    __shared__ double s[1024];
    double a = 0.0;
    for (int64_t i = threadIdx.x; i < n; i += 1024) a += x[i];
    s[threadIdx.x] = a;
    __syncthreads();
    for (int o = 512; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) s[threadIdx.x] += s[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) *out = s[0];
}

// the sum goes to the handle's scratch double (offset 8 of d_flag); the caller copies it out
int kst_sum_fixed_enqueue(kst_problem *h, const double *d_x, int64_t n)
{
    k_sum_fixed<<<1, 1024, 0, h->stream>>>(d_x, n, (double *)((char *)h->d_flag + 8));
    KST_CUDA(cudaGetLastError());
    return KST_OK;
}

int kst_sum_fixed(kst_problem *h, const double *d_x, int64_t n, double *host_out)
{
    double *d_out = (double *)((char *)h->d_flag + 8);  // the handle's scratch double: no allocation per call
    k_sum_fixed<<<1, 1024, 0, h->stream>>>(d_x, n, d_out);
    KST_CUDA(cudaGetLastError());
    KST_CUDA(cudaMemcpyAsync(host_out, d_out, sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    KST_CUDA(cudaStreamSynchronize(h->stream));
    return KST_OK;
}

// ------------------------------------------------------------------------------------------
// host-side dispatch
// ------------------------------------------------------------------------------------------

// d_scratch: N + 8 doubles owned by the caller (kst_problem_create frees it with its other upload temporaries)
int kst_launch_prep_draws(kst_problem *h, const double *d_theta_aos, const double *d_dt, double *d_scratch)
{
    double *woff = d_scratch;
    const int nt = 256;
    const unsigned nb = (unsigned)((h->N + nt - 1) / nt);
    KST_MODEL_SWITCH(h->model_id,
                     (k_prep_draws<MF><<<nb, nt, 0, h->stream>>>(d_theta_aos, h->pw, d_dt, h->N, h->ld, h->crit,
                                                                 h->trafo, h->t, h->mp, h->theta, h->dc, h->cs,
                                                                 h->aux, woff)));
    KST_CUDA(cudaGetLastError());
    int rc = kst_sum_fixed(h, woff, h->N, &h->offset);
    if (rc) return rc;
    // per-problem maxima of the hoisted constants (KstObjParams::dcmax)
    int ndc = 0;
    KST_MODEL_SWITCH(h->model_id, ndc = MF::NDC);
    unsigned long long *d_max = (unsigned long long *)(d_scratch + h->N), bits[8] = {};
    cudaError_t e = cudaMemsetAsync(d_max, 0, sizeof(bits), h->stream);
    if (e == cudaSuccess) {
        k_absmax<<<h->sm_count * 4, 256, 0, h->stream>>>(h->dc, ndc, h->ld, h->N, d_max);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(bits, d_max, sizeof(bits), cudaMemcpyDeviceToHost, h->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
    if (e != cudaSuccess) { kst_set_error("prep_draws: %s", cudaGetErrorString(e)); return KST_ECUDA; }
    for (int c = 0; c < 8; c++) memcpy(&h->dcmax[c], &bits[c], sizeof(double));
    return KST_OK;
}

// Grid of the evaluation kernel: tiles of *tile_p particles (gridDim.y) x n_blk draw splits (gridDim.x).
// Large problems use full tiles of KST_TILE_P particles (each draw's constants are loaded once per tile) and at
// least 8 waves.  Small problems -- the reference's own test and vignette sizes, a few dozen particles against a
// thousand draws -- cannot fill one wave that way and would run KST_TILE_P particles back to back on a handful of
// SMs; they get the largest tile size that still yields a full wave of CTAs (down to one particle per CTA).
int kst_objective_grid(const kst_problem *h, int S, int *n_blk, int *tile_p)
{
    const int64_t max_chunks = (h->N + KST_NT - 1) / KST_NT;
    const int cap = h->sm_count * KST_MINB;  // resident CTAs per wave
    int tp = KST_TILE_P;
    while (tp > 1 && (int64_t)((S + tp - 1) / tp) * max_chunks < cap) tp >>= 1;
    *tile_p = tp;
    const int tiles = (S + tp - 1) / tp;
    // Many short CTAs beat few long ones (measured on cfg 3 and cfg 5: the ragged end of the last wave costs more
    // than the per-CTA prologue): aim for >= 8 waves and pick the draw split whose last wave is fullest ...
    int64_t lo = (8LL * cap + tiles - 1) / tiles;
    if (lo < 1) lo = 1;
    int64_t best = lo;
    double best_waste = 1e30;
    for (int64_t nb = lo; nb < lo + cap; nb++) {
        const int64_t tot = nb * tiles;
        const int64_t waves = (tot + cap - 1) / cap;
        const double waste = (double)(waves * cap - tot) / (double)tot;
        if (waste < best_waste - 1e-12) { best_waste = waste; best = nb; }
    }
    // ... and when that split leaves each CTA only a few draws per thread, a coarser one with an (almost) full last
    // wave can be cheaper.  Cost in units of one draw per thread: waves * (F + dpt) + B * dpt, with F = 1 the fixed
    // part of a CTA (tables, particle tile, reduction; an SM runs below its FP64 rate while one of its two CTAs is in
    // it) and B = 0.3 the ragged end of the last wave.  Measured on cfg 5 (1024 particles): 125 000 draws per GPU run
    // 1.1 % faster on 23 splits (10 waves) than on 37 (16 waves), 250 000 draws and more are fastest on 37.
    {
        auto cost = [&](int64_t nb) {
            const double dpt = (double)h->N / ((double)nb * KST_NT);
            const int64_t waves = (nb * tiles + cap - 1) / cap;
            return (double)waves * (1.0 + dpt) + 0.3 * dpt;
        };
        double best_cost = cost(best);
        for (int64_t nb = best - 1; nb >= 1 && nb * tiles >= 6LL * cap - cap / 8; nb--) {
            const int64_t tot = nb * tiles, waves = (tot + cap - 1) / cap;
            if ((double)(waves * cap - tot) > 0.015 * (double)tot) continue;
            if (cost(nb) < best_cost * 0.998) { best_cost = cost(nb); best = nb; }
        }
    }
    // ... but not below ~4 draws per thread, where the prologue (math tables into shared memory, particle tile)
    // starts to dominate (measured on cfg 2), unless that would leave SMs without a CTA
    int64_t nb_dpt = h->N / (KST_NT * 4);
    const int64_t fill = cap / tiles > 0 ? cap / tiles : 1;  // the split that just fits one wave
    if (nb_dpt < fill) nb_dpt = fill;
    if (best > nb_dpt) best = nb_dpt;
    if (best > max_chunks) best = max_chunks;
    if (best < 1) best = 1;
    if (best > 65535) best = 65535;
    *n_blk = (int)best;
    return tiles;
}

// Argument and fixed-limit checks shared by every entry point that launches k_objective
static int kst_objective_check(const kst_problem *h, int S, int K, const char *who)
{
    if (S < 1 || K < 1 || K * h->m > KST_MAX_ROWS) {
        kst_set_error("%s: need S >= 1, K >= 1, K*m <= %d (S=%d K=%d m=%d)", who, KST_MAX_ROWS, S, K, h->m);
        return KST_EINVAL;
    }
    // 32-bit element indices into the per-draw SoA arrays inside the kernel: the widest array has
    // max(NDC, r(r+1)/2 with a regularisation matrix) rows of ld entries
    const int64_t rows = h->reg ? (int64_t)h->r * (h->r + 1) / 2 : 8;
    if (h->N > (1LL << 29) || (int64_t)h->ld * rows > (1LL << 32)) {
        kst_set_error("%s: at most 2^29 draws per GPU (2^32 / %lld with a regularisation matrix); shard the prior "
                      "sample over more GPUs", who, (long long)rows);
        return KST_ELIMIT;
    }
    if (S > 65535 * KST_TILE_P) {  // gridDim.y holds the particle tiles
        kst_set_error("%s: at most %d particles per call (split the swarm)", who, 65535 * KST_TILE_P);
        return KST_ELIMIT;
    }
    return KST_OK;
}

// Kernel arguments of one swarm evaluation.  gd: the Gateaux integrand at the prepared design instead of the criterion.
void kst_objective_params(const kst_problem *h, int S, int K, const double *d_pts, const double *d_w, double *partials,
                          int tile_p, int gd, KstObjParams *out)
{
    KstObjParams &prm = *out;
    prm.pw = h->pw; prm.dc = h->dc; prm.aux = gd ? h->gAp : h->aux; prm.reg = gd ? nullptr : h->reg; prm.ld = h->ld; prm.N = h->N;
    prm.pts = d_pts; prm.w = d_w; prm.S = S; prm.K = K; prm.partials = partials; prm.mp = h->mp;
    prm.w_uniform = h->w_uniform; prm.tile_p = tile_p; prm.gd = gd;
    for (int c = 0; c < 8; c++) prm.dcmax[c] = h->dcmax[c];
}

// what the finishing step adds to the fixed-order sum of the per-CTA partial sums
void kst_objective_finish_terms(const kst_problem *h, int gd, double *offset, int *nan_fill)
{
    *offset = gd ? -h->g_offset : h->offset;
    *nan_fill = gd ? h->g_singular : 0;
}

int kst_launch_objective(kst_problem *h, int S, int K, const double *d_pts, const double *d_w, double *d_fx,
                         int *defer_n_blk, int gd)
{
    {
        const int rc0 = kst_objective_check(h, S, K, gd ? "gateaux (swarm)" : "objective");
        if (rc0) return rc0;
    }
    if (gd && !h->g_ready) { kst_set_error("gateaux evaluation before gateaux_prepare"); return KST_ESTATE; }
    if (gd && K != 1) { kst_set_error("derivative only implemented for one-point directions"); return KST_EINVAL; }
    int n_blk = 1, tile_p = KST_TILE_P;
    const int tiles = kst_objective_grid(h, S, &n_blk, &tile_p);
    int rc = kst_ensure_partials(h, (size_t)S * n_blk);
    if (rc) return rc;
    KstObjParams prm;
    kst_objective_params(h, S, K, d_pts, d_w, h->partials, tile_p, gd, &prm);
    const dim3 grid(n_blk, tiles);
    rc = kst_prof_begin(h, KST_PROF_EVAL);
    if (rc) return rc;
    KST_MODEL_SWITCH(h->model_id, rc = kst_obj_launch<MF>(h, prm, grid, K));
    if (rc) return rc;
    KST_CUDA(cudaGetLastError());
    rc = kst_prof_end(h, KST_PROF_EVAL);
    if (rc) return rc;
    if (defer_n_blk) {
        *defer_n_blk = n_blk;
        return KST_OK;
    }
    double offset;
    int nan_fill;
    kst_objective_finish_terms(h, gd, &offset, &nan_fill);
    k_obj_finish<<<(S + 127) / 128, 128, 0, h->stream>>>(h->partials, n_blk, offset, d_fx, S, nan_fill);
    KST_CUDA(cudaGetLastError());
    return KST_OK;
}

// The persistent swarm kernel for the handle's swarm (kst_pso_fused.cuh).  Applicable on one GPU, without a
// regularisation matrix, when the evaluation grid kst_objective_grid picks fits one wave of resident CTAs.
int kst_launch_pso_fused(kst_problem *h, int n_steps, int first_is_init, int tr_slot0, int with_trace, int *d_status,
                         double *d_snap)
{
    KstPso &p = h->pso;
    const int gd = p.mode == KST_PSO_MODE_GATEAUX;
    const bool disabled = getenv("KST_NO_FUSED") != nullptr;  // A/B switch for tests and measurements (read per call)
    if (disabled || p.fused_off || h->nranks > 1 || (h->reg && !gd) || p.S > KST_FUSED_MAX_S || n_steps < 1)
        return KST_FUSED_NOFIT;
    int rc = kst_objective_check(h, p.S, p.K, "pso (persistent kernel)");
    if (rc) return rc;
    int n_blk = 1, tile_p = KST_TILE_P;
    const int tiles = kst_objective_grid(h, p.S, &n_blk, &tile_p);
    if ((long long)tiles * n_blk > (long long)h->sm_count * KST_MINB) { p.fused_off = 1; return KST_FUSED_NOFIT; }
    rc = kst_ensure_partials(h, (size_t)2 * p.S * n_blk);
    if (rc) return rc;
    KstFusedParams fp;
    kst_objective_params(h, p.S, p.K, nullptr, nullptr, nullptr, tile_p, gd, &fp.obj);
    kst_pso_make_dev(h, &fp.pso);
    fp.x2_pts = p.x2_pts; fp.x2_w = p.x2_w; fp.partials2 = h->partials;
    fp.bar = p.bar; fp.bar_base = p.bar_count;
    fp.n_steps = n_steps; fp.first_is_init = first_is_init; fp.iter0 = (uint32_t)p.iter;
    fp.c1 = p.c1; fp.c2 = p.c2;
    const double phi = p.c1 + p.c2;
    fp.chi = 2 / fabs(2 - phi - sqrt(phi * phi - 4 * phi));  // src/pso.jl:134-135
    kst_objective_finish_terms(h, gd, &fp.offset, &fp.nan_fill);
    fp.tr_fx = with_trace ? p.tr_fx : nullptr; fp.tr_pts = with_trace ? p.tr_pts : nullptr; fp.tr_w = with_trace ? p.tr_w : nullptr;
    fp.tr_slot0 = tr_slot0; fp.status = d_status; fp.snap = d_snap;
    const dim3 grid(n_blk, tiles);
    KST_MODEL_SWITCH(h->model_id, rc = kst_fused_launch<MF>(h, fp, grid));
    if (rc == KST_FUSED_NOFIT) p.fused_off = 1;
    if (rc) return rc;
    p.bar_count += (unsigned long long)n_steps * (unsigned long long)(n_blk * tiles);
    return KST_OK;
}

int kst_launch_infomatrix(kst_problem *h, int K, const double *d_pts, const double *d_w, int64_t draw, double *d_M)
{
    if (K < 1 || K * h->m > KST_MAX_ROWS || draw < 0 || draw >= h->N) {
        kst_set_error("informationmatrix: bad K or draw index");
        return KST_EINVAL;
    }
    KST_MODEL_SWITCH(h->model_id, {
        const size_t smem = KstTile<MF>::bytes(K * h->m);
        KST_OPT_IN(k_infomatrix<MF>);
        k_infomatrix<MF><<<1, 32, smem, h->stream>>>(h->dc, h->cs, h->reg, h->ld, draw, d_pts, d_w, K, h->mp, d_M);
    });
    KST_CUDA(cudaGetLastError());
    return KST_OK;
}

// ------------------------------------------------------------------------------------------
// Evaluation without any allocation (no counterpart in the reference)
// ------------------------------------------------------------------------------------------
// Who needs this: hosts that must not touch the CUDA allocator while they run.  Typical cases are a
// process whose device memory was handed in full to another pool up front (a framework's caching
// allocator, a garbage-collected runtime's CUDA pool) so that a later cudaMalloc can fail, and
// latency-critical loops that cannot afford the device-wide synchronisation cudaMalloc / cudaFree
// imply.  The regular entry points grow the per-handle scratch for the per-CTA partial sums with
// cudaMalloc when a larger swarm arrives; this one keeps that scratch in a fixed array inside the
// library image instead.
//
// Cost and limits: every GPU that loads the library holds KST_NOALLOC_MAX_PARTIALS doubles (2 MiB)
// until the process exits, used or not.  At most KST_NOALLOC_MAX_SWARM particles per call.  All
// calls on one GPU share the array, so two calls that overlap on that GPU (different handles or
// streams) race on it: serialise them.  Nothing in the array has to start at zero: within a call
// every entry the finishing kernel reads has been written by the evaluation kernel of that same
// call, so contents left over from the previous call are never observed.
#define KST_NOALLOC_MAX_PARTIALS 262144
#define KST_NOALLOC_MAX_SWARM 4096
#define KST_NOALLOC_MAX_DEVICES 64

static __device__ __align__(256) double kst_noalloc_scratch[KST_NOALLOC_MAX_PARTIALS];

// Inside host code `kst_noalloc_scratch` is only a host-side handle for the symbol, not a device
// address; the address valid on a given GPU comes from cudaGetSymbolAddress with that GPU current.
static double *g_noalloc_ptr[KST_NOALLOC_MAX_DEVICES];
static std::mutex g_noalloc_mutex;

static int kst_noalloc_scratch_ptr(int device, double **out)
{
    if (device < 0 || device >= KST_NOALLOC_MAX_DEVICES) {
        kst_set_error("objective_batch_noalloc: device %d beyond the %d devices this build tracks", device,
                      KST_NOALLOC_MAX_DEVICES);
        return KST_ELIMIT;
    }
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_ptr[device]) {
        void *ptr = nullptr;
        KST_CUDA(cudaSetDevice(device));
        KST_CUDA(cudaGetSymbolAddress(&ptr, kst_noalloc_scratch));
        g_noalloc_ptr[device] = static_cast<double *>(ptr);
    }
    *out = g_noalloc_ptr[device];
    return KST_OK;
}

// Same evaluation as kst_objective_batch_device (the same k_objective / k_obj_finish kernels), with the
// partial sums kept in the static array above.  Calls no cudaMalloc, cudaFree or any other allocating
// routine.  The first call on each GPU looks the array's device address up (cudaGetSymbolAddress); make
// that first call before starting any stream capture.  The CUDA runtime also loads a kernel's code the first
// time that kernel is launched, and k_objective is one kernel per (model family, criterion, transformation):
// so warm up once per such combination you are going to use, not just once per GPU.  Asynchronous on the
// handle's stream.  Only the
// draws held by this handle are summed: no collective is issued even when a communicator is attached.
extern "C" int kst_objective_batch_noalloc(kst_problem *h, int S, int K, const double *d_points,
                                           const double *d_weights, double *d_fx)
{
    if (!h) { kst_set_error("NULL problem handle"); return KST_EINVAL; }
    if (!d_points || !d_weights || !d_fx) { kst_set_error("objective_batch_noalloc: NULL buffer"); return KST_EINVAL; }
    int rc = kst_objective_check(h, S, K, "objective_batch_noalloc");
    if (rc) return rc;
    if (S > KST_NOALLOC_MAX_SWARM) {
        kst_set_error("objective_batch_noalloc: swarm of %d particles exceeds the fixed limit of %d; "
                      "use kst_objective_batch_device, which allocates its scratch", S, KST_NOALLOC_MAX_SWARM);
        return KST_ELIMIT;
    }
    KST_CUDA(cudaSetDevice(h->device));
    double *scratch = nullptr;
    rc = kst_noalloc_scratch_ptr(h->device, &scratch);
    if (rc) return rc;
    int n_blk = 1, tile_p = KST_TILE_P;
    const int tiles = kst_objective_grid(h, S, &n_blk, &tile_p);
    if ((size_t)n_blk * S > KST_NOALLOC_MAX_PARTIALS) n_blk = KST_NOALLOC_MAX_PARTIALS / S;  // >= 64 by the swarm limit
    KstObjParams prm;
    kst_objective_params(h, S, K, d_points, d_weights, scratch, tile_p, 0, &prm);
    const dim3 grid(n_blk, tiles);
    KST_MODEL_SWITCH(h->model_id, rc = kst_obj_launch<MF>(h, prm, grid, K));
    if (rc) return rc;
    KST_CUDA(cudaGetLastError());
    k_obj_finish<<<(S + 127) / 128, 128, 0, h->stream>>>(scratch, n_blk, h->offset, d_fx, S, 0);
    KST_CUDA(cudaGetLastError());
    return KST_OK;
}
