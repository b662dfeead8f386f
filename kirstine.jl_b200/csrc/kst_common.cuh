// Shared host-side definitions of libkirstine_b200.so (handle layout, error helpers,
// launch interfaces between the translation units).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <vector>

#include "../../include/kirstine_b200.h"
#include "kst_models.cuh"

#ifndef KST_NT
#define KST_NT 256          // threads per CTA of the evaluation kernels
#endif
#ifndef KST_MINB
#define KST_MINB 2          // minimum resident CTAs per SM the objective kernel is compiled for
#endif
#ifndef KST_TILE_P
#define KST_TILE_P 8        // particles per CTA tile (objective kernel)
#endif
#define KST_GD_NT 128       // threads (= directions) per CTA of the Gateaux kernel
#define KST_GD_CHUNK 128    // draws staged per shared-memory chunk in the Gateaux kernel (= KST_GD_NT: one draw per thread)
#define KST_PEER_CAP 16384  // doubles per peer-exchange buffer (larger all-reduces go through NCCL)
#define KST_PEER_MAX 16     // ranks in one peer-exchange group

void kst_set_error(const char *fmt, ...);

#define KST_CUDA(call)                                                                      \
    do {                                                                                    \
        cudaError_t e__ = (call);                                                           \
        if (e__ != cudaSuccess) {                                                           \
            kst_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__)); \
            return KST_ECUDA;                                                               \
        }                                                                                   \
    } while (0)

// kst_problem::model_id holds the public model id, except for KST_MODEL_LINLOGREG whose functor depends on the
// covariate dimension: there it is KST_FID_LINLOGREG + d
#define KST_FID_LINLOGREG 60

// dispatch on the registered model family; the body sees the functor type as MF
#define KST_MODEL_SWITCH(id, ...)                                        \
    switch (id) {                                                        \
    case KST_MODEL_EMAX: { using MF = KstEmax; __VA_ARGS__; } break;     \
    case KST_MODEL_SIGEMAX: { using MF = KstSigEmax; __VA_ARGS__; } break; \
    case KST_MODEL_TPC: { using MF = KstTpc; __VA_ARGS__; } break;       \
    case KST_MODEL_VU: { using MF = KstVu; __VA_ARGS__; } break;         \
    case KST_MODEL_DTR: { using MF = KstDtr; __VA_ARGS__; } break;       \
    case KST_FID_LINLOGREG + 1: { using MF = KstLinLogReg<1>; __VA_ARGS__; } break; \
    case KST_FID_LINLOGREG + 2: { using MF = KstLinLogReg<2>; __VA_ARGS__; } break; \
    case KST_FID_LINLOGREG + 3: { using MF = KstLinLogReg<3>; __VA_ARGS__; } break; \
    case KST_MODEL_SINLOGREG: { using MF = KstSinLogReg; __VA_ARGS__; } break; \
    case KST_MODEL_VUAR: { using MF = KstVuAr; __VA_ARGS__; } break;     \
    default: kst_set_error("unregistered model id %d", id); return KST_EINVAL; \
    }

struct KstPso {
    int S = 0, K = 0, mode = 0, iter = 0;
    int pending_n_blk = 0;  // > 0: fx still sits in h->partials as per-CTA sums, k_pso_best finishes it ...
    double pending_offset = 0.0;  // ... adding this
    int pending_nan = 0;          // ... or filling NaN (Gateaux mode at a singular design)
    double c1 = 0, c2 = 0;
    uint64_t seed = 0;
    int64_t n_eval = 0;
    // device state: points are [S][K][d], weights [S][K]; all carved out of `arena` (kst_pso_alloc)
    double *x_pts = nullptr, *x_w = nullptr, *p_pts = nullptr, *p_w = nullptr, *v_pts = nullptr,
           *v_w = nullptr, *g_pts = nullptr, *g_w = nullptr, *fx = nullptr, *fp = nullptr,
           *fg = nullptr;
    int *fixw = nullptr, *fixp = nullptr;
    double *best_val = nullptr; int *best_idx = nullptr; unsigned int *best_counter = nullptr;  // k_pso_best scratch (arena)
    // persistent swarm kernel (kst_pso_fused.cuh): published positions [2][S][K*d] / [2][S][K], grid barrier counter
    // (+ error word) and how far it has counted
    double *x2_pts = nullptr, *x2_w = nullptr;
    unsigned long long *bar = nullptr;
    unsigned long long bar_count = 0;
    int fused_off = 0;  // this swarm does not fit the persistent kernel: do not ask again
    double *arena = nullptr;
    size_t arena_cap = 0;  // doubles
    // what a run sends back, carved out of `out` (device) and mirrored in `h_out` (pinned host) by
    // kst_pso_ensure_trace: status word, trace of tr_n entries, packed snapshot of the initial state
    int *d_status = nullptr;
    double *tr_fx = nullptr, *tr_pts = nullptr, *tr_w = nullptr, *snap = nullptr;
    int tr_n = 0;
    double *out = nullptr, *h_out = nullptr;
    size_t out_cap = 0;    // doubles
};

struct kst_problem {
    int device = 0;
    int model_id = 0, crit = 0, trafo = 0;
    int r = 0, t = 0, m = 0, d = 0;
    int64_t N = 0;      // draws held by this handle (its shard)
    size_t ld = 0;      // padded leading dimension of the per-draw SoA arrays
    KstModelParams mp{};
    double lb[KST_MAX_D]{}, ub[KST_MAX_D]{};
    cudaStream_t stream = nullptr;
    int sm_count = 0;

    // per-draw SoA arrays, each [rows][ld]
    double *theta = nullptr;   // r   raw parameters
    double *pw = nullptr;      // 1   prior weights
    double *dc = nullptr;      // NDC hoisted per-draw constants of the model functor
    double *cs = nullptr;      // r   column scales
    double *aux = nullptr;     // A/Identity: r (1/cs^2); DeltaMethod: t*r (DT/cs); else unused
    int n_aux = 0;
    double dcmax[8] = {};      // max_j |dc[c][j]| over this handle's draws (NaN / Inf poison the entry)
    // regularised normal approximation (kst_problem_set_regularization): Rt_j = D_j^-1 R D_j^-1, packed upper
    // [r(r+1)/2][ld], added to every Mt_j; nullptr = FisherMatrix
    double *reg = nullptr;
    double offset = 0.0;       // sum_j pw_j * 2 sum_i log|cs_ij|  (D / Identity), else 0
    int uniform_w = 0;         // all prior weights of this handle are the same positive finite number ...
    double w_uniform = 0.0;    // ... namely this one

    // scratch for the objective kernel: partial sums [n_blk][S]
    double *partials = nullptr;
    size_t partials_cap = 0;
    // staging for the host-buffer entry points
    double *stage_pts = nullptr, *stage_w = nullptr, *stage_fx = nullptr;
    size_t stage_cap_particles = 0, stage_cap_rows = 0;

    // Gateaux constants at the last prepared design: At [r*r][ld] (scaled space), trB [ld]
    double *gA = nullptr, *gtrB = nullptr;
    double *g_wtrB = nullptr;  // scratch of gateaux_prepare: pw_j (trB_j - tr[A_j R]) per draw, summed into g_offset
    double *gAp = nullptr;     // what the evaluation kernel streams: packed upper of pw_j At_j, off-diagonals doubled, [r(r+1)/2][ld]
    int g_ready = 0, g_singular = 0;
    double g_offset = 0.0;     // sum_j pw_j trB_j
    int *d_flag = nullptr;     // 16 bytes: device flag word (singular detection) + one double of scratch at offset 8
    double *gd_dirs = nullptr, *gd_out = nullptr, *gd_partials = nullptr;
    size_t gd_cap = 0, gd_partials_cap = 0;
    double *g_out = nullptr;    // kst_gateaux_constants: A_j in the reference's scale, [N][r*r]
    double *gd_ones = nullptr;  // unit weights for directions evaluated through the objective kernel (dense covariance)
    size_t gd_ones_cap = 0;

    KstPso pso;

    // communicator (optional)
    void *comm = nullptr;
    int nranks = 1, rank = 0;
    // peer-memory exchange over NVLink (optional, kst_comm_peer_*): each rank owns one exchange block
    // {double data[2][KST_PEER_CAP]; unsigned long long flag[2]; int error;} that every peer maps through CUDA IPC
    void *peer_local = nullptr;          // this rank's block (cudaMalloc)
    void *peer_map[KST_PEER_MAX] = {};   // all ranks' blocks as seen from this GPU (own entry = peer_local)
    int peer_ready = 0;
    unsigned long long peer_seq = 0;     // collectives issued so far (identical on every rank)

    // profiling of the evaluation kernel
    int prof_on = 0;   // 0 off, 1 evaluation kernel, 2 swarm move, 3 best update / multi-GPU tail (kst_profile_enable)
    std::vector<cudaEvent_t> prof_events;  // start/stop pairs, resolved by kst_profile_read
};

// kst_objective.cu
int kst_launch_prep_draws(kst_problem *h, const double *d_theta_aos, const double *d_dt, double *d_scratch);
// defer_n_blk != nullptr: the per-CTA partial sums stay in h->partials ([*defer_n_blk][S]) and the caller finishes them
// (kst_api.cu fuses that step into the peer-memory all-reduce)
// gd = 1: the Gateaux derivative of the prepared design in the swarm's one-point directions (K = 1) instead of the
// objective: same kernel body, the draw's "criterion" is tr_U[A_j M(delta, theta_j)]
int kst_launch_objective(kst_problem *h, int S, int K, const double *d_pts, const double *d_w,
                         double *d_fx, int *defer_n_blk = nullptr, int gd = 0);
void kst_objective_finish_terms(const kst_problem *h, int gd, double *offset, int *nan_fill);
int kst_launch_infomatrix(kst_problem *h, int K, const double *d_pts, const double *d_w,
                          int64_t draw, double *d_M);
int kst_objective_grid(const kst_problem *h, int S, int *n_blk, int *tile_p);
// n_steps swarm evaluations (ticks; the first one optimizer_state's initial evaluation when first_is_init) in one
// cooperative launch.  Returns 1 when the problem does not fit the persistent kernel (nothing was launched).
// Trace slot tr_slot0 + i receives (fg, g) after evaluation i when with_trace; snap: see KstFusedParams.
int kst_launch_pso_fused(kst_problem *h, int n_steps, int first_is_init, int tr_slot0, int with_trace, int *d_status,
                         double *d_snap);
int kst_sum_fixed(kst_problem *h, const double *d_x, int64_t n, double *host_out);
int kst_sum_fixed_enqueue(kst_problem *h, const double *d_x, int64_t n);
// kst_gateaux.cu
int kst_launch_gateaux_prepare(kst_problem *h, int K, const double *d_pts, const double *d_w);
int kst_launch_gateaux_eval(kst_problem *h, int64_t G, const double *d_dirs, double *d_out);
// kst_pso.cu
int kst_pso_alloc(kst_problem *h, int S, int K);
void kst_pso_free(kst_problem *h);
int kst_pso_ensure_trace(kst_problem *h, int n);
int kst_pso_fetch(kst_problem *h, int with_snap, int *status, const double **tr_fx, const double **tr_pts,
                  const double **tr_w, const double **snap);
struct KstPsoDev;
void kst_pso_make_dev(const kst_problem *h, KstPsoDev *out);
int kst_launch_pso_init(kst_problem *h, const double *d_proto_pts, const double *d_proto_w, int P);
int kst_launch_pso_move(kst_problem *h);
int kst_launch_pso_best(kst_problem *h, int init, int trace_slot);
// kst_comm.cu
int kst_allreduce_sum(kst_problem *h, double *d_buf, size_t count);
int kst_objective_allreduce(kst_problem *h, int S, int K, const double *d_pts, const double *d_w, double *d_fx);
// One launch after the evaluation kernel of a swarm iteration on several GPUs: finish the per-CTA partial sums, sum
// over the ranks through the peer-memory exchange, pso_update_best!.  Returns 1 when not applicable (one GPU, no
// peer exchange, or S beyond its capacity).
int kst_launch_pso_tail(kst_problem *h, int trace_slot, int n_blk, double offset, int nan_fill);
void kst_comm_release(kst_problem *h);
// kst_api.cu
int kst_ensure_partials(kst_problem *h, size_t n);
int kst_ensure_stage(kst_problem *h, size_t particles, size_t rows);
int kst_pso_optimize_enqueue(kst_problem *h, int S, int K, int iterations, double c1, double c2, const double *proto_points,
                             const double *proto_weights, int P, const int *fixw, const int *fixp, uint64_t seed, int mode,
                             int want_snapshot);

// kst_profile_enable: a pair of CUDA events around the launches of one stage, resolved by kst_profile_read
#define KST_PROF_EVAL 1
#define KST_PROF_MOVE 2
#define KST_PROF_BEST 3
static inline int kst_prof_begin(kst_problem *h, int stage)
{
    if (h->prof_on != stage) return KST_OK;
    cudaEvent_t e0, e1;
    KST_CUDA(cudaEventCreate(&e0));
    KST_CUDA(cudaEventCreate(&e1));
    h->prof_events.push_back(e0);
    h->prof_events.push_back(e1);
    KST_CUDA(cudaEventRecord(e0, h->stream));
    return KST_OK;
}
static inline int kst_prof_end(kst_problem *h, int stage)
{
    if (h->prof_on != stage) return KST_OK;
    KST_CUDA(cudaEventRecord(h->prof_events.back(), h->stream));
    return KST_OK;
}
