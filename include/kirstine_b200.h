/*
 * kirstine_b200.h -- C ABI of the B200 engine behind Kirstine.jl's design-objective hot path.
 *
 * The reference (lsandig/Kirstine.jl v0.5.0) is pure Julia and has no FFI; its plugin surface
 * is Julia multiple dispatch.  Each entry point below names the reference function(s) it
 * stands in for (paths relative to the reference checkout).  INTEGRATION.md shows the ccall
 * stubs a Kirstine maintainer would add, and the ctypes binding used by this repo's tests.
 *
 * Conventions
 *  - every function returns 0 on success, a negative KST_E* code otherwise;
 *    kst_last_error() returns a thread-local message for the last failure;
 *  - no exceptions, Julia or Python objects cross the boundary; the caller owns every buffer
 *    it passes, the handle owns all device memory, streams and (optionally) one communicator;
 *  - one handle drives ONE GPU; a handle is single-threaded like the reference's Workspaces;
 *  - all floating point is IEEE double; matrices are column-major like Julia's Matrix{Float64};
 *  - "points" of a design measure are d x K column-major (DesignMeasure.points,
 *    src/designmeasure.jl:19-42), i.e. point k occupies points[k*d .. k*d+d-1];
 *  - there is no CPU fallback: every entry point needs a CUDA device.
 */
#ifndef KIRSTINE_B200_H
#define KIRSTINE_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define KST_VERSION 100

#if defined(__GNUC__)
#define KST_API __attribute__((visibility("default")))
#else
#define KST_API
#endif

/* error codes */
#define KST_OK 0
#define KST_EINVAL (-1)   /* bad argument / unsupported combination */
#define KST_ECUDA (-2)    /* CUDA runtime failure (message has the cudaError string) */
#define KST_ENOMEM (-3)
#define KST_ESTATE (-4)   /* call order violated (e.g. gateaux_eval before gateaux_prepare) */
#define KST_ENCCL (-5)    /* NCCL unavailable or failed */
#define KST_ELIMIT (-6)   /* size beyond a documented fixed limit */
#define KST_ESINGULAR (-7)  /* kst_exchange_run: the reference's SingularException (singular candidate design) */
#define KST_EOBJECTIVE (-8) /* kst_exchange_run: the reference's ErrorExceptions of optimizer_state (src/pso.jl:119-123) */

/* Registered model families = device functors standing in for the user's jacobianmatrix!,
 * update_model_vcov!, unit_length, dimension (src/user.jl:18-98).  The reference ships no
 * concrete models; these restate its test/doc models:
 *   KST_MODEL_EMAX     test/example-emax.jl:4-47         params {sigma_squared},  theta = (e0, emax, ec50)
 *   KST_MODEL_SIGEMAX  docs/src/tutorial.md:107-142      params {sigma},          theta = (e0, emax, ed50, h)
 *   KST_MODEL_TPC      test/example-compartment.jl:10-21 params {sigma},          theta = (a, e, s)
 *   KST_MODEL_VU       test/example-vector.jl:9-41       params {sigma_squared},  theta = (a, e, s), unit length m
 *   KST_MODEL_DTR      test/example-dtr.jl:4-47, docs/src/dtr.md:87-129  params {sigma}, theta = (a, e, e0, emax, ec50),
 *                      unit length m, design points (dose[, time or delta-t])
 * Unregistered model types must keep using stock Kirstine. */
#define KST_MODEL_EMAX 1
#define KST_MODEL_SIGEMAX 2
#define KST_MODEL_TPC 3
#define KST_MODEL_VU 4
#define KST_MODEL_DTR 5
/* Logistic regression, the user-defined Model supertype of docs/src/extend-model.md:59-106 (its own
 * average_fishermatrix!: M = sum_k w_k p_k (1 - p_k) J_k' J_k with p = expit(predictor), J = predictor Jacobian;
 * no unit covariance: model_params is ignored, pass one dummy value):
 *   KST_MODEL_LINLOGREG  docs/src/extend-model.md:123-179  linear predictor on a d-dimensional covariate copied
 *                        from the design point (CopyVector = KST_CP_COPY), d = 1..3, theta = (beta_0, ..., beta_d)
 *   KST_MODEL_SINLOGREG  docs/src/extend-model.md:330-355  predictor a sin(2 pi b (t - c)), theta = (a, b, c) */
#define KST_MODEL_LINLOGREG 6
#define KST_MODEL_SINLOGREG 7
/* A NonlinearRegression whose unit covariance Sigma(c) is NOT diagonal -- what update_model_workspace! /
 * average_fishermatrix! (src/model-nonlinear-regression.jl:39-49,60-86: potrf! of Sigma(c_k), trsm! of every unit's
 * Jacobian) allow for any user model, none of the reference's own examples uses, and the other registered families
 * (diagonal Sigma) do not exercise:
 *   KST_MODEL_VUAR  mean function, parameters and covariates of KST_MODEL_VU, errors within a unit AR(1)-correlated:
 *                   Sigma_ij = sigma_squared rho^|i-j| sqrt((1 + t_i)(1 + t_j)); params {sigma_squared, rho}, |rho| < 1;
 *                   rho = 0 is KST_MODEL_VU.  Covariate parameterization KST_CP_EQUITIME, unit length m <= KST_MAX_UNIT. */
#define KST_MODEL_VUAR 8

/* covariate parameterizations = map_to_covariate! (src/user.jl:68):
 *   KST_CP_COPY      CopyDose / CopyTime / JustCopy: covariate = design point (src/user.jl:185-216)
 *   KST_CP_EQUITIME  EquiTime (test/example-vector.jl:45-50): time_i = (i-1) * z[1], i = 1..m
 * for KST_MODEL_DTR (docs/src/dtr.md; z = design point, cp_params in brackets):
 *   KST_CP_DTR_COPYBOTH        CopyBoth :197-203          dose = z[1], time_1 = z[2]              (d = 2, m = 1)
 *   KST_CP_DTR_FIXEDDOSE       FixedDose :331-339         dose = [dose], time_1 = z[1]            (d = 1, m = 1)
 *   KST_CP_DTR_FIXEDTIMES      FixedTimes :401-409        dose = z[1], time = [time_1..time_m]    (d = 1)
 *   KST_CP_DTR_EQUIDISTANT     EquidistantTimes :484-495  dose = z[1], time_i = (i-1) z[2]        (d = 2)
 *   KST_CP_DTR_LOGEQUIDISTANT  LogEquidistantTimes :663-681 dose = z[1], time = 0, z[2], b z[2], ... [b] (d = 2, m >= 2) */
#define KST_CP_COPY 0
#define KST_CP_EQUITIME 1
#define KST_CP_DTR_COPYBOTH 2
#define KST_CP_DTR_FIXEDDOSE 3
#define KST_CP_DTR_FIXEDTIMES 4
#define KST_CP_DTR_EQUIDISTANT 5
#define KST_CP_DTR_LOGEQUIDISTANT 6

#define KST_CRIT_D 0 /* DCriterion, src/criterion-d.jl:29 */
#define KST_CRIT_A 1 /* ACriterion, src/criterion-a.jl:29 */
#define KST_CRIT_DEXP 2 /* det^(1/t), the user-extension DexpCriterion of docs/src/extend-criterion.md:100-183 */

#define KST_TRAFO_IDENTITY 0 /* Identity,    src/transformation.jl:11 */
#define KST_TRAFO_DELTA 1    /* DeltaMethod, src/transformation.jl:59-61 */

#define KST_PSO_MODE_OBJECTIVE 0 /* particles are K-point designs, f = objective!        */
#define KST_PSO_MODE_GATEAUX 1   /* particles are one-point designs, f = gateauxderivative! */

/* fixed limits of this build */
#define KST_MAX_R 5      /* parameter dimension */
#define KST_MAX_D 4      /* design-region dimension */
#define KST_MAX_UNIT 16  /* unit length m */
#define KST_MAX_ROWS 128 /* K * m per design measure */

typedef struct kst_problem kst_problem;

KST_API const char *kst_last_error(void);
KST_API int kst_version(void);

/* Number of visible CUDA devices (0 if none). */
KST_API int kst_device_count(int *count);

/*
 * DesignProblem (src/designproblem.jl:22-79) restricted to what the hot path reads: model,
 * covariate parameterization, criterion, transformation, prior sample, design region.
 * Replaces allocate_workspaces (src/designproblem.jl:130-138) and trafo_constants
 * (src/transformation.jl:71-92): the caller evaluates DeltaMethod's Jacobian on the host for
 * every draw (arbitrary user closure) and passes the result.
 *
 *  cp_params  parameters of the covariate parameterization (may be NULL when it has none)
 *  theta    N*r doubles, array-of-structs exactly like Julia's Vector{P} of isbits parameter
 *           structs (PriorSample.p, src/priorknowledge.jl:11-38); transposed to SoA on upload
 *  prior_w  N prior weights (PriorSample.weight)
 *  dt_jm    N*(t*r) doubles, per draw one column-major t x r Matrix{Float64}
 *           (TCDeltaMethod.jm, src/transformation.jl:67-69); NULL for KST_TRAFO_IDENTITY
 *  t        codomain dimension (== r for identity)
 *  lb, ub   d lower / upper bounds of the DesignInterval (src/designregion.jl:22-51)
 *
 * With several GPUs each process passes only ITS shard of (theta, prior_w, dt_jm); see
 * kst_comm_init.
 */
KST_API int kst_problem_create(int model_id, const double *model_params, int n_model_params, int cp_id,
                       const double *cp_params, int n_cp_params,
                       int criterion_id, int trafo_id, int r, int t, int m, int d, int64_t N,
                       const double *theta, const double *prior_w, const double *dt_jm,
                       const double *lb, const double *ub, int device_id, kst_problem **out);
KST_API int kst_problem_destroy(kst_problem *h);

/*
 * Multi-GPU: the prior sample is sharded over processes (one per GPU); per swarm evaluation the
 * S partial objectives are summed with one ncclAllReduce (double, sum) on the handle's stream.
 * The reference has no counterpart (it is serial: src/pso.jl:173-179, src/designproblem.jl:205-215).
 * kst_comm_unique_id fills a 128-byte ncclUniqueId on rank 0; the host broadcasts it by any
 * means (torch.distributed, MPI, a file) and every rank calls kst_comm_init.  NCCL is loaded
 * with dlopen("libnccl.so.2") on first use, so single-GPU users need no NCCL.
 */
KST_API int kst_comm_unique_id(void *id128);
KST_API int kst_comm_init(kst_problem *h, int nranks, int rank, const void *id128);
/*
 * The same all-reduce over NVLink peer memory instead of NCCL (one box, GPUs with peer access): every rank
 * publishes its S partial sums in an exchange block that all peers map through CUDA IPC, and ONE single-CTA
 * kernel per evaluation publishes, waits for the peers' sequence flags and sums all vectors in rank order
 * (identical bits on every rank; swarm iterations run the same exchange per slice of 128 particles inside the
 * kernel that updates the bests).  kst_comm_peer_handle allocates this rank's block and returns its 64-byte
 * cudaIpcMemHandle_t; the host gathers the nranks handles (rank order, 64 bytes each) and every rank calls
 * kst_comm_peer_open.  Vectors longer than 16384 doubles still go through NCCL (kst_comm_init).  A peer that
 * does not arrive within 10 s makes the result NaN and sets the error word read by kst_comm_peer_error.
 * The reference has no counterpart.
 */
KST_API int kst_comm_peer_handle(kst_problem *h, int nranks, int rank, void *handle64);
KST_API int kst_comm_peer_open(kst_problem *h, const void *handles);
KST_API int kst_comm_peer_error(kst_problem *h, int *error);
/* Unmap the peers' blocks and go back to NCCL (or to single-GPU operation); all ranks must do this together. */
KST_API int kst_comm_peer_close(kst_problem *h);

/*
 * pso_evaluate_objective! (src/pso.jl:173-179) calling objective! (src/designproblem.jl:192-224)
 * for S design measures of K points each, all at once.
 *  points   S*K*d doubles (per particle d x K column-major), weights S*K, fx S (out).
 * fx is -Inf where the reference returns -Inf (PosDefException / SingularException / log_det!
 * of a singular matrix); NaN propagates like in the reference (e.g. prior weight 0 on a
 * singular draw).  Host-buffer variant: copies in, evaluates, all-reduces if a communicator is
 * attached, copies out, synchronises.
 */
KST_API int kst_objective_batch(kst_problem *h, int S, int K, const double *points, const double *weights,
                        double *fx);
/* Same with buffers already in device memory of the handle's GPU; asynchronous on the handle's
 * stream (kst_stream_synchronize waits). */
KST_API int kst_objective_batch_device(kst_problem *h, int S, int K, const double *d_points,
                               const double *d_weights, double *d_fx);
/*
 * No counterpart in the reference.  The same evaluation as kst_objective_batch_device for callers that
 * must not allocate device memory while they run (all device memory already given to another pool, or a
 * loop that cannot take the device-wide synchronisation of cudaMalloc / cudaFree): the per-CTA partial
 * sums live in a fixed 2 MiB array inside the library image, which every GPU that loads the library
 * keeps until exit.  At most 4096 particles per call (KST_ELIMIT beyond).  The first call on each GPU
 * looks up the array's device address and should come before any stream capture; since the CUDA runtime
 * loads each kernel on its first launch and the evaluation kernel is compiled per (model family, criterion,
 * transformation), make one warm-up call per such combination in use, not only per GPU.  At most 2^29 draws
 * per handle, like the other entry points (KST_ELIMIT beyond).  Calls that overlap
 * on one GPU race on the shared array: serialise them.  Asynchronous on the handle's stream; sums only
 * the draws held by this handle (no collective even when a communicator is attached).
 */
KST_API int kst_objective_batch_noalloc(kst_problem *h, int S, int K, const double *d_points,
                                        const double *d_weights, double *d_fx);
KST_API int kst_stream_synchronize(kst_problem *h);
/* The handle's cudaStream_t, for callers that enqueue their own work or copies in order. */
KST_API void *kst_stream(kst_problem *h);

/*
 * normal_approximation (src/designproblem.jl:31,55; src/design-common.jl:74-76).  A new handle uses FisherMatrix
 * (the only NormalApproximation the reference ships).  This call switches it to the RegularizedFisherMatrix(R) of
 * docs/src/extend-approximation.md:69-79, whose informationmatrix! adds the constant matrix R to every average
 * Fisher matrix -- in objective!, in the Gateaux constants and in the direction's matrix
 * (src/designproblem.jl:207,254; src/design-common.jl:110).  R is r x r column-major; like every normalized
 * information matrix of the reference only its upper triangle is read.  R = NULL switches back to FisherMatrix.
 * A prepared Gateaux state is invalidated.
 */
KST_API int kst_problem_set_regularization(kst_problem *h, const double *R);

/* informationmatrix (src/design-common.jl:96-112) for draw `draw_index` (0-based) of the prior
 * sample: full symmetric r x r matrix. */
KST_API int kst_informationmatrix(kst_problem *h, int K, const double *points, const double *weights,
                          int64_t draw_index, double *M);

/*
 * gateaux_constants (src/criterion-d.jl:43-87, src/criterion-a.jl:47-77) at design `at`; the
 * constants stay on the device.  *singular = 1 where the reference throws SingularException
 * (then kst_gateaux_eval fills NaN like gateauxderivative, src/designproblem.jl:281-284).
 */
KST_API int kst_gateaux_prepare(kst_problem *h, int K, const double *points, const double *weights,
                        int *singular);
/* GCPriorSample copy-out (src/design-common.jl:29-32): A = N*(r*r) column-major dense, trB = N. */
KST_API int kst_gateaux_constants(kst_problem *h, double *A, double *tr_B);
/* gateauxderivative! (src/designproblem.jl:239-258) for G one-point directions dirs[G*d]. */
KST_API int kst_gateaux_eval(kst_problem *h, int64_t G, const double *dirs, double *gd);

/*
 * optimizer_state (src/pso.jl:83-125): particle i < P is prototype i, the others are prototype
 * 0 randomised by ap_rand! (src/designmeasure-abstractpoint.jl:73-107); v = 0; evaluate; best.
 * fixw / fixp: DesignConstraints (src/designmeasure-abstractpoint.jl:23-58), K ints each.
 * *status: 0 ok, 1 "infinite objective value at best initial point", 2 "objective evaluates to
 * NaN on all initial points" (the reference's two ErrorExceptions).
 * The swarm lives on the device from here on.
 */
KST_API int kst_pso_init(kst_problem *h, int S, int K, double c1, double c2, const double *proto_points,
                 const double *proto_weights, int P, const int *fixw, const int *fixp,
                 uint64_t seed, int mode, int *status);
/*
 * n_iters times tick! (src/pso.jl:127-141).  trace_fx[n_iters], trace_points[n_iters*K*d],
 * trace_weights[n_iters*K] receive maximum/maximizer after each tick (optimize's trace_fx /
 * trace_x, src/optimize.jl:54-57); any of them may be NULL.  With all three NULL the call only enqueues the
 * iterations on the handle's stream and returns (kst_stream_synchronize / kst_pso_get_state wait for them).
 */
KST_API int kst_pso_run(kst_problem *h, int n_iters, double *trace_fx, double *trace_points,
                double *trace_weights);
/* PsoState copy-out (src/pso.jl:63-76), what trace_state = true deep-copies; NULL skips. */
KST_API int kst_pso_get_state(kst_problem *h, double *x_points, double *x_weights, double *p_points,
                      double *p_weights, double *g_points, double *g_weights, double *v_points,
                      double *v_weights, double *fx, double *fp, double *fg, int64_t *n_eval);

/*
 * optimize (src/optimize.jl:42-70) with a Pso optimizer in ONE call: optimizer_state (src/pso.jl:83-125), then
 * iterations - 1 times tick! (src/pso.jl:127-141), with nothing returning to the host in between.  Arguments as
 * kst_pso_init.  The traces have `iterations` entries like OptimizationResult.trace_fx / trace_x
 * (src/optimize.jl:28-40): entry 0 is the initial state, entry i the maximum / maximizer after tick i; any of them
 * may be NULL.  state0 (or NULL): the PsoState optimize keeps as trace_state[1] (src/optimize.jl:51-53), packed as
 * x_points[S*K*d], x_weights[S*K], p_points, p_weights, v_points, v_weights, g_points[K*d], g_weights[K], fx[S],
 * fp[S], fg -- 3*S*K*(d+1) + K*(d+1) + 2*S + 1 doubles.  *status as kst_pso_init (the ticks run even when it is
 * non-zero; the reference throws instead, so discard the results then).  The swarm stays on the device:
 * kst_pso_get_state and kst_pso_run continue from it.
 * Problems whose evaluation grid fits one wave of CTAs run the whole call as one persistent kernel (tables staged
 * once, one grid barrier per iteration); larger ones use three launches per iteration.  The results are identical.
 */
KST_API int kst_pso_optimize(kst_problem *h, int S, int K, int iterations, double c1, double c2,
                             const double *proto_points, const double *proto_weights, int P, const int *fixw,
                             const int *fixp, uint64_t seed, int mode, int *status, double *trace_fx,
                             double *trace_points, double *trace_weights, double *state0);

/*
 * solve_with(dp, ::Exchange, trace_state = false) (src/solve-exchange.jl:116-179) in one call: `steps` times
 *   simplify (src/designproblem.jl:185-190) -> gateaux_constants -> optimize the Gateaux derivative over one-point
 *   directions (prototypes: the candidate's points) -> prepend the direction (mixture, or simplify_merge with distance 0
 *   when it is already in the support) -> optimize the weights with all points fixed.
 * Both optimizers are Pso and run on the device (kst_pso_optimize); per step only the singularity flag, the direction
 * and the new weights return to the host, where the K <= Kmax point bookkeeping is done.
 *
 * Exchange (src/solve-exchange.jl:15-69) as a plain struct.  maxweight / maxdist: simplify_args for simplify_drop /
 * simplify_merge; unique_a / unique_b: the two keyword arguments of the registered covariate parameterizations'
 * simplify_unique methods (docs/src/dtr.md:229-247: minposdose, minpostime; :498-523: minposdose, minposdelta), ignored
 * by the others.  seed: the optimizers of step i (0-based) use the 2i-th and (2i+1)-th value of the sequence
 * s -> s * 6364136223846793005 + 1442695040888963407 (mod 2^64) started at `seed`.
 */
typedef struct kst_exchange_options {
    int steps;
    int swarm_direction, iters_direction;
    double c1_direction, c2_direction;
    int swarm_weight, iters_weight;
    double c1_weight, c2_weight;
    double maxweight, maxdist, unique_a, unique_b;
    uint64_t seed;
} kst_exchange_options;
/*
 *  cand_points[K0*d], cand_weights[K0]  the strategy's candidate
 *  Kmax        capacity, in design points, of every per-step output slot (a candidate never has more than
 *              K0 + steps points); Kmax * m <= KST_MAX_ROWS
 *  K_out, points_out[Kmax*d], weights_out[Kmax]   solution(er): the last weight optimizer's maximizer
 *  dir_trace_fx[steps*iters_direction], dir_trace_x[steps*iters_direction*d]   trace_fx / trace_x of every direction run
 *  w_trace_fx[steps*iters_weight]; w_trace_points / w_trace_weights: per step a slot of iters_weight*Kmax*d resp.
 *              iters_weight*Kmax doubles holding iters_weight designs of w_K[step] points, packed
 *  dir_state0 / w_state0 (or NULL)  per step the packed initial PsoState (layout: kst_pso_optimize), slots of
 *              3*S*(d+1) + (d+1) + 2*S + 1 resp. 3*S*Kmax*(d+1) + Kmax*(d+1) + 2*S + 1 doubles
 *  steps_done  completed steps (== steps on success)
 * Any trace pointer may be NULL.  Errors: KST_ESINGULAR where the reference throws SingularException,
 * KST_EOBJECTIVE for optimizer_state's two ErrorExceptions; outputs are then undefined beyond steps_done.
 */
KST_API int kst_exchange_run(kst_problem *h, const kst_exchange_options *opt, int K0, const double *cand_points,
                             const double *cand_weights, int Kmax, int *K_out, double *points_out, double *weights_out,
                             double *dir_trace_fx, double *dir_trace_x, double *w_trace_fx, double *w_trace_points,
                             double *w_trace_weights, int *w_K, double *dir_state0, double *w_state0, int *steps_done);

/* measurement helpers (bench.py): CUDA-event timing of one stage on the handle's stream -- on = 1 the evaluation
 * kernel (k_objective / k_gateaux_eval), 2 the swarm move, 3 the best update (on several GPUs: the tail kernel with
 * the peer exchange), 0 off; kst_profile_read returns the summed time and the number of launches since the last
 * read -- and a DFMA micro-kernel that measures this GPU's FP64 pipe peak. */
KST_API int kst_profile_enable(kst_problem *h, int on);
KST_API int kst_profile_read(kst_problem *h, double *total_ms, int64_t *launches);
KST_API int kst_fp64_peak_probe(int device_id, double *tflops);
/* accuracy self-test of the engine's own FP64 exp, reciprocal and log (csrc/kst_math.cuh) on n host-supplied
 * arguments: exp_out[i] = exp(x[i]), rcp_out[i] = 1/x[i], log_out[i] = log|x[i]| (the latter two are only
 * meaningful for normal, non-zero x[i]). */
KST_API int kst_math_selftest(int device_id, int64_t n, const double *x, double *exp_out, double *rcp_out,
                              double *log_out);

#ifdef __cplusplus
}
#endif
#endif
