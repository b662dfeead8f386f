// C ABI of libkirstine_b200.so (declared in include/kirstine_b200.h): handle management,
// host<->device staging and the PSO driver loop (communication between ranks: kst_comm.cu).
#include <dlfcn.h>
#include <stdarg.h>
#include <stddef.h>
#include <string.h>

#include <cmath>
#include <mutex>
#include <new>

#include "kst_common.cuh"

static thread_local char g_err[512] = "";

void kst_set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" const char *kst_last_error(void) { return g_err; }
extern "C" int kst_version(void) { return KST_VERSION; }

extern "C" int kst_device_count(int *count)
{
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) {
        cudaGetLastError();
        n = 0;
    }
    if (count) *count = n;
    return KST_OK;
}


// ---------------------------------------------------------------------------------------
// problem handle
// ---------------------------------------------------------------------------------------
static int model_dims(int model_id, int d, int *r, int *ndc)
{
    switch (model_id) {
    case KST_MODEL_LINLOGREG:
        if (d < 1 || d > 3) return -1;
        *r = d + 1; *ndc = d + 1; return 0;
    case KST_MODEL_SINLOGREG: *r = KstSinLogReg::R; *ndc = KstSinLogReg::NDC; return 0;
    case KST_MODEL_EMAX: *r = KstEmax::R; *ndc = KstEmax::NDC; return 0;
    case KST_MODEL_SIGEMAX: *r = KstSigEmax::R; *ndc = KstSigEmax::NDC; return 0;
    case KST_MODEL_TPC: *r = KstTpc::R; *ndc = KstTpc::NDC; return 0;
    case KST_MODEL_VU: *r = KstVu::R; *ndc = KstVu::NDC; return 0;
    case KST_MODEL_DTR: *r = KstDtr::R; *ndc = KstDtr::NDC; return 0;
    case KST_MODEL_VUAR: *r = KstVuAr::R; *ndc = KstVuAr::NDC; return 0;
    }
    return -1;
}

extern "C" int kst_problem_destroy(kst_problem *h)
{
    if (!h) return KST_OK;
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    kst_comm_release(h);
    kst_pso_free(h);
    void *ptrs[] = {h->theta, h->pw, h->dc, h->cs, h->aux, h->reg, h->partials, h->stage_pts, h->stage_w, h->stage_fx,
                    h->gA, h->gAp, h->gtrB, h->g_wtrB, h->d_flag, h->gd_dirs, h->gd_out, h->gd_partials, h->gd_ones, h->g_out};
    for (void *p : ptrs)
        if (p) cudaFree(p);
    for (cudaEvent_t e : h->prof_events) cudaEventDestroy(e);
    if (h->stream) cudaStreamDestroy(h->stream);
    delete h;
    return KST_OK;
}

extern "C" int kst_problem_create(int model_id, const double *model_params, int n_model_params, int cp_id,
                                  const double *cp_params, int n_cp_params, int criterion_id, int trafo_id, int r, int t, int m, int d, int64_t N,
                                  const double *theta, const double *prior_w, const double *dt_jm, const double *lb,
                                  const double *ub, int device_id, kst_problem **out)
{
    if (!out) { kst_set_error("problem_create: out is NULL"); return KST_EINVAL; }
    *out = nullptr;
    int mr = 0, ndc = 0;
    if (model_dims(model_id, d, &mr, &ndc)) {
        if (model_id == KST_MODEL_LINLOGREG) kst_set_error("KST_MODEL_LINLOGREG is registered for covariate dimension 1..3, got d = %d", d);
        else kst_set_error("unregistered model id %d", model_id);
        return KST_EINVAL;
    }
    if (r != mr) { kst_set_error("model %d has parameter dimension %d, got r = %d", model_id, mr, r); return KST_EINVAL; }
    if (criterion_id != KST_CRIT_D && criterion_id != KST_CRIT_A && criterion_id != KST_CRIT_DEXP) { kst_set_error("unknown criterion %d", criterion_id); return KST_EINVAL; }
    if (trafo_id == KST_TRAFO_IDENTITY) {
        if (t != r) { kst_set_error("identity transformation needs t == r"); return KST_EINVAL; }
    } else if (trafo_id == KST_TRAFO_DELTA) {
        if (!dt_jm) { kst_set_error("DeltaMethod needs dt_jm"); return KST_EINVAL; }
        // more rows than columns makes every information matrix singular (src/transformation.jl:86-88)
        if (t < 1 || t > r) { kst_set_error("DeltaMethod needs 1 <= t <= r, got t = %d", t); return KST_EINVAL; }
    } else { kst_set_error("unknown transformation %d", trafo_id); return KST_EINVAL; }
    // which (model, covariate parameterization, unit length, region dimension) combinations are registered
    int want_d = 1, want_ncp = 0;
    bool ok_m = m == 1, ok_cp = cp_id == KST_CP_COPY;
    if (model_id == KST_MODEL_LINLOGREG) want_d = d;  // range checked by model_dims
    if (model_id == KST_MODEL_VU || model_id == KST_MODEL_VUAR) {
        ok_m = m >= 1 && m <= KST_MAX_UNIT;
        ok_cp = cp_id == KST_CP_EQUITIME;
        if (model_id == KST_MODEL_VUAR && (n_model_params < 2 || !model_params || !(fabs(model_params[1]) < 1.0) || !(model_params[0] > 0.0))) {
            kst_set_error("KST_MODEL_VUAR needs model_params = {sigma_squared > 0, rho with |rho| < 1}");
            return KST_EINVAL;
        }
    } else if (model_id == KST_MODEL_DTR) {
        ok_m = m >= 1 && m <= KST_MAX_UNIT;
        switch (cp_id) {
        case KST_CP_DTR_COPYBOTH: ok_cp = true; want_d = 2; ok_m = m == 1; break;
        case KST_CP_DTR_FIXEDDOSE: ok_cp = true; want_ncp = 1; ok_m = m == 1; break;
        case KST_CP_DTR_FIXEDTIMES: ok_cp = true; want_ncp = m; break;
        case KST_CP_DTR_EQUIDISTANT: ok_cp = true; want_d = 2; break;
        case KST_CP_DTR_LOGEQUIDISTANT: ok_cp = true; want_d = 2; want_ncp = 1; ok_m = ok_m && m >= 2; break;
        default: ok_cp = false;
        }
    }
    if (!ok_cp) { kst_set_error("covariate parameterization %d not registered for model %d", cp_id, model_id); return KST_EINVAL; }
    if (!ok_m) { kst_set_error("bad unit length m = %d for model %d with covariate parameterization %d", m, model_id, cp_id); return KST_EINVAL; }
    if (d != want_d) { kst_set_error("design-region dimension d = %d, model %d with covariate parameterization %d needs %d", d, model_id, cp_id, want_d); return KST_EINVAL; }
    if (want_ncp > 0 && (!cp_params || n_cp_params != want_ncp)) {
        kst_set_error("covariate parameterization %d needs %d parameter(s), got %d", cp_id, want_ncp, cp_params ? n_cp_params : 0);
        return KST_EINVAL;
    }
    if (N < 1 || !theta || !prior_w || !lb || !ub) { kst_set_error("problem_create: empty prior sample or NULL buffer"); return KST_EINVAL; }
    if (n_model_params < 1 || !model_params) { kst_set_error("problem_create: model_params missing"); return KST_EINVAL; }
    for (int j = 0; j < d; j++)
        if (!(ub[j] > lb[j])) { kst_set_error("upper bounds must be strictly larger than lower bounds"); return KST_EINVAL; }

    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1) {
        cudaGetLastError();
        kst_set_error("no CUDA device: this engine has no CPU fallback");
        return KST_ECUDA;
    }
    if (device_id < 0 || device_id >= ndev) { kst_set_error("device %d out of range (have %d)", device_id, ndev); return KST_EINVAL; }
    KST_CUDA(cudaSetDevice(device_id));

    kst_problem *h = new (std::nothrow) kst_problem();
    if (!h) { kst_set_error("out of host memory"); return KST_ENOMEM; }
    h->device = device_id;
    h->model_id = model_id == KST_MODEL_LINLOGREG ? KST_FID_LINLOGREG + d : model_id;
    h->crit = criterion_id; h->trafo = trafo_id;
    h->r = r; h->t = t; h->m = m; h->d = d; h->N = N;
    h->ld = ((size_t)N + 31) / 32 * 32;
    for (int i = 0; i < 4; i++) h->mp.v[i] = i < n_model_params ? model_params[i] : 0.0;
    for (int i = 0; i < 16; i++) h->mp.cpv[i] = (cp_params && i < n_cp_params) ? cp_params[i] : 0.0;
    h->mp.m = m; h->mp.cp = cp_id; h->mp.d = d;
    for (int j = 0; j < d; j++) { h->lb[j] = lb[j]; h->ub[j] = ub[j]; }
    h->n_aux = trafo_id == KST_TRAFO_DELTA ? t * r : (criterion_id == KST_CRIT_A ? r : (criterion_id == KST_CRIT_DEXP ? 1 : 0));
    h->uniform_w = prior_w[0] > 0.0 && prior_w[0] < 1e300;
    for (int64_t j = 1; j < N && h->uniform_w; j++) h->uniform_w = prior_w[j] == prior_w[0];
    h->w_uniform = prior_w[0];

    int rc = KST_OK;
    double *d_theta_aos = nullptr, *d_dt = nullptr, *d_scratch = nullptr;
#define KST_TRY(call)                        \
    do {                                     \
        cudaError_t e__ = (call);            \
        if (e__ != cudaSuccess) {            \
            kst_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__)); \
            rc = KST_ECUDA;                  \
            goto fail;                       \
        }                                    \
    } while (0)
    {
        cudaDeviceProp prop;
        KST_TRY(cudaGetDeviceProperties(&prop, device_id));
        h->sm_count = prop.multiProcessorCount;
    }
    KST_TRY(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
    KST_TRY(cudaMalloc(&h->theta, sizeof(double) * r * h->ld));
    KST_TRY(cudaMalloc(&h->pw, sizeof(double) * h->ld));
    KST_TRY(cudaMalloc(&h->dc, sizeof(double) * ndc * h->ld));
    KST_TRY(cudaMalloc(&h->cs, sizeof(double) * r * h->ld));
    if (h->n_aux) KST_TRY(cudaMalloc(&h->aux, sizeof(double) * h->n_aux * h->ld));
    KST_TRY(cudaMalloc(&h->d_flag, 16));  // flag word at offset 0, one double of scratch at offset 8 (kst_sum_fixed)
    KST_TRY(cudaMalloc(&d_theta_aos, sizeof(double) * N * r));
    KST_TRY(cudaMemsetAsync(h->pw, 0, sizeof(double) * h->ld, h->stream));
    KST_TRY(cudaMemcpyAsync(d_theta_aos, theta, sizeof(double) * N * r, cudaMemcpyHostToDevice, h->stream));
    KST_TRY(cudaMemcpyAsync(h->pw, prior_w, sizeof(double) * N, cudaMemcpyHostToDevice, h->stream));
    if (trafo_id == KST_TRAFO_DELTA) {
        KST_TRY(cudaMalloc(&d_dt, sizeof(double) * N * t * r));
        KST_TRY(cudaMemcpyAsync(d_dt, dt_jm, sizeof(double) * N * t * r, cudaMemcpyHostToDevice, h->stream));
    }
    KST_TRY(cudaMalloc(&d_scratch, sizeof(double) * (N + 8)));
    rc = kst_launch_prep_draws(h, d_theta_aos, d_dt, d_scratch);
    if (rc) goto fail;
    KST_TRY(cudaStreamSynchronize(h->stream));
    cudaFree(d_theta_aos);
    cudaFree(d_scratch);
    if (d_dt) cudaFree(d_dt);
    *out = h;
    return KST_OK;
fail:
    if (d_theta_aos) cudaFree(d_theta_aos);
    if (d_scratch) cudaFree(d_scratch);
    if (d_dt) cudaFree(d_dt);
    kst_problem_destroy(h);
    return rc;
#undef KST_TRY
}

int kst_ensure_partials(kst_problem *h, size_t n)
{
    if (n <= h->partials_cap) return KST_OK;
    if (h->partials) KST_CUDA(cudaFree(h->partials));
    h->partials = nullptr;
    h->partials_cap = 0;
    KST_CUDA(cudaMalloc(&h->partials, sizeof(double) * n));
    h->partials_cap = n;
    return KST_OK;
}

int kst_ensure_stage(kst_problem *h, size_t particles, size_t rows)
{
    if (particles > h->stage_cap_particles) {
        if (h->stage_fx) KST_CUDA(cudaFree(h->stage_fx));
        h->stage_fx = nullptr;
        h->stage_cap_particles = 0;  // stays 0 if the allocation below fails
        KST_CUDA(cudaMalloc(&h->stage_fx, sizeof(double) * particles));
        h->stage_cap_particles = particles;
    }
    if (rows > h->stage_cap_rows) {
        if (h->stage_pts) KST_CUDA(cudaFree(h->stage_pts));
        if (h->stage_w) KST_CUDA(cudaFree(h->stage_w));
        h->stage_pts = h->stage_w = nullptr;
        h->stage_cap_rows = 0;
        KST_CUDA(cudaMalloc(&h->stage_pts, sizeof(double) * rows * h->d));
        KST_CUDA(cudaMalloc(&h->stage_w, sizeof(double) * rows));
        h->stage_cap_rows = rows;
    }
    return KST_OK;
}

#define KST_CHECK_HANDLE(h)                                  \
    do {                                                     \
        if (!(h)) { kst_set_error("NULL problem handle"); return KST_EINVAL; } \
        KST_CUDA(cudaSetDevice((h)->device));                \
    } while (0)

extern "C" void *kst_stream(kst_problem *h) { return h ? (void *)h->stream : nullptr; }

extern "C" int kst_stream_synchronize(kst_problem *h)
{
    KST_CHECK_HANDLE(h);
    KST_CUDA(cudaStreamSynchronize(h->stream));
    return KST_OK;
}

extern "C" int kst_objective_batch_device(kst_problem *h, int S, int K, const double *d_points, const double *d_weights,
                                          double *d_fx)
{
    KST_CHECK_HANDLE(h);
    if (!d_points || !d_weights || !d_fx) { kst_set_error("objective_batch_device: NULL buffer"); return KST_EINVAL; }
    return kst_objective_allreduce(h, S, K, d_points, d_weights, d_fx);
}

extern "C" int kst_objective_batch(kst_problem *h, int S, int K, const double *points, const double *weights, double *fx)
{
    KST_CHECK_HANDLE(h);
    if (!points || !weights || !fx) { kst_set_error("objective_batch: NULL buffer"); return KST_EINVAL; }
    if (S < 1 || K < 1) { kst_set_error("objective_batch: need S >= 1 and K >= 1"); return KST_EINVAL; }
    int rc = kst_ensure_stage(h, (size_t)S, (size_t)S * K);
    if (rc) return rc;
    KST_CUDA(cudaMemcpyAsync(h->stage_pts, points, sizeof(double) * S * K * h->d, cudaMemcpyHostToDevice, h->stream));
    KST_CUDA(cudaMemcpyAsync(h->stage_w, weights, sizeof(double) * S * K, cudaMemcpyHostToDevice, h->stream));
    rc = kst_objective_batch_device(h, S, K, h->stage_pts, h->stage_w, h->stage_fx);
    if (rc) return rc;
    KST_CUDA(cudaMemcpyAsync(fx, h->stage_fx, sizeof(double) * S, cudaMemcpyDeviceToHost, h->stream));
    KST_CUDA(cudaStreamSynchronize(h->stream));
    return KST_OK;
}

// Rt_j = D_j^-1 R D_j^-1 (packed upper) for every draw of the handle
__global__ void k_prep_reg(const double *__restrict__ cs, const double *__restrict__ Rm, int r, int64_t N, size_t ld,
                           double *__restrict__ reg)
{
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= N) return;
    int idx = 0;
    for (int a = 0; a < r; a++)
        for (int b = a; b < r; b++, idx++)
            reg[(size_t)idx * ld + j] = Rm[a + b * r] / (cs[(size_t)a * ld + j] * cs[(size_t)b * ld + j]);
}

extern "C" int kst_problem_set_regularization(kst_problem *h, const double *R)
{
    KST_CHECK_HANDLE(h);
    KST_CUDA(cudaStreamSynchronize(h->stream));
    h->g_ready = 0;
    if (!R) {
        if (h->reg) cudaFree(h->reg);
        h->reg = nullptr;
        return KST_OK;
    }
    const int r = h->r;
    for (int i = 0; i < r * r; i++)
        if (!std::isfinite(R[i])) { kst_set_error("set_regularization: R must be finite"); return KST_EINVAL; }
    if (!h->reg) KST_CUDA(cudaMalloc(&h->reg, sizeof(double) * (size_t)(r * (r + 1) / 2) * h->ld));
    // R travels through the handle's staging buffer (r * r <= 25 doubles)
    int rcs = kst_ensure_stage(h, (size_t)(r * r), 1);
    if (rcs) return rcs;
    double *d_R = h->stage_fx;
    cudaError_t e = cudaMemcpyAsync(d_R, R, sizeof(double) * r * r, cudaMemcpyHostToDevice, h->stream);
    if (e == cudaSuccess) {
        k_prep_reg<<<(unsigned)((h->N + 255) / 256), 256, 0, h->stream>>>(h->cs, d_R, r, h->N, h->ld, h->reg);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
    if (e != cudaSuccess) {
        cudaFree(h->reg);
        h->reg = nullptr;
        kst_set_error("set_regularization: %s", cudaGetErrorString(e));
        return KST_ECUDA;
    }
    return KST_OK;
}

extern "C" int kst_informationmatrix(kst_problem *h, int K, const double *points, const double *weights,
                                     int64_t draw_index, double *M)
{
    KST_CHECK_HANDLE(h);
    if (!points || !weights || !M || K < 1) { kst_set_error("informationmatrix: bad arguments"); return KST_EINVAL; }
    int rc = kst_ensure_stage(h, (size_t)(h->r * h->r), (size_t)K);
    if (rc) return rc;
    KST_CUDA(cudaMemcpyAsync(h->stage_pts, points, sizeof(double) * K * h->d, cudaMemcpyHostToDevice, h->stream));
    KST_CUDA(cudaMemcpyAsync(h->stage_w, weights, sizeof(double) * K, cudaMemcpyHostToDevice, h->stream));
    rc = kst_launch_infomatrix(h, K, h->stage_pts, h->stage_w, draw_index, h->stage_fx);
    if (rc) return rc;
    KST_CUDA(cudaMemcpyAsync(M, h->stage_fx, sizeof(double) * h->r * h->r, cudaMemcpyDeviceToHost, h->stream));
    KST_CUDA(cudaStreamSynchronize(h->stream));
    return KST_OK;
}

// ---------------------------------------------------------------------------------------
// Gateaux derivative
// ---------------------------------------------------------------------------------------
extern "C" int kst_gateaux_prepare(kst_problem *h, int K, const double *points, const double *weights, int *singular)
{
    KST_CHECK_HANDLE(h);
    if (!points || !weights || K < 1) { kst_set_error("gateaux_prepare: bad arguments"); return KST_EINVAL; }
    int rc = kst_ensure_stage(h, 1, (size_t)K);
    if (rc) return rc;
    KST_CUDA(cudaMemcpyAsync(h->stage_pts, points, sizeof(double) * K * h->d, cudaMemcpyHostToDevice, h->stream));
    KST_CUDA(cudaMemcpyAsync(h->stage_w, weights, sizeof(double) * K, cudaMemcpyHostToDevice, h->stream));
    rc = kst_launch_gateaux_prepare(h, K, h->stage_pts, h->stage_w);
    if (rc) return rc;
    if (h->nranks > 1) {
        // every rank must agree on "singular": sum the per-shard flags (NCCL or the peer exchange, whichever is set up)
        double f = (double)h->g_singular, *d_f = h->stage_fx;
        KST_CUDA(cudaMemcpyAsync(d_f, &f, sizeof(double), cudaMemcpyHostToDevice, h->stream));
        rc = kst_allreduce_sum(h, d_f, 1);
        if (rc) return rc;
        KST_CUDA(cudaMemcpyAsync(&f, d_f, sizeof(double), cudaMemcpyDeviceToHost, h->stream));
        KST_CUDA(cudaStreamSynchronize(h->stream));
        h->g_singular = f != 0.0;
    }
    if (singular) *singular = h->g_singular;
    return KST_OK;
}

__global__ void k_unscale_gconst(const double *__restrict__ gA, const double *__restrict__ cs, size_t ld, int64_t N, int r,
                                 double *__restrict__ out)
{
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= N) return;
    for (int a = 0; a < r; a++)
        for (int b = 0; b < r; b++)
            out[j * r * r + a + b * r] = gA[(size_t)(a * r + b) * ld + j] / (cs[(size_t)a * ld + j] * cs[(size_t)b * ld + j]);
}

extern "C" int kst_gateaux_constants(kst_problem *h, double *A, double *tr_B)
{
    KST_CHECK_HANDLE(h);
    if (!h->g_ready) { kst_set_error("gateaux_constants before gateaux_prepare"); return KST_ESTATE; }
    if (A) {
        // the copy-out buffer belongs to the handle and is allocated on first use (debugging path of the reference)
        if (!h->g_out) KST_CUDA(cudaMalloc(&h->g_out, sizeof(double) * h->N * h->r * h->r));
        k_unscale_gconst<<<(unsigned)((h->N + 127) / 128), 128, 0, h->stream>>>(h->gA, h->cs, h->ld, h->N, h->r, h->g_out);
        KST_CUDA(cudaGetLastError());
        KST_CUDA(cudaMemcpyAsync(A, h->g_out, sizeof(double) * h->N * h->r * h->r, cudaMemcpyDeviceToHost, h->stream));
        KST_CUDA(cudaStreamSynchronize(h->stream));
    }
    if (tr_B) {
        KST_CUDA(cudaMemcpyAsync(tr_B, h->gtrB, sizeof(double) * h->N, cudaMemcpyDeviceToHost, h->stream));
        KST_CUDA(cudaStreamSynchronize(h->stream));
    }
    return KST_OK;
}

extern "C" int kst_gateaux_eval(kst_problem *h, int64_t G, const double *dirs, double *gd)
{
    KST_CHECK_HANDLE(h);
    if (!dirs || !gd || G < 1) { kst_set_error("gateaux_eval: bad arguments"); return KST_EINVAL; }
    if ((size_t)G > h->gd_cap) {
        if (h->gd_dirs) KST_CUDA(cudaFree(h->gd_dirs));
        if (h->gd_out) KST_CUDA(cudaFree(h->gd_out));
        h->gd_dirs = h->gd_out = nullptr;
        h->gd_cap = 0;
        KST_CUDA(cudaMalloc(&h->gd_dirs, sizeof(double) * G * h->d));
        KST_CUDA(cudaMalloc(&h->gd_out, sizeof(double) * G));
        h->gd_cap = (size_t)G;
    }
    KST_CUDA(cudaMemcpyAsync(h->gd_dirs, dirs, sizeof(double) * G * h->d, cudaMemcpyHostToDevice, h->stream));
    int rc = kst_launch_gateaux_eval(h, G, h->gd_dirs, h->gd_out);
    if (rc) return rc;
    rc = kst_allreduce_sum(h, h->gd_out, (size_t)G);
    if (rc) return rc;
    KST_CUDA(cudaMemcpyAsync(gd, h->gd_out, sizeof(double) * G, cudaMemcpyDeviceToHost, h->stream));
    KST_CUDA(cudaStreamSynchronize(h->stream));
    return KST_OK;
}

// ---------------------------------------------------------------------------------------
// PSO driver
// ---------------------------------------------------------------------------------------
static int pso_evaluate(kst_problem *h)
{
    KstPso &p = h->pso;
    const int gd = p.mode == KST_PSO_MODE_GATEAUX;
    int rc;
    if (h->nranks <= 1) {
        // one GPU: the finish step rides along in k_pso_best, which follows every evaluation of the swarm
        rc = kst_launch_objective(h, p.S, p.K, p.x_pts, p.x_w, p.fx, &p.pending_n_blk, gd);
        kst_objective_finish_terms(h, gd, &p.pending_offset, &p.pending_nan);
    } else if (!gd) {
        rc = kst_objective_allreduce(h, p.S, p.K, p.x_pts, p.x_w, p.fx);
    } else {
        // Gateaux mode: the swarm's one-point directions go through the evaluation kernel with the Gateaux integrand
        // as the per-draw value (lanes along draws), not through the grid kernel (one thread per direction), which
        // is laid out for thousands of directions
        rc = kst_launch_objective(h, p.S, p.K, p.x_pts, p.x_w, p.fx, nullptr, 1);
        if (rc) return rc;
        rc = kst_allreduce_sum(h, p.fx, (size_t)p.S);
    }
    p.n_eval += p.S;
    return rc;
}

// Pso constructor checks (src/pso.jl:35-43), prototype checks (src/pso.jl:89-91, check_compatible
// src/designmeasure-abstractpoint.jl:60-71), swarm allocation, upload, k_pso_init: everything of optimizer_state
// (src/pso.jl:83-125) up to the first evaluation.
static int pso_setup(kst_problem *h, int S, int K, double c1, double c2, const double *proto_points,
                     const double *proto_weights, int P, const int *fixw, const int *fixp, uint64_t seed, int mode)
{
    if (S <= 1) { kst_set_error("swarmsize must be at least 2"); return KST_EINVAL; }
    if (c1 < 0 || c2 < 0 || c1 + c2 <= 4) { kst_set_error("c1 and c2 must be non-negative and, c1 + c2 > 4"); return KST_EINVAL; }
    if (P < 1 || S < P) { kst_set_error("swarmsize must be >= number of prototype particles"); return KST_EINVAL; }
    if (K < 1 || K * h->m > KST_MAX_ROWS || !proto_points || !proto_weights || !fixw || !fixp) { kst_set_error("pso_init: bad arguments"); return KST_EINVAL; }
    if (mode == KST_PSO_MODE_GATEAUX) {
        if (K != 1) { kst_set_error("derivative only implemented for one-point directions"); return KST_EINVAL; }
        if (!h->g_ready) { kst_set_error("pso_init(gateaux) before gateaux_prepare"); return KST_ESTATE; }
    } else if (mode != KST_PSO_MODE_OBJECTIVE) { kst_set_error("unknown pso mode %d", mode); return KST_EINVAL; }
    for (int i = 0; i < P * K; i++)
        for (int j = 0; j < h->d; j++) {
            const double x = proto_points[(size_t)i * h->d + j];
            if (!(x >= h->lb[j] && x <= h->ub[j])) { kst_set_error("design point is outside design region"); return KST_EINVAL; }
        }
    int rc = kst_pso_alloc(h, S, K);
    if (rc) return rc;
    KstPso &p = h->pso;
    p.mode = mode; p.c1 = c1; p.c2 = c2; p.seed = seed; p.iter = 0; p.n_eval = 0; p.pending_n_blk = 0;
    // "fixing all weights but one is equivalent to fixing them all" (src/designmeasure-abstractpoint.jl:51-56)
    int fw[KST_MAX_ROWS], nfix = 0;
    for (int k = 0; k < K; k++) { fw[k] = fixw[k] != 0; nfix += fw[k]; }
    if (nfix == K - 1)
        for (int k = 0; k < K; k++) fw[k] = 1;
    // (a copy from pageable host memory is staged before cudaMemcpyAsync returns: the stack buffer may go away)
    KST_CUDA(cudaMemcpyAsync(p.fixw, fw, sizeof(int) * K, cudaMemcpyHostToDevice, h->stream));
    KST_CUDA(cudaMemcpyAsync(p.fixp, fixp, sizeof(int) * K, cudaMemcpyHostToDevice, h->stream));
    rc = kst_ensure_stage(h, 1, (size_t)P * K);
    if (rc) return rc;
    KST_CUDA(cudaMemcpyAsync(h->stage_pts, proto_points, sizeof(double) * P * K * h->d, cudaMemcpyHostToDevice, h->stream));
    KST_CUDA(cudaMemcpyAsync(h->stage_w, proto_weights, sizeof(double) * P * K, cudaMemcpyHostToDevice, h->stream));
    return kst_launch_pso_init(h, h->stage_pts, h->stage_w, P);
}

extern "C" int kst_pso_init(kst_problem *h, int S, int K, double c1, double c2, const double *proto_points,
                            const double *proto_weights, int P, const int *fixw, const int *fixp, uint64_t seed, int mode,
                            int *status)
{
    KST_CHECK_HANDLE(h);
    int rc = pso_setup(h, S, K, c1, c2, proto_points, proto_weights, P, fixw, fixp, seed, mode);
    if (rc) return rc;
    KstPso &p = h->pso;
    rc = kst_pso_ensure_trace(h, 1);
    if (rc) return rc;
    rc = pso_evaluate(h);
    if (rc) return rc;
    rc = kst_launch_pso_best(h, 1, -1);
    if (rc) return rc;
    int st = 0;
    KST_CUDA(cudaMemcpyAsync(&st, p.d_status, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    KST_CUDA(cudaStreamSynchronize(h->stream));
    p.iter = 1;
    if (status) *status = st;
    return KST_OK;
}

// after a synchronisation: did a persistent swarm kernel give up at its grid barrier?
static int pso_check_barrier(kst_problem *h)
{
    KstPso &p = h->pso;
    if (!p.bar || p.bar_count == 0) return KST_OK;
    unsigned long long err = 0;
    KST_CUDA(cudaMemcpy(&err, p.bar + 1, sizeof(err), cudaMemcpyDeviceToHost));
    if (err) {
        kst_set_error("persistent swarm kernel: grid barrier timed out (results invalid); set KST_NO_FUSED=1");
        return KST_ECUDA;
    }
    return KST_OK;
}

// n_iters times tick! on the three-launch path (move, evaluation, best update), trace slots slot0, slot0 + 1, ...
static int pso_ticks_unfused(kst_problem *h, int n_iters, int slot0)
{
    KstPso &p = h->pso;
    const int gd = p.mode == KST_PSO_MODE_GATEAUX;
    const bool tail = h->nranks > 1 && h->peer_ready && p.S <= KST_PEER_CAP;
    for (int it = 0; it < n_iters; it++) {
        int rc = kst_launch_pso_move(h);
        if (rc) return rc;
        if (tail) {
            // several GPUs with a peer exchange: evaluation + ONE launch that finishes the partial sums, sums them over
            // the ranks and updates the bests
            int n_blk = 0, nan_fill = 0;
            double offset = 0.0;
            rc = kst_launch_objective(h, p.S, p.K, p.x_pts, p.x_w, p.fx, &n_blk, gd);
            if (rc) return rc;
            p.n_eval += p.S;
            kst_objective_finish_terms(h, gd, &offset, &nan_fill);
            rc = kst_launch_pso_tail(h, slot0 + it, n_blk, offset, nan_fill);
            if (rc) return rc;
        } else {
            rc = pso_evaluate(h);
            if (rc) return rc;
            rc = kst_launch_pso_best(h, 0, slot0 + it);
            if (rc) return rc;
        }
        p.iter++;
    }
    return KST_OK;
}

extern "C" int kst_pso_run(kst_problem *h, int n_iters, double *trace_fx, double *trace_points, double *trace_weights)
{
    KST_CHECK_HANDLE(h);
    KstPso &p = h->pso;
    if (p.iter < 1) { kst_set_error("pso_run before pso_init"); return KST_ESTATE; }
    if (n_iters < 0) { kst_set_error("pso_run: n_iters < 0"); return KST_EINVAL; }
    if (n_iters == 0) return KST_OK;
    int rc = kst_pso_ensure_trace(h, n_iters);
    if (rc) return rc;
    // small problems: the whole run in one cooperative launch (kst_pso_fused.cuh); otherwise three launches per tick
    rc = kst_launch_pso_fused(h, n_iters, 0, 0, 1, nullptr, nullptr);
    if (rc == 1) rc = pso_ticks_unfused(h, n_iters, 0);
    else if (rc == KST_OK) { p.iter += n_iters; p.n_eval += (int64_t)p.S * n_iters; }
    if (rc) return rc;
    // without trace buffers nothing has to reach the host: the iterations stay queued on the handle's stream
    if (trace_fx || trace_points || trace_weights) {
        const double *fx, *pts, *w;
        rc = kst_pso_fetch(h, 0, nullptr, &fx, &pts, &w, nullptr);
        if (rc) return rc;
        if (trace_fx) memcpy(trace_fx, fx, sizeof(double) * n_iters);
        if (trace_points) memcpy(trace_points, pts, sizeof(double) * n_iters * p.K * h->d);
        if (trace_weights) memcpy(trace_weights, w, sizeof(double) * n_iters * p.K);
        return pso_check_barrier(h);
    }
    return KST_OK;
}

// Enqueues a whole optimize() (src/optimize.jl:42-70) -- optimizer_state, then iterations - 1 ticks -- without a host
// synchronisation.  Afterwards tr_fx / tr_pts / tr_w hold the `iterations` trace entries (entry 0 = initial state),
// h->d_flag the status word and, when want_snapshot, p.snap the packed state after the initial evaluation.
int kst_pso_optimize_enqueue(kst_problem *h, int S, int K, int iterations, double c1, double c2, const double *proto_points,
                             const double *proto_weights, int P, const int *fixw, const int *fixp, uint64_t seed, int mode,
                             int want_snapshot)
{
    if (iterations < 1) { kst_set_error("number of iterations must be positive"); return KST_EINVAL; }
    int rc = pso_setup(h, S, K, c1, c2, proto_points, proto_weights, P, fixw, fixp, seed, mode);
    if (rc) return rc;
    KstPso &p = h->pso;
    rc = kst_pso_ensure_trace(h, iterations);
    if (rc) return rc;
    p.iter = 1;  // the first move is iteration 1 (key of its random numbers), as after kst_pso_init
    rc = kst_launch_pso_fused(h, iterations, 1, 0, 1, p.d_status, want_snapshot ? p.snap : nullptr);
    if (rc == KST_OK) {
        p.iter = iterations;
        p.n_eval = (int64_t)S * iterations;
        return KST_OK;
    }
    if (rc != 1) return rc;
    p.iter = 0;
    rc = pso_evaluate(h);
    if (rc) return rc;
    rc = kst_launch_pso_best(h, 1, 0);
    if (rc) return rc;
    p.iter = 1;
    if (want_snapshot) {
        // the same packed layout the persistent kernel writes (KstFusedParams::snap), gathered with device copies
        const size_t np = (size_t)S * K * h->d, nw = (size_t)S * K;
        struct { const double *src; size_t n; } parts[] = {{p.x_pts, np}, {p.x_w, nw}, {p.p_pts, np}, {p.p_w, nw}, {p.v_pts, np},
                                                           {p.v_w, nw}, {p.g_pts, (size_t)K * h->d}, {p.g_w, (size_t)K},
                                                           {p.fx, (size_t)S}, {p.fp, (size_t)S}, {p.fg, 1}};
        double *dst = p.snap;
        for (auto &q : parts) {
            KST_CUDA(cudaMemcpyAsync(dst, q.src, sizeof(double) * q.n, cudaMemcpyDeviceToDevice, h->stream));
            dst += q.n;
        }
    }
    return pso_ticks_unfused(h, iterations - 1, 1);
}

extern "C" int kst_pso_optimize(kst_problem *h, int S, int K, int iterations, double c1, double c2,
                                const double *proto_points, const double *proto_weights, int P, const int *fixw,
                                const int *fixp, uint64_t seed, int mode, int *status, double *trace_fx,
                                double *trace_points, double *trace_weights, double *state0)
{
    KST_CHECK_HANDLE(h);
    int rc = kst_pso_optimize_enqueue(h, S, K, iterations, c1, c2, proto_points, proto_weights, P, fixw, fixp, seed, mode,
                                      state0 != nullptr);
    if (rc) return rc;
    int st = 0;
    const double *fx, *pts, *w, *snap;
    rc = kst_pso_fetch(h, state0 != nullptr, &st, &fx, &pts, &w, &snap);
    if (rc) return rc;
    if (trace_fx) memcpy(trace_fx, fx, sizeof(double) * iterations);
    if (trace_points) memcpy(trace_points, pts, sizeof(double) * iterations * K * h->d);
    if (trace_weights) memcpy(trace_weights, w, sizeof(double) * iterations * K);
    if (state0) memcpy(state0, snap, sizeof(double) * (3 * (size_t)S * K * (h->d + 1) + (size_t)K * (h->d + 1) + 2 * (size_t)S + 1));
    if (status) *status = st;
    return pso_check_barrier(h);
}

extern "C" int kst_pso_get_state(kst_problem *h, double *x_points, double *x_weights, double *p_points, double *p_weights,
                                 double *g_points, double *g_weights, double *v_points, double *v_weights, double *fx,
                                 double *fp, double *fg, int64_t *n_eval)
{
    KST_CHECK_HANDLE(h);
    KstPso &p = h->pso;
    if (p.iter < 1) { kst_set_error("pso_get_state before pso_init"); return KST_ESTATE; }
    const size_t np = (size_t)p.S * p.K * h->d, nw = (size_t)p.S * p.K;
    struct { double *dst; const double *src; size_t n; } cp[] = {
        {x_points, p.x_pts, np}, {x_weights, p.x_w, nw}, {p_points, p.p_pts, np}, {p_weights, p.p_w, nw},
        {g_points, p.g_pts, (size_t)p.K * h->d}, {g_weights, p.g_w, (size_t)p.K}, {v_points, p.v_pts, np},
        {v_weights, p.v_w, nw}, {fx, p.fx, (size_t)p.S}, {fp, p.fp, (size_t)p.S}, {fg, p.fg, 1}};
    for (auto &c : cp)
        if (c.dst) KST_CUDA(cudaMemcpyAsync(c.dst, c.src, sizeof(double) * c.n, cudaMemcpyDeviceToHost, h->stream));
    KST_CUDA(cudaStreamSynchronize(h->stream));
    if (n_eval) *n_eval = p.n_eval;
    return pso_check_barrier(h);
}

// ---------------------------------------------------------------------------------------
// measurement helpers
// ---------------------------------------------------------------------------------------
extern "C" int kst_profile_enable(kst_problem *h, int on)
{
    KST_CHECK_HANDLE(h);
    if (on < 0 || on > 3) { kst_set_error("profile_enable: stage must be 0 (off), 1 (evaluation), 2 (move) or 3 (best / tail)"); return KST_EINVAL; }
    h->prof_on = on;
    return KST_OK;
}

extern "C" int kst_profile_read(kst_problem *h, double *total_ms, int64_t *launches)
{
    KST_CHECK_HANDLE(h);
    KST_CUDA(cudaStreamSynchronize(h->stream));
    double tot = 0.0;
    for (size_t i = 0; i + 1 < h->prof_events.size(); i += 2) {
        float ms = 0.f;
        KST_CUDA(cudaEventElapsedTime(&ms, h->prof_events[i], h->prof_events[i + 1]));
        tot += ms;
    }
    if (total_ms) *total_ms = tot;
    if (launches) *launches = (int64_t)(h->prof_events.size() / 2);
    for (cudaEvent_t e : h->prof_events) cudaEventDestroy(e);
    h->prof_events.clear();
    return KST_OK;
}

// DFMA throughput probe: 8 independent FMA chains per thread, enough CTAs to fill every SM.
__global__ void __launch_bounds__(256) k_dfma_probe(double *out, int iters, double a, double b)
{
    double x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int u = 0; u < 16; u++) {
            x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
            x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
        }
    }
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
}

extern "C" int kst_fp64_peak_probe(int device_id, double *tflops)
{
    KST_CUDA(cudaSetDevice(device_id));
    cudaDeviceProp prop;
    KST_CUDA(cudaGetDeviceProperties(&prop, device_id));
    const int blocks = prop.multiProcessorCount * 8, nt = 256, iters = 2048;
    double *d_out = nullptr;
    KST_CUDA(cudaMalloc(&d_out, sizeof(double) * blocks * nt));
    cudaEvent_t e0, e1;
    KST_CUDA(cudaEventCreate(&e0));
    KST_CUDA(cudaEventCreate(&e1));
    double best = 0.0;
    for (int rep = 0; rep < 5; rep++) {
        KST_CUDA(cudaEventRecord(e0, 0));
        k_dfma_probe<<<blocks, nt>>>(d_out, iters, 0.999999, 1e-9);
        KST_CUDA(cudaEventRecord(e1, 0));
        KST_CUDA(cudaEventSynchronize(e1));
        float ms = 0.f;
        KST_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        const double flops = 2.0 * 8 * 16 * (double)iters * nt * blocks;
        const double tf = flops / (ms * 1e-3) / 1e12;
        if (rep > 0 && tf > best) best = tf;
    }
    KST_CUDA(cudaGetLastError());
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(d_out);
    if (tflops) *tflops = best;
    return KST_OK;
}

__global__ void k_math_selftest(int64_t n, const double *__restrict__ x, double *__restrict__ e, double *__restrict__ r,
                                double *__restrict__ l)
{
    __shared__ __align__(16) double s_mathtab[KST_MATH_SMEM];
    const KstTab tab = kst_math_init(s_mathtab);
    __syncthreads();
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    e[i] = kst_exp(x[i], tab);
    r[i] = kst_rcp(x[i]);
    l[i] = kst_log(fabs(x[i]), tab);
}

extern "C" int kst_math_selftest(int device_id, int64_t n, const double *x, double *exp_out, double *rcp_out,
                                 double *log_out)
{
    if (n < 1 || !x || !exp_out || !rcp_out || !log_out) { kst_set_error("math_selftest: bad arguments"); return KST_EINVAL; }
    KST_CUDA(cudaSetDevice(device_id));
    double *d = nullptr;
    KST_CUDA(cudaMalloc(&d, sizeof(double) * 4 * n));
    KST_CUDA(cudaMemcpy(d, x, sizeof(double) * n, cudaMemcpyHostToDevice));
    k_math_selftest<<<(unsigned)((n + 255) / 256), 256>>>(n, d, d + n, d + 2 * n, d + 3 * n);
    KST_CUDA(cudaGetLastError());
    KST_CUDA(cudaMemcpy(exp_out, d + n, sizeof(double) * n, cudaMemcpyDeviceToHost));
    KST_CUDA(cudaMemcpy(rcp_out, d + 2 * n, sizeof(double) * n, cudaMemcpyDeviceToHost));
    KST_CUDA(cudaMemcpy(log_out, d + 3 * n, sizeof(double) * n, cudaMemcpyDeviceToHost));
    KST_CUDA(cudaFree(d));
    return KST_OK;
}
