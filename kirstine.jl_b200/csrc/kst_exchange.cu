// The Exchange strategy behind one C-ABI call: solve_with(dp, ::Exchange, trace_state) of the reference
// (src/solve-exchange.jl:116-179) with both optimizers running on the device and only what the host logic needs
// coming back per step -- the singularity flag of the candidate, the direction found (d doubles) and the re-optimised
// weights -- plus the traces at the end of each optimizer run.  Per step:
//
//   candidate = simplify(candidate; simplify_args)            host (K <= a few dozen points): simplify_unique of the
//                                                             registered covariate parameterizations, simplify_merge
//                                                             (src/designmeasure.jl:443-469), simplify_drop (:372-381)
//   gateaux_constants(candidate)                              k_gateaux_prepare, constants stay on the device
//   optimize(gateauxderivative!, optimizer_direction, one-point prototypes at the candidate's points)
//                                                             kst_pso_optimize_enqueue, Gateaux mode
//   prepend the direction (mixture, src/designmeasure.jl:312-322; simplify_merge with distance 0 when the point is
//   already in the support)                                   host
//   optimize(objective!, optimizer_weight, candidate with all points fixed)
//                                                             kst_pso_optimize_enqueue, objective mode
//
// Small problems (the usual case: swarms of tens of particles against 10^3 draws) run every optimize as ONE persistent
// kernel launch (kst_pso_fused.cuh), so a step costs three short host round trips instead of ~60 kernel launches and
// two whole-swarm copies.
#include <math.h>
#include <string.h>

#include <vector>

#include "kst_common.cuh"

namespace {

struct Design {
    int K = 0, d = 1;
    std::vector<double> pts;  // [K][d]
    std::vector<double> w;    // [K]
};

// simplify_unique for the registered covariate parameterizations (the generic method is a no-op,
// src/designmeasure.jl:412-424; CopyBoth docs/src/dtr.md:229-247; EquidistantTimes :498-523, whose canonical
// representative LogEquidistantTimes shares).  ua, ub: the method's two keyword arguments (minposdose and
// minpostime / minposdelta), 0 when not given.
void simplify_unique(const kst_problem *h, Design &c, double ua, double ub)
{
    if (h->model_id != KST_MODEL_DTR) return;
    const int cp = h->mp.cp;
    if (cp == KST_CP_DTR_COPYBOTH) {
        if (h->lb[0] == 0.0 && h->lb[1] == 0.0)
            for (int k = 0; k < c.K; k++)
                if (c.pts[k * 2] < ua || c.pts[k * 2 + 1] < ub) c.pts[k * 2] = c.pts[k * 2 + 1] = 0.0;
    } else if (cp == KST_CP_DTR_EQUIDISTANT || cp == KST_CP_DTR_LOGEQUIDISTANT) {
        if (h->lb[0] == 0.0)
            for (int k = 0; k < c.K; k++) {
                bool drop = c.pts[k * 2] <= ua;
                if (h->lb[1] == 0.0) drop = drop || c.pts[k * 2 + 1] <= ub;
                if (drop) { c.pts[k * 2] = 0.0; c.pts[k * 2 + 1] = h->ub[1]; }
            }
    }
}

// simplify_merge (src/designmeasure.jl:443-469): repeatedly average the two points closest to each other (distances
// relative to the bounding box) with their weights until all are more than maxdist apart; the merged point moves
// to the front.  findmin scans column-major, so among equal distances the smallest column, then the smallest row wins.
void simplify_merge(const kst_problem *h, Design &c, double maxdist)
{
    if (c.K == 1) return;
    const int d = c.d;
    double cur = 0.0;
    while (cur <= maxdist) {
        const int n = c.K;
        if (n < 2) break;
        double best = INFINITY;
        int bi = -1, bj = -1;
        for (int j = 0; j < n; j++)
            for (int i = 0; i < n; i++) {
                if (i == j) continue;
                double s = 0.0;
                if (d == 1) {
                    s = fabs((c.pts[i] - c.pts[j]) / (h->lb[0] - h->ub[0]));
                } else {
                    for (int a = 0; a < d; a++) {
                        const double t = (c.pts[i * d + a] - c.pts[j * d + a]) / (h->lb[a] - h->ub[a]);
                        s += t * t;
                    }
                    s = sqrt(s);
                }
                if (s < best) { best = s; bi = i; bj = j; }
            }
        cur = best;
        if (!(cur <= maxdist)) break;
        Design o;
        o.d = d;
        o.K = n - 1;
        const double w_new = c.w[bi] + c.w[bj];
        for (int a = 0; a < d; a++) o.pts.push_back((c.w[bi] * c.pts[bi * d + a] + c.w[bj] * c.pts[bj * d + a]) / w_new);
        o.w.push_back(w_new);
        for (int k = 0; k < n; k++) {
            if (k == bi || k == bj) continue;
            for (int a = 0; a < d; a++) o.pts.push_back(c.pts[k * d + a]);
            o.w.push_back(c.w[k]);
        }
        c = o;
    }
}

// simplify_drop (src/designmeasure.jl:372-381)
void simplify_drop(Design &c, double maxweight)
{
    if (c.K == 1) return;
    Design o;
    o.d = c.d;
    double sum = 0.0;
    for (int k = 0; k < c.K; k++)
        if (c.w[k] > maxweight) {
            for (int a = 0; a < c.d; a++) o.pts.push_back(c.pts[k * c.d + a]);
            o.w.push_back(c.w[k]);
            sum += c.w[k];
        }
    o.K = (int)o.w.size();
    for (double &x : o.w) x /= sum;
    c = o;
}

// mixture(alpha, one-point design at z, c) (src/designmeasure.jl:312-322): z goes to the front
void prepend(Design &c, const double *z, double alpha)
{
    Design o;
    o.d = c.d;
    o.K = c.K + 1;
    for (int a = 0; a < c.d; a++) o.pts.push_back(z[a]);
    o.w.push_back(alpha * 1.0);
    o.pts.insert(o.pts.end(), c.pts.begin(), c.pts.end());
    for (int k = 0; k < c.K; k++) o.w.push_back((1 - alpha) * c.w[k]);
    c = o;
}

}  // namespace

extern "C" int kst_exchange_run(kst_problem *h, const kst_exchange_options *opt, int K0, const double *cand_points,
                                const double *cand_weights, int Kmax, int *K_out, double *points_out, double *weights_out,
                                double *dir_trace_fx, double *dir_trace_x, double *w_trace_fx, double *w_trace_points,
                                double *w_trace_weights, int *w_K, double *dir_state0, double *w_state0, int *steps_done)
{
    if (!h) { kst_set_error("NULL problem handle"); return KST_EINVAL; }
    KST_CUDA(cudaSetDevice(h->device));
    if (steps_done) *steps_done = 0;
    if (!opt || !cand_points || !cand_weights || !K_out || !points_out || !weights_out) { kst_set_error("exchange_run: NULL argument"); return KST_EINVAL; }
    if (opt->steps < 1) { kst_set_error("steps must be >= 1"); return KST_EINVAL; }
    if (K0 < 1 || Kmax < K0 + 1 || Kmax * h->m > KST_MAX_ROWS) { kst_set_error("exchange_run: need 1 <= K0 < Kmax and Kmax * m <= %d", KST_MAX_ROWS); return KST_EINVAL; }
    if (opt->iters_direction < 1 || opt->iters_weight < 1) { kst_set_error("number of iterations must be positive"); return KST_EINVAL; }
    const int d = h->d, itd = opt->iters_direction, itw = opt->iters_weight, Sd = opt->swarm_direction, Sw = opt->swarm_weight;
    Design c;
    c.d = d;
    c.K = K0;
    c.pts.assign(cand_points, cand_points + (size_t)K0 * d);
    c.w.assign(cand_weights, cand_weights + K0);
    // check_compatible (src/designmeasure-abstractpoint.jl:60-71)
    for (int k = 0; k < K0; k++)
        for (int a = 0; a < d; a++)
            if (!(c.pts[k * d + a] >= h->lb[a] && c.pts[k * d + a] <= h->ub[a])) { kst_set_error("design point is outside design region"); return KST_EINVAL; }
    uint64_t seed = opt->seed;
    auto next_seed = [&seed]() {
        const uint64_t s = seed;
        seed = s * 6364136223846793005ull + 1442695040888963407ull;
        return s;
    };
    std::vector<int> zeros(KST_MAX_ROWS, 0), ones(KST_MAX_ROWS, 1);
    std::vector<double> protos, proto_w;
    const double *tr_fx, *tr_x, *tr_w, *snap;  // into the handle's pinned mirror of the last run's output block
    const size_t snap_d = 3 * (size_t)Sd * (d + 1) + (size_t)(d + 1) + 2 * (size_t)Sd + 1;
    const size_t snap_w = 3 * (size_t)Sw * Kmax * (d + 1) + (size_t)Kmax * (d + 1) + 2 * (size_t)Sw + 1;
    for (int step = 0; step < opt->steps; step++) {
        // simplify the result of the previous iteration
        simplify_unique(h, c, opt->unique_a, opt->unique_b);
        simplify_merge(h, c, opt->maxdist);
        simplify_drop(c, opt->maxweight);
        if (c.K + 1 > Kmax) { kst_set_error("exchange_run: candidate grew beyond Kmax = %d points", Kmax); return KST_ELIMIT; }
        if (Sd < c.K) { kst_set_error("swarmsize must be >= number of prototype particles"); return KST_EINVAL; }
        // gateaux_constants at the candidate
        int rc = kst_ensure_stage(h, 1, (size_t)c.K);
        if (rc) return rc;
        KST_CUDA(cudaMemcpyAsync(h->stage_pts, c.pts.data(), sizeof(double) * c.K * d, cudaMemcpyHostToDevice, h->stream));
        KST_CUDA(cudaMemcpyAsync(h->stage_w, c.w.data(), sizeof(double) * c.K, cudaMemcpyHostToDevice, h->stream));
        rc = kst_launch_gateaux_prepare(h, c.K, h->stage_pts, h->stage_w);
        if (rc) return rc;
        if (h->g_singular) { kst_set_error("SingularException: the candidate's information matrix is singular (step %d)", step + 1); return KST_ESINGULAR; }
        // the direction in which the Gateaux derivative is largest: prototypes = one-point designs at the support
        protos.assign(c.pts.begin(), c.pts.end());
        proto_w.assign(c.K, 1.0);
        rc = kst_pso_optimize_enqueue(h, Sd, 1, itd, opt->c1_direction, opt->c2_direction, protos.data(), proto_w.data(), c.K,
                                      zeros.data(), zeros.data(), next_seed(), KST_PSO_MODE_GATEAUX, dir_state0 != nullptr);
        if (rc) return rc;
        int st = 0;
        rc = kst_pso_fetch(h, dir_state0 != nullptr, &st, &tr_fx, &tr_x, &tr_w, &snap);
        if (rc) return rc;
        if (dir_state0) memcpy(dir_state0 + (size_t)step * snap_d, snap, sizeof(double) * snap_d);
        if (st == 1) { kst_set_error("infinite objective value at best initial point"); return KST_EOBJECTIVE; }
        if (st == 2) { kst_set_error("objective evaluates to NaN on all initial points"); return KST_EOBJECTIVE; }
        if (dir_trace_fx) memcpy(dir_trace_fx + (size_t)step * itd, tr_fx, sizeof(double) * itd);
        if (dir_trace_x) memcpy(dir_trace_x + (size_t)step * itd * d, tr_x, sizeof(double) * itd * d);
        double dir[KST_MAX_D];
        memcpy(dir, tr_x + (size_t)(itd - 1) * d, sizeof(double) * d);
        // append (prepend) this point to the candidate
        bool in_support = false;
        for (int k = 0; k < c.K && !in_support; k++) in_support = memcmp(dir, &c.pts[k * d], sizeof(double) * d) == 0;
        if (!in_support)  // memcmp is bitwise: also accept -0.0 == 0.0 like the reference's `in`
            for (int k = 0; k < c.K && !in_support; k++) {
                bool eq = true;
                for (int a = 0; a < d; a++) eq = eq && dir[a] == c.pts[k * d + a];
                in_support = eq;
            }
        if (in_support) {
            prepend(c, dir, 0.0);
            simplify_merge(h, c, 0.0);
        } else {
            prepend(c, dir, 1.0 / (1 + c.K));
        }
        // re-optimize the weights: DirectMaximization with every point fixed
        rc = kst_pso_optimize_enqueue(h, Sw, c.K, itw, opt->c1_weight, opt->c2_weight, c.pts.data(), c.w.data(), 1,
                                      zeros.data(), ones.data(), next_seed(), KST_PSO_MODE_OBJECTIVE, w_state0 != nullptr);
        if (rc) return rc;
        rc = kst_pso_fetch(h, w_state0 != nullptr, &st, &tr_fx, &tr_x, &tr_w, &snap);
        if (rc) return rc;
        if (w_state0)
            memcpy(w_state0 + (size_t)step * snap_w, snap,
                   sizeof(double) * (3 * (size_t)Sw * c.K * (d + 1) + (size_t)c.K * (d + 1) + 2 * (size_t)Sw + 1));
        if (st == 1) { kst_set_error("infinite objective value at best initial point"); return KST_EOBJECTIVE; }
        if (st == 2) { kst_set_error("objective evaluates to NaN on all initial points"); return KST_EOBJECTIVE; }
        if (w_trace_fx) memcpy(w_trace_fx + (size_t)step * itw, tr_fx, sizeof(double) * itw);
        // per step the traces are packed with the step's own K (w_K[step]) inside a slot sized for Kmax
        if (w_trace_points) memcpy(w_trace_points + (size_t)step * itw * Kmax * d, tr_x, sizeof(double) * itw * c.K * d);
        if (w_trace_weights) memcpy(w_trace_weights + (size_t)step * itw * Kmax, tr_w, sizeof(double) * itw * c.K);
        if (w_K) w_K[step] = c.K;
        // the maximizer becomes the next candidate (its points are the fixed ones)
        memcpy(c.w.data(), tr_w + (size_t)(itw - 1) * c.K, sizeof(double) * c.K);
        memcpy(c.pts.data(), tr_x + (size_t)(itw - 1) * c.K * d, sizeof(double) * c.K * d);
        if (steps_done) *steps_done = step + 1;
    }
    *K_out = c.K;
    memcpy(points_out, c.pts.data(), sizeof(double) * c.K * d);
    memcpy(weights_out, c.w.data(), sizeof(double) * c.K);
    return KST_OK;
}
