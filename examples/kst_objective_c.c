/* Pure-C caller of libkirstine_b200.so: shows that the drop-in boundary (include/kirstine_b200.h) needs nothing
 * but a C compiler on the host side, and serves as the C-ABI parity test that does not go through Python's ctypes.
 *
 *   kst_objective_c <in.bin> <out.bin>
 *
 * in.bin  (little-endian, no padding):
 *   int32  model_id, criterion_id, r, K, S, pso_iters
 *   int64  N
 *   double model_param, lb, ub
 *   double theta[N*r]       one parameter struct after the other, like Julia's Vector{P}
 *   double prior_w[N]
 *   double points[S*K]      one-dimensional design points, particle by particle
 *   double weights[S*K]
 * out.bin:
 *   double fx[S]            objective of every particle              (kst_objective_batch)
 *   double M[r*r]           information matrix of particle 0, draw 0 (kst_informationmatrix)
 *   double trace_fx[pso_iters]   best objective per PSO iteration, swarm = the particles above as prototypes
 *
 * Build:  gcc -std=c99 -Iinclude examples/kst_objective_c.c -Lkirstine.jl_b200 -lkirstine_b200 -o kst_objective_c
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include "kirstine_b200.h"

#define CHECK(call)                                                         \
    do {                                                                    \
        int rc_ = (call);                                                   \
        if (rc_ != KST_OK) {                                                \
            fprintf(stderr, "%s failed (%d): %s\n", #call, rc_, kst_last_error()); \
            return 2;                                                       \
        }                                                                   \
    } while (0)

static int read_all(FILE *f, void *dst, size_t bytes) { return fread(dst, 1, bytes, f) == bytes ? 0 : -1; }

int main(int argc, char **argv)
{
    if (argc != 3) {
        fprintf(stderr, "usage: %s in.bin out.bin\n", argv[0]);
        return 1;
    }
    FILE *in = fopen(argv[1], "rb");
    if (!in) { perror(argv[1]); return 1; }
    int32_t hdr[6];
    int64_t N;
    double sc[3];
    if (read_all(in, hdr, sizeof hdr) || read_all(in, &N, sizeof N) || read_all(in, sc, sizeof sc)) return 1;
    const int model_id = hdr[0], crit = hdr[1], r = hdr[2], K = hdr[3], S = hdr[4], iters = hdr[5];
    double *theta = malloc(sizeof(double) * (size_t)N * r), *pw = malloc(sizeof(double) * (size_t)N);
    double *pts = malloc(sizeof(double) * (size_t)S * K), *w = malloc(sizeof(double) * (size_t)S * K);
    double *fx = malloc(sizeof(double) * (size_t)S), *M = malloc(sizeof(double) * (size_t)r * r);
    double *trace = malloc(sizeof(double) * (size_t)(iters > 0 ? iters : 1));
    int *nofix = calloc((size_t)K, sizeof(int));
    if (!theta || !pw || !pts || !w || !fx || !M || !trace || !nofix) return 1;
    if (read_all(in, theta, sizeof(double) * (size_t)N * r) || read_all(in, pw, sizeof(double) * (size_t)N) ||
        read_all(in, pts, sizeof(double) * (size_t)S * K) || read_all(in, w, sizeof(double) * (size_t)S * K))
        return 1;
    fclose(in);

    int ndev = 0;
    CHECK(kst_device_count(&ndev));
    if (ndev < 1) {
        fprintf(stderr, "no CUDA device: the engine has no CPU fallback\n");
        return 3;
    }
    kst_problem *h = NULL;
    CHECK(kst_problem_create(model_id, &sc[0], 1, KST_CP_COPY, NULL, 0, crit, KST_TRAFO_IDENTITY, r, r, 1, 1, N, theta,
                             pw, NULL, &sc[1], &sc[2], 0, &h));
    CHECK(kst_objective_batch(h, S, K, pts, w, fx));
    CHECK(kst_informationmatrix(h, K, pts, w, 0, M));
    if (iters > 0) {
        int status = 0;
        CHECK(kst_pso_init(h, S, K, 2.05, 2.05, pts, w, S, nofix, nofix, 12345u, KST_PSO_MODE_OBJECTIVE, &status));
        if (status != 0) { fprintf(stderr, "pso_init status %d\n", status); return 2; }
        /* the initial state counts as iteration 1 (src/optimize.jl:51-61): its best value is the state's fg */
        CHECK(kst_pso_get_state(h, NULL, NULL, NULL, NULL, NULL, NULL, NULL, NULL, NULL, NULL, &trace[0], NULL));
        if (iters > 1) CHECK(kst_pso_run(h, iters - 1, trace + 1, NULL, NULL));
    }
    CHECK(kst_problem_destroy(h));

    FILE *out = fopen(argv[2], "wb");
    if (!out) { perror(argv[2]); return 1; }
    fwrite(fx, sizeof(double), (size_t)S, out);
    fwrite(M, sizeof(double), (size_t)r * r, out);
    if (iters > 0) fwrite(trace, sizeof(double), (size_t)iters, out);
    fclose(out);
    printf("ok: %d particles x %lld draws, fx[0] = %.17g\n", S, (long long)N, fx[0]);
    return 0;
}
