// FP64-pipe micro-benchmarks (tuning aid, not part of the library): which instruction mixes reach the DFMA peak?
#include <cstdio>
#include <cuda_runtime.h>
__constant__ double cc[4] = {0.999999, 1e-9, 1.000001, -1e-9};
template <int MODE>
__global__ void __launch_bounds__(256) k(double *out, int iters, double a, double b)
{
    double x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
    int n0 = threadIdx.x, n1 = 1, n2 = 2, n3 = 3;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int u = 0; u < 8; u++) {
            if (MODE == 0) {  // reg, reg(const), reg(const)
                x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
                x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
            } else if (MODE == 1) {  // three varying register operands
                x0 = fma(x0, x4, x1); x1 = fma(x1, x5, x2); x2 = fma(x2, x6, x3); x3 = fma(x3, x7, x0);
                x4 = fma(x4, x0, x5); x5 = fma(x5, x1, x6); x6 = fma(x6, x2, x7); x7 = fma(x7, x3, x4);
            } else if (MODE == 2) {  // constant-bank operands
                x0 = fma(x0, cc[0], cc[1]); x1 = fma(x1, cc[2], cc[3]); x2 = fma(x2, cc[0], cc[1]); x3 = fma(x3, cc[2], cc[3]);
                x4 = fma(x4, cc[0], cc[1]); x5 = fma(x5, cc[2], cc[3]); x6 = fma(x6, cc[0], cc[1]); x7 = fma(x7, cc[2], cc[3]);
            } else if (MODE == 3) {  // DMUL / DADD mix
                x0 = x0 * a; x1 = x1 + b; x2 = x2 * a; x3 = x3 + b; x4 = x4 * a; x5 = x5 + b; x6 = x6 * a; x7 = x7 + b;
            } else if (MODE == 4) {  // DFMA with one integer op per FP64 op
                x0 = fma(x0, a, b); n0 = n0 * 3 + n1; x1 = fma(x1, a, b); n1 = n1 ^ n2; x2 = fma(x2, a, b); n2 = n2 * 5 + n3;
                x3 = fma(x3, a, b); n3 = n3 ^ n0; x4 = fma(x4, a, b); n0 = n0 * 7 + n2; x5 = fma(x5, a, b); n1 = n1 ^ n3;
                x6 = fma(x6, a, b); n2 = n2 * 9 + n0; x7 = fma(x7, a, b); n3 = n3 ^ n1;
            } else if (MODE == 5) {  // 4 chains only (less ILP)
                x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
                x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
            } else if (MODE == 6) {  // 2 chains
                x0 = fma(x0, a, b); x1 = fma(x1, a, b); x0 = fma(x0, a, b); x1 = fma(x1, a, b);
                x0 = fma(x0, a, b); x1 = fma(x1, a, b); x0 = fma(x0, a, b); x1 = fma(x1, a, b);
            } else if (MODE == 7) {  // 1 chain
                x0 = fma(x0, a, b); x0 = fma(x0, a, b); x0 = fma(x0, a, b); x0 = fma(x0, a, b);
                x0 = fma(x0, a, b); x0 = fma(x0, a, b); x0 = fma(x0, a, b); x0 = fma(x0, a, b);
            }
        }
    }
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7 + n0 + n1 + n2 + n3;
}
template <int MODE>
void run(const char *name, int blocks_per_sm, int nt)
{
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int blocks = p.multiProcessorCount * blocks_per_sm, iters = 4096;
    double *d; cudaMalloc(&d, sizeof(double) * blocks * nt);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int r = 0; r < 4; r++) {
        cudaEventRecord(e0); k<MODE><<<blocks, nt>>>(d, iters, 0.999999, 1e-9); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (r && ms < best) best = ms;
    }
    double inst = 64.0 * iters * nt * blocks;  // FP64 instructions (thread level)
    printf("%-34s blocks/SM %d x %3d thr: %7.2f Ginst/s/SM-clk-norm  %6.2f TFLOP/s-equiv (as FMA)\n", name, blocks_per_sm, nt,
           inst / (best * 1e-3) / 1e9, 2 * inst / (best * 1e-3) / 1e12);
    cudaFree(d);
}
int main()
{
    run<0>("DFMA r,const-reg,const-reg 8ch", 8, 256);
    run<1>("DFMA 3 varying regs 8ch", 8, 256);
    run<2>("DFMA const-bank operands 8ch", 8, 256);
    run<3>("DMUL/DADD mix 8ch", 8, 256);
    run<4>("DFMA + 1 int op each 8ch", 8, 256);
    run<5>("DFMA 4 chains", 8, 256);
    run<6>("DFMA 2 chains", 8, 256);
    run<7>("DFMA 1 chain", 8, 256);
    run<7>("DFMA 1 chain, 3 CTA/SM (24 warps)", 3, 256);
    run<6>("DFMA 2 chains, 3 CTA/SM (24 warps)", 3, 256);
    run<5>("DFMA 4 chains, 3 CTA/SM (24 warps)", 3, 256);
    run<7>("DFMA 1 chain, 1 CTA/SM (8 warps)", 1, 256);
    run<6>("DFMA 2 chains, 1 CTA/SM (8 warps)", 1, 256);
    return 0;
}
