// Issue-rate microbenchmark: scalar FFMA vs packed FFMA2 (fma.rn.f32x2) on sm_100a.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ffma2 ffma2.cu && ./ffma2
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE>
__global__ void __launch_bounds__(256) k(float *out, int iters, float a, float b)
{
    float2 x[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) x[i] = make_float2(threadIdx.x * 0.001f + i, threadIdx.x * 0.002f - i);
    const float2 A = make_float2(a, a), B = make_float2(b, b);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (MODE == 0) { x[i].x = fmaf(x[i].x, a, b); x[i].y = fmaf(x[i].y, a, b); }            // 2 FFMA (reg, reg, reg)
            else if (MODE == 1) x[i] = __ffma2_rn(x[i], A, B);                                          // 1 FFMA2
            else if (MODE == 2) { x[i] = __ffma2_rn(x[i], A, B); x[i].x = fmaf(x[i].x, a, b); }        // mix 1 FFMA2 + 1 FFMA
            else if (MODE == 3) { x[i].x = fmaf(x[i].x, 1.0001f, b); x[i].y = fmaf(x[i].y, 1.0001f, b); } // imm form
            else if (MODE == 4) x[i] = __fadd2_rn(__fmul2_rn(x[i], A), B);
            else if (MODE == 5) x[i] = __ffma2_rn(x[i], x[(i + 1) & 7], x[(i + 2) & 7]);                         // all-pair operands                              // FMUL2 + FADD2
        }
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += x[i].x + x[i].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE> void run(const char *name, int instr_per_iter)
{
    float *out; cudaMalloc(&out, 148 * 8 * 256 * 4);
    const int iters = 20000;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE><<<148 * 8, 256>>>(out, 100, 1.0001f, 0.5f);
    cudaEventRecord(e0);
    k<MODE><<<148 * 8, 256>>>(out, iters, 1.0001f, 0.5f);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double warp_instr = (double)148 * 8 * 8 * iters * instr_per_iter;   // CTAs * warps * iters * instr
    int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    printf("%-28s %.3f ms  %.2f warp-instr/clk/SM (at %d MHz nominal)  %.1f TFLOP/s fp32\n", name, ms,
           warp_instr / (ms * 1e-3) / (clk * 1e3) / 148, clk / 1000,
           (double)148 * 8 * 256 * iters * 8 * 2 * 2 * (MODE == 2 ? 1.5 : 1.0) / (ms * 1e-3) / 1e12);
    cudaFree(out);
}

int main()
{
    run<0>("2x FFMA rrr", 16);
    run<3>("2x FFMA imm", 16);
    run<1>("FFMA2", 8);
    run<2>("FFMA2 + FFMA", 16);
    run<4>("FMUL2 + FADD2", 16);
    run<5>("FFMA2 pair operands", 8);
    return 0;
}
