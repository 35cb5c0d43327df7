// microbench5.cu — wall-clock issue cost (SM cycles per warp-instruction per SM sub-partition at
// 1.965 GHz) of fp32 forms by operand kind: r = register, c = constant bank / uniform.
#include <cuda_runtime.h>
#include <cstdio>
#define ITERS 8192
#define ILP 8
template <int MODE>
__global__ void __launch_bounds__(256) k(float *out, const float *in, float ca, float cb)
{
    float x[ILP], y[ILP], z[ILP];
    float2 p[ILP], q[ILP], r[ILP];
#pragma unroll
    for (int i = 0; i < ILP; i++) {
        x[i] = in[threadIdx.x + i]; y[i] = in[threadIdx.x + 64 + i]; z[i] = in[threadIdx.x + 128 + i];
        p[i] = make_float2(x[i], y[i]); q[i] = make_float2(y[i], z[i]); r[i] = make_float2(z[i], x[i]);
    }
    const float2 c2a = make_float2(ca, ca), c2b = make_float2(cb, cb);
#pragma unroll 1
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < ILP; i++) {
            if (MODE == 0) x[i] = fmaf(x[i], y[i], z[i]);          // FFMA rrr
            if (MODE == 1) x[i] = fmaf(x[i], ca, z[i]);            // FFMA rcr
            if (MODE == 2) x[i] = fmaf(x[i], ca, cb);              // FFMA rcc
            if (MODE == 3) x[i] = x[i] + y[i];                     // FADD rr
            if (MODE == 4) x[i] = x[i] + ca;                       // FADD rc
            if (MODE == 5) p[i] = __ffma2_rn(p[i], q[i], r[i]);    // FFMA2 rrr
            if (MODE == 6) p[i] = __ffma2_rn(p[i], c2a, r[i]);     // FFMA2 r c r
            if (MODE == 7) p[i] = __ffma2_rn(p[i], c2a, c2b);      // FFMA2 r c c
            if (MODE == 8) p[i] = __fadd2_rn(p[i], q[i]);          // FADD2 rr
            if (MODE == 9) p[i] = __fadd2_rn(p[i], c2a);           // FADD2 rc
            if (MODE == 10) p[i] = __ffma2_rn(p[i], q[i], p[i]);   // FFMA2 with dst==src (2 distinct pairs)
            if (MODE == 11) x[i] = fmaf(x[i], y[i], x[i]);         // FFMA 2 distinct regs
            if (MODE == 12) x[i] = fmaf(y[i], z[i], x[i]);         // FFMA accumulate rrr (dst = src3)
        }
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < ILP; i++) s += x[i] + p[i].x + p[i].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int MODE>
void run(const char *name)
{
    float *out, *in;
    int bpsm = 4;   // 4 x 256 threads = 32 warps / SM = 8 per SMSP
    cudaMalloc(&out, 4 * 148 * bpsm * 256); cudaMalloc(&in, 4 * 1024); cudaMemset(in, 0, 4 * 1024);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE><<<148 * bpsm, 256>>>(out, in, 1.0001f, 0.5f);
    cudaEventRecord(e0);
    k<MODE><<<148 * bpsm, 256>>>(out, in, 1.0001f, 0.5f);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double inst_per_smsp = 8.0 * ITERS * ILP;      // 8 warps per SMSP
    printf("%-22s %6.3f cycles per warp-instruction per SMSP\n", name, ms * 1e-3 * 1.965e9 / inst_per_smsp);
    cudaFree(out); cudaFree(in);
}
int main()
{
    run<0>("FFMA rrr"); run<1>("FFMA rcr"); run<2>("FFMA rcc"); run<11>("FFMA r,r,same"); run<12>("FFMA acc rrr");
    run<3>("FADD rr"); run<4>("FADD rc");
    run<5>("FFMA2 rrr"); run<6>("FFMA2 rcr"); run<7>("FFMA2 rcc"); run<10>("FFMA2 r,r,same"); run<8>("FADD2 rr"); run<9>("FADD2 rc");
    return 0;
}
