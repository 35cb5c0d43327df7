// microbench6.cu — issue cost of FFMA2/FMUL2 forms whose coefficient is a per-lane SCALAR register
// broadcast to both halves (SASS operand `R.F32`), against the three-register-pair form.
// Same method as microbench5 (cycles per warp-instruction per SM sub-partition at 1.965 GHz).
#include <cuda_runtime.h>
#include <cstdio>
#define ITERS 8192
#define ILP 8
template <int MODE>
__global__ void __launch_bounds__(256) k(float *out, const float *in)
{
    float2 p[ILP], q[ILP], r[ILP];
    float s[ILP], t[ILP];
#pragma unroll
    for (int i = 0; i < ILP; i++) {
        const float x = in[threadIdx.x + i], y = in[threadIdx.x + 64 + i], z = in[threadIdx.x + 128 + i];
        p[i] = make_float2(x, y); q[i] = make_float2(y, z); r[i] = make_float2(z, x);
        s[i] = in[threadIdx.x + 192 + i]; t[i] = in[threadIdx.x + 256 + i];
    }
#pragma unroll 1
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < ILP; i++) {
            const float2 s2 = make_float2(s[i], s[i]), t2 = make_float2(t[i], t[i]);
            if (MODE == 0) p[i] = __ffma2_rn(p[i], q[i], r[i]);        // pair,pair,pair
            if (MODE == 1) p[i] = __ffma2_rn(s2, q[i], p[i]);          // scalar,pair,(dst)
            if (MODE == 2) p[i] = __ffma2_rn(s2, p[i], r[i]);          // scalar,pair,pair
            if (MODE == 3) p[i] = __ffma2_rn(p[i], q[i], s2);          // pair,pair,scalar
            if (MODE == 4) p[i] = __ffma2_rn(s2, p[i], t2);            // scalar,pair,scalar
            if (MODE == 5) p[i] = __fmul2_rn(p[i], q[i]);              // FMUL2 pair,pair
            if (MODE == 6) p[i] = __fmul2_rn(p[i], s2);                // FMUL2 pair,scalar
            if (MODE == 7) p[i] = __ffma2_rn(s2, q[(i + 1) % ILP], p[i]);  // scalar,pair,(dst), no shared operands
        }
    }
    float acc = 0;
#pragma unroll
    for (int i = 0; i < ILP; i++) acc += p[i].x + p[i].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}
template <int MODE>
void run(const char *name)
{
    float *out, *in;
    int bpsm = 4;
    cudaMalloc(&out, 4 * 148 * bpsm * 256); cudaMalloc(&in, 4 * 1024); cudaMemset(in, 0, 4 * 1024);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE><<<148 * bpsm, 256>>>(out, in);
    cudaEventRecord(e0);
    k<MODE><<<148 * bpsm, 256>>>(out, in);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double inst_per_smsp = 8.0 * ITERS * ILP;
    printf("%-28s %6.3f cycles per warp-instruction per SMSP\n", name, ms * 1e-3 * 1.965e9 / inst_per_smsp);
    cudaFree(out); cudaFree(in);
}
int main()
{
    run<0>("FFMA2 pair,pair,pair"); run<1>("FFMA2 scal,pair,(dst)"); run<7>("FFMA2 scal,pair,(dst) b");
    run<2>("FFMA2 scal,pair,pair"); run<3>("FFMA2 pair,pair,scal"); run<4>("FFMA2 scal,pair,scal");
    run<5>("FMUL2 pair,pair"); run<6>("FMUL2 pair,scal");
    return 0;
}
