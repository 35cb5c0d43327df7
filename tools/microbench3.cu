// microbench3.cu — latency (dependent chain, one warp per SM sub-partition) and throughput (many
// warps, independent chains) of the fp64 and packed-fp32 operations the step kernel issues.
#include <cuda_runtime.h>
#include <cstdio>
#define ITERS 4096
template <int MODE>
__global__ void lat(double *out, const double *in, long long *cycles)
{
    double x = in[threadIdx.x & 31], y = in[32], z = in[33];
    float2 p = make_float2((float)x, (float)y), q = make_float2((float)y, (float)z);
    float f = (float)x;
    long long t0 = clock64();
#pragma unroll 16
    for (int it = 0; it < ITERS; it++) {
        if (MODE == 0) x = x + y;
        if (MODE == 1) x = x * y;
        if (MODE == 2) x = fma(x, y, z);
        if (MODE == 3) p = __ffma2_rn(p, q, q);
        if (MODE == 4) p = __fadd2_rn(p, q);
        if (MODE == 5) f = fmaf(f, q.x, q.y);
        if (MODE == 6) f = __shfl_up_sync(0xffffffffu, f, 1);
    }
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = x + p.x + p.y + f;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}
template <int MODE>
__global__ void thr(double *out, const double *in, long long *cycles)
{
    double x[8];
    double y = in[32], z = in[33];
#pragma unroll
    for (int i = 0; i < 8; i++) x[i] = in[(threadIdx.x + i) & 31];
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS / 8; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            if (MODE == 0) x[i] = x[i] + y;
            if (MODE == 1) x[i] = x[i] * y;
            if (MODE == 2) x[i] = fma(x[i], y, z);
        }
    }
    long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s += x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}
int main()
{
    double *out, *in; long long *cyc, hc;
    cudaMalloc(&out, 8 * 148 * 1024); cudaMalloc(&in, 8 * 64); cudaMemset(in, 0, 8 * 64); cudaMalloc(&cyc, 8);
    const char *names[] = {"DADD", "DMUL", "DFMA", "FFMA2", "FADD2", "FFMA", "SHFL"};
#define LAT(M) lat<M><<<148, 32>>>(out, in, cyc); lat<M><<<148, 32>>>(out, in, cyc); cudaDeviceSynchronize(); \
    cudaMemcpy(&hc, cyc, 8, cudaMemcpyDeviceToHost); printf("latency %-6s %6.2f cycles\n", names[M], (double)hc / ITERS);
    LAT(0) LAT(1) LAT(2) LAT(3) LAT(4) LAT(5) LAT(6)
#define THR(M) thr<M><<<148, 1024>>>(out, in, cyc); thr<M><<<148, 1024>>>(out, in, cyc); cudaDeviceSynchronize(); \
    cudaMemcpy(&hc, cyc, 8, cudaMemcpyDeviceToHost); printf("throughput %-6s %6.3f warp-inst/clk/SM\n", names[M], 32.0 * ITERS / (double)hc);
    THR(0) THR(1) THR(2)
    return 0;
}
