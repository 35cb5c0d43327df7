// microbench.cu — measured issue rates of the pipes the fused kernel uses (B200, sm_100a).
// Prints warp-instructions per clock per SM for FFMA, FADD, FFMA+FADD mix, FMNMX, DFMA, DADD,
// SHFL, and the implied FP32 peak at the observed clock.  Used for the roofline denominator
// (MEASURED_PEAKS.json has no FP32 entry).
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>

#define ITERS 4096
#define ILP 8

template <int MODE>
__global__ void k(float *out, double *dout, float a, float b, long long *cycles)
{
    float x[ILP];
    double d[ILP];
#pragma unroll
    for (int i = 0; i < ILP; i++) { x[i] = threadIdx.x * 0.001f + i; d[i] = threadIdx.x * 0.001 + i; }
    long long t0 = clock64();
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < ILP; i++) {
            if (MODE == 0) x[i] = fmaf(x[i], a, b);
            if (MODE == 1) x[i] = x[i] + a;
            if (MODE == 2) { if (i & 1) x[i] = x[i] + a; else x[i] = fmaf(x[i], a, b); }
            if (MODE == 3) x[i] = fmaxf(x[i], a + i);
            if (MODE == 4) d[i] = fma(d[i], (double)a, (double)b);
            if (MODE == 5) d[i] = d[i] + (double)a;
            if (MODE == 6) x[i] = __shfl_up_sync(0xffffffffu, x[i], 1);
            if (MODE == 7) { x[i] = fmaf(x[i], a, b); if ((i & 3) == 0) x[i] = fmaxf(x[i], b); }
            if (MODE == 8) d[i] = d[i] / ((double)a + d[i]);
        }
    }
    long long t1 = clock64();
    float s = 0; double ds = 0;
#pragma unroll
    for (int i = 0; i < ILP; i++) { s += x[i]; ds += d[i]; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    dout[blockIdx.x * blockDim.x + threadIdx.x] = ds;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}

template <int MODE>
void run(const char *name, double ops_per_inst)
{
    int sms = 148, bpsm = 4, threads = 512;
    float *out; double *dout; long long *cyc, hc;
    cudaMalloc(&out, sizeof(float) * sms * bpsm * threads);
    cudaMalloc(&dout, sizeof(double) * sms * bpsm * threads);
    cudaMalloc(&cyc, sizeof(long long));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE><<<sms * bpsm, threads>>>(out, dout, 1.0001f, 0.5f, cyc);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    k<MODE><<<sms * bpsm, threads>>>(out, dout, 1.0001f, 0.5f, cyc);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    cudaMemcpy(&hc, cyc, sizeof(hc), cudaMemcpyDeviceToHost);
    double warp_insts = (double)sms * bpsm * (threads / 32) * ITERS * ILP;
    double per_clk_sm = warp_insts / sms / (double)hc;   // block 0's cycles ~ whole kernel
    double rate = warp_insts * 32 * ops_per_inst / (ms * 1e-3);
    printf("%-14s %8.3f ms  cycles %9lld  warp-inst/clk/SM %6.3f  %8.2f Tops/s  (eff clk %.0f MHz)\n", name, ms,
           hc, per_clk_sm, rate / 1e12, hc / (ms * 1e-3) / 1e6);
    cudaFree(out); cudaFree(dout); cudaFree(cyc);
}

int main()
{
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    printf("device %s, %d SMs, clock %d kHz\n", p.name, p.multiProcessorCount, p.clockRate);
    run<0>("FFMA", 2); run<1>("FADD", 1); run<2>("FFMA+FADD", 1.5); run<3>("FMNMX", 1);
    run<7>("FFMA+.25FMNMX", 1.6);
    run<4>("DFMA", 2); run<5>("DADD", 1); run<8>("DDIV", 1); run<6>("SHFL", 1);
    return 0;
}
