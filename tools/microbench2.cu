// microbench2.cu — issue rate of register-operand FFMA/FADD vs the packed sm_100 forms
// (FFMA2/FADD2, PTX fma.rn.f32x2 / add.rn.f32x2) the stencil could use.  One block of 512 threads
// per SM x 2 blocks; cycles from clock64 of block 0; prints warp-instructions per clock per SM.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/microbench2 tools/microbench2.cu
#include <cuda_runtime.h>
#include <cstdio>

#define ITERS 2048
#define ILP 8

template <int MODE>
__global__ void k(float *out, const float *in, long long *cycles)
{
    float x[ILP], y[ILP], z[ILP];
    float2 p[ILP], q[ILP], r[ILP];
#pragma unroll
    for (int i = 0; i < ILP; i++) {
        x[i] = in[threadIdx.x + i]; y[i] = in[threadIdx.x + 64 + i]; z[i] = in[threadIdx.x + 128 + i];
        p[i] = make_float2(x[i], y[i]); q[i] = make_float2(y[i], z[i]); r[i] = make_float2(z[i], x[i]);
    }
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < ILP; i++) {
            if (MODE == 0) x[i] = fmaf(x[i], y[i], z[i]);                       // FFMA 3-reg
            if (MODE == 1) x[i] = x[i] + y[i];                                  // FADD
            if (MODE == 2) p[i] = __ffma2_rn(p[i], q[i], r[i]);                 // FFMA2
            if (MODE == 3) p[i] = __fadd2_rn(p[i], q[i]);                       // FADD2
            if (MODE == 4) { p[i] = __ffma2_rn(p[i], q[i], r[i]); x[i] = fmaxf(x[i], y[i]); }   // FFMA2 + ALU op
            if (MODE == 5) { x[i] = fmaf(x[i], y[i], z[i]); y[i] = fmaxf(y[i], z[i]); }          // FFMA + ALU op
            if (MODE == 6) { p[i] = __ffma2_rn(p[i], q[i], r[i]); x[i] = fmaf(x[i], y[i], z[i]); } // FFMA2 + FFMA
        }
    }
    long long t1 = clock64();
    float s = 0;
#pragma unroll
    for (int i = 0; i < ILP; i++) s += x[i] + y[i] + p[i].x + p[i].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}

template <int MODE>
void run(const char *name, double inst_per_iter)
{
    int sms = 148, bpsm = 2, threads = 512;
    float *out, *in; long long *cyc, hc;
    cudaMalloc(&out, sizeof(float) * sms * bpsm * threads);
    cudaMalloc(&in, sizeof(float) * 1024);
    cudaMemset(in, 0, sizeof(float) * 1024);
    cudaMalloc(&cyc, sizeof(long long));
    for (int rep = 0; rep < 2; rep++) k<MODE><<<sms * bpsm, threads>>>(out, in, cyc);
    cudaDeviceSynchronize();
    cudaMemcpy(&hc, cyc, sizeof(hc), cudaMemcpyDeviceToHost);
    double per_sm = (double)bpsm * (threads / 32) * ITERS * ILP * inst_per_iter;
    printf("%-16s cycles %9lld  warp-inst/clk/SM %6.3f\n", name, hc, per_sm / (double)hc);
    cudaFree(out); cudaFree(in); cudaFree(cyc);
}

int main()
{
    run<0>("FFMA 3reg", 1); run<1>("FADD", 1); run<2>("FFMA2", 1); run<3>("FADD2", 1);
    run<4>("FFMA2+FMNMX", 2); run<5>("FFMA+FMNMX", 2); run<6>("FFMA2+FFMA", 2);
    return 0;
}
