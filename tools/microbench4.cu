// microbench4.cu — the quiet-run stencil loop in isolation (K = 12 cells per lane, one warp per
// wire), in several formulations, at 16 and 28 warps per SM.  Prints SM cycles per warp-iteration
// per SM sub-partition (= the cost of one env-µs of stencil on one scheduler).
#include <cuda_runtime.h>
#include <cstdio>
#define FULL 0xffffffffu
constexpr int K = 12, H = 6;
__device__ __forceinline__ float2 sub2(float2 a, float2 b) { return __fadd2_rn(a, make_float2(-b.x, -b.y)); }

template <int V>
__global__ void __launch_bounds__(128) k(float *out, const float *in, int n, long long *cycles)
{
    const int lane = threadIdx.x & 31;
    float spoolf = in[0];
    float ncL = in[1];
    long long t0 = 0, t1 = 0;
    if (V == 0) {                                   // scalar FFMA/FADD
        float T[K], cE[K], c0[K], cR[K];
#pragma unroll
        for (int j = 0; j < K; j++) { T[j] = in[2 + j + lane]; cE[j] = in[40 + j]; c0[j] = in[60 + j]; cR[j] = in[80 + j]; }
        t0 = clock64();
#pragma unroll 1
        for (int it = 0; it < n; it++) {
            float left = __shfl_up_sync(FULL, T[K - 1], 1);
            const float right = __shfl_down_sync(FULL, T[0], 1);
            if (lane == 0) left = spoolf;
            float dprev = T[0] - left;
#pragma unroll
            for (int j = 0; j < K; j++) {
                const float dnext = ((j + 1 < K) ? T[j + 1] : right) - T[j];
                float q = fmaf(cE[j], T[j], c0[j]);
                q = fmaf(cR[j], dnext, q);
                q = fmaf(ncL, dprev, q);
                dprev = dnext;
                T[j] = T[j] + q;
            }
        }
        t1 = clock64();
        float s = 0;
#pragma unroll
        for (int j = 0; j < K; j++) s += T[j];
        out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    } else {                                        // packed variants
        float2 T[H], cE[H], c0[H], cR[H];
#pragma unroll
        for (int j = 0; j < H; j++) {
            T[j] = make_float2(in[2 + j + lane], in[8 + j + lane]);
            cE[j] = make_float2(in[40 + j], in[46 + j]); c0[j] = make_float2(in[60 + j], in[66 + j]);
            cR[j] = make_float2(in[80 + j], in[86 + j]);
        }
        const float2 ncL2 = make_float2(ncL, ncL);
        t0 = clock64();
#pragma unroll 1
        for (int it = 0; it < n; it++) {
            float left = __shfl_up_sync(FULL, T[H - 1].y, 1);
            const float right = __shfl_down_sync(FULL, T[0].x, 1);
            if (lane == 0) left = spoolf;
            if (V == 1) {                           // as in step_kernel2.cuh
                float2 dprev = sub2(T[0], make_float2(left, T[H - 1].x));
                const float2 last_next = make_float2(T[0].y, right);
#pragma unroll
                for (int j = 0; j < H; j++) {
                    const float2 dnext = sub2((j + 1 < H) ? T[j + 1] : last_next, T[j]);
                    float2 q = __ffma2_rn(cE[j], T[j], c0[j]);
                    q = __ffma2_rn(cR[j], dnext, q);
                    q = __ffma2_rn(ncL2, dprev, q);
                    dprev = dnext;
                    T[j] = __fadd2_rn(T[j], q);
                }
            }
            if (V == 2) {                           // packed FMAs, scalar differences
                float dpx = T[0].x - left, dpy = T[0].y - T[H - 1].x;
                const float t0y = T[0].y;
#pragma unroll
                for (int j = 0; j < H; j++) {
                    const float nx = (j + 1 < H) ? T[j + 1].x : t0y;
                    const float ny = (j + 1 < H) ? T[j + 1].y : right;
                    const float dnx = nx - T[j].x, dny = ny - T[j].y;
                    float2 q = __ffma2_rn(cE[j], T[j], c0[j]);
                    q = __ffma2_rn(cR[j], make_float2(dnx, dny), q);
                    q = __ffma2_rn(ncL2, make_float2(dpx, dpy), q);
                    dpx = dnx; dpy = dny;
                    T[j] = __fadd2_rn(T[j], q);
                }
            }
            if (V == 3) {                           // first FMA scalar (2 FFMA), rest packed
                float2 dprev = sub2(T[0], make_float2(left, T[H - 1].x));
                const float2 last_next = make_float2(T[0].y, right);
#pragma unroll
                for (int j = 0; j < H; j++) {
                    const float2 dnext = sub2((j + 1 < H) ? T[j + 1] : last_next, T[j]);
                    float2 q = make_float2(fmaf(cE[j].x, T[j].x, c0[j].x), fmaf(cE[j].y, T[j].y, c0[j].y));
                    q = __ffma2_rn(cR[j], dnext, q);
                    q = __ffma2_rn(ncL2, dprev, q);
                    dprev = dnext;
                    T[j] = __fadd2_rn(T[j], q);
                }
            }
        }
        t1 = clock64();
        float s = 0;
#pragma unroll
        for (int j = 0; j < H; j++) s += T[j].x + T[j].y;
        out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    }
    if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}

template <int V>
void run(const char *name, int blocks_per_sm)
{
    float *out, *in; long long *cyc, hc;
    const int n = 20000;
    cudaMalloc(&out, 4 * 148 * 8 * 128); cudaMalloc(&in, 4 * 256); cudaMemset(in, 0, 4 * 256); cudaMalloc(&cyc, 8);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<V><<<148 * blocks_per_sm, 128>>>(out, in, n, cyc);
    cudaEventRecord(e0);
    k<V><<<148 * blocks_per_sm, 128>>>(out, in, n, cyc);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    cudaMemcpy(&hc, cyc, 8, cudaMemcpyDeviceToHost);
    printf("%-28s %2d warps/SM  %7.1f clock64 / %6.2f ns per warp-iteration per SMSP  (clock64 rate %.0f MHz)\n", name,
           blocks_per_sm * 4, (double)hc / n / blocks_per_sm, ms * 1e6 / n / blocks_per_sm, hc / (ms * 1e3));
    cudaFree(out); cudaFree(in); cudaFree(cyc);
}
int main()
{
    for (int b = 1; b <= 8; b++) { if (b == 2 || b == 3 || b == 6) continue;
        run<0>("scalar FFMA/FADD", b); run<1>("packed (kernel)", b); run<2>("packed FMA, scalar diffs", b);
        run<3>("scalar first FMA, packed rest", b);
    }
    return 0;
}
