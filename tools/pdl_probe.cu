// pdl_probe.cu — does programmatic dependent launch overlap a primary that triggers early with its
// secondary on this driver, on a created stream and on the legacy default stream?
#include <cuda_runtime.h>
#include <cstdio>
__device__ __forceinline__ unsigned long long gtime()
{
    unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t;
}
__global__ void primary(int trigger, unsigned long long ns, int *sink)
{
    if (trigger) asm volatile("griddepcontrol.launch_dependents;");
    const unsigned long long t0 = gtime();
    while (gtime() - t0 < ns) {}
    if (sink && threadIdx.x == 9999) *sink = 1;
}
__global__ void secondary(unsigned long long ns, int *sink)
{
    const unsigned long long t0 = gtime();
    while (gtime() - t0 < ns) {}
    if (sink && threadIdx.x == 9999) *sink = 1;
}
static float run(cudaStream_t st, int trigger, int pdl)
{
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e9f;
    for (int rep = 0; rep < 5; rep++) {
        cudaEventRecord(e0, st);
        primary<<<148, 32, 0, st>>>(trigger, 100000ull, nullptr);
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(2048); cfg.blockDim = dim3(64); cfg.stream = st;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
        cudaLaunchKernelEx(&cfg, secondary, 100000ull, (int *)nullptr);
        cudaEventRecord(e1, st);
        cudaStreamSynchronize(st);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    return best * 1000.f;
}
int main()
{
    cudaStream_t st; cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking);
    cudaStream_t st2; cudaStreamCreate(&st2);
    printf("two 100-us kernels back to back [us]:\n");
    printf("  created non-blocking stream, plain launches      %7.1f\n", run(st, 0, 0));
    printf("  created non-blocking stream, PDL, early trigger  %7.1f\n", run(st, 1, 1));
    printf("  created non-blocking stream, PDL, no trigger     %7.1f\n", run(st, 0, 1));
    printf("  created blocking stream, PDL, early trigger      %7.1f\n", run(st2, 1, 1));
    printf("  legacy default stream, PDL, early trigger        %7.1f\n", run(0, 1, 1));
    printf("  per-thread default stream handle, PDL            %7.1f\n", run(cudaStreamPerThread, 1, 1));
    printf("err: %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
