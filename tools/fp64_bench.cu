// fp64_bench.cu -- what the FP64 pipe of this GPU does: DFMA latency (one dependent chain), DFMA / MUFU.RCP64H throughput per SM
// with W warps per SM and C independent chains per thread.  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_bench fp64_bench.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int C>
__global__ void dfma_kernel(double* out, int iters, double a, double b)
{
    double x[C];
#pragma unroll
    for (int c = 0; c < C; c++) x[c] = a + c + threadIdx.x;
    const long long t0 = clock64();
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int c = 0; c < C; c++) x[c] = __fma_rn(x[c], b, a);
    }
    const long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int c = 0; c < C; c++) s += x[c];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s + (double)(t1 - t0) * 1e-300;
    if (threadIdx.x == 0 && blockIdx.x == 0) ((long long*)out)[gridDim.x * blockDim.x] = t1 - t0;
}

template <int C>
__global__ void rcp_kernel(double* out, int iters, double a)
{
    double x[C];
#pragma unroll
    for (int c = 0; c < C; c++) x[c] = a + c + threadIdx.x;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int c = 0; c < C; c++) { double r; asm volatile("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x[c])); x[c] = r; }
    }
    double s = 0;
#pragma unroll
    for (int c = 0; c < C; c++) s += x[c];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <class F> static float time_ms(F f)
{
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    f(); cudaDeviceSynchronize();
    cudaEventRecord(a); f(); cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b); return ms;
}

int main()
{
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int khz = 0; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    const int sms = p.multiProcessorCount;
    double* out; cudaMalloc(&out, (size_t)sms * 8 * 1024 * 8 + 64);
    printf("%s, %d SMs, %.0f MHz nominal\n", p.name, sms, khz / 1e3);
    const int iters = 20000;
    {   // latency: one warp, one chain
        dfma_kernel<1><<<1, 32>>>(out, iters, 1.0, 0.999999);
        cudaDeviceSynchronize();
        long long cyc; cudaMemcpy(&cyc, (long long*)out + 32, 8, cudaMemcpyDeviceToHost);
        printf("DFMA dependent-chain latency: %.2f cycles\n", (double)cyc / iters);
    }
#define TP(C, W)                                                                                                         \
    {                                                                                                                    \
        float ms = time_ms([&] { dfma_kernel<C><<<sms, 32 * W>>>(out, iters, 1.0, 0.999999); });                          \
        double inst = (double)sms * W * C * iters;                                                                       \
        printf("DFMA  chains/thread %d warps/SM %2d: %.3f warp-instr/ns/SM-equivalent -> %.2f warp-instr per SM per cycle at %.0f MHz\n", \
               C, W, inst / (ms * 1e6) / sms, inst / (ms * 1e-3) / sms / (khz * 1e3), khz / 1e3);                          \
    }
    TP(1, 4) TP(1, 8) TP(1, 16) TP(1, 32) TP(4, 4) TP(4, 8) TP(4, 16) TP(8, 4) TP(8, 8) TP(8, 16) TP(8, 32)
#define TR(C, W)                                                                                                         \
    {                                                                                                                    \
        float ms = time_ms([&] { rcp_kernel<C><<<sms, 32 * W>>>(out, iters, 1.5); });                                     \
        double inst = (double)sms * W * C * iters;                                                                       \
        printf("RCP64H chains/thread %d warps/SM %2d: %.3f warp-instr per SM per cycle\n", C, W,                          \
               inst / (ms * 1e-3) / sms / (khz * 1e3));                                                                  \
    }
    TR(4, 8) TR(8, 16) TR(8, 32)
    return 0;
}
