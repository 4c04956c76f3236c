// rf_bench.cu -- does the number of distinct register-pair operands bound one warp's FP64 issue rate?  One warp per SM
// sub-partition (4 warps per SM), 8 independent chains per thread, DFMA with 1 / 2 / 3 distinct 64-bit register operands.
#include <cstdio>
#include <cuda_runtime.h>
template <int MODE, int C>
__global__ void k(double* out, long long* cyc, int iters)
{
    double x[C], y[C], z[C];
#pragma unroll
    for (int c = 0; c < C; c++) { x[c] = 1.0 + threadIdx.x + c; y[c] = 0.999 + 1e-6 * c; z[c] = 1e-3 * (c + 1); }
    const double a = out[0], b = out[1];
    const long long t0 = clock64();
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int c = 0; c < C; c++) {
            if (MODE == 1) x[c] = __fma_rn(x[c], b, a);              // one distinct register operand per instruction (a, b reused)
            if (MODE == 2) x[c] = __fma_rn(x[c], y[c], a);           // two
            if (MODE == 3) x[c] = __fma_rn(x[c], y[c], z[c]);        // three
            if (MODE == 4) { x[c] = __fma_rn(x[c], y[c], z[c]); y[c] = __fma_rn(y[c], z[c], x[c]); }   // results feed other chains' operands
        }
    }
    const long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int c = 0; c < C; c++) s += x[c] + y[c] + z[c];
    out[2 + blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <int MODE, int C> void run(const char* name, int warps)
{
    double* out; long long* cyc; cudaMalloc(&out, (148 * 1024 + 2) * 8); cudaMalloc(&cyc, 8);
    double ab[2] = { 1e-3, 0.999 }; cudaMemcpy(out, ab, 16, cudaMemcpyHostToDevice);
    const int iters = 4000;
    k<MODE, C><<<148, 32 * warps>>>(out, cyc, iters); cudaDeviceSynchronize();
    long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
    const int per = MODE == 4 ? 2 * C : C;
    printf("%-44s warps/SM %2d: %5.2f cycles per DFMA per warp\n", name, warps, (double)c / iters / per);
    cudaFree(out); cudaFree(cyc);
}
int main()
{
    run<1, 8>("8 chains, 1 distinct register-pair operand", 4);
    run<2, 8>("8 chains, 2 distinct register-pair operands", 4);
    run<3, 8>("8 chains, 3 distinct register-pair operands", 4);
    run<4, 8>("16 chains, 3 distinct, cross-feeding", 4);
    run<3, 8>("8 chains, 3 distinct register-pair operands", 8);
    run<3, 8>("8 chains, 3 distinct register-pair operands", 16);
    run<3, 8>("8 chains, 3 distinct register-pair operands", 32);
    return 0;
}
