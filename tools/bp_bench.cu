// bp_bench.cu -- cycles of one check pass / variable pass / syndrome check of bp_ref.cuh for the n=300 geometry (m = 148 rows of 6,
// 300 columns of <= 3) with W warps per SM: what a lone warp needs, and what W warps sharing the FP64 pipe need.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I gr-dtl_b200/csrc -o tools/bp_bench tools/bp_bench.cu
#include <cstdio>
#include <vector>
#include "bp_ref.cuh"
using namespace dtl;

template <int VAR>
__global__ void bench_kernel(const uint4* rv2g, const uint4* vdg, long long* cyc, int reps, int n, int m, int nvr, int z)
{
    extern __shared__ __align__(16) unsigned char sm[];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, W = blockDim.x >> 5;
    uint4* rv2 = (uint4*)sm; uint4* vd = rv2 + m;
    for (int i = threadIdx.x; i < m; i += blockDim.x) rv2[i] = rv2g[i];
    for (int i = threadIdx.x; i < nvr * 32; i += blockDim.x) vd[i] = vdg[i];
    __syncthreads();
    FastS c; c.n = n; c.m = m; c.k = 152; c.rank = 148; c.rs = 6; c.nvr = nvr; c.max_row = 6; c.max_col = 3; c.uniform_rows = 1; c.dummy = m * 6;
    c.o_rv2 = 0; c.o_vdesc = 16 * m; c.o_permute = 0; c.smb = sm;
    unsigned char* base = sm + 16 * (m + nvr * 32) + (size_t)wid * fast_warp_bytes(n, m, 6);
    FastBuf w = fast_carve(base, n, m, 6);
    for (int i = lane; i < m * 6 + 2; i += 32) w.msg[i] = 0.5003 + 0.001 * i;
    for (int i = lane; i < n + 2; i += 32) { w.lr0[i] = 0.7 + 0.002 * i; w.est[i] = i & 1; }
    w.msg[c.dummy] = 1.0;
    __syncthreads();
    long long t[4] = { 0, 0, 0, 0 };
    for (int r = 0; r < reps; r++) {
        long long a = clock64();
        if (VAR == 0) chk_pass<6, 6, true, false>(c, w, lane); else chk_pass_b3<6, 6, true, false>(c, w, lane, z);
        __syncwarp();
        long long b = clock64();
        var_pass<3>(c, w, lane);
        __syncwarp();
        long long d = clock64();
        const bool bad = syn_nonzero<6>(c, w.est, lane);
        long long e = clock64();
        t[0] += b - a; t[1] += d - b; t[2] += e - d; t[3] += bad;
        for (int i = lane; i < m * 6; i += 32) w.msg[i] = 0.5003 + 0.001 * i;     // keep the values in range
        __syncwarp();
    }
    if (lane == 0 && blockIdx.x == 0) { cyc[wid * 4] = t[0]; cyc[wid * 4 + 1] = t[1]; cyc[wid * 4 + 2] = t[2]; cyc[wid * 4 + 3] = t[3]; }
}

// the same passes with NT-thread teams per codeword (NT / 32 warps share one codeword's rows and columns)
template <int NT>
__global__ void __launch_bounds__(256, 4) team_kernel(const uint4* rv2g, const uint4* vdg, long long* cyc, int reps, int n, int m, int nvr)
{
    extern __shared__ __align__(16) unsigned char sm[];
    const int team = threadIdx.x / NT, tid = threadIdx.x % NT, bar = team + 1;
    uint4* rv2 = (uint4*)sm; uint4* vd = rv2 + m;
    for (int i = threadIdx.x; i < m; i += blockDim.x) rv2[i] = rv2g[i];
    for (int i = threadIdx.x; i < nvr * 32; i += blockDim.x) vd[i] = vdg[i];
    __syncthreads();
    FastS c; c.n = n; c.m = m; c.k = 152; c.rank = 148; c.rs = 6; c.nvr = nvr; c.max_row = 6; c.max_col = 3; c.uniform_rows = 1; c.dummy = m * 6;
    c.o_rv2 = 0; c.o_vdesc = 16 * m; c.o_permute = 0; c.smb = sm;
    unsigned char* base = sm + 16 * (m + nvr * 32) + (size_t)team * fast_warp_bytes(n, m, 6);
    FastBuf w = fast_carve(base, n, m, 6);
    for (int i = tid; i < m * 6 + 2; i += NT) w.msg[i] = 0.5003 + 0.001 * i;
    for (int i = tid; i < n + 2; i += NT) { w.lr0[i] = 0.7 + 0.002 * i; w.est[i] = i & 1; }
    if (tid == 0) w.msg[c.dummy] = 1.0;
    __syncthreads();
    long long t[4] = { 0, 0, 0, 0 };
    for (int r = 0; r < reps; r++) {
        long long a = clock64();
        chk_pass<6, 6, true, false, NT>(c, w, tid);
        team_sync<NT>(bar);
        long long b = clock64();
        var_pass<3, NT>(c, w, tid);
        team_sync<NT>(bar);
        long long d = clock64();
        const bool bad = syn_nonzero<6, NT>(c, w.est, tid, bar);
        long long e = clock64();
        t[0] += b - a; t[1] += d - b; t[2] += e - d; t[3] += bad;
        for (int i = tid; i < m * 6; i += NT) w.msg[i] = 0.5003 + 0.001 * i;
        team_sync<NT>(bar);
    }
    if (tid == 0 && blockIdx.x == 0) { cyc[team * 4] = t[0]; cyc[team * 4 + 1] = t[1]; cyc[team * 4 + 2] = t[2]; cyc[team * 4 + 3] = t[3]; }
}

int main()
{
    const int n = 300, m = 148, nvr = 10;
    std::vector<uint32_t> rv2(m * 4), vd(nvr * 32 * 4);
    // a regular pseudo-random graph: row r touches variables (r*7 + j*53) % 300; variable v's slots follow
    std::vector<std::vector<int>> vs(n);
    for (int r = 0; r < m; r++) {
        uint16_t id[8];
        for (int j = 0; j < 8; j++) id[j] = n;
        for (int j = 0; j < 6; j++) { int v = (r * 7 + j * 53 + (r / 43)) % n; id[j] = v; if (vs[v].size() < 3) vs[v].push_back(r * 6 + j); }
        for (int j = 0; j < 4; j++) rv2[r * 4 + j] = id[2 * j] | (id[2 * j + 1] << 16);
    }
    for (int i = 0; i < nvr * 32; i++) {
        uint32_t s[4] = { (uint32_t)m * 6, (uint32_t)m * 6, (uint32_t)m * 6, (uint32_t)m * 6 };
        if (i < n) for (size_t j = 0; j < vs[i].size(); j++) s[j] = vs[i][j];
        vd[i * 4] = s[0] | (s[1] << 16); vd[i * 4 + 1] = s[2] | (s[3] << 16); vd[i * 4 + 2] = i < n ? i : n + 1; vd[i * 4 + 3] = 0;
    }
    uint4 *d_rv2, *d_vd; long long* d_c;
    cudaMalloc(&d_rv2, rv2.size() * 4); cudaMalloc(&d_vd, vd.size() * 4); cudaMalloc(&d_c, 64 * 4 * 8);
    cudaMemcpy(d_rv2, rv2.data(), rv2.size() * 4, cudaMemcpyHostToDevice); cudaMemcpy(d_vd, vd.data(), vd.size() * 4, cudaMemcpyHostToDevice);
    cudaFuncSetAttribute(bench_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    cudaFuncSetAttribute(bench_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    const int reps = 200;
    for (int var = 0; var < 2; var++)
    for (int W : { 1, 4, 8, 16, 20 }) {
        size_t smem = 16 * (m + nvr * 32) + (size_t)W * fast_warp_bytes(n, m, 6);
        if (smem > 220 * 1024) break;
        if (var == 0) bench_kernel<0><<<p.multiProcessorCount, 32 * W, smem>>>(d_rv2, d_vd, d_c, reps, n, m, nvr, 0);
        else bench_kernel<1><<<p.multiProcessorCount, 32 * W, smem>>>(d_rv2, d_vd, d_c, reps, n, m, nvr, 0);
        if (W == 1) printf(var ? "-- check pass in batches of three chains\n" : "-- check pass six chains at once\n");
        cudaError_t e = cudaDeviceSynchronize();
        long long c[4]; cudaMemcpy(c, d_c, 32, cudaMemcpyDeviceToHost);
        printf("W=%2d warps/SM: check pass %7.0f cycles, variable pass %7.0f, syndrome %6.0f  (per iteration %7.0f; per SM-iteration %6.0f) %s\n", W,
               (double)c[0] / reps, (double)c[1] / reps, (double)c[2] / reps, (double)(c[0] + c[1] + c[2]) / reps,
               (double)(c[0] + c[1] + c[2]) / reps / W, e == cudaSuccess ? "" : cudaGetErrorString(e));
    }
    cudaFuncSetAttribute(team_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    for (int blocks : { 4 }) {                       // 4 teams of 64 threads per CTA
        size_t smem = 16 * (m + nvr * 32) + (size_t)4 * fast_warp_bytes(n, m, 6);
        team_kernel<64><<<p.multiProcessorCount * blocks, 256, smem>>>(d_rv2, d_vd, d_c, reps, n, m, nvr);
        cudaError_t e = cudaDeviceSynchronize();
        long long c[4]; cudaMemcpy(c, d_c, 32, cudaMemcpyDeviceToHost);
        printf("teams of 64: %2d codewords/SM (%2d warps): check pass %7.0f cycles, variable pass %7.0f, syndrome %6.0f  (per iteration %7.0f; per SM-iteration %6.0f) %s\n",
               4 * blocks, 8 * blocks, (double)c[0] / reps, (double)c[1] / reps, (double)c[2] / reps, (double)(c[0] + c[1] + c[2]) / reps,
               (double)(c[0] + c[1] + c[2]) / reps / (4 * blocks), e == cudaSuccess ? "" : cudaGetErrorString(e));
    }
    return 0;
}
