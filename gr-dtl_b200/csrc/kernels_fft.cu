// kernels_fft.cu -- batched small-size FFT kernels (stage 1 and stage 5).
//
// Replaces, for many frames of many streams per launch:
//   digital.header_payload_demux(output_symbols=True)  CP strip        python/dtl/ofdm_receiver.py:79-86
//   fft.fft_vcc(fft_len, True, (), True)               forward, shift  python/dtl/ofdm_receiver.py:153
//   digital.ofdm_carrier_allocator_cvc + fft.fft_vcc(fft_len, False, (), True) + digital.ofdm_cyclic_prefixer
//                                                                      python/dtl/ofdm_transmitter.py:168-189
// Both directions are unnormalised; "shift" puts DC at index N/2 of the frequency-domain vector.
//
// Design: one OFDM symbol per group of N/4 threads, Stockham autosort radix-4 stages (one radix-2 stage
// when log2 N is odd) ping-ponging between two shared-memory buffers, twiddles from a global table
// computed in double on the host.  A CTA of 256 threads holds 256/(N/4) symbols (16 for N=64), so the
// global loads/stores of a warp cover whole 128-byte lines; the cyclic prefix is dropped (RX) or
// duplicated (TX) in the load/store indexing, never materialised.  HBM-bound: 8*(2N+cp) bytes/symbol.
#include "kernels.cuh"

namespace dtl {

__device__ __forceinline__ float2 cmul(float2 a, float2 b) { return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }

// One FFT of size N in shared memory by T = N/4 cooperating threads (t = 0..T-1).  `a` holds the input
// in natural order; returns the buffer that holds the result.  dir = -1 forward, +1 inverse: the table is
// tw[j] = exp(-2*pi*i*j/N) and the inverse conjugates it.  All groups of the CTA call this together.
__device__ float2* fft_smem(float2* a, float2* b, int N, int log2n, int t, const float2* __restrict__ tw, bool inverse)
{
    const int T = N >> 2;
    int Ns = 1;
    int stages4 = log2n >> 1;
    if (log2n & 1) {
        // radix-2 first stage (Ns = 1): out[2j + r] = in[j] +- in[j + N/2]; two butterflies per thread
        for (int j = t; j < (N >> 1); j += T) {
            float2 u = a[j], v = a[j + (N >> 1)];
            b[2 * j] = make_float2(u.x + v.x, u.y + v.y);
            b[2 * j + 1] = make_float2(u.x - v.x, u.y - v.y);
        }
        __syncthreads();
        float2* s = a; a = b; b = s;
        Ns = 2;
    }
    for (int st = 0; st < stages4; st++) {
        const int j = t, k = j & (Ns - 1);
        const int tstep = (N / (Ns * 4)) * k;          // twiddle index of w^k for this stage
        float2 v0 = a[j], v1 = a[j + T], v2 = a[j + 2 * T], v3 = a[j + 3 * T];
        if (k) {
            float2 w1 = __ldg(&tw[tstep]), w2 = __ldg(&tw[2 * tstep]), w3 = __ldg(&tw[3 * tstep]);
            if (inverse) { w1.y = -w1.y; w2.y = -w2.y; w3.y = -w3.y; }
            v1 = cmul(v1, w1); v2 = cmul(v2, w2); v3 = cmul(v3, w3);
        }
        const float2 s02 = make_float2(v0.x + v2.x, v0.y + v2.y), d02 = make_float2(v0.x - v2.x, v0.y - v2.y);
        const float2 s13 = make_float2(v1.x + v3.x, v1.y + v3.y), d13 = make_float2(v1.x - v3.x, v1.y - v3.y);
        // forward: multiply d13 by -i ; inverse: by +i
        const float2 r13 = inverse ? make_float2(-d13.y, d13.x) : make_float2(d13.y, -d13.x);
        const int j0 = ((j - k) << 2) + k;
        b[j0] = make_float2(s02.x + s13.x, s02.y + s13.y);
        b[j0 + Ns] = make_float2(d02.x + r13.x, d02.y + r13.y);
        b[j0 + 2 * Ns] = make_float2(s02.x - s13.x, s02.y - s13.y);
        b[j0 + 3 * Ns] = make_float2(d02.x - r13.x, d02.y - r13.y);
        __syncthreads();
        float2* s = a; a = b; b = s;
        Ns <<= 2;
    }
    return a;
}

// Forward shifted FFT of `n_sym` symbols.  Symbol q = frame * syms_per_frame + i is read from
//   in + frame * in_frame_stride + (in_sym0 + i) * in_sym_stride + in_skip   (N samples)
// and written to out + q * N.  Plain fft_vcc: syms_per_frame = 1, strides N, skip 0.
__global__ void __launch_bounds__(256) fft_fwd_kernel(const float2* __restrict__ in, float2* __restrict__ out,
                                                        long long n_sym, int syms_per_frame, long long in_frame_stride,
                                                        int in_sym0, int in_sym_stride, int in_skip, int N, int log2n,
                                                        const float2* __restrict__ tw)
{
    extern __shared__ float2 sm[];
    const int T = N >> 2, G = blockDim.x / T;          // threads per symbol, symbols per CTA
    const int g = threadIdx.x / T, t = threadIdx.x - g * T;
    float2* a = sm + (size_t)g * 2 * N;
    float2* b = a + N;
    const long long q = (long long)blockIdx.x * G + g;
    const bool live = q < n_sym;
    if (live) {
        const long long fr = q / syms_per_frame; const int i = (int)(q - fr * syms_per_frame);
        const float2* src = in + fr * in_frame_stride + (long long)(in_sym0 + i) * in_sym_stride + in_skip;
#pragma unroll
        for (int r = 0; r < 4; r++) a[t + r * T] = __ldcs(&src[t + r * T]);
    }
    __syncthreads();
    float2* res = fft_smem(a, b, N, log2n, t, tw, false);
    if (live) {
        float2* dst = out + q * N;
        const int h = N >> 1;
#pragma unroll
        for (int r = 0; r < 4; r++) { const int o = t + r * T; dst[o] = res[(o + h) & (N - 1)]; }   // out[(i+N/2)%N] = X[i]
    }
}

// ---- N = 64 (the reference's fft_len): 8 x 8 decomposition, 8 lanes per symbol, 8 points per lane ----------------------
// n = n0 + 8 n1, k = ka + 8 kb:  X[ka + 8 kb] = sum_n0 W8^(n0 kb) . W64^(n0 ka) . sum_n1 x[n0 + 8 n1] W8^(n1 ka).
// Lane n0 does the inner 8-point DFT in registers on global loads 64 bytes apart, applies W64^(n0 ka), and the 8 lanes of the
// symbol transpose through shared memory ONCE (one store + one load per point, __syncwarp only: a symbol never leaves
// its warp); lane ka then does the outer 8-point DFT and stores X[ka + 8 kb] -- 64 contiguous bytes per symbol and
// instruction on both sides.  No CTA barrier, no staging of input or output, twiddles live in registers across the
// grid-stride loop.
__global__ void __launch_bounds__(256) fft64_fwd_kernel(const float2* __restrict__ in, float2* __restrict__ out,
                                                          unsigned n_sym, unsigned syms_per_frame, long long in_frame_stride,
                                                          int in_sym0, int in_sym_stride, int in_skip,
                                                          const float2* __restrict__ tw)
{
    __shared__ float2 xch[8][4 * 64];                          // per warp: 4 symbols
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int s = lane >> 3, t = lane & 7;                     // symbol within the warp, lane within the symbol
    float2* xs = xch[w] + s * 64;
    const int flip = (s & 1) << 3;                             // odd symbols use the other 64-byte half: no bank conflicts
    float2 wt[8];
#pragma unroll
    for (int ka = 1; ka < 8; ka++) wt[ka] = __ldg(&tw[t * ka]);               // W64^(t ka)
    const unsigned per_cta = 32, step = gridDim.x * per_cta;
    for (unsigned q0 = blockIdx.x * per_cta; q0 < n_sym; q0 += step) {
        const unsigned q = q0 + w * 4 + s;
        const bool live = q < n_sym;
        float2 x[8];
        if (live) {
            const unsigned fr = q / syms_per_frame, i = q - fr * syms_per_frame;
            const float2* src = in + fr * in_frame_stride + (long long)(in_sym0 + (int)i) * in_sym_stride + in_skip;
#pragma unroll
            for (int r = 0; r < 8; r++) x[r] = __ldcs(&src[t + 8 * r]);
        } else {
#pragma unroll
            for (int r = 0; r < 8; r++) x[r] = make_float2(0.f, 0.f);
        }
        dft8_fwd(x);
#pragma unroll
        for (int ka = 0; ka < 8; ka++) {
            const float2 v = ka ? cmul(x[ka], wt[ka]) : x[0];
            xs[(ka * 8 + (t ^ ka)) ^ flip] = v;
        }
        __syncwarp();
#pragma unroll
        for (int n0 = 0; n0 < 8; n0++) x[n0] = xs[(t * 8 + (n0 ^ t)) ^ flip];
        __syncwarp();
        dft8_fwd(x);
        if (live) {
            float2* dst = out + (size_t)q * 64;
#pragma unroll
            for (int kb = 0; kb < 8; kb++) dst[t + 8 * (kb ^ 4)] = x[kb];       // out[(k + 32) % 64] = X[k], k = t + 8 kb
        }
    }
}

// Inverse shifted FFT, plain vectors (fft_vcc(N, False, (), True)): a[i] = in[(i+N/2)%N].
__global__ void __launch_bounds__(256) fft_inv_kernel(const float2* __restrict__ in, float2* __restrict__ out,
                                                        long long n_sym, int N, int log2n, const float2* __restrict__ tw)
{
    extern __shared__ float2 sm[];
    const int T = N >> 2, G = blockDim.x / T;
    const int g = threadIdx.x / T, t = threadIdx.x - g * T;
    float2* a = sm + (size_t)g * 2 * N;
    float2* b = a + N;
    const long long q = (long long)blockIdx.x * G + g;
    const bool live = q < n_sym;
    const int h = N >> 1;
    if (live) {
        const float2* src = in + q * N;
#pragma unroll
        for (int r = 0; r < 4; r++) { const int o = t + r * T; a[(o + h) & (N - 1)] = __ldcs(&src[o]); }
    }
    __syncthreads();
    float2* res = fft_smem(a, b, N, log2n, t, tw, true);
    if (live) {
        float2* dst = out + q * N;
#pragma unroll
        for (int r = 0; r < 4; r++) dst[t + r * T] = res[t + r * T];
    }
}

// Stage 5: per OFDM symbol of a frame: sync words, then header/payload symbols mapped onto the occupied
// carriers + pilots, inverse shifted FFT, cyclic prefix prepended.  Symbol index within frame:
// 0,1 = sync words; 2..2+H-1 header (BPSK chunks); then L payload symbols (constellation of the frame).
__global__ void __launch_bounds__(256) tx_mod_kernel(TxParams p)
{
    extern __shared__ float2 sm[];
    const int N = p.N, T = N >> 2, G = blockDim.x / T;
    const int g = threadIdx.x / T, t = threadIdx.x - g * T;
    float2* a = sm + (size_t)g * 2 * N;
    float2* b = a + N;
    const int spf = 2 + p.H + p.L;
    const long long q = (long long)blockIdx.x * G + g;
    const bool live = q < p.n_frames * spf;
    const int h = N >> 1;
    long long fr = 0; int i = 0;
    if (live) {
        fr = q / spf; i = (int)(q - fr * spf);
        if (i < 2) {
            const float2* sw = i == 0 ? p.sync1 : p.sync2;
#pragma unroll
            for (int r = 0; r < 4; r++) { const int o = t + r * T; a[(o + h) & (N - 1)] = sw[o]; }
        } else {
            const int s = i - 2;                           // data-symbol index (header first)
            const int oset = s % p.nocc;
            const int cnst = s < p.H ? DTL_CNST_BPSK : p.meta[fr].constellation;
            const int bps = p.cn.bps[cnst];
            const float2* pts = p.cn.points + p.cn.off[cnst];
#pragma unroll
            for (int r = 0; r < 4; r++) {
                const int o = t + r * T;                   // shifted bin
                float2 v = make_float2(0.f, 0.f);
                const int pos = p.ser_pos[oset * N + o];
                if (pos >= 0) {
                    unsigned c;
                    if (s < p.H) c = p.hdr_bits[fr * p.hdr_stride + p.tx_base[s] + pos];
                    else c = p.chunks[fr * p.cap + (p.tx_base[s] - p.tx_base[p.H]) + pos];
                    v = __ldg(&pts[c & ((1u << bps) - 1u)]);
                }
                if (p.tx_pil_mask[s * N + o]) v = p.tx_pil_val[s * N + o];
                a[(o + h) & (N - 1)] = v;
            }
        }
    }
    __syncthreads();
    float2* res = fft_smem(a, b, N, log2n_of(N), t, p.tw, true);
    if (live) {
        float2* dst = p.iq + fr * (long long)spf * (N + p.cp) + (long long)i * (N + p.cp);
        for (int o = t; o < N + p.cp; o += T) dst[o] = res[(o + N - p.cp) & (N - 1)];
    }
}

// ---- N = 1024 (SURVEY.md section 8d configuration 5): 32 x 32 decomposition, ONE WARP per symbol, 32 points per lane ------
// n = n0 + 32 n1, k = k1 + 32 k2:  X[k1 + 32 k2] = sum_n0 W32^(n0 k2) . W1024^(n0 k1) . sum_n1 x[n0 + 32 n1] W32^(n1 k1).
// Lane n0 loads x[n0 + 32 n1] (every load instruction of the warp is one contiguous 256-byte run), does the inner 32-point
// DFT in registers (radix-2 decimation in frequency, fully unrolled: twiddles are immediates, the bit-reversed result order
// is resolved at compile time), applies W1024^(n0 k1) from the table, and the warp transposes through its own 8.25 KB of
// shared memory ONCE (row stride 33: both sides conflict-free; __syncwarp only); lane k1 then does the outer 32-point DFT
// and stores X[k1 + 32 k2] -- again 256 contiguous bytes per instruction.  No CTA barrier, no staging of input or output:
// HBM traffic is the algorithmic 16 N bytes per symbol.
__device__ __forceinline__ constexpr float cos32(int k)             // cos(2 pi k / 32), k = 0 .. 16
{
    constexpr float t[17] = { 1.0f, 0.98078528040323043f, 0.92387953251128674f, 0.83146961230254524f, 0.70710678118654752f,
                              0.55557023301960218f, 0.38268343236508977f, 0.19509032201612825f, 0.0f, -0.19509032201612825f,
                              -0.38268343236508977f, -0.55557023301960218f, -0.70710678118654752f, -0.83146961230254524f,
                              -0.92387953251128674f, -0.98078528040323043f, -1.0f };
    return t[k];
}
__device__ __forceinline__ constexpr float sin32(int k) { return k <= 8 ? cos32(8 - k) : cos32(k - 8); }      // sin(2 pi k / 32), k = 0 .. 16
// d * W32^k (forward: exp(-2 pi i k / 32)), k a compile-time constant after unrolling
__device__ __forceinline__ float2 mul_w32(float2 d, int k)
{
    if (k == 0) return d;
    if (k == 8) return make_float2(d.y, -d.x);
    const float c = cos32(k), sn = sin32(k);
    return make_float2(d.x * c + d.y * sn, d.y * c - d.x * sn);
}
__device__ __forceinline__ constexpr int bitrev5(int i) { return ((i & 1) << 4) | ((i & 2) << 2) | (i & 4) | ((i & 8) >> 2) | ((i & 16) >> 4); }
// in place, natural order in, X[bitrev5(i)] in x[i] out
__device__ __forceinline__ void dft32_fwd(float2 (&x)[32])
{
#pragma unroll
    for (int half = 16; half >= 1; half >>= 1) {
#pragma unroll
        for (int base = 0; base < 32; base += 2 * half) {
#pragma unroll
            for (int j = 0; j < half; j++) {
                const float2 a = x[base + j], b = x[base + j + half];
                x[base + j] = make_float2(a.x + b.x, a.y + b.y);
                x[base + j + half] = mul_w32(make_float2(a.x - b.x, a.y - b.y), j * (16 / half));
            }
        }
    }
}

#define DTL_FFT1K_WARPS 4
#define DTL_FFT1K_ROW 33
__global__ void __launch_bounds__(DTL_FFT1K_WARPS * 32, 4) fft1024_fwd_kernel(const float2* __restrict__ in, float2* __restrict__ out,
                                                                             unsigned n_sym, unsigned syms_per_frame, long long in_frame_stride,
                                                                             int in_sym0, int in_sym_stride, int in_skip,
                                                                             const float2* __restrict__ tw)
{
    extern __shared__ float2 sm[];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    float2* xs = sm + (size_t)w * 32 * DTL_FFT1K_ROW;
    const unsigned step = gridDim.x * DTL_FFT1K_WARPS;
    for (unsigned q = blockIdx.x * DTL_FFT1K_WARPS + w; q < n_sym; q += step) {
        const unsigned fr = q / syms_per_frame, i = q - fr * syms_per_frame;
        const float2* src = in + fr * in_frame_stride + (long long)(in_sym0 + (int)i) * in_sym_stride + in_skip;
        float2 x[32];
#pragma unroll
        for (int r = 0; r < 32; r++) x[r] = __ldcs(&src[lane + 32 * r]);
        dft32_fwd(x);
#pragma unroll
        for (int r = 0; r < 32; r++) {                          // x[r] = A[k1], k1 = bitrev5(r): twiddle W1024^(lane k1), row k1, column lane
            const int k1 = bitrev5(r);
            const float2 v = k1 ? cmul(x[r], __ldg(&tw[lane * k1])) : x[r];
            xs[k1 * DTL_FFT1K_ROW + lane] = v;
        }
        __syncwarp();
#pragma unroll
        for (int n0 = 0; n0 < 32; n0++) x[n0] = xs[lane * DTL_FFT1K_ROW + n0];
        __syncwarp();
        dft32_fwd(x);
        float2* dst = out + (size_t)q * 1024;
#pragma unroll
        for (int r = 0; r < 32; r++) {                          // x[r] = X[lane + 32 k2], k2 = bitrev5(r); out[(k + 512) % 1024] = X[k]
            const int k2 = bitrev5(r);
            dst[lane + 32 * ((k2 + 16) & 31)] = x[r];
        }
    }
}

// ---- launchers -------------------------------------------------------------------------------------------
static inline int ilog2(int n) { int l = 0; while ((1 << l) < n) l++; return l; }

cudaError_t launch_fft_fwd(cudaStream_t st, const float2* in, float2* out, long long n_sym, int syms_per_frame,
                           long long in_frame_stride, int in_sym0, int in_sym_stride, int in_skip, int N, const float2* tw)
{
    if (n_sym <= 0) return cudaSuccess;
    if (N == 64 && n_sym < 0x7fffffffLL && syms_per_frame > 0) {
        static int sms = 0;
        if (!sms) { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev); }
        const long long need = (n_sym + 31) / 32, cap = (long long)sms * 4 * 4;       // 4 resident CTAs per SM (59 registers), 4 waves
        fft64_fwd_kernel<<<(unsigned)(need < cap ? need : cap), 256, 0, st>>>(in, out, (unsigned)n_sym, (unsigned)syms_per_frame,
                                                                           in_frame_stride, in_sym0, in_sym_stride, in_skip, tw);
        return cudaGetLastError();
    }
    if (N == 1024 && n_sym < 0x7fffffffLL && syms_per_frame > 0) {
        static int sms = 0;
        if (!sms) { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev); }
        const long long need = (n_sym + DTL_FFT1K_WARPS - 1) / DTL_FFT1K_WARPS, cap = (long long)sms * 4 * 8;   // 4 resident CTAs per SM, 8 waves
        const size_t smem = (size_t)DTL_FFT1K_WARPS * 32 * DTL_FFT1K_ROW * sizeof(float2);
        fft1024_fwd_kernel<<<(unsigned)(need < cap ? need : cap), DTL_FFT1K_WARPS * 32, smem, st>>>(
            in, out, (unsigned)n_sym, (unsigned)syms_per_frame, in_frame_stride, in_sym0, in_sym_stride, in_skip, tw);
        return cudaGetLastError();
    }
    const int T = N / 4, threads = T >= 256 ? T : 256, G = threads / T;
    const long long blocks = (n_sym + G - 1) / G;
    const size_t smem = (size_t)G * 2 * N * sizeof(float2);
    fft_fwd_kernel<<<(unsigned)blocks, threads, smem, st>>>(in, out, n_sym, syms_per_frame, in_frame_stride, in_sym0,
                                                          in_sym_stride, in_skip, N, ilog2(N), tw);
    return cudaGetLastError();
}

cudaError_t launch_fft_inv(cudaStream_t st, const float2* in, float2* out, long long n_sym, int N, const float2* tw)
{
    if (n_sym <= 0) return cudaSuccess;
    const int T = N / 4, threads = T >= 256 ? T : 256, G = threads / T;
    const long long blocks = (n_sym + G - 1) / G;
    const size_t smem = (size_t)G * 2 * N * sizeof(float2);
    fft_inv_kernel<<<(unsigned)blocks, threads, smem, st>>>(in, out, n_sym, N, ilog2(N), tw);
    return cudaGetLastError();
}

cudaError_t launch_tx_mod(cudaStream_t st, const TxParams& p)
{
    if (p.n_frames <= 0) return cudaSuccess;
    const int N = p.N, T = N / 4, threads = T >= 256 ? T : 256, G = threads / T;
    const long long n_sym = p.n_frames * (2 + p.H + p.L);
    const long long blocks = (n_sym + G - 1) / G;
    const size_t smem = (size_t)G * 2 * N * sizeof(float2);
    tx_mod_kernel<<<(unsigned)blocks, threads, smem, st>>>(p);
    return cudaGetLastError();
}

cudaError_t init_kernels_fft()
{
    cudaError_t e = set_max_dyn_smem(fft_fwd_kernel);
    if (e == cudaSuccess) e = set_max_dyn_smem(fft_inv_kernel);
    if (e == cudaSuccess) e = set_max_dyn_smem(tx_mod_kernel);
    if (e == cudaSuccess) e = set_max_dyn_smem(fft1024_fwd_kernel);
    return e;
}

} // namespace dtl
