// kernels_hdr.cu -- header branch of the receiver: the stream tags of a frame from its IQ alone.
//
// Replaces, one frame per work() call in the reference (python/dtl/ofdm_receiver.py:100-150), S*F frames per launch here:
//   digital.ofdm_chanest_vcvc(sync_word1, sync_word2, header_len)           (gr-digital; integer carrier offset, taps)
//   digital.ofdm_frame_equalizer_vcvc(dtl.ofdm_adaptive_header_equalizer)   lib/dtl/ofdm_adaptive_equalizer.cc:161-174,217-268
//   digital.ofdm_serializer_vcc + constellation_decoder_cb(BPSK)
//   packet_headerparser_b(ofdm_adaptive_packet_header::header_parser)        lib/dtl/ofdm_adaptive_packet_header.cc:232-312
// The FFT of the two sync symbols and the H header symbols is fft_fwd_kernel (same launcher as the payload symbols).
//
// header_kernel: one CTA per frame.  The (2+H) frequency-domain symbols are staged in shared memory once; the warps
// share the candidate carrier offsets of Schmidl & Cox's B(g) (one warp-reduced correlation each), every thread then owns
// carriers: taps = S2 / sync2, the decision-directed BPSK recurrence over the header symbols (same arithmetic contract
// as the payload equaliser: explicit IEEE single operations), hard decisions in serialiser order, and one thread
// unpacks the fields and checks the CRC-16.  header_fix_kernel walks each stream's frames in order for the parser's
// one piece of sequential state (the constellation it falls back to when the field is out of range).
#include "kernels.cuh"

namespace dtl {

__device__ __forceinline__ float2 h_div(float2 y, float2 h)
{
    const float den = __fmaf_rn(h.x, h.x, __fmul_rn(h.y, h.y));
    const float r = __frcp_rn(den);
    const float nr = __fmaf_rn(y.x, h.x, __fmul_rn(y.y, h.y));
    const float ni = __fmaf_rn(y.y, h.x, -__fmul_rn(y.x, h.y));
    return make_float2(__fmul_rn(nr, r), __fmul_rn(ni, r));
}
__device__ __forceinline__ float2 h_mul(float2 a, float2 b)
{
    return make_float2(__fmaf_rn(a.x, b.x, -__fmul_rn(a.y, b.y)), __fmaf_rn(a.x, b.y, __fmul_rn(a.y, b.x)));
}
__device__ __forceinline__ float2 h_conj(float2 a) { return make_float2(a.x, -a.y); }

__device__ __forceinline__ unsigned hdr_field(const uint8_t* bits, int nbits, int& k, int n)
{
    unsigned v = 0;
    for (int i = 0; i < n && k < nbits; i++, k++) v |= ((unsigned)bits[k] & 1u) << i;      // LSB first, one bit per item
    return v;
}

__global__ void __launch_bounds__(256) header_kernel(HdrParams p)
{
    extern __shared__ __align__(16) unsigned char hsm[];
    const HdrDev& hd = p.hd;
    const int N = hd.N, H = hd.H, T = blockDim.x, tid = threadIdx.x;
    const int lane = tid & 31, wid = tid >> 5, nwarps = T >> 5;
    const long long f = blockIdx.x;
    float2* sym = (float2*)hsm;                                // [2+H][N]: sync1, sync2, header symbols
    float2* taps = sym + (size_t)(2 + H) * N;                  // [N] channel state
    float2* tmp = taps + N;                                    // [H][N] header symbols after the carrier-offset shift
    float* cand = (float*)(tmp + (size_t)H * N);                           // [ncand] |B(g)|
    uint8_t* bits = (uint8_t*)(cand + ((N / 2 + 4) & ~3));     // [nbits]
    __shared__ int s_off;

    const float2* src = p.hfreq + f * (long long)(2 + H) * N;
    for (int i = tid; i < (2 + H) * N; i += T) sym[i] = __ldcs(&src[i]);
    __syncthreads();

    // ---- ofdm_chanest_vcvc::get_carr_offset: argmax over even g of |sum_k conj(S1[k+g]) conj(v[k]) S2[k+g]|
    const int ncand = (hd.max_pos - hd.max_neg) / 2 + 1;
    for (int c = wid; c < ncand; c += nwarps) {
        const int g = hd.max_neg + 2 * c;
        float2 acc = make_float2(0.f, 0.f);
        for (int k = lane; k < N; k += 32) {
            const float2 v = __ldg(&hd.corr_v[k]);
            if ((v.x != 0.f || v.y != 0.f) && k + g >= 0 && k + g < N) {
                const float2 t = h_mul(h_mul(h_conj(sym[k + g]), h_conj(v)), sym[N + k + g]);
                acc.x += t.x; acc.y += t.y;
            }
        }
#pragma unroll
        for (int o = 16; o; o >>= 1) { acc.x += __shfl_xor_sync(0xffffffffu, acc.x, o); acc.y += __shfl_xor_sync(0xffffffffu, acc.y, o); }
        if (lane == 0) cand[c] = __fsqrt_rn(__fadd_rn(__fmul_rn(acc.x, acc.x), __fmul_rn(acc.y, acc.y)));
    }
    __syncthreads();
    if (tid == 0) {
        int best = 0; float bg = 0.f;
        for (int c = 0; c < ncand; c++) if (cand[c] > bg) { bg = cand[c]; best = hd.max_neg + 2 * c; }
        s_off = best;
    }
    __syncthreads();
    const int off = s_off;

    // ---- get_chan_taps: taps[i - off] = S2[i] / sync2[i - off] where sync word 2 is non-zero
    {
        int lo = 0, hi = N;
        if (off > 0) lo = off; else if (off < 0) hi = N + off;
        for (int k = tid; k < N; k += T) {
            const float2 r = __ldg(&hd.ref[k]);
            const int i = k + off;
            float2 t = make_float2(0.f, 0.f);
            if (i >= lo && i < hi && (r.x != 0.f || r.y != 0.f)) t = h_div(sym[N + i], r);
            taps[k] = t;
        }
    }
    __syncthreads();

    // ---- ofdm_frame_equalizer_vcvc: integer-offset shift of the flattened header symbols, zero fill, per-symbol phase
    float2* hs = sym + (size_t)2 * N;                          // [H][N]
    if (off != 0 && off <= p.pl.max_off && off >= -p.pl.max_off) {
        const float2* ptab = p.pl.phase_tab + (size_t)(off + p.pl.max_off) * (p.pl.L + 1);
        const int total = H * N;
        for (int q = tid; q < total; q += T) {                 // out[q] = in[q + off] (zero outside), times the phase of its symbol
            const int j = q + off;
            float2 v = make_float2(0.f, 0.f);
            if (j >= 0 && j < total) v = hs[j];
            tmp[q] = h_mul(v, __ldg(&ptab[q / N]));
        }
        __syncthreads();
        hs = tmp;
    }

    // ---- frame_equalize with BPSK (ofdm_adaptive_equalizer.cc:217-268): serial over the header symbols, per carrier
    const float alpha = p.pl.alpha, oma = p.pl.oma;
    const float2* ipts = p.cn.inv_points + p.cn.off[DTL_CNST_BPSK];
    for (int k = tid; k < N; k += T) {
        float2 h = taps[k];
        for (int i = 0; i < H; i++) {
            const int kind = __ldg(&hd.kind[i * N + k]);
            if (!kind) continue;
            const float2 y = hs[i * N + k];
            const float2 se = h_div(y, h);
            float2 inv = kind == 2 ? __ldg(&hd.pinv[i * N + k]) : __ldg(&ipts[se.x > 0.f ? 1 : 0]);
            const float2 t = make_float2(__fmul_rn(oma, y.x), __fmul_rn(oma, y.y));
            const float2 u = h_mul(t, inv);
            h = make_float2(__fmaf_rn(alpha, h.x, u.x), __fmaf_rn(alpha, h.y, u.y));
            hs[i * N + k] = se;                                 // the decision below only needs the sign of Re
        }
        if (off != 0 && off <= p.pl.max_off && off >= -p.pl.max_off) h = h_mul(h, __ldg(&hd.rot[off + p.pl.max_off]));
        if (p.taps) __stcs(&p.taps[f * N + k], h);              // "ofdm_sync_chan_taps" of the payload branch
    }
    __syncthreads();

    // ---- ofdm_serializer_vcc + constellation_decoder_cb(BPSK)
    for (int j = tid; j < hd.nbits; j += T) bits[j] = hs[__ldg(&hd.ser[j])].x > 0.f ? 1 : 0;
    __syncthreads();

    // ---- header_parser (ofdm_adaptive_packet_header.cc:232-312): fields LSB first, CRC-16 over the packed bits
    if (tid == 0) {
        const int nbits = hd.nbits;
        int k = 0;
        dtl_frame_meta m;
        m.carrier_offset = off; m.reserved = 0; m.fec_id = 0; m.tb_no = 0; m.tb_offset = 0; m.tb_payload = 0; m.packet_len = 0;
        m.payload_len = (int)hdr_field(bits, nbits, k, 12);
        m.frame_no = (int)hdr_field(bits, nbits, k, 12);
        m.constellation = (int)hdr_field(bits, nbits, k, 4);     // raw field: header_fix_kernel applies the fallback
        const int fcnst = (int)hdr_field(bits, nbits, k, 4);
        int ffec = 0;
        if (hd.has_fec) {
            m.tb_no = (int)hdr_field(bits, nbits, k, 12); ffec = (int)hdr_field(bits, nbits, k, 4);
            m.tb_offset = (int)hdr_field(bits, nbits, k, 12); m.fec_id = (int)hdr_field(bits, nbits, k, 4);
            m.tb_payload = (int)hdr_field(bits, nbits, k, 16);
        }
        // pack_crc (:99-111, with the reference's `i*len + j < header_len` bound), crc(16, 0x1021, 0xFFFF, 0, false, true)
        int len = (nbits - 16) / 8;
        if ((nbits - 16) % 8) ++len;
        unsigned crc = 0xFFFFu;
        for (int i = 0; i < len; i++) {
            unsigned byte = 0;
            for (int j = 0; j < 8 && i * len + j < nbits; j++) byte = ((byte << 1) | (bits[i * 8 + j] & 1u)) & 0xffu;
            crc ^= byte << 8;
            for (int b = 0; b < 8; b++) crc = ((crc & 0x8000u) ? ((crc << 1) ^ 0x1021u) : (crc << 1)) & 0xFFFFu;
        }
        unsigned calc = 0;
        for (int i = 0; i < 16; i++) if ((crc >> i) & 1u) calc |= 1u << (15 - i);
        int ok = 1;
        for (int i = 0; i < 16 && k < nbits; i++, k++)
            if (((unsigned)bits[k] & 1u) != ((calc >> i) & 1u)) ok = 0;
        if (!ok) { m.payload_len = 0; m.frame_no = 0; m.constellation = 0; m.fec_id = 0; m.tb_no = 0; m.tb_offset = 0; m.tb_payload = 0; m.reserved = 1; }
        p.meta[f] = m;
        if (p.hres) { dtl_header_result r; r.crc_ok = ok; r.feedback_cnst = ok ? fcnst : 0; r.feedback_fec = ok ? ffec : 0; r.reserved = 0; p.hres[f] = r; }
    }
}

// The parser keeps the last valid constellation (ofdm_adaptive_packet_header.cc:287-289: a field of 0 or above the
// highest known constellation leaves it unchanged) and derives packet_len from it (:296-298): sequential per stream.
__global__ void header_fix_kernel(HdrParams p)
{
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= p.n_streams) return;
    int cur = p.st[s].hdr_cnst;
    for (int j = 0; j < p.F; j++) {
        dtl_frame_meta& m = p.meta[(long long)s * p.F + j];
        if (m.reserved & 1) continue;
        const int c = m.constellation;
        if (c >= 1 && c <= DTL_CNST_MAX) cur = c;
        const int bps = p.cn.bps[cur];
        m.constellation = cur;
        m.packet_len = (m.payload_len * 8 + bps - 1) / bps;
    }
    p.st[s].hdr_cnst = cur;
}

cudaError_t launch_header(cudaStream_t st, const HdrParams& p)
{
    if (p.n_frames <= 0) return cudaSuccess;
    const int N = p.hd.N, H = p.hd.H;
    const int threads = N >= 256 ? 256 : (N < 32 ? 32 : N);
    const size_t smem = (size_t)(2 + H) * N * 8 + (size_t)(1 + H) * N * 8 + (size_t)((N / 2 + 4) & ~3) * 4 + (size_t)((p.hd.nbits + 15) & ~15);
    if (smem > 200 * 1024 || p.n_frames > 0x7fffffffLL) return cudaErrorInvalidConfiguration;
    static size_t attr = 0;
    if (smem > 48 * 1024 && smem > attr) { cudaFuncSetAttribute(header_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); attr = smem; }
    header_kernel<<<(unsigned)p.n_frames, threads, smem, st>>>(p);
    header_fix_kernel<<<(p.n_streams + 127) / 128, 128, 0, st>>>(p);
    return cudaGetLastError();
}

} // namespace dtl
