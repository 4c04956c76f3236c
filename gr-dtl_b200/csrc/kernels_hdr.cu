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

// Field i of the header (ofdm_adaptive_packet_header.cc:232-286), LSB first, one bit per item; the 16 CRC bits follow the
// last field.  Without FEC: payload_len 12, frame_no 12, constellation 4, feedback constellation 4.  With FEC also
// tb_no 12, feedback FEC 4, tb_offset 12, fec_id 4, tb_payload 16.
__device__ __forceinline__ unsigned hdr_field(const uint8_t* bits, int nbits, int start, int n, int* got)
{
    unsigned v = 0;
    int i = 0;
    for (; i < n && start + i < nbits; i++) v |= ((unsigned)bits[start + i] & 1u) << i;
    *got = i;
    return v;
}

// G frames per CTA, TPF = blockDim.x / G threads per frame (all barriers are CTA-wide: every group runs the same steps).
__global__ void __launch_bounds__(256) header_kernel(HdrParams p, int G, int group_bytes)
{
    extern __shared__ __align__(16) unsigned char hsm[];
    __shared__ int s_off[8];
    const HdrDev& hd = p.hd;
    const int N = hd.N, H = hd.H, T = blockDim.x / G;
    const int grp = threadIdx.x / T, tid = threadIdx.x - grp * T;
    const int lane = tid & 31, wid = tid >> 5, nwarps = T >> 5;
    const long long f = (long long)blockIdx.x * G + grp;
    const bool valid = f < p.n_frames;
    float2* sym = (float2*)(hsm + (size_t)grp * group_bytes);  // [2+H][N]: sync1, sync2, header symbols
    float2* taps = sym + (size_t)(2 + H) * N;                  // [N] channel state
    float2* tmp = taps + N;                                    // [H][N] header symbols after the carrier-offset shift
    float* cand = (float*)(tmp + (size_t)H * N);               // [ncand] |B(g)|
    uint8_t* bits = (uint8_t*)(cand + ((N / 2 + 4) & ~3));     // [nbits]

    if (valid) {
        const float2* src = p.hfreq + f * (long long)(2 + H) * N;
        for (int i = tid; i < (2 + H) * N; i += T) sym[i] = __ldcs(&src[i]);
    }
    __syncthreads();

    // ---- ofdm_chanest_vcvc::get_carr_offset: argmax over even g of |sum_k conj(S1[k+g]) conj(v[k]) S2[k+g]|
    const int ncand = (hd.max_pos - hd.max_neg) / 2 + 1;
    if (valid)
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
    if (valid && wid == 0) {                                   // first maximum wins, as the serial scan's `>` does
        float bg = 0.f; int bc = -1;
        for (int c = lane; c < ncand; c += 32) if (cand[c] > bg) { bg = cand[c]; bc = c; }
#pragma unroll
        for (int o = 16; o; o >>= 1) {
            const float og = __shfl_xor_sync(0xffffffffu, bg, o);
            const int oc = __shfl_xor_sync(0xffffffffu, bc, o);
            if (oc >= 0 && (og > bg || (og == bg && (bc < 0 || oc < bc)))) { bg = og; bc = oc; }
        }
        if (lane == 0) s_off[grp] = bc >= 0 ? hd.max_neg + 2 * bc : 0;
    }
    __syncthreads();
    const int off = valid ? s_off[grp] : 0;
    const bool shifted = off != 0 && off <= p.pl.max_off && off >= -p.pl.max_off;

    // ---- get_chan_taps: taps[i - off] = S2[i] / sync2[i - off] where sync word 2 is non-zero
    float2* hs = sym + (size_t)2 * N;                          // [H][N]
    if (valid) {
        int lo = 0, hi = N;
        if (off > 0) lo = off; else if (off < 0) hi = N + off;
        for (int k = tid; k < N; k += T) {
            const float2 r = __ldg(&hd.ref[k]);
            const int i = k + off;
            float2 t = make_float2(0.f, 0.f);
            if (i >= lo && i < hi && (r.x != 0.f || r.y != 0.f)) t = h_div(sym[N + i], r);
            taps[k] = t;
        }
        // ---- ofdm_frame_equalizer_vcvc: integer-offset shift of the flattened header symbols, zero fill, per-symbol phase
        if (shifted) {
            const float2* ptab = hd.hphase + (size_t)(off + p.pl.max_off) * H;
            const int total = H * N;
            for (int q = tid; q < total; q += T) {             // out[q] = in[q + off] (zero outside), times the phase of its symbol
                const int j = q + off;
                float2 v = make_float2(0.f, 0.f);
                if (j >= 0 && j < total) v = hs[j];
                tmp[q] = h_mul(v, __ldg(&ptab[q / N]));
            }
        }
    }
    __syncthreads();
    if (shifted) hs = tmp;

    // ---- frame_equalize with BPSK (ofdm_adaptive_equalizer.cc:217-268): serial over the header symbols, per carrier
    if (valid) {
        const float alpha = p.pl.alpha, oma = p.pl.oma;
        const float2* ipts = p.cn.inv_points + p.cn.off[DTL_CNST_BPSK];
        const float2 ip0 = __ldg(&ipts[0]), ip1 = __ldg(&ipts[1]);
        for (int k = tid; k < N; k += T) {
            float2 h = taps[k];
            for (int i = 0; i < H; i++) {
                const int kind = __ldg(&hd.kind[i * N + k]);
                if (!kind) continue;
                const float2 y = hs[i * N + k];
                const float2 se = h_div(y, h);
                const float2 inv = kind == 2 ? __ldg(&hd.pinv[i * N + k]) : (se.x > 0.f ? ip1 : ip0);
                const float2 t = make_float2(__fmul_rn(oma, y.x), __fmul_rn(oma, y.y));
                const float2 u = h_mul(t, inv);
                h = make_float2(__fmaf_rn(alpha, h.x, u.x), __fmaf_rn(alpha, h.y, u.y));
                hs[i * N + k] = se;                             // the decision below only needs the sign of Re
            }
            if (shifted) h = h_mul(h, __ldg(&hd.rot[off + p.pl.max_off]));
            if (p.taps) __stcs(&p.taps[f * N + k], h);          // "ofdm_sync_chan_taps" of the payload branch
        }
    }
    __syncthreads();

    // ---- ofdm_serializer_vcc + constellation_decoder_cb(BPSK)
    if (valid)
        for (int j = tid; j < hd.nbits; j += T) bits[j] = hs[__ldg(&hd.ser[j])].x > 0.f ? 1 : 0;
    __syncthreads();

    // ---- header_parser (ofdm_adaptive_packet_header.cc:232-312) by the first warp of the group: one field per lane, the
    // CRC-16 as the XOR of per-bit contributions (the register update is linear over GF(2); crc_tab[j] = what header bit j
    // adds to the reflected CRC of pack_crc's bytes, reference bound `i*len + j < header_len` included, built in api.cc)
    if (valid && wid == 0) {
        const int nbits = hd.nbits, nf = hd.has_fec ? 9 : 4;
        unsigned part = 0;
        for (int j = lane; j < hd.crc_bits; j += 32) if (bits[j] & 1u) part ^= __ldg(&hd.crc_tab[j]);
        const unsigned calc = (__reduce_xor_sync(0xffffffffu, part) ^ hd.crc0) & 0xffffu;
        const int fw = lane < 4 ? ((0x44CC >> (4 * lane)) & 15)                      // 12,12,4,4
                     : lane < nf ? ((0x104C4C >> (4 * (lane - 4))) & 15) + ((lane == 8) ? 16 : 0)   // 12,4,12,4,16
                     : 16;                                                            // lane nf: the received CRC
        // start of field `lane`: prefix sum of the widths of the lanes below (lanes above nf hold nothing)
        int start = fw;
#pragma unroll
        for (int o = 1; o < 16; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, start, o); if (lane >= o) start += t; }
        start -= fw;
        int got = 0;
        unsigned v = 0;
        if (lane <= nf) v = hdr_field(bits, nbits, start, fw, &got);
        const unsigned want = got >= 16 ? calc : (calc & ((1u << got) - 1u));       // compared while k < header_len only
        const int ok = __shfl_sync(0xffffffffu, (int)(v == want), nf) && (!p.valid || p.valid[f]);
        // dtl_frame_meta word j <- field: constellation 2, carrier_offset -, frame_no 1, payload_len 0, packet_len -,
        // fec_id 7, tb_no 4, tb_offset 6, tb_payload 8, reserved -   (raw constellation: header_fix_kernel applies the fallback)
        const int srcmap = lane == 0 ? 2 : lane == 2 ? 1 : lane == 3 ? 0 : lane == 5 ? 7 : lane == 6 ? 4 : lane == 7 ? 6 : lane == 8 ? 8 : 31;
        unsigned w = __shfl_sync(0xffffffffu, v, srcmap);
        if (srcmap == 31 || (srcmap >= 4 && !hd.has_fec) || !ok) w = 0;
        if (lane == 1) w = (unsigned)off;
        if (lane == 9) w = ok ? 0u : 1u;
        if (lane < 10) ((int*)&p.meta[f])[lane] = (int)w;
        const unsigned fcnst = __shfl_sync(0xffffffffu, v, 3), ffec = __shfl_sync(0xffffffffu, v, 5);
        if (p.hres && lane < 4)
            ((int*)&p.hres[f])[lane] = lane == 0 ? ok : lane == 1 ? (ok ? (int)fcnst : 0) : lane == 2 ? ((ok && hd.has_fec) ? (int)ffec : 0) : 0;
    }
}

// The parser keeps the last valid constellation (ofdm_adaptive_packet_header.cc:287-289: a field of 0 or above the
// highest known constellation leaves it unchanged) and derives packet_len from it (:296-298): sequential per stream.
// One warp per stream, 32 frames per step: a lane's constellation is that of the nearest lane at or below it whose
// field is valid (ballot + count-leading-zeros), else the one carried in.
__global__ void __launch_bounds__(128) header_fix_kernel(HdrParams p)
{
    const int s = (int)((blockIdx.x * (unsigned)blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
    if (s >= p.n_streams) return;
    int cur = p.st[s].hdr_cnst;
    for (int j0 = 0; j0 < p.F; j0 += 32) {
        const int j = j0 + lane;
        dtl_frame_meta* m = p.meta + (long long)s * p.F + j;
        int c = 0, pay = 0;
        bool live = false;
        if (j < p.F) { live = !(m->reserved & 1); c = m->constellation; pay = m->payload_len; }
        const bool sets = live && c >= 1 && c <= DTL_CNST_MAX;
        const unsigned mask = __ballot_sync(0xffffffffu, sets);
        const unsigned upto = mask & (0xffffffffu >> (31 - lane));
        const int cv = __shfl_sync(0xffffffffu, c, upto ? 31 - __clz(upto) : 0);
        const int mine = upto ? cv : cur;
        if (live) {
            const int bps = p.cn.bps[mine];
            m->constellation = mine;
            m->packet_len = (pay * 8 + bps - 1) / bps;
        }
        const int last = __shfl_sync(0xffffffffu, c, mask ? 31 - __clz(mask) : 0);
        if (mask) cur = last;
    }
    if (lane == 0) p.st[s].hdr_cnst = cur;
}

cudaError_t launch_header(cudaStream_t st, const HdrParams& p)
{
    if (p.n_frames <= 0) return cudaSuccess;
    const int N = p.hd.N, H = p.hd.H;
    const int tpf = N >= 256 ? 256 : (N < 32 ? 32 : N);        // threads per frame; 256-thread CTAs hold 256 / tpf frames
    const int G = 256 / tpf;
    const size_t group = ((size_t)(2 + H) * N * 8 + (size_t)(1 + H) * N * 8 + (size_t)((N / 2 + 4) & ~3) * 4 + (size_t)p.hd.nbits + 15) & ~(size_t)15;
    const size_t smem = group * G;
    const long long ctas = (p.n_frames + G - 1) / G;
    if (smem > 200 * 1024 || ctas > 0x7fffffffLL) return cudaErrorInvalidConfiguration;
    header_kernel<<<(unsigned)ctas, tpf * G, smem, st>>>(p, G, (int)group);
    header_fix_kernel<<<(p.n_streams + 3) / 4, 128, 0, st>>>(p);
    return cudaGetLastError();
}

cudaError_t init_kernels_hdr()
{
    return set_max_dyn_smem(header_kernel);
}

} // namespace dtl
