// kernels_txf.cu -- TX FEC framer: user bytes -> frames of constellation chunks + tags.
//
// Replaces, one frame per work() call in the reference, S streams x F frames per launch here:
//   ofdm_adaptive_fec_frame_bvb_impl::general_work   lib/dtl/ofdm_adaptive_fec_frame_bvb_impl.cc:343-432
//       a TB in flight is continued at the start of the frame, a new TB starts behind it when there is room; TB size from the
//       frame capacity at the constellation of the frame that starts it (:369-370), CRC-32 appended (:394-399), tags (:405-428)
//   tb_encoder::encode / buf_out                     lib/dtl/tb_encoder.cc:32-74, 88-99
//       TB split into ncws shortened codewords (sizes ceil-first), each systematically encoded; the TB bit stream is
//       [parity | data(k_new)] per codeword; bits leave bps at a time, LSB first (repack 1 -> bps)
//   ldpc_enc::encode                                  lib/dtl/ldpc_enc.cc:58-68
// The reference framer is paced by wall clock and by what its input queue holds; here every frame is filled from the
// stream's input bytes as long as a whole TB payload is available (a TB is never started on less).
//
// Four kernels: txf_plan_kernel (one thread per stream: the framer's serial decisions -- which TB bits go to which frame,
// every tag -- kept in a per-stream state that survives between calls), txf_crc_kernel (CRC-32 per new TB),
// txf_encode_kernel (one warp per codeword: ballot the data bits into words, parity = P . data over GF(2), bits written at
// the codeword's place in the TB bit stream), txf_pack_kernel (one CTA per frame: bits -> chunks) and txf_carry_kernel
// (the bits of the TB still in flight move to the stream's carry buffer for the next call).
#include "kernels.cuh"

namespace dtl {

__device__ __forceinline__ int txf_syms(int nbytes, int bps) { return (nbytes * 8 + bps - 1) / bps; }

__global__ void __launch_bounds__(128) txf_plan_kernel(TxfParams p)
{
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= p.n_streams) return;
    TxStream st = p.st[s];
    const int cap = p.cap, in_len = p.in_len ? min(p.in_len[s], p.in_cap) : p.in_cap;
    int in_pos = 0, ntb = 0;
    int cur_tb = -1;                                    // index of the TB in flight among this call's TBs; -1: carried in
    for (int j = 0; j < p.F; j++) {
        const long long f = (long long)s * p.F + j;
        const int c = p.cnst[f], fec = p.fec[f];
        TxSeg* sg = p.segs + 2 * f;
        sg[0].nbits = 0; sg[1].nbits = 0;
        dtl_frame_meta m;
        m.constellation = c; m.carrier_offset = 0; m.frame_no = st.frame_count & 0xFFF; m.reserved = 0;
        st.frame_count++;
        if (c < 1 || c > DTL_CNST_MAX || fec < 1 || fec > p.n_codes) {
            atomicOr(p.err, 1);
            m.payload_len = m.packet_len = m.fec_id = m.tb_no = m.tb_offset = m.tb_payload = 0;
            p.meta[f] = m;
            continue;
        }
        const int bps = p.bps_tab[c];
        int used = 0, payload = 0, offset = 0;
        bool started = false;
        if (st.rem > 0) {                                // (1) continue the TB in flight
            const int avail = (cap - used) * bps / 8;
            const int out_bytes = min((st.rem + 7) / 8, avail);
            const int nb = min(out_bytes * 8, st.rem);
            sg[0].tb = cur_tb; sg[0].bit0 = st.total - st.rem; sg[0].nbits = nb; sg[0].chunk0 = used;
            st.rem -= nb;
            used += txf_syms(out_bytes, bps);
            payload += out_bytes;
        }
        if (st.rem == 0) {                               // (2) start the next TB behind it
            const int avail = (cap - used) * bps / 8;
            const CodeDev& code = p.codes[fec];
            const int frame_bits = (cap * bps / 8) * 8;
            const int ncws = frame_bits > code.n ? 1 + frame_bits / code.n : 1;          // compute_tb_len
            const int payload_max = (ncws * code.k) / 8 - 4;
            if (avail > 0 && payload_max <= 0) atomicOr(p.err, 128);
            else if (avail > 0 && in_pos + payload_max <= in_len) {
                offset = payload;
                TxTb tb;
                tb.fec_id = fec; tb.ncws = ncws; tb.payload_bytes = payload_max; tb.in_off = in_pos;
                tb.bits = ncws * (code.n - code.k) + (payload_max + 4) * 8;
                if (tb.bits > p.maxbits) { atomicOr(p.err, 64); tb.bits = p.maxbits; }
                p.tbs[(long long)s * p.F + ntb] = tb;
                cur_tb = ntb++;
                in_pos += payload_max;
                st.tb_count++;
                st.cur_fec = fec; st.cur_payload = (payload_max + 4) * 8;
                st.total = st.rem = tb.bits;
                started = true;
                int out_bytes = min((st.rem + 7) / 8, avail);
                const int extra = (offset * 8) % bps ? bps - (offset * 8) % bps : 0;
                while (out_bytes > 1 && extra && (out_bytes * 8) % bps && (out_bytes * 8) % bps <= extra) --out_bytes;
                const int nb = min(out_bytes * 8, st.rem);
                sg[1].tb = cur_tb; sg[1].bit0 = 0; sg[1].nbits = nb; sg[1].chunk0 = used;
                st.rem -= nb;
                used += txf_syms(out_bytes, bps);
                payload += out_bytes;
            }
        }
        m.payload_len = payload; m.packet_len = txf_syms(payload, bps);
        m.fec_id = st.cur_fec; m.tb_no = st.tb_count & 0xff; m.tb_offset = started ? offset : 0;
        m.tb_payload = st.cur_payload & 0xffff;
        p.meta[f] = m;
    }
    p.ntb[s] = ntb;
    p.keep[s] = st.rem > 0 ? cur_tb : -2;                 // TB whose bits move to the carry buffer (-1: already there)
    p.st[s] = st;
    if (p.consumed) p.consumed[s] = in_pos;
    if (p.tbs_started) p.tbs_started[s] = ntb;
}

// CRC-32 (crc_util::append_crc, lib/dtl/crc_util.cc:31-39) of every new TB's payload: one thread per TB.
__global__ void __launch_bounds__(128) txf_crc_kernel(TxfParams p)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)p.n_streams * p.F) return;
    const int s = (int)(i / p.F), t = (int)(i - (long long)s * p.F);
    if (t >= p.ntb[s]) return;
    const TxTb tb = p.tbs[i];
    const uint8_t* d = p.data + (size_t)s * p.in_cap + tb.in_off;
    uint32_t crc = 0xFFFFFFFFu;
    for (int q = 0; q < tb.payload_bytes; q++) crc = __ldg(&p.crc_table[(crc ^ d[q]) & 0xffu]) ^ (crc >> 8);
    p.tbcrc[i] = ~crc;
}

// One warp per codeword slot (stream, TB, codeword).
__global__ void __launch_bounds__(256) txf_encode_kernel(TxfParams p)
{
    extern __shared__ uint32_t encw[];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const long long item = (long long)blockIdx.x * (blockDim.x >> 5) + wid;
    const long long slot = item / p.max_ncws;
    const int i = (int)(item - slot * p.max_ncws);
    if (slot >= (long long)p.n_streams * p.F) return;
    const int s = (int)(slot / p.F), t = (int)(slot - (long long)s * p.F);
    if (t >= p.ntb[s]) return;
    const TxTb tb = p.tbs[slot];
    if (i >= tb.ncws) return;
    const CodeDev& cg = p.codes[tb.fec_id];
    const int cn = cg.n, ck = cg.k, kwords = cg.kwords;       // into registers: the stores below could alias the table
    const uint32_t* __restrict__ pmat = cg.pmat;
    uint32_t* dw = encw + wid * p.max_kwords;
    const int len = (tb.payload_bytes + 4) * 8, q = len / tb.ncws, r = len % tb.ncws;
    const int k_new = q + (i < r), read = i * q + min(i, r), ncheck = cn - ck;
    const uint8_t* __restrict__ d = p.data + (size_t)s * p.in_cap + tb.in_off;
    const uint32_t crc = p.tbcrc[slot];
    const int pay = tb.payload_bytes;
    uint8_t* __restrict__ out = p.tbbits + (size_t)slot * p.maxbits + (size_t)i * ncheck + read;
    if ((long long)i * ncheck + read + ncheck + k_new > p.maxbits) return;             // flagged by the planner
    for (int w = 0; w < kwords; w++) {
        const int j = w * 32 + lane;
        unsigned bit = 0;
        if (j < k_new) {
            const int b = read + j;                      // bit of the TB: payload bytes LSB first, then the CRC little-endian
            const int by = b >> 3;
            const unsigned v = by < pay ? d[by] : (crc >> (8 * (by - pay))) & 0xffu;
            bit = (v >> (b & 7)) & 1u;
            out[ncheck + j] = (uint8_t)bit;
        }
        const uint32_t word = __ballot_sync(0xffffffffu, bit);
        if (lane == 0) dw[w] = word;
    }
    __syncwarp();
    for (int row = lane; row < ncheck; row += 32) {
        const uint32_t* pr = pmat + (size_t)row * kwords;
        uint32_t acc = 0;
        for (int w = 0; w < kwords; w++) acc ^= __ldg(&pr[w]) & dw[w];
        out[row] = (uint8_t)(__popc(acc) & 1);
    }
}

// One CTA per frame: zero the frame, then pack each segment's bits into chunks (repack 1 -> bps, LSB first).
__global__ void __launch_bounds__(128) txf_pack_kernel(TxfParams p)
{
    const long long f = blockIdx.x;
    const int s = (int)(f / p.F);
    uint8_t* ch = p.chunks + (size_t)f * p.cap;
    const int c = p.cnst[f];
    const int bps = (c >= 1 && c <= DTL_CNST_MAX) ? p.bps_tab[c] : 1;
    const TxSeg sg0 = p.segs[2 * f], sg1 = p.segs[2 * f + 1];
    const int n0 = (sg0.nbits + bps - 1) / bps, n1 = (sg1.nbits + bps - 1) / bps;
    for (int q = threadIdx.x; q < p.cap; q += blockDim.x) {
        TxSeg sg; int k = -1;
        if (sg0.nbits > 0 && q >= sg0.chunk0 && q < sg0.chunk0 + n0) { sg = sg0; k = q - sg0.chunk0; }
        else if (sg1.nbits > 0 && q >= sg1.chunk0 && q < sg1.chunk0 + n1) { sg = sg1; k = q - sg1.chunk0; }
        unsigned v = 0;
        if (k >= 0) {
            const uint8_t* bits = sg.tb >= 0 ? p.tbbits + ((size_t)s * p.F + sg.tb) * p.maxbits : p.carry + (size_t)s * p.maxbits;
            const int b0 = k * bps, nb = min(bps, sg.nbits - b0);
            for (int b = 0; b < nb; b++) v |= (unsigned)(bits[sg.bit0 + b0 + b] & 1u) << b;
        }
        ch[q] = (uint8_t)v;
    }
}

__global__ void __launch_bounds__(256) txf_carry_kernel(TxfParams p)
{
    const int s = blockIdx.x;
    const int t = p.keep[s];
    if (t < 0) return;
    const int n = p.tbs[(long long)s * p.F + t].bits;
    const uint8_t* src = p.tbbits + ((size_t)s * p.F + t) * p.maxbits;
    uint8_t* dst = p.carry + (size_t)s * p.maxbits;
    for (int i = threadIdx.x; i < n; i += blockDim.x) dst[i] = src[i];
}

__global__ void txf_reset_kernel(TxStream* st, int n)
{
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s < n) { TxStream z; z.tb_count = 0; z.frame_count = 0; z.rem = 0; z.total = 0; z.cur_fec = 0; z.cur_payload = 0; st[s] = z; }
}

cudaError_t launch_txf_reset(cudaStream_t st, TxStream* s, int n)
{
    if (n > 0) txf_reset_kernel<<<(n + 127) / 128, 128, 0, st>>>(s, n);
    return cudaGetLastError();
}

cudaError_t launch_txf(cudaStream_t st, const TxfParams& p, int* launches)
{
    if (p.n_streams <= 0 || p.F <= 0) return cudaSuccess;
    const long long slots = (long long)p.n_streams * p.F, items = slots * p.max_ncws;
    if (slots > 0x7fffffffLL || (items + 7) / 8 > 0x7fffffffLL) return cudaErrorInvalidConfiguration;
    txf_plan_kernel<<<(p.n_streams + 127) / 128, 128, 0, st>>>(p);
    txf_crc_kernel<<<(unsigned)((slots + 127) / 128), 128, 0, st>>>(p);
    txf_encode_kernel<<<(unsigned)((items + 7) / 8), 256, (size_t)8 * p.max_kwords * 4, st>>>(p);
    txf_pack_kernel<<<(unsigned)slots, 128, 0, st>>>(p);
    txf_carry_kernel<<<p.n_streams, 256, 0, st>>>(p);
    if (launches) *launches += 5;
    return cudaGetLastError();
}

} // namespace dtl
