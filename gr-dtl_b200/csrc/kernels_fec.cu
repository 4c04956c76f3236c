// kernels_fec.cu -- stage 4: transport-block reassembly, LDPC belief-propagation decode, CRC-32, byte packing.
//
// Replaces (one frame / one codeword at a time in the reference):
//   ofdm_adaptive_fec_decoder_impl::general_work  lib/dtl/ofdm_adaptive_fec_decoder_impl.cc:80-217
//   make_fec_info / compute_tb_len                lib/dtl/fec_utils.cc:60-121
//   tb_decoder::process_frame / decode            lib/dtl/tb_decoder.cc:32-195
//   ldpc_dec::decode -> gr-fec awgn_bp::decode    lib/dtl/ldpc_dec.cc:62-71
//   ofdm_adaptive_fec_pack_bb                     lib/dtl/ofdm_adaptive_fec_pack_bb_impl.cc:48-69
//
// Pipeline of one call (on the handle's stream; the small tag-only kernels run on its side stream beside the heavy ones,
// nothing returns to the host in between):
//   1. tb_plan_kernel       one WARP per stream replays tb_decoder's frame-by-frame state machine on the frame tags only
//                           (32 frames' tags per coalesced fetch, shuffled to every lane), emitting for every released
//                           transport block (TB) the list of LLR segments it is made of (carry buffer from earlier calls +
//                           ranges of this call's frames) and one work item per codeword.
//   2. ldpc_screen_kernel   bit-sliced syndrome check of every codeword's hard decisions, read from the SIGN WORDS the
//                           equaliser kernel writes next to the LLRs (1 bit per LLR); clean codewords leave here, the others
//                           are queued on four lists by a coarse syndrome weight, hardest first, each with a record
//                           (BpItem) of where its bits go and where its received run lies: a run inside one LLR row is
//                           fetched IN PLACE by the decoder (the 16-byte-aligned superset of the run + a lead-in).
//   3. bp_gather_kernel     only the queued codewords whose run straddles two segments (or is cut short): pieces and
//                           defaults (shortened positions = -15 of tb_decoder.cc:145) -> one contiguous wire-order block.
//   4. ldpc_bp_ref_kernel   (bp_ref.cuh; DTL_LDPC_REFERENCE) or ldpc_bp_lay_kernel (bp_layered.cuh; DTL_LDPC_LAYERED):
//                           persistent grid, one codeword per warp from the queue, the NEXT codeword's run prefetched by a
//                           TMA bulk copy from the address in its record (sign flip of ldpc_dec.cc:64-66 on consumption);
//                           messages in shared memory; syndrome checks are warp votes, so a warp leaves as
//                           soon as its word converges.  (ldpc_bp_kernel below is the slot-major decoder of round 1, kept for
//                           codes without a row-major instance and for A/B: DTL_BP_OLD=1.)
//   5. tb_finish_kernel     one warp per TB: bits -> bytes LSB-first, mean iterations.
//   6. tb_gather_kernel     one CTA per stream: CRC-32 of its TBs, CRC-passing payload bytes concatenated in release order.
//   7. carry_save_kernel    LLRs (and sign words) of the TB still in flight move to the per-stream carry buffer.
//
// Decoder flavours: DTL_LDPC_REFERENCE is gr-fec's awgn_bp restated operation by operation (flooding
// sum-product on likelihood ratios, IEEE double, products in alist order) so payload bits AND iteration
// counts match the reference; DTL_LDPC_LAYERED is a layered normalised min-sum in fp32 (throughput mode).
#include "kernels.cuh"
#include "bp_ref.cuh"
#include "bp_layered.cuh"
#include <algorithm>
#include <cstdlib>

namespace dtl {

#define ERR_TAGS 4
#define ERR_FEC_TAGS 8
#define ERR_TB_SLOTS 16
#define ERR_CW_CAP 32
#define ERR_BITS_CAP 64
#define ERR_TB_META 128
#define ERR_SEGS 256
#define ERR_CARRY 512

__device__ __forceinline__ int tb_ncws(int cw_len, int frame_len) { return frame_len > cw_len ? 1 + frame_len / cw_len : 1; }

struct Rcv { TbSeg seg[DTL_TB_MAXSEG]; int nseg; int len; bool bad; };
__device__ __forceinline__ void rcv_clear(Rcv& r) { r.nseg = 0; r.len = 0; r.bad = false; }
__device__ __forceinline__ void rcv_push(Rcv& r, int src, int start, int len)
{
    if (len <= 0) return;
    if (r.nseg == DTL_TB_MAXSEG) { r.bad = true; return; }
    r.seg[r.nseg].src = src; r.seg[r.nseg].start = start; r.seg[r.nseg].len = len;
    r.nseg++; r.len += len;
}

// One WARP per stream.  Lanes fetch the tags of 32 frames at a time (coalesced); the state machine itself is serial over
// the frames of a stream and runs replicated in every lane on shuffled tags, so its loop touches no memory it has to
// wait for.  Released TBs are numbered inside the stream first (codeword index, bit offset); the stream then takes its
// ranges of the global work lists with ONE atomic per list, and the lanes fill in the TB list and the per-codeword
// work items (geometry of tb_decoder.cc:161-183 in closed form: the first payload % ncws codewords carry one more bit).
__global__ void __launch_bounds__(128) tb_plan_kernel(FecParams p)
{
    const int lane = threadIdx.x & 31;
    const int s = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (s >= p.n_streams) return;
    StreamState st = p.st[s];
    Rcv rcv; rcv_clear(rcv);
    if (st.carry_len > 0) rcv_push(rcv, -1, 0, st.carry_len);
    int ntb = 0, err = 0, cw_local = 0;
    long long bits_local = 0;

    auto release = [&](int fec_id, int ncws, int payload_len, int tb_no, int frame_index) {
        const CodeDev& c = p.codes[fec_id];
        if (rcv.bad) { err |= ERR_SEGS; return; }
        if (payload_len < 0 || payload_len > ncws * c.k) { err |= ERR_TB_META; return; }
        if (ntb >= p.tb_slots) { err |= ERR_TB_SLOTS; return; }
        if (lane == 0) {
            TbDesc& d = p.tbs[s * p.tb_slots + ntb];
            d.stream = s; d.fec_id = fec_id; d.ncws = ncws; d.payload_len = payload_len; d.tb_no = tb_no;
            d.frame_index = frame_index; d.nseg = rcv.nseg; d.cw_first = cw_local; d.bit_base = bits_local;
            for (int i = 0; i < rcv.nseg; i++) d.seg[i] = rcv.seg[i];
        }
        cw_local += ncws; bits_local += (payload_len + 3) & ~3;
        ntb++;
    };

    for (int j0 = 0; j0 < p.F; j0 += 32) {
        dtl_frame_meta mine;
        mine.constellation = 0; mine.payload_len = 0; mine.fec_id = 0; mine.tb_no = 0; mine.tb_offset = 0; mine.tb_payload = 0;
        if (j0 + lane < p.F) mine = p.meta[(long long)s * p.F + j0 + lane];
        const int jn = min(32, p.F - j0);
        for (int jj = 0; jj < jn; jj++) {
            const int j = j0 + jj;
            dtl_frame_meta m;
            m.payload_len = __shfl_sync(0xffffffffu, mine.payload_len, jj);
            if (m.payload_len == 0) continue;                                   // frame not emitted / empty (:141-143)
            m.constellation = __shfl_sync(0xffffffffu, mine.constellation, jj);
            m.fec_id = __shfl_sync(0xffffffffu, mine.fec_id, jj);
            m.tb_no = __shfl_sync(0xffffffffu, mine.tb_no, jj);
            m.tb_offset = __shfl_sync(0xffffffffu, mine.tb_offset, jj);
            m.tb_payload = __shfl_sync(0xffffffffu, mine.tb_payload, jj);
            const int bps = (m.constellation >= 1 && m.constellation <= DTL_CNST_MAX) ? p.bps_tab[m.constellation] : 0;
            if (!bps) { err |= ERR_TAGS; continue; }                            // "Tags missing" (:121-126)
            if (m.fec_id < 1 || m.fec_id > p.n_codes) { err |= ERR_FEC_TAGS; continue; }   // "FEC tags missing" (:136-138)
            const int frame_index = st.frame_index++;
            const CodeDev& fc = p.codes[m.fec_id];
            const int fpl = 8 * m.payload_len, tb_off = 8 * m.tb_offset;       // :147
            const int frame_len = 8 * ((p.cap * bps + 7) / 8);
            if (m.tb_no == st.tb_number) {                                      // tb_decoder.cc:57-70
                if (st.buf_idx + fpl > 2 * p.max_tb) { err |= ERR_CARRY; continue; }
                rcv_push(rcv, j, 0, fpl);
                st.buf_idx += fpl;
                const int ncws = tb_ncws(p.codes[st.info_fec].n, frame_len);
                const int expected = (fc.n - fc.k) * ncws + m.tb_payload;
                if (rcv.len >= expected) {
                    release(st.info_fec, ncws, st.info_tb_payload, m.tb_no, frame_index);
                    st.buf_idx = 0; rcv_clear(rcv);
                }
            } else if (rcv.len == 0 && tb_off == fpl) {                         // :75-89
                st.info_fec = m.fec_id; st.info_tb_payload = m.tb_payload; st.info_tb_number = m.tb_no; st.have_info = 1;
                st.tb_number = m.tb_no; st.tb_payload_len = m.tb_payload; st.buf_idx = 0;
                rcv_clear(rcv);
                rcv_push(rcv, j, 0, tb_off);
                release(m.fec_id, tb_ncws(fc.n, frame_len), m.tb_payload, m.tb_no, frame_index);
                rcv_clear(rcv);
            } else {                                                            // :92-133
                if (tb_off > 0) rcv_push(rcv, j, 0, tb_off);
                if (rcv.len > 0 && st.have_info)
                    release(st.info_fec, tb_ncws(p.codes[st.info_fec].n, frame_len), st.info_tb_payload, st.info_tb_number,
                            frame_index);
                st.info_fec = m.fec_id; st.info_tb_payload = m.tb_payload; st.info_tb_number = m.tb_no; st.have_info = 1;
                st.tb_number = m.tb_no; st.tb_payload_len = m.tb_payload; st.buf_idx = 0;
                rcv_clear(rcv);
                int extra = 0;
                if (tb_off % bps) extra = bps - tb_off % bps;
                const int new_off = tb_off ? tb_off + extra : 0;
                rcv_push(rcv, j, new_off, fpl + extra - new_off);
                st.buf_idx += fpl - tb_off;
                const int ncws = tb_ncws(fc.n, frame_len);
                const int tb_size = 8 * ((ncws * (fc.n - fc.k) + m.tb_payload + 7) / 8);
                if (fpl - tb_off == tb_size) {
                    release(m.fec_id, ncws, m.tb_payload, m.tb_no, frame_index);
                    st.buf_idx = 0; rcv_clear(rcv);
                }
            }
        }
    }
    // what is left belongs to the TB in flight: carry_save_kernel moves it to the carry buffer
    if (rcv.bad) err |= ERR_SEGS;
    const bool carry_over = rcv.len > p.carry_cap;
    if (carry_over) err |= ERR_CARRY;
    st.carry_len = carry_over ? 0 : rcv.len;
    if (lane == 0) {
        TbDesc& pd = p.pend[s];
        pd.nseg = carry_over ? 0 : rcv.nseg; pd.stream = s;
        for (int i = 0; i < rcv.nseg; i++) pd.seg[i] = rcv.seg[i];
        p.tb_count[s] = ntb;
        p.st[s] = st;
    }

    // ---- the stream's ranges of the global lists, then the work items
    int cw_base = 0, tb_base = 0;
    long long bit_base = 0;
    if (lane == 0 && ntb) {
        cw_base = atomicAdd(&p.cnt[0], cw_local);
        bit_base = (long long)atomicAdd(p.bits_used, (unsigned long long)bits_local);
        tb_base = atomicAdd(&p.cnt[2], ntb);
    }
    cw_base = __shfl_sync(0xffffffffu, cw_base, 0);
    bit_base = __shfl_sync(0xffffffffu, bit_base, 0);
    tb_base = __shfl_sync(0xffffffffu, tb_base, 0);
    if (ntb && cw_base + cw_local > p.cw_cap) err |= ERR_CW_CAP;
    if (ntb && bit_base + bits_local > p.bits_cap) { err |= ERR_BITS_CAP; bit_base = 0; bits_local = 0; }   // failing call: stay in bounds
    __syncwarp();
    for (int i = 0; i < ntb; i++) {
        const int gi = s * p.tb_slots + i;
        TbDesc& d = p.tbs[gi];
        const int ncws = d.ncws, payload_len = d.payload_len, first = cw_base + d.cw_first;
        const long long bb = bits_local ? bit_base + d.bit_base : 0;
        const CodeDev& c = p.codes[d.fec_id];
        const int nk = c.n - c.k, q = payload_len / ncws, r = payload_len - q * ncws;
        __syncwarp();
        if (lane == 0) { d.cw_first = first; d.bit_base = bb; p.tb_list[tb_base + i] = gi; }
        for (int cw = lane; cw < ncws; cw += 32) {
            const int it = first + cw;
            if (it >= p.cw_cap) break;
            const int mn = min(cw, r);
            p.cw_tb[it] = gi;
            p.cw_geo[3 * it] = cw * nk + cw * q + mn; p.cw_geo[3 * it + 1] = cw * q + mn; p.cw_geo[3 * it + 2] = q + (cw < r ? 1 : 0);
        }
    }
    if (err && lane == 0) atomicOr(p.err, err);
}

// ---- awgn_bp::decode for one codeword per warp ---------------------------------------------------------------
// Graph tables live in shared memory (staged once per CTA).  Edge storage is slot-major: edge j of check row r
// (alist order) is slot j*m + r, so the lanes of a warp, which own consecutive rows, touch consecutive doubles
// (no bank conflicts in the check pass); col_edge[] lists the slots of a variable's edges in alist order.
// Per warp: msg[slots] double (Q and R share storage: a check overwrites its Q's with R's, a variable its R's
// with Q's), lr0[n] double (channel likelihood ratios), est[n] u8.  The received values rx[n] (float) are only
// needed until lr0 is built and live at the start of the msg area.
struct CodeS {
    int n, m, k, rank, slots, edges, max_row, max_col;
    int o_row_len, o_row_var, o_col_ptr, o_col_edge, o_permute;     // byte offsets from the CTA's dynamic shared memory
    const unsigned char* smb;                                       // = the kernel's `sm` (set in the kernel body so the
                                                                    //   compiler keeps the accesses in the shared space)
    __device__ __forceinline__ const uint16_t* row_len() const { return (const uint16_t*)(smb + o_row_len); }
    __device__ __forceinline__ const uint16_t* row_var() const { return (const uint16_t*)(smb + o_row_var); }
    __device__ __forceinline__ const uint16_t* col_ptr() const { return (const uint16_t*)(smb + o_col_ptr); }
    __device__ __forceinline__ const uint16_t* col_edge() const { return (const uint16_t*)(smb + o_col_edge); }
    __device__ __forceinline__ const uint16_t* permute() const { return (const uint16_t*)(smb + o_permute); }
};
struct WarpBuf { double* msg; double* lr0; float* rx; uint8_t* est; };

static __host__ __device__ inline size_t code_tab_bytes(int n, int m, int slots, int edges)
{
    return (2 * ((size_t)m + slots + (n + 1) + edges + n) + 15) & ~(size_t)15;
}
size_t ldpc_tab_bytes(int n, int m, int slots, int edges) { return code_tab_bytes(n, m, slots, edges); }

static __host__ __device__ inline size_t warp_msg_bytes(int n, int slots)
{
    size_t msg = (size_t)slots * 8;
    if (msg < (size_t)n * 4) msg = (size_t)n * 4;
    return (msg + 7) & ~(size_t)7;
}
size_t ldpc_warp_bytes(int n, int slots)
{
    return (warp_msg_bytes(n, slots) + (size_t)n * 8 + (size_t)((n + 16) & ~15) + 15) & ~(size_t)15;   // est[n] = 0 pad byte
}

__device__ __forceinline__ WarpBuf carve(unsigned char* base, int n, int slots)
{
    const size_t msg = warp_msg_bytes(n, slots);
    WarpBuf w;
    w.msg = (double*)base;
    w.rx = (float*)base;
    w.lr0 = (double*)(base + msg);
    w.est = (uint8_t*)(base + msg + (size_t)n * 8);
    return w;
}

// Copies one code's tables from HBM into shared memory at `tab` (all threads of the CTA) and returns the
// shared-memory view.  The caller synchronises.
__device__ __forceinline__ CodeS stage_code(const CodeDev& g, unsigned char* sm, int tab_off)
{
    const int nthr = blockDim.x, slots = g.max_row * g.m;
    uint16_t* d = (uint16_t*)(sm + tab_off);
    uint16_t* row_len = d; d += g.m;
    uint16_t* row_var = d; d += slots;
    uint16_t* col_ptr = d; d += g.n + 1;
    uint16_t* col_edge = d; d += g.edges;
    uint16_t* permute = d;
    for (int i = threadIdx.x; i < g.m; i += nthr) row_len[i] = g.row_len[i];
    for (int i = threadIdx.x; i < slots; i += nthr) row_var[i] = g.row_var[i];
    for (int i = threadIdx.x; i <= g.n; i += nthr) col_ptr[i] = g.col_ptr[i];
    for (int i = threadIdx.x; i < g.edges; i += nthr) col_edge[i] = g.col_edge[i];
    for (int i = threadIdx.x; i < g.n; i += nthr) permute[i] = g.permute[i];
    CodeS c;
    c.n = g.n; c.m = g.m; c.k = g.k; c.rank = g.rank; c.slots = slots; c.edges = g.edges; c.max_row = g.max_row; c.max_col = g.max_col;
    c.o_row_len = tab_off; c.o_row_var = c.o_row_len + 2 * g.m; c.o_col_ptr = c.o_row_var + 2 * slots;
    c.o_col_edge = c.o_col_ptr + 2 * (g.n + 1); c.o_permute = c.o_col_edge + 2 * g.edges;
    c.smb = sm;
    return c;
}

__device__ __forceinline__ bool syndrome_nonzero(const CodeS& c, const uint8_t* est, int lane)
{
    bool bad = false;
    for (int r = lane; r < c.m; r += 32) {
        const int deg = c.row_len()[r];
        unsigned par = 0;
        for (int j = 0; j < deg; j++) par ^= est[c.row_var()[j * c.m + r]];
        bad |= (par & 1u) != 0;
    }
    return __any_sync(0xffffffffu, bad);
}

// The same check with every row run at max_row edges, unrolled, no per-lane predicate: the pad slots of row_var point at
// est[n], a byte the caller keeps at 0.
template <int RMAX>
__device__ __forceinline__ bool syndrome_nonzero_pad(const CodeS& c, const uint8_t* est, int lane)
{
    const uint16_t* rv = c.row_var();
    const int m = c.m;
    bool bad = false;
    for (int r = lane; r < m; r += 32) {
        unsigned par = 0;
#pragma unroll
        for (int j = 0; j < RMAX; j++) par ^= est[rv[j * m + r]];   // RMAX == max_row
        bad |= (par & 1u) != 0;
    }
    return __any_sync(0xffffffffu, bad);
}

// One flooding iteration, any degrees (awgn_bp::update_chks / update_vars, products in alist order).
__device__ __forceinline__ void bp_iter_generic(const CodeS& c, const WarpBuf& w, int lane)
{
    for (int r = lane; r < c.m; r += 32) {                                         // update_chks
        const int deg = c.row_len()[r];
        double prod = 1.0;
        for (int j = 0; j < deg; j++)
            prod = __dmul_rn(prod, __dsub_rn(__ddiv_rn(2.0, __dadd_rn(1.0, w.msg[j * c.m + r])), 1.0));
        for (int j = 0; j < deg; j++) {
            const double t = __dsub_rn(__ddiv_rn(2.0, __dadd_rn(1.0, w.msg[j * c.m + r])), 1.0);
            const double pp = __ddiv_rn(prod, t);
            w.msg[j * c.m + r] = __ddiv_rn(__dsub_rn(1.0, pp), __dadd_rn(1.0, pp));
        }
    }
    __syncwarp();
    for (int v = lane; v < c.n; v += 32) {                                         // update_vars + decision
        const int q0 = c.col_ptr()[v], q1 = c.col_ptr()[v + 1];
        double sacc = w.lr0[v];
        for (int q = q0; q < q1; q++) sacc = __dmul_rn(sacc, w.msg[c.col_edge()[q]]);
        for (int q = q0; q < q1; q++) { const int e = c.col_edge()[q]; w.msg[e] = __ddiv_rn(sacc, w.msg[e]); }
        w.est[v] = sacc > 1.0 ? 1 : 0;
    }
    __syncwarp();
}

// The same iteration for row weight <= RMAX and column weight <= CMAX with the per-edge work unrolled, branch-free
// and predicated: the divisions of a row (column) are independent instruction chains the scheduler interleaves, and
// each t = 2/(1+Q) - 1 is computed once (identical value, identical roundings).
template <int RMAX, int CMAX>
__device__ __forceinline__ void bp_iter_unrolled(const CodeS& c, const WarpBuf& w, int lane)
{
    // RMAX == max_row.  The pad slots of rows lighter than max_row hold Q = 0.0 (bp_reference_run): t = 2/(1+0) - 1 = 1
    // exactly and prod * 1 = prod exactly, so every row runs RMAX edges with no per-lane select; only the stores look at deg.
    for (int r = lane; r < c.m; r += 32) {
        const int deg = c.row_len()[r];
        double den[RMAX], t[RMAX];
        bool ok[RMAX], all_ok = true;
#pragma unroll
        for (int j = 0; j < RMAX; j++) den[j] = __dadd_rn(1.0, w.msg[j * c.m + r]);
#pragma unroll
        for (int j = 0; j < RMAX; j++) { t[j] = ddiv_try(2.0, den[j], ok[j]); all_ok &= ok[j]; }
        if (__builtin_expect(!all_ok, 0)) {
#pragma unroll
            for (int j = 0; j < RMAX; j++) if (!ok[j]) t[j] = __ddiv_rn(2.0, den[j]);
        }
#pragma unroll
        for (int j = 0; j < RMAX; j++) t[j] = __dsub_rn(t[j], 1.0);
        double prod = 1.0;
#pragma unroll
        for (int j = 0; j < RMAX; j++) prod = __dmul_rn(prod, t[j]);
        double pp[RMAX];
        all_ok = true;
#pragma unroll
        for (int j = 0; j < RMAX; j++) { pp[j] = ddiv_try(prod, t[j], ok[j]); all_ok &= ok[j]; }
        if (__builtin_expect(!all_ok, 0)) {
#pragma unroll
            for (int j = 0; j < RMAX; j++) if (!ok[j]) pp[j] = __ddiv_rn(prod, t[j]);
        }
        double o[RMAX];
        all_ok = true;
#pragma unroll
        for (int j = 0; j < RMAX; j++) { o[j] = ddiv_try(__dsub_rn(1.0, pp[j]), __dadd_rn(1.0, pp[j]), ok[j]); all_ok &= ok[j]; }
        if (__builtin_expect(!all_ok, 0)) {
#pragma unroll
            for (int j = 0; j < RMAX; j++) if (!ok[j]) o[j] = __ddiv_rn(__dsub_rn(1.0, pp[j]), __dadd_rn(1.0, pp[j]));
        }
#pragma unroll
        for (int j = 0; j < RMAX; j++) if (j < deg) w.msg[j * c.m + r] = o[j];
    }
    __syncwarp();
    for (int v = lane; v < c.n; v += 32) {
        const int q0 = c.col_ptr()[v], deg = c.col_ptr()[v + 1] - q0;
        int e[CMAX]; double mv[CMAX];
#pragma unroll
        for (int j = 0; j < CMAX; j++) { e[j] = j < deg ? c.col_edge()[q0 + j] : 0; mv[j] = j < deg ? w.msg[e[j]] : 1.0; }
        double sacc = w.lr0[v];
#pragma unroll
        for (int j = 0; j < CMAX; j++) if (j < deg) sacc = __dmul_rn(sacc, mv[j]);
        double o[CMAX];
        bool ok[CMAX], all_ok = true;
#pragma unroll
        for (int j = 0; j < CMAX; j++) { o[j] = ddiv_try(sacc, mv[j], ok[j]); all_ok &= ok[j]; }
        if (__builtin_expect(!all_ok, 0)) {
#pragma unroll
            for (int j = 0; j < CMAX; j++) if (!ok[j]) o[j] = __ddiv_rn(sacc, mv[j]);
        }
#pragma unroll
        for (int j = 0; j < CMAX; j++) if (j < deg) w.msg[e[j]] = o[j];
        w.est[v] = sacc > 1.0 ? 1 : 0;
    }
    __syncwarp();
}

template <int RMAX, int CMAX>
__device__ __forceinline__ int bp_loop(const CodeS& c, const WarpBuf& w, int lane, int max_it)
{
    int nit = 0;
    while (nit < max_it) {
        nit++;
        if (RMAX > 0) bp_iter_unrolled<(RMAX > 0 ? RMAX : 1), (CMAX > 0 ? CMAX : 1)>(c, w, lane);
        else bp_iter_generic(c, w, lane);
        if (RMAX > 0 ? !syndrome_nonzero_pad<(RMAX > 0 ? RMAX : 1)>(c, w.est, lane) : !syndrome_nonzero(c, w.est, lane)) break;
    }
    return nit;
}

// rx[] (H column order, rx < 0 => bit 1) and est[] (hard decisions of rx, known to fail the syndrome check) must be
// filled; returns the iteration count, est[] = final decisions.
__device__ __forceinline__ int bp_reference_run(const CodeS& c, const WarpBuf& w, int lane, int max_it, double sigma2)
{
    if (lane == 0) w.est[c.n] = 0;                                                     // pad byte of the syndrome check
    for (int v = lane; v < c.n; v += 32) {                                             // rx_lr_calc
        const float m2 = -1 * 2 * w.rx[v];
        bool ok;
        double q = ddiv_try((double)m2, sigma2, ok);
        if (__builtin_expect(!ok, 0)) q = __ddiv_rn((double)m2, sigma2);
        w.lr0[v] = exp(q);
    }
    __syncwarp();                                                                      // rx (aliases msg) is dead now
    const bool unroll_init = c.max_col <= 4;
    if (unroll_init) {                                                                 // spa_initialize, edges unrolled
        for (int v = lane; v < c.n; v += 32) {
            const double lr = w.lr0[v];
            const int q0 = c.col_ptr()[v], deg = c.col_ptr()[v + 1] - q0;
#pragma unroll
            for (int j = 0; j < 4; j++) if (j < deg) w.msg[c.col_edge()[q0 + j]] = lr;
        }
    } else {
        for (int v = lane; v < c.n; v += 32) {
            const double lr = w.lr0[v];
            for (int q = c.col_ptr()[v]; q < c.col_ptr()[v + 1]; q++) w.msg[c.col_edge()[q]] = lr;
        }
    }
    __syncwarp();
    if (c.slots != c.edges) {                                                          // pad slots: Q = 0 (see bp_iter_unrolled)
        for (int e = lane; e < c.slots; e += 32) if (c.row_var()[e] == c.n) w.msg[e] = 0.0;
        __syncwarp();
    }
    // the unrolled instance whose row count IS the code's max_row (pad slots exist for exactly that many edges per row)
    if (c.max_col <= 3) {
        switch (c.max_row) {
        case 3: return bp_loop<3, 3>(c, w, lane, max_it);
        case 4: return bp_loop<4, 3>(c, w, lane, max_it);
        case 5: return bp_loop<5, 3>(c, w, lane, max_it);
        case 6: return bp_loop<6, 3>(c, w, lane, max_it);
        case 7: return bp_loop<7, 3>(c, w, lane, max_it);
        case 8: return bp_loop<8, 3>(c, w, lane, max_it);
        default: break;
        }
    } else if (c.max_col <= 4) {
        switch (c.max_row) {
        case 4: return bp_loop<4, 4>(c, w, lane, max_it);
        case 5: return bp_loop<5, 4>(c, w, lane, max_it);
        case 6: return bp_loop<6, 4>(c, w, lane, max_it);
        case 8: return bp_loop<8, 4>(c, w, lane, max_it);
        default: break;
        }
    }
    return bp_loop<0, 0>(c, w, lane, max_it);
}

// Layered normalised min-sum (fp32, LLR domain, rx > 0 => bit 0): check rows are processed in order, the lanes of
// the warp cooperating over one row's edges (row weight <= 32).  est[] must hold the hard decisions of rx.
__device__ __forceinline__ int bp_layered_run(const CodeS& c, const WarpBuf& w, int lane, int max_it)
{
    float* post = (float*)w.lr0;                  // posterior LLR per variable
    float* cmsg = (float*)w.msg;                  // check->variable message per slot
    for (int v = lane; v < c.n; v += 32) post[v] = w.rx[v];
    __syncwarp();                                 // rx (aliases cmsg) is dead now
    for (int e = lane; e < c.slots; e += 32) cmsg[e] = 0.f;
    __syncwarp();
    const float norm = 0.8125f;
    int nit = 0;
    while (nit < max_it) {
        nit++;
        for (int r = 0; r < c.m; r++) {
            const int deg = c.row_len()[r], slot = lane * c.m + r;
            float q = 1e30f; int v = 0; float old = 0.f;
            if (lane < deg) { v = c.row_var()[slot]; old = cmsg[slot]; q = post[v] - old; }
            float a = fabsf(q);
            unsigned sgn = __ballot_sync(0xffffffffu, lane < deg && q < 0.f);
            float m1 = a; int arg = lane;
#pragma unroll
            for (int o = 16; o; o >>= 1) {
                const float om = __shfl_xor_sync(0xffffffffu, m1, o); const int oa = __shfl_xor_sync(0xffffffffu, arg, o);
                if (om < m1 || (om == m1 && oa < arg)) { m1 = om; arg = oa; }
            }
            float a2 = lane == arg ? 1e30f : a, m2 = a2;
#pragma unroll
            for (int o = 16; o; o >>= 1) m2 = fminf(m2, __shfl_xor_sync(0xffffffffu, m2, o));
            if (lane < deg) {
                const float mag = norm * (lane == arg ? m2 : m1);
                const bool neg = ((__popc(sgn) & 1) != 0) != (q < 0.f);
                const float nm = neg ? -mag : mag;
                cmsg[slot] = nm;
                post[v] = q + nm;
            }
            __syncwarp();
        }
        for (int v = lane; v < c.n; v += 32) w.est[v] = post[v] < 0.f ? 1 : 0;
        __syncwarp();
        if (!syndrome_nonzero(c, w.est, lane)) break;
    }
    return nit;
}

// Gather of one codeword: n LLRs of the TB's receive buffer (reads past the received length = 0, shortened positions
// = -15, tb_decoder.cc:145,161-183), negated and permuted into H column order (ldpc_dec.cc:64-66), plus the hard
// decisions (awgn_bp::compute_init_estimate).  Segment-major: each segment that overlaps the codeword's LLR range
// is copied by the whole warp with consecutive lanes on consecutive floats.
__device__ __forceinline__ void gather_cw(const FecParams& p, const TbDesc& d, const CodeS& c, int it, float* rx, uint8_t* est, int lane)
{
    const int rcv_idx = p.cw_geo[3 * it], len = (c.n - c.k) + p.cw_geo[3 * it + 2];
    for (int q = lane; q < c.n; q += 32) {
        const float x = q < len ? -0.0f : 15.0f;          // -1 * 0 and -1 * -15
        const int col = c.permute()[q];
        rx[col] = x;
        est[col] = 0;
    }
    __syncwarp();
    const int nseg = d.nseg;
    int base = 0;
    for (int i = 0; i < nseg; i++) {
        const TbSeg sg = d.seg[i];
        const int lo = max(rcv_idx, base), hi = min(rcv_idx + len, base + sg.len);
        if (lo < hi) {
            const float* src = sg.src < 0 ? p.carry + (size_t)d.stream * p.carry_cap + sg.start
                                          : p.llr + ((size_t)d.stream * p.F + sg.src) * p.llr_stride + sg.start;
            for (int g0 = lo + lane; g0 < hi; g0 += 128) {           // four loads in flight per lane
                float xv[4];
#pragma unroll
                for (int u = 0; u < 4; u++) { const int g = g0 + 32 * u; xv[u] = g < hi ? __ldcs(&src[g - base]) : 0.f; }
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    const int g = g0 + 32 * u;
                    if (g < hi) {
                        const float x = -1 * xv[u];
                        const int col = c.permute()[g - rcv_idx];
                        rx[col] = x;
                        est[col] = x < 0.f ? 1 : 0;
                    }
                }
            }
        }
        base += sg.len;
    }
    __syncwarp();
}

#define DTL_STAGE_ALL_CODES(p, scodes, sm)                                                         \
    {                                                                                              \
        size_t off_ = 0;                                                                           \
        for (int ci_ = 1; ci_ <= (p).n_codes; ci_++) {                                             \
            const CodeDev g_ = (p).codes[ci_];                                                     \
            const CodeS sc_ = stage_code(g_, (sm), (int)off_);                                         \
            if (threadIdx.x == 0) (scodes)[ci_] = sc_;                                             \
            off_ += code_tab_bytes(g_.n, g_.m, g_.max_row * g_.m, g_.edges);                       \
        }                                                                                          \
        __syncthreads();                                                                           \
    }

// Pass 1: awgn_bp's exit before the first iteration (zero syndrome of the hard decisions, nit = 0), BIT-SLICED: a warp
// screens 32 codewords at a time.
//   * lane l fetches the descriptors of codeword l of the batch (one round of dependent loads for 32 codewords);
//   * the hard decision of wire position q (ldpc_dec.cc:64-66: rx = -llr, bit = rx < 0, i.e. llr > 0; shortened and
//     missing positions are 0) is one ballot bit of a coalesced load: est[cw][ceil(n/32)] words in shared memory;
//   * a 32x32 bit transpose per word column turns them into planes: plane[q] holds bit q of all 32 codewords, so one
//     XOR per graph edge (CodeDev::row_wire: the graph with cldpc's column permutation folded in) checks a parity
//     equation of 32 codewords at once, and the OR over the rows is the mask of codewords that need belief propagation;
//   * the systematic bits est[permute[rank + i]] of a clean codeword are simply wire bits rank.. (get_systematic_bits).
// Batches that mix codes, and codewords that straddle two LLR segments, take the per-codeword path with the same tables.
struct ScreenCode { int n, m, k, rank, max_row, tab; };

__device__ __forceinline__ unsigned transpose32(unsigned x, int lane)
{
    unsigned m = 0x0000ffffu;
#pragma unroll
    for (int j = 16; j != 0; j >>= 1, m ^= (m << j)) {
        const unsigned y = __shfl_xor_sync(0xffffffffu, x, j);
        x = (lane & j) ? ((x & (m << j)) | ((y >> j) & m)) : ((x & m) | ((y & m) << j));
    }
    return x;
}

// The record of a codeword the screening pass queues for belief propagation (BpItem), written at queue position `pos` (class *
// cw_cap + rank in the class).  A codeword whose received run lies inside one segment (one frame's LLR row, or the carry buffer)
// is not copied at all: the record points the decoder's bulk copy at the 16-byte-aligned superset of the run (`shift` floats of
// lead-in) as long as that superset stays inside the row.  The others (runs that straddle two frames, TBs cut short) take a
// block of cwbuf, which bp_gather_kernel fills.
__device__ __forceinline__ void bp_make_item(const FecParams& p, int it, size_t pos)
{
    const TbDesc& d = p.tbs[p.cw_tb[it]];
    const CodeDev& c = p.codes[d.fec_id];
    const int rcv_idx = p.cw_geo[3 * it], k_ = p.cw_geo[3 * it + 2], len = (c.n - c.k) + k_;
    BpItem r;
    r.out_off = d.bit_base + p.cw_geo[3 * it + 1]; r.it = it; r.fec_id = d.fec_id; r.k_ = k_;
    r.src = nullptr; r.shift = 0; r.nbytes = 0; r.len = len;
    const int nseg = d.nseg;
    int base = 0;
    for (int i = 0; i < nseg; i++) {
        const TbSeg sg = d.seg[i];
        if (rcv_idx >= base && rcv_idx + len <= base + sg.len) {
            const float* row = sg.src < 0 ? p.carry + (size_t)d.stream * p.carry_cap
                                          : p.llr + ((size_t)d.stream * p.F + sg.src) * p.llr_stride;
            const size_t row_bytes = (size_t)(sg.src < 0 ? p.carry_cap : p.llr_stride) * 4;
            const uintptr_t a = (uintptr_t)(row + sg.start + (rcv_idx - base));
            const uintptr_t a0 = a & ~(uintptr_t)15, a1 = (a + (size_t)len * 4 + 15) & ~(uintptr_t)15;
            if (a0 >= (uintptr_t)row && a1 <= (uintptr_t)row + row_bytes) {
                r.src = (const float*)a0; r.shift = (int)((a - a0) >> 2); r.nbytes = (int)(a1 - a0);
            }
            break;
        }
        base += sg.len;
    }
    if (!r.src) {
        const int g = atomicAdd(&p.cnt[10], 1);
        p.glist[g] = (int)pos;
        r.src = p.cwbuf + (size_t)g * p.cw_stride; r.nbytes = p.cw_stride * 4; r.len = c.n;
    }
    p.items[pos] = r;
}

#ifndef DTL_SCREEN_UNROLL
#define DTL_SCREEN_UNROLL 5      // loads in flight per lane (A/B: 4 -> 2.143 ms LDPC stage, 5 -> 2.090, 10 -> 2.099)
#endif
__global__ void __launch_bounds__(256) ldpc_screen_kernel(FecParams p)
{
    extern __shared__ __align__(16) unsigned char sm[];
    __shared__ ScreenCode s_code[DTL_MAX_SMEM_CODES];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, warps = blockDim.x >> 5;
    {
        int off = 0;
        for (int ci = 1; ci <= p.n_codes; ci++) {
            const CodeDev g = p.codes[ci];
            const int slots = g.max_row * g.m;
            uint16_t* rl = (uint16_t*)(sm + off); uint16_t* rw = rl + g.m;
            for (int i = threadIdx.x; i < g.m; i += blockDim.x) rl[i] = g.row_len[i];
            for (int i = threadIdx.x; i < slots; i += blockDim.x) rw[i] = g.row_wire[i];      // pad slots hold 32*ceil(n/32)
            if (threadIdx.x == 0) { ScreenCode sc; sc.n = g.n; sc.m = g.m; sc.k = g.k; sc.rank = g.rank; sc.max_row = g.max_row; sc.tab = off; s_code[ci] = sc; }
            off += (2 * (g.m + slots) + 15) & ~15;
        }
        __syncthreads();
    }
    const int total = min(p.cnt[0], p.cw_cap);
    const int nwp = p.screen_words + 1;                                   // est row: ceil(n/32) words + one that stays 0
    uint32_t* estw = (uint32_t*)(sm + p.screen_tab_bytes + (size_t)wid * p.screen_bytes);      // [32][nwp]
    uint32_t* plane = estw + 32 * nwp;                                    // [32*screen_words + 1]
    for (int it0 = (blockIdx.x * warps + wid) * 32; it0 < total; it0 += gridDim.x * warps * 32) {
        const int cnt = min(32, total - it0);
        int my_fec = 0, my_k = 0;
        long long my_out = 0;
        if (p.sgn) {
            // ---- hard decisions from the sign words the equaliser kernel wrote next to the LLRs (16 MB instead of the 508 MB
            // of LLRs this pass used to read): lane l builds the row of codeword it0 + l on its own -- per segment that
            // overlaps the codeword, 32 stream bits per word with one funnel shift, masked to the segment's range
            uint32_t* row = estw + lane * nwp;
            for (int w = 0; w < nwp; w++) row[w] = 0u;
            if (lane < cnt) {
                const int it = it0 + lane;
                const TbDesc& d = p.tbs[p.cw_tb[it]];
                my_fec = d.fec_id;
                const ScreenCode sc = s_code[my_fec];
                const int rcv_idx = p.cw_geo[3 * it], nw = (sc.n + 31) >> 5;
                my_k = p.cw_geo[3 * it + 2];
                my_out = d.bit_base + p.cw_geo[3 * it + 1];
                const int len = (sc.n - sc.k) + my_k, nseg = d.nseg;
                int base = 0;
                for (int i = 0; i < nseg; i++) {
                    const TbSeg sg = d.seg[i];
                    const int lo = max(rcv_idx, base), hi = min(rcv_idx + len, base + sg.len);
                    if (lo < hi) {
                        const uint32_t* sw = sg.src < 0 ? p.csgn + (size_t)d.stream * p.csgn_stride
                                                        : p.sgn + ((size_t)d.stream * p.F + sg.src) * p.sgn_stride;
                        const int nwords = sg.src < 0 ? p.csgn_stride : p.sgn_stride;
                        const int e0 = sg.start + (rcv_idx - base);          // stream index of codeword position 0
                        const int p0 = lo - rcv_idx, p1 = hi - rcv_idx;
                        for (int w = p0 >> 5; w < nw && 32 * w < p1; w++) {
                            const int bitpos = e0 + 32 * w, wi = bitpos >> 5, sh = bitpos & 31;
                            const uint32_t a = (wi >= 0 && wi < nwords) ? __ldg(&sw[wi]) : 0u;
                            const uint32_t b = (wi + 1 >= 0 && wi + 1 < nwords) ? __ldg(&sw[wi + 1]) : 0u;
                            const int l0 = max(p0 - 32 * w, 0), l1 = min(p1 - 32 * w, 32);
                            const uint32_t m = (l1 >= 32 ? 0xffffffffu : (1u << l1) - 1u) & ~((1u << l0) - 1u);
                            row[w] |= __funnelshift_r(a, b, sh) & m;
                        }
                    }
                    base += sg.len;
                }
            }
        } else {
        // ---- descriptors of codeword it0 + lane
        int my_p0 = 0, my_p1 = 0, my_single = 1;
        const float* my_src = nullptr;
        if (lane < cnt) {
            const int it = it0 + lane;
            const TbDesc& d = p.tbs[p.cw_tb[it]];
            my_fec = d.fec_id;
            const ScreenCode sc = s_code[my_fec];
            const int rcv_idx = p.cw_geo[3 * it];
            my_k = p.cw_geo[3 * it + 2];
            my_out = d.bit_base + p.cw_geo[3 * it + 1];
            const int len = (sc.n - sc.k) + my_k, nseg = d.nseg;
            int base = 0, hits = 0;
            for (int i = 0; i < nseg; i++) {
                const TbSeg sg = d.seg[i];
                const int lo = max(rcv_idx, base), hi = min(rcv_idx + len, base + sg.len);
                if (lo < hi) {
                    hits++;
                    const float* src = sg.src < 0 ? p.carry + (size_t)d.stream * p.carry_cap + sg.start
                                                  : p.llr + ((size_t)d.stream * p.F + sg.src) * p.llr_stride + sg.start;
                    my_src = src + (rcv_idx - base);                     // my_src[pos]: LLR of codeword position pos
                    my_p0 = lo - rcv_idx; my_p1 = hi - rcv_idx;
                }
                base += sg.len;
            }
            my_single = hits <= 1;
        }
        // ---- hard decisions, one ballot word per 32 wire positions
        for (int l = 0; l < cnt; l++) {
            const int fec = __shfl_sync(0xffffffffu, my_fec, l);
            const int nw = (s_code[fec].n + 31) >> 5;
            uint32_t* row = estw + l * nwp;
            if (__shfl_sync(0xffffffffu, my_single, l)) {
                const float* src = (const float*)__shfl_sync(0xffffffffu, (unsigned long long)my_src, l);
                const int p0 = __shfl_sync(0xffffffffu, my_p0, l), p1 = __shfl_sync(0xffffffffu, my_p1, l);
                for (int cb = 0; cb < nw; cb += DTL_SCREEN_UNROLL) {     // independent loads in flight per lane
                    float x[DTL_SCREEN_UNROLL];
#pragma unroll
                    for (int u = 0; u < DTL_SCREEN_UNROLL; u++) {
                        const int pos = (cb + u) * 32 + lane;
                        x[u] = (pos >= p0 && pos < p1) ? __ldcs(&src[pos]) : 0.f;
                    }
#pragma unroll
                    for (int u = 0; u < DTL_SCREEN_UNROLL; u++) {
                        const unsigned word = __ballot_sync(0xffffffffu, x[u] > 0.f);
                        if (lane == 0 && cb + u < nw) row[cb + u] = word;
                    }
                }
                if (lane == 0) row[nw] = 0u;
            } else {                                                      // the codeword straddles LLR segments
                const int it = it0 + l;
                const TbDesc& d = p.tbs[p.cw_tb[it]];
                const ScreenCode sc = s_code[fec];
                const int rcv_idx = p.cw_geo[3 * it], len = (sc.n - sc.k) + p.cw_geo[3 * it + 2];
                for (int w = lane; w <= nw; w += 32) row[w] = 0u;
                __syncwarp();
                int base = 0;
                for (int i = 0; i < d.nseg; i++) {
                    const TbSeg sg = d.seg[i];
                    const int lo = max(rcv_idx, base), hi = min(rcv_idx + len, base + sg.len);
                    if (lo < hi) {
                        const float* src = sg.src < 0 ? p.carry + (size_t)d.stream * p.carry_cap + sg.start
                                                      : p.llr + ((size_t)d.stream * p.F + sg.src) * p.llr_stride + sg.start;
                        src += rcv_idx - base;
                        const int p0 = lo - rcv_idx, p1 = hi - rcv_idx;
                        for (int cb = p0 & ~31; cb < p1; cb += 32) {
                            const int pos = cb + lane;
                            const bool bit = pos >= p0 && pos < p1 && __ldcs(&src[pos]) > 0.f;
                            const unsigned word = __ballot_sync(0xffffffffu, bit);
                            if (lane == 0 && word) row[cb >> 5] |= word;
                        }
                    }
                    base += sg.len;
                }
            }
        }
        }
        const int fec0 = __shfl_sync(0xffffffffu, my_fec, 0);
        const bool uniform = __all_sync(0xffffffffu, lane >= cnt || my_fec == fec0);
        __syncwarp();
        // ---- syndromes
        unsigned badw = 0;
        int heavy = 32;
        if (uniform) {
            const ScreenCode sc = s_code[fec0];
            const int nw = (sc.n + 31) >> 5;
            for (int w = 0; w < nw; w++) plane[w * 32 + lane] = transpose32(lane < cnt ? estw[lane * nwp + w] : 0u, lane);
            if (lane == 0) plane[nw * 32] = 0u;
            __syncwarp();
            const uint16_t* rw = (const uint16_t*)(sm + sc.tab) + sc.m;
            for (int r = lane; r < sc.m; r += 32) {
                const uint16_t* rwr = rw + r;
                unsigned par = 0;
#pragma unroll 3
                for (int j = 0; j < sc.max_row; j++) par ^= plane[rwr[j * sc.m]];
                badw |= par;
            }
            // how hard each codeword looks: the number of lanes that saw a failing check of it (a coarse syndrome weight)
#pragma unroll
            for (int c = 0; c < 32; c++) {
                const unsigned b = __ballot_sync(0xffffffffu, (badw >> c) & 1u);
                if (lane == c) heavy = __popc(b);
            }
            badw = __reduce_or_sync(0xffffffffu, badw);
        } else {
            for (int l = 0; l < cnt; l++) {
                const ScreenCode sc = s_code[__shfl_sync(0xffffffffu, my_fec, l)];
                const uint16_t* rw = (const uint16_t*)(sm + sc.tab) + sc.m;
                const uint32_t* row = estw + l * nwp;
                bool bad = false;
                for (int r = lane; r < sc.m; r += 32) {
                    unsigned par = 0;
                    for (int j = 0; j < sc.max_row; j++) { const unsigned v = rw[j * sc.m + r]; par ^= row[v >> 5] >> (v & 31u); }
                    bad |= (par & 1u) != 0;
                }
                if (__any_sync(0xffffffffu, bad)) badw |= 1u << l;
            }
        }
        if (cnt < 32) badw &= (1u << cnt) - 1u;
        // ---- clean codewords: systematic bits out; the rest: queued for the belief-propagation pass
        for (int l = 0; l < cnt; l++) {
            if ((badw >> l) & 1u) continue;
            const int rank = s_code[__shfl_sync(0xffffffffu, my_fec, l)].rank, k_ = __shfl_sync(0xffffffffu, my_k, l);
            uint8_t* out = p.bits + __shfl_sync(0xffffffffu, my_out, l);
            const uint32_t* row = estw + l * nwp;
            for (int q = lane; q < k_; q += 32) { const unsigned v = rank + q; out[q] = (uint8_t)((row[v >> 5] >> (v & 31u)) & 1u); }
        }
        if (lane < cnt && !((badw >> lane) & 1u)) p.cw_nit[it0 + lane] = 0;
        // four lists, hardest-looking first: the decoder takes them in that order, so the codewords that run all
        // max_it iterations start early and the short ones fill the end of the launch
        if (badw) {
            const bool mine = (badw >> lane) & 1u;
            const int bucket = heavy >= p.bp_th[0] ? 0 : heavy >= p.bp_th[1] ? 1 : heavy >= p.bp_th[2] ? 2 : 3;
#pragma unroll
            for (int b = 0; b < 4; b++) {
                const unsigned mk = __ballot_sync(0xffffffffu, mine && bucket == b);
                if (!mk) continue;
                int base = 0;
                if (lane == 0) base = atomicAdd(&p.cnt[b == 0 ? 3 : 4 + b], __popc(mk));
                base = __shfl_sync(0xffffffffu, base, 0);
                if ((mk >> lane) & 1u) {
                    const size_t pos = (size_t)b * p.cw_cap + base + __popc(mk & ((1u << lane) - 1u));
                    p.bp_list[pos] = it0 + lane;
                    if (p.items) bp_make_item(p, it0 + lane, pos);
                }
            }
        }
        __syncwarp();
    }
}

// Pass 2, persistent: each warp pulls queued codewords from a device counter and runs the decoder proper.
#ifndef DTL_BP_WARPS
#define DTL_BP_WARPS 4
#define DTL_BP_MINB 5
#endif
__global__ void __launch_bounds__(32 * DTL_BP_WARPS, DTL_BP_MINB) ldpc_bp_kernel(FecParams p)
{
    extern __shared__ __align__(16) unsigned char sm[];
    __shared__ CodeS scodes[DTL_MAX_SMEM_CODES];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    DTL_STAGE_ALL_CODES(p, scodes, sm)
    unsigned char* base = sm + p.tab_bytes + (size_t)wid * p.warp_bytes;
    const int n0 = p.cnt[3], n1 = n0 + p.cnt[5], n2 = n1 + p.cnt[6], total = n2 + p.cnt[7];
    for (;;) {
        int j = 0;
        if (lane == 0) j = atomicAdd(&p.cnt[1], 1);
        j = __shfl_sync(0xffffffffu, j, 0);
        if (j >= total) break;
        const int it = j < n0 ? p.bp_list[j] : j < n1 ? p.bp_list[(size_t)p.cw_cap + (j - n0)]
                     : j < n2 ? p.bp_list[2 * (size_t)p.cw_cap + (j - n1)] : p.bp_list[3 * (size_t)p.cw_cap + (j - n2)];
        const TbDesc& d = p.tbs[p.cw_tb[it]];
        CodeS c = scodes[d.fec_id];
        c.smb = sm;
        const WarpBuf w = carve(base, c.n, c.slots);
        gather_cw(p, d, c, it, w.rx, w.est, lane);
        const int nit = p.mode == DTL_LDPC_LAYERED ? bp_layered_run(c, w, lane, p.max_it)
                                                   : bp_reference_run(c, w, lane, p.max_it, p.sigma2);
        const int d_idx = p.cw_geo[3 * it + 1], k_ = p.cw_geo[3 * it + 2];
        uint8_t* out = p.bits + d.bit_base + d_idx;
        for (int q = lane; q < k_; q += 32) out[q] = w.est[c.permute()[c.rank + q]];
        if (lane == 0) p.cw_nit[it] = nit;
        __syncwarp();
    }
}


// ---- reference mode, row-major layout (bp_ref.cuh) ------------------------------------------------------------------
// Gather of one codeword straight into what awgn_bp starts from: lr0[col] = exp(-2 rx / sigma^2) with rx = -llr
// (ldpc_dec.cc:64-66, rx_lr_calc) and est[col] = rx < 0 (compute_init_estimate); positions past the received length read
// as LLR 0 (lr = exp(0) = 1), shortened positions as -15 (tb_decoder.cc:145,161-183).
// sigma2 > 0: divide; sigma2 < 0: -sigma2 is an exact power of two and 1/sigma2 = -1/(-sigma2) is passed instead -- the quotient
// by a power of two is the product by its reciprocal, bit for bit (the default sigma^2 = 2: one multiplication, no division)
__device__ __forceinline__ double lr_of_rx(float x, double sigma2, const double2* __restrict__ T)
{
    const float m2 = -1 * 2 * x;
    double q;
    if (sigma2 < 0.0) q = __dmul_rn((double)m2, -sigma2);
    else {
        bool ok;
        q = ddiv_try((double)m2, sigma2, ok);
        if (__builtin_expect(!ok, 0)) q = __ddiv_rn((double)m2, sigma2);
    }
    return exp_tab(q, T);
}

// Gather of the queued codewords, one warp each, grid-stride: the n received values of the TB's receive buffer (reads past
// the received length = 0, shortened positions = -15, tb_decoder.cc:145,161-183), negated and permuted into H column order
// (ldpc_dec.cc:64-66), written as one contiguous block per queued codeword -- the decoder then fetches a codeword with ONE
// bulk copy issued a whole decode ahead, instead of walking descriptors -> segments -> LLRs with every load waiting for the
// one before it (profiles/r02a: a third of the decoder's warps sat in that chain at any time).
__global__ void __launch_bounds__(256) bp_gather_kernel(FecParams p)
{
    // one WARP per queued codeword that needs gathering (bp_make_item): the pieces go into its block of cwbuf, wire order,
    // coalesced: defaults first (positions past the received length read as LLR 0, shortened ones as -15), then every segment
    // that overlaps the codeword copies its run.  Values as received: the decoder negates (ldpc_dec.cc:64) and applies the
    // column permutation when it consumes the block
    const int lane = threadIdx.x & 31, warps = blockDim.x >> 5;
    const int nl = p.cnt[10];
    for (int li = blockIdx.x * warps + (threadIdx.x >> 5); li < nl; li += gridDim.x * warps) {
        const int it = p.bp_list[p.glist[li]];
        const TbDesc& d = p.tbs[p.cw_tb[it]];
        const CodeDev& c = p.codes[d.fec_id];
        const int rcv_idx = p.cw_geo[3 * it], len = (c.n - c.k) + p.cw_geo[3 * it + 2];
        float* dst = p.cwbuf + (size_t)li * p.cw_stride;
        for (int q = lane; q < p.cw_stride; q += 32) dst[q] = q < len ? 0.0f : -15.0f;
        __syncwarp();
        const int nseg = d.nseg;
        int base = 0;
        for (int i = 0; i < nseg; i++) {
            const TbSeg sg = d.seg[i];
            const int lo = max(rcv_idx, base), hi = min(rcv_idx + len, base + sg.len);
            if (lo < hi) {
                const float* src = sg.src < 0 ? p.carry + (size_t)d.stream * p.carry_cap + sg.start
                                              : p.llr + ((size_t)d.stream * p.F + sg.src) * p.llr_stride + sg.start;
                for (int g = lo + lane; g < hi; g += 32) dst[g - rcv_idx] = __ldcs(&src[g - base]);
            }
            base += sg.len;
        }
    }
}

// bp_ref_run behind a call: the decoder's inner loops get a register allocation of their own, the kernel's per-codeword
// state (queue index, item record, pointers) is saved around ONE call per codeword instead of competing with the
// division chains for the 64 registers a thread has when 32 warps are resident
template <int NT>
__device__ __noinline__ int bp_ref_call(const FastS* cp, unsigned char* base, const unsigned char* smb, int tid, int max_it, int bar_id)
{
    FastS c = *cp;
    c.smb = smb;
    const FastBuf w = fast_carve(base, c.n, c.m, c.rs);
    return bp_ref_run<NT>(c, w, tid, max_it, bar_id);
}

// queue position j (hardest class first) -> index of the codeword's record: class * cw_cap + rank in the class
__device__ __forceinline__ size_t bp_item_pos(const int* q, int j, int cap)
{
    return j < q[0] ? (size_t)j : j < q[1] ? (size_t)cap + (j - q[0]) : j < q[2] ? 2 * (size_t)cap + (j - q[1]) : 3 * (size_t)cap + (j - q[2]);
}

// Pass 2, reference mode: persistent, one codeword per TEAM of NT threads (NT / 32 warps behind a named barrier) from the
// queue.  What bounds this kernel is how many warps the SM can choose from, not instruction count: one warp issues an FP64
// instruction with three register-pair operands about every six cycles however many independent chains it holds (tools/
// div_bench.cu, tools/bp_bench.cu on a B200), the FP64 pipe takes one every two, and the shared memory a codeword needs
// (11 KB) caps the codewords in flight at 19 per SM.  Default: NT = 32, ONE CTA of 19 warps per SM (the code tables staged
// once).  NT = 64 (DTL_BP_TEAM=64: two warps share a codeword's rows and columns, 64 registers per thread keep 32 warps
// resident) reaches the pipe's peak in the microbenchmark and loses in the kernel to spills in the division chains.
// Each team owns a staging buffer and an mbarrier; the run of the NEXT codeword is fetched by one bulk copy
// (cp.async.bulk -> mbarrier) from the address in its record (BpItem: the LLR row itself, or the gathered block) as soon as
// the current one has been turned into likelihood ratios.
template <int NT, int TEAMS, int MINB>
__global__ void __launch_bounds__(NT * TEAMS, MINB) ldpc_bp_ref_kernel(FecParams p)
{
    extern __shared__ __align__(16) unsigned char sm[];
    __shared__ FastS scodes[DTL_MAX_SMEM_CODES];
    __shared__ __align__(8) uint64_t bars[TEAMS];
    __shared__ int s_next[TEAMS];
    __shared__ int s_q[3];
    const int team = threadIdx.x / NT, tid = threadIdx.x - team * NT, bar_id = team + 1;
    {
        int off = 0;
        for (int ci = 1; ci <= p.n_codes; ci++) {
            const CodeDev g = p.codes[ci];
            const FastS sc = stage_fast(g, sm, off);
            if (threadIdx.x == 0) scodes[ci] = sc;
            off += (int)fast_tab_bytes(g.n, g.m, g.nvr);
        }
        if (tid == 0) { mbar_init(&bars[team], 1); s_next[team] = atomicAdd(&p.cnt[1], 1); }
        if (threadIdx.x == 0) { s_q[0] = p.cnt[3]; s_q[1] = s_q[0] + p.cnt[5]; s_q[2] = s_q[1] + p.cnt[6]; }
        __syncthreads();
    }
    const int stg_bytes = p.cw_stride * 4 + DTL_STG_SLACK;
    unsigned char* base = sm + p.fast_tab_bytes + (size_t)team * (p.fast_warp_bytes + stg_bytes);
    float* stg = (float*)(base + p.fast_warp_bytes);
    uint64_t* bar = &bars[team];
    const int total = p.cnt[3] + p.cnt[5] + p.cnt[6] + p.cnt[7];
    int j = s_next[team];
    BpItem rec;
    rec.out_off = 0; rec.src = nullptr; rec.it = 0; rec.fec_id = 1; rec.k_ = 0; rec.shift = 0; rec.nbytes = 0; rec.len = 0;
    if (j < total) {
        rec = p.items[bp_item_pos(s_q, j, p.cw_cap)];
        if (tid == 0) { mbar_expect_tx(bar, (uint32_t)rec.nbytes); bulk_g2s(stg, rec.src, (uint32_t)rec.nbytes, bar); }
    }
    team_sync<NT>(bar_id);                                             // everybody has read s_next
    uint32_t phase = 0;
    while (j < total) {
        // the claim of the next codeword: issued here, first looked at after the likelihood-ratio pass (one warp per codeword:
        // through a shuffle, not through shared memory -- a store of the atomic's result would wait for it right here)
        int nx = 0;
        if (tid == 0) { nx = atomicAdd(&p.cnt[1], 1); if (NT != 32) s_next[team] = nx; }
        FastS c = scodes[rec.fec_id];
        c.smb = sm;
        const FastBuf w = fast_carve(base, c.n, c.m, c.rs);
        mbar_wait(bar, phase);
        phase ^= 1u;
        for (int q = tid; q < c.n; q += NT) {                          // ldpc_dec.cc:64-66, rx_lr_calc, compute_init_estimate
            const float x = q < rec.len ? -1 * stg[rec.shift + q] : 15.0f;      // shortened positions: -1 * -15
            const int col = c.permute()[q];
            w.lr0[col] = lr_of_rx(x, p.sigma2_lr, p.exp_tab);
            w.est[col] = x < 0.f ? 1 : 0;
        }
        team_sync<NT>(bar_id);
        const int nxt = NT == 32 ? __shfl_sync(0xffffffffu, nx, 0) : s_next[team];
        BpItem nrec = rec;
        if (nxt < total) {
            nrec = p.items[bp_item_pos(s_q, nxt, p.cw_cap)];
            if (tid == 0) { mbar_expect_tx(bar, (uint32_t)nrec.nbytes); bulk_g2s(stg, nrec.src, (uint32_t)nrec.nbytes, bar); }
        }
        const int nit = NT == 32 ? bp_ref_run<NT>(c, w, tid, p.max_it, bar_id) : bp_ref_call<NT>(&scodes[rec.fec_id], base, sm, tid, p.max_it, bar_id);
        uint8_t* out = p.bits + rec.out_off;
        for (int q = tid; q < rec.k_; q += NT) out[q] = w.est[c.permute()[c.rank + q]];
        if (tid == 0) p.cw_nit[rec.it] = nit;
        team_sync<NT>(bar_id);
        j = nxt; rec = nrec;
    }
}

// Pass 2, layered mode (bp_layered.cuh): the same queue, records and bulk-copy prefetch as ldpc_bp_ref_kernel, one
// codeword per warp; the block's received values ARE the initial posteriors (no likelihood-ratio pass).
template <int W, int MINB>
__global__ void __launch_bounds__(32 * W, MINB) ldpc_bp_lay_kernel(FecParams p)
{
    extern __shared__ __align__(16) unsigned char sm[];
    __shared__ LayS scodes[DTL_MAX_SMEM_CODES];
    __shared__ __align__(8) uint64_t bars[W];
    __shared__ int s_q[3];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    {
        int off = 0;
        for (int ci = 1; ci <= p.n_codes; ci++) {
            const CodeDev g = p.codes[ci];
            const LayS sc = stage_lay(g, sm, off);
            if (threadIdx.x == 0) scodes[ci] = sc;
            off += (int)lay_tab_bytes(g.n, g.m, g.lay_steps);
        }
        if (lane == 0) mbar_init(&bars[wid], 1);
        if (threadIdx.x == 0) { s_q[0] = p.cnt[3]; s_q[1] = s_q[0] + p.cnt[5]; s_q[2] = s_q[1] + p.cnt[6]; }
        __syncthreads();
    }
    const int stg_bytes = p.cw_stride * 4 + DTL_STG_SLACK;
    unsigned char* base = sm + p.lay_tab_bytes + (size_t)wid * (p.lay_warp_bytes + stg_bytes);
    float* stg = (float*)(base + p.lay_warp_bytes);
    uint64_t* bar = &bars[wid];
    const int total = p.cnt[3] + p.cnt[5] + p.cnt[6] + p.cnt[7];
    int j = 0;
    if (lane == 0) j = atomicAdd(&p.cnt[1], 1);
    j = __shfl_sync(0xffffffffu, j, 0);
    BpItem rec;
    rec.out_off = 0; rec.src = nullptr; rec.it = 0; rec.fec_id = 1; rec.k_ = 0; rec.shift = 0; rec.nbytes = 0; rec.len = 0;
    if (j < total) {
        rec = p.items[bp_item_pos(s_q, j, p.cw_cap)];
        if (lane == 0) { mbar_expect_tx(bar, (uint32_t)rec.nbytes); bulk_g2s(stg, rec.src, (uint32_t)rec.nbytes, bar); }
    }
    uint32_t phase = 0;
    while (j < total) {
        int nxt = 0;
        if (lane == 0) nxt = atomicAdd(&p.cnt[1], 1);
        LayS c = scodes[rec.fec_id];
        c.smb = sm;
        const LayBuf w = lay_carve(base, c.ns, c.nsteps);
        mbar_wait(bar, phase);
        phase ^= 1u;
        if (lane == 0) w.post[c.ns] = 1.0e30f;                                      // pad edge of the lighter rows
        for (int q = lane; q < c.n; q += 32) w.post[c.permute()[q]] = q < rec.len ? -1 * stg[rec.shift + q] : 15.0f;   // ldpc_dec.cc:64-66
        __syncwarp();
        nxt = __shfl_sync(0xffffffffu, nxt, 0);
        BpItem nrec = rec;
        if (nxt < total) {
            nrec = p.items[bp_item_pos(s_q, nxt, p.cw_cap)];
            if (lane == 0) { mbar_expect_tx(bar, (uint32_t)nrec.nbytes); bulk_g2s(stg, nrec.src, (uint32_t)nrec.nbytes, bar); }
        }
        const int nit = bp_lay_run(c, w, lane, p.max_it);
        uint8_t* out = p.bits + rec.out_off;
        for (int q = lane; q < rec.k_; q += 32) out[q] = w.post[c.permute()[c.rank + q]] < 0.f ? 1 : 0;
        if (lane == 0) p.cw_nit[rec.it] = nit;
        __syncwarp();
        j = nxt; rec = nrec;
    }
}

__global__ void __launch_bounds__(256, 4) ldpc_plain_lay_kernel(LdpcParams p)
{
    extern __shared__ __align__(16) unsigned char sm[];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    LayS c = stage_lay(p.code, sm, 0);
    c.smb = sm;
    __syncthreads();
    const LayBuf w = lay_carve(sm + p.lay_tab_bytes + (size_t)wid * p.lay_warp_bytes, c.ns, c.nsteps);
    for (;;) {
        long long it = 0;
        if (lane == 0) it = atomicAdd(p.cursor, 1);
        it = __shfl_sync(0xffffffffu, it, 0);
        if (it >= p.count) break;
        const float* in = p.llr + it * c.n;
        for (int q = lane; q < c.n; q += 32) w.post[c.permute()[q]] = -1 * in[q];
        if (lane == 0) w.post[c.ns] = 1.0e30f;
        __syncwarp();
        int nit = 0;
        if (bp_lay_syndrome(c, w, lane)) nit = bp_lay_run(c, w, lane, p.max_it);
        uint8_t* out = p.bits + it * c.k;
        for (int q = lane; q < c.k; q += 32) out[q] = w.post[c.permute()[c.rank + q]] < 0.f ? 1 : 0;
        if (lane == 0 && p.nit) p.nit[it] = nit;
        __syncwarp();
    }
}

__global__ void __launch_bounds__(128, 4) ldpc_plain_ref_kernel(LdpcParams p)
{
    extern __shared__ __align__(16) unsigned char sm[];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    FastS c = stage_fast(p.code, sm, 0);
    c.smb = sm;
    __syncthreads();
    const FastBuf w = fast_carve(sm + p.fast_tab_bytes + (size_t)wid * p.fast_warp_bytes, c.n, c.m, c.rs);
    if (lane == 0) w.est[c.n] = 0;
    for (;;) {
        long long it = 0;
        if (lane == 0) it = atomicAdd(p.cursor, 1);
        it = __shfl_sync(0xffffffffu, it, 0);
        if (it >= p.count) break;
        const float* in = p.llr + it * c.n;
        for (int q = lane; q < c.n; q += 32) {                                      // ldpc_dec.cc:64-66
            const float x = -1 * in[q];
            const int col = c.permute()[q];
            w.lr0[col] = lr_of_rx(x, p.sigma2_lr, p.exp_tab);
            w.est[col] = x < 0.f ? 1 : 0;
        }
        __syncwarp();
        int nit = 0;
        if (syn_nonzero<8>(c, w.est, lane)) nit = bp_ref_run(c, w, lane, p.max_it);
        uint8_t* out = p.bits + it * c.k;
        for (int q = lane; q < c.k; q += 32) out[q] = w.est[c.permute()[c.rank + q]];
        if (lane == 0 && p.nit) p.nit[it] = nit;
        __syncwarp();
    }
}

__global__ void __launch_bounds__(128, 4) ldpc_plain_kernel(LdpcParams p)
{
    extern __shared__ __align__(16) unsigned char sm[];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const CodeS c = stage_code(p.code, sm, 0);
    __syncthreads();
    const WarpBuf w = carve(sm + p.tab_bytes + (size_t)wid * p.warp_bytes, c.n, c.slots);
    for (;;) {
        long long it = 0;
        if (lane == 0) it = atomicAdd(p.cursor, 1);
        it = __shfl_sync(0xffffffffu, it, 0);
        if (it >= p.count) break;
        const float* in = p.llr + it * c.n;
        for (int q = lane; q < c.n; q += 32) {                                      // ldpc_dec.cc:64-66
            const float x = -1 * in[q];
            const int col = c.permute()[q];
            w.rx[col] = x; w.est[col] = x < 0.f ? 1 : 0;
        }
        __syncwarp();
        int nit = 0;
        if (syndrome_nonzero(c, w.est, lane))
            nit = p.mode == DTL_LDPC_LAYERED ? bp_layered_run(c, w, lane, p.max_it) : bp_reference_run(c, w, lane, p.max_it, p.sigma2);
        uint8_t* out = p.bits + it * c.k;
        for (int q = lane; q < c.k; q += 32) out[q] = w.est[c.permute()[c.rank + q]];
        if (lane == 0 && p.nit) p.nit[it] = nit;
        __syncwarp();
    }
}

// ---- TB epilogue: bits -> bytes LSB-first, CRC-32 verify (fec_decoder_impl.cc:161-204) -------------------------
// One warp per released TB.  The TB's bits (one per byte) are packed in place, 32 bits per lane and round: a lane
// loads its 32 bit-bytes as eight words, folds each into a nibble with one multiply ((x & 0x01010101) * 0x01020408
// puts bit-byte i at bit 24+i without carries) and writes the packed word after the warp has finished reading the
// round (the bytes a round overwrites only hold bits of this or earlier rounds).  CRC-32 runs four bytes per step
// (slicing-by-4 tables in shared memory) on the words the lanes already hold.
__global__ void __launch_bounds__(256) tb_finish_kernel(FecParams p)
{
    const int lane = threadIdx.x & 31;
    const int t = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (t >= p.cnt[2]) return;
    const int gi = p.tb_list[t];
    const TbDesc& d = p.tbs[gi];
    uint8_t* bits = p.bits + d.bit_base;                       // 4-byte aligned (bit offsets are multiples of 4)
    const int user = d.payload_len - 32;
    int nit_sum = 0;
    for (int i = lane; i < d.ncws; i += 32) nit_sum += p.cw_nit[d.cw_first + i];
#pragma unroll
    for (int o = 16; o; o >>= 1) nit_sum += __shfl_xor_sync(0xffffffffu, nit_sum, o);
    dtl_tb_result r;
    r.tb_no = d.tb_no; r.crc_ok = 0; r.nbytes = -1; r.avg_it = nit_sum / d.ncws; r.frame_index = d.frame_index;
    if (user >= 0) {
        const uint32_t stored = __ballot_sync(0xffffffffu, bits[user + lane] & 1);   // little-endian CRC field
        __syncwarp();
        const int words = (user + 31) >> 5;
        const uint32_t* bw = (const uint32_t*)bits;
        for (int w0 = 0; w0 < words; w0 += 32) {
            const int w = w0 + lane;
            uint32_t word = 0;
            if (w < words) {
                const int nb = min(32, user - 32 * w);                          // valid bits of this word
#pragma unroll
                for (int q = 0; q < 8; q++) {
                    uint32_t x = 4 * q < nb ? __ldcg(&bw[8 * w + q]) & 0x01010101u : 0u;
                    if (nb - 4 * q < 4 && 4 * q < nb) x &= 0xffffffffu >> (8 * (4 - (nb - 4 * q)));
                    word |= (((x * 0x01020408u) >> 24) & 0xfu) << (4 * q);
                }
            }
            __syncwarp();
            if (w < words) ((uint32_t*)bits)[w] = word;
            __syncwarp();
        }
        r.crc_ok = (int)stored; r.nbytes = user;                // tb_gather_kernel (tb_crc_one) turns these into the verdict
    }
    if (lane == 0) p.tres[gi] = r;
}

// CRC-32 of a released TB by ONE THREAD over the packed words tb_finish_kernel left (slicing-by-4 tables in shared
// memory).  A warp that runs the byte-serial recurrence together does 32 times the work for the same chain length; in
// tb_gather_kernel the TBs of a stream share the CTA instead.  r.nbytes holds the TB's user bits (-1: none), r.crc_ok the
// stored CRC on entry; both become the verdict.
__device__ __forceinline__ void tb_crc_one(const FecParams& p, int gi, dtl_tb_result& r, const uint32_t (*crc_tab)[256])
{
    const int user = r.nbytes;
    if (user < 0) { r.crc_ok = 0; r.nbytes = 0; return; }
    const uint32_t stored = (uint32_t)r.crc_ok;
    const uint32_t* bw = (const uint32_t*)(p.bits + p.tbs[gi].bit_base);
    const int n0 = (user + 7) >> 3, full = n0 >> 2;
    uint32_t crc = 0xFFFFFFFFu;
    // the recurrence is serial but the message is not: sixteen words are fetched ahead (all loads in flight at once) while
    // the sixteen before them go through the tables -- a 2 KB TB of the scaled PHY otherwise pays one global-memory latency
    // per four words (0.18 ms per call on configuration 5)
    constexpr int B = 16;
    uint32_t nxt[B];
#pragma unroll
    for (int u = 0; u < B; u++) nxt[u] = u < full ? __ldcs(&bw[u]) : 0u;
    for (int i0 = 0; i0 < full; i0 += B) {
        uint32_t cur[B];
#pragma unroll
        for (int u = 0; u < B; u++) cur[u] = nxt[u];
#pragma unroll
        for (int u = 0; u < B; u++) nxt[u] = i0 + B + u < full ? __ldcs(&bw[i0 + B + u]) : 0u;
#pragma unroll
        for (int u = 0; u < B; u++) {
            if (i0 + u < full) {
                const uint32_t y = crc ^ cur[u];
                crc = crc_tab[3][y & 0xffu] ^ crc_tab[2][(y >> 8) & 0xffu] ^ crc_tab[1][(y >> 16) & 0xffu] ^ crc_tab[0][y >> 24];
            }
        }
    }
    if (n0 & 3) {
        const uint32_t x = bw[full];
        for (int b = 0; b < (n0 & 3); b++) crc = crc_tab[0][(crc ^ (x >> (8 * b))) & 0xffu] ^ (crc >> 8);
    }
    const int ok = ~crc == stored;
    r.crc_ok = ok;
    r.nbytes = ok ? user >> 3 : 0;
}

// One CTA per stream: concatenate the payload bytes of its CRC-passing TBs in release order
// (ofdm_adaptive_fec_pack_bb) and compact the per-TB monitor records.  Threads fetch the TB records in parallel, one
// thread lays the bytes out, then every warp copies its share of the TBs.
#define DTL_GATHER_MAXTB 1024
__global__ void __launch_bounds__(512) tb_gather_kernel(FecParams p)
{
    __shared__ int s_nb[DTL_GATHER_MAXTB], s_pos[DTL_GATHER_MAXTB + 1];
    __shared__ long long s_bb[DTL_GATHER_MAXTB];
    const int s = blockIdx.x;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, warps = blockDim.x >> 5;
    const int ntb = min(p.tb_count[s], DTL_GATHER_MAXTB);
    __shared__ uint32_t crc_tab[4][256];
    for (int i = threadIdx.x; i < 256; i += blockDim.x) {
        uint32_t c = p.crc_table[i];
        crc_tab[0][i] = c;
        for (int k = 1; k < 4; k++) { c = p.crc_table[c & 0xffu] ^ (c >> 8); crc_tab[k][i] = c; }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < ntb; i += blockDim.x) {
        const int gi = s * p.tb_slots + i;
        dtl_tb_result r = p.tres[gi];
        tb_crc_one(p, gi, r, crc_tab);
        p.tres[gi] = r;
        if (p.tb_out && i < p.tb_cap) p.tb_out[(size_t)s * p.tb_cap + i] = r;
        s_nb[i] = r.nbytes; s_bb[i] = p.tbs[gi].bit_base;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        int pos = 0, err = 0;
        for (int i = 0; i < ntb; i++) {
            s_pos[i] = pos;
            if (s_nb[i] > 0 && p.bytes && pos + s_nb[i] > p.bytes_cap) { err = ERR_BITS_CAP; s_nb[i] = 0 - s_nb[i]; }
            pos += s_nb[i] < 0 ? -s_nb[i] : s_nb[i];
        }
        s_pos[ntb] = pos;
        if (p.tb_out && p.tb_count[s] > p.tb_cap) err |= ERR_TB_SLOTS;
        if (p.tb_count[s] > DTL_GATHER_MAXTB) err |= ERR_TB_SLOTS;
        if (p.nbytes) p.nbytes[s] = pos;
        if (p.ntbs) p.ntbs[s] = p.tb_count[s];
        if (err) atomicOr(p.err, err);
    }
    __syncthreads();
    if (!p.bytes) return;
    for (int i = wid; i < ntb; i += warps) {
        const int n = s_nb[i];
        if (n <= 0) continue;
        const uint8_t* src = p.bits + s_bb[i];                 // 4-byte aligned (bit offsets are multiples of 4)
        uint8_t* dst = p.bytes + (size_t)s * p.bytes_cap + s_pos[i];
        // word stores whatever the destination's alignment (TB payloads have odd byte counts, so most start unaligned): up to
        // three head bytes, then destination word j = source bytes [head + 4j, head + 4j + 4) funnel-shifted out of two aligned
        // source words, then the tail bytes.  (Byte-wise copies of a 2 KB TB cost 61 dependent round trips per lane.)
        const uint32_t* sw = (const uint32_t*)src;
        const int head = min(n, (int)((4 - ((uintptr_t)dst & 3)) & 3));
        if (lane < head) dst[lane] = src[lane];
        const int nw = (n - head) >> 2;
        uint32_t* dw = (uint32_t*)(dst + head);
        const int sh = 8 * head;
#pragma unroll 4
        for (int w = lane; w < nw; w += 32) {
            const uint32_t lo = sw[w], hi = head ? sw[w + 1] : 0u;       // sw[w + 1] is inside the TB: bytes up to head + 4w + 3 < n
            dw[w] = __funnelshift_r(lo, hi, sh);
        }
        for (int b = head + (nw << 2) + lane; b < n; b += 32) dst[b] = src[b];
    }
}

// LLRs of the TB still in flight move behind what the carry buffer already holds.  grid = (streams, chunks): block (s, c)
// copies floats [c * CHUNK, (c + 1) * CHUNK) of the stream's pending run (a TB of the scaled PHY is 31 k LLRs: one warp per
// stream took 0.48 ms on 256 streams).  The sign words of the carry buffer (screening pass of the next call) follow in
// carry_sign_kernel, after every block of this one is done.
#define DTL_CARRY_CHUNK 4096
__global__ void __launch_bounds__(256) carry_save_kernel(FecParams p)
{
    const int s = blockIdx.x, c0 = blockIdx.y * DTL_CARRY_CHUNK;
    if (s >= p.n_streams) return;
    const TbDesc& d = p.pend[s];
    float* carry = p.carry + (size_t)s * p.carry_cap;
    int base = 0;
    for (int i = 0; i < d.nseg; i++) {
        const TbSeg sg = d.seg[i];
        if (sg.src >= 0) {
            const float* src = p.llr + ((size_t)s * p.F + sg.src) * p.llr_stride + sg.start;
            const int lo = max(base, c0), hi = min(base + sg.len, c0 + DTL_CARRY_CHUNK);
            for (int q = lo + threadIdx.x; q < hi; q += blockDim.x) carry[q] = src[q - base];
        }                                      // src < 0: the carry segment is always first and already in place
        base += sg.len;
    }
}

__global__ void __launch_bounds__(256) carry_sign_kernel(FecParams p)
{
    const int s = blockIdx.x, lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (s >= p.n_streams) return;
    const TbDesc& d = p.pend[s];
    int len = 0;
    for (int i = 0; i < d.nseg; i++) len += d.seg[i].len;
    const float* carry = p.carry + (size_t)s * p.carry_cap;
    uint32_t* cs = p.csgn + (size_t)s * p.csgn_stride;
    for (int q0 = (blockIdx.y * (blockDim.x >> 5) + wid) * 32; q0 < len; q0 += gridDim.y * blockDim.x) {
        const bool bit = q0 + lane < len && carry[q0 + lane] > 0.f;
        const unsigned word = __ballot_sync(0xffffffffu, bit);
        if (lane == 0) cs[q0 >> 5] = word;
    }
}

__global__ void fec_clear_kernel(int* cnt, unsigned long long* bits_used)
{
    cnt[0] = cnt[1] = cnt[2] = cnt[3] = 0; cnt[5] = cnt[6] = cnt[7] = 0; cnt[10] = 0;     // cnt[4] is the stand-alone decoder's cursor
    *bits_used = 0ull;
}

static void launch_carry_save(cudaStream_t st, const FecParams& p)
{
    const int chunks = (p.carry_cap + DTL_CARRY_CHUNK - 1) / DTL_CARRY_CHUNK;
    carry_save_kernel<<<dim3(p.n_streams, chunks), 256, 0, st>>>(p);
    if (p.csgn) carry_sign_kernel<<<dim3(p.n_streams, (chunks + 1) / 2), 256, 0, st>>>(p);
}

cudaError_t launch_fec(cudaStream_t st, const FecParams& p, int sm_count, int* launches, Prof* prof, int phase,
                       cudaStream_t side, cudaEvent_t ev)
{
    if (p.n_streams <= 0 || p.F <= 0) return cudaSuccess;
    Prof none; if (!prof) prof = &none;
    if (phase != 2) {
        prof->begin(DTL_STAGE_TBPLAN, 2);
        fec_clear_kernel<<<1, 1, 0, st>>>(p.cnt, p.bits_used);
        tb_plan_kernel<<<(p.n_streams + 3) / 4, 128, 0, st>>>(p);
        prof->end();
        if (launches) *launches += 2;
        if (phase == 1) return cudaGetLastError();
    }
    {   // pass 1: screening, grid-stride over the codeword work list
        const size_t smem = (size_t)p.screen_tab_bytes + 8 * (size_t)p.screen_bytes;
        prof->begin(DTL_STAGE_LDPC, 2);
        static int per_sm_screen = 0;                // CTAs per SM of the grid-stride screening pass (6 are resident at 40 registers)
        if (!per_sm_screen) { const char* e = getenv("DTL_SCREEN_CTAS"); per_sm_screen = e && atoi(e) > 0 ? atoi(e) : 8; }
        // a call with little work (the one-frame latency mode) does not launch -- and stage tables in -- a full machine's CTAs
        const long long scr_need = ((long long)p.cw_cap + 255) / 256;               // 32 codewords per warp, 8 warps per CTA
        const int scr_grid = (int)std::max(1ll, std::min((long long)sm_count * per_sm_screen, scr_need));
        ldpc_screen_kernel<<<scr_grid, 256, smem, st>>>(p);
    }
    if (p.fast) {   // pass 2, reference mode on codes with a row-major instance (bp_ref.cuh)
        static int wsel = 0;                         // DTL_BP_TEAM=32|64: threads per codeword (A/B; 64 = two warps per codeword)
        if (!wsel) { const char* e = getenv("DTL_BP_TEAM"); wsel = e && atoi(e) > 0 ? atoi(e) : 32; }
        bp_gather_kernel<<<(int)std::max(1ll, std::min((long long)sm_count * 8, ((long long)p.cw_cap + 255) / 256)), 256, 0, st>>>(p);
        const size_t per_warp = (size_t)p.fast_warp_bytes + (size_t)p.cw_stride * 4 + DTL_STG_SLACK;
        auto launch = [&](auto kern, int teams, int threads, int max_ctas) {
            const size_t sm_bytes = (size_t)p.fast_tab_bytes + per_warp * teams;
            int per_sm = (int)((227 * 1024) / (sm_bytes + 1024 + 256));
            if (per_sm < 1) per_sm = 1;
            if (per_sm > max_ctas) per_sm = max_ctas;
            if (sm_bytes + 256 > 227 * 1024) return cudaErrorInvalidConfiguration;
            const long long need = ((long long)p.cw_cap + teams - 1) / teams;
            kern<<<(int)std::max(1ll, std::min((long long)sm_count * per_sm, need)), threads, sm_bytes, st>>>(p);
            return cudaSuccess;
        };
        // One CTA of 19 warps per SM where 19 codewords fit (n = 300: 221 KB) -- the code tables are staged once per SM instead of
        // three times, which buys the nineteenth warp (LDPC stage 1.780 -> 1.749 ms); otherwise three CTAs of six.
        // DTL_BP_BIGCTA=0 forces the latter (A/B).
        static int big = -1;
        if (big < 0) { const char* e = getenv("DTL_BP_BIGCTA"); big = e ? (atoi(e) ? 1 : 0) : 1; }
        const bool fits19 = (size_t)p.fast_tab_bytes + per_warp * 19 + 2048 <= 227 * 1024;
        cudaError_t le = wsel == 32 ? (big && fits19 ? launch(ldpc_bp_ref_kernel<32, 19, 1>, 19, 608, 1) : launch(ldpc_bp_ref_kernel<32, 6, 3>, 6, 192, 3))
                                    : launch(ldpc_bp_ref_kernel<64, 4, 4>, 4, 256, 4);
        if (le != cudaSuccess) return le;
        if (launches) *launches += 1;
        prof->end();
    } else if (p.lay) {   // pass 2, layered mode on codes with row-major tables (bp_layered.cuh)
        bp_gather_kernel<<<(int)std::max(1ll, std::min((long long)sm_count * 8, ((long long)p.cw_cap + 255) / 256)), 256, 0, st>>>(p);
        const int W = 8;
        const size_t sm_bytes = (size_t)p.lay_tab_bytes + ((size_t)p.lay_warp_bytes + (size_t)p.cw_stride * 4 + DTL_STG_SLACK) * W;
        if (sm_bytes + 256 > 227 * 1024) return cudaErrorInvalidConfiguration;
        int per_sm = (int)((227 * 1024) / (sm_bytes + 1024 + 256));
        if (per_sm < 1) per_sm = 1;
        if (per_sm > 4) per_sm = 4;                  // 64 registers per thread: four CTAs of eight warps (5 / 6 CTAs at 48 / 40 registers measured 4 % slower)
        ldpc_bp_lay_kernel<8, 4><<<(int)std::max(1ll, std::min((long long)sm_count * per_sm, ((long long)p.cw_cap + W - 1) / W)), W * 32, sm_bytes, st>>>(p);
        if (launches) *launches += 1;
        prof->end();
    } else {   // pass 2: belief propagation on what did not pass, persistent
        const int warps = DTL_BP_WARPS;
        const size_t smem = (size_t)p.tab_bytes + (size_t)p.warp_bytes * warps;
        if (smem > 227 * 1024) return cudaErrorInvalidConfiguration;
        int per_sm = (int)((226 * 1024) / (smem + 1024));
        if (per_sm < 1) per_sm = 1;
        if (per_sm > 8) per_sm = 8;
        ldpc_bp_kernel<<<sm_count * per_sm, warps * 32, smem, st>>>(p);
        prof->end();
    }
    if (side && ev) {
        cudaEventRecord(ev, st);
        cudaStreamWaitEvent(side, ev, 0);
        launch_carry_save(side, p);
    }
    const int tb_max = p.n_streams * p.tb_slots;
    prof->begin(DTL_STAGE_TBFINISH, 3);
    tb_finish_kernel<<<(tb_max + 7) / 8, 256, 0, st>>>(p);
    tb_gather_kernel<<<p.n_streams, p.max_tb > 8192 ? 512 : 128, 0, st>>>(p);      // big TBs (scaled PHY): more warps for the copies
    if (!(side && ev)) launch_carry_save(st, p);
    prof->end();
    if (launches) *launches += p.csgn ? 6 : 5;
    return cudaGetLastError();
}

__global__ void cursor_clear_kernel(int* c) { *c = 0; }

// Function attributes are per device: dtl_b200_create calls this after cudaSetDevice for every handle.
cudaError_t init_kernels_fec()
{
    cudaError_t e = set_max_dyn_smem(ldpc_screen_kernel);
    if (e == cudaSuccess) e = set_max_dyn_smem(ldpc_bp_kernel);
    if (e == cudaSuccess) e = set_max_dyn_smem(ldpc_plain_kernel);
    if (e == cudaSuccess) e = set_max_dyn_smem(ldpc_bp_ref_kernel<32, 6, 3>);
    if (e == cudaSuccess) e = set_max_dyn_smem(ldpc_bp_ref_kernel<64, 4, 4>);
    if (e == cudaSuccess) e = set_max_dyn_smem(ldpc_bp_ref_kernel<32, 19, 1>);
    if (e == cudaSuccess) e = set_max_dyn_smem(ldpc_plain_ref_kernel);
    if (e == cudaSuccess) e = set_max_dyn_smem(ldpc_bp_lay_kernel<8, 4>);
    if (e == cudaSuccess) e = set_max_dyn_smem(ldpc_plain_lay_kernel);
    return e;
}

cudaError_t launch_ldpc(cudaStream_t st, const LdpcParams& p, int sm_count)
{
    if (p.count <= 0) return cudaSuccess;
    const int warps = 4;
    const size_t smem = (size_t)p.tab_bytes + (size_t)p.warp_bytes * warps;
    if (smem > 227 * 1024) return cudaErrorInvalidConfiguration;
    int per_sm = (int)((226 * 1024) / (smem + 1024));
    if (per_sm < 1) per_sm = 1;
    if (per_sm > 8) per_sm = 8;
    long long want = (p.count + warps - 1) / warps;
    int blocks = (int)(want < (long long)sm_count * per_sm ? want : (long long)sm_count * per_sm);
    cursor_clear_kernel<<<1, 1, 0, st>>>(p.cursor);
    if (p.fast) {
        const size_t fsmem = (size_t)p.fast_tab_bytes + (size_t)p.fast_warp_bytes * warps;
        if (fsmem > 227 * 1024) return cudaErrorInvalidConfiguration;
        ldpc_plain_ref_kernel<<<blocks, warps * 32, fsmem, st>>>(p);
    } else if (p.lay) {
        const size_t lsmem = (size_t)p.lay_tab_bytes + (size_t)p.lay_warp_bytes * 8;
        if (lsmem > 227 * 1024) return cudaErrorInvalidConfiguration;
        long long lb = (p.count + 7) / 8;
        if (lb > (long long)sm_count * 4) lb = (long long)sm_count * 4;
        ldpc_plain_lay_kernel<<<(unsigned)lb, 256, lsmem, st>>>(p);
    } else ldpc_plain_kernel<<<blocks, warps * 32, smem, st>>>(p);
    return cudaGetLastError();
}

// ---- ldpc_enc::encode (lib/dtl/ldpc_enc.cc:58-68; gr-fec cldpc::encode) -------------------------------------------
// Systematic encode of `count` blocks of k bits (one bit per byte): parity[i] = P[i] . data over GF(2); the wire
// order after gr-dtl's un-permute (ldpc_enc.cc:65-67) is [parity(rank) | data(k)].  One warp per codeword: the k
// data bits become kwords ballot words in shared memory, each lane then owns parity rows lane, lane+32, ...
__global__ void __launch_bounds__(256) ldpc_encode_kernel(CodeDev c, const uint8_t* __restrict__ data, long long count,
                                                           uint8_t* __restrict__ wire)
{
    extern __shared__ uint32_t encw[];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint32_t* dw = encw + wid * c.kwords;
    const long long cw = (long long)blockIdx.x * (blockDim.x >> 5) + wid;
    if (cw >= count) return;
    const uint8_t* in = data + cw * c.k;
    uint8_t* out = wire + cw * c.n;
    for (int w = 0; w < c.kwords; w++) {
        const int j = w * 32 + lane;
        const unsigned bit = j < c.k ? (in[j] & 1u) : 0u;
        const uint32_t word = __ballot_sync(0xffffffffu, bit);
        if (lane == 0) dw[w] = word;
        if (j < c.k) out[c.rank + j] = (uint8_t)bit;
    }
    __syncwarp();
    for (int i = lane; i < c.rank; i += 32) {
        const uint32_t* row = c.pmat + (size_t)i * c.kwords;
        uint32_t acc = 0;
        for (int w = 0; w < c.kwords; w++) acc ^= __ldg(&row[w]) & dw[w];
        out[i] = (uint8_t)(__popc(acc) & 1);
    }
}

cudaError_t launch_ldpc_encode(cudaStream_t st, const CodeDev& code, const uint8_t* data, long long count, uint8_t* wire)
{
    if (count <= 0) return cudaSuccess;
    const int warps = 8;
    const long long blocks = (count + warps - 1) / warps;
    if (blocks > 0x7fffffffLL) return cudaErrorInvalidValue;
    ldpc_encode_kernel<<<(unsigned)blocks, warps * 32, (size_t)warps * code.kwords * 4, st>>>(code, data, count, wire);
    return cudaGetLastError();
}

} // namespace dtl
