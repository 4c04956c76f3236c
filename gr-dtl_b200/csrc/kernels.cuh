// kernels.cuh -- kernel parameter blocks and launcher prototypes (internal).
#pragma once
#include "ctx.hpp"

namespace dtl {

#define DTL_MAX_DYN_SMEM (227 * 1024)   // opt-in dynamic shared memory per CTA on sm_100
// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-device attribute: each init_kernels_* sets it for the kernels of its
// translation unit on the CURRENT device; dtl_b200_create calls them after cudaSetDevice, so handles on several GPUs of
// one process all find it set.
cudaError_t init_kernels_eq();
cudaError_t init_kernels_fec();
cudaError_t init_kernels_hdr();
cudaError_t init_kernels_fft();

#define DTL_MAX_SMEM_CODES 5     // codes 1..4 get their graph tables staged in shared memory

// reference-mode decoder, row-major layout (bp_ref.cuh): shared-memory bytes of one code's tables / of one warp's buffers, and
// whether bp_ref_run has an instance for a code with these maximum degrees
__host__ __device__ inline size_t fast_tab_bytes(int n, int m, int nvr)
{
    return 16 * (size_t)m + 16 * (size_t)nvr * 32 + ((2 * (size_t)n + 15) & ~(size_t)15);
}
__host__ __device__ inline size_t fast_warp_bytes(int n, int m, int rs)
{
    return ((size_t)m * rs + 2) * 8 + ((((size_t)n + 2) * 8 + 15) & ~(size_t)15) + (((size_t)n + 2 + 15) & ~(size_t)15);   // msg, lr0, est: 16-byte multiples (rs is even)
}
// layered decoder (bp_layered.cuh)
__host__ __device__ inline size_t lay_tab_bytes(int n, int m, int steps)
{
    (void)m;
    return 512 * (size_t)steps + ((2 * (size_t)n + 15) & ~(size_t)15);      // row descriptors + wire position -> slot
}
__host__ __device__ inline size_t lay_warp_bytes(int ns, int steps)
{
    return (((size_t)steps * 32 * 3 * 4 + 15) & ~(size_t)15) + ((((size_t)ns + 1) * 4 + 15) & ~(size_t)15);   // compressed messages + posteriors by slot
}
inline bool bp_ref_has_instance(int max_row, int max_col)
{
    return max_row >= 3 && max_row <= 8 && max_col >= 1 && max_col <= 4 && !(max_row == 3 && max_col > 3);
}

__host__ __device__ inline int log2n_of(int n) { int l = 0; while ((1 << l) < n) l++; return l; }

// ---- stage 5 ------------------------------------------------------------------------------------------
struct TxParams {
    int N, cp, H, L, cap, nocc, hdr_stride;
    long long n_frames;
    const uint8_t* hdr_bits; const uint8_t* chunks; const dtl_frame_meta* meta;
    const float2* sync1; const float2* sync2;      // [N] shifted layout
    const int16_t* ser_pos;                        // [nocc][N]
    const int* tx_base;                            // [H+L+1] data symbols before OFDM data symbol s
    const uint8_t* tx_pil_mask;                    // [H+L][N]
    const float2* tx_pil_val;                      // [H+L][N]
    const float2* tw; CnstDev cn; float2* iq;
};

// ---- stages 2+3 -----------------------------------------------------------------------------------------
struct EqParams {
    PlanDev pl; CnstDev cn;
    const float2* freq; const float2* taps; const dtl_frame_meta* meta;
    long long n_frames;
    float2* out_hard; float2* out_soft; float2* taps_out;
    dtl_frame_result* res;                         // snr_db, noise, emitted filled here
    float* llr; int llr_stride; int* n_llr;
    uint8_t* bytes; int bytes_stride; int* nbytes; int* crc_ok;
    const uint32_t* crc_table;                     // [256] reflected CRC-32
    int* err;                                      // device error word (bit 0: bad constellation)
    const float2* tw;
    int use_tma;                                   // set by the launcher: source 16-byte aligned
    float* mag;                                    // [n_frames][npil_frame] scratch: |pilot_eq| (eq_fast_kernel -> snr_tags_kernel)
    uint32_t* sgn; int sgn_stride;                 // [n_frames][sgn_stride] hard decisions of the LLR stream, bit p % 32 of word p / 32 =
                                                   //   (llr[p] > 0): what the decoder's screening pass reads instead of the LLRs (or null)
    // fused CP strip + FFT (eq_fast_kernel, fft_len 64): time-domain frames instead of `freq`
    const float2* iq; long long iq_frame_stride; int iq_sym0, iq_sym_stride, iq_skip;
};
// true when launch_eq would run the lean kernel with the FFT fused in for these parameters (p.iq set, p.freq unused)
bool eq_can_fuse_fft(const EqParams& p);
// true when launch_eq would fill p.sgn for these parameters (the lean kernel, fft_len >= 32); otherwise the caller runs launch_llr_sign
bool eq_emits_signs(const EqParams& p);

// ---- per-stream sequential bookkeeping ------------------------------------------------------------------
struct StreamParams {
    StreamState* st; int n_streams, F;
    const dtl_frame_meta* meta; dtl_frame_result* res;
    int n_mcs; const double* mcs_snr; const int* mcs_cnst; const int* mcs_fec; double hysteresis; int decision_th;
};

// One queued codeword (bp_gather_kernel -> ldpc_bp_ref_kernel / ldpc_bp_lay_kernel): where its bits go, and what the decoder's
// bulk copy fetches -- `nbytes` (a multiple of 16) from `src` (16-byte aligned: the LLR row itself, or the codeword's gathered
// block in cwbuf); wire position q of the codeword is float shift + q of the copy for q < len, a shortened position past it.
struct BpItem { long long out_off; const float* src; int it, fec_id, k_, shift, nbytes, len; };
#define DTL_STG_SLACK 32                             // bytes a staging block has beyond cw_stride floats: lead-in + round-up of an unaligned run

// ---- stage 4 ------------------------------------------------------------------------------------------------
struct FecParams {
    StreamState* st; int n_streams, F;
    const dtl_frame_meta* meta;
    const float* llr; int llr_stride;
    const uint32_t* sgn; int sgn_stride;           // sign words of the LLR stream (EqParams::sgn / llr_sign_kernel), [S*F][sgn_stride]
    uint32_t* csgn; int csgn_stride;               // the same for the carry buffer, [S][csgn_stride] (carry_save_kernel)
    float* carry; int carry_cap;                   // [S][carry_cap]
    int n_codes; const CodeDev* codes;             // [n_codes+1], entry 0 unused
    int cap; int bps_tab[DTL_CNST_MAX + 1];
    int max_tb;                                    // tb_decoder capacity (max_n * ncws(max_bps))
    TbDesc* tbs; int tb_slots;                     // [S][tb_slots] released TBs of this call
    TbDesc* pend;                                  // [S] segments of the TB still in flight
    dtl_tb_result* tres;                           // [S][tb_slots]
    int* tb_count;                                 // [S]
    int* tb_list;                                  // compact list of released TB indices
    int* cw_tb; int cw_cap;                        // codeword work list: TB index of work item
    int* cnt;                                      // [0] codewords, [1] work cursor, [2] released TBs, [3],[5],[6],[7] queued for BP, [10] of them to be gathered
    unsigned long long* bits_used;                 // bump allocator of the bit scratch
    uint8_t* bits; long long bits_cap;             // decoded TB bits, one byte per bit, then packed in place
    int* cw_nit;                                   // [cw_cap]
    int* cw_geo;                                   // [cw_cap][3]: first LLR of the codeword in the TB, first data bit, k_
    int* bp_list;                                  // [4][cw_cap] codewords that failed the initial syndrome check, hardest-looking list
                                                   // first (cnt[3], cnt[5], cnt[6], cnt[7] of them)
    int bp_th[3];                                  // list 0: >= bp_th[0] lanes saw a failing check, list 1: >= bp_th[1], ...
    int screen_words;                              // ceil(max n / 32)
    int tab_bytes, screen_bytes, screen_tab_bytes; // shared-memory graph tables; per-warp bytes and table bytes of the screening pass
    int max_it; double sigma2; int mode; int warp_bytes;
    // reference mode on codes bp_ref.cuh has an instance for (all codes of the handle): ldpc_bp_ref_kernel
    int fast; int fast_tab_bytes, fast_warp_bytes; const double2* exp_tab;
    int lay; int lay_tab_bytes, lay_warp_bytes;    // layered mode on codes with row-major tables: ldpc_bp_lay_kernel
    double sigma2_lr;                              // sigma2, or -(1/sigma2) when sigma2 is a power of two (lr_of_rx)
    float* cwbuf; int cw_stride;                   // [queued codewords][cw_stride] received values in H column order (bp_gather_kernel)
    BpItem* items;                                 // [4][cw_cap] records of the queued codewords, same positions as bp_list (ldpc_screen_kernel)
    int* glist;                                    // [cnt[10]] positions of the queued codewords whose runs bp_gather_kernel copies into cwbuf
    // outputs
    uint8_t* bytes; int bytes_cap; int* nbytes; dtl_tb_result* tb_out; int tb_cap; int* ntbs;
    const uint32_t* crc_table; int* err;
};

struct LdpcParams {                                // stand-alone fec_dec::decode batches
    CodeDev code; const float* llr; long long count; uint8_t* bits; int* nit; int max_it; double sigma2; int mode;
    int* cursor; int warp_bytes; int tab_bytes;
    int fast; int fast_tab_bytes, fast_warp_bytes; const double2* exp_tab; double sigma2_lr;
    int lay; int lay_tab_bytes, lay_warp_bytes;
};

// ---- TX FEC framer --------------------------------------------------------------------------------
struct TxfParams {
    TxStream* st; int n_streams, F, cap;
    const uint8_t* data; int in_cap; const int* in_len;      // [S][in_cap] user bytes, [S] valid bytes (or null: in_cap)
    const int* cnst; const int* fec;                          // [S*F] constellation / FEC id of the frame (the MCS feedback)
    int n_codes; const CodeDev* codes; int bps_tab[DTL_CNST_MAX + 1];
    TxTb* tbs; int* ntb; int* keep; TxSeg* segs;              // [S][F], [S], [S], [S*F][2]
    uint32_t* tbcrc;                                          // [S][F]
    uint8_t* tbbits; int maxbits;                             // [S][F][maxbits] encoded TB bit streams, one bit per byte
    uint8_t* carry;                                           // [S][maxbits] bits of the TB in flight across calls
    int max_ncws, max_kwords;
    uint8_t* chunks; dtl_frame_meta* meta; int* consumed; int* tbs_started;
    const uint32_t* crc_table; int* err;
};
cudaError_t launch_txf(cudaStream_t st, const TxfParams& p, int* launches);
cudaError_t launch_txf_reset(cudaStream_t st, TxStream* s, int n);

// ---- synchronisation front end ----------------------------------------------------------------------
struct SyncParams {
    SyncState* st; int n_streams; int T, N, cp, FS; float threshold;
    const float2* x;                               // [S][T] the call's new samples
    const float2* w_old; float2* w_new; int w_cap, keep_cap;     // per-stream work buffers [S][w_cap], double-buffered
    uint32_t* flags; int flag_words;               // [S][flag_words] metric >= threshold, one bit per position
    SyncSeg* segs; int seg_cap; int* n_seg;        // [S][seg_cap] phase segments of this call (carried ones first)
    long long* raw; long long* fixed; int trig_cap; int* n_raw; int* n_fixed;   // [S][trig_cap] raw / regularised triggers
    long long* starts; long long* starts_out; int frames_cap; int* n_frames;    // [S][frames_cap] frame starts (delayed coordinates)
    float2* frames; int* valid;                    // [S][frames_cap][FS], [S][frames_cap]
    int* err;
};
cudaError_t launch_sync(cudaStream_t st, const SyncParams& p, int* launches);
cudaError_t launch_sync_reset(cudaStream_t st, SyncState* state, int n);
cudaError_t init_kernels_sync();
size_t sync_metric_smem(int N);

// ---- header branch ----------------------------------------------------------------------------------------------
struct HdrParams {
    HdrDev hd; PlanDev pl; CnstDev cn;
    const float2* hfreq;                           // [n_frames][2+H][N] frequency-domain sync + header symbols
    long long n_frames; int n_streams, F;
    dtl_frame_meta* meta; float2* taps; dtl_header_result* hres;
    StreamState* st;
    const int* valid;                              // [n_frames] or null: 0 = an empty frame slot (dtl_b200_rx_stream), dropped like a CRC failure
};

struct DemapParams {
    CnstDev cn; const float2* syms; int stride_syms; const dtl_frame_meta* meta; const double* noise;
    long long n_frames; float* llr; int llr_stride; uint8_t* chunks; const uint8_t* in_chunks; float2* out_syms; int mode;
};

cudaError_t launch_fft_fwd(cudaStream_t st, const float2* in, float2* out, long long n_sym, int syms_per_frame,
                           long long in_frame_stride, int in_sym0, int in_sym_stride, int in_skip, int N, const float2* tw);
cudaError_t launch_fft_inv(cudaStream_t st, const float2* in, float2* out, long long n_sym, int N, const float2* tw);
cudaError_t launch_tx_mod(cudaStream_t st, const TxParams& p);
cudaError_t launch_eq(cudaStream_t st, const EqParams& p, int* launches = nullptr, bool* tags_deferred = nullptr);   // *launches += kernels launched
// tags_deferred != nullptr: the lean kernel's snr_tags_kernel is left to the caller (launch_snr_tags, any stream ordered
// after `st`), *tags_deferred says whether one is due
cudaError_t launch_snr_tags(cudaStream_t st, const EqParams& p);
// eq_fast_kernel + snr_tags_kernel with p.mag replaced by the library's static device scratch (no allocation; kernels_eq.cu)
cudaError_t launch_eq_noalloc(cudaStream_t st, const EqParams& p, int* why, long long* need, long long* have);
// sign words of n_frames rows of llr_stride floats: sgn[f][w] bit b = llr[f][32 w + b] > 0 (rows padded with zeros to sgn_stride words)
cudaError_t launch_llr_sign(cudaStream_t st, const float* llr, int llr_stride, long long n_frames, uint32_t* sgn, int sgn_stride);
cudaError_t launch_stream_reset(cudaStream_t st, StreamState* state, int n, int initial_mcs);
cudaError_t launch_stream_state(cudaStream_t st, const StreamParams& p);
// phase 0: everything; 1: TB planning only (needs the tags, not the LLRs); 2: what follows the planning
// side/ev: when given, carry_save_kernel (needs the decoder to be done with the carry buffer, nothing waits for it) runs on
// `side` behind an event recorded after the BP kernel, beside the TB epilogue; the caller joins `side` afterwards
cudaError_t launch_fec(cudaStream_t st, const FecParams& p, int sm_count, int* launches, Prof* prof, int phase = 0,
                       cudaStream_t side = nullptr, cudaEvent_t ev = nullptr);
size_t ldpc_warp_bytes(int n, int slots);
size_t ldpc_tab_bytes(int n, int m, int slots, int edges);
cudaError_t launch_ldpc(cudaStream_t st, const LdpcParams& p, int sm_count);
cudaError_t launch_ldpc_encode(cudaStream_t st, const CodeDev& code, const uint8_t* data, long long count, uint8_t* wire);
cudaError_t launch_header(cudaStream_t st, const HdrParams& p);      // header_kernel + header_fix_kernel
cudaError_t launch_demap(cudaStream_t st, const DemapParams& p);   // mode 0 soft, 1 hard, 2 map
cudaError_t launch_frame_gather(cudaStream_t st, const uint8_t* fbytes, int bytes_stride, const int* fn, int n_streams, int F,
                                uint8_t* out, int cap, int* out_n, int* err);

} // namespace dtl

#ifdef __CUDACC__
// ---- TMA bulk copy (cp.async.bulk, SASS UBLKCP) + mbarrier helpers ------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t phase)
{
    asm volatile("{\n .reg .pred p;\nW: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n @p bra D;\n bra W;\nD:\n}"
                 ::"r"(smem_u32(bar)), "r"(phase) : "memory");
}

// opt-in dynamic shared memory of kernel `k` on the current device: everything the SM allows beyond the kernel's static part
template <class K> static inline cudaError_t set_max_dyn_smem(K k)
{
    cudaFuncAttributes a;
    cudaError_t e = cudaFuncGetAttributes(&a, k);
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, DTL_MAX_DYN_SMEM - (int)a.sharedSizeBytes);
}
// 8-point forward DFT in registers (radix-2 then two radix-4), used by the 64-point FFT (8 x 8) in kernels_fft.cu and by
// the equaliser kernel when the FFT is fused into it.
__device__ __forceinline__ void dft8_fwd(float2 (&x)[8])
{
    const float r = 0.70710678118654752440f;
#define DTL_ADD(a, b) make_float2((a).x + (b).x, (a).y + (b).y)
#define DTL_SUB(a, b) make_float2((a).x - (b).x, (a).y - (b).y)
    const float2 a0 = DTL_ADD(x[0], x[4]), a1 = DTL_ADD(x[1], x[5]), a2 = DTL_ADD(x[2], x[6]), a3 = DTL_ADD(x[3], x[7]);
    const float2 b0 = DTL_SUB(x[0], x[4]), t1 = DTL_SUB(x[1], x[5]), t2 = DTL_SUB(x[2], x[6]), t3 = DTL_SUB(x[3], x[7]);
    const float2 b1 = make_float2((t1.x + t1.y) * r, (t1.y - t1.x) * r);        // * W8^1 = (1 - i)/sqrt 2
    const float2 b2 = make_float2(t2.y, -t2.x);                                 // * W8^2 = -i
    const float2 b3 = make_float2((t3.y - t3.x) * r, -(t3.x + t3.y) * r);       // * W8^3 = (-1 - i)/sqrt 2
    {   // 4-point DFT of a -> X0, X2, X4, X6
        const float2 s0 = DTL_ADD(a0, a2), d0 = DTL_SUB(a0, a2), s1 = DTL_ADD(a1, a3), u = DTL_SUB(a1, a3);
        const float2 d1 = make_float2(u.y, -u.x);
        x[0] = DTL_ADD(s0, s1); x[2] = DTL_ADD(d0, d1); x[4] = DTL_SUB(s0, s1); x[6] = DTL_SUB(d0, d1);
    }
    {   // 4-point DFT of b -> X1, X3, X5, X7
        const float2 s0 = DTL_ADD(b0, b2), d0 = DTL_SUB(b0, b2), s1 = DTL_ADD(b1, b3), u = DTL_SUB(b1, b3);
        const float2 d1 = make_float2(u.y, -u.x);
        x[1] = DTL_ADD(s0, s1); x[3] = DTL_ADD(d0, d1); x[5] = DTL_SUB(s0, s1); x[7] = DTL_SUB(d0, d1);
    }
#undef DTL_ADD
#undef DTL_SUB
}
#endif
