// ctx.hpp -- internal definitions shared by the host side and the kernels of libdtl_b200.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <vector>
#include "../../include/dtl_b200.h"

namespace dtl {

// ---- device-side tables -----------------------------------------------------------------------------
// Constellation point tables live in global memory (divergent index -> __ldg, not __constant__).
struct CnstDev {
    const float2* points;       // concatenated, see off[]
    const float2* inv_points;   // 1/point (fp32 of the double quotient)
    int off[DTL_CNST_MAX + 2];  // start of constellation c in points[]
    int bps[DTL_CNST_MAX + 1];
    float level[DTL_CNST_MAX + 1], inv2l[DTL_CNST_MAX + 1];   // square-QAM level, 1/(2*level)
};

// Carrier plan of the payload equaliser + serialiser, in shifted-bin coordinates.
struct PlanDev {
    int N, L, cap;               // fft_len, payload symbols per frame, frame capacity (data symbols)
    int npc, nps, nocc;          // pilot-carrier sets, pilot-symbol sets, occupied sets
    int pcs0;                    // symbols_skipped % npc
    int symbols_skipped;
    int npil_frame;              // pilots per frame (pilot_prefix[L])
    int max_off;                 // |carrier offset| bound of phase_tab
    float alpha, oma;            // alpha, 1 - alpha (fp32)
    const uint8_t* occupied;     // [N]
    const uint8_t* pilot_mask;   // [npc][N]
    const uint16_t* pilot_rank;  // [npc][N] rank of bin among the pilots of the set
    const int* pilot_prefix;     // [L+1] pilots before symbol i
    const float2* pilot_sym;     // [nps][N]
    const float2* pilot_inv;     // [nps][N]
    const int16_t* ser_pos;      // [nocc][N] serialiser position of bin in set, -1 = not in set
    const int* ser_base;         // [L+1] serialised symbols before OFDM symbol i
    const float2* phase_tab;     // [2*max_off+1][L+1]: per-symbol correction i<L, taps rotation at L
    const uint32_t* sym_kind;    // [L][N] bits 0-1: 0 unused / 1 data / 2 pilot; bits 16-31: pilot slot in the frame
    const uint16_t* ser_src;     // [cap] serialised symbol idx -> i*N + bin
    const double* rcp_tab;       // [npil_frame] RN(1/(i+1)) for the SNR recurrence
    // the tables eq_demap_kernel touches per symbol, as one 16-byte aligned blob (copied to shared memory by TMA):
    const unsigned char* blob; int blob_bytes;
    int blob_inv;                // float2[n_points] 1/point of every constellation (= CnstDev::inv_points)
    int blob_ser;                // u16[cap]         = ser_src
    int blob_kind;               // u16[L*N]         kind | pilot slot << 2
    int blob_pinv, blob_psym;    // float2[npil_frame] 1/pilot and pilot of pilot slot q
    // eq_fast_kernel (one pilot-carrier set, one occupied set: a carrier's role does not change from symbol to symbol)
    int fast_ok;                 // the plan qualifies
    int pil_per_sym;             // pilots per OFDM symbol
    const uint8_t* carrier_kind; // [N] 0 unused / 1 data / 2 pilot
    const unsigned char* fblob; int fblob_bytes;   // small table blob, one TMA copy per CTA:
    int fb_inv;                  // float2[30]        1/point of BPSK..16QAM (QAM64/256 come from CnstDev::inv_points)
    int fb_ser;                  // u16[cap]          = ser_src
    int fb_pinv;                 // float2[npil_frame] 1/pilot of pilot slot q
    int fb_ppos;                 // u16[npil_frame]   i*N + bin of pilot slot q
    int fb_rcp;                  // end of the blob (the SNR recurrence's reciprocals are read by snr_tags_kernel from rcp_tab)
};

// Header branch (python/dtl/ofdm_receiver.py:100-150): channel estimator, header equaliser, header parser.
struct HdrDev {
    int N, H, nbits, has_fec;    // fft_len, header symbols, header bits (H * |occupied[0]|), FEC header fields present
    int max_neg, max_pos;        // range of the even integer carrier offsets ofdm_chanest_vcvc searches
    const float2* ref;           // [N] sync word 2, shifted layout
    const float2* corr_v;        // [N] sync2 / sync1 on the carriers sync word 1 uses, else 0
    const uint8_t* kind;         // [H][N] 0 unused / 1 data / 2 pilot (header symbol i: pilot-carrier set i % npc)
    const float2* pinv;          // [H][N] 1 / pilot symbol (pilot-symbol set i % nps)
    const uint16_t* ser;         // [nbits] i*N + bin of header bit j (ofdm_serializer_vcc order, symbols_skipped 0)
    const float2* rot;           // [2*max_off+1] channel-state rotation exp(j 2 pi off cp/N H) after the header symbols
    const float2* hphase;        // [2*max_off+1][H] per-symbol phase of the header symbols exp(-j 2 pi off cp/N (i+1))
    const uint16_t* crc_tab;     // [crc_bits] contribution of header bit j to the CRC-16 the parser computes (linear over GF(2))
    int crc_bits;                // bits pack_crc reads: 8 * ceil((nbits - 16) / 8)
    unsigned crc0;               // that CRC of an all-zero header
};

// One LDPC code (gr-fec alist + cldpc permutation) in device memory.
struct CodeDev {
    int n, m, k, rank, edges, max_row, max_col;
    const uint16_t* row_len;     // [m] row weight; edge j of row r (mlist order) lives in slot j*m + r
    const uint16_t* row_var;     // [max_row*m] variable of slot
    const uint16_t* col_ptr;     // [n+1]
    const uint16_t* col_edge;    // [edges] slots of variable v's edges in nlist order
    const uint16_t* permute;     // [n]
    const uint16_t* row_wire;    // [max_row*m] WIRE position (inverse permutation) of slot's variable: screening pass
    const uint32_t* pmat;        // [rank][kwords] parity part of the systematic generator, bit j of row i = P[i][j]
    int kwords;                  // ceil(k / 32)
    // reference-mode decoder, row-major layout (bp_ref.cuh): message of edge j of row r lives at r*rs + j
    int fast;                    // 1: max_row <= 8 and max_col <= 4 -- the tables below exist
    int rs;                      // row stride of the message array in doubles (max_row rounded up to even)
    int uniform_rows;            // every row has exactly max_row edges
    int nvr;                     // variable rounds: ceil(n / 32)
    const uint4* rv2;            // [m] variables of the row's edges as 8 x u16 in mlist order (pad: n)
    int lay_steps;               // layered decoder (bp_layered.cuh): groups of variable-disjoint rows, one step each
    const uint16_t* lay_rows;    // [lay_steps][32][8] per (step, lane): the row's variable SLOTS (pad: lay_ns), [7] = row id (0xffff = idle)
    int lay_ns;                  // posterior slots (>= n; bank-conflict-aware placement of the variables, api.cc: lay_assign_slots)
    const uint16_t* lay_perm;    // [n] wire position -> slot
    const uint4* vdesc;          // [nvr*32] one entry per (round, lane): x = slot0 | slot1 << 16, y = slot2 | slot3 << 16 in nlist
                                 //   order (pad: the dummy slot m*rs, which holds 1.0), z = variable (idle lane: n + 1)
};

// Per-stream receiver state that survives between calls.
struct StreamState {
    // ofdm_adaptive_feedback_decision
    int fb_last, fb_next, fb_counter;
    // frame equaliser block counters
    int expected_frame_no; int lost_frames; int frames_count;
    // tb_decoder
    int tb_number, tb_payload_len, have_info;
    int info_fec, info_tb_payload, info_tb_number;
    int carry_len;               // LLRs of the TB in flight held in the carry buffer
    int buf_idx;                 // tb_decoder's d_buf_idx (capacity check only)
    int frame_index;             // frames seen by the FEC decoder
    int hdr_cnst;                // header parser's fallback constellation (ofdm_adaptive_packet_header.cc:70,287-289)
};

// Synchronisation front end (kernels_sync.cu): per-stream state that survives between calls
#define DTL_SYNC_PEND 4
#define DTL_SYNC_SEGC 6
struct SyncSeg { long long t; double ph; float f; int pad; };   // oscillator: phase[t] = ph, slope -2/N * f per sample from t on
struct SyncState {
    long long total, scan_pos, scan_lo, chunk_pos, busy_until, base_next, base_cur;
    int keep_off, keep_len, len_cur, npend, nsegc;
    long long pend[DTL_SYNC_PEND];
    SyncSeg segc[DTL_SYNC_SEGC];
    // ofdm_adaptive_frame_detect_bb_impl members (lib/dtl/ofdm_adaptive_frame_detect_bb_impl.h)
    int d_remainder, d_missing_count, d_correction_count, d_acc_error, d_trigger_counter, d_synced, d_in_sync, d_error_count;
};

// TX FEC framer (kernels_txf.cu): per-stream state that survives between calls, and the per-call plan
struct TxStream { int tb_count, frame_count; int rem, total; int cur_fec, cur_payload; };   // rem/total: bits of the TB in flight
struct TxTb { int fec_id, ncws, payload_bytes, in_off, bits; };
struct TxSeg { int tb, bit0, nbits, chunk0; };        // tb: index among this call's TBs of the stream, -1 = carry buffer

struct TbSeg { int src; int start; int len; };   // src: -1 = carry buffer, else frame index in this call
#define DTL_TB_MAXSEG 6
struct TbDesc {
    int stream, fec_id, ncws, payload_len, tb_no, frame_index;
    int nseg; TbSeg seg[DTL_TB_MAXSEG];
    int cw_first;                // index of the first codeword work item
    long long bit_base;          // offset of the TB's bits in the bit scratch
};

struct Code {                    // host copy
    int n = 0, m = 0, k = 0, rank = 0, edges = 0, max_row = 0, max_col = 0;
    std::vector<std::vector<int>> nlist, mlist;
    std::vector<int> permute;
    std::vector<std::vector<uint8_t>> P;   // rank x k parity part (for the TX-side encoder)
    CodeDev dev{};
    int lay_shared = 0;          // variable occurrences beyond the first inside one step of the layered schedule (api.cc: schedule_rows)
    int bank_extra = 0;          // shared-memory wavefronts beyond one per access the variable order leaves (api.cc: order_variables)
    std::string path;
};

bool load_alist(const std::string& path, Code& c, std::string& err);
void exp_table(std::vector<double>& tab);   // [128][2]: tail, scale bits (bp_ref.cuh)

// Per-stage device timer (dtl_b200_profile / dtl_b200_stage_times): event pairs on the handle's stream.
struct Prof {
    bool on = false;
    cudaStream_t st = nullptr;
    struct Rec { int stage; int launches; cudaEvent_t a, b; };
    std::vector<Rec> pending;
    std::vector<cudaEvent_t> pool;
    double ms[DTL_STAGE_COUNT] = { 0 };
    long long n[DTL_STAGE_COUNT] = { 0 };
    cudaEvent_t get()
    {
        if (!pool.empty()) { cudaEvent_t e = pool.back(); pool.pop_back(); return e; }
        cudaEvent_t e = nullptr; cudaEventCreate(&e); return e;
    }
    void begin(int stage, int launches)
    {
        if (!on) return;
        Rec r; r.stage = stage; r.launches = launches; r.a = get(); r.b = get();
        cudaEventRecord(r.a, st);
        pending.push_back(r);
    }
    void end() { if (on && !pending.empty()) cudaEventRecord(pending.back().b, st); }
    void collect()                       // the stream must be idle
    {
        for (auto& r : pending) {
            float t = 0.f;
            if (cudaEventElapsedTime(&t, r.a, r.b) == cudaSuccess) { ms[r.stage] += t; n[r.stage] += r.launches; }
            pool.push_back(r.a); pool.push_back(r.b);
        }
        pending.clear();
    }
    void release()
    {
        for (auto& r : pending) { cudaEventDestroy(r.a); cudaEventDestroy(r.b); }
        for (auto e : pool) cudaEventDestroy(e);
        pending.clear(); pool.clear();
    }
};

// host constellation tables (lib/dtl/ofdm_adaptive_utils.cc:41-70, lib/dtl/constellation.cc:18-24)
int cnst_bps(int c);
void cnst_points(int c, std::vector<float2>& pts, std::vector<float2>& inv, float* level, float* inv2l);

} // namespace dtl
