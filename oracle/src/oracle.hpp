// oracle/src/oracle.hpp -- TEST INFRASTRUCTURE ONLY.
//
// CPU restatement ("oracle") of gr-dtl's adaptive-OFDM receive hot path and of the
// GNU Radio 3.10 pieces it calls.  Only tests/, __graft_entry__.smoke() and
// bench.py's cpu_baseline / --impl reference legs may load this library; the product
// (gr-dtl_b200/) never includes, links or calls anything in oracle/.
//
// PARITY STATUS: the gr-dtl code under /root/reference cannot be compiled here
// (every translation unit needs GNU Radio 3.10 headers/libs, which are absent), so
// this is a restatement, pinned against the reference's own golden vectors where they
// exist (tests/test_oracle_golden.py): CRC-32 + LSB-first repack bytes
// (python/dtl/qa_ofdm_adaptive_frame_pack_bb.py:39-42), CRC-16 + header wire format
// (python/dtl/qa_ofdm_adaptive_packet_header.py:127-135,233-241), the 8PSK/BPSK metric
// constant 6.568538665771484 (python/dtl/qa_ofdm_adaptive_constellation_metric_vcvf.py:86),
// the feedback-decision sequence (python/dtl/qa_ofdm_adaptive_feedback_decision.py:50-59),
// alist edge counts/ranks, and the equaliser + FEC round-trip scenarios.
// Floating-point values of third-party GNU Radio arithmetic (calc_soft_dec,
// mpsk_snr_est_simple, awgn_bp, GF2Mat::get_G pivot order, FFTW rounding) are RECALLED
// from GNU Radio 3.10 and are "parity unpinned" -- see DESIGN.md section 3.
//
// Arithmetic contract: every fp32 operation of the equaliser recurrence is written as
// an explicit IEEE-754 single operation (fmaf / * / + / 1.0f/x) and this library is
// compiled with -ffp-contract=off, so the CUDA kernels (which use __fmaf_rn/__fmul_rn/
// __frcp_rn) can match it bit for bit.  `eq_literal` keeps the reference's literal
// std::complex<float> expressions for cross-checking that contract.
#pragma once
#include <complex>
#include <cstdint>
#include <cstddef>
#include <string>
#include <vector>
#include <functional>
#include <memory>

namespace orc {

typedef std::complex<float> cf32;

// ---- include/gnuradio/dtl/ofdm_adaptive_utils.h:22-28 (+ extension values 5,6) ----
enum cnst_t { UNKNOWN = 0, BPSK = 1, QPSK = 2, PSK8 = 3, QAM16 = 4, QAM64 = 5, QAM256 = 6 };
static const int MAX_CNST = 6;

// lib/dtl/ofdm_adaptive_utils.cc:64-70
int bits_per_symbol(int cnst);

struct constellation {
    int type = 0;
    int bps = 0;
    std::vector<cf32> points;     // index -> point (map_to_points)
    std::vector<cf32> inv_points; // 1/point, fp32 (used by the equaliser contract)
    float level = 0, inv2l = 0;   // square-QAM level and 1/(2*level) (QAM16/64/256)
    unsigned decide(cf32 s) const;               // decision_maker
    // gr-digital constellation::calc_soft_dec (RECALLED): returns MSB-first LLRs
    void calc_soft_dec(cf32 sample, float npwr, float* out_msb_first) const;
    double min_dist2() const;
};
const constellation& get_constellation(int cnst);

// ---- gr::digital::crc (RECALLED, Rocksoft model) ----
struct crc_engine {
    unsigned nbits; uint64_t poly, init, xorout; bool in_refl, out_refl;
    uint64_t table[256];
    crc_engine(unsigned nbits, uint64_t poly, uint64_t init, uint64_t xorout, bool in_refl, bool out_refl);
    uint64_t compute(const uint8_t* d, size_t n) const;
};
// lib/dtl/crc_util.cc
struct crc_util {
    crc_engine eng; size_t crc_len; size_t count = 0, failed = 0;
    crc_util(size_t len, uint64_t poly, uint64_t init, uint64_t xorout);
    uint64_t append_crc(uint8_t* buf, size_t len);
    bool verify_crc(const uint8_t* buf, size_t len);
};

// lib/testbed/repack.cc
struct repack {
    uint8_t out_index = 0, in_index = 0, bits_in, bits_out;
    repack(uint8_t bi, uint8_t bo) : bits_in(bi), bits_out(bo) {}
    int repack_lsb_first(const uint8_t* in, size_t n_in, uint8_t* out, bool full_output_symbols = false);
    int repack_msb_first(const uint8_t* in, size_t n_in, uint8_t* out, bool full_output_symbols = false);
};

// lib/dtl/ofdm_adaptive_feedback_decision.cc
struct feedback_decision {
    std::vector<double> snr_th; std::vector<int> cnst; std::vector<int> fec;
    double hysteresis; int decision_th; int counter = 0; int last, next;
    feedback_decision(double hyst, int th, const std::vector<double>& snr,
                      const std::vector<int>& c, const std::vector<int>& f, int initial);
    void get_feedback(double snr, int* c, int* f);
  private:
    void update_decision(int id);
};

// gr-digital mpsk_snr_est_simple behind include/gnuradio/dtl/ofdm_adaptive_frame_snr.h (RECALLED)
struct snr_simple {
    double y1 = 0, y2 = 0, counter = 1, signal = 0, noise = 0;
    void reset() { *this = snr_simple(); }
    void update(cf32 x);
    double snr();       // dB; also refreshes signal/noise
    double noise_db() const;
};

// gr-fft fft_vcc(N, forward, (), shift=True) (RECALLED)
void fft_shifted(const cf32* in, cf32* out, int n, bool forward);

// ---- equaliser: lib/dtl/ofdm_adaptive_equalizer.cc ----
struct ofdm_cfg {
    int fft_len = 64, cp_len = 16;
    std::vector<std::vector<int>> occupied_carriers;
    std::vector<std::vector<int>> pilot_carriers;
    std::vector<std::vector<cf32>> pilot_symbols;
    int symbols_skipped = 0;
    float alpha = 0.1f;
    bool input_is_shifted = true;
};

struct equalizer {
    ofdm_cfg cfg;
    std::vector<char> occupied;                    // [N]
    std::vector<std::vector<char>> pilot_mask;     // [sets][N]
    std::vector<std::vector<cf32>> pilot_sym;      // [symsets][N]
    std::vector<std::vector<cf32>> pilot_inv;      // [symsets][N] 1/pilot (contract)
    std::vector<cf32> channel_state;
    int pilot_carr_set = 0;
    snr_simple snr_est;
    bool literal = false;                          // true: reference's std::complex expressions
    explicit equalizer(const ofdm_cfg& c);
    void reset();
    // ofdm_adaptive_equalizer_base::frame_equalize (equalizer.cc:217-268)
    void frame_equalize(cf32* frame, cf32* frame_soft, int n_sym, const std::vector<cf32>& taps, int cnst);
    double get_snr() { return snr_est.snr(); }
    double get_noise();
};

struct frame_meta {   // POD mirror of the stream tags (see include/dtl_b200.h)
    int32_t constellation, carrier_offset, frame_no, payload_len, packet_len;
    int32_t fec_id, tb_no, tb_offset, tb_payload, reserved;
};
struct frame_result {
    double snr_db, noise; int32_t fb_cnst, fb_fec; int32_t emitted; int32_t lost_frames;
};

// lib/dtl/ofdm_adaptive_frame_equalizer_vcvc_impl.cc:99-255
struct frame_equalizer {
    equalizer eq; feedback_decision fb; int cp_len;
    std::vector<cf32> channel_state;
    int expected_frame_no = 0; long lost_frames = 0, frames_count = 0;
    frame_equalizer(const ofdm_cfg& c, const feedback_decision& f);
    // returns n_sym emitted (0 when payload == 0). out_hard/out_soft are [n_sym][N].
    int work(const cf32* in, int n_sym, const cf32* taps /*nullable*/, const frame_meta& m,
             cf32* out_hard, cf32* out_soft, frame_result* res);
};

// gr-digital ofdm_serializer_vcc (RECALLED)
int serialize(const ofdm_cfg& c, const cf32* frame, int n_sym, int carr_offset, int packet_len,
              int symbols_skipped, cf32* out);

// ---- gr-fec alist / cldpc / awgn_bp (RECALLED) + lib/dtl/ldpc_{enc,dec}.cc ----
struct ldpc_code {
    int N = 0, M = 0, K = 0, rank = 0;
    std::vector<std::vector<int>> nlist, mlist;   // 0-based
    std::vector<std::vector<std::pair<int, int>>> vedge;   // per var: (check, slot in mlist[check])
    std::vector<int> permute;                     // GF2Mat::get_G column permutation
    std::vector<std::vector<uint8_t>> P;          // rank x K parity part of G
    double sigma2 = 2.0;                          // awgn_bp sigma^2; gr-dtl never sets it
    int max_it = 15;
    bool load(const std::string& path);
    // ldpc_enc::encode (ldpc_enc.cc:58-68): out in wire order [parity(rank) | data(K)]
    void encode(const uint8_t* data_k, uint8_t* out_n) const;
    // ldpc_dec::decode (ldpc_dec.cc:62-71): in = n LLRs in wire order; out = K bits
    int decode(const float* in, int* nit, uint8_t* out) const;
    // awgn_bp::decode on a codeword in H column order (rx<0 => bit 1)
    void bp_decode(const std::vector<float>& rx, int* nit, std::vector<uint8_t>& est) const;
};

// lib/dtl/fec_utils.cc
int compute_tb_len(int cw_len, int frame_len);
int align_bits_to_bytes(int nbits);

struct fec_info {
    const ldpc_code* code = nullptr; int frame_payload = 0, tb_offset = 0, tb_number = 0, tb_payload_len = 0, ncheck = 0;
    int n() const { return code ? code->N : 0; } int k() const { return code ? code->K : 0; }
};

// lib/dtl/tb_encoder.cc
struct tb_encoder {
    std::vector<uint8_t> tb; int payload = 0; size_t buf_idx = 0;
    int encode(const uint8_t* in_bits, int len, const ldpc_code& c, int ncws);
    int buf_out(uint8_t* out, int len, int bps);
    int remaining() const { return (int)(tb.size() - buf_idx); }
    bool ready() const { return tb.empty() || tb.size() == buf_idx; }
};

// lib/dtl/tb_decoder.cc
struct tb_decoder {
    std::vector<float> rcv, full; std::vector<uint8_t> data; size_t max_tb;
    int tb_number = -1, tb_payload_len = 0; size_t buf_idx = 0; bool have_info = false; fec_info info;
    explicit tb_decoder(int max_tb_len);
    typedef std::function<void(const std::vector<uint8_t>&, const fec_info&, int)> ready_fn;
    bool process_frame(const float* in, int frame_len, int bps, const fec_info& fi, const ready_fn& on_ready);
  private:
    int decode(int tb_len, int& avg_it);
};

// lib/dtl/ofdm_adaptive_fec_decoder_impl.cc + ofdm_adaptive_fec_pack_bb_impl.cc
struct tb_result { int tb_no, crc_ok, nbytes, avg_it, frame_index; };
struct fec_decoder {
    std::vector<const ldpc_code*> decoders;   // [0] = nullptr
    int frame_capacity; crc_util crc; std::unique_ptr<tb_decoder> tbd;
    fec_decoder(const std::vector<const ldpc_code*>& d, int frame_capacity, int max_bps);
    // one frame of LLRs; appends packed payload bytes of CRC-passing TBs to out_bytes.
    void frame(const float* llr, const frame_meta& m, int frame_index, std::vector<uint8_t>& out_bytes, std::vector<tb_result>& tbs);
};

} // namespace orc
