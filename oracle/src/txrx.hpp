// oracle/src/txrx.hpp -- TEST INFRASTRUCTURE ONLY (see oracle.hpp header).
#pragma once
#include "oracle.hpp"

namespace orc {

struct rng { uint64_t s; uint64_t next(); double uniform(); double normal(); };

// lib/dtl/ofdm_adaptive_packet_header.cc (bits_per_header_sym = 1, no scrambling)
struct packet_header {
    int header_len, payload_syms; bool has_fec; crc_engine crc;
    int header_number = 0; int constellation = QAM16 /* :70 */; int max_cnst = QAM16;
    packet_header(int header_len_bits, int payload_syms, bool has_fec);
    void pack_crc(const uint8_t* bits, std::vector<uint8_t>& out) const;
    void format(const frame_meta& m, int feedback_cnst, int feedback_fec, uint8_t* out_bits);
    bool parse(const uint8_t* in_bits, frame_meta* m, int* feedback_cnst, int* feedback_fec);
};

// python/dtl/ofdm_adaptive_config.py:10-42 + ofdm_receiver.py:29-57
struct stream_cfg {
    ofdm_cfg ofdm;
    int frame_len = 20;
    bool fec = false;
    bool eq_literal = false;                       // payload equaliser runs the reference's literal std::complex expressions
    std::vector<const ldpc_code*> codes;           // [0] = nullptr (ldpc_dec.cc:25)
    std::vector<int> constellations;               // allowed set
    std::vector<double> mcs_snr; std::vector<int> mcs_cnst, mcs_fec; int initial_mcs = 0;
    std::vector<cf32> sync_word1, sync_word2;      // [N], shifted layout
    int capacity() const;                          // python/dtl/ofdm_adaptive.py:2-6
    int header_syms() const;                       // ofdm_receiver.py:66-68
    int frame_samples() const;
};

struct tx_framer {
    stream_cfg cfg; rng r; crc_util crc; tb_encoder enc;
    int tb_count = 0, frame_count = 0, cur_ncws = 0, cur_tb_fec = 0;
    std::vector<uint8_t> cur_payload;
    // user bytes: from `src` when set (a TB is only started while a whole TB payload is left), else pseudo-random
    const uint8_t* src = nullptr; int src_len = 0, src_pos = 0;
    tx_framer(const stream_cfg& c, uint64_t seed);
    void next_frame(int cnst, int fec_id, uint8_t* chunks, frame_meta* m, std::vector<std::vector<uint8_t>>& done);
  private:
    bool new_tb(int cnst, int fec_id);
};

void map_chunks(int cnst, const uint8_t* chunks, int n, cf32* out);
void tx_modulate(const stream_cfg& cfg, const cf32* hdr_syms, const cf32* payload_syms, cf32* iq_out);

// gr-digital ofdm_chanest_vcvc (RECALLED; header.cc)
struct chanest {
    int N, n_data_symbols, first_active, last_active, max_neg, max_pos;
    std::vector<cf32> ref, corr_v;                 // sync word 2; sync2/sync1 on the carriers sync word 1 uses
    chanest(const std::vector<cf32>& sync1, const std::vector<cf32>& sync2, int n_data);
    int carr_offset(const cf32* y1, const cf32* y2) const;
    void chan_taps(const cf32* y2, int off, cf32* taps) const;
};

// header branch of the receiver (python/dtl/ofdm_receiver.py:100-150; header.cc): one instance per stream
struct rx_header {
    stream_cfg cfg; chanest est; frame_equalizer feq; packet_header parser;
    explicit rx_header(const stream_cfg& c);
    // iq: one frame; fills the tags the payload branch receives.  false: header CRC failed (frame dropped, m->reserved = 1)
    bool frame(const cf32* iq, frame_meta* m, cf32* taps_out, int* fb_cnst, int* fb_fec);
};

// synchronisation front end (python/dtl/ofdm_receiver.py:70-97; sync.cc): one instance per continuous IQ stream
struct stream_sync {
    struct seg { long long t; double ph; float f; };       // oscillator: phase[t] = ph, slope -2/N * f from t on
    int N, cp, FS; float threshold;
    std::vector<cf32> buf; long long base = 0, total = 0, scan_pos = 0, chunk_pos = 0, busy_until = 0;
    std::vector<seg> segs; std::vector<long long> pending;
    // ofdm_adaptive_frame_detect_bb_impl members
    int d_remainder = 0, d_missing_count = 0, d_correction_count = 0, d_acc_error = 0, d_trigger_counter = 0, d_synced = 0,
        d_error_count = 0;
    bool d_in_sync = false;
    stream_sync(int fft_len, int cp_len, int frame_samples, float thr);
    cf32 x_at(long long n) const;
    void metric_at(long long n, float* m, cf32* p) const;
    double phase_at(long long n) const;
    std::vector<long long> fix_sync(long long c0, long long c1, const std::vector<long long>& trig);
    void push(const cf32* x, int T, std::vector<cf32>& frames, std::vector<long long>& starts,
              std::vector<long long>* raw_out = nullptr, std::vector<long long>* fixed_out = nullptr);
};

struct rx_frame_out {
    frame_result res; const cf32* hard = nullptr; const cf32* soft = nullptr;
    const float* llr = nullptr; int n_llr = 0; int crc_ok = -1;
    std::vector<uint8_t> bytes; std::vector<tb_result> tbs;
};

struct rx_stream {
    stream_cfg cfg; frame_equalizer feq; std::unique_ptr<fec_decoder> fdec; crc_util crc;
    std::vector<cf32> freq, hard, soft, ser; std::vector<float> llr; int frame_index = 0;
    explicit rx_stream(const stream_cfg& c);
    static ofdm_cfg make_payload_cfg(const stream_cfg& c);
    void demod(const cf32* iq, cf32* out) const;
    void frame(const cf32* iq, const cf32* taps, const frame_meta& m, rx_frame_out* o);
    void frame_freq(const cf32* freq, const cf32* taps, const frame_meta& m, rx_frame_out* o);
};

} // namespace orc
