// oracle/src/tb.cc -- TEST INFRASTRUCTURE ONLY (see oracle.hpp header).
// Transport-block (TB) encoder / decoder and the FEC decoder block:
//   lib/dtl/tb_encoder.cc:32-99, lib/dtl/tb_decoder.cc:21-201,
//   lib/dtl/ofdm_adaptive_fec_decoder_impl.cc:80-217, lib/dtl/ofdm_adaptive_fec_pack_bb_impl.cc:48-69,
//   lib/dtl/fec_utils.cc:60-121.
#include "oracle.hpp"
#include <algorithm>
#include <cstring>
#include <stdexcept>

namespace orc {

// ---- tb_encoder.cc:32-74 -------------------------------------------------------------------
int tb_encoder::encode(const uint8_t* in, int len, const ldpc_code& c, int ncws)
{
    int read_index = 0;
    tb.clear(); payload = 0; buf_idx = 0;
    const int n = c.N, k = c.K, ncheck = n - k;
    std::vector<uint8_t> cw_in(n), cw_out(n);
    for (int i = 0; i < ncws; ++i) {
        int k_new = (len - payload) / (ncws - i);
        if ((len - payload) % (ncws - i)) ++k_new;
        payload += k_new;
        std::fill(cw_in.begin(), cw_in.end(), 0);
        memcpy(&cw_in[0], &in[read_index], k_new);
        c.encode(&cw_in[0], &cw_out[0]);
        read_index += k_new;
        tb.insert(tb.end(), cw_out.begin(), cw_out.begin() + ncheck);
        tb.insert(tb.end(), cw_out.begin() + ncheck, cw_out.begin() + ncheck + k_new);
    }
    return (int)tb.size();
}

// tb_encoder.cc:88-99
int tb_encoder::buf_out(uint8_t* out, int len, int bps)
{
    repack r(1, (uint8_t)bps);
    int syms = r.repack_lsb_first(&tb[buf_idx], len, out);
    buf_idx += len;
    return syms;
}

// ---- tb_decoder.cc ---------------------------------------------------------------------------
tb_decoder::tb_decoder(int max_tb_len) : max_tb((size_t)max_tb_len)
{
    rcv.reserve(2 * max_tb); full.reserve(2 * max_tb); data.reserve(2 * max_tb);
}

// tb_decoder.cc:32-141
bool tb_decoder::process_frame(const float* in, int frame_len, int bps, const fec_info& fi, const ready_fn& on_ready)
{
    if (!fi.code) return false;
    const int frame_payload_len = fi.frame_payload;
    int avg_it = 0;
    if (fi.tb_number == tb_number) {                                       // :57-70
        if (buf_idx + frame_payload_len > 2 * max_tb) throw std::runtime_error("rcv buffer does not have enough capacity!");
        rcv.insert(rcv.end(), in, in + frame_payload_len);
        buf_idx += frame_payload_len;
        int ncws = compute_tb_len(info.n(), frame_len);
        size_t expected = (size_t)((fi.n() - fi.k()) * ncws + fi.tb_payload_len);   // expected_tb_len(fec_info, ncws)
        if (rcv.size() >= expected) {
            decode(ncws, avg_it);
            buf_idx = 0;
            rcv.clear();
            on_ready(data, fi, avg_it);
        }
    } else {
        if (rcv.size() == 0 && fi.tb_offset == frame_payload_len) {        // :75-89
            info = fi; have_info = true;
            tb_number = fi.tb_number; tb_payload_len = fi.tb_payload_len;
            buf_idx = 0;
            rcv.clear();
            rcv.insert(rcv.end(), in, in + fi.tb_offset);
            int tb_len = compute_tb_len(info.n(), frame_len);
            decode(tb_len, avg_it);
            rcv.clear();
            on_ready(data, fi, avg_it);
        } else {                                                           // :92-133
            if (fi.tb_offset > 0) rcv.insert(rcv.end(), in, in + fi.tb_offset);
            if (rcv.size() > 0 && have_info) {
                int tb_len = compute_tb_len(info.n(), frame_len);
                decode(tb_len, avg_it);
                on_ready(data, info, avg_it);
            }
            info = fi; have_info = true;
            tb_number = fi.tb_number; tb_payload_len = fi.tb_payload_len;
            buf_idx = 0;
            rcv.clear();
            int extra_bits = 0;
            if (fi.tb_offset % bps) extra_bits = bps - fi.tb_offset % bps;
            int new_tb_offset = 0;
            if (fi.tb_offset) new_tb_offset = fi.tb_offset + extra_bits;
            rcv.insert(rcv.end(), in + new_tb_offset, in + frame_payload_len + extra_bits);
            buf_idx += frame_payload_len - fi.tb_offset;
            int tb_len = compute_tb_len(info.n(), frame_len);
            int tb_size = 8 * align_bits_to_bytes(tb_len * fi.ncheck + fi.tb_payload_len);
            if (frame_payload_len - fi.tb_offset == tb_size) {
                decode(tb_len, avg_it);
                on_ready(data, info, avg_it);
                buf_idx = 0;
                rcv.clear();
            }
        }
    }
    return true;
}

// tb_decoder.cc:143-195.  The reference reads RCV_BUF past size() when a TB is incomplete
// (undefined behaviour); here missing LLRs read as 0.0 (erasure).  Output: the first k_ bits
// of every codeword, concatenated (the k-byte overrun of ldpc_dec::decode is overwritten by
// the next codeword in the reference, :185-187).
int tb_decoder::decode(int tb_len, int& avg_it)
{
    static const float SHORTENED_VALUE = -15;
    int payload_len = info.tb_payload_len;
    const int n = info.n(), k = info.k(), ncheck = n - k;
    int n_iterations = 0;
    avg_it = 0;
    data.assign((size_t)info.tb_payload_len, 0);
    full.assign((size_t)tb_len * n, SHORTENED_VALUE);
    std::vector<uint8_t> cw_bits(k);
    auto rcv_at = [&](int idx) { return (size_t)idx < rcv.size() ? rcv[idx] : 0.0f; };
    int rcv_idx = 0;
    for (int i = 0, d_idx = 0; i < tb_len; ++i) {
        int k_ = payload_len / (tb_len - i);
        if (payload_len % (tb_len - i)) ++k_;
        payload_len -= k_;
        for (int j = 0; j < ncheck; j++) full[(size_t)i * n + j] = rcv_at(rcv_idx + j);
        rcv_idx += ncheck;
        for (int j = 0; j < k_; j++) full[(size_t)i * n + ncheck + j] = rcv_at(rcv_idx + j);
        rcv_idx += k_;
        info.code->decode(&full[(size_t)i * n], &n_iterations, &cw_bits[0]);
        for (int j = 0; j < k_; j++) data[d_idx + j] = cw_bits[j];
        d_idx += k_;
        avg_it += n_iterations;
    }
    avg_it /= tb_len;
    return tb_len * ncheck + info.tb_payload_len;
}

// ---- ofdm_adaptive_fec_decoder_impl.cc ---------------------------------------------------------
fec_decoder::fec_decoder(const std::vector<const ldpc_code*>& d, int cap, int max_bps)
    : decoders(d), frame_capacity(cap), crc(4, 0x04C11DB7, 0xFFFFFFFF, 0xFFFFFFFF)
{
    int max_n = 0;
    for (size_t i = 1; i < decoders.size(); i++) if (decoders[i]) max_n = std::max(max_n, decoders[i]->N);
    if (!max_n) throw std::runtime_error("Misconfiguration: decoder not found!");
    int ncws = compute_tb_len(max_n, frame_capacity * max_bps);
    tbd.reset(new tb_decoder(max_n * ncws));
}

// general_work :80-217 for ONE frame, followed by fec_pack_bb (bits -> bytes LSB-first)
void fec_decoder::frame(const float* llr, const frame_meta& m, int frame_index,
                        std::vector<uint8_t>& out_bytes, std::vector<tb_result>& tbs)
{
    const int bps = bits_per_symbol(m.constellation);
    if (!bps) throw std::runtime_error("Tags missing");
    if (m.fec_id <= 0 || (size_t)m.fec_id >= decoders.size() || !decoders[m.fec_id]) throw std::runtime_error("FEC tags missing");
    fec_info fi;                                                          // make_fec_info, fec_utils.cc:60-102
    fi.code = decoders[m.fec_id];
    fi.ncheck = fi.code->N - fi.code->K;
    fi.tb_number = m.tb_no; fi.tb_offset = m.tb_offset; fi.tb_payload_len = m.tb_payload;
    fi.frame_payload = 8 * m.payload_len;
    if (fi.frame_payload == 0) return;                                    // :141-143 empty frame
    fi.tb_offset *= 8;                                                    // :147
    auto on_ready = [&](const std::vector<uint8_t>& bits, const fec_info& tbi, int avg_it) {   // :161-204
        int user_data_len = (int)bits.size() - 32;
        tb_result r{ tbi.tb_number, 0, 0, avg_it, frame_index };
        if (user_data_len >= 0) {
            std::vector<uint8_t> buf(bits.size() / 8 + 2, 0);
            repack to_bytes(1, 8);
            int n0 = to_bytes.repack_lsb_first(bits.data(), user_data_len, buf.data());
            to_bytes.repack_lsb_first(bits.data() + user_data_len, 32, buf.data() + n0);
            bool ok = crc.verify_crc(buf.data(), n0 + 4);
            r.crc_ok = ok;
            if (ok) {
                // fec_decoder emits user_data_len unpacked bits; fec_pack_bb packs 8 -> 1 LSB-first
                r.nbytes = user_data_len / 8;
                out_bytes.insert(out_bytes.end(), buf.begin(), buf.begin() + r.nbytes);
            }
        }
        tbs.push_back(r);
    };
    tbd->process_frame(llr, 8 * align_bits_to_bytes(frame_capacity * bps), bps, fi, on_ready);
}

} // namespace orc
