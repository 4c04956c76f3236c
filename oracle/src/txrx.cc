// oracle/src/txrx.cc -- TEST INFRASTRUCTURE ONLY (see oracle.hpp header).
// Per-stream RX chain (stages 1-4) and the synthetic loopback TX (stage 5 + framing):
//   RX wiring: python/dtl/ofdm_receiver.py:152-239
//   soft demapper block: lib/dtl/ofdm_adaptive_constellation_soft_cf_impl.cc:68-156
//   hard demapper block: lib/dtl/ofdm_adaptive_constellation_decoder_cb_impl.cc:69-93
//   frame pack:          lib/dtl/ofdm_adaptive_frame_pack_bb_impl.cc:59-123
//   TX mapper:           lib/dtl/ofdm_adaptive_chunks_to_symbols_bc_impl.cc:59-81
//   header wire format:  lib/dtl/ofdm_adaptive_packet_header.cc:78-312
//   gr-digital ofdm_carrier_allocator_cvc / ofdm_cyclic_prefixer / header_payload_demux (RECALLED).
// The TX *framers* (ofdm_adaptive_fec_frame_bvb, ofdm_adaptive_frame_bb) are out of scope
// (wall-clock paced state machines, random padding); `tx_framer`
// below is this repo's deterministic generator of the frame layout tb_decoder consumes
// (full frames, back-to-back maximum-size TBs, seeded payload, no pacing).
#include "txrx.hpp"
#include <cmath>
#include <cstring>
#include <stdexcept>

namespace orc {

// ---- deterministic PRNG for payloads/noise (splitmix64 + Box-Muller) ------------------------
uint64_t rng::next()
{
    uint64_t z = (s += 0x9e3779b97f4a7c15ull);
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
    return z ^ (z >> 31);
}
double rng::uniform() { return (double)(next() >> 11) * (1.0 / 9007199254740992.0); }
double rng::normal()
{
    double u1 = uniform(), u2 = uniform();
    if (u1 < 1e-300) u1 = 1e-300;
    return std::sqrt(-2.0 * std::log(u1)) * std::cos(2.0 * M_PI * u2);
}

// ---- header formatter / parser (packet_header.cc) ---------------------------------------------
packet_header::packet_header(int header_len_bits, int payload_syms_, bool has_fec_)
    : header_len(header_len_bits), payload_syms(payload_syms_), has_fec(has_fec_),
      crc(16, 0x1021, 0xFFFF, 0, false, true) {}

static int add_field(uint8_t* buf, int header_len, int offset, unsigned long long val, int n_bits)
{
    int k = offset;
    for (int i = 0; i < n_bits && k < header_len; i++, k++) buf[k] = (uint8_t)((val >> i) & 1);   // bits_per_byte = 1
    return k;
}

// packet_header.cc:99-111 (note the reference's `i*len+j < d_header_len` bound)
void packet_header::pack_crc(const uint8_t* bits, std::vector<uint8_t>& out) const
{
    int len = (header_len - 16) / 8;
    if ((header_len - 16) % 8) ++len;
    out.assign(len, 0);
    for (int i = 0; i < len; ++i)
        for (int j = 0; j < 8 && i * len + j < header_len; j++)
            out[i] = (uint8_t)((out[i] << 1) | (bits[i * 8 + j] & 0x01));
}

// packet_header.cc:141-200 (scrambling disabled: scramble_bits defaults to False)
void packet_header::format(const frame_meta& m, int feedback_cnst, int feedback_fec, uint8_t* out)
{
    memset(out, 0, header_len);
    int k = 0;
    k = add_field(out, header_len, k, (unsigned)m.payload_len & 0xFFF, 12);
    k = add_field(out, header_len, k, (unsigned)header_number, 12);
    k = add_field(out, header_len, k, (unsigned)m.constellation & 0xF, 4);
    k = add_field(out, header_len, k, (unsigned)feedback_cnst & 0xF, 4);
    if (has_fec) {                                              // add_fec_header :114-139
        add_field(out, header_len, k + 0, (unsigned)m.tb_no, 12);
        add_field(out, header_len, k + 12, (unsigned)feedback_fec, 4);
        add_field(out, header_len, k + 16, (unsigned)m.tb_offset, 12);
        add_field(out, header_len, k + 28, (unsigned)m.fec_id, 4);
        add_field(out, header_len, k + 32, (unsigned)m.tb_payload, 16);
        k += 48;
    }
    std::vector<uint8_t> buf;
    pack_crc(out, buf);
    unsigned c = (unsigned)crc.compute(buf.data(), buf.size());
    add_field(out, header_len, k, c, 16);
    header_number = (header_number + 1) & 0x0FFF;
}

// packet_header.cc:232-312.  Returns false on CRC failure / unknown constellation.
bool packet_header::parse(const uint8_t* in, frame_meta* m, int* feedback_cnst, int* feedback_fec)
{
    auto get = [&](int& k, int nbits) { unsigned v = 0; for (int i = 0; i < nbits && k < header_len; i++, k++) v |= ((unsigned)in[k] & 1u) << i; return v; };
    int k = 0;
    unsigned payload_len = get(k, 12), frame_no = get(k, 12), cnst = get(k, 4), fcnst = get(k, 4);
    memset(m, 0, sizeof(*m));
    *feedback_fec = 0;
    if (has_fec) {
        m->tb_no = (int)get(k, 12); *feedback_fec = (int)get(k, 4); m->tb_offset = (int)get(k, 12);
        m->fec_id = (int)get(k, 4); m->tb_payload = (int)get(k, 16);
    }
    std::vector<uint8_t> buf;
    pack_crc(in, buf);
    unsigned calc = (unsigned)crc.compute(buf.data(), buf.size());
    for (int i = 0; i < 16 && k < header_len; i++, k++)
        if (((unsigned)in[k] & 1u) != ((calc >> i) & 1u)) return false;
    if (cnst && cnst <= (unsigned)max_cnst) constellation = (int)cnst;
    int bps = bits_per_symbol(constellation);
    if (bps == 0) return false;
    unsigned nsym = payload_len * 8 / bps;
    if (payload_len * 8 % bps) nsym++;
    m->payload_len = (int)payload_len; m->packet_len = (int)nsym; m->frame_no = (int)frame_no;
    m->constellation = constellation;
    *feedback_cnst = (int)fcnst;
    return true;
}

// ---- stream configuration ------------------------------------------------------------------------
int stream_cfg::capacity() const
{
    int n = 0;
    for (int i = 0; i < frame_len; i++) n += (int)ofdm.occupied_carriers[i % ofdm.occupied_carriers.size()].size();
    return n;
}
int stream_cfg::header_syms() const { return fec ? 2 : 1; }
int stream_cfg::frame_samples() const { return (frame_len + header_syms() + 2) * (ofdm.fft_len + ofdm.cp_len); }

// ---- TX framer (this repo's generator; see file header) -------------------------------------------
tx_framer::tx_framer(const stream_cfg& c, uint64_t seed) : cfg(c), r{ seed }, crc(4, 0x04C11DB7, 0xFFFFFFFF, 0xFFFFFFFF) {}

bool tx_framer::new_tb(int cnst, int fec_id)
{
    const ldpc_code& code = *cfg.codes[fec_id];
    const int bps = bits_per_symbol(cnst), cap = cfg.capacity();
    const int frame_bytes = cap * bps / 8;
    cur_ncws = compute_tb_len(code.N, frame_bytes * 8);
    int payload_max = (cur_ncws * code.K) / 8 - 4;             // fec_frame_bvb_impl.cc:369-370
    if (payload_max <= 0) throw std::runtime_error("Misconfiguration: TB should be able to carry at least 1 user data bit");
    std::vector<uint8_t> bytes(payload_max + 4);
    if (src) {
        if (src_pos + payload_max > src_len) return false;
        memcpy(bytes.data(), src + src_pos, payload_max);
        src_pos += payload_max;
    } else {
        for (int i = 0; i < payload_max; i++) bytes[i] = (uint8_t)(r.next() & 0xff);
    }
    crc.append_crc(bytes.data(), payload_max);                 // :394-399
    std::vector<uint8_t> bits((size_t)(payload_max + 4) * 8);
    repack to_bits(8, 1);
    to_bits.repack_lsb_first(bytes.data(), payload_max + 4, bits.data());
    enc.encode(bits.data(), (int)bits.size(), code, cur_ncws);
    tb_count++;
    cur_payload.assign(bytes.begin(), bytes.begin() + payload_max);
    cur_tb_fec = fec_id;
    return true;
}

// One frame of `capacity` chunks + its tags.  TBs whose last bit went out in this frame are
// appended to `done` (their user payload bytes), in order.
void tx_framer::next_frame(int cnst, int fec_id, uint8_t* chunks, frame_meta* m, std::vector<std::vector<uint8_t>>& done)
{
    const int cap = cfg.capacity(), bps = bits_per_symbol(cnst);
    memset(chunks, 0, cap);
    memset(m, 0, sizeof(*m));
    m->constellation = cnst; m->frame_no = frame_count & 0xFFF; frame_count++;
    if (!cfg.fec) {
        // no-FEC frame (ofdm_adaptive_frame_bb_impl.cc:236-262, full frames only)
        int in_bytes = cap * bps / 8 - 4;
        std::vector<uint8_t> bytes(in_bytes + 4);
        for (int i = 0; i < in_bytes; i++) bytes[i] = (uint8_t)(r.next() & 0xff);
        crc.append_crc(bytes.data(), in_bytes);
        repack rp(8, (uint8_t)bps);
        int nsym = rp.repack_lsb_first(bytes.data(), in_bytes + 4, chunks);
        m->payload_len = in_bytes + 4; m->packet_len = nsym;
        done.push_back(std::vector<uint8_t>(bytes.begin(), bytes.begin() + in_bytes));
        return;
    }
    auto bytes_to_syms = [&](int nbytes) { int s = nbytes * 8 / bps; if (nbytes * 8 % bps) ++s; return s; };
    int used = 0, payload = 0, offset = 0;
    bool started_new = false;
    // (1) finish / continue the TB in flight
    if (!enc.ready()) {
        int avail = (cap - used) * bps / 8;
        int out_bytes = std::min(align_bits_to_bytes(enc.remaining()), avail);
        enc.buf_out(&chunks[used], std::min(out_bytes * 8, enc.remaining()), bps);
        used += bytes_to_syms(out_bytes);
        payload += out_bytes;
        if (enc.ready()) done.push_back(cur_payload);
    }
    // (2) start the next TB at the next symbol boundary if there is room (tb_decoder.cc:112-124)
    if (enc.ready()) {
        int avail = (cap - used) * bps / 8;
        if (avail > 0 && new_tb(cnst, fec_id)) {
            offset = payload;
            started_new = true;
            int out_bytes = std::min(align_bits_to_bytes(enc.remaining()), avail);
            // keep ceil(payload*8/bps) (what the header parser derives, packet_header.cc:281-284)
            // large enough to cover the symbol-alignment pad bits
            int extra = (offset * 8) % bps ? bps - (offset * 8) % bps : 0;
            while (out_bytes > 1 && extra && (out_bytes * 8) % bps && (out_bytes * 8) % bps <= extra) --out_bytes;
            enc.buf_out(&chunks[used], std::min(out_bytes * 8, enc.remaining()), bps);
            used += bytes_to_syms(out_bytes);
            payload += out_bytes;
            if (enc.ready()) done.push_back(cur_payload);
        }
    }
    m->payload_len = payload;
    m->packet_len = bytes_to_syms(payload);
    m->fec_id = cur_tb_fec; m->tb_no = tb_count & 0xff; m->tb_offset = started_new ? offset : 0;
    m->tb_payload = enc.payload & 0xffff;
}

// ---- TX: map + allocate + IFFT + CP (stage 5) -------------------------------------------------------
void map_chunks(int cnst, const uint8_t* chunks, int n, cf32* out)    // chunks_to_symbols_bc_impl.cc:75-79
{
    const constellation& c = get_constellation(cnst);
    for (int i = 0; i < n; i++) out[i] = c.points[chunks[i] & ((1u << c.bps) - 1)];
}

// gr-digital ofdm_carrier_allocator_cvc + fft_vcc(inverse, shift) + ofdm_cyclic_prefixer (RECALLED)
void tx_modulate(const stream_cfg& cfg, const cf32* hdr_syms, const cf32* payload_syms, cf32* iq_out)
{
    const int N = cfg.ofdm.fft_len, cp = cfg.ofdm.cp_len, H = cfg.header_syms(), L = cfg.frame_len;
    const int shift = N / 2;
    auto bin = [&](int c) { if (c < 0) c += N; return (c + shift) % N; };
    std::vector<cf32> freq(N), time(N);
    int sym_out = 0;
    auto emit = [&](const cf32* f) {
        fft_shifted(f, time.data(), N, false);
        cf32* o = iq_out + (size_t)sym_out * (N + cp);
        for (int i = 0; i < cp; i++) o[i] = time[N - cp + i];
        for (int i = 0; i < N; i++) o[cp + i] = time[i];
        sym_out++;
    };
    emit(cfg.sync_word1.data());
    emit(cfg.sync_word2.data());
    const auto& occ = cfg.ofdm.occupied_carriers;
    const auto& pc = cfg.ofdm.pilot_carriers;
    const auto& ps = cfg.ofdm.pilot_symbols;
    int hi = 0, pi = 0;
    for (int s = 0; s < H + L; s++) {
        std::fill(freq.begin(), freq.end(), cf32(0, 0));
        const auto& set = occ[s % occ.size()];
        for (int c : set) freq[bin(c)] = (s < H) ? hdr_syms[hi++] : payload_syms[pi++];
        if (!pc.empty()) {
            const auto& pset = pc[s % pc.size()];
            const auto& psym = ps[s % ps.size()];
            for (size_t k = 0; k < pset.size(); k++) freq[bin(pset[k])] = psym[k];
        }
        emit(freq.data());
    }
}

// ---- RX stream ---------------------------------------------------------------------------------------
rx_stream::rx_stream(const stream_cfg& c)
    : cfg(c), feq(make_payload_cfg(c), feedback_decision(1.0, 5, c.mcs_snr, c.mcs_cnst, c.mcs_fec, c.initial_mcs)),
      crc(4, 0x04C11DB7, 0xFFFFFFFF, 0xFFFFFFFF)
{
    feq.eq.literal = cfg.eq_literal;
    if (cfg.fec) {
        int max_bps = 0;
        for (int x : cfg.constellations) max_bps = std::max(max_bps, bits_per_symbol(x));
        fdec.reset(new fec_decoder(cfg.codes, cfg.capacity(), max_bps));
    }
    const int N = cfg.ofdm.fft_len, L = cfg.frame_len;
    freq.resize((size_t)L * N); hard.resize((size_t)L * N); soft.resize((size_t)L * N);
    ser.resize(cfg.capacity()); llr.resize((size_t)cfg.capacity() * 8);
}

ofdm_cfg rx_stream::make_payload_cfg(const stream_cfg& c)
{
    ofdm_cfg o = c.ofdm;
    o.symbols_skipped = c.header_syms();       // ofdm_receiver.py:162
    o.alpha = 0.1f;                            // ofdm_receiver.py:163
    o.input_is_shifted = true;
    return o;
}

// header_payload_demux (CP strip) + fft_vcc on the payload symbols of one frame
void rx_stream::demod(const cf32* iq, cf32* out) const
{
    const int N = cfg.ofdm.fft_len, cp = cfg.ofdm.cp_len, L = cfg.frame_len, skip = cfg.header_syms() + 2;
    for (int i = 0; i < L; i++) fft_shifted(iq + (size_t)(skip + i) * (N + cp) + cp, out + (size_t)i * N, N, true);
}

// one frame: stages 1-4.  `out` gets the decoded payload bytes this frame released.
void rx_stream::frame(const cf32* iq, const cf32* taps, const frame_meta& m, rx_frame_out* o)
{
    const int L = cfg.frame_len;
    demod(iq, freq.data());
    frame_freq(freq.data(), taps, m, o);
    (void)L;
}

void rx_stream::frame_freq(const cf32* fr, const cf32* taps, const frame_meta& m, rx_frame_out* o)
{
    const int N = cfg.ofdm.fft_len, L = cfg.frame_len;
    o->bytes.clear(); o->tbs.clear(); o->n_llr = 0; o->crc_ok = -1;
    int emitted = feq.work(fr, L, taps, m, hard.data(), soft.data(), &o->res);
    o->hard = hard.data(); o->soft = soft.data();
    if (!emitted) return;
    // (payload_eq, 1) -> ofdm_serializer_vcc(..., packet_length, symbols_skipped = 1)
    int n = serialize(feq.eq.cfg, soft.data(), L, 0, m.packet_len, 1, ser.data());
    const constellation& cn = get_constellation(m.constellation);
    if (cfg.fec) {
        // soft_cf_impl.cc:84-148: sigma = sqrt(noise) is handed to calc_soft_dec as "npwr"
        double sigma = std::sqrt(o->res.noise);
        float tmp[8];
        for (int i = 0; i < n; i++) {
            cn.calc_soft_dec(ser[i], (float)sigma, tmp);
            for (int b = 0; b < cn.bps; b++) llr[(size_t)i * cn.bps + b] = tmp[cn.bps - 1 - b];   // std::reverse
        }
        o->n_llr = n * cn.bps; o->llr = llr.data();
        fdec->frame(llr.data(), m, frame_index, o->bytes, o->tbs);
    } else {
        // constellation_decoder_cb + frame_pack_bb (:73-123)
        std::vector<uint8_t> chunks(n);
        for (int i = 0; i < n; i++) chunks[i] = (uint8_t)cn.decide(ser[i]);
        std::vector<uint8_t> bytes((size_t)n * cn.bps / 8 + 2);
        repack rp((uint8_t)cn.bps, 8);
        int n_written = rp.repack_lsb_first(chunks.data(), n, bytes.data());
        if (n * cn.bps % 8) --n_written;
        bool ok = n_written >= 4 && crc.verify_crc(bytes.data(), n_written);
        o->crc_ok = ok;
        if (n_written >= 4) o->bytes.assign(bytes.begin(), bytes.begin() + (n_written - 4));
    }
    frame_index++;
    (void)N;
}

} // namespace orc
