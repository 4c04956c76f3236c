// oracle/src/header.cc -- TEST INFRASTRUCTURE ONLY (see oracle.hpp header).
//
// Header branch of gr-dtl's receiver (python/dtl/ofdm_receiver.py:100-150), SURVEY.md section 8(f) rows 1 and 2:
//   fft_vcc on the two sync symbols and the H header symbols           ofdm_receiver.py:101
//   digital.ofdm_chanest_vcvc(sync_word1, sync_word2, header_len)       ofdm_receiver.py:102-103   (gr-digital, RECALLED)
//   digital.ofdm_frame_equalizer_vcvc(ofdm_adaptive_header_equalizer(BPSK, symbols_skipped=0, alpha=0.1), cp_len,
//           frame_length_tag_key, propagate_channel_state=True, header_len)   ofdm_receiver.py:104-121
//           equaliser arithmetic: lib/dtl/ofdm_adaptive_equalizer.cc:161-174 -> frame_equalize :217-268
//   digital.ofdm_serializer_vcc + constellation_decoder_cb(BPSK)       ofdm_receiver.py:122-127
//   digital.packet_headerparser_b(ofdm_adaptive_packet_header)          ofdm_receiver.py:129-141,
//           lib/dtl/ofdm_adaptive_packet_header.cc:232-312
// What the payload branch then sees as stream tags: the parsed header fields, "ofdm_sync_carr_offset" from the channel
// estimator and "ofdm_sync_chan_taps" = the header equaliser's channel state AFTER the header symbols
// (propagate_channel_state).  A header whose CRC-16 fails is dropped by header_payload_demux: no payload is emitted.
//
// ofdm_chanest_vcvc is third-party (GNU Radio 3.10 gr-digital, not under /root/reference); restated from its published
// algorithm: integer carrier offset by Schmidl & Cox's B(g) over even g (argmax of |sum_k conj(S1[k+g]) conj(v[k]) S2[k+g]|,
// v = sync2/sync1 on the carriers sync1 uses), taps = S2[i] / sync2[i - offset] where sync2 is non-zero.  Parity unpinned
// (the reference has no test at this boundary); complex divisions follow this repo's arithmetic contract (oracle.hpp).
#include "txrx.hpp"
#include "contract.hpp"
#include <cstring>

namespace orc {

chanest::chanest(const std::vector<cf32>& s1, const std::vector<cf32>& s2, int n_data)
    : N((int)s1.size()), n_data_symbols(n_data), ref(s2), corr_v(s2)
{
    first_active = 0; last_active = N - 1;
    for (int i = 0; i < N; i++) if (ref[i] != cf32(0, 0)) { first_active = i; break; }
    for (int i = N - 1; i >= 0; i--) if (ref[i] != cf32(0, 0)) { last_active = i; break; }
    max_neg = -first_active;
    max_pos = N - last_active - 1;
    if (max_neg % 2) max_neg++;                                  // carrier offsets must be even
    if (max_pos % 2) max_pos--;
    for (int i = 0; i < N; i++) {
        if (s1[i] == cf32(0, 0)) corr_v[i] = cf32(0, 0);
        else corr_v[i] = c_div(corr_v[i], s1[i]);
    }
}

int chanest::carr_offset(const cf32* y1, const cf32* y2) const
{
    int best = 0;
    float bg_max = 0;
    for (int g = max_neg; g <= max_pos; g += 2) {
        cf32 tmp(0, 0);
        for (int k = 0; k < N; k++) {
            if (corr_v[k] != cf32(0, 0) && k + g >= 0 && k + g < N)
                tmp += c_mul(c_mul(std::conj(y1[k + g]), std::conj(corr_v[k])), y2[k + g]);
        }
        const float a = std::sqrt(tmp.real() * tmp.real() + tmp.imag() * tmp.imag());
        if (a > bg_max) { bg_max = a; best = g; }
    }
    return best;
}

void chanest::chan_taps(const cf32* y2, int off, cf32* taps) const
{
    for (int i = 0; i < N; i++) taps[i] = cf32(0, 0);
    int lo = 0, hi = N;
    if (off > 0) lo = off; else if (off < 0) hi = N + off;
    for (int i = lo; i < hi; i++)
        if (ref[i - off] != cf32(0, 0)) taps[i - off] = c_div(y2[i], ref[i - off]);
}

static ofdm_cfg make_header_cfg(const stream_cfg& c)
{
    ofdm_cfg o = c.ofdm;
    o.symbols_skipped = 0; o.alpha = 0.1f;                       // ofdm_receiver.py:113-114
    return o;
}

static feedback_decision dummy_fb()
{
    return feedback_decision(0.0, 1, std::vector<double>{ 0.0 }, std::vector<int>{ BPSK }, std::vector<int>{ 0 }, 0);
}

rx_header::rx_header(const stream_cfg& c)
    : cfg(c), est(c.sync_word1, c.sync_word2, c.header_syms()), feq(make_header_cfg(c), dummy_fb()),
      parser(c.header_syms() * (int)c.ofdm.occupied_carriers[0].size(), c.frame_len, c.fec)
{
    parser.max_cnst = MAX_CNST;
}

bool rx_header::frame(const cf32* iq, frame_meta* m, cf32* taps_out, int* fb_cnst, int* fb_fec)
{
    const int N = cfg.ofdm.fft_len, cp = cfg.ofdm.cp_len, H = cfg.header_syms();
    std::vector<cf32> f((size_t)(H + 2) * N), eqd((size_t)H * N), ser((size_t)H * N), taps(N);
    for (int i = 0; i < H + 2; i++) fft_shifted(iq + (size_t)i * (N + cp) + cp, f.data() + (size_t)i * N, N, true);
    const int off = est.carr_offset(f.data(), f.data() + N);
    est.chan_taps(f.data() + N, off, taps.data());
    frame_meta hm;
    memset(&hm, 0, sizeof hm);
    hm.constellation = BPSK; hm.carrier_offset = off; hm.payload_len = 1; hm.frame_no = feq.expected_frame_no;
    feq.work(f.data() + (size_t)2 * N, H, taps.data(), hm, eqd.data(), nullptr, nullptr);
    const int n = serialize(feq.eq.cfg, eqd.data(), H, 0, 0, 0, ser.data());
    std::vector<uint8_t> bits(n);
    for (int i = 0; i < n; i++) bits[i] = (uint8_t)get_constellation(BPSK).decide(ser[i]);
    int fc = 0, ff = 0;
    const bool ok = parser.parse(bits.data(), m, &fc, &ff);
    m->carrier_offset = off;
    if (!ok) { const int o = m->carrier_offset; memset(m, 0, sizeof *m); m->carrier_offset = o; m->reserved = 1; }
    if (taps_out) memcpy((void*)taps_out, feq.channel_state.data(), sizeof(cf32) * (size_t)N);
    if (fb_cnst) *fb_cnst = ok ? fc : 0;                        // no header_data message for a header that failed
    if (fb_fec) *fb_fec = ok ? ff : 0;
    return ok;
}

} // namespace orc
