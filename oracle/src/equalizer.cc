// oracle/src/equalizer.cc -- TEST INFRASTRUCTURE ONLY (see oracle.hpp header).
// Restates lib/dtl/ofdm_adaptive_equalizer.cc:177-287 (ofdm_adaptive_equalizer_base),
// lib/dtl/ofdm_adaptive_frame_equalizer_vcvc_impl.cc:99-255 (work()), the gr-digital base
// class ofdm_equalizer_1d_pilots (RECALLED, Appendix A.3) and ofdm_serializer_vcc (A.10).
#include "contract.hpp"
#include "oracle.hpp"
#include <cmath>
#include <cstring>
#include <stdexcept>

namespace orc {


// ---- ofdm_equalizer_1d_pilots ctor (RECALLED) + equalizer.cc:177-214 ---------------------
equalizer::equalizer(const ofdm_cfg& c) : cfg(c)
{
    const int N = cfg.fft_len;
    const int shift = cfg.input_is_shifted ? N / 2 : 0;
    auto bin = [&](int carrier) { if (carrier < 0) carrier += N; return (carrier + shift) % N; };
    occupied.assign(N, 0);
    for (auto& set : cfg.occupied_carriers) for (int cidx : set) occupied[bin(cidx)] = 1;
    pilot_mask.assign(cfg.pilot_carriers.size(), std::vector<char>(N, 0));
    for (size_t s = 0; s < cfg.pilot_carriers.size(); s++)
        for (int cidx : cfg.pilot_carriers[s]) pilot_mask[s][bin(cidx)] = 1;
    // equalizer.cc:196-213: load ALL pilot symbol sets
    pilot_sym.assign(cfg.pilot_symbols.size(), std::vector<cf32>(N, cf32(0, 0)));
    pilot_inv.assign(cfg.pilot_symbols.size(), std::vector<cf32>(N, cf32(0, 0)));
    for (size_t i = 0; i < cfg.pilot_symbols.size(); i++) {
        if (cfg.pilot_carriers.empty()) break;
        size_t set = i % cfg.pilot_carriers.size();
        for (size_t k = 0; k < cfg.pilot_carriers[set].size() && k < cfg.pilot_symbols[i].size(); k++) {
            cf32 p = cfg.pilot_symbols[i][k];
            int b = bin(cfg.pilot_carriers[set][k]);
            pilot_sym[i][b] = p;
            double re = p.real(), im = p.imag(), d = re * re + im * im;
            pilot_inv[i][b] = cf32((float)(re / d), (float)(-im / d));
        }
    }
    channel_state.assign(N, cf32(1, 0));
    reset();
}

void equalizer::reset()
{
    pilot_carr_set = cfg.pilot_carriers.empty() ? 0 : cfg.symbols_skipped % (int)cfg.pilot_carriers.size();
    std::fill(channel_state.begin(), channel_state.end(), cf32(1, 0));
}

// equalizer.cc:217-268
void equalizer::frame_equalize(cf32* frame, cf32* frame_soft, int n_sym, const std::vector<cf32>& taps, int cnst)
{
    const int N = cfg.fft_len;
    if (!taps.empty()) channel_state = taps;
    const constellation& cn = get_constellation(cnst);
    const float alpha = cfg.alpha;
    const float oma = 1.0f - alpha;
    snr_est.reset();
    for (int i = 0; i < n_sym; i++) {
        const int pset = pilot_sym.empty() ? 0 : (cfg.symbols_skipped + i) % (int)pilot_sym.size();
        for (int k = 0; k < N; k++) {
            const bool is_pilot = !pilot_mask.empty() && pilot_mask[pilot_carr_set][k];
            if (!occupied[k] && !is_pilot) continue;
            cf32& y = frame[i * N + k];
            if (is_pilot) {
                if (literal) {
                    cf32 pilot_eq = y / channel_state[k];
                    snr_est.update(pilot_eq);
                    channel_state[k] = alpha * channel_state[k] + (1 - alpha) * y / pilot_sym[pset][k];
                } else {
                    cf32 pilot_eq = c_div(y, channel_state[k]);
                    snr_est.update(pilot_eq);
                    cf32 t(oma * y.real(), oma * y.imag());
                    channel_state[k] = c_axpy(alpha, channel_state[k], c_mul(t, pilot_inv[pset][k]));
                }
                y = pilot_sym[pset][k];
            } else {
                cf32 sym_eq, sym_est;
                if (literal) {
                    sym_eq = y / channel_state[k];
                    sym_est = cn.points[cn.decide(sym_eq)];
                    channel_state[k] = alpha * channel_state[k] + (1 - alpha) * y / sym_est;
                } else {
                    sym_eq = c_div(y, channel_state[k]);
                    unsigned idx = cn.decide(sym_eq);
                    sym_est = cn.points[idx];
                    cf32 t(oma * y.real(), oma * y.imag());
                    channel_state[k] = c_axpy(alpha, channel_state[k], c_mul(t, cn.inv_points[idx]));
                }
                y = sym_est;
                if (frame_soft) frame_soft[i * N + k] = sym_eq;
            }
        }
        if (!pilot_mask.empty()) pilot_carr_set = (pilot_carr_set + 1) % (int)pilot_mask.size();
    }
}

// equalizer.cc:281-287
double equalizer::get_noise()
{
    snr_est.snr();
    double noise_db = snr_est.noise_db();
    return std::pow(10, noise_db / 10.0);
}

// ---- frame equaliser block -----------------------------------------------------------------
frame_equalizer::frame_equalizer(const ofdm_cfg& c, const feedback_decision& f)
    : eq(c), fb(f), cp_len(c.cp_len), channel_state(c.fft_len, cf32(1, 0)) {}

int frame_equalizer::work(const cf32* in, int n_sym, const cf32* taps, const frame_meta& m,
                          cf32* out, cf32* out_soft, frame_result* res)
{
    const int N = eq.cfg.fft_len;
    const int carrier_offset = m.carrier_offset;
    // :114-145 tags
    if (taps) channel_state.assign(taps, taps + N);
    {
        int cur = m.frame_no, lost = 0;
        if (cur != expected_frame_no) {
            lost = (cur < expected_frame_no) ? 4096 + cur - expected_frame_no : cur - expected_frame_no;
        }
        lost_frames += lost;
        frames_count += lost + 1;
        expected_frame_no = (cur + 1) % 4096;
    }
    if (m.constellation == UNKNOWN) throw std::invalid_argument("Missing constellation tag.");   // :147-150

    // :154-166 integer carrier-offset shift of the flattened frame
    const long total = (long)N * n_sym;
    if (carrier_offset < 0) {
        memset((void*)out, 0, sizeof(cf32) * (size_t)(-carrier_offset));
        memcpy((void*)&out[-carrier_offset], (const void*)in, sizeof(cf32) * (size_t)(total + carrier_offset));
    } else {
        memset((void*)(out + total - carrier_offset), 0, sizeof(cf32) * (size_t)carrier_offset);
        memcpy((void*)out, (const void*)(in + carrier_offset), sizeof(cf32) * (size_t)(total - carrier_offset));
    }
    // :168-177 per-symbol phase correction (identity for offset 0, skipped: see DESIGN.md)
    if (carrier_offset != 0) {
        for (int i = 0; i < n_sym; i++) {
            double ph = -(2.0 * M_PI) * carrier_offset * cp_len / N * (i + 1);
            cf32 pc = eq.literal ? cf32(cosf((float)ph), sinf((float)ph)) : expj_contract(ph);
            for (int k = 0; k < N; k++)
                out[i * N + k] = eq.literal ? out[i * N + k] * pc : c_mul(out[i * N + k], pc);
        }
    }
    // :180-186
    eq.reset();
    bool failed = false;
    try {
        int c = m.constellation;
        if (c < 1 || c > MAX_CNST) throw std::invalid_argument("Invalid constellation");
        eq.frame_equalize(out, out_soft, n_sym, channel_state, c);
    } catch (const std::exception&) { failed = true; }
    channel_state = eq.channel_state;
    // :189-193
    if (carrier_offset != 0) {
        double ph = (2.0 * M_PI) * carrier_offset * cp_len / N * n_sym;
        cf32 pc = eq.literal ? cf32(cosf((float)ph), sinf((float)ph)) : expj_contract(ph);
        for (int k = 0; k < N; k++) channel_state[k] = eq.literal ? channel_state[k] * pc : c_mul(channel_state[k], pc);
    }
    // :196-216
    int fc = 0, ff = 0;
    double snr = eq.get_snr();
    fb.get_feedback(snr, &fc, &ff);
    if (res) {
        res->snr_db = snr; res->noise = eq.get_noise(); res->fb_cnst = fc; res->fb_fec = ff;
        res->emitted = m.payload_len ? n_sym : 0;
        res->lost_frames = (int32_t)lost_frames;
    }
    (void)failed;
    return m.payload_len ? n_sym : 0;   // :218-254
}

// gr-digital ofdm_serializer_vcc::work (RECALLED): occupied carriers in list order, stop at
// packet_len.  carr_offset is 0 in gr-dtl's wiring (python/dtl/ofdm_receiver.py:175-180 passes
// no carr_offset_key).
int serialize(const ofdm_cfg& c, const cf32* frame, int n_sym, int carr_offset, int packet_len,
              int symbols_skipped, cf32* out)
{
    const int N = c.fft_len, shift = c.input_is_shifted ? N / 2 : 0;
    const int nsets = (int)c.occupied_carriers.size();
    int set = symbols_skipped % nsets, n_out = 0;
    for (int i = 0; i < n_sym; i++) {
        for (int cidx : c.occupied_carriers[set]) {
            int b = cidx < 0 ? cidx + N : cidx;
            b = (b + shift) % N;
            out[n_out++] = frame[i * N + b + carr_offset];
            if (packet_len && n_out == packet_len) return n_out;
        }
        set = (set + 1) % nsets;
    }
    return n_out;
}

} // namespace orc
