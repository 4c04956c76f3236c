// Configuration objects of the payload branch: parameters only, exported into dtl_config.
//   ofdm_adaptive_payload_equalizer   lib/dtl/ofdm_adaptive_equalizer.cc:176-215 (ctor checks restated below)
//   ofdm_adaptive_feedback_decision   lib/dtl/ofdm_adaptive_feedback_decision.cc:33-53
#include "shim_common.h"

namespace gr {
namespace dtl {

std::size_t get_bits_per_symbol(constellation_type_t c)
{
    switch (c) {
    case constellation_type_t::BPSK: return 1;
    case constellation_type_t::QPSK: return 2;
    case constellation_type_t::PSK8: return 3;
    case constellation_type_t::QAM16: return 4;
    case constellation_type_t::QAM64: return 6;
    case constellation_type_t::QAM256: return 8;
    default: throw std::invalid_argument("Unknown constellation");
    }
}

void ofdm_adaptive_equalizer_base::export_config(dtl_config* c) const
{
    c->fft_len = d_fft_len;
    c->n_occupied_sets = (int32_t)d_occ_len.size(); c->occupied_set_len = d_occ_len.data(); c->occupied_carriers = d_occ.data();
    c->n_pilot_carrier_sets = (int32_t)d_pc_len.size(); c->pilot_carrier_set_len = d_pc_len.data(); c->pilot_carriers = d_pc.data();
    c->n_pilot_symbol_sets = (int32_t)d_ps_len.size(); c->pilot_symbol_set_len = d_ps_len.data(); c->pilot_symbols = d_ps.data();
    c->symbols_skipped = d_symbols_skipped; c->alpha = d_alpha; c->input_is_shifted = d_input_is_shifted ? 1 : 0;
    int mb = 1;
    for (auto t : d_cnsts) mb = std::max<int>(mb, (int)get_bits_per_symbol(t));
    c->max_bps = mb;
}

ofdm_adaptive_payload_equalizer::sptr ofdm_adaptive_payload_equalizer::make(
    int fft_len, const std::vector<constellation_type_t>& constellations, const std::shared_ptr<ofdm_adaptive_frame_snr_base> snr_est,
    const std::vector<std::vector<int>>& occupied_carriers, const std::vector<std::vector<int>>& pilot_carriers,
    const std::vector<std::vector<gr_complex>>& pilot_symbols, int symbols_skipped, float alpha, bool input_is_shifted)
{
    if (!snr_est) throw std::invalid_argument("SNR estimator missing");
    if (dynamic_cast<ofdm_adaptive_frame_snr_simple*>(snr_est.get()) == nullptr)
        throw std::invalid_argument("libdtl_b200 implements ofdm_adaptive_frame_snr_simple only");
    // digital::ofdm_equalizer_1d_pilots ctor: symbol set i goes with carrier set (symbols_skipped + i) % n_carrier_sets
    for (size_t i = 0; i < pilot_symbols.size(); i++) {
        if (pilot_carriers.empty()) throw std::invalid_argument("pilot symbols without pilot carriers");
        if (pilot_carriers[(symbols_skipped + i) % pilot_carriers.size()].size() != pilot_symbols[i].size())
            throw std::invalid_argument("pilot carriers and symbols do not match");
    }
    sptr e(new ofdm_adaptive_payload_equalizer());
    e->d_self = e;
    e->d_fft_len = fft_len; e->d_cnsts = constellations; e->d_symbols_skipped = symbols_skipped; e->d_alpha = alpha;
    e->d_input_is_shifted = input_is_shifted;
    for (auto t : constellations) get_bits_per_symbol(t);                       // "Unknown constellation"
    for (auto& v : occupied_carriers) { e->d_occ_len.push_back((int32_t)v.size()); for (int k : v) e->d_occ.push_back(k); }
    for (auto& v : pilot_carriers) { e->d_pc_len.push_back((int32_t)v.size()); for (int k : v) e->d_pc.push_back(k); }
    for (auto& v : pilot_symbols) { e->d_ps_len.push_back((int32_t)v.size()); for (auto z : v) { e->d_ps.push_back(z.real()); e->d_ps.push_back(z.imag()); } }
    if (e->d_occ_len.empty()) {                                                  // gr-digital default plan: every carrier but DC... (ofdm_equalizer_base)
        e->d_occ_len.push_back(fft_len - 1);
        for (int k = -fft_len / 2; k < fft_len / 2; k++) if (k) e->d_occ.push_back(k);
    }
    return e;
}

ofdm_adaptive_feedback_decision::sptr ofdm_adaptive_feedback_decision::make(
    double hysterisis, int decision_th, const std::vector<std::pair<double, ofdm_adaptive_feedback_t>>& lut, mcs_id_t initial_mcs)
{
    return std::make_shared<ofdm_adaptive_feedback_decision>(hysterisis, decision_th, lut, initial_mcs);
}

ofdm_adaptive_feedback_decision::ofdm_adaptive_feedback_decision(
    double hysterisis, int decision_counter, const std::vector<std::pair<double, ofdm_adaptive_feedback_t>>& lut, mcs_id_t initial_mcs)
    : d_hysteresis(hysterisis), d_decision_th(decision_counter), d_initial(initial_mcs)
{
    if (lut.empty()) throw std::runtime_error("Feedback lookup table is empty");             // feedback_decision.cc:45-47
    for (auto& e : lut) { d_snr.push_back(e.first); d_cnst.push_back((int32_t)e.second.first); d_fec.push_back(e.second.second); }
}

void ofdm_adaptive_feedback_decision::export_config(dtl_config* c) const
{
    c->n_mcs = (int32_t)d_snr.size(); c->mcs_snr = d_snr.data(); c->mcs_cnst = d_cnst.data(); c->mcs_fec = d_fec.data();
    c->initial_mcs = d_initial; c->hysteresis = d_hysteresis; c->decision_th = d_decision_th;
}

} // namespace dtl
} // namespace gr
