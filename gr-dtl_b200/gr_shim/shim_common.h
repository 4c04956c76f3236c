// shim_common.h -- what the block shims share: tag keys (lib/dtl/ofdm_adaptive_utils.cc:23-39), stream tags -> dtl_frame_meta,
// dtl_status -> the exception the reference throws at that point.
#pragma once
#include "dtl_b200_blocks.h"
#include <stdexcept>
#include <string>

namespace gr {
namespace dtl {
namespace b200 {

#define DTL_KEY(fn, str) inline const pmt::pmt_t& fn() { static const pmt::pmt_t k = pmt::string_to_symbol(str); return k; }
DTL_KEY(constellation_tag_key, "constellation_tag_key")
DTL_KEY(estimated_snr_tag_key, "estimated_snr_tag_key")
DTL_KEY(noise_tag_key, "noise_tag_key")
DTL_KEY(feedback_constellation_key, "feedback_constellation_key")
DTL_KEY(payload_length_key, "payload_length_key")
DTL_KEY(fec_key, "fec_key")
DTL_KEY(fec_feedback_key, "fec_feedback_key")
DTL_KEY(fec_tb_key, "fec_tb_key")
DTL_KEY(fec_offset_key, "fec_offset_key")
DTL_KEY(fec_tb_payload_key, "fec_tb_payload_key")
DTL_KEY(carr_offset_key, "ofdm_sync_carr_offset")
DTL_KEY(chan_taps_key, "ofdm_sync_chan_taps")
#undef DTL_KEY

// first-item tags of one frame -> POD.  `seen` gets one bit per field found, in dtl_frame_meta order (bit 0 = constellation).
inline dtl_frame_meta meta_from_tags(const std::vector<gr::tag_t>& tags, const pmt::pmt_t& len_key, const pmt::pmt_t& frame_no_key, unsigned* seen = nullptr)
{
    dtl_frame_meta m{};
    unsigned s = 0;
    for (auto& t : tags) {
        if (t.key == constellation_tag_key()) { m.constellation = (int32_t)pmt::to_long(t.value); s |= 1; }
        else if (t.key == carr_offset_key()) { m.carrier_offset = (int32_t)pmt::to_long(t.value); s |= 2; }
        else if (frame_no_key && t.key == frame_no_key) { m.frame_no = (int32_t)pmt::to_long(t.value); s |= 4; }
        else if (t.key == payload_length_key()) { m.payload_len = (int32_t)pmt::to_long(t.value); s |= 8; }
        else if (len_key && t.key == len_key) { m.packet_len = (int32_t)pmt::to_long(t.value); s |= 16; }
        else if (t.key == fec_key()) { m.fec_id = (int32_t)pmt::to_long(t.value); s |= 32; }
        else if (t.key == fec_tb_key()) { m.tb_no = (int32_t)pmt::to_long(t.value); s |= 64; }
        else if (t.key == fec_offset_key()) { m.tb_offset = (int32_t)pmt::to_long(t.value); s |= 128; }
        else if (t.key == fec_tb_payload_key()) { m.tb_payload = (int32_t)pmt::to_long(t.value); s |= 256; }
    }
    if (seen) *seen = s;
    return m;
}

// no exception crosses the C ABI; the shim turns the status back into what the reference block throws
inline void check(dtl_status st, const dtl_b200* h, const char* what)
{
    if (st == DTL_OK) return;
    const char* e = dtl_b200_last_error(h);
    const std::string text = std::string(what) + ": " + (e && *e ? e : "dtl_b200 error");
    if (st == DTL_ERR_INVALID_ARG) throw std::invalid_argument(text);
    throw std::runtime_error(text);
}

// minimal dtl_config for the stand-alone blocks, whose make() carries no carrier plan (soft_cf, fec_decoder, fec_dec):
// `capacity` data carriers per frame factored into frame_len symbols of D carriers of a 64-point plan.
struct plain_config {
    dtl_config cfg{};
    std::vector<int32_t> occ, occ_len, mcs_cnst, mcs_fec;
    std::vector<double> mcs_snr;
    std::vector<const char*> paths;
    plain_config(int capacity, int max_bps, const std::vector<std::string>& alists)
    {
        int D = 0;
        for (int d = 62; d >= 1; d--) if (capacity % d == 0 && capacity / d <= 40) { D = d; break; }
        if (!D) throw std::invalid_argument("frame_capacity cannot be laid out");
        for (int i = 0; i < D; i++) occ.push_back(i < D / 2 ? i - D / 2 : i - D / 2 + 1);
        occ_len = { D };
        mcs_snr = { 0.0 }; mcs_cnst = { DTL_CNST_BPSK }; mcs_fec = { alists.empty() ? 0 : 1 };
        for (auto& a : alists) paths.push_back(a.c_str());
        cfg.fft_len = 64; cfg.cp_len = 16; cfg.frame_len = capacity / D; cfg.header_syms = 1;
        cfg.n_occupied_sets = 1; cfg.occupied_set_len = occ_len.data(); cfg.occupied_carriers = occ.data();
        cfg.alpha = 0.1f; cfg.input_is_shifted = 1;
        cfg.n_mcs = 1; cfg.mcs_snr = mcs_snr.data(); cfg.mcs_cnst = mcs_cnst.data(); cfg.mcs_fec = mcs_fec.data();
        cfg.hysteresis = 1; cfg.decision_th = 5;
        cfg.n_codes = (int32_t)alists.size(); cfg.alist_paths = paths.empty() ? nullptr : paths.data();
        cfg.fec_enabled = alists.empty() ? 0 : 1; cfg.max_bps = max_bps;
        cfg.ldpc_max_iterations = 15; cfg.ldpc_mode = DTL_LDPC_REFERENCE; cfg.ldpc_sigma2 = 2.0;
    }
};

} // namespace b200
} // namespace dtl
} // namespace gr
