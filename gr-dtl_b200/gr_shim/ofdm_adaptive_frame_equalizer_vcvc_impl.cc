// ofdm_adaptive_frame_equalizer_vcvc -- lib/dtl/ofdm_adaptive_frame_equalizer_vcvc_impl.cc with the arithmetic of work()
// (:154-196: carrier-offset shift, phase correction, frame_equalize, channel-state rotation, get_feedback) replaced by ONE call
// of dtl_b200_equalize; tag parsing (:112-152), messages (:196-216) and tag emission (:218-254) stay on the host as they were.
#include "shim_common.h"
#include <cstring>

namespace gr {
namespace dtl {
using namespace b200;

static const pmt::pmt_t FEEDBACK_PORT = pmt::mp("feedback_port");
static const pmt::pmt_t MONITOR_PORT = pmt::mp("monitor");

class ofdm_adaptive_frame_equalizer_vcvc_impl : public ofdm_adaptive_frame_equalizer_vcvc
{
public:
    ofdm_adaptive_frame_equalizer_vcvc_impl(ofdm_adaptive_equalizer_base::sptr equalizer,
                                            ofdm_adaptive_feedback_decision_base::sptr feedback_decision, int cp_len,
                                            const std::string& len_tag_key, const std::string& frame_no_key,
                                            bool propagate_channel_state, bool propagate_feedback_tags)
        : tagged_stream_block("ofdm_adaptive_frame_equalizer_vcvc",
                              io_signature::make(1, 1, sizeof(gr_complex) * equalizer->fft_len()),
                              io_signature::make(2, 2, sizeof(gr_complex) * equalizer->fft_len()), len_tag_key),
          d_fft_len(equalizer->fft_len()), d_cp_len(cp_len), d_eq(equalizer), d_decision_feedback(feedback_decision),
          d_frame_no_key(pmt::string_to_symbol(frame_no_key)), d_len_key(pmt::string_to_symbol(len_tag_key)),
          d_propagate_channel_state(propagate_channel_state), d_propagate_feedback_tags(propagate_feedback_tags),
          d_channel_state(equalizer->fft_len(), gr_complex(1, 0))
    {
        if (!d_decision_feedback) throw std::invalid_argument("feedback decision missing");
        message_port_register_out(FEEDBACK_PORT);
        message_port_register_out(MONITOR_PORT);
        set_tag_propagation_policy(TPP_DONT);
    }
    ~ofdm_adaptive_frame_equalizer_vcvc_impl() override { if (d_gpu) dtl_b200_destroy(d_gpu); }

    int work(int, gr_vector_int& ninput_items, gr_vector_const_void_star& input_items, gr_vector_void_star& output_items) override
    {
        const gr_complex* in = (const gr_complex*)input_items[0];
        gr_complex* out = (gr_complex*)output_items[0];
        gr_complex* out_soft = (gr_complex*)output_items[1];
        const int n_ofdm_sym = ninput_items[0];
        std::vector<tag_t> tags;
        get_tags_in_window(tags, 0, 0, 1);
        for (auto& t : tags) if (t.key == chan_taps_key()) d_channel_state = pmt::c32vector_elements(t.value);      // :116-118
        unsigned seen = 0;
        dtl_frame_meta m = meta_from_tags(tags, d_len_key, d_frame_no_key, &seen);
        if (!(seen & 1)) throw std::invalid_argument("Missing constellation tag.");                                // :150-152
        handle_for(n_ofdm_sym);
        dtl_frame_result r{};
        std::vector<gr_complex> taps_out(d_fft_len);
        // the reference's expected-frame-number bookkeeping (:119-138) lives in the handle's per-stream state; a frame without
        // the frame_no tag leaves it untouched there, which frame_no = expected reproduces
        if (!(seen & 4)) m.frame_no = d_expected_frame_no;
        const dtl_status st = dtl_b200_equalize(d_gpu, (const float*)in, (const float*)d_channel_state.data(), &m, 1, 1, (float*)out,
                                                (float*)out_soft, (float*)taps_out.data(), &r, nullptr, nullptr, nullptr, nullptr,
                                                nullptr, DTL_MEM_HOST);
        check(st, d_gpu, "ofdm_adaptive_frame_equalizer_vcvc");
        d_channel_state = taps_out;
        d_expected_frame_no = (m.frame_no + 1) % 4096;
        d_frames_seen++;
        const ofdm_adaptive_feedback_t feedback((constellation_type_t)r.fb_cnst, (unsigned char)r.fb_fec);

        pmt::pmt_t feedback_msg = pmt::dict_add(pmt::make_dict(), feedback_constellation_key(), pmt::from_long((unsigned char)feedback.first));
        feedback_msg = pmt::dict_add(feedback_msg, fec_feedback_key(), pmt::from_long(feedback.second));
        message_port_pub(FEEDBACK_PORT, feedback_msg);
        // monitor message (:210-216): the reference serialises these pairs with its protobuf builder; a PMT dict with the same keys here
        pmt::pmt_t mon = pmt::make_dict();
        mon = pmt::dict_add(mon, pmt::mp("constellation_key"), pmt::from_long((unsigned char)feedback.first));
        mon = pmt::dict_add(mon, pmt::mp("fec_key"), pmt::from_long(feedback.second));
        mon = pmt::dict_add(mon, pmt::mp("estimated_snr_tag_key"), pmt::from_double(r.snr_db));
        mon = pmt::dict_add(mon, pmt::mp("noise_tag_key"), pmt::from_double(r.noise));
        mon = pmt::dict_add(mon, pmt::mp("lost_frames_rate"), pmt::from_double(100.0 * (double)r.lost_frames / (double)(r.lost_frames + d_frames_seen)));
        message_port_pub(MONITOR_PORT, mon);

        if (!m.payload_len) return 0;                                                                               // :218
        for (int oi = 0; oi < 2; ++oi) {
            for (auto& t : tags)
                if (t.key != chan_taps_key() && t.key != d_len_key) add_item_tag(oi, nitems_written(0), t.key, t.value);
            if (d_propagate_channel_state) add_item_tag(oi, nitems_written(0), chan_taps_key(), pmt::init_c32vector(d_fft_len, d_channel_state));
            add_item_tag(oi, nitems_written(0), noise_tag_key(), pmt::from_double(r.noise));
            if (d_propagate_feedback_tags) {
                add_item_tag(oi, nitems_written(0), estimated_snr_tag_key(), pmt::from_double(r.snr_db));
                add_item_tag(oi, nitems_written(0), feedback_constellation_key(), pmt::from_long((unsigned char)feedback.first));
                add_item_tag(oi, nitems_written(0), fec_feedback_key(), pmt::from_long(feedback.second));
            }
        }
        return n_ofdm_sym;
    }

private:
    // the reference block takes whatever frame length the tagged stream has; the handle is sized for one (all frames of a
    // gr-dtl flowgraph have config.frame_length symbols) and rebuilt if that ever changes -- per-stream state restarts then
    void handle_for(int n_sym)
    {
        if (d_gpu && n_sym == d_frame_len) return;
        if (d_gpu) { dtl_b200_destroy(d_gpu); d_gpu = nullptr; }
        dtl_config cfg{};
        d_eq->export_config(&cfg);
        d_decision_feedback->export_config(&cfg);
        cfg.cp_len = d_cp_len; cfg.frame_len = n_sym; cfg.header_syms = cfg.symbols_skipped > 0 ? cfg.symbols_skipped : 1;
        cfg.fec_enabled = 0; cfg.n_codes = 0; cfg.ldpc_max_iterations = 15; cfg.ldpc_sigma2 = 2.0; cfg.device = 0;
        check(dtl_b200_create(&cfg, &d_gpu), nullptr, "dtl_b200_create");
        check(dtl_b200_reset(d_gpu, 1), d_gpu, "dtl_b200_reset");               // one stream per block instance
        d_frame_len = n_sym;
    }

    int d_fft_len, d_cp_len;
    ofdm_adaptive_equalizer_base::sptr d_eq;
    ofdm_adaptive_feedback_decision_base::sptr d_decision_feedback;
    pmt::pmt_t d_frame_no_key, d_len_key;
    bool d_propagate_channel_state, d_propagate_feedback_tags;
    std::vector<gr_complex> d_channel_state;
    dtl_b200* d_gpu = nullptr;
    int d_frame_len = 0, d_expected_frame_no = 0;
    long d_frames_seen = 0;
};

ofdm_adaptive_frame_equalizer_vcvc::sptr ofdm_adaptive_frame_equalizer_vcvc::make(
    ofdm_adaptive_equalizer_base::sptr equalizer, ofdm_adaptive_feedback_decision_base::sptr feedback_decision, int cp_len,
    const std::string& len_tag_key, const std::string& frame_no_key, bool propagate_channel_state, bool propagate_feedback_tags)
{
    return gnuradio::make_block_sptr<ofdm_adaptive_frame_equalizer_vcvc_impl>(
        equalizer, feedback_decision, cp_len, len_tag_key, frame_no_key, propagate_channel_state, propagate_feedback_tags);
}

} // namespace dtl
} // namespace gr
