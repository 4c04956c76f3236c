// ofdm_adaptive_constellation_soft_cf -- lib/dtl/ofdm_adaptive_constellation_soft_cf_impl.cc: the per-symbol calc_soft_dec +
// reverse loop (:143-148) becomes one dtl_b200_soft_demap call per frame; the frame walk, tag checks, tag re-emission and the
// scheduler hints (:76-141) are the reference's.
#include "shim_common.h"
#include <cmath>

namespace gr {
namespace dtl {
using namespace b200;

class ofdm_adaptive_constellation_soft_cf_impl : public ofdm_adaptive_constellation_soft_cf
{
public:
    ofdm_adaptive_constellation_soft_cf_impl(const std::vector<constellation_type_t>& constellations, const std::string& len_key)
        : gr::block("ofdm_adaptive_constellation_soft_cf", gr::io_signature::make(1, 1, sizeof(gr_complex)),
                    gr::io_signature::make(1, 1, sizeof(float))),
          d_len_key(pmt::string_to_symbol(len_key))
    {
        int mb = 1;
        for (auto c : constellations) { mb = std::max<int>(mb, (int)get_bits_per_symbol(c)); d_known.push_back(c); }   // "Unknown constellation" (:46-48)
        plain_config pc(48, mb, {});
        check(dtl_b200_create(&pc.cfg, &d_gpu), nullptr, "dtl_b200_create");
        set_max_noutput_items(65536);
        set_tag_propagation_policy(TPP_DONT);
    }
    ~ofdm_adaptive_constellation_soft_cf_impl() override { if (d_gpu) dtl_b200_destroy(d_gpu); }

    void forecast(int, gr_vector_int& ninput_items_required) override { ninput_items_required[0] = 1; }

    int general_work(int noutput_items, gr_vector_int& ninput_items, gr_vector_const_void_star& input_items,
                     gr_vector_void_star& output_items) override
    {
        auto in = static_cast<const gr_complex*>(input_items[0]);
        auto out = static_cast<float*>(output_items[0]);
        int read_index = 0, write_index = 0;
        while (read_index < ninput_items[0]) {
            std::vector<tag_t> tags;
            get_tags_in_range(tags, 0, nitems_read(0) + read_index, nitems_read(0) + read_index + 1);
            int test = 0, len = 0;
            constellation_type_t cnst = constellation_type_t::UNKNOWN;
            double noise = -1.0;                                            // no noise tag: sigma = 0.001 (:89)
            for (auto& tag : tags) {
                if (tag.key == constellation_tag_key()) { cnst = (constellation_type_t)pmt::to_long(tag.value); test |= 1; }
                else if (tag.key == d_len_key) { len = (int)pmt::to_long(tag.value); test |= 2; }
                else if (tag.key == noise_tag_key()) { noise = pmt::to_double(tag.value); test |= 4; }
                if (test == 7) break;
            }
            if (test != 7 && test != 3) throw std::runtime_error("Tags missing");
            bool known = false;
            for (auto c : d_known) known |= c == cnst;
            if (!known) throw std::invalid_argument("Unknown constellation");    // the reference dereferences a null constellation here
            const int bps = (int)get_bits_per_symbol(cnst);
            set_relative_rate(bps, 1);
            set_output_multiple(len * bps);
            if (read_index + len > ninput_items[0]) break;
            if (write_index + bps * len > noutput_items) { set_min_noutput_items(bps * len); break; }
            set_min_noutput_items(1);
            for (auto& tag : tags) {
                if (tag.key == d_len_key) add_item_tag(0, nitems_written(0) + write_index, d_len_key, pmt::from_long(len * bps));
                else add_item_tag(0, nitems_written(0) + write_index, tag.key, tag.value);
            }
            if (len > 0) {
                dtl_frame_meta m{};
                m.constellation = (int32_t)cnst; m.packet_len = len;
                check(dtl_b200_soft_demap(d_gpu, (const float*)(in + read_index), len, &m, &noise, 1, out + write_index, len * bps, DTL_MEM_HOST),
                      d_gpu, "ofdm_adaptive_constellation_soft_cf");
            }
            read_index += len;
            write_index += len * bps;
        }
        if (read_index > 0) consume_each(read_index);
        return write_index;
    }

private:
    pmt::pmt_t d_len_key;
    std::vector<constellation_type_t> d_known;
    dtl_b200* d_gpu = nullptr;
};

ofdm_adaptive_constellation_soft_cf::sptr ofdm_adaptive_constellation_soft_cf::make(const std::vector<constellation_type_t>& constellations,
                                                                                   const std::string& len_key)
{
    return gnuradio::make_block_sptr<ofdm_adaptive_constellation_soft_cf_impl>(constellations, len_key);
}

} // namespace dtl
} // namespace gr
