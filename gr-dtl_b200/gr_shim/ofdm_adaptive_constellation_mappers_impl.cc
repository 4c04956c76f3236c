// ofdm_adaptive_constellation_decoder_cb / ofdm_adaptive_chunks_to_symbols_bc -- the per-symbol loops of
// lib/dtl/ofdm_adaptive_constellation_decoder_cb_impl.cc:87-91 (decision_maker) and
// lib/dtl/ofdm_adaptive_chunks_to_symbols_bc_impl.cc:75-79 (map_to_points) become one dtl_b200_hard_demap / dtl_b200_map
// call per tagged stream; constructors, tag look-up and error texts are the reference's.
#include "shim_common.h"

namespace gr {
namespace dtl {
using namespace b200;

namespace {
constellation_type_t find_constellation_type(const std::vector<gr::tag_t>& tags)
{
    for (auto& t : tags) if (t.key == constellation_tag_key()) return (constellation_type_t)pmt::to_long(t.value);
    return constellation_type_t::UNKNOWN;
}
dtl_b200* make_handle(const std::vector<constellation_type_t>& constellations)
{
    int mb = 1;
    for (auto c : constellations) mb = std::max<int>(mb, (int)get_bits_per_symbol(c));          // "Unknown constellation"
    plain_config pc(48, mb, {});
    dtl_b200* h = nullptr;
    check(dtl_b200_create(&pc.cfg, &h), nullptr, "dtl_b200_create");
    return h;
}
} // namespace

class ofdm_adaptive_constellation_decoder_cb_impl : public ofdm_adaptive_constellation_decoder_cb
{
public:
    ofdm_adaptive_constellation_decoder_cb_impl(const std::vector<constellation_type_t>& constellations, const std::string& tsb_tag_key)
        : gr::tagged_stream_block("ofdm_adaptive_constellation_decoder_cb", gr::io_signature::make(1, 1, sizeof(gr_complex)),
                                  gr::io_signature::make(1, 1, sizeof(unsigned char)), tsb_tag_key),
          d_known(constellations), d_gpu(make_handle(constellations)) {}
    ~ofdm_adaptive_constellation_decoder_cb_impl() override { if (d_gpu) dtl_b200_destroy(d_gpu); }
    int work(int, gr_vector_int& ninput_items, gr_vector_const_void_star& input_items, gr_vector_void_star& output_items) override
    {
        std::vector<tag_t> tags;
        get_tags_in_range(tags, 0, nitems_read(0), nitems_read(0) + ninput_items[0]);
        const constellation_type_t c = find_constellation_type(tags);
        if (c == constellation_type_t::UNKNOWN) throw std::invalid_argument("Constellation type not found in tags");
        bool known = false;
        for (auto k : d_known) known |= k == c;
        if (!known) throw std::invalid_argument("Unknown constellation");
        if (ninput_items[0] > 0) {
            dtl_frame_meta m{};
            m.constellation = (int32_t)c; m.packet_len = ninput_items[0];
            check(dtl_b200_hard_demap(d_gpu, (const float*)input_items[0], ninput_items[0], &m, 1, (uint8_t*)output_items[0], DTL_MEM_HOST),
                  d_gpu, "ofdm_adaptive_constellation_decoder_cb");
        }
        return ninput_items[0];
    }
private:
    std::vector<constellation_type_t> d_known;
    dtl_b200* d_gpu;
};

ofdm_adaptive_constellation_decoder_cb::sptr ofdm_adaptive_constellation_decoder_cb::make(const std::vector<constellation_type_t>& constellations,
                                                                                         const std::string& tsb_tag_key)
{
    return gnuradio::make_block_sptr<ofdm_adaptive_constellation_decoder_cb_impl>(constellations, tsb_tag_key);
}

class ofdm_adaptive_chunks_to_symbols_bc_impl : public ofdm_adaptive_chunks_to_symbols_bc
{
public:
    ofdm_adaptive_chunks_to_symbols_bc_impl(const std::vector<constellation_type_t>& constellations, const std::string& tsb_tag_key)
        : gr::tagged_stream_block("ofdm_adaptive_chunks_to_symbols_bc", gr::io_signature::make(1, 1, sizeof(unsigned char)),
                                  gr::io_signature::make(1, 1, sizeof(gr_complex)), tsb_tag_key),
          d_known(constellations), d_gpu(make_handle(constellations)) {}
    ~ofdm_adaptive_chunks_to_symbols_bc_impl() override { if (d_gpu) dtl_b200_destroy(d_gpu); }
    int work(int, gr_vector_int& ninput_items, gr_vector_const_void_star& input_items, gr_vector_void_star& output_items) override
    {
        std::vector<tag_t> tags;
        get_tags_in_range(tags, 0, nitems_read(0), nitems_read(0) + 1);
        const constellation_type_t c = find_constellation_type(tags);
        if (c == constellation_type_t::UNKNOWN) throw std::invalid_argument("Constellation not found");
        bool known = false;
        for (auto k : d_known) known |= k == c;
        if (!known) throw std::invalid_argument("Unknown constellation");
        if (ninput_items[0] > 0) {
            dtl_frame_meta m{};
            m.constellation = (int32_t)c; m.packet_len = ninput_items[0];
            check(dtl_b200_map(d_gpu, (const uint8_t*)input_items[0], ninput_items[0], &m, 1, (float*)output_items[0], DTL_MEM_HOST),
                  d_gpu, "ofdm_adaptive_chunks_to_symbols_bc");
        }
        return ninput_items[0];
    }
private:
    std::vector<constellation_type_t> d_known;
    dtl_b200* d_gpu;
};

ofdm_adaptive_chunks_to_symbols_bc::sptr ofdm_adaptive_chunks_to_symbols_bc::make(const std::vector<constellation_type_t>& constellations,
                                                                                 const std::string& tsb_tag_key)
{
    return gnuradio::make_block_sptr<ofdm_adaptive_chunks_to_symbols_bc_impl>(constellations, tsb_tag_key);
}

} // namespace dtl
} // namespace gr
