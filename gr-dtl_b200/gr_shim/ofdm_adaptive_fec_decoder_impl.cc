// ofdm_adaptive_fec_decoder -- lib/dtl/ofdm_adaptive_fec_decoder_impl.cc: tb_decoder::process_frame + ldpc_dec::decode + CRC
// (:141-204, lib/dtl/tb_decoder.cc:32-195) become one dtl_b200_fec_decode call per frame (the TB in flight is per-stream state
// of the handle); the frame walk, tag checks and the monitor message are the reference's.  The block emits UNPACKED bits
// (one per byte, LSB first, :160) with a len_key tag per transport block, as the reference does before ofdm_adaptive_fec_pack_bb.
#include "shim_common.h"

namespace gr {
namespace dtl {
using namespace b200;
std::string b200_alist_of(const fec_dec::sptr& d);

static const pmt::pmt_t MONITOR_PORT = pmt::mp("monitor");

class ofdm_adaptive_fec_decoder_impl : public ofdm_adaptive_fec_decoder
{
public:
    ofdm_adaptive_fec_decoder_impl(const std::vector<fec_dec::sptr>& decoders, int frame_capacity, int max_bps, const std::string& len_key)
        : gr::block("ofdm_adaptive_fec_decoder", gr::io_signature::make(1, 1, sizeof(float)), gr::io_signature::make(1, 1, sizeof(unsigned char))),
          d_len_key(pmt::string_to_symbol(len_key)), d_decoders(decoders), d_frame_capacity(frame_capacity)
    {
        if (decoders.size() < 2) throw std::runtime_error("Misconfiguration: decoder not found!");              // :66-68
        std::vector<std::string> alists;
        int max_k = 0;
        for (size_t i = 1; i < decoders.size(); i++) {
            const std::string a = decoders[i] ? b200_alist_of(decoders[i]) : std::string();
            if (a.empty()) throw std::invalid_argument("ofdm_adaptive_fec_decoder: decoders must come from make_ldpc_decoders");
            alists.push_back(a);
            max_k = std::max(max_k, decoders[i]->get_k());
        }
        plain_config pc(frame_capacity, max_bps, alists);
        check(dtl_b200_create(&pc.cfg, &d_gpu), nullptr, "dtl_b200_create");
        check(dtl_b200_reset(d_gpu, 1), d_gpu, "dtl_b200_reset");
        d_bytes.resize((size_t)frame_capacity * max_bps / 8 + 64 + (size_t)max_k);                               // a TB never carries more than its frames
        message_port_register_out(MONITOR_PORT);
        set_tag_propagation_policy(TPP_DONT);
    }
    ~ofdm_adaptive_fec_decoder_impl() override { if (d_gpu) dtl_b200_destroy(d_gpu); }

    int general_work(int, gr_vector_int& ninput_items, gr_vector_const_void_star& input_items, gr_vector_void_star& output_items) override
    {
        auto in = static_cast<const float*>(input_items[0]);
        auto out = static_cast<unsigned char*>(output_items[0]);
        int read_index = 0, write_index = 0;
        while (read_index < ninput_items[0]) {
            std::vector<tag_t> tags;
            get_tags_in_range(tags, 0, nitems_read(0) + read_index, nitems_read(0) + read_index + 1);
            unsigned seen = 0;
            dtl_frame_meta m = meta_from_tags(tags, d_len_key, pmt::pmt_t(), &seen);
            if ((seen & 17) != 17) throw std::runtime_error("Tags missing");                                    // constellation + len_key (:118-123)
            const int bps = (int)get_bits_per_symbol((constellation_type_t)m.constellation);
            const int frame_len = m.packet_len;                                                                   // LLRs of this frame
            if (read_index + frame_len > ninput_items[0]) break;
            if ((seen & (8 | 32 | 64 | 128 | 256)) != (8 | 32 | 64 | 128 | 256)) throw std::runtime_error("FEC tags missing");   // make_fec_info (:133-137)
            if (m.payload_len == 0) { read_index += frame_len; continue; }                                         // empty frame (:140-142)
            if (m.fec_id <= 0 || (size_t)m.fec_id >= d_decoders.size()) throw std::runtime_error("FEC tags missing");
            dtl_tb_result tb[8];
            int32_t ntb = 0, nbytes = 0;
            m.packet_len = frame_len / bps;                          // the ABI's packet_len counts constellation symbols
            check(dtl_b200_fec_decode(d_gpu, in + read_index, frame_len, &m, 1, 1, d_bytes.data(), (int32_t)d_bytes.size(), &nbytes,
                                      tb, 8, &ntb, DTL_MEM_HOST), d_gpu, "ofdm_adaptive_fec_decoder");
            size_t off = 0;
            for (int t = 0; t < ntb; t++) {
                d_crc_ok += tb[t].crc_ok ? 1 : 0;
                d_crc_fail += tb[t].crc_ok ? 0 : 1;
                if (tb[t].crc_ok) {
                    const int user_data_len = tb[t].nbytes * 8;
                    for (int b = 0; b < user_data_len; b++) out[write_index + b] = (d_bytes[off + b / 8] >> (b & 7)) & 1;
                    add_item_tag(0, nitems_written(0) + write_index, d_len_key, pmt::from_long(user_data_len));
                    write_index += user_data_len;
                    off += tb[t].nbytes;
                }
                auto dec = d_decoders[m.fec_id];
                pmt::pmt_t msg = pmt::make_dict();
                msg = pmt::dict_add(msg, pmt::mp("tb_no"), pmt::from_long(tb[t].tb_no));
                msg = pmt::dict_add(msg, pmt::mp("tb_code_k"), pmt::from_long(dec->get_k()));
                msg = pmt::dict_add(msg, pmt::mp("tb_code_n"), pmt::from_long(dec->get_n()));
                msg = pmt::dict_add(msg, pmt::mp("bps"), pmt::from_long(bps));
                msg = pmt::dict_add(msg, pmt::mp("crc_ok_count"), pmt::from_long(d_crc_ok));
                msg = pmt::dict_add(msg, pmt::mp("crc_fail_count"), pmt::from_long(d_crc_fail));
                msg = pmt::dict_add(msg, pmt::mp("tber"), pmt::from_double(100.0 * d_crc_fail / (double)(d_crc_fail + d_crc_ok)));
                msg = pmt::dict_add(msg, pmt::mp("avg_it"), pmt::from_long(tb[t].avg_it));
                message_port_pub(MONITOR_PORT, msg);
            }
            read_index += frame_len;
        }
        consume_each(read_index);
        return write_index;
    }

private:
    pmt::pmt_t d_len_key;
    std::vector<fec_dec::sptr> d_decoders;
    int d_frame_capacity;
    dtl_b200* d_gpu = nullptr;
    std::vector<uint8_t> d_bytes;
    long d_crc_ok = 0, d_crc_fail = 0;
};

ofdm_adaptive_fec_decoder::sptr ofdm_adaptive_fec_decoder::make(const std::vector<fec_dec::sptr>& decoders, int frame_capacity, int max_bps,
                                                               const std::string& len_key)
{
    return gnuradio::make_block_sptr<ofdm_adaptive_fec_decoder_impl>(decoders, frame_capacity, max_bps, len_key);
}

} // namespace dtl
} // namespace gr
