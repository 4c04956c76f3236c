// shim_driver -- TEST HARNESS: plays the scheduler for the block shims.  Reads one scenario file (records written by
// tests/test_gpu_shim.py), builds the payload branch of python/dtl/ofdm_receiver.py:152-207 through the blocks' make() functions,
// drives it ONE FRAME PER work() CALL with the stream tags header_payload_demux / the header parser would attach, and writes
// what comes out of every port (items, tags, messages) to the result file.
// usage: shim_driver scenario.bin result.bin
#include "../dtl_b200_blocks.h"
#include <cstdio>
#include <cstring>
#include <map>

using namespace gr::dtl;
typedef std::map<std::string, std::vector<char>> blob_t;

static blob_t read_blob(const char* path)
{
    blob_t b;
    FILE* f = fopen(path, "rb");
    if (!f) throw std::runtime_error(std::string("cannot open ") + path);
    for (;;) {
        uint32_t nl; uint64_t nb;
        if (fread(&nl, 4, 1, f) != 1) break;
        std::string name(nl, 0);
        if (fread(&name[0], 1, nl, f) != nl || fread(&nb, 8, 1, f) != 1) throw std::runtime_error("truncated scenario");
        std::vector<char> d(nb);
        if (nb && fread(d.data(), 1, nb, f) != nb) throw std::runtime_error("truncated scenario");
        b[name].swap(d);
    }
    fclose(f);
    return b;
}
static void write_rec(FILE* f, const std::string& name, const void* p, size_t nb)
{
    const uint32_t nl = (uint32_t)name.size(); const uint64_t n = nb;
    fwrite(&nl, 4, 1, f); fwrite(name.data(), 1, nl, f); fwrite(&n, 8, 1, f); if (nb) fwrite(p, 1, nb, f);
}
template <class T> static std::vector<T> vec(const blob_t& b, const char* k)
{
    auto it = b.find(k);
    if (it == b.end()) return {};
    std::vector<T> v(it->second.size() / sizeof(T));
    memcpy(v.data(), it->second.data(), v.size() * sizeof(T));
    return v;
}
static std::vector<std::vector<int>> sets(const std::vector<int32_t>& len, const std::vector<int32_t>& flat)
{
    std::vector<std::vector<int>> s; size_t o = 0;
    for (int32_t l : len) { s.emplace_back(flat.begin() + o, flat.begin() + o + l); o += l; }
    return s;
}
static long tag_long(const std::vector<gr::tag_t>& tags, const char* key, long dflt)
{
    for (auto& t : tags) if (t.key == pmt::mp(key)) return pmt::to_long(t.value);
    return dflt;
}

int main(int argc, char** argv)
{
    if (argc < 3) { fprintf(stderr, "usage: shim_driver scenario.bin result.bin\n"); return 2; }
    try {
        const blob_t sc = read_blob(argv[1]);
        const auto cfg = vec<int32_t>(sc, "cfg");           // fft_len, cp_len, frame_len, symbols_skipped, n_frames, fec, initial_mcs, decision_th
        const int N = cfg[0], cp = cfg[1], L = cfg[2], skipped = cfg[3], F = cfg[4], fec = cfg[5], initial_mcs = cfg[6], decision_th = cfg[7];
        const double hyst = vec<double>(sc, "hysteresis")[0];
        std::vector<constellation_type_t> cnsts;
        for (int32_t c : vec<int32_t>(sc, "cnsts")) cnsts.push_back((constellation_type_t)c);
        const auto occ = sets(vec<int32_t>(sc, "occ_len"), vec<int32_t>(sc, "occ"));
        const auto pcs = sets(vec<int32_t>(sc, "pc_len"), vec<int32_t>(sc, "pc"));
        std::vector<std::vector<gr_complex>> pss;
        {
            const auto len = vec<int32_t>(sc, "ps_len"); const auto flat = vec<gr_complex>(sc, "ps"); size_t o = 0;
            for (int32_t l : len) { pss.emplace_back(flat.begin() + o, flat.begin() + o + l); o += l; }
        }
        std::vector<std::pair<double, ofdm_adaptive_feedback_t>> lut;
        {
            const auto s = vec<double>(sc, "mcs_snr"); const auto c = vec<int32_t>(sc, "mcs_cnst"), f = vec<int32_t>(sc, "mcs_fec");
            for (size_t i = 0; i < s.size(); i++) lut.push_back({ s[i], { (constellation_type_t)c[i], (unsigned char)f[i] } });
        }
        std::vector<std::string> alists;
        {
            const auto raw = vec<char>(sc, "alists"); std::string cur;
            for (char ch : raw) { if (ch == '\n') { if (!cur.empty()) alists.push_back(cur); cur.clear(); } else cur.push_back(ch); }
            if (!cur.empty()) alists.push_back(cur);
        }
        const auto freq = vec<gr_complex>(sc, "freq");
        const auto taps = vec<gr_complex>(sc, "taps");
        const auto meta = vec<int32_t>(sc, "meta");
        const std::string frame_len_key = "frame_length", packet_len_key = "packet_length", frame_no_key = "frame_no";

        // ---- the flowgraph of ofdm_receiver.py:155-207
        auto payload_equalizer = ofdm_adaptive_payload_equalizer::make(N, cnsts, std::make_shared<ofdm_adaptive_frame_snr_simple>(0.1), occ, pcs, pss, skipped, 0.1f);
        auto payload_eq = ofdm_adaptive_frame_equalizer_vcvc::make(payload_equalizer->base(), ofdm_adaptive_feedback_decision::make(hyst, decision_th, lut, initial_mcs),
                                                                   cp, frame_len_key, frame_no_key, false, false);
        ofdm_adaptive_constellation_soft_cf::sptr payload_demod;
        ofdm_adaptive_fec_decoder::sptr fec_dec_blk;
        std::vector<fec_dec::sptr> ldpc_decs;
        int capacity = 0, max_bps = 1;
        for (int i = 0; i < L; i++) capacity += (int)occ[(1 + i) % occ.size()].size();
        for (auto c : cnsts) max_bps = std::max<int>(max_bps, (int)get_bits_per_symbol(c));
        if (fec) {
            ldpc_decs = make_ldpc_decoders(alists);
            payload_demod = ofdm_adaptive_constellation_soft_cf::make(cnsts, packet_len_key);
            fec_dec_blk = ofdm_adaptive_fec_decoder::make(ldpc_decs, capacity, max_bps, packet_len_key);
        }

        std::vector<gr_complex> all_hard, all_soft;
        std::vector<double> snr, noise, lost_rate;
        std::vector<int32_t> fb, emitted, n_llr, tb_rec, bits_len;
        std::vector<float> all_llr;
        std::vector<unsigned char> all_bits;
        std::string error;
        for (int f = 0; f < F && error.empty(); f++) {
            const int32_t* m = &meta[(size_t)f * 10];
            // tags on the first item of the frame, as header_payload_demux attaches them from the header_data message
            auto& tin = payload_eq->in_tags[0];
            const uint64_t at = payload_eq->nitems_read(0);
            auto tag = [&](const std::string& k, pmt::pmt_t v) { gr::tag_t t; t.offset = at; t.key = pmt::mp(k); t.value = v; tin.push_back(t); };
            tag("ofdm_sync_chan_taps", pmt::init_c32vector(N, std::vector<gr_complex>(taps.begin() + (size_t)f * N, taps.begin() + (size_t)(f + 1) * N)));
            tag("ofdm_sync_carr_offset", pmt::from_long(m[1]));
            tag(frame_len_key, pmt::from_long(L));
            if (m[0]) tag("constellation_tag_key", pmt::from_long(m[0]));
            tag(frame_no_key, pmt::from_long(m[2]));
            tag("payload_length_key", pmt::from_long(m[3]));
            tag(packet_len_key, pmt::from_long(m[4]));
            tag("fec_key", pmt::from_long(m[5]));
            tag("fec_tb_key", pmt::from_long(m[6]));
            tag("fec_offset_key", pmt::from_long(m[7]));
            tag("fec_tb_payload_key", pmt::from_long(m[8]));
            std::vector<gr_complex> hard((size_t)L * N), soft((size_t)L * N);
            gr_vector_int nin{ L };
            gr_vector_const_void_star in{ &freq[(size_t)f * L * N] };
            gr_vector_void_star out{ hard.data(), soft.data() };
            payload_eq->out_tags[0].clear(); payload_eq->out_tags[1].clear();
            int r;
            try { r = payload_eq->run(L, nin, in, out); } catch (const std::exception& e) { error = e.what(); break; }
            emitted.push_back(r);
            all_hard.insert(all_hard.end(), hard.begin(), hard.end());
            all_soft.insert(all_soft.end(), soft.begin(), soft.end());
            const auto& mon = payload_eq->messages["monitor"].back();
            const auto& fbm = payload_eq->messages["feedback_port"].back();
            snr.push_back(pmt::to_double(pmt::dict_ref(mon, pmt::mp("estimated_snr_tag_key"), nullptr)));
            noise.push_back(pmt::to_double(pmt::dict_ref(mon, pmt::mp("noise_tag_key"), nullptr)));
            lost_rate.push_back(pmt::to_double(pmt::dict_ref(mon, pmt::mp("lost_frames_rate"), nullptr)));
            fb.push_back((int32_t)pmt::to_long(pmt::dict_ref(fbm, pmt::mp("feedback_constellation_key"), nullptr)));
            fb.push_back((int32_t)pmt::to_long(pmt::dict_ref(fbm, pmt::mp("fec_feedback_key"), nullptr)));
            if (!fec || r == 0) continue;
            // digital.ofdm_serializer_vcc(fft_len, occupied_carriers, frame_length, packet_length, 1) on port 1 (test harness):
            // data carriers of every symbol in order, truncated to packet_length items; tags re-attached to the first item
            const std::vector<gr::tag_t> eq_tags = payload_eq->out_tags[1];
            const int packet_len = (int)tag_long(eq_tags, packet_len_key.c_str(), 0);
            std::vector<gr_complex> ser;
            for (int i = 0; i < r && (int)ser.size() < packet_len; i++)
                for (int k : occ[(1 + i) % occ.size()]) { if ((int)ser.size() < packet_len) ser.push_back(soft[(size_t)i * N + (k + N / 2)]); }
            auto& din = payload_demod->in_tags[0];
            for (auto t : eq_tags) { if (t.key == pmt::mp(frame_len_key)) continue; t.offset = payload_demod->nitems_read(0); din.push_back(t); }
            std::vector<float> llr((size_t)packet_len * max_bps);
            gr_vector_int nin2{ packet_len }; gr_vector_const_void_star in2{ ser.data() }; gr_vector_void_star out2{ llr.data() };
            payload_demod->out_tags[0].clear();
            int nl;
            try { nl = payload_demod->run((int)llr.size(), nin2, in2, out2); } catch (const std::exception& e) { error = e.what(); break; }
            n_llr.push_back(nl);
            all_llr.insert(all_llr.end(), llr.begin(), llr.begin() + nl);
            auto& fin = fec_dec_blk->in_tags[0];
            for (auto t : payload_demod->out_tags[0]) { t.offset = fec_dec_blk->nitems_read(0); fin.push_back(t); }
            std::vector<unsigned char> bits((size_t)capacity * max_bps * 4 + 4096);
            gr_vector_int nin3{ nl }; gr_vector_const_void_star in3{ llr.data() }; gr_vector_void_star out3{ bits.data() };
            fec_dec_blk->out_tags[0].clear();
            const size_t msgs_before = fec_dec_blk->messages["monitor"].size();
            int nb;
            try { nb = fec_dec_blk->run((int)bits.size(), nin3, in3, out3); } catch (const std::exception& e) { error = e.what(); break; }
            all_bits.insert(all_bits.end(), bits.begin(), bits.begin() + nb);
            for (auto& t : fec_dec_blk->out_tags[0]) bits_len.push_back((int32_t)pmt::to_long(t.value));
            auto& mm = fec_dec_blk->messages["monitor"];
            for (size_t i = msgs_before; i < mm.size(); i++)
                for (const char* k : { "tb_no", "crc_ok_count", "crc_fail_count", "avg_it" }) tb_rec.push_back((int32_t)pmt::to_long(pmt::dict_ref(mm[i], pmt::mp(k), nullptr)));
        }

        // ---- fec_dec::decode / fec_enc::encode objects one codeword at a time (include/gnuradio/dtl/fec.h:17-31)
        std::vector<unsigned char> cw_bits, cw_wire;
        std::vector<int32_t> cw_nit, code_nk;
        const auto cw_llr = vec<float>(sc, "cw_llr");
        const auto cw_fec = vec<int32_t>(sc, "cw_fec");
        if (!cw_fec.empty()) {
            if (ldpc_decs.empty()) ldpc_decs = make_ldpc_decoders(alists);
            auto encs = make_ldpc_encoders(alists);
            size_t o = 0;
            for (int32_t id : cw_fec) {
                auto d = ldpc_decs.at(id);
                const int n = d->get_n(), k = d->get_k();
                code_nk.push_back(n); code_nk.push_back(k);
                std::vector<unsigned char> out(k), wire(n);
                int nit = -1;
                if (d->decode(&cw_llr[o], &nit, out.data()) != k) throw std::runtime_error("decode() must return k");
                o += n;
                cw_bits.insert(cw_bits.end(), out.begin(), out.end());
                cw_nit.push_back(nit);
                encs.at(id)->encode(out.data(), k, wire.data());
                cw_wire.insert(cw_wire.end(), wire.begin(), wire.end());
            }
        }

        // ---- python/dtl/qa_ofdm_adaptive_chunks_to_symbols_bc.py:39-63: vector_source_b -> chunks_to_symbols_bc -> decoder_cb, one
        // tagged stream per constellation ("md_chunks" u8 [n_md][md_len], "md_cnst" i32 [n_md])
        std::vector<gr_complex> md_syms;
        std::vector<unsigned char> md_out;
        std::string md_error;
        {
            const auto mc = vec<unsigned char>(sc, "md_chunks");
            const auto mk = vec<int32_t>(sc, "md_cnst");
            if (!mk.empty()) {
                const int len = (int)(mc.size() / mk.size());
                std::vector<constellation_type_t> all;
                for (int32_t c : vec<int32_t>(sc, "cnsts")) all.push_back((constellation_type_t)c);
                auto mod = ofdm_adaptive_chunks_to_symbols_bc::make(all, "len_tag");
                auto demod = ofdm_adaptive_constellation_decoder_cb::make(all, "len_tag");
                for (size_t k = 0; k < mk.size() && md_error.empty(); k++) {
                    auto tagit = [&](gr::block& b, bool with_cnst) {
                        gr::tag_t t; t.offset = b.nitems_read(0);
                        if (with_cnst) { t.key = pmt::mp("constellation_tag_key"); t.value = pmt::from_long(mk[k]); b.in_tags[0].push_back(t); }
                        t.key = pmt::mp("len_tag"); t.value = pmt::from_long(len); b.in_tags[0].push_back(t);
                    };
                    std::vector<gr_complex> syms(len);
                    std::vector<unsigned char> back(len);
                    gr_vector_int nin{ len };
                    try {
                        tagit(*mod, mk[k] != 0);
                        gr_vector_const_void_star i1{ &mc[k * len] }; gr_vector_void_star o1{ syms.data() };
                        if (mod->run(len, nin, i1, o1) != len) throw std::runtime_error("chunks_to_symbols: short output");
                        tagit(*demod, true);
                        gr_vector_const_void_star i2{ syms.data() }; gr_vector_void_star o2{ back.data() };
                        if (demod->run(len, nin, i2, o2) != len) throw std::runtime_error("decoder_cb: short output");
                    } catch (const std::exception& e) { md_error = e.what(); break; }
                    md_syms.insert(md_syms.end(), syms.begin(), syms.end());
                    md_out.insert(md_out.end(), back.begin(), back.end());
                }
            }
        }

        FILE* fo = fopen(argv[2], "wb");
        if (!fo) throw std::runtime_error("cannot open result file");
        write_rec(fo, "error", error.data(), error.size());
        write_rec(fo, "hard", all_hard.data(), all_hard.size() * sizeof(gr_complex));
        write_rec(fo, "soft", all_soft.data(), all_soft.size() * sizeof(gr_complex));
        write_rec(fo, "snr", snr.data(), snr.size() * 8);
        write_rec(fo, "noise", noise.data(), noise.size() * 8);
        write_rec(fo, "lost_rate", lost_rate.data(), lost_rate.size() * 8);
        write_rec(fo, "fb", fb.data(), fb.size() * 4);
        write_rec(fo, "emitted", emitted.data(), emitted.size() * 4);
        write_rec(fo, "n_llr", n_llr.data(), n_llr.size() * 4);
        write_rec(fo, "llr", all_llr.data(), all_llr.size() * 4);
        write_rec(fo, "bits", all_bits.data(), all_bits.size());
        write_rec(fo, "bits_len", bits_len.data(), bits_len.size() * 4);
        write_rec(fo, "tb", tb_rec.data(), tb_rec.size() * 4);
        write_rec(fo, "cw_bits", cw_bits.data(), cw_bits.size());
        write_rec(fo, "cw_nit", cw_nit.data(), cw_nit.size() * 4);
        write_rec(fo, "cw_wire", cw_wire.data(), cw_wire.size());
        write_rec(fo, "code_nk", code_nk.data(), code_nk.size() * 4);
        write_rec(fo, "md_syms", md_syms.data(), md_syms.size() * sizeof(gr_complex));
        write_rec(fo, "md_out", md_out.data(), md_out.size());
        write_rec(fo, "md_error", md_error.data(), md_error.size());
        fclose(fo);
        return 0;
    } catch (const std::exception& e) {
        fprintf(stderr, "shim_driver: %s\n", e.what());
        return 1;
    }
}
