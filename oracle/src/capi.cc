// oracle/src/capi.cc -- TEST INFRASTRUCTURE ONLY (see oracle.hpp header).
// C entry points for ctypes (tests/, bench.py cpu_baseline / --impl reference, smoke()).
#include "txrx.hpp"
#include <chrono>
#include <cmath>
#include <cstring>
#include <thread>

using namespace orc;

extern "C" {

// ---- primitives ---------------------------------------------------------------------------------
int orc_bps(int cnst) { return bits_per_symbol(cnst); }

int orc_points(int cnst, float* out)
{
    const constellation& c = get_constellation(cnst);
    for (size_t i = 0; i < c.points.size(); i++) { out[2 * i] = c.points[i].real(); out[2 * i + 1] = c.points[i].imag(); }
    return (int)c.points.size();
}

double orc_min_dist2(int cnst) { return get_constellation(cnst).min_dist2(); }

void orc_decide(int cnst, const float* sym, int n, uint8_t* out)
{
    const constellation& c = get_constellation(cnst);
    for (int i = 0; i < n; i++) out[i] = (uint8_t)c.decide(cf32(sym[2 * i], sym[2 * i + 1]));
}

void orc_map(int cnst, const uint8_t* chunks, int n, float* out) { map_chunks(cnst, chunks, n, (cf32*)out); }

// soft_cf_impl.cc:143-148: calc_soft_dec then std::reverse (LSB-first on the wire)
void orc_soft_dec(int cnst, const float* sym, int n, float npwr, float* out)
{
    const constellation& c = get_constellation(cnst);
    float tmp[8];
    for (int i = 0; i < n; i++) {
        c.calc_soft_dec(cf32(sym[2 * i], sym[2 * i + 1]), npwr, tmp);
        for (int b = 0; b < c.bps; b++) out[(size_t)i * c.bps + b] = tmp[c.bps - 1 - b];
    }
}

uint64_t orc_crc(unsigned nbits, uint64_t poly, uint64_t init, uint64_t xorout, int in_refl, int out_refl,
                 const uint8_t* d, size_t n)
{
    return crc_engine(nbits, poly, init, xorout, in_refl, out_refl).compute(d, n);
}

int orc_repack(const uint8_t* in, size_t n, int bits_in, int bits_out, int lsb_first, uint8_t* out)
{
    repack r((uint8_t)bits_in, (uint8_t)bits_out);
    return lsb_first ? r.repack_lsb_first(in, n, out) : r.repack_msb_first(in, n, out);
}

// frame_pack_bb_impl.cc:73-123 for one frame of chunks; returns bytes emitted (n_written - 4)
int orc_frame_pack(int cnst, const uint8_t* chunks, int n, uint8_t* out, int* crc_ok)
{
    int bps = bits_per_symbol(cnst);
    repack r((uint8_t)bps, 8);
    std::vector<uint8_t> buf((size_t)n * bps / 8 + 2);
    int n_written = r.repack_lsb_first(chunks, n, buf.data());
    if (n * bps % 8) --n_written;
    crc_util crc(4, 0x04C11DB7, 0xFFFFFFFF, 0xFFFFFFFF);
    *crc_ok = n_written >= 4 ? crc.verify_crc(buf.data(), n_written) : 0;
    if (n_written < 4) return 0;
    memcpy(out, buf.data(), n_written - 4);
    return n_written - 4;
}

void* orc_fb_create(double hyst, int th, int n, const double* snr, const int* cnst, const int* fec, int initial)
{
    return new feedback_decision(hyst, th, std::vector<double>(snr, snr + n), std::vector<int>(cnst, cnst + n),
                                 std::vector<int>(fec, fec + n), initial);
}
void orc_fb_get(void* h, double snr, int* c, int* f) { ((feedback_decision*)h)->get_feedback(snr, c, f); }
void orc_fb_destroy(void* h) { delete (feedback_decision*)h; }

void orc_snr_simple(const float* x, int n, double* snr_db, double* noise_db, double* noise_lin)
{
    snr_simple s;
    for (int i = 0; i < n; i++) s.update(cf32(x[2 * i], x[2 * i + 1]));
    *snr_db = s.snr(); *noise_db = s.noise_db(); *noise_lin = std::pow(10, *noise_db / 10.0);
}

void orc_fft(const float* in, float* out, int n, int forward, int count)
{
    for (int i = 0; i < count; i++) fft_shifted((const cf32*)in + (size_t)i * n, (cf32*)out + (size_t)i * n, n, forward);
}

// ---- configuration --------------------------------------------------------------------------------
struct cfg_box { stream_cfg c; std::vector<std::unique_ptr<ldpc_code>> owned; };

void* orc_cfg_create(int fft_len, int cp_len, int frame_len, int n_occ, const int* occ,
                     int n_pcsets, int n_pil, const int* pil, int n_psets, const float* psym)
{
    cfg_box* b = new cfg_box;
    b->c.ofdm.fft_len = fft_len; b->c.ofdm.cp_len = cp_len; b->c.frame_len = frame_len;
    b->c.ofdm.occupied_carriers = { std::vector<int>(occ, occ + n_occ) };
    for (int s = 0; s < n_pcsets; s++) b->c.ofdm.pilot_carriers.push_back(std::vector<int>(pil + s * n_pil, pil + (s + 1) * n_pil));
    for (int s = 0; s < n_psets; s++) {
        std::vector<cf32> v;
        for (int k = 0; k < n_pil; k++) v.push_back(cf32(psym[2 * (s * n_pil + k)], psym[2 * (s * n_pil + k) + 1]));
        b->c.ofdm.pilot_symbols.push_back(v);
    }
    b->c.constellations = { BPSK, QPSK, PSK8, QAM16 };
    b->c.mcs_snr = { 0 }; b->c.mcs_cnst = { BPSK }; b->c.mcs_fec = { 0 };
    b->c.codes = { nullptr };
    b->c.sync_word1.assign(fft_len, cf32(0, 0)); b->c.sync_word2.assign(fft_len, cf32(0, 0));
    return b;
}
void orc_cfg_destroy(void* h) { delete (cfg_box*)h; }

// variable-length pilot sets (e.g. qa_ofdm_adaptive_frame_equalizer_vcvc.py:81-84: ((), (), (1,2,6,7), ()))
void orc_cfg_set_pilots(void* h, int n_csets, const int* cset_len, const int* carriers,
                        int n_ssets, const int* sset_len, const float* syms)
{
    ofdm_cfg& o = ((cfg_box*)h)->c.ofdm;
    o.pilot_carriers.clear(); o.pilot_symbols.clear();
    for (int s = 0, p = 0; s < n_csets; s++) { o.pilot_carriers.push_back(std::vector<int>(carriers + p, carriers + p + cset_len[s])); p += cset_len[s]; }
    for (int s = 0, p = 0; s < n_ssets; s++) {
        std::vector<cf32> v;
        for (int k = 0; k < sset_len[s]; k++, p++) v.push_back(cf32(syms[2 * p], syms[2 * p + 1]));
        o.pilot_symbols.push_back(v);
    }
}

int orc_cfg_add_code(void* h, const char* alist_path)
{
    cfg_box* b = (cfg_box*)h;
    std::unique_ptr<ldpc_code> c(new ldpc_code);
    if (!c->load(alist_path)) return -1;
    b->c.codes.push_back(c.get());
    b->owned.push_back(std::move(c));
    b->c.fec = true;
    return (int)b->c.codes.size() - 1;
}
void orc_cfg_set_fec(void* h, int fec) { ((cfg_box*)h)->c.fec = fec != 0; }
void orc_cfg_set_literal(void* h, int lit) { ((cfg_box*)h)->c.eq_literal = lit != 0; }
void orc_cfg_set_mcs(void* h, int n, const double* snr, const int* cnst, const int* fec, int initial)
{
    stream_cfg& c = ((cfg_box*)h)->c;
    c.mcs_snr.assign(snr, snr + n); c.mcs_cnst.assign(cnst, cnst + n); c.mcs_fec.assign(fec, fec + n); c.initial_mcs = initial;
}
void orc_cfg_set_constellations(void* h, int n, const int* ids) { ((cfg_box*)h)->c.constellations.assign(ids, ids + n); }
void orc_cfg_set_sync(void* h, const float* sw1, const float* sw2)
{
    stream_cfg& c = ((cfg_box*)h)->c;
    for (int i = 0; i < c.ofdm.fft_len; i++) { c.sync_word1[i] = cf32(sw1[2 * i], sw1[2 * i + 1]); c.sync_word2[i] = cf32(sw2[2 * i], sw2[2 * i + 1]); }
}
int orc_cfg_capacity(void* h) { return ((cfg_box*)h)->c.capacity(); }
int orc_cfg_frame_samples(void* h) { return ((cfg_box*)h)->c.frame_samples(); }
int orc_cfg_header_syms(void* h) { return ((cfg_box*)h)->c.header_syms(); }

// ---- LDPC ---------------------------------------------------------------------------------------------
int orc_code_info(void* h, int fec_id, int* N, int* K, int* M, int* permute /*[N] or null*/)
{
    const ldpc_code* c = ((cfg_box*)h)->c.codes.at(fec_id);
    if (!c) return -1;
    *N = c->N; *K = c->K; *M = c->M;
    if (permute) for (int i = 0; i < c->N; i++) permute[i] = c->permute[i];
    int edges = 0; for (auto& r : c->mlist) edges += (int)r.size();
    return edges;
}
void orc_ldpc_encode(void* h, int fec_id, const uint8_t* data, int count, uint8_t* out)
{
    const ldpc_code* c = ((cfg_box*)h)->c.codes.at(fec_id);
    for (int i = 0; i < count; i++) c->encode(data + (size_t)i * c->K, out + (size_t)i * c->N);
}
void orc_ldpc_decode(void* h, int fec_id, const float* llr, int count, uint8_t* out, int* nit)
{
    const ldpc_code* c = ((cfg_box*)h)->c.codes.at(fec_id);
    for (int i = 0; i < count; i++) c->decode(llr + (size_t)i * c->N, nit + i, out + (size_t)i * c->K);
}
// syndrome of a word given in WIRE order [parity | data]; returns number of unsatisfied checks
int orc_ldpc_syndrome(void* h, int fec_id, const uint8_t* wire)
{
    const ldpc_code* c = ((cfg_box*)h)->c.codes.at(fec_id);
    std::vector<uint8_t> y(c->N);
    for (int i = 0; i < c->N; i++) y[c->permute[i]] = wire[i];
    int bad = 0;
    for (auto& row : c->mlist) { uint8_t s = 0; for (int v : row) s ^= y[v]; bad += s; }
    return bad;
}

// ---- frame equaliser ------------------------------------------------------------------------------------
void* orc_feq_create(void* cfgh, int literal, int symbols_skipped, float alpha)
{
    stream_cfg& c = ((cfg_box*)cfgh)->c;
    ofdm_cfg o = c.ofdm; o.symbols_skipped = symbols_skipped; o.alpha = alpha; o.input_is_shifted = true;
    frame_equalizer* f = new frame_equalizer(o, feedback_decision(1.0, 5, c.mcs_snr, c.mcs_cnst, c.mcs_fec, c.initial_mcs));
    f->eq.literal = literal != 0;
    return f;
}
void orc_feq_destroy(void* h) { delete (frame_equalizer*)h; }
int orc_feq_work(void* h, const float* in, int n_sym, const float* taps, const frame_meta* m,
                 float* out_hard, float* out_soft, frame_result* res, float* taps_out)
{
    frame_equalizer* f = (frame_equalizer*)h;
    int r = f->work((const cf32*)in, n_sym, (const cf32*)taps, *m, (cf32*)out_hard, (cf32*)out_soft, res);
    if (taps_out) memcpy(taps_out, f->channel_state.data(), sizeof(cf32) * f->channel_state.size());
    return r;
}
int orc_serialize(void* cfgh, const float* frame, int n_sym, int packet_len, float* out)
{
    stream_cfg& c = ((cfg_box*)cfgh)->c;
    return serialize(c.ofdm, (const cf32*)frame, n_sym, 0, packet_len, 1, (cf32*)out);
}

// ---- FEC decoder block ------------------------------------------------------------------------------------
void* orc_fdec_create(void* cfgh)
{
    stream_cfg& c = ((cfg_box*)cfgh)->c;
    int max_bps = 0; for (int x : c.constellations) max_bps = std::max(max_bps, bits_per_symbol(x));
    return new fec_decoder(c.codes, c.capacity(), max_bps);
}
void orc_fdec_destroy(void* h) { delete (fec_decoder*)h; }
// returns number of bytes appended to out; tbs gets up to tb_cap results
int orc_fdec_frame(void* h, const float* llr, const frame_meta* m, int frame_index, uint8_t* out, int out_cap,
                   tb_result* tbs, int tb_cap, int* n_tbs)
{
    std::vector<uint8_t> bytes; std::vector<tb_result> r;
    ((fec_decoder*)h)->frame(llr, *m, frame_index, bytes, r);
    int n = std::min((int)bytes.size(), out_cap);
    memcpy(out, bytes.data(), n);
    *n_tbs = std::min((int)r.size(), tb_cap);
    for (int i = 0; i < *n_tbs; i++) tbs[i] = r[i];
    return n;
}

// ---- header ---------------------------------------------------------------------------------------------------
void* orc_hdr_create(int header_len_bits, int payload_syms, int has_fec, int max_cnst)
{
    packet_header* p = new packet_header(header_len_bits, payload_syms, has_fec != 0);
    p->max_cnst = max_cnst;
    return p;
}
void orc_hdr_destroy(void* h) { delete (packet_header*)h; }
void orc_hdr_format(void* h, const frame_meta* m, int fb_cnst, int fb_fec, uint8_t* out) { ((packet_header*)h)->format(*m, fb_cnst, fb_fec, out); }
int orc_hdr_parse(void* h, const uint8_t* in, frame_meta* m, int* fb_cnst, int* fb_fec) { return ((packet_header*)h)->parse(in, m, fb_cnst, fb_fec); }

// ---- synthetic TX for many streams ---------------------------------------------------------------------------
// Layout: frames are stream-major, index f = s*F + j.  Any output pointer may be null.
//   chunks   u8 [S*F][cap]           meta frame_meta[S*F]
//   iq       c64[S*F][frame_samples] (TX waveform through h_s = (1+1j)e^{j phi_s} plus AWGN)
//   taps     c64[S*F][N]             (N*h_s + estimation noise: stands in for ofdm_chanest_vcvc)
//   payload  u8 [S][payload_cap]     user bytes of every TB/frame completed inside the F frames
// cnst_per_frame (nullable, [S*F]) overrides `cnst`.
// TX FEC framer alone: user bytes of each stream -> chunks + tags, F consecutive frames per stream.
// state != nullptr: [S] opaque framer handles kept between calls (created on first use, freed by orc_frame_free).
int orc_frame_batch(void* cfgh, int S, int F, const int* cnst_per_frame, const int* fec_per_frame, const uint8_t* data,
                    int in_cap, const int* in_len, uint8_t* chunks, frame_meta* meta, int* consumed, int* started, void** state)
{
    const stream_cfg& cfg = ((cfg_box*)cfgh)->c;
    const int cap = cfg.capacity();
    try {
        for (int s = 0; s < S; s++) {
            tx_framer local(cfg, 1);
            tx_framer* fr = &local;
            if (state) { if (!state[s]) state[s] = new tx_framer(cfg, 1); fr = (tx_framer*)state[s]; }
            fr->src = data + (size_t)s * in_cap; fr->src_len = in_len ? std::min(in_len[s], in_cap) : in_cap; fr->src_pos = 0;
            const int tb0 = fr->tb_count;
            std::vector<std::vector<uint8_t>> done;
            for (int j = 0; j < F; j++) {
                const size_t f = (size_t)s * F + j;
                fr->next_frame(cnst_per_frame[f], fec_per_frame[f], chunks + f * cap, &meta[f], done);
            }
            if (consumed) consumed[s] = fr->src_pos;
            if (started) started[s] = fr->tb_count - tb0;
        }
    } catch (const std::exception&) { return -1; }
    return 0;
}
void orc_frame_free(void** state, int S) { for (int s = 0; s < S; s++) { delete (tx_framer*)state[s]; state[s] = nullptr; } }

int orc_tx_batch(void* cfgh, int S, int F, uint64_t seed, int cnst, int fec_id, double snr_db, int nthreads,
                 const int* cnst_per_frame, uint8_t* chunks, frame_meta* meta, float* iq, float* taps,
                 uint8_t* payload, int payload_cap, int* payload_len, int* payload_tbs)
{
    const stream_cfg& cfg = ((cfg_box*)cfgh)->c;
    const int cap = cfg.capacity(), N = cfg.ofdm.fft_len, H = cfg.header_syms(), fs = cfg.frame_samples();
    const int D0 = (int)cfg.ofdm.occupied_carriers[0].size();
    if (nthreads < 1) nthreads = 1;
    std::vector<int> status(S, 0);
    auto worker = [&](int t) {
        std::vector<uint8_t> ch(cap), hbits((size_t)H * D0);
        std::vector<cf32> psym(cap), hsym((size_t)H * D0), wave(fs);
        for (int s = t; s < S; s += nthreads) {
            tx_framer fr(cfg, seed * 1000003ull + (uint64_t)s * 7919ull + 1);
            rng nr{ seed * 69069ull + (uint64_t)s * 104729ull + 17 };
            packet_header hdr(H * D0, cfg.frame_len, cfg.fec);
            hdr.max_cnst = MAX_CNST;
            double phi = 2 * M_PI * nr.uniform();
            std::complex<double> h = std::complex<double>(1, 1) * std::exp(std::complex<double>(0, phi));
            double sigma_t = std::sqrt((double)N * std::norm(h) / std::pow(10.0, snr_db / 10.0));
            std::vector<std::vector<uint8_t>> done;
            for (int j = 0; j < F; j++) {
                size_t f = (size_t)s * F + j;
                frame_meta m;
                int c = cnst_per_frame ? cnst_per_frame[f] : cnst;
                fr.next_frame(c, fec_id, ch.data(), &m, done);
                if (chunks) memcpy(chunks + f * cap, ch.data(), cap);
                if (meta) meta[f] = m;
                if (iq) {
                    hdr.format(m, 0, 0, hbits.data());
                    map_chunks(BPSK, hbits.data(), H * D0, hsym.data());
                    map_chunks(c, ch.data(), cap, psym.data());
                    tx_modulate(cfg, hsym.data(), psym.data(), wave.data());
                    cf32* o = (cf32*)iq + f * fs;
                    const double sd = sigma_t / std::sqrt(2.0);
                    for (int i = 0; i < fs; i++) {
                        std::complex<double> v = std::complex<double>(wave[i]) * h;
                        o[i] = cf32((float)(v.real() + sd * nr.normal()), (float)(v.imag() + sd * nr.normal()));
                    }
                }
                if (taps) {
                    cf32* o = (cf32*)taps + f * N;
                    const double sd = sigma_t * std::sqrt((double)N) / std::sqrt(2.0);
                    for (int k = 0; k < N; k++)
                        o[k] = cf32((float)(N * h.real() + sd * nr.normal()), (float)(N * h.imag() + sd * nr.normal()));
                }
            }
            int pos = 0;
            for (auto& p : done) {
                if (payload && pos + (int)p.size() <= payload_cap) memcpy(payload + (size_t)s * payload_cap + pos, p.data(), p.size());
                else if (payload) status[s] = 1;
                pos += (int)p.size();
            }
            if (payload_len) payload_len[s] = pos;
            if (payload_tbs) payload_tbs[s] = (int)done.size();
        }
    };
    std::vector<std::thread> th;
    for (int t = 0; t < nthreads; t++) th.emplace_back(worker, t);
    for (auto& t : th) t.join();
    for (int s = 0; s < S; s++) if (status[s]) return -1;
    return 0;
}

// frequency-domain variant used by the stage-2 tests: Y = N*h*X + noise on [S*F][L][N] (payload symbols only)
int orc_tx_modulate_frame(void* cfgh, const float* hdr_syms, const float* payload_syms, float* iq)
{
    tx_modulate(((cfg_box*)cfgh)->c, (const cf32*)hdr_syms, (const cf32*)payload_syms, (cf32*)iq);
    return 0;
}

// ---- RX for many streams (the CPU baseline and the parity reference for the full chain) -------------------------
// in_kind: 0 = time-domain IQ [S*F][frame_samples]; 1 = frequency-domain payload symbols [S*F][L][N].
// Outputs (nullable): res[S*F]; llr [S*F][cap*max_bps] + n_llr[S*F]; hard/soft [S*F][L*N];
// bytes [S][bytes_cap] + nbytes[S]; tbs [S][tb_cap] + ntbs[S]; frame_crc[S*F] (no-FEC path, -1 if n/a).
// Returns wall seconds spent in the RX loop (threads joined).
double orc_rx_batch(void* cfgh, int in_kind, const float* in, const float* taps, const frame_meta* meta, int S, int F,
                    int nthreads, frame_result* res, float* llr, int llr_stride, int* n_llr, float* hard, float* soft,
                    uint8_t* bytes, int bytes_cap, int* nbytes, tb_result* tbs, int tb_cap, int* ntbs, int* frame_crc)
{
    const stream_cfg& cfg = ((cfg_box*)cfgh)->c;
    const int N = cfg.ofdm.fft_len, L = cfg.frame_len, fs = cfg.frame_samples();
    if (nthreads < 1) nthreads = 1;
    auto t0 = std::chrono::steady_clock::now();
    auto worker = [&](int t) {
        for (int s = t; s < S; s += nthreads) {
            rx_stream rx(cfg);
            rx_frame_out o;
            int pos = 0, nt = 0;
            for (int j = 0; j < F; j++) {
                size_t f = (size_t)s * F + j;
                const cf32* tp = taps ? (const cf32*)taps + f * N : nullptr;
                if (meta[f].reserved & 1) {        // header CRC failed: header_payload_demux emits no payload for this frame
                    if (res) { memset(&res[f], 0, sizeof res[f]); res[f].lost_frames = (int32_t)rx.feq.lost_frames; }
                    if (n_llr) n_llr[f] = 0;
                    if (frame_crc) frame_crc[f] = -1;
                    continue;
                }
                if (in_kind == 0) rx.frame((const cf32*)in + f * fs, tp, meta[f], &o);
                else rx.frame_freq((const cf32*)in + f * (size_t)L * N, tp, meta[f], &o);
                if (res) res[f] = o.res;
                if (hard) memcpy(hard + f * (size_t)L * N * 2, o.hard, sizeof(cf32) * (size_t)L * N);
                if (soft) memcpy(soft + f * (size_t)L * N * 2, o.soft, sizeof(cf32) * (size_t)L * N);
                if (n_llr) n_llr[f] = o.n_llr;
                if (llr && o.n_llr) memcpy(llr + f * (size_t)llr_stride, o.llr, sizeof(float) * std::min(o.n_llr, llr_stride));
                if (frame_crc) frame_crc[f] = o.crc_ok;
                if (bytes) { int n = std::min((int)o.bytes.size(), bytes_cap - pos); if (n > 0) memcpy(bytes + (size_t)s * bytes_cap + pos, o.bytes.data(), n); }
                pos += (int)o.bytes.size();
                for (auto& r : o.tbs) { if (tbs && nt < tb_cap) tbs[(size_t)s * tb_cap + nt] = r; nt++; }
            }
            if (nbytes) nbytes[s] = pos;
            if (ntbs) ntbs[s] = nt;
        }
    };
    std::vector<std::thread> th;
    for (int t = 0; t < nthreads; t++) th.emplace_back(worker, t);
    for (auto& t : th) t.join();
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

// CP strip + FFT only (stage 1) for `count` frames: in [count][frame_samples] -> out [count][L][N]
// header branch for S streams x F frames of IQ (header.cc): meta/taps as the payload branch receives them, ok[f] = header
// CRC passed, fb[f][2] = feedback fields the peer put in the header
int orc_rx_headers(void* cfgh, const float* iq, int S, int F, int nthreads, frame_meta* meta, float* taps, int* ok, int* fb)
{
    const stream_cfg& cfg = ((cfg_box*)cfgh)->c;
    if (cfg.sync_word1.empty() || cfg.sync_word2.empty()) return -1;
    const int N = cfg.ofdm.fft_len, fs = cfg.frame_samples();
    if (nthreads < 1) nthreads = 1;
    auto worker = [&](int t) {
        for (int s = t; s < S; s += nthreads) {
            rx_header rh(cfg);
            for (int j = 0; j < F; j++) {
                const size_t f = (size_t)s * F + j;
                int fc = 0, ff = 0;
                const bool good = rh.frame((const cf32*)iq + f * fs, &meta[f], taps ? (cf32*)taps + f * N : nullptr, &fc, &ff);
                if (ok) ok[f] = good ? 1 : 0;
                if (fb) { fb[2 * f] = fc; fb[2 * f + 1] = ff; }
            }
        }
    };
    std::vector<std::thread> th;
    for (int t = 0; t < nthreads; t++) th.emplace_back(worker, t);
    for (auto& t : th) t.join();
    return 0;
}

// ---- synchronisation front end (sync.cc) ------------------------------------------------------------------------------
void* orc_sync_create(void* cfgh, float threshold)
{
    const stream_cfg& cfg = ((cfg_box*)cfgh)->c;
    return new stream_sync(cfg.ofdm.fft_len, cfg.ofdm.cp_len, cfg.frame_samples(), threshold);
}
void orc_sync_destroy(void* h) { delete (stream_sync*)h; }
// one chunk of T samples; returns the number of frames completed (<= frame_cap), their samples in frames[n][FS], their start
// positions in starts; raw / fixed: trigger positions of this chunk (counts in n_raw / n_fixed, capacity trig_cap)
int orc_sync_push(void* h, const float* x, int T, float* frames, long long* starts, int frame_cap,
                  long long* raw, int* n_raw, long long* fixed, int* n_fixed, int trig_cap)
{
    stream_sync* s = (stream_sync*)h;
    std::vector<cf32> fr; std::vector<long long> st, rw, fx;
    s->push((const cf32*)x, T, fr, st, &rw, &fx);
    const int n = std::min((int)st.size(), frame_cap);
    if (frames) memcpy(frames, fr.data(), (size_t)n * s->FS * sizeof(cf32));
    for (int i = 0; i < n; i++) starts[i] = st[i];
    if (n_raw) { *n_raw = std::min((int)rw.size(), trig_cap); for (int i = 0; i < *n_raw; i++) raw[i] = rw[i]; }
    if (n_fixed) { *n_fixed = std::min((int)fx.size(), trig_cap); for (int i = 0; i < *n_fixed; i++) fixed[i] = fx[i]; }
    return (int)st.size();
}

void orc_rx_demod(void* cfgh, const float* iq, int count, float* out)
{
    const stream_cfg& cfg = ((cfg_box*)cfgh)->c;
    rx_stream rx(cfg);
    for (int f = 0; f < count; f++)
        rx.demod((const cf32*)iq + (size_t)f * cfg.frame_samples(), (cf32*)out + (size_t)f * cfg.frame_len * cfg.ofdm.fft_len);
}

} // extern "C"
