// dtl_b200_blocks.h -- gr-dtl's hot-path blocks with their bodies replaced by calls into libdtl_b200.so (include/dtl_b200.h).
// Same namespace, class names, make() signatures, io signatures, stream tags and message ports as the reference:
//   ofdm_adaptive_payload_equalizer::make       include/gnuradio/dtl/ofdm_adaptive_equalizer.h:106-117
//   ofdm_adaptive_feedback_decision::make       include/gnuradio/dtl/ofdm_adaptive_feedback_decision.h:49-52
//   ofdm_adaptive_frame_equalizer_vcvc::make    include/gnuradio/dtl/ofdm_adaptive_frame_equalizer_vcvc.h:30-36
//   ofdm_adaptive_constellation_soft_cf::make   include/gnuradio/dtl/ofdm_adaptive_constellation_soft_cf.h:28
//   ofdm_adaptive_fec_decoder::make             include/gnuradio/dtl/ofdm_adaptive_fec_decoder.h:27
//   ofdm_adaptive_constellation_decoder_cb::make include/gnuradio/dtl/ofdm_adaptive_constellation_decoder_cb.h:34-35
//   ofdm_adaptive_chunks_to_symbols_bc::make     include/gnuradio/dtl/ofdm_adaptive_chunks_to_symbols_bc.h:32-35
//   fec_enc / fec_dec / make_ldpc_{en,de}coders include/gnuradio/dtl/fec.h:17-35
// In a gr-dtl tree these declarations ARE the reference's public headers (unchanged); only lib/dtl/*_impl.cc change.  Here the
// configuration objects (equaliser, feedback decision, SNR estimator) carry parameters only: their arithmetic runs on the GPU.
#pragma once
#ifdef DTL_B200_HAVE_GNURADIO
#include <gnuradio/block.h>
#include <gnuradio/tagged_stream_block.h>
#include <pmt/pmt.h>
#else
#include "standin/gr_standin.h"
#endif
#include "dtl_b200.h"

namespace gr {
namespace dtl {

enum class constellation_type_t { UNKNOWN = 0, BPSK, QPSK, PSK8, QAM16, QAM64, QAM256 };   // ofdm_adaptive_utils.h:22-28 (+ dtl_b200.h extensions)
std::size_t get_bits_per_symbol(constellation_type_t c);
typedef std::pair<constellation_type_t, unsigned char> ofdm_adaptive_feedback_t;
typedef int mcs_id_t;

// ---- SNR estimator handle (include/gnuradio/dtl/ofdm_adaptive_frame_snr.h:17-46): the estimator is reset per frame
// (ofdm_adaptive_equalizer.cc:226), so only its kind matters; the kernel implements the `simple` one ofdm_receiver.py:158 uses.
class ofdm_adaptive_frame_snr_base
{
public:
    virtual ~ofdm_adaptive_frame_snr_base() {}
    virtual double alpha() const = 0;
};
class ofdm_adaptive_frame_snr_simple : public ofdm_adaptive_frame_snr_base
{
public:
    explicit ofdm_adaptive_frame_snr_simple(double alpha) : d_alpha(alpha) {}
    double alpha() const override { return d_alpha; }
private:
    double d_alpha;
};

// ---- equaliser configuration (ofdm_adaptive_equalizer.h:28-117)
class ofdm_adaptive_equalizer_base
{
public:
    typedef std::shared_ptr<ofdm_adaptive_equalizer_base> sptr;
    virtual ~ofdm_adaptive_equalizer_base() {}
    // the accessor INTEGRATION.md section 4 adds to the reference class: protected members -> POD (pointers stay owned by *this)
    void export_config(dtl_config* cfg) const;
    int fft_len() const { return d_fft_len; }
    const std::vector<constellation_type_t>& constellations() const { return d_cnsts; }
protected:
    int d_fft_len = 0, d_symbols_skipped = 0;
    float d_alpha = 0.1f;
    bool d_input_is_shifted = true;
    std::vector<constellation_type_t> d_cnsts;
    std::vector<int32_t> d_occ_len, d_occ, d_pc_len, d_pc, d_ps_len;
    std::vector<float> d_ps;
};
class ofdm_adaptive_payload_equalizer : public ofdm_adaptive_equalizer_base
{
public:
    typedef std::shared_ptr<ofdm_adaptive_payload_equalizer> sptr;
    static sptr make(int fft_len,
                     const std::vector<constellation_type_t>& constellations,
                     const std::shared_ptr<ofdm_adaptive_frame_snr_base> snr_est,
                     const std::vector<std::vector<int>>& occupied_carriers = std::vector<std::vector<int>>(),
                     const std::vector<std::vector<int>>& pilot_carriers = std::vector<std::vector<int>>(),
                     const std::vector<std::vector<gr_complex>>& pilot_symbols = std::vector<std::vector<gr_complex>>(),
                     int symbols_skipped = 0,
                     float alpha = 0.1,
                     bool input_is_shifted = true);
    ofdm_adaptive_equalizer_base::sptr base() { return d_self.lock(); }
private:
    std::weak_ptr<ofdm_adaptive_payload_equalizer> d_self;
};

// ---- feedback decision configuration (ofdm_adaptive_feedback_decision.h:23-69)
class ofdm_adaptive_feedback_decision_base
{
public:
    typedef std::shared_ptr<ofdm_adaptive_feedback_decision_base> sptr;
    virtual ~ofdm_adaptive_feedback_decision_base() {}
    virtual void export_config(dtl_config* cfg) const = 0;
};
class ofdm_adaptive_feedback_decision : public ofdm_adaptive_feedback_decision_base
{
public:
    typedef std::shared_ptr<ofdm_adaptive_feedback_decision> sptr;
    static sptr make(double hysterisis, int decision_th,
                     const std::vector<std::pair<double, ofdm_adaptive_feedback_t>>& lut, mcs_id_t initial_mcs = 0);
    ofdm_adaptive_feedback_decision(double hysterisis, int decision_counter,
                                    const std::vector<std::pair<double, ofdm_adaptive_feedback_t>>& lut, mcs_id_t initial_mcs = 0);
    void export_config(dtl_config* cfg) const override;
private:
    std::vector<double> d_snr;
    std::vector<int32_t> d_cnst, d_fec;
    double d_hysteresis;
    int d_decision_th;
    mcs_id_t d_initial;
};

// ---- blocks
class ofdm_adaptive_frame_equalizer_vcvc : virtual public ::gr::tagged_stream_block
{
public:
    typedef std::shared_ptr<ofdm_adaptive_frame_equalizer_vcvc> sptr;
    static sptr make(::gr::dtl::ofdm_adaptive_equalizer_base::sptr equalizer,
                     ofdm_adaptive_feedback_decision_base::sptr feedback_decision,
                     int cp_len,
                     const std::string& len_tag_key = "len_key",
                     const std::string& frame_no_key = "frame_no_key",
                     bool propagate_channel_state = false,
                     bool propagate_feedback_tags = false);
};

class ofdm_adaptive_constellation_soft_cf : virtual public gr::block
{
public:
    typedef std::shared_ptr<ofdm_adaptive_constellation_soft_cf> sptr;
    static sptr make(const std::vector<constellation_type_t>& constellations, const std::string& len_key);
};

// lib/dtl side: include/gnuradio/dtl/ofdm_adaptive_constellation_decoder_cb.h:23-37, ofdm_adaptive_chunks_to_symbols_bc.h:23-38
class ofdm_adaptive_constellation_decoder_cb : virtual public gr::tagged_stream_block
{
public:
    typedef std::shared_ptr<ofdm_adaptive_constellation_decoder_cb> sptr;
    static sptr make(const std::vector<constellation_type_t>& constellations, const std::string& tsb_tag_key);
};
class ofdm_adaptive_chunks_to_symbols_bc : virtual public gr::tagged_stream_block
{
public:
    typedef std::shared_ptr<ofdm_adaptive_chunks_to_symbols_bc> sptr;
    static sptr make(const std::vector<constellation_type_t>& constellations, const std::string& tsb_tag_key);
};

class fec_enc
{
public:
    typedef std::shared_ptr<fec_enc> sptr;
    virtual ~fec_enc() {}
    virtual void encode(const unsigned char* in_data, int len, unsigned char* out_data) = 0;
    virtual int get_k() = 0;
    virtual int get_n() = 0;
};
class fec_dec
{
public:
    typedef std::shared_ptr<fec_dec> sptr;
    virtual ~fec_dec() {}
    virtual int decode(const float* in_data, int* nit, unsigned char* out_data) = 0;
    virtual int get_k() = 0;
    virtual int get_n() = 0;
};
std::vector<fec_enc::sptr> make_ldpc_encoders(const std::vector<std::string>& alist_fnames);
std::vector<fec_dec::sptr> make_ldpc_decoders(const std::vector<std::string>& alist_fnames);

class ofdm_adaptive_fec_decoder : virtual public gr::block
{
public:
    typedef std::shared_ptr<ofdm_adaptive_fec_decoder> sptr;
    static sptr make(const std::vector<fec_dec::sptr>& decoders, int frame_capacity, int max_bps, const std::string& len_key);
};

} // namespace dtl
} // namespace gr
