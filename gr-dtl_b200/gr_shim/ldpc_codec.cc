// ldpc_dec / ldpc_enc -- lib/dtl/ldpc_dec.cc:23-77, lib/dtl/ldpc_enc.cc:21-80: fec_dec::decode / fec_enc::encode through
// dtl_b200_ldpc_decode / dtl_b200_ldpc_encode, one codeword per call as the interface asks (include/gnuradio/dtl/fec.h:17-31).
// The objects of one make_ldpc_*coders() call share one handle (fec id = position in the returned vector, entry 0 = nullptr).
#include "shim_common.h"

namespace gr {
namespace dtl {
using namespace b200;

namespace {
struct codec_handle {
    dtl_b200* h = nullptr;
    std::vector<std::string> alists;
    explicit codec_handle(const std::vector<std::string>& a) : alists(a)
    {
        plain_config pc(48, 1, alists);
        check(dtl_b200_create(&pc.cfg, &h), nullptr, "make_ldpc_decoders");          // unreadable alist -> DTL_ERR_IO -> runtime_error
    }
    ~codec_handle() { if (h) dtl_b200_destroy(h); }
};
} // namespace

class ldpc_dec_b200 : public fec_dec
{
public:
    ldpc_dec_b200(std::shared_ptr<codec_handle> ch, int fec_id) : d_ch(ch), d_id(fec_id)
    {
        int32_t n = 0, k = 0;
        check(dtl_b200_code_info(ch->h, fec_id, &n, &k, nullptr), ch->h, "ldpc_dec");
        d_n = n; d_k = k;
    }
    int decode(const float* in_data, int* nit, unsigned char* out_data) override
    {
        int32_t it = 0;
        check(dtl_b200_ldpc_decode(d_ch->h, d_id, in_data, 1, out_data, &it, DTL_MEM_HOST), d_ch->h, "ldpc_dec::decode");
        if (nit) *nit = it;
        return d_k;
    }
    int get_k() override { return d_k; }
    int get_n() override { return d_n; }
    const std::string& alist() const { return d_ch->alists[d_id - 1]; }
private:
    std::shared_ptr<codec_handle> d_ch;
    int d_id, d_n = 0, d_k = 0;
};

class ldpc_enc_b200 : public fec_enc
{
public:
    ldpc_enc_b200(std::shared_ptr<codec_handle> ch, int fec_id) : d_ch(ch), d_id(fec_id)
    {
        int32_t n = 0, k = 0;
        check(dtl_b200_code_info(ch->h, fec_id, &n, &k, nullptr), ch->h, "ldpc_enc");
        d_n = n; d_k = k;
    }
    void encode(const unsigned char* in_data, int len, unsigned char* out_data) override
    {
        if (len != d_k) throw std::invalid_argument("ldpc_enc::encode: len != k");
        check(dtl_b200_ldpc_encode(d_ch->h, d_id, in_data, 1, out_data, DTL_MEM_HOST), d_ch->h, "ldpc_enc::encode");
    }
    int get_k() override { return d_k; }
    int get_n() override { return d_n; }
private:
    std::shared_ptr<codec_handle> d_ch;
    int d_id, d_n = 0, d_k = 0;
};

std::vector<fec_dec::sptr> make_ldpc_decoders(const std::vector<std::string>& alist_fnames)
{
    std::vector<fec_dec::sptr> decoders{ nullptr };
    if (alist_fnames.empty()) return decoders;
    auto ch = std::make_shared<codec_handle>(alist_fnames);
    for (size_t i = 0; i < alist_fnames.size(); i++) decoders.push_back(std::make_shared<ldpc_dec_b200>(ch, (int)i + 1));
    return decoders;
}

std::vector<fec_enc::sptr> make_ldpc_encoders(const std::vector<std::string>& alist_fnames)
{
    std::vector<fec_enc::sptr> encoders{ nullptr };
    if (alist_fnames.empty()) return encoders;
    auto ch = std::make_shared<codec_handle>(alist_fnames);
    for (size_t i = 0; i < alist_fnames.size(); i++) encoders.push_back(std::make_shared<ldpc_enc_b200>(ch, (int)i + 1));
    return encoders;
}

// for ofdm_adaptive_fec_decoder: the alist behind a decoder object (empty if it is not one of ours)
std::string b200_alist_of(const fec_dec::sptr& d)
{
    auto p = std::dynamic_pointer_cast<ldpc_dec_b200>(d);
    return p ? p->alist() : std::string();
}

} // namespace dtl
} // namespace gr
