// oracle/src/util.cc -- TEST INFRASTRUCTURE ONLY (see oracle.hpp header).
// CRC, bit repacker, feedback decision, SNR estimator, FFT.
#include "oracle.hpp"
#include <cmath>
#include <limits>
#include <stdexcept>

namespace orc {

// ------------------------------------------------------------------------------------------
// gr::digital::crc(num_bits, poly, initial_value, final_xor, input_reflected, result_reflected)
// RECALLED (Rocksoft model).  Call sites: lib/dtl/crc_util.cc:19-27 (32, 0x04C11DB7, ~0, ~0,
// true, true) and lib/dtl/ofdm_adaptive_packet_header.cc:72 (16, 0x1021, 0xFFFF, 0, false, true).
// ------------------------------------------------------------------------------------------
static uint64_t reflect(uint64_t v, unsigned n)
{
    uint64_t r = 0;
    for (unsigned i = 0; i < n; i++) if (v & (1ull << i)) r |= 1ull << (n - 1 - i);
    return r;
}

crc_engine::crc_engine(unsigned nb, uint64_t p, uint64_t i, uint64_t x, bool ir, bool orf)
    : nbits(nb), poly(p), init(i), xorout(x), in_refl(ir), out_refl(orf)
{
    const uint64_t top = 1ull << (nbits - 1), mask = (nbits == 64) ? ~0ull : ((1ull << nbits) - 1);
    for (unsigned b = 0; b < 256; b++) {
        uint64_t r = (uint64_t)b << (nbits - 8);
        for (int k = 0; k < 8; k++) r = (r & top) ? ((r << 1) ^ poly) : (r << 1);
        table[b] = r & mask;
    }
}

uint64_t crc_engine::compute(const uint8_t* d, size_t n) const
{
    const uint64_t mask = (nbits == 64) ? ~0ull : ((1ull << nbits) - 1);
    uint64_t rem = init & mask;
    for (size_t i = 0; i < n; i++) {
        uint8_t byte = in_refl ? (uint8_t)reflect(d[i], 8) : d[i];
        rem = ((rem << 8) ^ table[((rem >> (nbits - 8)) ^ byte) & 0xff]) & mask;
    }
    if (out_refl) rem = reflect(rem, nbits);
    return (rem ^ xorout) & mask;
}

// lib/dtl/crc_util.cc:19-60
crc_util::crc_util(size_t len, uint64_t poly, uint64_t init, uint64_t xorout)
    : eng((unsigned)len * 8, poly, init, xorout, true, true), crc_len(len) {}

uint64_t crc_util::append_crc(uint8_t* buf, size_t len)
{
    uint64_t v = eng.compute(buf, len);
    for (unsigned i = 0; i < crc_len; ++i) buf[len + i] = (v >> (i * 8)) & 0xFF;
    return v;
}

bool crc_util::verify_crc(const uint8_t* buf, size_t len)
{
    size_t payload_len = len - crc_len;
    uint64_t v = eng.compute(buf, payload_len);
    bool ok = true;
    for (unsigned i = 0; i < crc_len; ++i) if (buf[payload_len + i] != ((v >> (i * 8)) & 0xFF)) ok = false;
    ++count;
    if (count == 0) failed = 0; else if (!ok) ++failed;
    return ok;
}

// ------------------------------------------------------------------------------------------
// lib/testbed/repack.cc:31-112 (stateful across calls, exactly as the reference)
// ------------------------------------------------------------------------------------------
int repack::repack_lsb_first(const uint8_t* in, size_t n_in, uint8_t* out, bool full_output_symbols)
{
    int bi = bits_in, bo = bits_out;
    size_t bytes_to_write = n_in * bi / bo;
    if (((n_in * bi) % bo) != 0) bytes_to_write += (int)(!full_output_symbols);
    size_t n_read = 0, n_written = 0;
    while (n_written < bytes_to_write && n_read < n_in) {
        if (out_index == 0) out[n_written] = 0;
        out[n_written] |= ((in[n_read] >> in_index) & 0x01) << out_index;
        in_index = (in_index + 1) % bi;
        out_index = (out_index + 1) % bo;
        if (in_index == 0) n_read++;
        if (out_index == 0) n_written++;
    }
    if (out_index) { n_written++; out_index = 0; }
    return (int)n_written;
}

int repack::repack_msb_first(const uint8_t* in, size_t n_in, uint8_t* out, bool full_output_symbols)
{
    int bi = bits_in, bo = bits_out;
    size_t bytes_to_write = n_in * bi / bo;
    if (((n_in * bi) % bo) != 0) bytes_to_write += (int)(!full_output_symbols);
    size_t n_read = 0, n_written = 0;
    while (n_written < bytes_to_write && n_read < n_in) {
        if (out_index == 0) out[n_written] = 0;
        out[n_written] |= ((in[n_read] >> (bi - 1 - in_index)) & 0x01) << (bo - 1 - out_index);
        in_index = (in_index + 1) % bi;
        out_index = (out_index + 1) % bo;
        if (in_index == 0) n_read++;
        if (out_index == 0) n_written++;
    }
    if (out_index) { n_written++; out_index = 0; }
    return (int)n_written;
}

// ------------------------------------------------------------------------------------------
// lib/dtl/ofdm_adaptive_feedback_decision.cc:33-96
// ------------------------------------------------------------------------------------------
feedback_decision::feedback_decision(double hyst, int th, const std::vector<double>& snr,
                                     const std::vector<int>& c, const std::vector<int>& f, int initial)
    : snr_th(snr), cnst(c), fec(f), hysteresis(hyst), decision_th(th), last(initial), next(initial)
{
    if (snr_th.empty()) throw std::runtime_error("Feedback lookup table is empty");
    snr_th[0] = std::numeric_limits<double>::min();   // :50
}

void feedback_decision::get_feedback(double snr, int* c, int* f)
{
    int id = last;
    if (snr < snr_th[id]) {
        // :70 `assert(current_mcs_id > 0)`: undefined in a release build when the estimate is <= 0 dB at MCS 0; this
        // restatement (and the CUDA path) define it as "no change, counter cleared"
        if (id > 0) update_decision(id - 1);
        else counter = 0;
    } else if ((size_t)(id + 1) < snr_th.size()) {
        if (snr > snr_th[id + 1] + hysteresis) update_decision(id + 1);
        else counter = 0;
    } else {
        counter = 0;
    }
    *c = cnst[last]; *f = fec[last];
}

void feedback_decision::update_decision(int id)
{
    if (id != next) { counter = 0; next = id; }
    else if (++counter >= decision_th) { counter = 0; last = next; }
}

// ------------------------------------------------------------------------------------------
// gr-digital mpsk_snr_est_simple (RECALLED; parity unpinned):
// Welford running mean / sum of squared deviations of |x| in double; signal = mean^2,
// noise = 2 * M2 / n; snr() in dB; noise() in dB.  Instantiated through
// include/gnuradio/dtl/ofdm_adaptive_frame_snr.h:27-46 and re-created on every reset().
// ------------------------------------------------------------------------------------------
void snr_simple::update(cf32 x)
{
    // |x| as glibc hypotf computes it: in double, rounded to float
    double re = x.real(), im = x.imag();
    float a = (float)std::sqrt(re * re + im * im);
    double v = a;
    double t = v - y1;
    double m = y1 + t / counter;
    y2 = y2 + t * (v - m);
    y1 = m;
    counter += 1.0;
}

double snr_simple::snr()
{
    signal = y1 * y1;
    noise = 2.0 * y2 / (counter - 1.0);
    return 10.0 * std::log10(signal / noise);
}

double snr_simple::noise_db() const { return 10.0 * std::log10(noise); }

// ------------------------------------------------------------------------------------------
// gr-fft fft_vcc(N, forward, window=(), shift=True) (RECALLED, Appendix A.1): forward =
// unnormalised DFT followed by a swap of the output halves; inverse = swap of the input
// halves followed by an unnormalised inverse DFT.  Radix-2 in fp32 with double twiddles
// (FFTW's own rounding is not reproducible; compare with a tolerance).
// ------------------------------------------------------------------------------------------
static void fft_radix2(std::vector<cf32>& a, bool forward)
{
    const int n = (int)a.size();
    for (int i = 1, j = 0; i < n; i++) {
        int bit = n >> 1;
        for (; j & bit; bit >>= 1) j ^= bit;
        j ^= bit;
        if (i < j) std::swap(a[i], a[j]);
    }
    for (int len = 2; len <= n; len <<= 1) {
        double ang = 2 * M_PI / len * (forward ? -1 : 1);
        for (int i = 0; i < n; i += len)
            for (int k = 0; k < len / 2; k++) {
                cf32 w((float)std::cos(ang * k), (float)std::sin(ang * k));
                cf32 u = a[i + k], v = a[i + k + len / 2] * w;
                a[i + k] = u + v; a[i + k + len / 2] = u - v;
            }
    }
}

void fft_shifted(const cf32* in, cf32* out, int n, bool forward)
{
    std::vector<cf32> a(n);
    if (forward) {
        for (int i = 0; i < n; i++) a[i] = in[i];
        fft_radix2(a, true);
        for (int i = 0; i < n; i++) out[(i + n / 2) % n] = a[i];
    } else {
        for (int i = 0; i < n; i++) a[i] = in[(i + n / 2) % n];
        fft_radix2(a, false);
        for (int i = 0; i < n; i++) out[i] = a[i];
    }
}

int compute_tb_len(int cw_len, int frame_len)
{
    int ncws = 1;
    if (frame_len > cw_len) ncws = 1 + frame_len / cw_len;
    return ncws;
}

int align_bits_to_bytes(int nbits)
{
    int nbytes = nbits / 8;
    if (nbits % 8) ++nbytes;
    return nbytes;
}

} // namespace orc
