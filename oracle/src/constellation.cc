// oracle/src/constellation.cc -- TEST INFRASTRUCTURE ONLY (see oracle.hpp header).
// Constellation tables, hard decisions and soft decisions.
//   gr-dtl:     lib/dtl/ofdm_adaptive_utils.cc:41-70 (type -> constellation, bits/symbol),
//               lib/dtl/constellation.cc:18-24 (QPSK points scaled by 0.5)
//   GNU Radio:  gr-digital constellation_{bpsk,qpsk,8psk,16qam}, constellation::calc_soft_dec
//               (RECALLED from 3.10 -- SURVEY.md Appendix A.4/A.5; parity unpinned except
//               8PSK[0], BPSK[0] which the metric golden 6.568538665771484 pins).
//   Extension:  QAM64 / QAM256 (enum 5/6) do not exist in the reference; Gray-coded
//               square QAM, unit average energy, index = (gray(im level) << b/2) | gray(re level).
#include <vector>
#include "oracle.hpp"
#include <cmath>
#include <stdexcept>

namespace orc {

int bits_per_symbol(int cnst)
{
    switch (cnst) {
    case BPSK: return 1; case QPSK: return 2; case PSK8: return 3; case QAM16: return 4;
    case QAM64: return 6; case QAM256: return 8; default: return 0;
    }
}

static constellation make(int type)
{
    constellation c; c.type = type; c.bps = bits_per_symbol(type);
    switch (type) {
    case BPSK:
        c.points = { cf32(-1, 0), cf32(1, 0) };
        break;
    case QPSK: {
        // gr-digital: #define SQRT_TWO 0.707107 ; gr-dtl scales by d_factor = 0.5 (double)
        const double s = 0.707107;
        cf32 base[4] = { cf32(-s, -s), cf32(s, -s), cf32(-s, s), cf32(s, s) };
        for (auto& b : base) c.points.push_back(cf32(b.real() * 0.5, b.imag() * 0.5));
        break; }
    case PSK8: {
        const float angle = (float)(M_PI / 8.0);
        static const int mult[8] = { 1, 7, 15, 9, 3, 5, 13, 11 };
        for (int i = 0; i < 8; i++) c.points.push_back(cf32(std::cos(mult[i] * angle), std::sin(mult[i] * angle)));
        break; }
    case QAM16: {
        const float level = std::sqrt(float(0.1));
        c.level = level; c.inv2l = 1.0f / (2.0f * level);
        for (int i = 0; i < 16; i++) {
            float re = ((i & 2) ? 3.0f : 1.0f) * level; if (i & 4) re = -re;
            float im = ((i & 1) ? 3.0f : 1.0f) * level; if (!(i & 8)) im = -im;
            c.points.push_back(cf32(re, im));
        }
        break; }
    case QAM64: case QAM256: {
        const int hb = c.bps / 2, m = 1 << hb;
        const float level = (float)(1.0 / std::sqrt(type == QAM64 ? 42.0 : 170.0));
        c.level = level; c.inv2l = 1.0f / (2.0f * level);
        c.points.resize((size_t)1 << c.bps);
        for (int ai = 0; ai < m; ai++) for (int ar = 0; ar < m; ar++) {
            int gi = ai ^ (ai >> 1), gr = ar ^ (ar >> 1);
            c.points[(gi << hb) | gr] = cf32((float)(2 * ar - (m - 1)) * level, (float)(2 * ai - (m - 1)) * level);
        }
        break; }
    default: throw std::invalid_argument("Unknown constellation");
    }
    for (auto& p : c.points) {
        double re = p.real(), im = p.imag(), d = re * re + im * im;
        c.inv_points.push_back(cf32((float)(re / d), (float)(-im / d)));
    }
    return c;
}

const constellation& get_constellation(int cnst)
{
    // built once under the compiler's thread-safe static initialisation: the batch entry points call this from several
    // worker threads at once, and a hand-rolled "init" flag here let a second thread rebuild the tables under a reader
    static const std::vector<constellation> all = [] {
        std::vector<constellation> v(MAX_CNST + 1);
        for (int i = 1; i <= MAX_CNST; i++) v[i] = make(i);
        return v;
    }();
    if (cnst < 1 || cnst > MAX_CNST) throw std::invalid_argument("Invalid constellation");
    return all[cnst];
}

static inline int qam_axis(float x, float inv2l, int m)
{
    float t = x * inv2l; t = t + (float)(m / 2);
    int a = (int)std::floor(t);
    if (a < 0) a = 0; if (a > m - 1) a = m - 1;
    return a;
}

unsigned constellation::decide(cf32 s) const
{
    const float re = s.real(), im = s.imag();
    switch (type) {
    case BPSK: return re > 0;
    case QPSK: return 2 * (im > 0) + (re > 0);
    case PSK8: {
        unsigned ret = 0;
        if (std::fabs(re) <= std::fabs(im)) ret = 4;
        if (re <= 0) ret |= 1;
        if (im <= 0) ret |= 2;
        return ret; }
    case QAM16: {
        const float level = std::sqrt(float(0.1));
        unsigned ret = 0;
        if (im > 0) ret |= 8;
        if (re < 0) ret |= 4;
        if (std::fabs(re) > 2.0f * level) ret |= 2;
        if (std::fabs(im) > 2.0f * level) ret |= 1;
        return ret; }
    case QAM64: case QAM256: {
        const int hb = bps / 2, m = 1 << hb;
        int ar = qam_axis(re, inv2l, m), ai = qam_axis(im, inv2l, m);
        return (unsigned)(((ai ^ (ai >> 1)) << hb) | (ar ^ (ar >> 1))); }
    }
    return 0;
}

// gr-digital constellation::calc_soft_dec(sample, npwr) (RECALLED, 3.10): squared Euclidean
// distance, d_i = expf(-dist/npwr), per-bit sums over the point index (no pre-diff code),
// out[k-1-j] = logf(one_j) - logf(zero_j), scale factor 1.
void constellation::calc_soft_dec(cf32 sample, float npwr, float* s) const
{
    const int M = (int)points.size(), k = bps;
    float tmp[2 * 8] = { 0 };
    for (int i = 0; i < M; i++) {
        float dist = std::norm(sample - points[i]);
        float d = expf(-dist / npwr);
        for (int j = 0; j < k; j++) tmp[2 * j + ((i >> j) & 1)] += d;
    }
    for (int i = 0; i < k; i++) s[k - 1 - i] = logf(tmp[2 * i + 1]) - logf(tmp[2 * i + 0]);
}

double constellation::min_dist2() const
{
    double best = 1e300;
    for (size_t i = 0; i < points.size(); i++) for (size_t j = i + 1; j < points.size(); j++) {
        double d = std::norm(std::complex<double>(points[i]) - std::complex<double>(points[j]));
        if (d < best) best = d;
    }
    return best;
}

} // namespace orc
