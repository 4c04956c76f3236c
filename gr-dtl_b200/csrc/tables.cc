// tables.cc -- host-side construction of the tables the kernels read: constellation points and the
// LDPC code (alist parse + systematic-form column permutation).
//
// Constellations: lib/dtl/ofdm_adaptive_utils.cc:41-70 (type -> gr-digital constellation, bits/symbol),
// lib/dtl/constellation.cc:18-24 (QPSK scaled by 0.5); point tables of gr-digital
// constellation_{bpsk,qpsk,8psk,16qam} (GNU Radio 3.10).  QAM64/QAM256 (ids 5, 6) are this repo's
// extension: Gray-coded square QAM, unit mean energy, index = gray(im level) << bps/2 | gray(re level).
// LDPC: gr-fec alist::read + cldpc::set_alist (GF2Mat::get_G) behind lib/dtl/ldpc_dec.cc:34-59.
#include "ctx.hpp"
#include <algorithm>
#include <cmath>
#include <cstring>
#include <fstream>
#include <sstream>

namespace dtl {

int cnst_bps(int c)
{
    static const int b[DTL_CNST_MAX + 1] = { 0, 1, 2, 3, 4, 6, 8 };
    return (c >= 0 && c <= DTL_CNST_MAX) ? b[c] : 0;
}

void cnst_points(int c, std::vector<float2>& pts, std::vector<float2>& inv, float* level, float* inv2l)
{
    pts.clear(); inv.clear(); *level = 0.f; *inv2l = 0.f;
    const int bps = cnst_bps(c);
    switch (c) {
    case DTL_CNST_BPSK:
        pts = { make_float2(-1.f, 0.f), make_float2(1.f, 0.f) };
        break;
    case DTL_CNST_QPSK: {
        const float a = (float)0.707107;            // gr-digital SQRT_TWO
        const float sgn[4][2] = { { -1, -1 }, { 1, -1 }, { -1, 1 }, { 1, 1 } };
        for (auto& s : sgn) pts.push_back(make_float2((float)((double)(s[0] * a) * 0.5), (float)((double)(s[1] * a) * 0.5)));
        break; }
    case DTL_CNST_PSK8: {
        const float step = (float)(M_PI / 8.0);
        const int mult[8] = { 1, 7, 15, 9, 3, 5, 13, 11 };
        for (int m : mult) pts.push_back(make_float2(cosf(m * step), sinf(m * step)));
        break; }
    case DTL_CNST_QAM16: {
        const float l = sqrtf(0.1f);
        *level = l; *inv2l = 1.0f / (2.0f * l);
        for (int i = 0; i < 16; i++) {
            float re = ((i & 2) ? 3.0f : 1.0f) * l, im = ((i & 1) ? 3.0f : 1.0f) * l;
            pts.push_back(make_float2((i & 4) ? -re : re, (i & 8) ? im : -im));
        }
        break; }
    case DTL_CNST_QAM64: case DTL_CNST_QAM256: {
        const int hb = bps / 2, m = 1 << hb;
        const float l = (float)(1.0 / std::sqrt(c == DTL_CNST_QAM64 ? 42.0 : 170.0));
        *level = l; *inv2l = 1.0f / (2.0f * l);
        pts.assign((size_t)1 << bps, make_float2(0, 0));
        for (int ai = 0; ai < m; ai++)
            for (int ar = 0; ar < m; ar++)
                pts[((ai ^ (ai >> 1)) << hb) | (ar ^ (ar >> 1))] =
                    make_float2((float)(2 * ar - (m - 1)) * l, (float)(2 * ai - (m - 1)) * l);
        break; }
    default: return;
    }
    for (auto& p : pts) {
        const double re = p.x, im = p.y, d = re * re + im * im;
        inv.push_back(make_float2((float)(re / d), (float)(-im / d)));
    }
}

// Table of the decoder's exp() (bp_ref.cuh: exp_tab): for i in [0, 128) the pair (tail_i, bits_i) with 2^(i/128) = H_i (1 + tail_i),
// H_i the nearest double, bits_i = the bit pattern of H_i minus (i << 45) stored in a double's 8 bytes -- the layout of the
// table-driven exp of glibc >= 2.28 (sysdeps/ieee754/dbl-64/e_exp_data.c, N = 128).  Computed here in 113-bit arithmetic:
// 2^(1/128) by seven square roots (Newton), powers by multiplication.
void exp_table(std::vector<double>& tab)
{
    typedef __float128 q;
    q step = 2;
    for (int k = 0; k < 7; k++) {
        q y = (q)std::sqrt((double)step);
        for (int it = 0; it < 6; it++) y = (y + step / y) / 2;
        step = y;
    }
    tab.assign(256, 0.0);
    q t = 1;
    for (int i = 0; i < 128; i++) {
        const double hi = (double)t, tail = (double)((t - (q)hi) / (q)hi);   // relative: 2^(i/128) = H (1 + tail)
        uint64_t bits;
        memcpy(&bits, &hi, 8);
        bits -= (uint64_t)i << 45;
        tab[2 * i] = tail;
        memcpy(&tab[2 * i + 1], &bits, 8);
        t *= step;
    }
}

static std::vector<int> parse_ints(const std::string& line)
{
    std::vector<int> v; std::istringstream is(line); int x;
    while (is >> x) v.push_back(x);
    return v;
}

bool load_alist(const std::string& path, Code& c, std::string& err)
{
    std::ifstream f(path);
    if (!f) { err = "cannot open alist file " + path; return false; }
    std::vector<std::vector<int>> rows;
    std::string line;
    while (std::getline(f, line)) { auto v = parse_ints(line); if (!v.empty()) rows.push_back(v); }
    if (rows.size() < 4 || rows[0].size() < 2) { err = "malformed alist header in " + path; return false; }
    c.n = rows[0][0]; c.m = rows[0][1];
    if (c.n <= 0 || c.m <= 0 || c.n > 65535 || (int)rows.size() < 4 + c.n + c.m ||
        (int)rows[2].size() != c.n || (int)rows[3].size() != c.m) { err = "malformed alist " + path; return false; }
    c.nlist.assign(c.n, {}); c.mlist.assign(c.m, {});
    for (int v = 0; v < c.n; v++) {
        for (int x : rows[4 + v]) if (x > 0) c.nlist[v].push_back(x - 1);
        if ((int)c.nlist[v].size() != rows[2][v]) { err = "alist column weight mismatch in " + path; return false; }
    }
    for (int r = 0; r < c.m; r++) {
        for (int x : rows[4 + c.n + r]) if (x > 0) c.mlist[r].push_back(x - 1);
        if ((int)c.mlist[r].size() != rows[3][r]) { err = "alist row weight mismatch in " + path; return false; }
    }
    c.edges = 0; c.max_row = 0; c.max_col = 0;
    for (auto& r : c.mlist) { c.edges += (int)r.size(); c.max_row = std::max(c.max_row, (int)r.size()); }
    for (auto& v : c.nlist) c.max_col = std::max(c.max_col, (int)v.size());
    if (c.edges > 65535) { err = "alist too large"; return false; }

    // GF2Mat::get_G: bring H to [I | P] by row reduction with column swaps; the swaps are cldpc's
    // `permute`.  Row `d` is scanned left to right from column d for its pivot; an all-zero row is
    // exchanged with the last live row.
    const int n = c.n, m = c.m, W = (n + 63) / 64;
    std::vector<std::vector<uint64_t>> H(m, std::vector<uint64_t>(W, 0));
    for (int r = 0; r < m; r++) for (int v : c.mlist[r]) H[r][v >> 6] |= 1ull << (v & 63);
    auto get = [&](int r, int col) { return (int)((H[r][col >> 6] >> (col & 63)) & 1); };
    auto swap_cols = [&](int a, int b) {
        for (int r = 0; r < m; r++) {
            int x = get(r, a), y = get(r, b);
            if (x != y) { H[r][a >> 6] ^= 1ull << (a & 63); H[r][b >> 6] ^= 1ull << (b & 63); }
        }
    };
    c.permute.resize(n);
    for (int i = 0; i < n; i++) c.permute[i] = i;
    int d = 0, live = m;
    while (d < live) {
        int piv = -1;
        for (int col = d; col < n; col++) if (get(d, col)) { piv = col; break; }
        if (piv < 0) { std::swap(H[d], H[live - 1]); live--; continue; }
        if (piv != d) { swap_cols(d, piv); std::swap(c.permute[d], c.permute[piv]); }
        for (int r = 0; r < live; r++)
            if (r != d && get(r, d)) for (int w = 0; w < W; w++) H[r][w] ^= H[d][w];
        d++;
    }
    c.rank = d; c.k = n - d;
    c.P.assign(c.rank, std::vector<uint8_t>(c.k));
    for (int i = 0; i < c.rank; i++) for (int j = 0; j < c.k; j++) c.P[i][j] = (uint8_t)get(i, c.rank + j);
    c.path = path;
    return true;
}

} // namespace dtl
