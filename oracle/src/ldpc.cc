// oracle/src/ldpc.cc -- TEST INFRASTRUCTURE ONLY (see oracle.hpp header).
// gr-fec alist / GF2Mat::get_G / cldpc / awgn_bp restated from GNU Radio 3.10 (RECALLED,
// SURVEY.md Appendix A.7-A.9; un-vendored dependency "Gnuradio 3.10", CMakeLists.txt:38)
// behind gr-dtl's lib/dtl/ldpc_dec.cc:34-71 and lib/dtl/ldpc_enc.cc:32-68.
//
// sigma: gr-dtl never calls awgn_bp::set_sigma (ldpc_dec.cc:55-57), so the reference's
// channel-LR scaling is formally undefined.  This repo DEFINES sigma^2 = 2, i.e. the
// demapper output is taken as a true LLR: lr = exp(-2*y/sigma^2) = exp(llr).  
#include "oracle.hpp"
#include <cmath>
#include <fstream>
#include <sstream>

namespace orc {

bool ldpc_code::load(const std::string& path)
{
    std::ifstream f(path);
    if (!f) return false;
    auto ints = [](const std::string& line) {
        std::vector<int> v; std::istringstream is(line); int x;
        while (is >> x) v.push_back(x);
        return v;
    };
    std::string line;
    auto next = [&]() { while (std::getline(f, line)) { auto v = ints(line); if (!v.empty()) return v; } return std::vector<int>(); };
    auto h = next(); if (h.size() < 2) return false;
    N = h[0]; M = h[1];
    next();                       // max col / row weights
    auto cw = next(), rw = next();
    if ((int)cw.size() != N || (int)rw.size() != M) return false;
    nlist.assign(N, {}); mlist.assign(M, {});
    for (int i = 0; i < N; i++) { auto v = next(); for (int x : v) if (x > 0) nlist[i].push_back(x - 1); if ((int)nlist[i].size() != cw[i]) return false; }
    for (int i = 0; i < M; i++) { auto v = next(); for (int x : v) if (x > 0) mlist[i].push_back(x - 1); if ((int)mlist[i].size() != rw[i]) return false; }

    // GF2Mat::get_G (RECALLED): for each diagonal position scan row `diag` for a 1 in
    // columns >= diag, swap columns (recorded in permute), eliminate the column from all
    // other rows; an all-zero row is swapped to the bottom and the row limit shrinks.
    std::vector<std::vector<uint8_t>> G(M, std::vector<uint8_t>(N, 0));
    for (int r = 0; r < M; r++) for (int c : mlist[r]) G[r][c] = 1;
    permute.resize(N);
    for (int c = 0; c < N; c++) permute[c] = c;
    int diag = 0, limit = M;
    while (diag < limit) {
        bool non_zero = false;
        for (int col = diag; col < N; col++) {
            if (G[diag][col]) {
                non_zero = true;
                std::swap(permute[diag], permute[col]);
                if (col != diag) for (int r = 0; r < M; r++) std::swap(G[r][diag], G[r][col]);
                break;
            }
        }
        if (non_zero) {
            for (int r = 0; r < limit; r++)
                if (r != diag && G[r][diag]) for (int c = 0; c < N; c++) G[r][c] ^= G[diag][c];
            diag++;
        } else {
            std::swap(G[diag], G[limit - 1]);
            limit--;
        }
    }
    vedge.assign(N, {});                       // (check, position in mlist[check]) in nlist order
    for (int v = 0; v < N; v++)
        for (int c : nlist[v]) {
            int pos = -1;
            for (size_t i = 0; i < mlist[c].size(); i++) if (mlist[c][i] == v) pos = (int)i;
            if (pos < 0) return false;
            vedge[v].push_back({ c, pos });
        }
    rank = diag;
    K = N - rank;
    P.assign(rank, std::vector<uint8_t>(K));
    for (int i = 0; i < rank; i++) for (int j = 0; j < K; j++) P[i][j] = G[i][rank + j];
    return true;
}

// cldpc::encode + ldpc_enc::encode un-permute: wire order x = [parity(rank) | data(K)]
void ldpc_code::encode(const uint8_t* data, uint8_t* out) const
{
    for (int i = 0; i < rank; i++) {
        uint8_t s = 0;
        for (int j = 0; j < K; j++) s ^= (P[i][j] & data[j]);
        out[i] = s & 1;
    }
    for (int j = 0; j < K; j++) out[rank + j] = data[j] & 1;
}

// awgn_bp::decode (RECALLED): flooding sum-product in the likelihood-ratio domain, double.
void ldpc_code::bp_decode(const std::vector<float>& rx, int* nit, std::vector<uint8_t>& est) const
{
    est.assign(N, 0);
    auto is_codeword = [&]() {
        for (int c = 0; c < M; c++) { uint8_t s = 0; for (int v : mlist[c]) s ^= est[v]; if (s) return false; }
        return true;
    };
    *nit = 0;
    for (int i = 0; i < N; i++) est[i] = rx[i] < 0 ? 1 : 0;          // compute_init_estimate
    if (is_codeword()) return;
    const double s2 = sigma2;
    std::vector<double> rx_lr(N), lr(N);
    for (int i = 0; i < N; i++) rx_lr[i] = std::exp((double)(-1 * 2 * rx[i]) / s2);   // rx_lr_calc
    // Q[c][v], R[c][v] stored per edge in mlist order; var side indexes them through vedge
    std::vector<std::vector<double>> Q(M), R(M);
    for (int c = 0; c < M; c++) { Q[c].assign(mlist[c].size(), 0); R[c].assign(mlist[c].size(), 0); }
    for (int v = 0; v < N; v++) for (auto& e : vedge[v]) Q[e.first][e.second] = rx_lr[v];   // spa_initialize
    while (*nit < max_it) {
        *nit += 1;
        for (int c = 0; c < M; c++) {                                    // update_chks
            double product = 1.0;
            for (size_t i = 0; i < mlist[c].size(); i++) product = product * (2 / (1 + Q[c][i]) - 1);
            for (size_t i = 0; i < mlist[c].size(); i++) {
                double p = product / (2 / (1 + Q[c][i]) - 1);
                R[c][i] = (1 - p) / (1 + p);
            }
        }
        for (int v = 0; v < N; v++) {                                    // update_vars
            double s = rx_lr[v];
            for (auto& e : vedge[v]) s = s * R[e.first][e.second];
            lr[v] = s;
            for (auto& e : vedge[v]) Q[e.first][e.second] = s / R[e.first][e.second];
        }
        for (int v = 0; v < N; v++) est[v] = lr[v] > 1 ? 1 : 0;          // decision
        if (is_codeword()) break;
    }
}

// ldpc_dec::decode (ldpc_dec.cc:62-71)
int ldpc_code::decode(const float* in, int* nit, uint8_t* out) const
{
    std::vector<float> cw(N);
    for (int i = 0; i < N; i++) cw[permute[i]] = -1 * in[i];
    std::vector<uint8_t> est;
    bp_decode(cw, nit, est);
    for (int i = 0; i < K; i++) out[i] = est[permute[rank + i]];        // get_systematic_bits
    return K;
}

} // namespace orc
