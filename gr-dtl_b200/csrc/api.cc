// api.cc -- host side of libdtl_b200.so: the C ABI of include/dtl_b200.h.
//
// A handle owns one CUDA stream, the device-resident tables the kernels read (carrier plan, pilot tables,
// constellation points, twiddles, LDPC graphs), the per-stream receiver state that survives between calls
// (feedback decision, frame counters, transport-block carry) and grow-on-demand device scratch.  Every entry
// point only enqueues copies and kernels on that stream; DTL_MEM_HOST calls additionally wait for the outputs.
// There is no CPU implementation behind this file: without a CUDA device dtl_b200_create fails.
#include "kernels.cuh"
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <random>
#include <mutex>

using namespace dtl;

namespace {

std::mutex g_err_mu;
std::string g_create_err;

unsigned long long g_alloc_gen = 0;      // bumped whenever a device buffer moves: captured graphs hold raw pointers (RxGraph)

struct DevBuf {
    void* p = nullptr; size_t cap = 0;
    cudaError_t ensure(size_t n)
    {
        if (n <= cap) return cudaSuccess;
        g_alloc_gen++;
        if (p) { cudaError_t e = cudaFree(p); p = nullptr; cap = 0; if (e != cudaSuccess) return e; }
        size_t want = n + n / 8 + 256;
        cudaError_t e = cudaMalloc(&p, want);
        if (e != cudaSuccess) { p = nullptr; return e; }
        cap = want;
        return cudaSuccess;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};

enum Slot {
    SL_IQ, SL_FREQ_IN, SL_TAPS, SL_META, SL_HARD, SL_SOFT, SL_TAPS_OUT, SL_RES, SL_LLR, SL_NLLR, SL_FBYTES, SL_FNBYTES,
    SL_FCRC, SL_FREQ, SL_BYTES, SL_NBYTES, SL_TBS_OUT, SL_NTBS, SL_NOISE, SL_SYMS, SL_CHUNKS, SL_BITS_OUT, SL_NIT,
    SL_HDR, SL_TXIQ, SL_FFT_IN, SL_FFT_OUT, SL_ENC_IN, SL_ENC_OUT,
    // FEC scratch
    SL_CWBUF, SL_BPITEMS, SL_GLIST, SL_SGN, SL_SYX, SL_SYFLAGS, SL_SYSEGS, SL_SYNSEG, SL_SYRAW, SL_SYFIXED, SL_SYSTARTS, SL_SYSTARTS_OUT, SL_SYNFR, SL_SYFRAMES, SL_SYVALID, SL_TBDESC, SL_TRES, SL_TBCOUNT, SL_TBLIST, SL_CWTB, SL_CWNIT, SL_CWGEO, SL_BPLIST, SL_BITS, SL_MAG, SL_HFREQ, SL_HMETA, SL_HTAPS, SL_HRES,
    SL_TXDATA, SL_TXLEN, SL_TXCNST, SL_TXFEC, SL_TXTBS, SL_TXNTB, SL_TXKEEP, SL_TXSEGS, SL_TXCRC, SL_TXBITS, SL_TXCHUNKS, SL_TXMETA, SL_TXCONS, SL_TXSTARTED, SL_COUNT
};

} // namespace

// The one-frame call (S = 1, F = 1, DTL_MEM_HOST: what a block shim does per work()) as a CUDA graph: ~15 dependent launches
// and the stream forks / joins around them become one cudaGraphLaunch.  Captured from the ordinary code path after two warm-up
// calls (every scratch buffer has its size by then); replays stage through page-locked buffers of the handle.
struct RxGraph {
    cudaGraphExec_t exec = nullptr;
    int in_kind = -1, bytes_cap = 0, tb_cap = 0, mask = 0, warm = 0;
    unsigned long long gen = 0;
    bool failed = false;
    uint64_t launches = 0;
    unsigned char* pin = nullptr; size_t pin_bytes = 0;
    size_t o_in = 0, o_taps = 0, o_meta = 0, o_res = 0, o_bytes = 0, o_nb = 0, o_tbs = 0, o_nt = 0, o_fcrc = 0;
    void drop() { if (exec) cudaGraphExecDestroy(exec); exec = nullptr; warm = 0; }
};

struct dtl_b200 {
    int device = 0, sm_count = 148;
    cudaStream_t stream = nullptr;
    std::string err;
    uint64_t launches = 0;
    RxGraph rxg;
    Prof prof;
    // configuration (deep copies)
    int N = 0, cp = 0, L = 0, H = 1, cap = 0, max_bps = 0, fec_enabled = 0;
    int llr_stride = 0, bytes_stride = 0, frame_samples = 0;
    int ldpc_max_it = 15, ldpc_mode = 0; double sigma2 = 2.0;
    int n_mcs = 0, initial_mcs = 0, decision_th = 5; double hysteresis = 1.0;
    bool have_sync = false;
    std::vector<Code> codes;          // [n_codes + 1], entry 0 unused
    // device tables
    std::vector<void*> owned;
    CnstDev cn{}; PlanDev pl{};
    const float2* tw = nullptr; const uint32_t* crc_table = nullptr;
    const CodeDev* d_codes = nullptr;
    const double* d_mcs_snr = nullptr; const int* d_mcs_cnst = nullptr; const int* d_mcs_fec = nullptr;
    const float2* d_sync1 = nullptr; const float2* d_sync2 = nullptr;
    const int* d_tx_base = nullptr; const uint8_t* d_tx_pil_mask = nullptr; const float2* d_tx_pil_val = nullptr;
    int hdr_stride = 0;
    HdrDev hd{}; bool have_hdr = false;   // header branch tables (need the sync words)
    int lay_tab_bytes = 0, lay_warp_bytes = 0; bool all_lay = true;         // layered decoder (bp_layered.cuh)
    int fast_tab_bytes = 0, fast_warp_bytes = 0; bool all_fast = true;   // reference-mode decoder, row-major layout (bp_ref.cuh)
    const double2* d_exp_tab = nullptr;
    int warp_bytes = 0, max_tb = 0, max_ncws = 1, max_k = 0, max_n = 0, tab_bytes = 0, screen_bytes = 16, screen_tab_bytes = 0, screen_words = 0;
    // per-stream state
    int n_streams = 0;
    DevBuf st_buf, carry_buf, pend_buf, csgn_buf;   // csgn: sign words of the carry buffer (screening pass)
    int sgn_stride = 0, csgn_stride = 0;
    DevBuf txst_buf, txcarry_buf; int tx_streams = -1;   // TX FEC framer state (allocated by its first call)
    int carry_cap = 0;
    int* d_err = nullptr; int* d_cnt = nullptr; unsigned long long* d_bits_used = nullptr; int* d_cursor = nullptr;
    int* h_err = nullptr;             // pinned mirror of d_err
    DevBuf slot[SL_COUNT];
    // pipelined DTL_MEM_HOST receive path: chunks of streams, H2D of chunk c+1 under the kernels of chunk c
    cudaStream_t copy_stream = nullptr, out_stream = nullptr;
    cudaEvent_t ev_h2d[2] = { nullptr, nullptr }, ev_done[2] = { nullptr, nullptr };
    DevBuf pipe_in[2][3];
    int s_off = 0;                    // first stream of the chunk the device-level stages work on
    // side stream: the small latency-bound kernels that nothing heavy waits for run beside the heavy ones --
    // TB planning (needs the tags only) under FFT + equaliser, SNR tags + per-stream feedback state under the LDPC decoder.
    // Off while stage profiling is on (the stage events sit on the main stream) and with DTL_NO_OVERLAP=1.
    cudaStream_t side_stream = nullptr;
    cudaEvent_t ev_fork = nullptr, ev_plan = nullptr, ev_eq = nullptr, ev_side = nullptr, ev_hdr = nullptr;
    bool side_pending = false, hdr_pending = false;
    // synchronisation front end (dtl_b200_rx_stream): per-stream state, double-buffered work buffers
    DevBuf sync_st, sync_w[2]; int sync_streams = -1, sync_cur = 0, sync_wcap = 0, sync_keep = 0, sync_tmax = 0; float sync_threshold = 0.95f;
    const int* hdr_valid = nullptr;   // set for one dev_header call: frame slots that hold a frame
    bool no_fast_bp = false;          // DTL_BP_OLD=1: the first (slot-major) reference-mode decoder, for A/B runs
    const float2* fuse_iq = nullptr;  // set by rx_device for one dev_equalize call: FFT fused into the equaliser kernel   // hdr_pending: the header branch of this call runs on the side stream
    int overlap = -1;

    template <class T> T* upload(const std::vector<T>& v, cudaError_t& e)
    {
        if (e != cudaSuccess) return nullptr;
        void* p = nullptr;
        e = cudaMalloc(&p, std::max<size_t>(v.size() * sizeof(T), 16));
        if (e != cudaSuccess) return nullptr;
        owned.push_back(p);
        if (!v.empty()) e = cudaMemcpy(p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice);
        return (T*)p;
    }
};

namespace {

dtl_status fail(dtl_b200* h, dtl_status s, const std::string& msg)
{
    if (h) h->err = msg;
    else { std::lock_guard<std::mutex> g(g_err_mu); g_create_err = msg; }
    return s;
}
dtl_status cuda_fail(dtl_b200* h, cudaError_t e, const char* where)
{
    return fail(h, DTL_ERR_CUDA, std::string(where) + ": " + cudaGetErrorString(e));
}
#define CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return cuda_fail(h, e_, #call); } while (0)

int bin_of(int carrier, int N, bool shifted)
{
    if (carrier < 0) carrier += N;
    return (carrier + (shifted ? N / 2 : 0)) % N;
}

// gr_expj(float(phase)) evaluated in double and rounded to float (arithmetic contract)
float2 expj_contract(double phase)
{
    const float p32 = (float)phase;
    return make_float2((float)std::cos((double)p32), (float)std::sin((double)p32));
}

int tb_ncws_host(int cw_len, int frame_len) { return frame_len > cw_len ? 1 + frame_len / cw_len : 1; }

// sigma^2 as lr_of_rx (kernels_fec.cu) takes it: a power of two in a safe exponent range is passed as -(1/sigma^2)
double sigma2_for_lr(double s2)
{
    int ex = 0;
    const double m = std::frexp(s2, &ex);
    return (m == 0.5 && ex > -500 && ex < 500) ? -(1.0 / s2) : s2;
}

// Steps of the layered decoder (bp_layered.cuh): groups of check rows that share no variable, at most 32 rows each, one lane
// per row.  A step costs the same whether 5 or 32 of its lanes work, so the colouring is BALANCED: the smallest number of groups
// K >= ceil(m / 32) for which a randomised most-constrained-first colouring with capacity 32 succeeds (5 groups for the 148
// rows of the n = 300 code; first-fit in row order gave 13).  Fixed seed: the schedule is the same in every process.
std::vector<std::vector<int>> schedule_rows(const std::vector<std::vector<int>>& mlist, int n, int* shared_out)
{
    const int m = (int)mlist.size();
    if (shared_out) *shared_out = 0;
    std::vector<std::vector<int>> rows_of(n);
    for (int r = 0; r < m; r++) for (int v : mlist[r]) rows_of[v].push_back(r);
    std::vector<std::vector<int>> adj(m);
    for (int r = 0; r < m; r++) {
        for (int v : mlist[r]) for (int r2 : rows_of[v]) if (r2 != r) adj[r].push_back(r2);
        std::sort(adj[r].begin(), adj[r].end());
        adj[r].erase(std::unique(adj[r].begin(), adj[r].end()), adj[r].end());
    }
    std::mt19937 rng(0x1A7E12u);
    for (int K = std::max(1, (m + 31) / 32); K <= m; K++) {
        for (int trial = 0; trial < 600; trial++) {
            std::vector<int> colour(m, -1), size(K, 0);
            std::vector<std::vector<char>> banned(m, std::vector<char>(K, 0));
            bool ok = true;
            for (int done = 0; done < m && ok; done++) {
                // the uncoloured row with the fewest feasible groups (ties: random)
                int best = -1, best_feas = K + 1, ties = 0;
                for (int r = 0; r < m; r++) {
                    if (colour[r] >= 0) continue;
                    int feas = 0;
                    for (int k = 0; k < K; k++) feas += !banned[r][k] && size[k] < 32;
                    if (feas < best_feas) { best = r; best_feas = feas; ties = 1; }
                    else if (feas == best_feas && (int)(rng() % (unsigned)++ties) == 0) best = r;
                }
                if (best_feas == 0) { ok = false; break; }
                // into the feasible group with the fewest rows (keeps the groups level), ties random
                int g = -1, gs = 33, gt = 0;
                for (int k = 0; k < K; k++) {
                    if (banned[best][k] || size[k] >= 32) continue;
                    if (size[k] < gs) { g = k; gs = size[k]; gt = 1; }
                    else if (size[k] == gs && (int)(rng() % (unsigned)++gt) == 0) g = k;
                }
                colour[best] = g; size[g]++;
                for (int r2 : adj[best]) banned[r2][g] = 1;
            }
            if (!ok) continue;
            std::vector<std::vector<int>> groups(K);
            for (int r = 0; r < m; r++) groups[colour[r]].push_back(r);
            groups.erase(std::remove_if(groups.begin(), groups.end(), [](const std::vector<int>& g) { return g.empty(); }), groups.end());
            return groups;
        }
    }
    std::vector<std::vector<int>> single;                       // unreachable (K = m always succeeds)
    for (int r = 0; r < m; r++) single.push_back({ r });
    return single;
}

// Placement of the layered decoder's posteriors (bp_layered.cuh).  Every variable gets a slot (bank = slot % 32) such that, per
// step, a bank holds at most `rmax` of the step's variables wherever the graph allows (greedy start + hill climbing on the
// overflow count); each step's rows then order their edges by a bipartite edge colouring rows x banks with rmax colours, so the
// j-th access of the step's lanes goes to distinct banks (Koenig: always possible without overflow).  Returns the slot of every
// variable, the slot count ns, per row the variable at each edge position (-1 = pad) and the wavefronts beyond one per access
// that are left (0 = conflict-free).
struct LayPlacement { std::vector<int> slot; int ns = 0; std::vector<std::vector<int>> row_edges; int extra = 0; };
LayPlacement lay_assign_slots(const std::vector<std::vector<int>>& groups, const std::vector<std::vector<int>>& mlist, int n, int rmax)
{
    const int NB = 32, S = (int)groups.size();
    LayPlacement out;
    out.ns = ((n + NB - 1) / NB + 1) * NB;
    const int cap = out.ns / NB;
    std::vector<std::vector<int>> steps_of(n);
    for (int s = 0; s < S; s++) for (int r : groups[s]) for (int v : mlist[r]) steps_of[v].push_back(s);
    std::mt19937 rng(0xB4A2C5u);
    std::vector<int> bank(n, -1), fill(NB, 0);
    std::vector<std::vector<int>> cnt(S, std::vector<int>(NB, 0));
    std::vector<int> order(n);
    for (int v = 0; v < n; v++) order[v] = v;
    std::shuffle(order.begin(), order.end(), rng);
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return steps_of[a].size() > steps_of[b].size(); });
    for (int v : order) {
        int best = -1, bc = 1 << 30, ties = 0;
        for (int b = 0; b < NB; b++) {
            if (fill[b] >= cap) continue;
            int c = 0;
            for (int s : steps_of[v]) c += cnt[s][b] * 16 + (cnt[s][b] >= rmax ? 1000 : 0);
            c += fill[b];
            if (c < bc) { best = b; bc = c; ties = 1; }
            else if (c == bc && (int)(rng() % (unsigned)++ties) == 0) best = b;
        }
        bank[v] = best; fill[best]++;
        for (int s : steps_of[v]) cnt[s][best]++;
    }
    auto overflow_of = [&](int v, int b) { int o = 0; for (int s : steps_of[v]) o += cnt[s][b] > rmax; return o; };   // v counted in cnt
    for (int it = 0; it < 200000; it++) {
        // a variable that sits in an overfull (step, bank) moves to the bank where it overflows least (or swaps with one there)
        int total = 0;
        for (int s = 0; s < S; s++) for (int b = 0; b < NB; b++) total += std::max(0, cnt[s][b] - rmax);
        if (!total) break;
        const int v = (int)(rng() % (unsigned)n);
        if (!overflow_of(v, bank[v])) continue;
        const int b0 = bank[v], b1 = (int)(rng() % (unsigned)NB);
        if (b1 == b0) continue;
        auto move = [&](int x, int from, int to) { for (int s : steps_of[x]) { cnt[s][from]--; cnt[s][to]++; } bank[x] = to; fill[from]--; fill[to]++; };
        int u = -1;
        if (fill[b1] >= cap) {                                  // full: swap with a random variable of b1
            std::vector<int> cand;
            for (int x = 0; x < n; x++) if (bank[x] == b1) cand.push_back(x);
            u = cand[rng() % cand.size()];
        }
        move(v, b0, b1);
        if (u >= 0) move(u, b1, b0);
        int after = 0;
        for (int s = 0; s < S; s++) for (int b = 0; b < NB; b++) after += std::max(0, cnt[s][b] - rmax);
        if (after > total) { if (u >= 0) move(u, b0, b1); move(v, b1, b0); }
    }
    out.slot.assign(n, 0);
    std::vector<int> next(NB, 0);
    for (int v = 0; v < n; v++) out.slot[v] = bank[v] + NB * next[bank[v]]++;
    // edge order per step: bipartite edge colouring (rows of the step x banks) with rmax colours, alternating-path recolouring
    out.row_edges.assign(mlist.size(), std::vector<int>(rmax, -1));
    for (int s = 0; s < S; s++) {
        const int R = (int)groups[s].size();
        std::vector<int> eu, eb, ev, ecol;
        for (int l = 0; l < R; l++) for (int v : mlist[groups[s][l]]) { eu.push_back(l); eb.push_back(bank[v]); ev.push_back(v); ecol.push_back(-1); }
        std::vector<std::vector<int>> colU(R, std::vector<int>(rmax, -1)), colB(NB, std::vector<int>(rmax, -1));
        for (int e = 0; e < (int)eu.size(); e++) {
            int cu = -1, cb = -1;
            for (int c = 0; c < rmax && cu < 0; c++) if (colU[eu[e]][c] < 0) cu = c;
            for (int c = 0; c < rmax && cb < 0; c++) if (colB[eb[e]][c] < 0) cb = c;
            if (cu < 0) continue;                               // row heavier than rmax: cannot happen (rmax = max row weight)
            if (cb >= 0 && colB[eb[e]][cu] >= 0 && cb != cu) {  // free cu at the bank: flip the cu/cb path that starts there
                std::vector<int> path;
                int node = eb[e], col = cu;
                bool at_bank = true;
                while (path.size() < eu.size()) {
                    const int e2 = at_bank ? colB[node][col] : colU[node][col];
                    if (e2 < 0) break;
                    path.push_back(e2);
                    node = at_bank ? eu[e2] : eb[e2];
                    at_bank = !at_bank;
                    col = col == cu ? cb : cu;
                }
                for (int e2 : path) { colU[eu[e2]][ecol[e2]] = -1; if (colB[eb[e2]][ecol[e2]] == e2) colB[eb[e2]][ecol[e2]] = -1; }
                for (int e2 : path) { ecol[e2] = ecol[e2] == cu ? cb : cu; colU[eu[e2]][ecol[e2]] = e2; colB[eb[e2]][ecol[e2]] = e2; }
            }
            ecol[e] = cu; colU[eu[e]][cu] = e; colB[eb[e]][cu] = e;
        }
        bool valid = true;
        for (int l = 0; l < R && valid; l++) {
            std::vector<int> seen(rmax, 0);
            int have = 0;
            for (int e = 0; e < (int)eu.size(); e++) if (eu[e] == l) { if (ecol[e] < 0 || seen[ecol[e]]++) valid = false; have++; }
            if (have != (int)mlist[groups[s][l]].size()) valid = false;
        }
        for (int l = 0; l < R; l++) {
            auto& re = out.row_edges[groups[s][l]];
            if (valid) { for (int e = 0; e < (int)eu.size(); e++) if (eu[e] == l) re[ecol[e]] = ev[e]; }
            else for (size_t j = 0; j < mlist[groups[s][l]].size(); j++) re[j] = mlist[groups[s][l]][j];
        }
        for (int j = 0; j < rmax; j++) {
            int c[32] = { 0 }, mx = 1;
            for (int l = 0; l < R; l++) { const int v = out.row_edges[groups[s][l]][j]; if (v >= 0) mx = std::max(mx, ++c[bank[v]]); }
            out.extra += mx - 1;
        }
    }
    return out;
}

// Assignment of the variables to (round, lane) positions of the decoder's column pass (bp_ref.cuh: var_pass).  The pass is
// a flooding update -- every variable is independent -- so any assignment gives the same messages; the one chosen keeps the
// 8-byte shared-memory accesses of a half warp (16 lanes, 16 bank pairs) on distinct banks wherever the graph allows:
// lane's j-th access goes to slot[v][j], bank pair = slot % 16.  Greedy fill of the half-warp groups with the variable that
// collides least with what the group already holds, a few hundred seeded restarts, best total kept.  Returns the order
// (position -> variable, -1 = idle) and the wavefronts beyond one per access it leaves (0 = conflict-free).
std::vector<int> order_variables(const std::vector<std::vector<int>>& slots, int n, int nvr, int cmax, int dummy, int* extra_out)
{
    const int groups = nvr * 2;
    auto cost_of = [&](const std::vector<int>& ord) {
        int extra = 0;
        for (int g = 0; g < groups; g++)
            for (int j = 0; j < cmax; j++) {
                int cnt[16] = { 0 }, mx = 1;
                for (int l = 0; l < 16; l++) {
                    const int v = ord[g * 16 + l];
                    if (v < 0 || j >= (int)slots[v].size()) continue;       // dummy slot: one address, a broadcast
                    mx = std::max(mx, ++cnt[slots[v][j] & 15]);
                }
                extra += mx - 1;
            }
        return extra;
    };
    std::vector<int> best(groups * 16, -1);
    for (int v = 0; v < n; v++) best[v] = v;
    int best_cost = cost_of(best);
    uint64_t rng = 0x9e3779b97f4a7c15ull;
    auto next = [&]() { rng ^= rng << 13; rng ^= rng >> 7; rng ^= rng << 17; return rng; };
    for (int attempt = 0; attempt < 400 && best_cost > 0; attempt++) {
        std::vector<int> pool(n), ord(groups * 16, -1);
        for (int v = 0; v < n; v++) pool[v] = v;
        for (int i = n - 1; i > 0; i--) std::swap(pool[i], pool[next() % (uint64_t)(i + 1)]);
        for (int g = 0; g < groups && !pool.empty(); g++) {
            int cnt[4][16] = { { 0 } };
            for (int l = 0; l < 16 && !pool.empty(); l++) {
                int pick = 0, pick_c = 1 << 30;
                for (size_t q = 0; q < pool.size() && pick_c > 0; q++) {
                    const int v = pool[q];
                    int c = 0;
                    for (int j = 0; j < (int)slots[v].size() && j < 4; j++) c += cnt[j][slots[v][j] & 15];
                    if (c < pick_c) { pick_c = c; pick = (int)q; }
                }
                const int v = pool[pick];
                pool[pick] = pool.back(); pool.pop_back();
                ord[g * 16 + l] = v;
                for (int j = 0; j < (int)slots[v].size() && j < 4; j++) cnt[j][slots[v][j] & 15]++;
            }
        }
        const int c = cost_of(ord);
        if (c < best_cost) { best_cost = c; best = ord; }
    }
    (void)dummy;
    if (extra_out) *extra_out = best_cost;
    return best;
}

} // namespace

// ---- lifecycle ---------------------------------------------------------------------------------------------------
extern "C" int dtl_b200_abi_version(void) { return DTL_B200_ABI_VERSION; }

extern "C" const char* dtl_b200_last_error(const dtl_b200* h)
{
    if (h) return h->err.c_str();
    std::lock_guard<std::mutex> g(g_err_mu);
    static thread_local std::string copy;
    copy = g_create_err;
    return copy.c_str();
}

extern "C" void dtl_b200_destroy(dtl_b200* h)
{
    if (!h) return;
    if (cudaSetDevice(h->device) != cudaSuccess) { cudaGetLastError(); delete h; return; }   // runtime unloading: nothing to free
    if (h->stream) cudaStreamSynchronize(h->stream);
    h->rxg.drop();
    if (h->rxg.pin) cudaFreeHost(h->rxg.pin);
    for (void* p : h->owned) cudaFree(p);
    for (auto& b : h->slot) b.release();
    h->st_buf.release(); h->carry_buf.release(); h->pend_buf.release(); h->csgn_buf.release();
    h->sync_st.release(); h->sync_w[0].release(); h->sync_w[1].release(); h->txst_buf.release(); h->txcarry_buf.release();
    h->prof.release();
    for (int i = 0; i < 2; i++) {
        if (h->ev_h2d[i]) cudaEventDestroy(h->ev_h2d[i]);
        if (h->ev_done[i]) cudaEventDestroy(h->ev_done[i]);
        for (auto& b : h->pipe_in[i]) b.release();
    }
    if (h->copy_stream) cudaStreamDestroy(h->copy_stream);
    if (h->out_stream) cudaStreamDestroy(h->out_stream);
    if (h->side_stream) { cudaStreamSynchronize(h->side_stream); cudaStreamDestroy(h->side_stream); }
    for (cudaEvent_t e : { h->ev_fork, h->ev_plan, h->ev_eq, h->ev_side, h->ev_hdr }) if (e) cudaEventDestroy(e);
    if (h->h_err) cudaFreeHost(h->h_err);
    if (h->stream) cudaStreamDestroy(h->stream);
    delete h;
}

static dtl_status build_tables(dtl_b200* h, const dtl_config* c)
{
    const int N = h->N, L = h->L, H = h->H;
    const bool shifted = c->input_is_shifted != 0;
    cudaError_t e = cudaSuccess;
    // ---- carrier sets
    std::vector<std::vector<int>> occ, pc; std::vector<std::vector<float2>> ps;
    for (int s = 0, p = 0; s < c->n_occupied_sets; s++) {
        occ.emplace_back(c->occupied_carriers + p, c->occupied_carriers + p + c->occupied_set_len[s]);
        p += c->occupied_set_len[s];
    }
    for (int s = 0, p = 0; s < c->n_pilot_carrier_sets; s++) {
        pc.emplace_back(c->pilot_carriers + p, c->pilot_carriers + p + c->pilot_carrier_set_len[s]);
        p += c->pilot_carrier_set_len[s];
    }
    for (int s = 0, p = 0; s < c->n_pilot_symbol_sets; s++) {
        std::vector<float2> v;
        for (int k = 0; k < c->pilot_symbol_set_len[s]; k++, p++) v.push_back(make_float2(c->pilot_symbols[2 * p], c->pilot_symbols[2 * p + 1]));
        ps.push_back(v);
    }
    if (pc.empty()) ps.clear();
    for (auto& set : occ) for (int x : set) if (x < -N / 2 - N || x >= N) return fail(h, DTL_ERR_INVALID_ARG, "occupied carrier out of range");
    const int nocc = (int)occ.size(), npc = (int)pc.size(), nps = (int)ps.size();
    PlanDev& pl = h->pl;
    pl.N = N; pl.L = L; pl.npc = npc; pl.nps = nps; pl.nocc = nocc;
    pl.symbols_skipped = c->symbols_skipped;
    pl.pcs0 = npc ? c->symbols_skipped % npc : 0;
    pl.alpha = c->alpha; pl.oma = 1.0f - c->alpha;
    // frame capacity (python/dtl/ofdm_adaptive.py:2-6)
    h->cap = 0;
    for (int i = 0; i < L; i++) h->cap += (int)occ[i % nocc].size();
    pl.cap = h->cap;
    // occupied mask, serialiser positions (list order), per-symbol bases for RX (ofdm_serializer_vcc, symbols_skipped = 1)
    std::vector<uint8_t> occupied(N, 0);
    std::vector<int16_t> ser_pos((size_t)nocc * N, -1);
    for (int s = 0; s < nocc; s++)
        for (size_t k = 0; k < occ[s].size(); k++) {
            const int b = bin_of(occ[s][k], N, shifted);
            occupied[b] = 1; ser_pos[(size_t)s * N + b] = (int16_t)k;
        }
    std::vector<int> ser_base(L + 1, 0);
    for (int i = 0; i < L; i++) ser_base[i + 1] = ser_base[i] + (int)occ[(1 + i) % nocc].size();
    // pilots of the payload equaliser (ofdm_equalizer_1d_pilots ctor + ofdm_adaptive_equalizer.cc:196-213)
    std::vector<uint8_t> pmask((size_t)std::max(npc, 1) * N, 0);
    std::vector<uint16_t> prank((size_t)std::max(npc, 1) * N, 0);
    for (int s = 0; s < npc; s++) {
        for (int x : pc[s]) pmask[(size_t)s * N + bin_of(x, N, shifted)] = 1;
        int r = 0;
        for (int b = 0; b < N; b++) if (pmask[(size_t)s * N + b]) prank[(size_t)s * N + b] = (uint16_t)r++;
    }
    std::vector<int> pprefix(L + 1, 0);
    for (int i = 0; i < L; i++) {
        int cnt = 0;
        if (npc) { const int s = (pl.pcs0 + i) % npc; for (int b = 0; b < N; b++) cnt += pmask[(size_t)s * N + b]; }
        pprefix[i + 1] = pprefix[i] + cnt;
    }
    pl.npil_frame = pprefix[L];
    std::vector<float2> psym((size_t)std::max(nps, 1) * N, make_float2(0, 0)), pinv((size_t)std::max(nps, 1) * N, make_float2(0, 0));
    for (int i = 0; i < nps; i++) {
        const int set = i % npc;
        for (size_t k = 0; k < pc[set].size() && k < ps[i].size(); k++) {
            const int b = bin_of(pc[set][k], N, shifted);
            const float2 p = ps[i][k];
            psym[(size_t)i * N + b] = p;
            const double re = p.x, im = p.y, d = re * re + im * im;
            pinv[(size_t)i * N + b] = make_float2((float)(re / d), (float)(-im / d));
        }
    }
    // carrier-offset phase table (frame_equalizer_vcvc_impl.cc:168-177,189-193)
    pl.max_off = N / 2;                                  // every offset ofdm_chanest_vcvc can report (|g| <= N/2)
    std::vector<float2> ptab((size_t)(2 * pl.max_off + 1) * (L + 1));
    for (int off = -pl.max_off; off <= pl.max_off; off++) {
        float2* row = &ptab[(size_t)(off + pl.max_off) * (L + 1)];
        for (int i = 0; i < L; i++) row[i] = expj_contract(-(2.0 * M_PI) * off * h->cp / N * (i + 1));
        row[L] = expj_contract((2.0 * M_PI) * off * h->cp / N * L);
    }
    pl.occupied = h->upload(occupied, e); pl.pilot_mask = h->upload(pmask, e); pl.pilot_rank = h->upload(prank, e);
    pl.pilot_prefix = h->upload(pprefix, e); pl.pilot_sym = h->upload(psym, e); pl.pilot_inv = h->upload(pinv, e);
    pl.ser_pos = h->upload(ser_pos, e); pl.ser_base = h->upload(ser_base, e); pl.phase_tab = h->upload(ptab, e);
    // per (symbol, bin) role and the inverse of the serialiser (eq_demap_kernel phases 1 and 2)
    if ((size_t)L * N > 65535 || pprefix[L] > 65535) return fail(h, DTL_ERR_UNSUPPORTED, "frame_len * fft_len > 65535");
    std::vector<uint32_t> sym_kind((size_t)L * N, 0u);
    std::vector<uint16_t> ser_src((size_t)std::max(h->cap, 1), 0);
    for (int i = 0; i < L; i++) {
        const int pcs = npc ? (pl.pcs0 + i) % npc : 0;
        for (int b = 0; b < N; b++) {
            uint32_t v = 0;
            if (npc && pmask[(size_t)pcs * N + b]) v = 2u | ((uint32_t)(pprefix[i] + prank[(size_t)pcs * N + b]) << 16);
            else if (occupied[b]) v = 1u;
            sym_kind[(size_t)i * N + b] = v;
        }
        const int oset = (1 + i) % nocc;
        for (size_t k = 0; k < occ[oset].size(); k++) ser_src[ser_base[i] + k] = (uint16_t)(i * N + bin_of(occ[oset][k], N, shifted));
    }
    pl.sym_kind = h->upload(sym_kind, e); pl.ser_src = h->upload(ser_src, e);
    std::vector<uint32_t> h_sym_kind = sym_kind; std::vector<uint16_t> h_ser_src = ser_src;
    std::vector<double> rcp_tab((size_t)std::max(pprefix[L], 1));
    for (size_t i = 0; i < rcp_tab.size(); i++) rcp_tab[i] = 1.0 / (double)(i + 1);
    pl.rcp_tab = h->upload(rcp_tab, e);

    // ---- TX allocator tables (digital.ofdm_carrier_allocator_cvc): data symbol s = 0 is the first header symbol
    std::vector<int> tx_base(H + L + 1, 0);
    std::vector<uint8_t> tx_mask((size_t)(H + L) * N, 0);
    std::vector<float2> tx_val((size_t)(H + L) * N, make_float2(0, 0));
    for (int s = 0; s < H + L; s++) {
        tx_base[s + 1] = tx_base[s] + (int)occ[s % nocc].size();
        if (npc && nps) {
            const auto& cset = pc[s % npc]; const auto& sset = ps[s % nps];
            for (size_t k = 0; k < cset.size() && k < sset.size(); k++) {
                const int b = bin_of(cset[k], N, true);
                tx_mask[(size_t)s * N + b] = 1; tx_val[(size_t)s * N + b] = sset[k];
            }
        }
    }
    h->hdr_stride = tx_base[H];
    h->d_tx_base = h->upload(tx_base, e); h->d_tx_pil_mask = h->upload(tx_mask, e); h->d_tx_pil_val = h->upload(tx_val, e);
    if (c->sync_word1 && c->sync_word2) {
        std::vector<float2> s1(N), s2(N);
        for (int i = 0; i < N; i++) { s1[i] = make_float2(c->sync_word1[2 * i], c->sync_word1[2 * i + 1]); s2[i] = make_float2(c->sync_word2[2 * i], c->sync_word2[2 * i + 1]); }
        h->d_sync1 = h->upload(s1, e); h->d_sync2 = h->upload(s2, e); h->have_sync = true;
        // ---- header branch (python/dtl/ofdm_receiver.py:100-150): ofdm_chanest_vcvc(sync_word1, sync_word2, header_len),
        // ofdm_adaptive_header_equalizer (symbols_skipped 0), ofdm_serializer_vcc (symbols_skipped 0), header parser
        HdrDev& hd = h->hd;
        hd.N = N; hd.H = H; hd.has_fec = h->fec_enabled;
        hd.nbits = H * (int)occ[0].size();                  // ofdm_receiver.py:129-130: [occupied_carriers[0]] * header_len
        int first = 0, last = N - 1;
        for (int i = 0; i < N; i++) if (s2[i].x != 0.f || s2[i].y != 0.f) { first = i; break; }
        for (int i = N - 1; i >= 0; i--) if (s2[i].x != 0.f || s2[i].y != 0.f) { last = i; break; }
        hd.max_neg = -first; hd.max_pos = N - last - 1;
        if (hd.max_neg % 2) hd.max_neg++;                    // carrier offsets must be even
        if (hd.max_pos % 2) hd.max_pos--;
        std::vector<float2> corr(N, make_float2(0, 0));
        for (int i = 0; i < N; i++)
            if (s1[i].x != 0.f || s1[i].y != 0.f) {          // sync2 / sync1 in this library's fp32 division contract
                const float den = std::fmaf(s1[i].x, s1[i].x, s1[i].y * s1[i].y), r = 1.0f / den;
                const float nr = std::fmaf(s2[i].x, s1[i].x, s2[i].y * s1[i].y), ni = std::fmaf(s2[i].y, s1[i].x, -(s2[i].x * s1[i].y));
                corr[i] = make_float2(nr * r, ni * r);
            }
        std::vector<uint8_t> hkind((size_t)H * N, 0);
        std::vector<float2> hpinv((size_t)H * N, make_float2(0, 0));
        std::vector<uint16_t> hser;
        for (int i = 0; i < H; i++) {
            const int pcs = npc ? i % npc : 0, pss = nps ? i % nps : 0;
            for (int b = 0; b < N; b++) {
                if (npc && pmask[(size_t)pcs * N + b]) { hkind[(size_t)i * N + b] = 2; if (nps) hpinv[(size_t)i * N + b] = pinv[(size_t)pss * N + b]; }
                else if (occupied[b]) hkind[(size_t)i * N + b] = 1;
            }
            const auto& oset = occ[i % nocc];
            for (size_t k = 0; k < oset.size(); k++) hser.push_back((uint16_t)(i * N + bin_of(oset[k], N, shifted)));
        }
        hser.resize((size_t)hd.nbits, 0);
        std::vector<float2> hrot((size_t)2 * pl.max_off + 1);
        for (int off = -pl.max_off; off <= pl.max_off; off++) hrot[(size_t)(off + pl.max_off)] = expj_contract((2.0 * M_PI) * off * h->cp / N * H);
        hd.ref = h->d_sync2; hd.corr_v = h->upload(corr, e); hd.kind = h->upload(hkind, e); hd.pinv = h->upload(hpinv, e);
        // per-symbol phase of the HEADER symbols (frame_equalizer: -2 pi off cp / N (i + 1), i < H): its own table, the payload
        // table's row has L + 1 entries and entry L is the channel-state rotation, not a symbol phase (H may exceed L)
        std::vector<float2> hph((size_t)(2 * pl.max_off + 1) * H);
        for (int off = -pl.max_off; off <= pl.max_off; off++)
            for (int i = 0; i < H; i++) hph[(size_t)(off + pl.max_off) * H + i] = expj_contract(-(2.0 * M_PI) * off * h->cp / N * (i + 1));
        hd.ser = h->upload(hser, e); hd.rot = h->upload(hrot, e); hd.hphase = h->upload(hph, e);
        // CRC-16 of the parser (ofdm_adaptive_packet_header.cc:99-111 pack_crc with its `i*len + j < header_len` bound, then
        // crc(16, 0x1021, 0xFFFF, 0, false, true)) as an affine map of the header bits: crc0 ^ XOR of crc_tab[j] over set bits
        {
            const int nb = hd.nbits, len = nb > 16 ? (nb - 16 + 7) / 8 : 0;
            auto crc_of = [&](int set_bit) {
                unsigned crc = 0xFFFFu;
                for (int i = 0; i < len; i++) {
                    unsigned byte = 0;
                    for (int j = 0; j < 8 && i * len + j < nb; j++) byte = ((byte << 1) | (unsigned)(i * 8 + j == set_bit)) & 0xffu;
                    crc ^= byte << 8;
                    for (int b = 0; b < 8; b++) crc = ((crc & 0x8000u) ? ((crc << 1) ^ 0x1021u) : (crc << 1)) & 0xFFFFu;
                }
                unsigned calc = 0;
                for (int i = 0; i < 16; i++) if ((crc >> i) & 1u) calc |= 1u << (15 - i);
                return calc;
            };
            hd.crc_bits = std::min(len * 8, nb);
            hd.crc0 = crc_of(-1);
            std::vector<uint16_t> tab((size_t)std::max(hd.crc_bits, 1), 0);
            for (int j = 0; j < hd.crc_bits; j++) tab[(size_t)j] = (uint16_t)(crc_of(j) ^ hd.crc0);
            hd.crc_tab = h->upload(tab, e);
        }
        h->have_hdr = (size_t)H * N <= 65535;
    }

    // ---- constellations
    std::vector<float2> pts_all, inv_all;
    h->cn.off[0] = 0; h->cn.bps[0] = 0; h->cn.level[0] = 0; h->cn.inv2l[0] = 0;
    for (int k = 1; k <= DTL_CNST_MAX; k++) {
        std::vector<float2> pts, inv;
        cnst_points(k, pts, inv, &h->cn.level[k], &h->cn.inv2l[k]);
        h->cn.off[k] = (int)pts_all.size(); h->cn.bps[k] = cnst_bps(k);
        pts_all.insert(pts_all.end(), pts.begin(), pts.end()); inv_all.insert(inv_all.end(), inv.begin(), inv.end());
    }
    h->cn.off[0] = h->cn.off[1];
    h->cn.off[DTL_CNST_MAX + 1] = (int)pts_all.size();
    h->cn.points = h->upload(pts_all, e); h->cn.inv_points = h->upload(inv_all, e);
    {   // table blob of eq_demap_kernel (PlanDev::blob)
        auto al = [](size_t x) { return (x + 15) & ~(size_t)15; };
        const size_t npil = (size_t)pl.npil_frame;
        const size_t o_inv = 0, o_ser = al(o_inv + inv_all.size() * 8), o_kind = al(o_ser + (size_t)std::max(h->cap, 1) * 2);
        const size_t o_pinv = al(o_kind + (size_t)L * N * 2), o_psym = al(o_pinv + npil * 8), total = al(o_psym + npil * 8);
        std::vector<unsigned char> blob(total, 0);
        memcpy(&blob[o_inv], inv_all.data(), inv_all.size() * 8);
        memcpy(&blob[o_ser], h_ser_src.data(), h_ser_src.size() * 2);
        uint16_t* k16 = (uint16_t*)&blob[o_kind];
        float2* bpinv = (float2*)&blob[o_pinv]; float2* bpsym = (float2*)&blob[o_psym];
        for (int i = 0; i < L; i++) {
            const int pset = nps ? (c->symbols_skipped + i) % nps : 0;
            for (int b = 0; b < N; b++) {
                const uint32_t g = h_sym_kind[(size_t)i * N + b];
                k16[(size_t)i * N + b] = (uint16_t)((g & 3u) | ((g >> 16) << 2));
                if ((g & 3u) == 2u && nps) { bpinv[g >> 16] = pinv[(size_t)pset * N + b]; bpsym[g >> 16] = psym[(size_t)pset * N + b]; }
            }
        }
        if (pl.npil_frame >= 16384) return fail(h, DTL_ERR_UNSUPPORTED, "more than 16383 pilots per frame");
        pl.blob = h->upload(blob, e); pl.blob_bytes = (int)total;
        pl.blob_inv = (int)o_inv; pl.blob_ser = (int)o_ser; pl.blob_kind = (int)o_kind; pl.blob_pinv = (int)o_pinv; pl.blob_psym = (int)o_psym;
        // tables of eq_fast_kernel: only when a carrier keeps its role for the whole frame
        bool uniform = npc >= 1 && nocc >= 1 && nps >= 1 && npil > 0 && npil % L == 0;
        for (int s = 1; s < nocc && uniform; s++) uniform = occ[s] == occ[0];
        for (int s = 1; s < npc && uniform; s++)
            for (int b = 0; b < N; b++)
                if (pmask[(size_t)s * N + b] != pmask[b] || prank[(size_t)s * N + b] != prank[b]) uniform = false;
        for (int i = 1; i < L && uniform; i++)
            for (int b = 0; b < N; b++)
                if ((h_sym_kind[(size_t)i * N + b] & 3u) != (h_sym_kind[b] & 3u)) uniform = false;
        pl.fast_ok = uniform ? 1 : 0;
        pl.pil_per_sym = pl.fast_ok ? (int)npil / L : 0;
        std::vector<uint8_t> ckind(N, 0);
        for (int b = 0; b < N; b++) ckind[b] = (uint8_t)(h_sym_kind[b] & 3u);
        pl.carrier_kind = h->upload(ckind, e);
        const size_t n_small = (size_t)h->cn.off[DTL_CNST_QAM16 + 1];
        const size_t f_inv = 0, f_ser = al(f_inv + n_small * 8), f_pinv = al(f_ser + (size_t)std::max(h->cap, 1) * 2);
        const size_t f_ppos = al(f_pinv + npil * 8), f_rcp = al(f_ppos + npil * 2), f_total = f_rcp;
        std::vector<unsigned char> fb(f_total, 0);
        memcpy(&fb[f_inv], inv_all.data(), n_small * 8);
        memcpy(&fb[f_ser], h_ser_src.data(), h_ser_src.size() * 2);
        memcpy(&fb[f_pinv], bpinv, npil * 8);
        uint16_t* ppos = (uint16_t*)&fb[f_ppos];
        for (int i = 0; i < L; i++)
            for (int b = 0; b < N; b++) {
                const uint32_t g = h_sym_kind[(size_t)i * N + b];
                if ((g & 3u) == 2u) ppos[g >> 16] = (uint16_t)(i * N + b);
            }
        pl.fblob = h->upload(fb, e); pl.fblob_bytes = (int)f_total;
        pl.fb_inv = (int)f_inv; pl.fb_ser = (int)f_ser; pl.fb_pinv = (int)f_pinv; pl.fb_ppos = (int)f_ppos; pl.fb_rcp = (int)f_rcp;
    }

    // ---- twiddles exp(-2 pi i j / N), CRC-32 table (poly 0x04C11DB7 reflected, crc_util.cc:23)
    std::vector<float2> tw(N);
    for (int j = 0; j < N; j++) { const double a = -2.0 * M_PI * j / N; tw[j] = make_float2((float)std::cos(a), (float)std::sin(a)); }
    h->tw = h->upload(tw, e);
    // [0]: the byte table; [1..4]: the CRC register after 16 zero bytes, per byte of the register (eq_fast_kernel MODE 2:
    // crc32_warp strings 16-byte chunk CRCs together)
    std::vector<uint32_t> crc(5 * 256);
    for (uint32_t i = 0; i < 256; i++) { uint32_t r = i; for (int k = 0; k < 8; k++) r = (r & 1) ? (r >> 1) ^ 0xEDB88320u : r >> 1; crc[i] = r; }
    for (int j = 0; j < 4; j++)
        for (uint32_t b = 0; b < 256; b++) {
            uint32_t r = b << (8 * j);
            for (int z = 0; z < 16; z++) r = crc[r & 0xffu] ^ (r >> 8);
            crc[256 * (1 + j) + b] = r;
        }
    h->crc_table = h->upload(crc, e);

    // ---- MCS look-up table (ofdm_adaptive_feedback_decision.cc:33-53: lut[0].first = DBL_MIN)
    std::vector<double> snr(c->mcs_snr, c->mcs_snr + c->n_mcs);
    if (!snr.empty()) snr[0] = 2.2250738585072014e-308;
    h->d_mcs_snr = h->upload(snr, e);
    h->d_mcs_cnst = h->upload(std::vector<int>(c->mcs_cnst, c->mcs_cnst + c->n_mcs), e);
    h->d_mcs_fec = h->upload(std::vector<int>(c->mcs_fec, c->mcs_fec + c->n_mcs), e);

    // ---- LDPC codes (make_ldpc_decoders, ldpc_dec.cc:23-59)
    h->codes.assign(1, Code());
    std::vector<CodeDev> cd(1);
    memset(&cd[0], 0, sizeof(CodeDev));
    int max_n = 0;
    for (int i = 0; i < c->n_codes; i++) {
        Code code; std::string err;
        if (!load_alist(c->alist_paths[i], code, err)) return fail(h, DTL_ERR_IO, err);
        // slot-major edge storage: edge j of row r (mlist order) is slot j*m + r (kernels_fec.cu)
        const int slots = code.max_row * code.m;
        if (slots > 65535) return fail(h, DTL_ERR_UNSUPPORTED, "LDPC graph too large (max_row * m > 65535)");
        // pad slots of row_var hold n: est[n] is a byte that is always 0, so the syndrome check runs every row at max_row edges
        std::vector<uint16_t> row_len(code.m, 0), row_var((size_t)slots, (uint16_t)code.n), col_ptr(code.n + 1, 0), col_edge, perm(code.n);
        for (int r = 0; r < code.m; r++) {
            row_len[r] = (uint16_t)code.mlist[r].size();
            for (size_t j = 0; j < code.mlist[r].size(); j++) row_var[j * code.m + r] = (uint16_t)code.mlist[r][j];
        }
        for (int v = 0; v < code.n; v++) {
            for (int r : code.nlist[v]) {
                int pos = -1;
                for (size_t q = 0; q < code.mlist[r].size(); q++) if (code.mlist[r][q] == v) pos = (int)q;
                if (pos < 0) return fail(h, DTL_ERR_IO, "alist row/column lists disagree in " + code.path);
                col_edge.push_back((uint16_t)(pos * code.m + r));
            }
            col_ptr[v + 1] = (uint16_t)col_edge.size();
        }
        for (int q = 0; q < code.n; q++) perm[q] = (uint16_t)code.permute[q];
        const int kwords = (code.k + 31) / 32;
        std::vector<uint32_t> pmat((size_t)code.rank * kwords, 0u);
        for (int r = 0; r < code.rank; r++)
            for (int j = 0; j < code.k; j++) if (code.P[r][j]) pmat[(size_t)r * kwords + (j >> 5)] |= 1u << (j & 31);
        if (code.max_row > 32) return fail(h, DTL_ERR_UNSUPPORTED, "LDPC row weight > 32");
        CodeDev& d = code.dev;
        d.n = code.n; d.m = code.m; d.k = code.k; d.rank = code.rank; d.edges = code.edges; d.max_row = code.max_row; d.max_col = code.max_col;
        d.row_len = h->upload(row_len, e); d.row_var = h->upload(row_var, e); d.col_ptr = h->upload(col_ptr, e);
        d.col_edge = h->upload(col_edge, e); d.permute = h->upload(perm, e);
        std::vector<uint16_t> inv_perm(code.n, 0), row_wire((size_t)slots, (uint16_t)(32 * ((code.n + 31) / 32)));
        for (int q = 0; q < code.n; q++) inv_perm[perm[q]] = (uint16_t)q;
        for (int r = 0; r < code.m; r++)
            for (size_t j = 0; j < code.mlist[r].size(); j++) row_wire[j * code.m + r] = inv_perm[code.mlist[r][j]];
        d.row_wire = h->upload(row_wire, e);
        h->screen_tab_bytes += (2 * (code.m + slots) + 15) & ~15;
        d.pmat = h->upload(pmat, e); d.kwords = kwords;
        {   // row-major tables of the reference-mode decoder (bp_ref.cuh)
            const int rs = (code.max_row + 1) & ~1, nvr = (code.n + 31) / 32, dummy = code.m * rs;
            d.rs = rs; d.nvr = nvr; d.uniform_rows = 1;
            for (int r = 0; r < code.m; r++) if ((int)code.mlist[r].size() != code.max_row) d.uniform_rows = 0;
            d.fast = bp_ref_has_instance(code.max_row, code.max_col) && dummy + 1 <= 65535 && code.n + 2 <= 65535;
            const bool have_rv2 = code.max_row <= 7 && dummy + 1 <= 65535 && code.n + 2 <= 65535;
            if (have_rv2) {                      // layered decoder: row groups (bp_layered.cuh)
                const auto groups = schedule_rows(code.mlist, code.n, nullptr);
                const int rmx = std::max(code.max_row, 1);
                const LayPlacement pl_ = lay_assign_slots(groups, code.mlist, code.n, rmx);
                code.lay_shared = pl_.extra;                      // wavefronts beyond one per access the placement leaves
                if (getenv("DTL_LAY_DEBUG"))
                    fprintf(stderr, "layered schedule n=%d m=%d: %zu steps, %d slots, %d extra wavefronts per iteration\n", code.n, code.m, groups.size(), pl_.ns, pl_.extra);
                // one 16-byte descriptor per (step, lane): the slots of the row's variables (pad: ns) and the row id (0xffff = idle lane)
                std::vector<uint16_t> lrows(groups.size() * 32 * 8, (uint16_t)pl_.ns);
                for (size_t g = 0; g < groups.size(); g++)
                    for (size_t l = 0; l < 32; l++) {
                        uint16_t* ds = &lrows[(g * 32 + l) * 8];
                        ds[7] = 0xffff;
                        if (l < groups[g].size()) {
                            const int r = groups[g][l];
                            for (int j = 0; j < rmx && j < 7; j++) if (pl_.row_edges[r][j] >= 0) ds[j] = (uint16_t)pl_.slot[pl_.row_edges[r][j]];
                            ds[7] = (uint16_t)r;
                        }
                    }
                std::vector<uint16_t> lperm(code.n);
                for (int q = 0; q < code.n; q++) lperm[q] = (uint16_t)pl_.slot[perm[q]];
                d.lay_steps = (int)groups.size(); d.lay_rows = h->upload(lrows, e);
                d.lay_ns = pl_.ns; d.lay_perm = h->upload(lperm, e);
                h->lay_tab_bytes += (int)lay_tab_bytes(code.n, code.m, d.lay_steps);
                h->lay_warp_bytes = std::max(h->lay_warp_bytes, (int)lay_warp_bytes(d.lay_ns, d.lay_steps));
            } else h->all_lay = false;
            if (have_rv2 && !d.fast) {           // the layered decoder needs the row table even where the reference one has no instance
                std::vector<uint16_t> rv2l((size_t)code.m * 8, (uint16_t)code.n);
                for (int r = 0; r < code.m; r++)
                    for (size_t j = 0; j < code.mlist[r].size(); j++) rv2l[(size_t)r * 8 + j] = (uint16_t)code.mlist[r][j];
                d.rv2 = (const uint4*)h->upload(rv2l, e);
            }
            if (d.fast) {
                std::vector<uint16_t> rv2((size_t)code.m * 8, (uint16_t)code.n);
                std::vector<std::vector<int>> vslots(code.n);
                for (int r = 0; r < code.m; r++)
                    for (size_t j = 0; j < code.mlist[r].size(); j++) rv2[(size_t)r * 8 + j] = (uint16_t)code.mlist[r][j];
                for (int v = 0; v < code.n; v++)
                    for (int r : code.nlist[v]) {
                        int pos = 0;
                        for (size_t q = 0; q < code.mlist[r].size(); q++) if (code.mlist[r][q] == v) pos = (int)q;
                        vslots[v].push_back(r * rs + pos);
                    }
                int extra = 0;
                const std::vector<int> ord = order_variables(vslots, code.n, nvr, code.max_col, dummy, &extra);
                code.bank_extra = extra;
                std::vector<uint32_t> vdesc((size_t)nvr * 32 * 4, 0u);
                for (int i = 0; i < nvr * 32; i++) {
                    const int v = ord[i];
                    uint32_t sl[4] = { (uint32_t)dummy, (uint32_t)dummy, (uint32_t)dummy, (uint32_t)dummy };
                    if (v >= 0) for (size_t j = 0; j < vslots[v].size(); j++) sl[j] = (uint32_t)vslots[v][j];
                    vdesc[(size_t)i * 4] = sl[0] | (sl[1] << 16); vdesc[(size_t)i * 4 + 1] = sl[2] | (sl[3] << 16);
                    vdesc[(size_t)i * 4 + 2] = v >= 0 ? (uint32_t)v : (uint32_t)(code.n + 1);
                }
                d.rv2 = (const uint4*)h->upload(rv2, e); d.vdesc = (const uint4*)h->upload(vdesc, e);
                h->fast_tab_bytes += (int)fast_tab_bytes(code.n, code.m, nvr);
                h->fast_warp_bytes = std::max(h->fast_warp_bytes, (int)fast_warp_bytes(code.n, code.m, rs));
            } else h->all_fast = false;
        }
        cd.push_back(d);
        max_n = std::max(max_n, code.n); h->max_k = std::max(h->max_k, code.k);
        h->warp_bytes = std::max(h->warp_bytes, (int)ldpc_warp_bytes(code.n, slots));
        h->screen_words = std::max(h->screen_words, (code.n + 31) / 32);
        h->screen_bytes = (4 * (32 * (h->screen_words + 1) + 32 * h->screen_words + 1) + 15) & ~15;   // est[32][nw+1] + planes
        h->tab_bytes += (int)ldpc_tab_bytes(code.n, code.m, slots, code.edges);
        if (i + 1 >= DTL_MAX_SMEM_CODES || h->tab_bytes > 96 * 1024)
            return fail(h, DTL_ERR_UNSUPPORTED, "too many / too large LDPC codes for the shared-memory graph tables");
        h->codes.push_back(std::move(code));
    }
    h->d_codes = h->upload(cd, e);
    {
        std::vector<double> et;
        exp_table(et);
        h->d_exp_tab = (const double2*)h->upload(et, e);
    }
    if (max_n) {
        h->max_n = max_n;
        h->max_ncws = tb_ncws_host(max_n, h->cap * h->max_bps);       // fec_decoder_impl.cc:54-72
        h->max_tb = max_n * h->max_ncws;
        // codes with a smaller n split a frame into more codewords
        for (size_t i = 1; i < h->codes.size(); i++) h->max_ncws = std::max(h->max_ncws, tb_ncws_host(h->codes[i].n, 8 * ((h->cap * h->max_bps + 7) / 8)));
        h->carry_cap = 2 * h->max_tb;
    }
    if (e != cudaSuccess) return cuda_fail(h, e, "table upload");
    return DTL_OK;
}

extern "C" dtl_status dtl_b200_create(const dtl_config* c, dtl_b200** out)
{
    if (!out) return DTL_ERR_INVALID_ARG;
    *out = nullptr;
    if (!c) return fail(nullptr, DTL_ERR_INVALID_ARG, "null config");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev <= 0)
        return fail(nullptr, DTL_ERR_NO_DEVICE, std::string("no CUDA device (") + cudaGetErrorString(e) + "); libdtl_b200 has no CPU path");
    if (c->device < 0 || c->device >= ndev) return fail(nullptr, DTL_ERR_INVALID_ARG, "device ordinal out of range");
    const int N = c->fft_len;
    if (N < 8 || N > 1024 || (N & (N - 1))) return fail(nullptr, DTL_ERR_UNSUPPORTED, "fft_len must be a power of two in [8, 1024]");
    if (c->cp_len < 0 || c->cp_len > N) return fail(nullptr, DTL_ERR_INVALID_ARG, "cp_len out of range");
    if (c->frame_len < 1 || c->frame_len > (N > 256 ? 20 : 40)) return fail(nullptr, DTL_ERR_UNSUPPORTED, "frame_len out of range");
    if (c->n_occupied_sets < 1 || !c->occupied_set_len || !c->occupied_carriers) return fail(nullptr, DTL_ERR_INVALID_ARG, "occupied carriers missing");
    if (c->n_mcs < 1 || !c->mcs_snr || !c->mcs_cnst || !c->mcs_fec || c->initial_mcs < 0 || c->initial_mcs >= c->n_mcs)
        return fail(nullptr, DTL_ERR_INVALID_ARG, "MCS look-up table missing (\"Feedback lookup table is empty\")");
    if (c->max_bps < 1 || c->max_bps > 8) return fail(nullptr, DTL_ERR_INVALID_ARG, "max_bps out of range");
    if (c->fec_enabled && c->n_codes < 1) return fail(nullptr, DTL_ERR_INVALID_ARG, "Misconfiguration: decoder not found!");
    dtl_b200* h = new dtl_b200;
    h->device = c->device;
    auto bail = [&](dtl_status s) { { std::lock_guard<std::mutex> g(g_err_mu); g_create_err = h->err; } dtl_b200_destroy(h); return s; };
    if ((e = cudaSetDevice(h->device)) != cudaSuccess) { cuda_fail(h, e, "cudaSetDevice"); return bail(DTL_ERR_CUDA); }
    cudaDeviceProp prop;
    if ((e = cudaGetDeviceProperties(&prop, h->device)) != cudaSuccess) { cuda_fail(h, e, "cudaGetDeviceProperties"); return bail(DTL_ERR_CUDA); }
    if (prop.major != 10) { fail(h, DTL_ERR_UNSUPPORTED, "libdtl_b200 is built for sm_100a only"); return bail(DTL_ERR_UNSUPPORTED); }
    h->sm_count = prop.multiProcessorCount;
    // per-device kernel attributes (opt-in dynamic shared memory), set for THIS handle's device
    for (cudaError_t (*init)() : { init_kernels_fft, init_kernels_eq, init_kernels_fec, init_kernels_hdr, init_kernels_sync })
        if ((e = init()) != cudaSuccess) { cuda_fail(h, e, "cudaFuncSetAttribute"); return bail(DTL_ERR_CUDA); }
    if ((e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking)) != cudaSuccess) { cuda_fail(h, e, "cudaStreamCreate"); return bail(DTL_ERR_CUDA); }
    h->prof.st = h->stream;
    h->N = N; h->cp = c->cp_len; h->L = c->frame_len; h->H = c->header_syms > 0 ? c->header_syms : (c->fec_enabled ? 2 : 1);
    h->max_bps = c->max_bps; h->fec_enabled = c->fec_enabled;
    h->ldpc_max_it = c->ldpc_max_iterations > 0 ? c->ldpc_max_iterations : 15;
    h->ldpc_mode = c->ldpc_mode; h->sigma2 = c->ldpc_sigma2 > 0 ? c->ldpc_sigma2 : 2.0;
    h->sync_threshold = c->sync_threshold > 0.f ? c->sync_threshold : 0.95f;
    h->n_mcs = c->n_mcs; h->initial_mcs = c->initial_mcs; h->decision_th = c->decision_th; h->hysteresis = c->hysteresis;
    dtl_status s = build_tables(h, c);
    if (s != DTL_OK) return bail(s);
    { const char* ev = getenv("DTL_BP_OLD"); h->no_fast_bp = ev && atoi(ev); }
    h->llr_stride = h->cap * h->max_bps;
    h->sgn_stride = (h->llr_stride + 31) / 32 + 1;
    h->bytes_stride = (((h->cap * h->max_bps + 7) / 8) + 3) & ~3;
    h->frame_samples = (h->L + h->H + 2) * (h->N + h->cp);
    void* p = nullptr;
    if ((e = cudaMalloc(&p, 64)) != cudaSuccess) { cuda_fail(h, e, "cudaMalloc"); return bail(DTL_ERR_CUDA); }
    h->owned.push_back(p);
    cudaMemset(p, 0, 64);
    h->d_err = (int*)p; h->d_cnt = (int*)p + 4; h->d_cursor = (int*)p + 8; h->d_bits_used = (unsigned long long*)((char*)p + 48);
    if ((e = cudaHostAlloc((void**)&h->h_err, 64, cudaHostAllocDefault)) != cudaSuccess) { cuda_fail(h, e, "cudaHostAlloc"); return bail(DTL_ERR_CUDA); }
    *h->h_err = 0;
    *out = h;
    return DTL_OK;
}

// ---- state, sync, getters ---------------------------------------------------------------------------------------------
extern "C" dtl_status dtl_b200_reset(dtl_b200* h, int32_t n_streams)
{
    if (!h || n_streams < 0) return DTL_ERR_INVALID_ARG;
    CK(cudaSetDevice(h->device));
    // growing a buffer frees the old one: wait for the kernels that may still read it (no-op when sizes are unchanged,
    // so a reset between batches only enqueues one small kernel on the handle's stream)
    const size_t st_bytes = std::max<size_t>((size_t)n_streams * sizeof(StreamState), 16);
    const size_t carry_bytes = h->carry_cap ? (size_t)n_streams * h->carry_cap * sizeof(float) + 16 : 0;
    const size_t pend_bytes = (size_t)n_streams * sizeof(TbDesc) + 16;
    if (st_bytes > h->st_buf.cap || carry_bytes > h->carry_buf.cap || pend_bytes > h->pend_buf.cap ||
        (carry_bytes && (size_t)n_streams * (h->carry_cap / 32 + 2) * 4 + 16 > h->csgn_buf.cap)) CK(cudaStreamSynchronize(h->stream));
    CK(h->st_buf.ensure(st_bytes));
    if (carry_bytes) CK(h->carry_buf.ensure(carry_bytes));
    if (carry_bytes) { h->csgn_stride = h->carry_cap / 32 + 2; CK(h->csgn_buf.ensure((size_t)n_streams * h->csgn_stride * 4 + 16)); }
    CK(h->pend_buf.ensure(pend_bytes));
    CK(launch_stream_reset(h->stream, (StreamState*)h->st_buf.p, n_streams, h->initial_mcs));
    h->launches += n_streams > 0 ? 1 : 0;
    CK(cudaMemsetAsync(h->d_err, 0, sizeof(int), h->stream));
    h->n_streams = n_streams;
    h->sync_streams = -1;                        // and the synchronisation front end
    h->tx_streams = -1;                          // the TX framer starts over as well (state re-created by its next call)
    return DTL_OK;
}

static dtl_status check_device_error(dtl_b200* h)
{
    const int e = *h->h_err;
    if (!e) return DTL_OK;
    *h->h_err = 0;
    cudaMemsetAsync(h->d_err, 0, sizeof(int), h->stream);
    char buf[160];
    snprintf(buf, sizeof buf, "device error word 0x%x", e);
    if (e & 1) return fail(h, DTL_ERR_MISSING_TAG, std::string(buf) + ": Missing constellation tag / unknown constellation");
    if (e & (4 | 8)) return fail(h, DTL_ERR_MISSING_TAG, std::string(buf) + ": Tags missing / FEC tags missing");
    if (e & 2) return fail(h, DTL_ERR_UNSUPPORTED, std::string(buf) + ": carrier offset beyond the phase table");
    if (e & (16 | 32 | 64 | 512)) return fail(h, DTL_ERR_CAPACITY, std::string(buf) + ": output buffer or transport-block table too small");
    return fail(h, DTL_ERR_INVALID_ARG, std::string(buf) + ": inconsistent transport-block tags");
}

extern "C" dtl_status dtl_b200_sync(dtl_b200* h)
{
    if (!h) return DTL_ERR_INVALID_ARG;
    CK(cudaSetDevice(h->device));
    CK(cudaMemcpyAsync(h->h_err, h->d_err, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return check_device_error(h);
}

extern "C" void* dtl_b200_cuda_stream(dtl_b200* h) { return h ? (void*)h->stream : nullptr; }
extern "C" int32_t dtl_b200_frame_capacity(const dtl_b200* h) { return h ? h->cap : 0; }
extern "C" int32_t dtl_b200_frame_samples(const dtl_b200* h) { return h ? h->frame_samples : 0; }
extern "C" int32_t dtl_b200_llr_stride(const dtl_b200* h) { return h ? h->llr_stride : 0; }
extern "C" int32_t dtl_b200_bytes_stride(const dtl_b200* h) { return h ? h->bytes_stride : 0; }
extern "C" uint64_t dtl_b200_launch_count(const dtl_b200* h) { return h ? h->launches : 0; }

extern "C" dtl_status dtl_b200_code_info(const dtl_b200* h, int32_t fec_id, int32_t* n, int32_t* k, int32_t* permute)
{
    if (!h || fec_id < 1 || fec_id >= (int)h->codes.size()) return DTL_ERR_INVALID_ARG;
    const Code& c = h->codes[fec_id];
    if (n) *n = c.n;
    if (k) *k = c.k;
    if (permute) for (int i = 0; i < c.n; i++) permute[i] = c.permute[i];
    return DTL_OK;
}

extern "C" dtl_status dtl_b200_profile(dtl_b200* h, int enable)
{
    if (!h) return DTL_ERR_INVALID_ARG;
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->stream));
    h->prof.collect();
    h->prof.on = enable != 0;
    return DTL_OK;
}

extern "C" dtl_status dtl_b200_stage_times(dtl_b200* h, double* ms, int64_t* launches, int32_t n, int reset)
{
    if (!h || n < 0 || n > DTL_STAGE_COUNT) return DTL_ERR_INVALID_ARG;
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->stream));
    h->prof.collect();
    for (int i = 0; i < n; i++) { if (ms) ms[i] = h->prof.ms[i]; if (launches) launches[i] = h->prof.n[i]; }
    if (reset) for (int i = 0; i < DTL_STAGE_COUNT; i++) { h->prof.ms[i] = 0; h->prof.n[i] = 0; }
    return DTL_OK;
}

extern "C" dtl_status dtl_b200_host_alloc(void** out, size_t bytes)
{
    if (!out) return DTL_ERR_INVALID_ARG;
    return cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocDefault) == cudaSuccess ? DTL_OK : DTL_ERR_CUDA;
}
extern "C" void dtl_b200_host_free(void* p) { if (p) cudaFreeHost(p); }

// ---- staging: user pointer -> device pointer ----------------------------------------------------------------------------------
namespace {

// Side stream (see the handle): created on first use.
bool side_on(dtl_b200* h)
{
    if (h->overlap < 0) {
        const char* e = getenv("DTL_NO_OVERLAP");
        h->overlap = (e && atoi(e)) ? 0 : 1;
        if (h->overlap) {
            bool ok = cudaStreamCreateWithFlags(&h->side_stream, cudaStreamNonBlocking) == cudaSuccess;
            for (cudaEvent_t* ev : { &h->ev_fork, &h->ev_plan, &h->ev_eq, &h->ev_side, &h->ev_hdr })
                ok = ok && cudaEventCreateWithFlags(ev, cudaEventDisableTiming) == cudaSuccess;
            if (!ok) { cudaGetLastError(); h->overlap = 0; }
        }
    }
    return h->overlap == 1 && !h->prof.on;
}
// The main stream waits for what was put on the side stream (no-op when nothing is pending).
cudaError_t side_join(dtl_b200* h)
{
    if (!h->side_pending) return cudaSuccess;
    h->side_pending = false;
    return cudaStreamWaitEvent(h->stream, h->ev_side, 0);
}

struct Stage {
    dtl_b200* h; int mem; cudaError_t e = cudaSuccess;
    struct Out { void* user; void* dev; size_t bytes; };
    std::vector<Out> outs;
    Stage(dtl_b200* h_, int mem_) : h(h_), mem(mem_) {}
    template <class T> const T* in(int slot, const T* user, size_t count)
    {
        if (!user || mem == DTL_MEM_DEVICE || e != cudaSuccess) return user;
        DevBuf& b = h->slot[slot];
        if ((e = b.ensure(count * sizeof(T) + 16)) != cudaSuccess) return nullptr;
        e = cudaMemcpyAsync(b.p, user, count * sizeof(T), cudaMemcpyHostToDevice, h->stream);
        return (const T*)b.p;
    }
    template <class T> T* out(int slot, T* user, size_t count)
    {
        if (!user || mem == DTL_MEM_DEVICE || e != cudaSuccess) return user;
        DevBuf& b = h->slot[slot];
        if ((e = b.ensure(count * sizeof(T) + 16)) != cudaSuccess) return nullptr;
        outs.push_back({ (void*)user, b.p, count * sizeof(T) });
        return (T*)b.p;
    }
    // device scratch that is needed whether or not the caller wants the data back
    template <class T> T* scratch(int slot, size_t count)
    {
        if (e != cudaSuccess) return nullptr;
        DevBuf& b = h->slot[slot];
        if ((e = b.ensure(count * sizeof(T) + 16)) != cudaSuccess) return nullptr;
        return (T*)b.p;
    }
    dtl_status finish()
    {
        if (e != cudaSuccess) return cuda_fail(h, e, "staging");
        if ((e = side_join(h)) != cudaSuccess) return cuda_fail(h, e, "side stream join");
        if (mem == DTL_MEM_DEVICE) return DTL_OK;
        for (auto& o : outs)
            if ((e = cudaMemcpyAsync(o.user, o.dev, o.bytes, cudaMemcpyDeviceToHost, h->stream)) != cudaSuccess) return cuda_fail(h, e, "cudaMemcpyAsync D2H");
        return dtl_b200_sync(h);
    }
};

dtl_status ensure_streams(dtl_b200* h, int n_streams)
{
    if (h->n_streams == n_streams) return DTL_OK;
    return dtl_b200_reset(h, n_streams);
}

// host-side tag validation for DTL_MEM_HOST calls (the device re-checks and reports through the error word)
dtl_status check_meta_host(dtl_b200* h, const dtl_frame_meta* meta, long long n, int mem)
{
    if (mem != DTL_MEM_HOST || !meta) return DTL_OK;
    for (long long f = 0; f < n; f++) {
        if (meta[f].constellation == DTL_CNST_UNKNOWN) return fail(h, DTL_ERR_MISSING_TAG, "Missing constellation tag.");
        if (meta[f].constellation < 0 || meta[f].constellation > DTL_CNST_MAX) return fail(h, DTL_ERR_INVALID_ARG, "Unknown constellation");
    }
    return DTL_OK;
}

} // namespace

// ---- device-level stages (all pointers are device pointers) ---------------------------------------------------------------------
namespace {

dtl_status dev_demod(dtl_b200* h, const float2* iq, long long n_frames, float2* freq)
{
    h->prof.begin(DTL_STAGE_FFT, 1);
    CK(launch_fft_fwd(h->stream, iq, freq, n_frames * h->L, h->L, h->frame_samples, h->H + 2, h->N + h->cp, h->cp, h->N, h->tw));
    h->prof.end();
    h->launches += 1;
    return DTL_OK;
}

dtl_status dev_equalize(dtl_b200* h, Stage& sg, const float2* freq, const float2* taps, const dtl_frame_meta* meta, int S, int F,
                        float2* out_hard, float2* out_soft, float2* taps_out, dtl_frame_result* res,
                        float* llr, int* n_llr, uint8_t* fbytes, int* fnbytes, int* fcrc, uint32_t* sgn = nullptr)
{
    const long long n = (long long)S * F;
    if (!res) res = sg.scratch<dtl_frame_result>(SL_RES, (size_t)n);
    if (sg.e != cudaSuccess) return cuda_fail(h, sg.e, "scratch");
    EqParams p;
    memset(&p, 0, sizeof p);
    p.pl = h->pl; p.cn = h->cn; p.freq = freq; p.taps = taps; p.meta = meta; p.n_frames = n;
    p.out_hard = out_hard; p.out_soft = out_soft; p.taps_out = taps_out; p.res = res;
    p.llr = llr; p.llr_stride = h->llr_stride; p.n_llr = n_llr;
    p.sgn = sgn; p.sgn_stride = h->sgn_stride;
    p.bytes = fbytes; p.bytes_stride = h->bytes_stride; p.nbytes = fnbytes; p.crc_ok = fcrc;
    p.crc_table = h->crc_table; p.err = h->d_err; p.tw = h->tw;
    if (h->fuse_iq) {
        p.iq = h->fuse_iq; p.iq_frame_stride = h->frame_samples; p.iq_sym0 = h->H + 2; p.iq_sym_stride = h->N + h->cp; p.iq_skip = h->cp;
        h->fuse_iq = nullptr;
    }
    p.mag = sg.scratch<float>(SL_MAG, (size_t)n * (size_t)std::max(h->pl.npil_frame, 1));     // |pilot_eq| for snr_tags_kernel
    if (sg.e != cudaSuccess) return cuda_fail(h, sg.e, "scratch");
    int eq_launches = 0;
    const bool ov = side_on(h);
    bool tags_due = false;
    if (h->hdr_pending) { h->hdr_pending = false; CK(cudaStreamWaitEvent(h->stream, h->ev_hdr, 0)); }   // tags + taps come from the side stream
    h->prof.begin(DTL_STAGE_EQ, 1);
    CK(launch_eq(h->stream, p, &eq_launches, &tags_due));       // the fast kernel leaves the SNR tags to snr_tags_kernel, launched below
    if (sgn && !eq_emits_signs(p)) {             // the general kernel / tiny FFTs: sign words by a pass over the LLRs
        CK(launch_llr_sign(h->stream, llr, h->llr_stride, n, sgn, h->sgn_stride));
        eq_launches++;
    }
    h->prof.end();
    cudaStream_t ss = h->stream;
    if (ov) {                                   // SNR tags and the per-stream feedback state: nothing on the main stream waits for them
        CK(cudaEventRecord(h->ev_eq, h->stream));
        CK(cudaStreamWaitEvent(h->side_stream, h->ev_eq, 0));
        ss = h->side_stream;
    }
    StreamParams sp;
    sp.st = (StreamState*)h->st_buf.p + h->s_off; sp.n_streams = S; sp.F = F; sp.meta = meta; sp.res = res;
    sp.n_mcs = h->n_mcs; sp.mcs_snr = h->d_mcs_snr; sp.mcs_cnst = h->d_mcs_cnst; sp.mcs_fec = h->d_mcs_fec;
    sp.hysteresis = h->hysteresis; sp.decision_th = h->decision_th;
    h->prof.begin(DTL_STAGE_STATE, tags_due ? 2 : 1);           // tags + feedback decision: the per-frame / per-stream bookkeeping stage
    if (tags_due) CK(launch_snr_tags(ss, p));
    CK(launch_stream_state(ss, sp));
    h->prof.end();
    if (ov) { CK(cudaEventRecord(h->ev_side, h->side_stream)); h->side_pending = true; }
    h->launches += eq_launches + 1 + (tags_due ? 1 : 0);
    return DTL_OK;
}

dtl_status fec_prepare(dtl_b200* h, Stage& sg, const float* llr, int llr_stride, const dtl_frame_meta* meta, int S, int F,
                       uint8_t* bytes, int bytes_cap, int* nbytes, dtl_tb_result* tbs, int tb_cap, int* ntbs, FecParams& p,
                       const uint32_t* sgn = nullptr, int sgn_stride = 0)
{
    if (h->codes.size() < 2) return fail(h, DTL_ERR_INVALID_ARG, "Misconfiguration: decoder not found!");
    memset(&p, 0, sizeof p);
    const size_t tb_slots = 2 * (size_t)F + 1, budget = (size_t)S * ((size_t)F + 2);
    p.st = (StreamState*)h->st_buf.p + h->s_off; p.n_streams = S; p.F = F; p.meta = meta; p.llr = llr; p.llr_stride = llr_stride;
    p.carry = (float*)h->carry_buf.p + (size_t)h->s_off * h->carry_cap; p.carry_cap = h->carry_cap;
    p.sgn = sgn; p.sgn_stride = sgn_stride;
    p.csgn = sgn ? (uint32_t*)h->csgn_buf.p + (size_t)h->s_off * h->csgn_stride : nullptr; p.csgn_stride = h->csgn_stride;
    p.n_codes = (int)h->codes.size() - 1; p.codes = h->d_codes; p.cap = h->cap;
    for (int c = 0; c <= DTL_CNST_MAX; c++) p.bps_tab[c] = cnst_bps(c);
    p.max_tb = h->max_tb;
    p.tb_slots = (int)tb_slots;
    p.tbs = sg.scratch<TbDesc>(SL_TBDESC, (size_t)S * tb_slots);
    p.pend = (TbDesc*)h->pend_buf.p + h->s_off;
    p.tres = sg.scratch<dtl_tb_result>(SL_TRES, (size_t)S * tb_slots);
    p.tb_count = sg.scratch<int>(SL_TBCOUNT, (size_t)S);
    p.tb_list = sg.scratch<int>(SL_TBLIST, (size_t)S * tb_slots);
    p.cw_cap = (int)std::min<size_t>(budget * h->max_ncws, 0x7fffffff);
    p.cw_tb = sg.scratch<int>(SL_CWTB, (size_t)p.cw_cap);
    p.cw_nit = sg.scratch<int>(SL_CWNIT, (size_t)p.cw_cap);
    p.cw_geo = sg.scratch<int>(SL_CWGEO, (size_t)p.cw_cap * 3);
    p.bp_list = sg.scratch<int>(SL_BPLIST, 4 * (size_t)p.cw_cap);
    {
        static int th[3] = { -1, 0, 0 };
        if (th[0] < 0) {
            th[0] = 6; th[1] = 4; th[2] = 2;
            if (const char* e = getenv("DTL_BP_BUCKETS")) sscanf(e, "%d,%d,%d", &th[0], &th[1], &th[2]);
        }
        p.bp_th[0] = th[0]; p.bp_th[1] = th[1]; p.bp_th[2] = th[2];
    }
    p.tab_bytes = h->tab_bytes; p.screen_bytes = h->screen_bytes; p.screen_tab_bytes = h->screen_tab_bytes; p.screen_words = h->screen_words;
    p.bits_cap = (long long)budget * ((long long)h->max_ncws * h->max_k + 4);
    p.bits = sg.scratch<uint8_t>(SL_BITS, (size_t)p.bits_cap + 64);
    if (sg.e != cudaSuccess) return cuda_fail(h, sg.e, "FEC scratch");
    p.cnt = h->d_cnt; p.bits_used = h->d_bits_used;
    p.max_it = h->ldpc_max_it; p.sigma2 = h->sigma2; p.mode = h->ldpc_mode; p.warp_bytes = h->warp_bytes;
    p.fast = h->ldpc_mode == DTL_LDPC_REFERENCE && h->all_fast && !h->no_fast_bp;
    p.fast_tab_bytes = h->fast_tab_bytes; p.fast_warp_bytes = h->fast_warp_bytes; p.exp_tab = h->d_exp_tab;
    p.sigma2_lr = sigma2_for_lr(h->sigma2);
    p.lay = h->ldpc_mode == DTL_LDPC_LAYERED && h->all_lay && !h->no_fast_bp;
    p.lay_tab_bytes = h->lay_tab_bytes; p.lay_warp_bytes = h->lay_warp_bytes;
    if (p.fast || p.lay) {
        p.cw_stride = (h->max_n + 3) & ~3;           // 16-byte multiples: one bulk copy per codeword
        p.cwbuf = sg.scratch<float>(SL_CWBUF, (size_t)p.cw_cap * p.cw_stride);
        p.items = sg.scratch<BpItem>(SL_BPITEMS, 4 * (size_t)p.cw_cap);
        p.glist = sg.scratch<int>(SL_GLIST, (size_t)p.cw_cap);
        if (sg.e != cudaSuccess) return cuda_fail(h, sg.e, "FEC scratch");
    }
    p.bytes = bytes; p.bytes_cap = bytes_cap; p.nbytes = nbytes; p.tb_out = tbs; p.tb_cap = tb_cap; p.ntbs = ntbs;
    p.crc_table = h->crc_table; p.err = h->d_err;
    return DTL_OK;
}

dtl_status dev_fec(dtl_b200* h, Stage& sg, const float* llr, int llr_stride, const dtl_frame_meta* meta, int S, int F,
                   uint8_t* bytes, int bytes_cap, int* nbytes, dtl_tb_result* tbs, int tb_cap, int* ntbs)
{
    FecParams p;
    const int sstride = (llr_stride + 31) / 32 + 1;
    uint32_t* sgn = sg.scratch<uint32_t>(SL_SGN, (size_t)S * F * sstride + 4);
    if (sg.e != cudaSuccess) return cuda_fail(h, sg.e, "FEC scratch");
    dtl_status s = fec_prepare(h, sg, llr, llr_stride, meta, S, F, bytes, bytes_cap, nbytes, tbs, tb_cap, ntbs, p, sgn, sstride);
    if (s != DTL_OK) return s;
    int launches = 1;
    h->prof.begin(DTL_STAGE_LDPC, 1);
    CK(launch_llr_sign(h->stream, llr, llr_stride, (long long)S * F, sgn, sstride));
    h->prof.end();
    CK(launch_fec(h->stream, p, h->sm_count, &launches, &h->prof));
    h->launches += launches;
    return DTL_OK;
}

} // namespace

// ---- stage 1 -----------------------------------------------------------------------------------------------------------------------
extern "C" dtl_status dtl_b200_demod(dtl_b200* h, const float* iq, int64_t n_frames, float* freq, int mem)
{
    if (!h || !iq || !freq || n_frames < 0) return DTL_ERR_INVALID_ARG;
    CK(cudaSetDevice(h->device));
    Stage sg(h, mem);
    const float2* d_iq = sg.in(SL_IQ, (const float2*)iq, (size_t)n_frames * h->frame_samples);
    float2* d_freq = sg.out(SL_FREQ, (float2*)freq, (size_t)n_frames * h->L * h->N);
    if (sg.e != cudaSuccess) return cuda_fail(h, sg.e, "staging");
    dtl_status s = dev_demod(h, d_iq, n_frames, d_freq);
    if (s != DTL_OK) return s;
    return sg.finish();
}

extern "C" dtl_status dtl_b200_fft(dtl_b200* h, const float* in, int64_t n_vec, int forward, float* out, int mem)
{
    if (!h || !in || !out || n_vec < 0) return DTL_ERR_INVALID_ARG;
    CK(cudaSetDevice(h->device));
    Stage sg(h, mem);
    const float2* d_in = sg.in(SL_FFT_IN, (const float2*)in, (size_t)n_vec * h->N);
    float2* d_out = sg.out(SL_FFT_OUT, (float2*)out, (size_t)n_vec * h->N);
    if (sg.e != cudaSuccess) return cuda_fail(h, sg.e, "staging");
    h->prof.begin(DTL_STAGE_FFT, 1);
    if (forward) CK(launch_fft_fwd(h->stream, d_in, d_out, n_vec, 1, h->N, 0, h->N, 0, h->N, h->tw));
    else CK(launch_fft_inv(h->stream, d_in, d_out, n_vec, h->N, h->tw));
    h->prof.end();
    h->launches += 1;
    return sg.finish();
}

// ---- stages 2+3 ----------------------------------------------------------------------------------------------------------------------
extern "C" dtl_status dtl_b200_equalize(dtl_b200* h, const float* freq, const float* taps, const dtl_frame_meta* meta,
                                        int32_t n_streams, int32_t frames_per_stream,
                                        float* out_hard, float* out_soft, float* taps_out, dtl_frame_result* res,
                                        float* llr, int32_t* n_llr, uint8_t* bytes, int32_t* nbytes, int32_t* crc_ok, int mem)
{
    if (!h || !freq || !taps || !meta || n_streams < 0 || frames_per_stream < 0) return DTL_ERR_INVALID_ARG;
    CK(cudaSetDevice(h->device));
    const size_t n = (size_t)n_streams * frames_per_stream, LN = (size_t)h->L * h->N;
    dtl_status s = check_meta_host(h, meta, (long long)n, mem);
    if (s != DTL_OK) return s;
    if ((s = ensure_streams(h, n_streams)) != DTL_OK) return s;
    Stage sg(h, mem);
    const float2* d_freq = sg.in(SL_FREQ_IN, (const float2*)freq, n * LN);
    const float2* d_taps = sg.in(SL_TAPS, (const float2*)taps, n * h->N);
    const dtl_frame_meta* d_meta = sg.in(SL_META, meta, n);
    float2* d_hard = sg.out(SL_HARD, (float2*)out_hard, n * LN);
    float2* d_soft = sg.out(SL_SOFT, (float2*)out_soft, n * LN);
    float2* d_tout = sg.out(SL_TAPS_OUT, (float2*)taps_out, n * h->N);
    dtl_frame_result* d_res = sg.out(SL_RES, res, n);
    float* d_llr = sg.out(SL_LLR, llr, n * h->llr_stride);
    int* d_nllr = sg.out(SL_NLLR, n_llr, n);
    uint8_t* d_bytes = sg.out(SL_FBYTES, bytes, n * h->bytes_stride);
    int* d_nb = sg.out(SL_FNBYTES, nbytes, n);
    int* d_crc = sg.out(SL_FCRC, crc_ok, n);
    if (sg.e != cudaSuccess) return cuda_fail(h, sg.e, "staging");
    if (mem == DTL_MEM_HOST && d_soft) CK(cudaMemsetAsync(d_soft, 0, n * LN * sizeof(float2), h->stream));
    if ((nbytes || crc_ok) && !bytes) d_bytes = sg.scratch<uint8_t>(SL_FBYTES, n * h->bytes_stride);
    s = dev_equalize(h, sg, d_freq, d_taps, d_meta, n_streams, frames_per_stream, d_hard, d_soft, d_tout, d_res, d_llr, d_nllr,
                     d_bytes, d_nb, d_crc);
    if (s != DTL_OK) return s;
    return sg.finish();
}

// The lean equaliser pass for callers that must not allocate (include/dtl_b200.h).  Device pointers only; nothing here or below
// calls cudaMalloc / cudaFree: no Stage (its scratch slots grow with cudaMalloc), no stream resize, the |pilot_eq| scratch is the
// static device array of kernels_eq.cu, everything on the handle's main stream.
extern "C" dtl_status dtl_b200_equalize_noalloc(dtl_b200* h, const float* freq, const float* taps, const dtl_frame_meta* meta,
                                                int32_t n_streams, int32_t frames_per_stream,
                                                dtl_frame_result* res, float* llr, int32_t* n_llr)
{
    if (!h || !freq || !taps || !meta || !res || n_streams < 0 || frames_per_stream < 0) return DTL_ERR_INVALID_ARG;
    CK(cudaSetDevice(h->device));
    if (h->n_streams != n_streams)
        return fail(h, DTL_ERR_INVALID_ARG, "dtl_b200_equalize_noalloc: per-stream state is sized for " + std::to_string(h->n_streams) +
                    " streams, the call has " + std::to_string(n_streams) + ": call dtl_b200_reset(n_streams) first (this entry point does not allocate)");
    const long long n = (long long)n_streams * frames_per_stream;
    if (n == 0) return DTL_OK;
    cudaError_t je = side_join(h);                  // whatever an earlier call left on the side stream is ordered in front
    if (je != cudaSuccess) return cuda_fail(h, je, "side stream join");
    if (h->hdr_pending) { h->hdr_pending = false; CK(cudaStreamWaitEvent(h->stream, h->ev_hdr, 0)); }
    EqParams p;
    memset(&p, 0, sizeof p);
    p.pl = h->pl; p.cn = h->cn; p.freq = (const float2*)freq; p.taps = (const float2*)taps; p.meta = meta; p.n_frames = n;
    p.res = res; p.llr = llr; p.llr_stride = h->llr_stride; p.n_llr = n_llr; p.sgn_stride = h->sgn_stride;
    p.bytes_stride = h->bytes_stride; p.crc_table = h->crc_table; p.err = h->d_err; p.tw = h->tw;
    int why = 0;
    long long need = 0, have = 0;
    CK(launch_eq_noalloc(h->stream, p, &why, &need, &have));
    if (why == 1) return fail(h, DTL_ERR_UNSUPPORTED, "dtl_b200_equalize_noalloc: this carrier plan needs the general equaliser kernel (use dtl_b200_equalize)");
    if (why == 2)
        return fail(h, DTL_ERR_CAPACITY, "dtl_b200_equalize_noalloc: " + std::to_string(n) + " frames x " + std::to_string(std::max(h->pl.npil_frame, 1)) +
                    " pilots = " + std::to_string(need) + " floats of scratch, the static array holds " + std::to_string(have) +
                    " (DTL_B200_NOALLOC_MAX_PILOT_FLOATS): split the call or use dtl_b200_equalize");
    if (why == 3) return fail(h, DTL_ERR_CUDA, "dtl_b200_equalize_noalloc: static scratch not set up on this device");
    StreamParams sp;
    sp.st = (StreamState*)h->st_buf.p + h->s_off; sp.n_streams = n_streams; sp.F = frames_per_stream; sp.meta = meta; sp.res = res;
    sp.n_mcs = h->n_mcs; sp.mcs_snr = h->d_mcs_snr; sp.mcs_cnst = h->d_mcs_cnst; sp.mcs_fec = h->d_mcs_fec;
    sp.hysteresis = h->hysteresis; sp.decision_th = h->decision_th;
    CK(launch_stream_state(h->stream, sp));
    h->launches += 3;
    return DTL_OK;
}

// ---- stage 3 / 5a stand-alone blocks -------------------------------------------------------------------------------------------------
static dtl_status demap_common(dtl_b200* h, int mode, const float* syms, const uint8_t* in_chunks, int32_t stride_syms,
                               const dtl_frame_meta* meta, const double* noise, int64_t n_frames, float* llr, int32_t llr_stride,
                               uint8_t* chunks, float* out_syms, int mem)
{
    if (!h || !meta || n_frames < 0 || stride_syms <= 0) return DTL_ERR_INVALID_ARG;
    CK(cudaSetDevice(h->device));
    dtl_status s = check_meta_host(h, meta, n_frames, mem);
    if (s != DTL_OK) return s;
    const size_t n = (size_t)n_frames;
    Stage sg(h, mem);
    DemapParams p;
    memset(&p, 0, sizeof p);
    p.cn = h->cn; p.stride_syms = stride_syms; p.n_frames = n_frames; p.llr_stride = llr_stride; p.mode = mode;
    p.meta = sg.in(SL_META, meta, n);
    p.syms = sg.in(SL_SYMS, (const float2*)syms, n * stride_syms);
    p.in_chunks = sg.in(SL_CHUNKS, in_chunks, n * stride_syms);
    p.noise = sg.in(SL_NOISE, noise, n);
    p.llr = sg.out(SL_LLR, llr, n * (size_t)llr_stride);
    p.chunks = sg.out(SL_BITS_OUT, chunks, n * stride_syms);
    p.out_syms = sg.out(SL_SOFT, (float2*)out_syms, n * stride_syms);
    if (sg.e != cudaSuccess) return cuda_fail(h, sg.e, "staging");
    h->prof.begin(DTL_STAGE_DEMAP, 1);
    CK(launch_demap(h->stream, p));
    h->prof.end();
    h->launches += 1;
    return sg.finish();
}

extern "C" dtl_status dtl_b200_soft_demap(dtl_b200* h, const float* syms, int32_t stride_syms, const dtl_frame_meta* meta,
                                          const double* noise, int64_t n_frames, float* llr, int32_t llr_stride, int mem)
{
    if (!syms || !llr || llr_stride < stride_syms) return DTL_ERR_INVALID_ARG;
    return demap_common(h, 0, syms, nullptr, stride_syms, meta, noise, n_frames, llr, llr_stride, nullptr, nullptr, mem);
}
extern "C" dtl_status dtl_b200_hard_demap(dtl_b200* h, const float* syms, int32_t stride_syms, const dtl_frame_meta* meta,
                                          int64_t n_frames, uint8_t* chunks, int mem)
{
    if (!syms || !chunks) return DTL_ERR_INVALID_ARG;
    return demap_common(h, 1, syms, nullptr, stride_syms, meta, nullptr, n_frames, nullptr, 0, chunks, nullptr, mem);
}
extern "C" dtl_status dtl_b200_map(dtl_b200* h, const uint8_t* chunks, int32_t stride, const dtl_frame_meta* meta,
                                   int64_t n_frames, float* syms, int mem)
{
    if (!chunks || !syms) return DTL_ERR_INVALID_ARG;
    return demap_common(h, 2, nullptr, chunks, stride, meta, nullptr, n_frames, nullptr, 0, nullptr, syms, mem);
}

// ---- stage 4 ---------------------------------------------------------------------------------------------------------------------------
extern "C" dtl_status dtl_b200_fec_decode(dtl_b200* h, const float* llr, int32_t llr_stride, const dtl_frame_meta* meta,
                                          int32_t n_streams, int32_t frames_per_stream,
                                          uint8_t* bytes, int32_t bytes_cap, int32_t* nbytes,
                                          dtl_tb_result* tbs, int32_t tb_cap, int32_t* ntbs, int mem)
{
    if (!h || !llr || !meta || n_streams < 0 || frames_per_stream < 0 || llr_stride <= 0) return DTL_ERR_INVALID_ARG;
    if ((bytes && bytes_cap <= 0) || (tbs && tb_cap <= 0)) return DTL_ERR_INVALID_ARG;
    CK(cudaSetDevice(h->device));
    const size_t n = (size_t)n_streams * frames_per_stream;
    dtl_status s = ensure_streams(h, n_streams);
    if (s != DTL_OK) return s;
    Stage sg(h, mem);
    const float* d_llr = sg.in(SL_LLR, llr, n * (size_t)llr_stride);
    const dtl_frame_meta* d_meta = sg.in(SL_META, meta, n);
    uint8_t* d_bytes = sg.out(SL_BYTES, bytes, (size_t)n_streams * bytes_cap);
    int* d_nb = sg.out(SL_NBYTES, nbytes, (size_t)n_streams);
    dtl_tb_result* d_tbs = sg.out(SL_TBS_OUT, tbs, (size_t)n_streams * tb_cap);
    int* d_nt = sg.out(SL_NTBS, ntbs, (size_t)n_streams);
    if (sg.e != cudaSuccess) return cuda_fail(h, sg.e, "staging");
    s = dev_fec(h, sg, d_llr, llr_stride, d_meta, n_streams, frames_per_stream, d_bytes, bytes_cap, d_nb, d_tbs, tb_cap, d_nt);
    if (s != DTL_OK) return s;
    return sg.finish();
}

extern "C" dtl_status dtl_b200_ldpc_decode(dtl_b200* h, int32_t fec_id, const float* llr, int64_t count,
                                           uint8_t* bits, int32_t* nit, int mem)
{
    if (!h || !llr || !bits || count < 0) return DTL_ERR_INVALID_ARG;
    if (fec_id < 1 || fec_id >= (int)h->codes.size()) return fail(h, DTL_ERR_MISSING_TAG, "FEC tags missing");
    CK(cudaSetDevice(h->device));
    const Code& c = h->codes[fec_id];
    Stage sg(h, mem);
    LdpcParams p;
    memset(&p, 0, sizeof p);
    p.code = c.dev; p.count = count; p.max_it = h->ldpc_max_it; p.sigma2 = h->sigma2; p.mode = h->ldpc_mode;
    p.cursor = h->d_cursor; p.warp_bytes = (int)ldpc_warp_bytes(c.n, c.max_row * c.m);
    p.tab_bytes = (int)ldpc_tab_bytes(c.n, c.m, c.max_row * c.m, c.edges);
    p.fast = h->ldpc_mode == DTL_LDPC_REFERENCE && c.dev.fast && !h->no_fast_bp;
    p.fast_tab_bytes = (int)fast_tab_bytes(c.n, c.m, c.dev.nvr); p.fast_warp_bytes = (int)fast_warp_bytes(c.n, c.m, c.dev.rs);
    p.exp_tab = h->d_exp_tab; p.sigma2_lr = sigma2_for_lr(h->sigma2);
    p.lay = h->ldpc_mode == DTL_LDPC_LAYERED && c.dev.lay_steps > 0 && !h->no_fast_bp;
    p.lay_tab_bytes = (int)lay_tab_bytes(c.n, c.m, c.dev.lay_steps); p.lay_warp_bytes = (int)lay_warp_bytes(c.dev.lay_ns, c.dev.lay_steps);
    p.llr = sg.in(SL_LLR, llr, (size_t)count * c.n);
    p.bits = sg.out(SL_BITS_OUT, bits, (size_t)count * c.k);
    p.nit = sg.out(SL_NIT, nit, (size_t)count);
    if (sg.e != cudaSuccess) return cuda_fail(h, sg.e, "staging");
    h->prof.begin(DTL_STAGE_LDPC, 2);
    CK(launch_ldpc(h->stream, p, h->sm_count));
    h->prof.end();
    h->launches += 2;
    return sg.finish();
}

// ldpc_enc::encode for `count` blocks (TX side of the loopback; lib/dtl/ldpc_enc.cc:58-68)
extern "C" dtl_status dtl_b200_ldpc_encode(dtl_b200* h, int32_t fec_id, const uint8_t* data, int64_t count, uint8_t* wire, int mem)
{
    if (!h || !data || !wire || count < 0) return DTL_ERR_INVALID_ARG;
    if (fec_id < 1 || fec_id >= (int)h->codes.size()) return fail(h, DTL_ERR_MISSING_TAG, "FEC tags missing");
    CK(cudaSetDevice(h->device));
    const Code& c = h->codes[fec_id];
    Stage sg(h, mem);
    const uint8_t* d_in = sg.in(SL_ENC_IN, data, (size_t)count * c.k);
    uint8_t* d_out = sg.out(SL_ENC_OUT, wire, (size_t)count * c.n);
    if (sg.e != cudaSuccess) return cuda_fail(h, sg.e, "staging");
    h->prof.begin(DTL_STAGE_TX, 1);
    CK(launch_ldpc_encode(h->stream, c.dev, d_in, count, d_out));
    h->prof.end();
    h->launches += 1;
    return sg.finish();
}

// ---- the whole receive path ---------------------------------------------------------------------------------------------------------------
namespace {

// stages 1-4 on device pointers for S streams x F frames (the streams [h->s_off, h->s_off + S) of the handle's state)
dtl_status rx_device(dtl_b200* h, Stage& sg, int in_kind, const float2* d_in, const float2* d_taps, const dtl_frame_meta* d_meta,
                     int S, int F, dtl_frame_result* d_res, uint8_t* d_bytes, int bytes_cap, int* d_nb,
                     dtl_tb_result* d_tbs, int tb_cap, int* d_nt, int* d_fcrc)
{
    const size_t n = (size_t)S * F, LN = (size_t)h->L * h->N;
    dtl_status s;
    const float2* d_freq = d_in;
    const bool ov = h->fec_enabled && side_on(h);
    FecParams fp;
    float* d_llr = nullptr;
    uint32_t* d_sgn = nullptr;
    int launches = 0;
    if (h->fec_enabled) {
        d_llr = sg.scratch<float>(SL_LLR, n * h->llr_stride);
        d_sgn = sg.scratch<uint32_t>(SL_SGN, n * h->sgn_stride + 4);
        if (sg.e != cudaSuccess) return cuda_fail(h, sg.e, "staging");
        if ((s = fec_prepare(h, sg, d_llr, h->llr_stride, d_meta, S, F, d_bytes, bytes_cap, d_nb, d_tbs, tb_cap, d_nt, fp, d_sgn, h->sgn_stride)) != DTL_OK) return s;
        if (ov) {                               // TB planning reads the tags only: beside the FFT and the equaliser
            CK(cudaEventRecord(h->ev_fork, h->stream));
            CK(cudaStreamWaitEvent(h->side_stream, h->ev_fork, 0));
            CK(launch_fec(h->side_stream, fp, h->sm_count, &launches, nullptr, 1));
            CK(cudaEventRecord(h->ev_plan, h->side_stream));
        }
    }
    bool fused = false;
    if (in_kind == 0 && h->fec_enabled && !h->prof.on) {     // stage profiling keeps FFT and equaliser apart
        EqParams probe;
        memset(&probe, 0, sizeof probe);
        probe.pl = h->pl; probe.iq = d_in; probe.llr = d_llr;
        fused = eq_can_fuse_fft(probe);
    }
    if (in_kind == 0 && !fused) {
        float2* fq = sg.scratch<float2>(SL_FREQ, n * LN);
        if (sg.e != cudaSuccess) return cuda_fail(h, sg.e, "staging");
        if ((s = dev_demod(h, d_in, (long long)n, fq)) != DTL_OK) return s;
        d_freq = fq;
    }
    if (h->fec_enabled) {
        if (fused) { h->fuse_iq = d_in; d_freq = nullptr; }
        s = dev_equalize(h, sg, d_freq, d_taps, d_meta, S, F, nullptr, nullptr, nullptr, d_res, d_llr, nullptr, nullptr, nullptr, nullptr, d_sgn);
        if (s != DTL_OK) return s;
        if (ov) CK(cudaStreamWaitEvent(h->stream, h->ev_plan, 0));
        CK(launch_fec(h->stream, fp, h->sm_count, &launches, &h->prof, ov ? 2 : 0, ov ? h->side_stream : nullptr, ov ? h->ev_eq : nullptr));
        h->launches += launches;
        if (ov) { CK(cudaEventRecord(h->ev_side, h->side_stream)); h->side_pending = true; }
        CK(side_join(h));
        return DTL_OK;
    }
    uint8_t* fb = sg.scratch<uint8_t>(SL_FBYTES, n * h->bytes_stride);
    int* fnb = sg.scratch<int>(SL_FNBYTES, n);
    if (!d_fcrc) d_fcrc = sg.scratch<int>(SL_FCRC, n);
    if (sg.e != cudaSuccess) return cuda_fail(h, sg.e, "staging");
    s = dev_equalize(h, sg, d_freq, d_taps, d_meta, S, F, nullptr, nullptr, nullptr, d_res, nullptr, nullptr, fb, fnb, d_fcrc);
    if (s != DTL_OK) return s;
    if (d_bytes || d_nb) {
        h->prof.begin(DTL_STAGE_TBFINISH, 1);
        CK(launch_frame_gather(h->stream, fb, h->bytes_stride, fnb, S, F, d_bytes, d_bytes ? bytes_cap : 0, d_nb, h->d_err));
        h->prof.end();
        h->launches += 1;
    }
    if (d_nt) CK(cudaMemsetAsync(d_nt, 0, (size_t)S * sizeof(int), h->stream));
    // the SNR tags and the per-stream feedback state were put on the side stream by dev_equalize: everything that reads
    // `res` after this call (the D2H copies of rx_host_pipelined included) is ordered behind the main stream
    CK(side_join(h));
    return DTL_OK;
}

// DTL_MEM_HOST with many streams: the batch is cut into chunks of whole streams (streams are independent; the
// per-stream state is indexed by stream).  Chunk c+1's host->device copies run on the copy stream under chunk c's
// kernels, the small device->host result copies of chunk c on a third stream; two input staging buffers alternate.
dtl_status rx_iq_device(dtl_b200* h, Stage& sg, const float2* d_iq, int S, int F, dtl_frame_meta* d_meta, float2* d_taps,
                        dtl_header_result* d_hres, dtl_frame_result* d_res, uint8_t* d_bytes, int bytes_cap, int* d_nb,
                        dtl_tb_result* d_tbs, int tb_cap, int* d_nt, int* d_fcrc, const int* valid);

// in_kind 0 / 1: IQ / frequency-domain symbols with the caller's taps and tags; in_kind 2: IQ ALONE (dtl_b200_rx_iq): nothing
// but the samples is uploaded, the header branch of every chunk produces its taps and tags on the device (meta_out / hres come back).
dtl_status rx_host_pipelined(dtl_b200* h, int in_kind, const float* in, const float* taps, const dtl_frame_meta* meta,
                             int S, int F, int Sc, dtl_frame_result* res, uint8_t* bytes, int bytes_cap, int* nbytes,
                             dtl_tb_result* tbs, int tb_cap, int* ntbs, int* frame_crc,
                             dtl_frame_meta* meta_out = nullptr, dtl_header_result* hres = nullptr)
{
    const bool iq_only = in_kind == 2;
    if (!h->copy_stream) {
        CK(cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
        CK(cudaStreamCreateWithFlags(&h->out_stream, cudaStreamNonBlocking));
        for (int i = 0; i < 2; i++) {
            CK(cudaEventCreateWithFlags(&h->ev_h2d[i], cudaEventDisableTiming));
            CK(cudaEventCreateWithFlags(&h->ev_done[i], cudaEventDisableTiming));
        }
    }
    const size_t in_per_frame = in_kind != 1 ? (size_t)h->frame_samples : (size_t)h->L * h->N;
    const size_t nmax = (size_t)Sc * F;
    Stage sg(h, DTL_MEM_DEVICE);
    // full-batch device outputs (small), per-chunk inputs (double-buffered)
    dtl_frame_result* d_res = res ? sg.scratch<dtl_frame_result>(SL_RES, (size_t)S * F) : nullptr;
    uint8_t* d_bytes = bytes ? sg.scratch<uint8_t>(SL_BYTES, (size_t)S * bytes_cap) : nullptr;
    int* d_nb = nbytes ? sg.scratch<int>(SL_NBYTES, (size_t)S) : nullptr;
    dtl_tb_result* d_tbs = tbs ? sg.scratch<dtl_tb_result>(SL_TBS_OUT, (size_t)S * tb_cap) : nullptr;
    int* d_nt = ntbs ? sg.scratch<int>(SL_NTBS, (size_t)S) : nullptr;
    int* d_fcrc = frame_crc ? sg.scratch<int>(SL_FCRC, (size_t)S * F) : nullptr;
    dtl_frame_meta* d_meta_all = iq_only ? sg.scratch<dtl_frame_meta>(SL_HMETA, (size_t)S * F) : nullptr;
    dtl_header_result* d_hres_all = iq_only && hres ? sg.scratch<dtl_header_result>(SL_HRES, (size_t)S * F) : nullptr;
    if (sg.e != cudaSuccess) return cuda_fail(h, sg.e, "staging");
    for (int i = 0; i < 2; i++) {
        CK(h->pipe_in[i][0].ensure(nmax * in_per_frame * sizeof(float2) + 16));
        CK(h->pipe_in[i][1].ensure(nmax * h->N * sizeof(float2) + 16));
        CK(h->pipe_in[i][2].ensure(nmax * sizeof(dtl_frame_meta) + 16));
    }
    // the copy stream must not overtake work already queued on the handle's stream (e.g. dtl_b200_reset)
    CK(cudaEventRecord(h->ev_done[0], h->stream));
    CK(cudaStreamWaitEvent(h->copy_stream, h->ev_done[0], 0));
    CK(cudaStreamWaitEvent(h->out_stream, h->ev_done[0], 0));
    dtl_status s = DTL_OK;
    int c = 0;
    for (int s0 = 0; s0 < S && s == DTL_OK; s0 += Sc, c++) {
        const int Sn = std::min(Sc, S - s0), b = c & 1;
        const size_t f0 = (size_t)s0 * F, n = (size_t)Sn * F;
        if (c >= 2) CK(cudaStreamWaitEvent(h->copy_stream, h->ev_done[b], 0));     // staging buffer b is free again
        CK(cudaMemcpyAsync(h->pipe_in[b][0].p, (const float2*)in + f0 * in_per_frame, n * in_per_frame * sizeof(float2), cudaMemcpyHostToDevice, h->copy_stream));
        if (!iq_only) {
            CK(cudaMemcpyAsync(h->pipe_in[b][1].p, (const float2*)taps + f0 * h->N, n * h->N * sizeof(float2), cudaMemcpyHostToDevice, h->copy_stream));
            CK(cudaMemcpyAsync(h->pipe_in[b][2].p, meta + f0, n * sizeof(dtl_frame_meta), cudaMemcpyHostToDevice, h->copy_stream));
        }
        CK(cudaEventRecord(h->ev_h2d[b], h->copy_stream));
        CK(cudaStreamWaitEvent(h->stream, h->ev_h2d[b], 0));
        h->s_off = s0;
        if (iq_only)
            s = rx_iq_device(h, sg, (const float2*)h->pipe_in[b][0].p, Sn, F, d_meta_all + f0, (float2*)h->pipe_in[b][1].p,
                             d_hres_all ? d_hres_all + f0 : nullptr, d_res ? d_res + f0 : nullptr,
                             d_bytes ? d_bytes + (size_t)s0 * bytes_cap : nullptr, bytes_cap, d_nb ? d_nb + s0 : nullptr,
                             d_tbs ? d_tbs + (size_t)s0 * tb_cap : nullptr, tb_cap, d_nt ? d_nt + s0 : nullptr, d_fcrc ? d_fcrc + f0 : nullptr, nullptr);
        else
            s = rx_device(h, sg, in_kind, (const float2*)h->pipe_in[b][0].p, (const float2*)h->pipe_in[b][1].p,
                          (const dtl_frame_meta*)h->pipe_in[b][2].p, Sn, F, d_res ? d_res + f0 : nullptr,
                          d_bytes ? d_bytes + (size_t)s0 * bytes_cap : nullptr, bytes_cap, d_nb ? d_nb + s0 : nullptr,
                          d_tbs ? d_tbs + (size_t)s0 * tb_cap : nullptr, tb_cap, d_nt ? d_nt + s0 : nullptr, d_fcrc ? d_fcrc + f0 : nullptr);
        h->s_off = 0;
        if (s != DTL_OK) break;
        CK(cudaEventRecord(h->ev_done[b], h->stream));
        CK(cudaStreamWaitEvent(h->out_stream, h->ev_done[b], 0));
        if (res) CK(cudaMemcpyAsync(res + f0, d_res + f0, n * sizeof(dtl_frame_result), cudaMemcpyDeviceToHost, h->out_stream));
        if (bytes) CK(cudaMemcpyAsync(bytes + (size_t)s0 * bytes_cap, d_bytes + (size_t)s0 * bytes_cap, (size_t)Sn * bytes_cap, cudaMemcpyDeviceToHost, h->out_stream));
        if (nbytes) CK(cudaMemcpyAsync(nbytes + s0, d_nb + s0, (size_t)Sn * sizeof(int), cudaMemcpyDeviceToHost, h->out_stream));
        if (tbs) CK(cudaMemcpyAsync(tbs + (size_t)s0 * tb_cap, d_tbs + (size_t)s0 * tb_cap, (size_t)Sn * tb_cap * sizeof(dtl_tb_result), cudaMemcpyDeviceToHost, h->out_stream));
        if (ntbs) CK(cudaMemcpyAsync(ntbs + s0, d_nt + s0, (size_t)Sn * sizeof(int), cudaMemcpyDeviceToHost, h->out_stream));
        if (frame_crc) CK(cudaMemcpyAsync(frame_crc + f0, d_fcrc + f0, n * sizeof(int), cudaMemcpyDeviceToHost, h->out_stream));
        if (iq_only && meta_out) CK(cudaMemcpyAsync(meta_out + f0, d_meta_all + f0, n * sizeof(dtl_frame_meta), cudaMemcpyDeviceToHost, h->out_stream));
        if (iq_only && hres) CK(cudaMemcpyAsync(hres + f0, d_hres_all + f0, n * sizeof(dtl_header_result), cudaMemcpyDeviceToHost, h->out_stream));
    }
    cudaError_t e1 = cudaStreamSynchronize(h->copy_stream), e2 = cudaStreamSynchronize(h->out_stream);
    dtl_status s2 = dtl_b200_sync(h);
    if (e1 != cudaSuccess) return cuda_fail(h, e1, "copy stream");
    if (e2 != cudaSuccess) return cuda_fail(h, e2, "result stream");
    return s != DTL_OK ? s : s2;
}

} // namespace

namespace {
// Returns true when the call was served by the graph (status in *st); false: the caller runs the ordinary path.
bool rx_one_frame_graph(dtl_b200* h, int in_kind, const float* in, const float* taps, const dtl_frame_meta* meta,
                        dtl_frame_result* res, uint8_t* bytes, int bytes_cap, int* nbytes, dtl_tb_result* tbs, int tb_cap, int* ntbs,
                        int* frame_crc, dtl_status* st)
{
    static int off = -1;                             // DTL_NO_GRAPH=1: always the ordinary launches (A/B)
    if (off < 0) { const char* e = getenv("DTL_NO_GRAPH"); off = e && atoi(e) ? 1 : 0; }
    RxGraph& g = h->rxg;
    if (off || g.failed || h->prof.on) return false;
    const int mask = (res ? 1 : 0) | (bytes ? 2 : 0) | (nbytes ? 4 : 0) | (tbs ? 8 : 0) | (ntbs ? 16 : 0) | (frame_crc ? 32 : 0);
    if (g.in_kind != in_kind || g.bytes_cap != bytes_cap || g.tb_cap != tb_cap || g.mask != mask || (g.exec && g.gen != g_alloc_gen)) {
        g.drop();
        g.in_kind = in_kind; g.bytes_cap = bytes_cap; g.tb_cap = tb_cap; g.mask = mask;
    }
    if (!g.exec && g.warm < 2) { g.warm++; return false; }          // two ordinary calls first: every buffer has its final size
    const size_t in_bytes = (in_kind == 0 ? (size_t)h->frame_samples : (size_t)h->L * h->N) * sizeof(float2);
    const size_t taps_bytes = (size_t)h->N * sizeof(float2);
    auto up = [](size_t x) { return (x + 63) & ~(size_t)63; };
    if (!g.exec) {
        g.o_in = 0; g.o_taps = up(in_bytes); g.o_meta = g.o_taps + up(taps_bytes); g.o_res = g.o_meta + 64;
        g.o_bytes = g.o_res + 64; g.o_nb = g.o_bytes + up((size_t)std::max(bytes_cap, 1)); g.o_tbs = g.o_nb + 64;
        g.o_nt = g.o_tbs + up((size_t)std::max(tb_cap, 1) * sizeof(dtl_tb_result)); g.o_fcrc = g.o_nt + 64;
        const size_t need = g.o_fcrc + 64;
        if (need > g.pin_bytes) {
            if (g.pin) cudaFreeHost(g.pin);
            g.pin = nullptr; g.pin_bytes = 0;
            if (cudaHostAlloc((void**)&g.pin, need, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); g.failed = true; return false; }
            g.pin_bytes = need;
        }
        // ---- capture the ordinary code path with the page-locked buffers in the caller's place
        cudaStreamSynchronize(h->stream);
        if (h->side_stream) cudaStreamSynchronize(h->side_stream);
        const unsigned long long gen0 = g_alloc_gen;
        const uint64_t l0 = h->launches;
        cudaGraph_t graph = nullptr;
        bool ok = cudaStreamBeginCapture(h->stream, cudaStreamCaptureModeThreadLocal) == cudaSuccess;
        if (ok) {
            Stage sg(h, DTL_MEM_HOST);
            const float2* d_in = sg.in(in_kind == 0 ? SL_IQ : SL_FREQ_IN, (const float2*)(g.pin + g.o_in), in_bytes / sizeof(float2));
            const float2* d_taps = sg.in(SL_TAPS, (const float2*)(g.pin + g.o_taps), (size_t)h->N);
            const dtl_frame_meta* d_meta = sg.in(SL_META, (const dtl_frame_meta*)(g.pin + g.o_meta), 1);
            dtl_frame_result* d_res = res ? sg.out(SL_RES, (dtl_frame_result*)(g.pin + g.o_res), 1) : nullptr;
            uint8_t* d_bytes = bytes ? sg.out(SL_BYTES, (uint8_t*)(g.pin + g.o_bytes), (size_t)bytes_cap) : nullptr;
            int* d_nb = nbytes ? sg.out(SL_NBYTES, (int*)(g.pin + g.o_nb), 1) : nullptr;
            dtl_tb_result* d_tbs = tbs ? sg.out(SL_TBS_OUT, (dtl_tb_result*)(g.pin + g.o_tbs), (size_t)tb_cap) : nullptr;
            int* d_nt = ntbs ? sg.out(SL_NTBS, (int*)(g.pin + g.o_nt), 1) : nullptr;
            int* d_fcrc = frame_crc ? sg.out(SL_FCRC, (int*)(g.pin + g.o_fcrc), 1) : nullptr;
            ok = sg.e == cudaSuccess &&
                 rx_device(h, sg, in_kind, d_in, d_taps, d_meta, 1, 1, d_res, d_bytes, bytes_cap, d_nb, d_tbs, tb_cap, d_nt, d_fcrc) == DTL_OK;
            if (ok) ok = side_join(h) == cudaSuccess;
            for (auto& o : sg.outs)
                if (ok) ok = cudaMemcpyAsync(o.user, o.dev, o.bytes, cudaMemcpyDeviceToHost, h->stream) == cudaSuccess;
            if (ok) ok = cudaMemcpyAsync(h->h_err, h->d_err, sizeof(int), cudaMemcpyDeviceToHost, h->stream) == cudaSuccess;
            const cudaError_t ee = cudaStreamEndCapture(h->stream, &graph);
            ok = ok && ee == cudaSuccess && graph != nullptr && gen0 == g_alloc_gen;
        }
        if (ok) ok = cudaGraphInstantiate(&g.exec, graph, 0) == cudaSuccess;
        if (graph) cudaGraphDestroy(graph);
        g.launches = h->launches - l0;
        h->launches = l0;
        if (!ok) {                                   // nothing ran: the ordinary path serves this call and every later one
            cudaGetLastError();
            h->side_pending = false; h->hdr_pending = false;
            if (g.exec) { cudaGraphExecDestroy(g.exec); g.exec = nullptr; }
            g.failed = true;
            return false;
        }
        g.gen = g_alloc_gen;
    }
    // ---- replay
    memcpy(g.pin + g.o_in, in, in_bytes);
    memcpy(g.pin + g.o_taps, taps, taps_bytes);
    memcpy(g.pin + g.o_meta, meta, sizeof(dtl_frame_meta));
    cudaError_t e = cudaGraphLaunch(g.exec, h->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
    if (e != cudaSuccess) { *st = cuda_fail(h, e, "cudaGraphLaunch"); return true; }
    h->launches += g.launches;
    if (res) memcpy(res, g.pin + g.o_res, sizeof(dtl_frame_result));
    if (bytes) memcpy(bytes, g.pin + g.o_bytes, (size_t)bytes_cap);
    if (nbytes) memcpy(nbytes, g.pin + g.o_nb, sizeof(int));
    if (tbs) memcpy(tbs, g.pin + g.o_tbs, (size_t)tb_cap * sizeof(dtl_tb_result));
    if (ntbs) memcpy(ntbs, g.pin + g.o_nt, sizeof(int));
    if (frame_crc) memcpy(frame_crc, g.pin + g.o_fcrc, sizeof(int));
    *st = check_device_error(h);
    return true;
}
} // namespace

extern "C" dtl_status dtl_b200_rx(dtl_b200* h, int in_kind, const float* in, const float* taps, const dtl_frame_meta* meta,
                                  int32_t n_streams, int32_t frames_per_stream, dtl_frame_result* res,
                                  uint8_t* bytes, int32_t bytes_cap, int32_t* nbytes,
                                  dtl_tb_result* tbs, int32_t tb_cap, int32_t* ntbs, int32_t* frame_crc, int mem)
{
    if (!h || !in || !taps || !meta || n_streams < 0 || frames_per_stream < 0 || (in_kind != 0 && in_kind != 1)) return DTL_ERR_INVALID_ARG;
    if ((bytes && bytes_cap <= 0) || (tbs && tb_cap <= 0)) return DTL_ERR_INVALID_ARG;
    CK(cudaSetDevice(h->device));
    const int S = n_streams, F = frames_per_stream;
    const size_t n = (size_t)S * F, LN = (size_t)h->L * h->N;
    dtl_status s = check_meta_host(h, meta, (long long)n, mem);
    if (s != DTL_OK) return s;
    if ((s = ensure_streams(h, S)) != DTL_OK) return s;
    if (mem == DTL_MEM_HOST && S == 1 && F == 1 &&
        rx_one_frame_graph(h, in_kind, in, taps, meta, res, bytes, bytes_cap, nbytes, tbs, tb_cap, ntbs, frame_crc, &s)) return s;
    if (mem == DTL_MEM_HOST && F > 0) {
        // chunks of about 48 MB of input, at least 4 of them
        const size_t per_stream = (size_t)F * (in_kind == 0 ? (size_t)h->frame_samples : LN) * sizeof(float2);
        static size_t chunk_mb = 0;
        if (!chunk_mb) { const char* e = getenv("DTL_HOST_CHUNK_MB"); chunk_mb = e && atoi(e) > 0 ? (size_t)atoi(e) : 48; }
        int Sc = (int)std::max<size_t>(1, (chunk_mb << 20) / std::max<size_t>(per_stream, 1));
        if (S >= 4 * Sc || (S >= 8 && per_stream * S >= ((size_t)16 << 20))) {
            Sc = std::min(Sc, (S + 3) / 4);
            return rx_host_pipelined(h, in_kind, in, taps, meta, S, F, Sc, res, bytes, bytes_cap, nbytes, tbs, tb_cap, ntbs, frame_crc);
        }
    }
    Stage sg(h, mem);
    const float2* d_in = in_kind == 0 ? sg.in(SL_IQ, (const float2*)in, n * h->frame_samples) : sg.in(SL_FREQ_IN, (const float2*)in, n * LN);
    const float2* d_taps = sg.in(SL_TAPS, (const float2*)taps, n * h->N);
    const dtl_frame_meta* d_meta = sg.in(SL_META, meta, n);
    dtl_frame_result* d_res = sg.out(SL_RES, res, n);
    uint8_t* d_bytes = sg.out(SL_BYTES, bytes, (size_t)S * bytes_cap);
    int* d_nb = sg.out(SL_NBYTES, nbytes, (size_t)S);
    dtl_tb_result* d_tbs = sg.out(SL_TBS_OUT, tbs, (size_t)S * tb_cap);
    int* d_nt = sg.out(SL_NTBS, ntbs, (size_t)S);
    int* d_fcrc = sg.out(SL_FCRC, frame_crc, n);
    if (sg.e != cudaSuccess) return cuda_fail(h, sg.e, "staging");
    s = rx_device(h, sg, in_kind, d_in, d_taps, d_meta, S, F, d_res, d_bytes, bytes_cap, d_nb, d_tbs, tb_cap, d_nt, d_fcrc);
    if (s != DTL_OK) return s;
    return sg.finish();
}

// ---- header branch ---------------------------------------------------------------------------------------------------------------------
namespace {
dtl_status dev_header(dtl_b200* h, Stage& sg, const float2* d_iq, int S, int F, dtl_frame_meta* d_meta, float2* d_taps,
                      dtl_header_result* d_hres)
{
    if (!h->have_hdr) return fail(h, DTL_ERR_INVALID_ARG, "header branch needs the sync words (and header_syms * fft_len <= 65535)");
    const long long n = (long long)S * F;
    float2* hfreq = sg.scratch<float2>(SL_HFREQ, (size_t)n * (2 + h->H) * h->N);
    if (sg.e != cudaSuccess) return cuda_fail(h, sg.e, "scratch");
    h->prof.begin(DTL_STAGE_HEADER, 1);
    CK(launch_fft_fwd(h->stream, d_iq, hfreq, n * (2 + h->H), 2 + h->H, h->frame_samples, 0, h->N + h->cp, h->cp, h->N, h->tw));
    HdrParams p;
    memset(&p, 0, sizeof p);
    p.hd = h->hd; p.pl = h->pl; p.cn = h->cn; p.hfreq = hfreq; p.n_frames = n; p.n_streams = S; p.F = F;
    p.meta = d_meta; p.taps = d_taps; p.hres = d_hres; p.st = (StreamState*)h->st_buf.p + h->s_off;
    p.valid = h->hdr_valid; h->hdr_valid = nullptr;
    CK(launch_header(h->stream, p));
    h->prof.end();
    h->launches += 3;
    return DTL_OK;
}
} // namespace

extern "C" dtl_status dtl_b200_rx_headers(dtl_b200* h, const float* iq, int32_t n_streams, int32_t frames_per_stream,
                                          dtl_frame_meta* meta, float* taps, dtl_header_result* hres, int mem)
{
    if (!h || !iq || !meta || n_streams < 0 || frames_per_stream < 0) return DTL_ERR_INVALID_ARG;
    CK(cudaSetDevice(h->device));
    const int S = n_streams, F = frames_per_stream;
    const size_t n = (size_t)S * F;
    dtl_status s = ensure_streams(h, S);
    if (s != DTL_OK) return s;
    Stage sg(h, mem);
    const float2* d_iq = sg.in(SL_IQ, (const float2*)iq, n * h->frame_samples);
    dtl_frame_meta* d_meta = sg.out(SL_HMETA, meta, n);
    float2* d_taps = sg.out(SL_HTAPS, (float2*)taps, n * h->N);
    dtl_header_result* d_hres = sg.out(SL_HRES, hres, n);
    if (sg.e != cudaSuccess) return cuda_fail(h, sg.e, "staging");
    if ((s = dev_header(h, sg, d_iq, S, F, d_meta, d_taps, d_hres)) != DTL_OK) return s;
    return sg.finish();
}

namespace {
// header branch, then the payload branch on the tags it produced (device pointers).  valid: frame slots that hold a frame (or null)
dtl_status rx_iq_device(dtl_b200* h, Stage& sg, const float2* d_iq, int S, int F, dtl_frame_meta* d_meta, float2* d_taps,
                        dtl_header_result* d_hres, dtl_frame_result* d_res, uint8_t* d_bytes, int bytes_cap, int* d_nb,
                        dtl_tb_result* d_tbs, int tb_cap, int* d_nt, int* d_fcrc, const int* valid)
{
    dtl_status s;
    h->hdr_valid = valid;
    if (side_on(h)) {
        // the header branch (instruction-bound) runs on the side stream beside the payload FFT (HBM-bound), which needs neither
        // tags nor taps; TB planning follows it on the same stream, the equaliser waits for it
        CK(cudaEventRecord(h->ev_fork, h->stream));
        CK(cudaStreamWaitEvent(h->side_stream, h->ev_fork, 0));
        std::swap(h->stream, h->side_stream);
        s = dev_header(h, sg, d_iq, S, F, d_meta, d_taps, d_hres);
        std::swap(h->stream, h->side_stream);
        if (s != DTL_OK) return s;
        CK(cudaEventRecord(h->ev_hdr, h->side_stream));
        CK(cudaEventRecord(h->ev_side, h->side_stream));
        h->hdr_pending = true; h->side_pending = true;
    } else if ((s = dev_header(h, sg, d_iq, S, F, d_meta, d_taps, d_hres)) != DTL_OK) return s;
    s = rx_device(h, sg, 0, d_iq, d_taps, d_meta, S, F, d_res, d_bytes, bytes_cap, d_nb, d_tbs, tb_cap, d_nt, d_fcrc);
    if (s != DTL_OK) return s;
    return DTL_OK;
}
} // namespace


extern "C" dtl_status dtl_b200_rx_iq(dtl_b200* h, const float* iq, int32_t n_streams, int32_t frames_per_stream,
                                     dtl_frame_meta* meta_out, dtl_header_result* hres, dtl_frame_result* res,
                                     uint8_t* bytes, int32_t bytes_cap, int32_t* nbytes,
                                     dtl_tb_result* tbs, int32_t tb_cap, int32_t* ntbs, int32_t* frame_crc, int mem)
{
    if (!h || !iq || n_streams < 0 || frames_per_stream < 0) return DTL_ERR_INVALID_ARG;
    if ((bytes && bytes_cap <= 0) || (tbs && tb_cap <= 0)) return DTL_ERR_INVALID_ARG;
    CK(cudaSetDevice(h->device));
    const int S = n_streams, F = frames_per_stream;
    const size_t n = (size_t)S * F;
    dtl_status s = ensure_streams(h, S);
    if (s != DTL_OK) return s;
    if (mem == DTL_MEM_HOST && F > 0) {              // many streams from host memory: chunks of whole streams, copies under compute
        const size_t per_stream = (size_t)F * (size_t)h->frame_samples * sizeof(float2);
        static size_t chunk_mb = 0;
        if (!chunk_mb) { const char* e = getenv("DTL_HOST_CHUNK_MB"); chunk_mb = e && atoi(e) > 0 ? (size_t)atoi(e) : 48; }
        int Sc = (int)std::max<size_t>(1, (chunk_mb << 20) / std::max<size_t>(per_stream, 1));
        if (S >= 4 * Sc || (S >= 8 && per_stream * S >= ((size_t)16 << 20))) {
            Sc = std::min(Sc, (S + 3) / 4);
            return rx_host_pipelined(h, 2, iq, nullptr, nullptr, S, F, Sc, res, bytes, bytes_cap, nbytes, tbs, tb_cap, ntbs, frame_crc, meta_out, hres);
        }
    }
    Stage sg(h, mem);
    const float2* d_iq = sg.in(SL_IQ, (const float2*)iq, n * h->frame_samples);
    dtl_frame_meta* d_meta = meta_out ? sg.out(SL_HMETA, meta_out, n) : sg.scratch<dtl_frame_meta>(SL_HMETA, n);
    float2* d_taps = sg.scratch<float2>(SL_HTAPS, n * h->N);
    dtl_header_result* d_hres = hres ? sg.out(SL_HRES, hres, n) : nullptr;
    dtl_frame_result* d_res = sg.out(SL_RES, res, n);
    uint8_t* d_bytes = sg.out(SL_BYTES, bytes, (size_t)S * bytes_cap);
    int* d_nb = sg.out(SL_NBYTES, nbytes, (size_t)S);
    dtl_tb_result* d_tbs = sg.out(SL_TBS_OUT, tbs, (size_t)S * tb_cap);
    int* d_nt = sg.out(SL_NTBS, ntbs, (size_t)S);
    int* d_fcrc = sg.out(SL_FCRC, frame_crc, n);
    if (sg.e != cudaSuccess) return cuda_fail(h, sg.e, "staging");
    s = rx_iq_device(h, sg, d_iq, S, F, d_meta, d_taps, d_hres, d_res, d_bytes, bytes_cap, d_nb, d_tbs, tb_cap, d_nt, d_fcrc, nullptr);
    if (s != DTL_OK) return s;
    return sg.finish();
}

// ---- TX FEC framer ---------------------------------------------------------------------------------------
extern "C" dtl_status dtl_b200_fec_frame(dtl_b200* h, const uint8_t* data, int32_t in_cap, const int32_t* in_len,
                                         const int32_t* cnst, const int32_t* fec_id, int32_t n_streams, int32_t frames_per_stream,
                                         uint8_t* chunks, dtl_frame_meta* meta, int32_t* consumed, int32_t* tbs_started, int mem)
{
    if (!h || !data || !cnst || !fec_id || !chunks || !meta || n_streams < 0 || frames_per_stream < 0 || in_cap < 0) return DTL_ERR_INVALID_ARG;
    if (!h->fec_enabled || h->codes.size() < 2) return fail(h, DTL_ERR_INVALID_ARG, "the FEC framer needs a FEC-enabled configuration");
    CK(cudaSetDevice(h->device));
    const int S = n_streams, F = frames_per_stream;
    const size_t n = (size_t)S * F;
    const int maxbits = (h->max_ncws * h->max_n + 15) & ~15;
    if (h->tx_streams != S) {                    // first call (or after dtl_b200_reset): framer state starts empty
        CK(cudaStreamSynchronize(h->stream));
        CK(h->txst_buf.ensure(std::max<size_t>((size_t)S * sizeof(TxStream), 16)));
        CK(h->txcarry_buf.ensure((size_t)S * maxbits + 16));
        CK(launch_txf_reset(h->stream, (TxStream*)h->txst_buf.p, S));
        h->tx_streams = S;
    }
    Stage sg(h, mem);
    TxfParams p;
    memset(&p, 0, sizeof p);
    p.st = (TxStream*)h->txst_buf.p; p.n_streams = S; p.F = F; p.cap = h->cap;
    p.data = sg.in(SL_TXDATA, data, (size_t)S * in_cap); p.in_cap = in_cap;
    p.in_len = in_len ? sg.in(SL_TXLEN, in_len, (size_t)S) : nullptr;
    p.cnst = sg.in(SL_TXCNST, cnst, n); p.fec = sg.in(SL_TXFEC, fec_id, n);
    p.n_codes = (int)h->codes.size() - 1; p.codes = h->d_codes;
    for (int c = 0; c <= DTL_CNST_MAX; c++) p.bps_tab[c] = cnst_bps(c);
    p.tbs = sg.scratch<TxTb>(SL_TXTBS, n); p.ntb = sg.scratch<int>(SL_TXNTB, (size_t)S); p.keep = sg.scratch<int>(SL_TXKEEP, (size_t)S);
    p.segs = sg.scratch<TxSeg>(SL_TXSEGS, 2 * n); p.tbcrc = sg.scratch<uint32_t>(SL_TXCRC, n);
    p.tbbits = sg.scratch<uint8_t>(SL_TXBITS, n * (size_t)maxbits); p.maxbits = maxbits;
    p.carry = (uint8_t*)h->txcarry_buf.p;
    p.max_ncws = h->max_ncws; p.max_kwords = (h->max_k + 31) / 32;
    p.chunks = sg.out(SL_TXCHUNKS, chunks, n * (size_t)h->cap);
    p.meta = sg.out(SL_TXMETA, meta, n);
    p.consumed = consumed ? sg.out(SL_TXCONS, consumed, (size_t)S) : nullptr;
    p.tbs_started = tbs_started ? sg.out(SL_TXSTARTED, tbs_started, (size_t)S) : nullptr;
    p.crc_table = h->crc_table; p.err = h->d_err;
    if (sg.e != cudaSuccess) return cuda_fail(h, sg.e, "staging");
    int launches = 0;
    CK(launch_txf(h->stream, p, &launches));
    h->launches += launches;
    return sg.finish();
}

// ---- stage 5 --------------------------------------------------------------------------------------------------------------------------------
extern "C" dtl_status dtl_b200_tx(dtl_b200* h, const uint8_t* hdr_bits, const uint8_t* chunks, const dtl_frame_meta* meta,
                                  int64_t n_frames, float* iq, int mem)
{
    if (!h || !hdr_bits || !chunks || !meta || !iq || n_frames < 0) return DTL_ERR_INVALID_ARG;
    if (!h->have_sync) return fail(h, DTL_ERR_INVALID_ARG, "sync words were not configured");
    CK(cudaSetDevice(h->device));
    dtl_status s = check_meta_host(h, meta, n_frames, mem);
    if (s != DTL_OK) return s;
    const size_t n = (size_t)n_frames;
    Stage sg(h, mem);
    TxParams p;
    memset(&p, 0, sizeof p);
    p.N = h->N; p.cp = h->cp; p.H = h->H; p.L = h->L; p.cap = h->cap; p.nocc = h->pl.nocc; p.hdr_stride = h->hdr_stride;
    p.n_frames = n_frames;
    p.hdr_bits = sg.in(SL_HDR, hdr_bits, n * h->hdr_stride);
    p.chunks = sg.in(SL_CHUNKS, chunks, n * h->cap);
    p.meta = sg.in(SL_META, meta, n);
    p.iq = sg.out(SL_TXIQ, (float2*)iq, n * h->frame_samples);
    if (sg.e != cudaSuccess) return cuda_fail(h, sg.e, "staging");
    p.sync1 = h->d_sync1; p.sync2 = h->d_sync2; p.ser_pos = h->pl.ser_pos; p.tx_base = h->d_tx_base;
    p.tx_pil_mask = h->d_tx_pil_mask; p.tx_pil_val = h->d_tx_pil_val; p.tw = h->tw; p.cn = h->cn;
    h->prof.begin(DTL_STAGE_TX, 1);
    CK(launch_tx_mod(h->stream, p));
    h->prof.end();
    h->launches += 1;
    return sg.finish();
}

// ---- synchronisation front end + receive path ---------------------------------------------
extern "C" dtl_status dtl_b200_rx_stream(dtl_b200* h, const float* iq, int32_t n_streams, int64_t samples_per_stream, int32_t frames_cap,
                                         int32_t* n_frames, int64_t* frame_start, float* frames_out,
                                         dtl_frame_meta* meta_out, dtl_header_result* hres, dtl_frame_result* res,
                                         uint8_t* bytes, int32_t bytes_cap, int32_t* nbytes,
                                         dtl_tb_result* tbs, int32_t tb_cap, int32_t* ntbs, int32_t* frame_crc, int mem)
{
    if (!h || !iq || !n_frames || n_streams <= 0 || samples_per_stream <= 0 || frames_cap <= 0) return DTL_ERR_INVALID_ARG;
    if (samples_per_stream > 0x3fffffff) return DTL_ERR_INVALID_ARG;
    if ((bytes && bytes_cap <= 0) || (tbs && tb_cap <= 0)) return DTL_ERR_INVALID_ARG;
    if (!h->have_hdr) return fail(h, DTL_ERR_INVALID_ARG, "dtl_b200_rx_stream needs the sync words");
    CK(cudaSetDevice(h->device));
    const int S = n_streams, F = frames_cap, T = (int)samples_per_stream, FS = h->frame_samples, N = h->N;
    const size_t n = (size_t)S * F;
    dtl_status s = ensure_streams(h, S);
    if (s != DTL_OK) return s;
    if (h->sync_streams != S) {                  // first call (or after dtl_b200_reset): detector / regulariser state starts empty
        CK(cudaStreamSynchronize(h->stream));
        h->sync_keep = FS + 2 * (N + h->cp) + 2 * N + 64;
        h->sync_tmax = T;
        h->sync_wcap = h->sync_keep + T;
        CK(h->sync_st.ensure(std::max<size_t>((size_t)S * sizeof(SyncState), 16)));
        for (int b = 0; b < 2; b++) CK(h->sync_w[b].ensure((size_t)S * h->sync_wcap * sizeof(float2) + 16));
        CK(launch_sync_reset(h->stream, (SyncState*)h->sync_st.p, S));
        h->sync_cur = 0;
        h->sync_streams = S;
    }
    if (T > h->sync_tmax) return fail(h, DTL_ERR_CAPACITY, "samples_per_stream larger than in the first call after dtl_b200_reset");
    Stage sg(h, mem);
    SyncParams p;
    memset(&p, 0, sizeof p);
    p.st = (SyncState*)h->sync_st.p; p.n_streams = S; p.T = T; p.N = N; p.cp = h->cp; p.FS = FS; p.threshold = h->sync_threshold;
    p.x = sg.in(SL_SYX, (const float2*)iq, (size_t)S * T);
    p.w_old = (const float2*)h->sync_w[h->sync_cur].p; p.w_new = (float2*)h->sync_w[h->sync_cur ^ 1].p;
    p.w_cap = h->sync_wcap; p.keep_cap = h->sync_keep;
    p.flag_words = (T + h->sync_keep) / 32 + 64;
    p.flags = sg.scratch<uint32_t>(SL_SYFLAGS, (size_t)S * p.flag_words);
    p.seg_cap = 2 * F + 16 + DTL_SYNC_SEGC; p.trig_cap = 2 * F + 16;
    p.segs = sg.scratch<SyncSeg>(SL_SYSEGS, (size_t)S * p.seg_cap);
    p.n_seg = sg.scratch<int>(SL_SYNSEG, (size_t)S * 3); p.n_raw = p.n_seg + S; p.n_fixed = p.n_seg + 2 * S;
    p.raw = sg.scratch<long long>(SL_SYRAW, (size_t)S * p.trig_cap);
    p.fixed = sg.scratch<long long>(SL_SYFIXED, (size_t)S * p.trig_cap);
    p.frames_cap = F;
    p.starts = sg.scratch<long long>(SL_SYSTARTS, n);
    p.starts_out = frame_start ? (long long*)sg.out(SL_SYSTARTS_OUT, (long long*)frame_start, n) : nullptr;
    p.n_frames = sg.out(SL_SYNFR, n_frames, (size_t)S);
    float2* d_frames = frames_out ? sg.out(SL_SYFRAMES, (float2*)frames_out, n * FS) : sg.scratch<float2>(SL_SYFRAMES, n * FS);
    p.frames = d_frames;
    p.valid = sg.scratch<int>(SL_SYVALID, n);
    p.err = h->d_err;
    dtl_frame_meta* d_meta = meta_out ? sg.out(SL_HMETA, meta_out, n) : sg.scratch<dtl_frame_meta>(SL_HMETA, n);
    float2* d_taps = sg.scratch<float2>(SL_HTAPS, n * N);
    dtl_header_result* d_hres = hres ? sg.out(SL_HRES, hres, n) : nullptr;
    dtl_frame_result* d_res = sg.out(SL_RES, res, n);
    uint8_t* d_bytes = sg.out(SL_BYTES, bytes, (size_t)S * bytes_cap);
    int* d_nb = sg.out(SL_NBYTES, nbytes, (size_t)S);
    dtl_tb_result* d_tbs = sg.out(SL_TBS_OUT, tbs, (size_t)S * tb_cap);
    int* d_nt = sg.out(SL_NTBS, ntbs, (size_t)S);
    int* d_fcrc = sg.out(SL_FCRC, frame_crc, n);
    if (sg.e != cudaSuccess) return cuda_fail(h, sg.e, "staging");
    int launches = 0;
    h->prof.begin(DTL_STAGE_SYNC, 4);
    CK(launch_sync(h->stream, p, &launches));
    h->prof.end();
    h->launches += launches;
    h->sync_cur ^= 1;
    s = rx_iq_device(h, sg, d_frames, S, F, d_meta, d_taps, d_hres, d_res, d_bytes, bytes_cap, d_nb, d_tbs, tb_cap, d_nt, d_fcrc, p.valid);
    if (s != DTL_OK) return s;
    return sg.finish();
}
