// kernels_sync.cu -- synchronisation front end: S continuous IQ streams -> frames cut at their
// boundaries, fine carrier-frequency offset removed.  Replaces, per stream (python/dtl/ofdm_receiver.py:70-97):
//   digital.ofdm_sync_sc_cfb(fft_len, cp_len, threshold)   Schmidl & Cox timing metric, plateau detector, fine-CFO estimate
//   blocks.delay(fft_len + cp_len) + analog.frequency_modulator_fc(-2 / fft_len) + blocks.multiply_cc
//   dtl.ofdm_adaptive_frame_detect_bb(frame samples)       lib/dtl/ofdm_adaptive_frame_detect_bb_impl.cc:64-156 (trigger regulariser)
//   digital.header_payload_demux (its trigger input)       frames start at the regularised triggers
//
// Three kernels per call over the stream's work buffer W = [samples kept from earlier calls | the call's new samples]:
//   sync_metric_kernel  per tile of 1024 positions: c[m] = -x[m] conj(x[m - N/2]) and e[m] = 0.5 |x[m]|^2 into shared memory,
//                       window sums over N/2 (complex) and N (real) samples as suffix-of-previous-block + prefix-of-current-block
//                       (blocks aligned to the global sample index, fixed summation order: the same floats whatever the caller's
//                       chunking), metric |P|^2 / E^2 >= threshold as one bit per position;
//   sync_scan_kernel    one thread per stream, sequential but sparse: plateau detector on the bit words, fine-CFO phase segments
//                       (arg P at the raw triggers), the regulariser's byte loop walked event by event, the demux's busy rule;
//   sync_cut_kernel     every completed frame: delayed samples times exp(j phase), phase piecewise linear between raw triggers.
// The tail each stream still needs (metric history, delay line of frames not complete yet) is copied to the other work
// buffer together with the next call's samples (sync_append_kernel).
#include "kernels.cuh"

namespace dtl {

__device__ __forceinline__ float2 ld_w(const float2* w, long long n, long long base, int len)
{
    const long long i = n - base;
    return (i >= 0 && i < len) ? w[i] : make_float2(0.f, 0.f);            // before the stream began: zeros (the delay lines' initial state)
}

// new work buffer = [kept tail of the old one | new samples]; also moves the per-stream bookkeeping on
__global__ void __launch_bounds__(256) sync_append_kernel(SyncParams p)
{
    const int s = blockIdx.y;
    const SyncState st = p.st[s];
    const float2* src = p.w_old + (size_t)s * p.w_cap + st.keep_off;
    float2* dst = p.w_new + (size_t)s * p.w_cap;
    const float2* x = p.x + (size_t)s * p.T;
    const int keep = st.keep_len, total = keep + p.T;
    const int stride = gridDim.x * blockDim.x;
    for (int i0 = blockIdx.x * blockDim.x + threadIdx.x; i0 < total; i0 += 4 * stride) {       // four loads in flight per thread
        float2 v[4];
#pragma unroll
        for (int u = 0; u < 4; u++) { const int i = i0 + u * stride; if (i < total) v[u] = i < keep ? src[i] : __ldcs(&x[i - keep]); }
#pragma unroll
        for (int u = 0; u < 4; u++) { const int i = i0 + u * stride; if (i < total) dst[i] = v[u]; }
    }
}

// ---- Schmidl & Cox metric, one bit per position ------------------------------------------------------------------------
// Tile of TILE positions starting at a multiple of TILE (global index).  Shared memory: x[TILE + N + N/2], c / prefix / suffix
// of c [TILE + N] (float2), e / prefix / suffix of e [TILE + N] (float).
// NFIX: fft_len known at compile time (64: the reference's PHY) so that the chains below unroll; 0 = any fft_len
template <int TILE, int NFIX>
__global__ void __launch_bounds__(256) sync_metric_kernel(SyncParams p)
{
    extern __shared__ __align__(16) unsigned char sm[];
    const int N = NFIX ? NFIX : p.N, W1 = N / 2, W2 = N, s = blockIdx.y;
    const SyncState st = p.st[s];
    const int len = st.keep_len + p.T;
    const long long base = st.base_next;                        // global index of W[0] for this call
    const long long lo = st.scan_lo & ~(long long)(TILE - 1);   // first tile that holds a position to examine
    const long long p0 = lo + (long long)blockIdx.x * TILE;
    const long long total = st.total + p.T;
    if (p0 >= total) return;
    const float2* w = p.w_new + (size_t)s * p.w_cap;
    // The scan arrays are padded by one element per 32 (PAD): the per-block chains below start 32 or 64 elements apart, which
    // without the padding is ONE bank for every thread of the warp (a 32-way conflict on every access).
#define PAD(i) ((i) + ((i) >> 5))
    float2* xs = (float2*)sm;                                   // x[p0 - N - W1 .. p0 + TILE)
    const int nx = TILE + N + W1, nc = TILE + N, ncp = PAD(nc) + 1;
    const int na = nx > ncp ? nx : ncp;
    float2* cs = xs + na;                                       // c[p0 - N .. p0 + TILE), padded like the ones below
    float2* pc = xs;                                            // the prefix sums take the samples' place (dead once c and e exist)
    float2* sc = cs + ncp;
    float* es = (float*)(sc + ncp);
    float* pe = es + ncp;
    float* se = pe + ncp;
    for (int i0 = threadIdx.x; i0 < nx; i0 += 4 * blockDim.x) {               // four loads in flight per thread
        float2 v[4];
#pragma unroll
        for (int u = 0; u < 4; u++) { const int i = i0 + u * blockDim.x; v[u] = i < nx ? ld_w(w, p0 - N - W1 + i, base, len) : make_float2(0.f, 0.f); }
#pragma unroll
        for (int u = 0; u < 4; u++) { const int i = i0 + u * blockDim.x; if (i < nx) xs[i] = v[u]; }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < nc; i += blockDim.x) {
        const float2 a = xs[i + W1], b = xs[i];                 // x[m], x[m - N/2], m = p0 - N + i
        cs[PAD(i)] = make_float2(-__fmaf_rn(a.x, b.x, __fmul_rn(a.y, b.y)), -__fmaf_rn(a.y, b.x, -__fmul_rn(a.x, b.y)));
        es[PAD(i)] = __fmul_rn(0.5f, __fmaf_rn(a.x, a.x, __fmul_rn(a.y, a.y)));
    }
    __syncthreads();
    // prefix (ascending) and suffix (descending) sums inside every aligned block: one thread per (block, direction) chain
    const int nb1 = nc / W1, nb2 = nc / W2;
    if (NFIX == 64 && blockDim.x == 256) {
        // fft_len 64: the same chains, laid out so that no warp mixes chain kinds and both directions run the same code (a
        // per-thread step of +-1 element): the 34 E chains (64 long) on warps 0-1, the P chains as 136 chains of ONE float
        // component (32 long) on warps 2-6.  The other threads wait at the barrier.
        const int tid = threadIdx.x;
        if (tid < 2 * nb2) {
            const int dir = tid & 1, bb = tid >> 1, o = bb * 64 + 2 * bb;                 // PAD(bb * 64)
            float* dst = dir ? se : pe;
            float acc = 0.f;
#pragma unroll
            for (int r = 0; r < 2; r++) {
                const int at = o + (dir ? 63 - 32 * r + (1 - r) : 33 * r), step = dir ? -1 : 1;
                const float* sp = es + at; float* dp = dst + at;
                float v[32];
#pragma unroll
                for (int ii = 0; ii < 32; ii++) v[ii] = sp[step * ii];
#pragma unroll
                for (int ii = 0; ii < 32; ii++) { acc = (r == 0 && ii == 0) ? v[0] : __fadd_rn(acc, v[ii]); dp[step * ii] = acc; }
            }
        } else if (tid >= 64 && tid < 64 + 4 * nb1) {
            const int c = tid - 64, comp = c & 1, dir = (c >> 1) & 1, bb = c >> 2;
            const int at = 2 * (33 * bb + (dir ? 31 : 0)) + comp, step = dir ? -2 : 2;   // float index into the float2 arrays
            const float* sp = (const float*)cs + at; float* dp = (float*)(dir ? sc : pc) + at;
            float v[32];
#pragma unroll
            for (int ii = 0; ii < 32; ii++) v[ii] = sp[step * ii];
            float acc = v[0];
            dp[0] = acc;
#pragma unroll
            for (int ii = 1; ii < 32; ii++) { acc = __fadd_rn(acc, v[ii]); dp[step * ii] = acc; }
        }
    } else
    for (int b = threadIdx.x; b < 2 * (nb1 + nb2); b += blockDim.x) {
        const int dir = b & 1, bb = b >> 1;
        if (bb < nb1) {
            const int o = bb * W1;
            if (!dir) {
                float2 acc = cs[PAD(o)];
                pc[PAD(o)] = acc;
#pragma unroll 16
                for (int i = 1; i < W1; i++) { const float2 c = cs[PAD(o + i)]; acc.x = __fadd_rn(acc.x, c.x); acc.y = __fadd_rn(acc.y, c.y); pc[PAD(o + i)] = acc; }
            } else {
                float2 acc = cs[PAD(o + W1 - 1)];
                sc[PAD(o + W1 - 1)] = acc;
#pragma unroll 16
                for (int i = W1 - 2; i >= 0; i--) { const float2 c = cs[PAD(o + i)]; acc.x = __fadd_rn(acc.x, c.x); acc.y = __fadd_rn(acc.y, c.y); sc[PAD(o + i)] = acc; }
            }
        } else {
            const int o = (bb - nb1) * W2;
            if (!dir) {
                float acc = es[PAD(o)];
                pe[PAD(o)] = acc;
#pragma unroll 16
                for (int i = 1; i < W2; i++) { acc = __fadd_rn(acc, es[PAD(o + i)]); pe[PAD(o + i)] = acc; }
            } else {
                float acc = es[PAD(o + W2 - 1)];
                se[PAD(o + W2 - 1)] = acc;
#pragma unroll 16
                for (int i = W2 - 2; i >= 0; i--) { acc = __fadd_rn(acc, es[PAD(o + i)]); se[PAD(o + i)] = acc; }
            }
        }
    }
    __syncthreads();
    // window sums and the comparison: position n = p0 + j, index N + j in the c / e arrays
    uint32_t* flags = p.flags + (size_t)s * p.flag_words;
    for (int j = threadIdx.x; j < TILE; j += blockDim.x) {
        const int i = N + j;
        float2 P = pc[PAD(i)];
        if ((i + 1) % W1) { const float2 t = sc[PAD(i - W1 + 1)]; P.x = __fadd_rn(P.x, t.x); P.y = __fadd_rn(P.y, t.y); }
        float E = pe[PAD(i)];
        if ((i + 1) % W2) E = __fadd_rn(E, se[PAD(i - W2 + 1)]);
        const float m = __fdiv_rn(__fmaf_rn(P.x, P.x, __fmul_rn(P.y, P.y)), __fmul_rn(E, E));
        const long long n = p0 + j;
        const unsigned word = __ballot_sync(0xffffffffu, n < total && m >= p.threshold);
        if ((threadIdx.x & 31) == 0) flags[(n - lo) >> 5] = word;
    }
}

// P[n] with the summation order of sync_metric_kernel (the fine-CFO estimate is arg P at a raw trigger)
__device__ float2 sync_p_at(const float2* w, long long n, long long base, int len, int N)
{
    const int W1 = N / 2;
    auto c_at = [&](long long m) {
        const float2 a = ld_w(w, m, base, len), b = ld_w(w, m - W1, base, len);
        return make_float2(-__fmaf_rn(a.x, b.x, __fmul_rn(a.y, b.y)), -__fmaf_rn(a.y, b.x, -__fmul_rn(a.x, b.y)));
    };
    const long long b0 = n - (n % W1);
    float2 pre = c_at(b0);
    for (long long m = b0 + 1; m <= n; m++) { const float2 c = c_at(m); pre.x = __fadd_rn(pre.x, c.x); pre.y = __fadd_rn(pre.y, c.y); }
    if (n == b0 + W1 - 1) return pre;
    float2 suf = c_at(b0 - 1);
    for (long long m = b0 - 2; m >= n - W1 + 1; m--) { const float2 c = c_at(m); suf.x = __fadd_rn(suf.x, c.x); suf.y = __fadd_rn(suf.y, c.y); }
    return make_float2(__fadd_rn(pre.x, suf.x), __fadd_rn(pre.y, suf.y));
}

// sync_p_at by a whole warp: the lanes fetch 32 terms at a time, the additions run in the same strict order (every lane does them
// on shuffled terms, so the result is uniform) -- 32 loads in flight instead of one thread's dependent chain.
__device__ float2 sync_p_at_warp(const float2* w, long long n, long long base, int len, int N, int lane)
{
    const int W1 = N / 2;
    auto c_at = [&](long long m) {
        const float2 a = ld_w(w, m, base, len), b = ld_w(w, m - W1, base, len);
        return make_float2(-__fmaf_rn(a.x, b.x, __fmul_rn(a.y, b.y)), -__fmaf_rn(a.y, b.x, -__fmul_rn(a.x, b.y)));
    };
    const long long b0 = n - (n % W1);
    float2 pre = make_float2(0.f, 0.f);
    bool first = true;
    for (long long m0 = b0; m0 <= n; m0 += 32) {                 // ascending from the block's first element
        const int cnt = (int)min(32ll, n - m0 + 1);
        float2 c = make_float2(0.f, 0.f);
        if (lane < cnt) c = c_at(m0 + lane);
        for (int l = 0; l < cnt; l++) {
            const float cx = __shfl_sync(0xffffffffu, c.x, l), cy = __shfl_sync(0xffffffffu, c.y, l);
            if (first) { pre = make_float2(cx, cy); first = false; } else { pre.x = __fadd_rn(pre.x, cx); pre.y = __fadd_rn(pre.y, cy); }
        }
    }
    if (n == b0 + W1 - 1) return pre;
    float2 suf = make_float2(0.f, 0.f);
    first = true;
    for (long long m0 = b0 - 1; m0 >= n - W1 + 1; m0 -= 32) {    // descending from the previous block's last element
        const int cnt = (int)min(32ll, m0 - (n - W1 + 1) + 1);
        float2 c = make_float2(0.f, 0.f);
        if (lane < cnt) c = c_at(m0 - lane);
        for (int l = 0; l < cnt; l++) {
            const float cx = __shfl_sync(0xffffffffu, c.x, l), cy = __shfl_sync(0xffffffffu, c.y, l);
            if (first) { suf = make_float2(cx, cy); first = false; } else { suf.x = __fadd_rn(suf.x, cx); suf.y = __fadd_rn(suf.y, cy); }
        }
    }
    return make_float2(__fadd_rn(pre.x, suf.x), __fadd_rn(pre.y, suf.y));
}

// first flagged position in [i, end), else end: the lanes look at 32 flag words (1024 positions) per round
__device__ long long sync_next_flag(const uint32_t* flags, long long flo, long long i, long long end, int lane)
{
    if (i >= end) return end;
    const long long w0 = (i - flo) >> 5, we = (end - 1 - flo) >> 5;
    const int sh = (int)((i - flo) & 31);
    for (long long wb = w0; wb <= we; wb += 32) {
        const long long wj = wb + lane;
        uint32_t word = wj <= we ? flags[wj] : 0u;
        if (wj == w0) word &= 0xffffffffu << sh;                // positions below i
        const unsigned m = __ballot_sync(0xffffffffu, word != 0u);
        if (m) {
            const int l = __ffs(m) - 1;
            const uint32_t wv = __shfl_sync(0xffffffffu, word, l);
            const long long n = flo + ((wb + l) << 5) + (__ffs(wv) - 1);
            return n < end ? n : end;
        }
    }
    return end;
}

__device__ __forceinline__ double sync_phase(const SyncSeg* segs, int nseg, long long n, int N)
{
    int k = -1;
    for (int i = 0; i < nseg; i++) { if (segs[i].t <= n) k = i; else break; }
    if (k < 0) return 0.0;
    return segs[k].ph + (double)(n - segs[k].t) * (-2.0 / N) * (double)segs[k].f;
}

// One WARP per stream: everything that is sequential in the stream, on the sparse events only.  The state machines run
// replicated in every lane (uniform control flow; identical values written to the per-stream lists by every lane); the lanes
// share the two things that touch memory in bulk: the search for the next flagged position and the window sum at a trigger.
__global__ void __launch_bounds__(128) sync_scan_kernel(SyncParams p)
{
    const int s = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (s >= p.n_streams) return;
    SyncState st = p.st[s];
    const int N = p.N, cp = p.cp, FS = p.FS, D = N + cp;
    const int len = st.keep_len + p.T;
    const long long base = st.base_next;
    const float2* w = p.w_new + (size_t)s * p.w_cap;
    const uint32_t* flags = p.flags + (size_t)s * p.flag_words;
    const long long flo = st.scan_lo & ~1023ll;
    st.total += p.T;
    const long long total = st.total, c0 = st.chunk_pos, scan_end = total - cp;
    auto flag = [&](long long n) -> bool { return n < total && ((flags[(n - flo) >> 5] >> ((n - flo) & 31)) & 1u); };
    SyncSeg* segs = p.segs + (size_t)s * p.seg_cap;              // [0] = the segment carried in, then this call's raw triggers
    int nseg = 0;
    for (int a = 0; a < st.nsegc; a++) segs[nseg++] = st.segc[a];
    long long* raw = p.raw + (size_t)s * p.trig_cap;
    int nraw = 0, err = 0;
    // ---- plateau_detector_fb(max_len = cp, threshold): skip the zero words, then walk the run
    long long i = st.scan_pos > c0 ? st.scan_pos : c0;
    while (i < scan_end) {
        i = sync_next_flag(flags, flo, i, scan_end, lane);       // never past scan_end: what the reference's i++ reaches
        if (i >= scan_end) break;
        int k = 1;
        for (; k < cp; k++) if (!flag(i + k)) break;
        if (k > 1) {
            const long long t = i + k / 2;
            if (nraw < p.trig_cap) raw[nraw++] = t; else err = 1;
            const float2 P = sync_p_at_warp(w, t, base, len, N, lane);
            const float f = (float)atan2((double)P.y, (double)P.x);
            double ph = sync_phase(segs, nseg, t - 1, N) + (-2.0 / N) * (double)f;
            ph = fmod(ph + 3.14159265358979323846, 6.28318530717958647692);
            if (ph < 0) ph += 6.28318530717958647692;
            if (nseg < p.seg_cap) { segs[nseg].t = t; segs[nseg].ph = ph - 3.14159265358979323846; segs[nseg].f = f; nseg++; } else err = 1;
        }
        i += k + cp;
    }
    st.scan_pos = i;
    long long c1 = i > scan_end ? i : scan_end;
    if (c1 < c0) c1 = c0;
    // ---- ofdm_adaptive_frame_detect_bb_impl::fix_sync (:64-156) on the chunk [c0, c1): the byte loop, event by event.
    // `out` of the reference is the raw trigger list plus the triggers the loop itself sets / clears ahead of its position.
    long long* fixed = p.fixed + (size_t)s * p.trig_cap;
    int nfixed = 0;
    {
        const int clen = (int)(c1 - c0);
        int last = -st.d_remainder, idx0 = st.d_remainder;      // index = idx0 + i while no event resets it
        bool found = false;
        int r = 0;                                              // next raw trigger
        int ipos = 0;                                           // loop position
        while (ipos < clen) {
            // next position where something happens: a raw trigger, or the missing-trigger rule on a zero
            while (r < nraw && (int)(raw[r] - c0) < ipos) r++;     // (cannot happen: triggers behind the loop position were consumed)
            const int tpos = r < nraw ? (int)(raw[r] - c0) : 0x7fffffff;
            int mpos = 0x7fffffff;
            if (st.d_in_sync) {
                // index > FS + 10  <=>  i > FS + 10 - idx0;  or index > FS + 1 at i == clen - 1
                long long a = (long long)FS + 11 - idx0;
                if (a < ipos) a = ipos;
                if (a < clen) mpos = (int)a;
                else if ((long long)(clen - 1) + idx0 > FS + 1 && clen - 1 >= ipos) mpos = clen - 1;
            }
            if (tpos >= clen && mpos >= clen) break;
            if (tpos <= mpos) {
                const int ii = tpos;
                r++;
                const int diff = ii - last - FS;
                if (st.d_acc_error) ++st.d_trigger_counter;
                const int inst = diff < 0 ? -diff : diff;
                if (inst == 1) st.d_acc_error += diff;
                st.d_error_count += inst != 0;
                if (inst <= 1) { if (st.d_synced < 3) ++st.d_synced; else st.d_in_sync = 1; } else { st.d_synced = 0; st.d_in_sync = 0; }
                int keep_at = ii;
                if (st.d_acc_error > 1 || (inst > 1 && inst < 10)) {
                    if (last + FS < clen && last + FS >= 0) {    // >= 0: the reference writes before its buffer here
                        last += FS;                              // out[i] = 0; out[last] = 1; i = last
                        keep_at = last;
                        st.d_acc_error = 0;
                        ++st.d_correction_count;
                        // a jump forward passes over raw triggers: they stay in the output as they came in, unvisited
                        while (r < nraw && (int)(raw[r] - c0) < keep_at) { if (nfixed < p.trig_cap) fixed[nfixed++] = raw[r]; else err = 1; r++; }
                        if (r < nraw && (int)(raw[r] - c0) == keep_at) r++;
                    } else last = ii;
                } else last = ii;
                if (nfixed < p.trig_cap) fixed[nfixed++] = c0 + keep_at; else err = 1;
                ipos = keep_at + 1;                              // ++i after `i = last` (or after the trigger itself)
                idx0 = -keep_at;                                 // index = 0 at the trigger: index = i - keep_at from here on
                st.d_remainder = clen - last;
                found = true;
                if (st.d_trigger_counter > 10) st.d_acc_error = 0;
            } else {
                const int ii = mpos;
                last += FS;
                if (last < 0) last = ii;                         // same guard
                if (nfixed < p.trig_cap) fixed[nfixed++] = c0 + last; else err = 1;
                ipos = last + 1;
                idx0 = -last;
                found = true;
                st.d_remainder = clen - last;
                ++st.d_missing_count;
                if (--st.d_synced < 3 - 5) { st.d_in_sync = 0; st.d_synced = 0; }
            }
        }
        if (!found) st.d_remainder += clen;
    }
    st.chunk_pos = c1;
    // `fixed` is in visiting order; triggers moved backwards can break the ascending order: sort (insertion, a handful of entries)
    // (one lane: the only place where the list is MOVED in memory; everywhere else the lanes write identical values)
    __syncwarp();
    if (lane == 0)
        for (int a = 1; a < nfixed; a++) { const long long v = fixed[a]; int b = a - 1; while (b >= 0 && fixed[b] > v) { fixed[b + 1] = fixed[b]; b--; } fixed[b + 1] = v; }
    __syncwarp();
    // ---- header_payload_demux: a trigger starts a frame unless one is being copied; frames complete when their last sample is in
    long long* starts = p.starts + (size_t)s * p.frames_cap;
    int nfr = 0, npend = 0;
    long long pend[DTL_SYNC_PEND];
    long long cand[DTL_SYNC_PEND];
    for (int a = 0; a < st.npend; a++) cand[a] = st.pend[a];
    int ncand = st.npend;
    auto place = [&](long long t) {
        if (t - D + FS <= total) { if (nfr < p.frames_cap) starts[nfr++] = t; else err = 2; }
        else if (npend < DTL_SYNC_PEND) pend[npend++] = t; else err = 1;
    };
    for (int a = 0; a < ncand; a++) place(cand[a]);
    for (int a = 0; a < nfixed; a++) { const long long t = fixed[a]; if (t >= st.busy_until) { st.busy_until = t + FS; place(t); } }
    st.npend = npend;
    for (int a = 0; a < npend; a++) st.pend[a] = pend[a];
    p.n_frames[s] = nfr;
    p.n_seg[s] = nseg;
    if (p.n_raw) p.n_raw[s] = nraw;
    if (p.n_fixed) p.n_fixed[s] = nfixed;
    // ---- what the next call still needs from the work buffer
    long long keep = st.scan_pos - (N + N / 2);
    if (c1 - D < keep) keep = c1 - D;
    if (npend && pend[0] - D < keep) keep = pend[0] - D;
    if (keep < base) keep = base;
    if (keep > base + len) keep = base + len;
    st.keep_off = (int)(keep - base);
    st.keep_len = (int)(base + len - keep);
    st.base_next = keep;
    st.scan_lo = st.scan_pos > c1 ? st.scan_pos : c1;
    // the last segment at or before the oldest position a later frame can ask for, and every later one
    const long long seg_keep = (npend ? pend[0] : c1) - 1;
    int ks = 0;
    for (int a = 0; a < nseg; a++) if (segs[a].t <= seg_keep) ks = a;
    st.nsegc = 0;
    for (int a = ks; a < nseg; a++) { if (st.nsegc < DTL_SYNC_SEGC) st.segc[st.nsegc++] = segs[a]; else err = 1; }
    st.base_cur = base; st.len_cur = len;
    if (lane == 0) {
        p.st[s] = st;
        if (err) atomicOr(p.err, err == 2 ? 64 : 1024);
    }
}

// Frame j of stream s: out[q] = x[t + q - D] * exp(j phase[t + q]).  Unused frame slots are zero-filled and flagged invalid.
__global__ void __launch_bounds__(256) sync_cut_kernel(SyncParams p)
{
    const int s = blockIdx.z, j = blockIdx.y;
    const int FS = p.FS, N = p.N, D = N + p.cp;
    float2* out = p.frames + ((size_t)s * p.frames_cap + j) * FS;
    const int nfr = p.n_frames[s];
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        p.valid[(size_t)s * p.frames_cap + j] = j < nfr;
        if (p.starts_out) p.starts_out[(size_t)s * p.frames_cap + j] = j < nfr ? p.starts[(size_t)s * p.frames_cap + j] : -1;
    }
    const int q0 = blockIdx.x * 1024;
    if (j >= nfr) {
        for (int q = q0 + threadIdx.x; q < min(q0 + 1024, FS); q += blockDim.x) out[q] = make_float2(0.f, 0.f);
        return;
    }
    const SyncState& st = p.st[s];                               // after sync_scan_kernel: base_prev / len of THIS call's buffer
    const long long base = st.base_cur, t = p.starts[(size_t)s * p.frames_cap + j];
    const int len = st.len_cur, nseg = p.n_seg[s];
    const float2* w = p.w_new + (size_t)s * p.w_cap;
    // the oscillator's segments that can matter for this CTA's samples, in shared memory (a handful per call)
    __shared__ SyncSeg sseg[16];
    __shared__ int s_n;
    // this thread's four delayed samples are requested before the segment search below: the two latencies overlap
    const int qe = min(q0 + 1024, FS);
    float2 yv[4];
#pragma unroll
    for (int u = 0; u < 4; u++) { const int q = q0 + threadIdx.x + 256 * u; yv[u] = q < qe ? ld_w(w, t + q - D, base, len) : make_float2(0.f, 0.f); }
    if (threadIdx.x < 32) {                                     // the list is sorted by t: counting replaces searching
        const SyncSeg* segs = p.segs + (size_t)s * p.seg_cap;
        const long long lo = t + q0, hi = t + min(q0 + 1024, FS) - 1;
        int c_lo = 0, c_hi = 0;
        for (int i = threadIdx.x; i < nseg; i += 32) { const long long ti = segs[i].t; c_lo += ti <= lo; c_hi += ti <= hi; }
        c_lo = __reduce_add_sync(0xffffffffu, c_lo); c_hi = __reduce_add_sync(0xffffffffu, c_hi);
        const int start = max(c_lo - 1, 0), n = c_hi - start;   // from the segment in force at the first sample to the last one that starts inside
        for (int i = threadIdx.x; i < min(n, 16); i += 32) sseg[i] = segs[start + i];
        if (threadIdx.x == 0) s_n = n > 16 ? -1 : n;            // more than 16 segments inside 1024 samples (a burst of false triggers): search the global list
    }
    __syncthreads();
    const int ns = s_n;
    const double w0 = -2.0 / N, inv2pi = 0.15915494309189533577, twopi = 6.28318530717958647692;
#pragma unroll
    for (int u = 0; u < 4; u++) {
        const int q = q0 + threadIdx.x + 256 * u;
        if (q >= qe) break;
        const float2 y = yv[u];
        const long long n = t + q;
        int k = -1;
        const SyncSeg* sl = sseg;
        int nsl = ns;
        if (ns < 0) { sl = p.segs + (size_t)s * p.seg_cap; nsl = nseg; }
        for (int i = 0; i < nsl; i++) { if (sl[i].t <= n) k = i; else break; }
        // phase in double as the oscillator accumulates it, reduced to [-pi, pi] in double; sine / cosine in fp32 (the samples are
        // fp32: 1e-6 absolute, the tolerance on the mixed samples is 1e-4)
        float c = 1.f, sf = 0.f;
        if (k >= 0) {
            const double ph = sl[k].ph + (double)(n - sl[k].t) * w0 * (double)sl[k].f;
            const float r = (float)(ph - twopi * rint(ph * inv2pi));
            __sincosf(r, &sf, &c);
        }
        out[q] = make_float2(__fmaf_rn(y.x, c, -__fmul_rn(y.y, sf)), __fmaf_rn(y.x, sf, __fmul_rn(y.y, c)));
    }
}

__global__ void sync_reset_kernel(SyncState* st, int n)
{
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    SyncState z;
    memset(&z, 0, sizeof z);
    st[s] = z;
}

cudaError_t launch_sync_reset(cudaStream_t st, SyncState* state, int n)
{
    if (n <= 0) return cudaSuccess;
    sync_reset_kernel<<<(n + 255) / 256, 256, 0, st>>>(state, n);
    return cudaGetLastError();
}

size_t sync_metric_smem(int N)
{
    const int T = N > 1024 ? N : 1024, nc = T + N, ncp = nc + (nc >> 5) + 1;      // PAD of sync_metric_kernel
    const int nx = T + N + N / 2;
    return (size_t)(nx > ncp ? nx : ncp) * 8 + 2 * (size_t)ncp * 8 + 3 * (size_t)ncp * 4 + 16;
}

cudaError_t launch_sync(cudaStream_t st, const SyncParams& p, int* launches)
{
    if (p.n_streams <= 0 || p.T <= 0) return cudaSuccess;
    const int max_len = p.w_cap;
    sync_append_kernel<<<dim3((unsigned)((max_len + 2047) / 2048 < 64 ? (max_len + 2047) / 2048 : 64), p.n_streams), 256, 0, st>>>(p);
    const int tiles = (int)((p.T + p.keep_cap + 1023) / 1024 + 1);
    if (p.N == 64) sync_metric_kernel<1024, 64><<<dim3(tiles, p.n_streams), 256, sync_metric_smem(p.N), st>>>(p);
    else sync_metric_kernel<1024, 0><<<dim3(tiles, p.n_streams), 256, sync_metric_smem(p.N), st>>>(p);
    sync_scan_kernel<<<(p.n_streams + 3) / 4, 128, 0, st>>>(p);
    sync_cut_kernel<<<dim3((p.FS + 1023) / 1024, p.frames_cap, p.n_streams), 256, 0, st>>>(p);
    if (launches) *launches += 4;
    return cudaGetLastError();
}

cudaError_t init_kernels_sync()
{
    cudaError_t e = set_max_dyn_smem(sync_metric_kernel<1024, 0>);
    if (e == cudaSuccess) e = set_max_dyn_smem(sync_metric_kernel<1024, 64>);
    return e;
}

} // namespace dtl
