// oracle/src/sync.cc -- TEST INFRASTRUCTURE ONLY (see oracle.hpp header).
// The synchronisation front end of python/dtl/ofdm_receiver.py:70-97, one continuous IQ stream at a time:
//   digital.ofdm_sync_sc_cfb(fft_len, cp_len, threshold)      Schmidl & Cox timing metric + fine frequency estimate
//   blocks.delay(fft_len + cp_len), analog.frequency_modulator_fc(-2/fft_len), blocks.multiply_cc   (fine CFO removed)
//   dtl.ofdm_adaptive_frame_detect_bb(frame samples)           trigger regulariser, lib/dtl/ofdm_adaptive_frame_detect_bb_impl.cc:64-156
//   digital.header_payload_demux                               frames cut at the (regularised) triggers
// gr-digital / gr-blocks / gr-analog are not under /root/reference: ofdm_sync_sc_cfb, plateau_detector_fb, sample_and_hold_ff
// and frequency_modulator_fc are RECALLED from GNU Radio 3.10 (parity unpinned at that boundary); fix_sync is restated line
// by line from the reference.  What the scheduler's chunking decides in the reference (fix_sync's `len`) is decided here by
// the caller's chunks: one push() = one work() call of the regulariser over the trigger positions the detector finished.
#include "txrx.hpp"
#include <cmath>
#include <cstring>

namespace orc {

stream_sync::stream_sync(int fft_len, int cp_len, int frame_samples, float thr)
    : N(fft_len), cp(cp_len), FS(frame_samples), threshold(thr) {}

cf32 stream_sync::x_at(long long n) const
{
    if (n < base || n >= base + (long long)buf.size()) return cf32(0, 0);      // before the stream began: the delay lines hold zeros
    return buf[(size_t)(n - base)];
}

// ofdm_sync_sc_cfb (RECALLED): delay_corr = x[n] * conj(x[n - N/2]); delay_ma = fir_filter_ccf(taps -1 x N/2) (use_even_carriers
// false); normalizer_ma = fir_filter_fff(taps 0.5 x N) on |x|^2; metric = |P|^2 / E^2.  Every fp32 operation explicit.
// Summation order of the two moving sums (GNU Radio's FIR kernels sum in whatever order VOLK picks for the machine; this is
// the order of the CUDA path): the window ending at n = the SUFFIX of the previous aligned block (accumulated descending from
// the block's last element) + the PREFIX of n's own block (accumulated ascending), blocks of the window's length aligned to
// the stream's sample index -- the same floats whatever the caller's chunking.
void stream_sync::metric_at(long long n, float* m, cf32* p) const
{
    auto c_at = [&](long long k) {
        const cf32 a = x_at(k), b = x_at(k - N / 2);
        return cf32(-fmaf(a.real(), b.real(), a.imag() * b.imag()), -fmaf(a.imag(), b.real(), -(a.real() * b.imag())));
    };
    auto e_at = [&](long long k) { const cf32 a = x_at(k); return 0.5f * fmaf(a.real(), a.real(), a.imag() * a.imag()); };
    const int W1 = N / 2, W2 = N;
    const long long b1 = n - n % W1, b2 = n - n % W2;
    float pr, pi, e;
    {
        cf32 pre = c_at(b1);
        float ar = pre.real(), ai = pre.imag();
        for (long long k = b1 + 1; k <= n; k++) { const cf32 c = c_at(k); ar = ar + c.real(); ai = ai + c.imag(); }
        if (n != b1 + W1 - 1) {
            const cf32 s0 = c_at(b1 - 1);
            float sr = s0.real(), si = s0.imag();
            for (long long k = b1 - 2; k >= n - W1 + 1; k--) { const cf32 c = c_at(k); sr = sr + c.real(); si = si + c.imag(); }
            ar = ar + sr; ai = ai + si;
        }
        pr = ar; pi = ai;
    }
    {
        float a = e_at(b2);
        for (long long k = b2 + 1; k <= n; k++) a = a + e_at(k);
        if (n != b2 + W2 - 1) {
            float sfx = e_at(b2 - 1);
            for (long long k = b2 - 2; k >= n - W2 + 1; k--) sfx = sfx + e_at(k);
            a = a + sfx;
        }
        e = a;
    }
    const float num = fmaf(pr, pr, pi * pi), den = e * e;
    *m = num / den;                                   // 0/0 = NaN: never >= threshold
    *p = cf32(pr, pi);
}

double stream_sync::phase_at(long long n) const
{
    // frequency_modulator_fc: phase[n] = phase[n-1] + sensitivity * f[n], f = sample_and_hold of arg(P) at the raw triggers
    const seg* s = nullptr;
    for (const seg& g : segs) if (g.t <= n) s = &g; else break;
    if (!s) return 0.0;
    return s->ph + (double)(n - s->t) * (-2.0 / N) * (double)s->f;
}

// lib/dtl/ofdm_adaptive_frame_detect_bb_impl.cc:64-156 on the chunk [c0, c1) whose raw triggers are `trig` (global positions,
// ascending).  Event-driven walk of the byte loop: between triggers only `index` moves.  Returns the regularised triggers.
std::vector<long long> stream_sync::fix_sync(long long c0, long long c1, const std::vector<long long>& trig)
{
    const int len = (int)(c1 - c0);
    std::vector<char> out((size_t)std::max(len, 0), 0);
    for (long long t : trig) out[(size_t)(t - c0)] = 1;
    int last_trigger_index = -d_remainder;
    bool trigger_found = false;
    for (int i = 0, index = d_remainder; i < len; ++i, ++index) {
        if (out[i]) {
            int frame_len_detected = i - last_trigger_index;
            int diff = frame_len_detected - FS;
            index = 0;
            if (d_acc_error) ++d_trigger_counter;
            int inst_error = std::abs(diff);
            if (inst_error == 1) d_acc_error += diff;
            d_error_count += (inst_error != 0);
            if (inst_error == 0 || inst_error == 1) {
                if (d_synced < 3) ++d_synced; else d_in_sync = true;
            } else { d_synced = 0; d_in_sync = false; }
            if (d_acc_error > 1 || (inst_error > 1 && inst_error < 10)) {
                if (last_trigger_index + FS < len && last_trigger_index + FS >= 0) {      // >= 0: the reference writes before its buffer here
                    out[i] = 0;
                    last_trigger_index += FS;
                    out[last_trigger_index] = 1;
                    i = last_trigger_index;
                    d_acc_error = 0;
                    ++d_correction_count;
                } else last_trigger_index = i;
            } else last_trigger_index = i;
            d_remainder = len - last_trigger_index;
            trigger_found = true;
            if (d_trigger_counter > 10) d_acc_error = 0;
        } else if (d_in_sync && (index > FS + 10 || (index > FS + 1 && i == len - 1))) {
            last_trigger_index += FS;
            if (last_trigger_index < 0) last_trigger_index = i;                              // same guard
            out[last_trigger_index] = 1;
            i = last_trigger_index;
            index = 0;
            trigger_found = true;
            d_remainder = len - last_trigger_index;
            ++d_missing_count;
            if (--d_synced < 3 - 5) { d_in_sync = false; d_synced = 0; }
        }
    }
    if (!trigger_found) d_remainder += len;
    std::vector<long long> res;
    for (int i = 0; i < len; i++) if (out[i]) res.push_back(c0 + i);
    return res;
}

// One chunk of T samples.  Frames completed by it are appended to `frames` (FS samples each) and their start positions
// (delayed-signal coordinates) to `starts`; raw / regularised trigger positions to raw_out / fixed_out when given.
void stream_sync::push(const cf32* x, int T, std::vector<cf32>& frames, std::vector<long long>& starts,
                       std::vector<long long>* raw_out, std::vector<long long>* fixed_out)
{
    buf.insert(buf.end(), x, x + T);
    total += T;
    const int D = N + cp;
    // ---- plateau_detector_fb(max_len = cp, threshold) (RECALLED) over the positions whose look-ahead is complete
    const long long c0 = chunk_pos, scan_end = total - cp;
    std::vector<long long> raw;
    long long i = std::max(scan_pos, c0);
    while (i < scan_end) {
        float m; cf32 p;
        metric_at(i, &m, &p);
        if (m >= threshold) {
            int k = 1;
            for (; k < cp; k++) { float mk; cf32 pk; metric_at(i + k, &mk, &pk); if (!(mk >= threshold)) break; }
            if (k > 1) {
                const long long t = i + k / 2;
                raw.push_back(t);
                float mt; cf32 pt;
                metric_at(t, &mt, &pt);
                // complex_to_arg + sample_and_hold_ff + frequency_modulator_fc(-2/N): a new phase segment starts at t
                const float f = (float)std::atan2((double)pt.imag(), (double)pt.real());
                double ph = phase_at(t - 1) + (-2.0 / N) * (double)f;
                ph = std::fmod(ph + M_PI, 2.0 * M_PI);
                if (ph < 0) ph += 2.0 * M_PI;
                segs.push_back({ t, ph - M_PI, f });
            }
            i += k + cp;
        } else i++;
    }
    scan_pos = i;
    const long long c1 = std::max(i, std::max(scan_end, c0));
    // ---- trigger regulariser on this chunk, then header_payload_demux: a trigger starts a frame unless one is being copied
    const std::vector<long long> fixed = fix_sync(c0, c1, raw);
    chunk_pos = c1;
    for (long long t : fixed) if (t >= busy_until) { pending.push_back(t); busy_until = t + FS; }
    if (raw_out) raw_out->insert(raw_out->end(), raw.begin(), raw.end());
    if (fixed_out) fixed_out->insert(fixed_out->end(), fixed.begin(), fixed.end());
    // ---- frames whose samples have all arrived: delay(N + cp), mixer
    size_t done = 0;
    for (; done < pending.size(); done++) {
        const long long t = pending[done];
        if (t - D + FS > total) break;
        starts.push_back(t);
        for (int q = 0; q < FS; q++) {
            const cf32 y = x_at(t + q - D);
            const double ph = phase_at(t + q);
            const float c = (float)std::cos(ph), s = (float)std::sin(ph);
            frames.push_back(cf32(fmaf(y.real(), c, -(y.imag() * s)), fmaf(y.real(), s, y.imag() * c)));
        }
    }
    pending.erase(pending.begin(), pending.begin() + done);
    // ---- what the next chunk still needs: metric history, the delay line of frames not cut yet
    long long keep = std::min(scan_pos - (N + N / 2), c1 - D);
    if (!pending.empty()) keep = std::min(keep, pending.front() - D);
    keep = std::max(keep, base);
    if (keep > base) { buf.erase(buf.begin(), buf.begin() + (size_t)(keep - base)); base = keep; }
    const long long seg_keep = (pending.empty() ? c1 : pending.front()) - 1;
    while (segs.size() > 1 && segs[1].t <= seg_keep) segs.erase(segs.begin());
}

} // namespace orc
