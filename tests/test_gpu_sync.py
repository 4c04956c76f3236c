"""Synchronisation front end on the GPU: dtl_b200_rx_stream on CONTINUOUS IQ.

Oracle: oracle/src/sync.cc (gr-digital ofdm_sync_sc_cfb / plateau_detector_fb / sample_and_hold / frequency_modulator_fc restated
from GNU Radio 3.10, parity unpinned at that boundary; ofdm_adaptive_frame_detect_bb line by line from
lib/dtl/ofdm_adaptive_frame_detect_bb_impl.cc:64-156), fed with the same chunks.
Bar: raw and regularised trigger positions and frame starts bit-exact; the frames handed to the demux 1e-4 relative
(phase accumulation in double on both sides, sincos of different libraries); decoded payload bytes, TB records and header tags
equal to dtl_b200_rx_iq / the oracle on the same frames cut at their boundaries by the caller."""
import numpy as np
import pytest

import pyoracle as orc
from common import ALIST_300, make_oracle_cfg, make_receiver

pytestmark = pytest.mark.gpu


def _streams(ocfg, S, F, cnst, snr, seed, eps_max=0.3, drop=(), gaps=False):
    """S continuous streams: noise lead-in, F frames back to back (the reference transmitter's pacing; gaps=True: random gaps
    after some frames, which the in-sync regulariser answers with triggers at the frame period), per-stream fractional CFO"""
    t = ocfg.tx_batch(S, F, seed=seed, cnst=cnst, fec_id=1, snr_db=snr)
    fs, N = ocfg.frame_samples, ocfg.fft_len
    rng = np.random.default_rng(seed)
    noise = lambda n: (0.02 * (rng.standard_normal(n) + 1j * rng.standard_normal(n))).astype(np.complex64)
    xs, starts = [], []
    for s in range(S):
        parts, st, pos = [noise(int(rng.integers(100, 900)))], [], 0
        pos = parts[0].size
        for j in range(F):
            if (s, j) in drop:
                parts.append(noise(fs))
            else:
                st.append(pos)
                parts.append(t["iq"][s * F + j])
            pos += fs
            if gaps and rng.random() < 0.3:
                g = noise(int(rng.integers(1, 500)))
                parts.append(g)
                pos += g.size
        parts.append(noise(4000))
        x = np.concatenate(parts)
        eps = rng.uniform(-eps_max, eps_max)
        xs.append((x * np.exp(2j * np.pi * eps * np.arange(x.size) / N)).astype(np.complex64))
        starts.append(st)
    T = max(x.size for x in xs)
    X = np.zeros((S, T), dtype=np.complex64)
    for s, x in enumerate(xs):
        X[s, :x.size] = x
        X[s, x.size:] = noise(T - x.size)
    return t, X, starts


@pytest.mark.parametrize("chunk", [6000, 25000])
def test_rx_stream_matches_the_oracle_front_end(chunk):
    ocfg = make_oracle_cfg(alists=(ALIST_300,), initial_mcs=1)
    rx = make_receiver(alists=(ALIST_300,), initial_mcs=1)
    S, F = 6, 9
    t, X, true_starts = _streams(ocfg, S, F, orc.QPSK, 24.0, seed=21)
    fs, D = ocfg.frame_samples, ocfg.fft_len + ocfg.cp_len
    syncs = [orc.StreamSync(ocfg, 0.95) for _ in range(S)]
    rx.reset(S)
    o_frames = [[] for _ in range(S)]
    o_starts = [[] for _ in range(S)]
    g_frames = [[] for _ in range(S)]
    g_bytes = [[] for _ in range(S)]
    g_tbs = [[] for _ in range(S)]
    g_meta = [[] for _ in range(S)]
    Fc = chunk // fs + 3
    for c0 in range(0, X.shape[1], chunk):
        xc = np.ascontiguousarray(X[:, c0:c0 + chunk])
        if xc.shape[1] < chunk:                                     # the last call may be shorter
            xc = np.ascontiguousarray(xc)
        got = rx.rx_stream(xc, S, Fc, want_frames=True)
        for s in range(S):
            r = syncs[s].push(xc[s])
            n = int(got["n_frames"][s])
            assert n == r["starts"].size, "stream %d: %d frames, oracle %d" % (s, n, r["starts"].size)
            assert np.array_equal(got["frame_start"][s * Fc:s * Fc + n], r["starts"]), "frame starts (stream %d)" % s
            assert (got["frame_start"][s * Fc + n:(s + 1) * Fc] == -1).all()
            for j in range(n):
                gf, of = got["frames"][s * Fc + j], r["frames"][j]
                err = np.abs(gf - of).max() / max(np.abs(of).max(), 1e-6)
                assert err < 1e-4, "mixed frame samples differ: %g" % err
                o_frames[s].append(of)
                o_starts[s].append(int(r["starts"][j]))
                g_frames[s].append(gf)
                g_meta[s].append(got["meta"][s * Fc + j].copy())
            assert (got["meta"]["reserved"][s * Fc + n:(s + 1) * Fc] == 1).all(), "empty slots must read as dropped frames"
            g_bytes[s].append(got["bytes"][s, :got["nbytes"][s]].copy())
            g_tbs[s].append(got["tbs"][s, :got["ntbs"][s]].copy())
    # every transmitted frame was found, at the position the Schmidl & Cox plateau puts it (middle of the cyclic prefix)
    for s in range(S):
        assert len(o_starts[s]) >= F                         # the in-sync regulariser keeps triggering at the frame period after the
        off = np.array(o_starts[s][:F]) - D - np.array(true_starts[s])     # last frame: frames of noise, dropped by their header CRC
        assert (np.abs(off + ocfg.cp_len // 2) <= 3).all(), off
    # the payload that comes out is what the oracle decodes from ITS frames, and what was sent
    for s in range(S):
        fr = np.array(o_frames[s])
        nf = fr.shape[0]
        h = ocfg.rx_headers(fr, 1, nf)
        assert h["ok"][:F].all()
        ref = ocfg.rx_batch(fr, h["taps"], h["meta"], 1, nf)
        gb = np.concatenate(g_bytes[s])
        assert np.array_equal(gb, ref["bytes"][0]), "stream %d: decoded payload differs from the oracle" % s
        assert gb.size > 0 and np.array_equal(gb, t["payload"][s][:gb.size])
        gt = np.concatenate(g_tbs[s])
        for k in ("tb_no", "crc_ok", "nbytes"):
            assert np.array_equal(gt[k], ref["tbs"][0][k]), k
        gm = np.array(g_meta[s])
        for k in ("constellation", "frame_no", "payload_len", "packet_len", "fec_id", "tb_no", "tb_offset", "tb_payload", "carrier_offset"):
            assert np.array_equal(gm[k], h["meta"][k]), k
    rx.close()


def test_rx_stream_irregular_gaps_triggers_and_frames_equal_the_oracle():
    """random gaps between frames: the regulariser moves / inserts triggers (frames are then cut where the REFERENCE would cut
    them, which is not always where they are); positions and frame samples equal the oracle's"""
    ocfg = make_oracle_cfg(alists=(ALIST_300,), initial_mcs=1)
    rx = make_receiver(alists=(ALIST_300,), initial_mcs=1)
    S, F = 8, 10
    t, X, _ = _streams(ocfg, S, F, orc.QPSK, 24.0, seed=45, gaps=True)
    syncs = [orc.StreamSync(ocfg, 0.95) for _ in range(S)]
    rx.reset(S)
    chunk, Fc = 7000, 8
    moved = 0
    for c0 in range(0, X.shape[1], chunk):
        xc = np.ascontiguousarray(X[:, c0:c0 + chunk])
        got = rx.rx_stream(xc, S, Fc, want_frames=True)
        for s in range(S):
            r = syncs[s].push(xc[s])
            n = int(got["n_frames"][s])
            assert n == r["starts"].size and np.array_equal(got["frame_start"][s * Fc:s * Fc + n], r["starts"]), (s, c0)
            for j in range(n):
                of = r["frames"][j]
                assert np.abs(got["frames"][s * Fc + j] - of).max() / max(np.abs(of).max(), 1e-6) < 1e-4
            moved += int(len(set(r["fixed"]) ^ set(r["raw"])) > 0)
    assert moved > 0, "the regulariser never changed a trigger: vacuous"
    rx.close()


def test_rx_stream_lost_frames_and_regulariser():
    """frames replaced by noise (no trigger there): once in sync the regulariser inserts the missing trigger at the frame
    period (frame_detect_bb_impl.cc:123-137), the header of that slot fails its CRC and the frame is dropped downstream;
    trigger lists equal the oracle's."""
    ocfg = make_oracle_cfg(alists=(ALIST_300,), initial_mcs=1)
    rx = make_receiver(alists=(ALIST_300,), initial_mcs=1)
    S, F = 4, 12
    t, X, _ = _streams(ocfg, S, F, orc.QPSK, 26.0, seed=33, drop={(0, 6), (1, 8), (2, 5), (2, 6)})
    syncs = [orc.StreamSync(ocfg, 0.95) for _ in range(S)]
    rx.reset(S)
    chunk, Fc = 9000, 8
    for c0 in range(0, X.shape[1], chunk):
        xc = np.ascontiguousarray(X[:, c0:c0 + chunk])
        got = rx.rx_stream(xc, S, Fc)
        for s in range(S):
            r = syncs[s].push(xc[s])
            n = int(got["n_frames"][s])
            assert n == r["starts"].size and np.array_equal(got["frame_start"][s * Fc:s * Fc + n], r["starts"]), (s, c0)
    rx.close()
