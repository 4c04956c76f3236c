"""Parity tests proper: the CUDA path (through the C ABI, via the dtl_b200 mirror) against the CPU oracle on
the same seeded inputs.  Bars: bit-exact for hard decisions, channel taps, feedback MCS, packed bytes, CRC flags,
decoded payload bits and LDPC iteration counts; equalised symbols within 1e-6 relative (the recurrence is written
as explicit IEEE operations on both sides, so they are normally identical); SNR/noise within 1e-9 relative
(libm vs CUDA log10/pow); LLRs and FFT outputs within 1e-4 relative (ex2/lg2.approx, different FFT factorisation)."""
import numpy as np
import pytest

import pyoracle as orc
from common import ALIST_100_23, ALIST_100_27, ALIST_300, assert_llr_close, make_oracle_cfg, make_receiver

pytestmark = pytest.mark.gpu

RT_SYM, RT_SNR, RT_FFT = 1e-6, 1e-9, 1e-4


def _data_mask(ocfg):
    m = np.zeros(ocfg.fft_len, dtype=bool)
    for c in ocfg.occupied:
        m[(c + ocfg.fft_len // 2) % ocfg.fft_len if c >= 0 else (c + ocfg.fft_len + ocfg.fft_len // 2) % ocfg.fft_len] = True
    return m


def _rel(a, b):
    return np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-30)


# ---- stage 1 / 5b -----------------------------------------------------------------------------------------
@pytest.mark.parametrize("n", [8, 16, 64, 128, 1024])
def test_fft_shifted_both_directions(n):
    rng = np.random.default_rng(n)
    x = (rng.standard_normal((37, n)) + 1j * rng.standard_normal((37, n))).astype(np.complex64)
    rx = make_receiver(fft_len=64) if n == 64 else None
    if rx is None:
        import dtl_b200
        cfg = dtl_b200.ofdm_adaptive_rx_config(fft_len=n, cp_len=n // 4, occupied_carriers=([1, 2, -1, -2],),
                                               pilot_carriers=((3,), (3,)), frame_length=4)
        rx = dtl_b200.ofdm_receiver(cfg)
    for fwd in (True, False):
        got = rx.fft(x, forward=fwd)
        ref = orc.fft_shifted(x, n, forward=fwd)
        assert _rel(got, ref.reshape(got.shape)) < RT_FFT
    # numpy cross-check of the convention: forward = fftshift(fft(x))
    assert _rel(rx.fft(x, True), np.fft.fftshift(np.fft.fft(x, axis=1), axes=1)) < RT_FFT
    rx.close()


@pytest.mark.parametrize("fec", [False, True])
def test_demod_cp_strip_fft(fec):
    alists = (ALIST_300,) if fec else ()
    ocfg = make_oracle_cfg(alists=alists)
    rx = make_receiver(alists=alists)
    t = ocfg.tx_batch(3, 2, seed=11, cnst=orc.QPSK, fec_id=1 if fec else 0, snr_db=25.0)
    got = rx.demod(t["iq"])
    ref = ocfg.rx_demod(t["iq"]).reshape(got.shape)
    assert _rel(got, ref) < RT_FFT
    rx.close()


# ---- stages 2+3, frequency-domain input so both sides see identical symbols -----------------------------------
@pytest.mark.parametrize("cnst,snr", [(1, 8.0), (2, 14.0), (3, 19.0), (4, 24.0), (5, 28.0), (6, 34.0)])
def test_equalizer_demapper_no_fec(cnst, snr):
    cn = (1, 2, 3, 4, 5, 6)
    ocfg = make_oracle_cfg(cnsts=cn)
    rx = make_receiver(cnsts=cn)
    S, F = 5, 7
    t = ocfg.tx_batch(S, F, seed=100 + cnst, cnst=cnst, snr_db=snr)
    freq = ocfg.rx_demod(t["iq"])
    ref = ocfg.rx_batch(freq, t["taps"], t["meta"], S, F, freq_domain=True)
    got = rx.equalize(freq.reshape(S * F, ocfg.frame_len, 64), t["taps"], t["meta"], S, F,
                      want=("hard", "soft", "taps", "res", "bytes"))
    mask = _data_mask(ocfg)
    L = ocfg.frame_len
    gh, rh = got["hard"].reshape(S * F, L, 64), ref["hard"].reshape(S * F, L, 64)
    active = np.abs(rh).sum(axis=(0, 1)) > 0
    assert np.array_equal(np.ascontiguousarray(gh[:, :, active]).view(np.uint32), np.ascontiguousarray(rh[:, :, active]).view(np.uint32)), "hard decisions differ"
    gs, rs = got["soft"].reshape(S * F, L, 64)[:, :, mask], ref["soft"].reshape(S * F, L, 64)[:, :, mask]
    assert _rel(gs, rs) < RT_SYM
    assert np.allclose(got["res"]["snr_db"], ref["res"]["snr_db"], rtol=RT_SNR, atol=1e-9)
    assert np.allclose(got["res"]["noise"], ref["res"]["noise"], rtol=RT_SNR)
    for k in ("fb_cnst", "fb_fec", "emitted", "lost_frames"):
        assert np.array_equal(got["res"][k], ref["res"][k]), k
    assert np.array_equal(got["crc_ok"], ref["frame_crc"])
    for s in range(S):
        cat = np.concatenate([got["bytes"][s * F + j, :got["nbytes"][s * F + j]] for j in range(F)])
        assert np.array_equal(cat, ref["bytes"][s])
        if np.all(ref["frame_crc"][s * F:(s + 1) * F] == 1):
            assert np.array_equal(cat, t["payload"][s])      # CRC-passing frames carry what was sent
    rx.close()


@pytest.mark.parametrize("cnst,snr", [(1, 6.0), (2, 12.0), (3, 17.0), (4, 20.0), (5, 26.0), (6, 32.0)])
def test_equalizer_soft_demapper_llr(cnst, snr):
    cn = (1, 2, 3, 4, 5, 6)
    ocfg = make_oracle_cfg(cnsts=cn, alists=(ALIST_300,))
    rx = make_receiver(cnsts=cn, alists=(ALIST_300,))
    S, F = 4, 5
    t = ocfg.tx_batch(S, F, seed=200 + cnst, cnst=cnst, fec_id=1, snr_db=snr)
    freq = ocfg.rx_demod(t["iq"])
    ref = ocfg.rx_batch(freq, t["taps"], t["meta"], S, F, freq_domain=True)
    got = rx.equalize(freq.reshape(S * F, ocfg.frame_len, 64), t["taps"], t["meta"], S, F, want=("res", "llr", "taps"))
    assert np.array_equal(got["n_llr"], ref["n_llr"])
    for f in range(S * F):
        n = ref["n_llr"][f]
        assert n > 0
        assert_llr_close(got["llr"][f, :n], ref["llr"][f, :n])
    rx.close()


def test_equalizer_carrier_offset_and_taps_out():
    ocfg = make_oracle_cfg()
    rx = make_receiver()
    S, F = 3, 4
    t = ocfg.tx_batch(S, F, seed=7, cnst=orc.QPSK, snr_db=20.0)
    freq = ocfg.rx_demod(t["iq"]).reshape(S * F, ocfg.frame_len, 64)
    meta = t["meta"].copy()
    offs = np.array([0, 1, -1, 2, -3, 0, 5, -2, 1, 0, -1, 3], dtype=np.int32)
    meta["carrier_offset"] = offs
    # emulate the un-corrected integer CFO the frame equaliser undoes: shift the flattened frame back
    shifted = np.zeros_like(freq)
    for f in range(S * F):
        shifted[f] = np.roll(freq[f].reshape(-1), offs[f]).reshape(freq[f].shape)
    got = rx.equalize(shifted, t["taps"], meta, S, F, want=("hard", "soft", "taps", "res"))
    L = ocfg.frame_len
    for f in range(S * F):
        feq = orc.FrameEqualizer(ocfg, symbols_skipped=ocfg.header_syms)
        emitted, hard, soft, res, taps_out = feq.work(shifted[f], t["taps"][f], meta[f])
        active = np.abs(hard).sum(axis=0) > 0
        assert np.array_equal(np.ascontiguousarray(got["hard"][f][:, active]).view(np.uint32), np.ascontiguousarray(hard.reshape(L, 64)[:, active]).view(np.uint32))
        assert np.array_equal(got["taps"][f].view(np.uint32), taps_out.view(np.uint32)), "channel state differs"
        assert np.isclose(got["res"]["snr_db"][f], res["snr_db"], rtol=RT_SNR)
    rx.close()


def test_missing_constellation_tag_is_an_error():
    import dtl_b200
    ocfg = make_oracle_cfg()
    rx = make_receiver()
    t = ocfg.tx_batch(1, 2, seed=3, cnst=orc.BPSK, snr_db=20.0)
    freq = ocfg.rx_demod(t["iq"]).reshape(2, ocfg.frame_len, 64)
    meta = t["meta"].copy()
    meta["constellation"][1] = 0
    with pytest.raises(dtl_b200.DtlError) as ei:
        rx.equalize(freq, t["taps"], meta, 1, 2)
    assert ei.value.status == dtl_b200._capi.ERR_MISSING_TAG
    rx.close()


# ---- stand-alone stage-3 / 5a blocks ----------------------------------------------------------------------------
@pytest.mark.parametrize("cnst", [1, 2, 3, 4, 5, 6])
def test_standalone_map_hard_soft(cnst):
    rx = make_receiver(cnsts=(1, 2, 3, 4, 5, 6))
    rng = np.random.default_rng(cnst)
    bps = orc.bps(cnst)
    chunks = rng.integers(0, 1 << bps, size=(6, 96), dtype=np.uint8)
    meta = np.zeros(6, dtype=orc.frame_meta_dtype)
    meta["constellation"] = cnst
    meta["packet_len"] = [96, 96, 50, 1, 96, 77]
    syms = rx.chunks_to_symbols(chunks, meta)
    ref = orc.map_chunks(cnst, chunks.reshape(-1)).reshape(6, 96)
    for f in range(6):
        n = meta["packet_len"][f]
        assert np.array_equal(syms[f, :n].view(np.uint32), ref[f, :n].view(np.uint32))
    noisy = (ref + 0.05 * (rng.standard_normal(ref.shape) + 1j * rng.standard_normal(ref.shape))).astype(np.complex64)
    hard = rx.hard_demap(noisy, meta)
    noise = np.full(6, 0.01)
    noise[3] = -1.0                                    # no noise tag: sigma = 0.001 (soft_cf_impl.cc:89)
    llr = rx.soft_demap(noisy, meta, noise)
    for f in range(6):
        n = meta["packet_len"][f]
        assert np.array_equal(hard[f, :n], orc.decide(cnst, noisy[f, :n]))
        sigma = np.float32(np.sqrt(noise[f])) if noise[f] >= 0 else np.float32(0.001)
        assert_llr_close(llr[f, :n * bps], orc.soft_dec(cnst, noisy[f, :n], sigma))
    rx.close()


# ---- stage 4 ------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("alist", [ALIST_300, ALIST_100_23, ALIST_100_27])
def test_ldpc_decode_matches_awgn_bp(alist):
    ocfg = make_oracle_cfg(alists=(alist,))
    rx = make_receiver(alists=(alist,))
    info = ocfg.code_info(1)
    assert rx.code_info(1)["K"] == info["K"] and np.array_equal(rx.code_info(1)["permute"], info["permute"])
    rng = np.random.default_rng(5)
    count = 300
    data = rng.integers(0, 2, size=(count, info["K"]), dtype=np.uint8)
    cw = ocfg.ldpc_encode(1, data)
    sig = np.repeat(np.array([0.55, 0.75, 0.95, 1.2]), count // 4)[:, None]
    y = (2.0 * cw - 1.0) + sig * rng.standard_normal(cw.shape)
    llr = (2.0 * y / sig ** 2).astype(np.float32)          # + = bit 1 (soft_cf convention)
    ref_bits, ref_nit = ocfg.ldpc_decode(1, llr)
    bits, nit = rx.ldpc_decode(1, llr)
    assert np.array_equal(nit, ref_nit), "iteration counts differ"
    assert np.array_equal(bits, ref_bits)
    assert 0 < (ref_nit == 15).sum() < count                # the set covers converged and non-converged words
    rx.close()


def test_ldpc_layered_decodes_what_reference_decodes():
    import dtl_b200
    ocfg = make_oracle_cfg(alists=(ALIST_300,))
    rx = make_receiver(alists=(ALIST_300,), ldpc_mode=dtl_b200.LDPC_LAYERED)
    info = ocfg.code_info(1)
    rng = np.random.default_rng(9)
    data = rng.integers(0, 2, size=(256, info["K"]), dtype=np.uint8)
    cw = ocfg.ldpc_encode(1, data)
    sig = 0.7
    llr = (2.0 * ((2.0 * cw - 1.0) + sig * rng.standard_normal(cw.shape)) / sig ** 2).astype(np.float32)
    ref_bits, ref_nit = ocfg.ldpc_decode(1, llr)
    bits, nit = rx.ldpc_decode(1, llr)
    ok_ref = (ref_bits == data).all(axis=1)
    ok = (bits == data).all(axis=1)
    assert ok_ref.sum() > 200
    assert ok.sum() >= ok_ref.sum() - 3                      # same codewords up to a few near-threshold words
    rx.close()


@pytest.mark.parametrize("cnst,snr", [(2, 17.0), (4, 20.0), (1, 6.0), (3, 15.0)])
def test_fec_chain_bytes_and_tb_records(cnst, snr):
    """config 2: QPSK / 16QAM (+ BPSK, 8PSK) with n=300 k=152; SNRs low enough that BP iterates."""
    ocfg = make_oracle_cfg(alists=(ALIST_300,))
    rx = make_receiver(alists=(ALIST_300,))
    S, F = 6, 9
    t = ocfg.tx_batch(S, F, seed=300 + cnst, cnst=cnst, fec_id=1, snr_db=snr)
    ref = ocfg.rx_batch(t["iq"], t["taps"], t["meta"], S, F)
    # stage 4 alone on the oracle's LLRs: identical inputs -> identical bits and iteration counts
    llr = np.ascontiguousarray(ref["llr"][:, :rx.llr_stride])
    got = rx.fec_decode(llr, t["meta"], S, F)
    assert any(tb["avg_it"].max() > 0 for tb in ref["tbs"]), "test SNR too high: BP never iterated"
    for s in range(S):
        g = got["tbs"][s, :got["ntbs"][s]]
        assert np.array_equal(g, ref["tbs"][s]), "TB records differ (tb_no, crc_ok, nbytes, avg_it, frame_index)"
        assert np.array_equal(got["bytes"][s, :got["nbytes"][s]], ref["bytes"][s])
    rx.close()


@pytest.mark.parametrize("cnst,snr", [(2, 19.0), (4, 24.0)])
def test_rx_whole_path_config2(cnst, snr):
    ocfg = make_oracle_cfg(alists=(ALIST_300,), initial_mcs=1)
    rx = make_receiver(alists=(ALIST_300,), initial_mcs=1)
    S, F = 8, 12
    t = ocfg.tx_batch(S, F, seed=400 + cnst, cnst=cnst, fec_id=1, snr_db=snr)
    ref = ocfg.rx_batch(t["iq"], t["taps"], t["meta"], S, F)
    got = rx.rx(t["iq"], t["taps"], t["meta"], S, F)
    for k in ("fb_cnst", "fb_fec", "emitted", "lost_frames"):
        assert np.array_equal(got["res"][k], ref["res"][k]), k
    assert np.allclose(got["res"]["snr_db"], ref["res"]["snr_db"], rtol=1e-4)    # FFT rounding differs upstream
    for s in range(S):
        gb = got["bytes"][s, :got["nbytes"][s]]
        assert np.array_equal(gb, ref["bytes"][s]), "decoded payload differs from the oracle"
        assert np.array_equal(gb, t["payload"][s][:gb.size]) and gb.size > 0
        g = got["tbs"][s, :got["ntbs"][s]]
        assert np.array_equal(g["crc_ok"], ref["tbs"][s]["crc_ok"]) and g["crc_ok"].sum() >= g.size - 1
    rx.close()


def test_rx_whole_path_config1_no_fec():
    ocfg = make_oracle_cfg()
    rx = make_receiver()
    S, F = 6, 10
    rng = np.random.default_rng(1)
    cpf = rng.integers(1, 5, size=S * F).astype(np.int32)          # adaptive: constellation changes per frame
    t = ocfg.tx_batch(S, F, seed=41, cnst=1, snr_db=34.0, cnst_per_frame=cpf)
    ref = ocfg.rx_batch(t["iq"], t["taps"], t["meta"], S, F)
    got = rx.rx(t["iq"], t["taps"], t["meta"], S, F)
    assert np.array_equal(got["frame_crc"], ref["frame_crc"]) and ref["frame_crc"].all()
    for s in range(S):
        gb = got["bytes"][s, :got["nbytes"][s]]
        assert np.array_equal(gb, ref["bytes"][s])
        assert np.array_equal(gb, t["payload"][s])
    rx.close()


def test_state_carries_across_calls():
    """TB reassembly, feedback counters and frame counters survive between calls: F frames in one call ==
    the same frames in three calls."""
    ocfg = make_oracle_cfg(alists=(ALIST_300,))
    S, F = 5, 12
    t = ocfg.tx_batch(S, F, seed=77, cnst=3, fec_id=1, snr_db=18.0)
    rx = make_receiver(alists=(ALIST_300,))
    one = rx.rx(t["iq"], t["taps"], t["meta"], S, F)
    rx.reset(S)
    iq = t["iq"].reshape(S, F, -1)
    taps = t["taps"].reshape(S, F, -1)
    meta = t["meta"].reshape(S, F)
    parts, cuts = [], [(0, 5), (5, 6), (6, 12)]
    for a, b in cuts:
        parts.append(rx.rx(np.ascontiguousarray(iq[:, a:b]).reshape(S * (b - a), -1),
                           np.ascontiguousarray(taps[:, a:b]).reshape(S * (b - a), -1),
                           np.ascontiguousarray(meta[:, a:b]).reshape(-1), S, b - a))
    for s in range(S):
        cat = np.concatenate([p["bytes"][s, :p["nbytes"][s]] for p in parts])
        assert np.array_equal(cat, one["bytes"][s, :one["nbytes"][s]])
        assert cat.size > 0
    fb = np.concatenate([p["res"]["fb_cnst"].reshape(S, -1) for p in parts], axis=1)
    assert np.array_equal(fb, one["res"]["fb_cnst"].reshape(S, F))
    rx.close()


# ---- stage 5 ----------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("fec", [False, True])
def test_tx_modulator(fec):
    alists = (ALIST_300,) if fec else ()
    ocfg = make_oracle_cfg(alists=alists)
    rx = make_receiver(alists=alists)
    S, F = 2, 3
    t = ocfg.tx_batch(S, F, seed=5, cnst=orc.QAM16, fec_id=1 if fec else 0, snr_db=200.0)
    H, D = ocfg.header_syms, 48
    hdr = orc.PacketHeader(H * D, ocfg.frame_len, fec, max_cnst=6)
    hb = np.zeros((S * F, H * D), dtype=np.uint8)
    ref = np.zeros((S * F, ocfg.frame_samples), dtype=np.complex64)
    for s in range(S):
        hdr = orc.PacketHeader(H * D, ocfg.frame_len, fec, max_cnst=6)
        for j in range(F):
            f = s * F + j
            hb[f] = hdr.format(t["meta"][f])
            ref[f] = ocfg.tx_modulate(orc.map_chunks(orc.BPSK, hb[f]), orc.map_chunks(orc.QAM16, t["chunks"][f]))
    got = rx.tx(hb, t["chunks"], t["meta"])
    assert _rel(got, ref) < RT_FFT
    # loopback: our TX -> our RX demod -> equaliser with flat taps decodes the chunks
    freq = rx.demod(got)
    taps = np.full((S * F, 64), 64.0 + 0j, dtype=np.complex64)
    out = rx.equalize(freq, taps, t["meta"], S, F, want=("soft",))
    for f in range(S * F):
        ser = ocfg.serialize(out["soft"][f], t["meta"]["packet_len"][f])
        assert np.array_equal(orc.decide(orc.QAM16, ser), t["chunks"][f][:ser.size])
    rx.close()


@pytest.mark.gpu
def test_rx_host_pipelined_matches_oracle_and_device_path():
    """Large DTL_MEM_HOST batches go through the chunked copy/compute pipeline (api.cc rx_host_pipelined): same bytes,
    TB records and per-frame results as the oracle and as the single-shot device-buffer path, across two calls
    (the TB in flight at the end of call 1 is finished in call 2)."""
    import torch
    ocfg = make_oracle_cfg(alists=(ALIST_300,), initial_mcs=3)
    rx = make_receiver(alists=(ALIST_300,), initial_mcs=3)
    rx_d = make_receiver(alists=(ALIST_300,), initial_mcs=3)
    S, F = 50, 24                                   # 18 MB of IQ per call -> 4 chunks
    t = ocfg.tx_batch(S, 2 * F, seed=77, cnst=4, fec_id=1, snr_db=23.0)
    fs = ocfg.frame_samples
    iq = t["iq"].reshape(S, 2 * F, fs)
    taps = t["taps"].reshape(S, 2 * F, 64)
    meta = t["meta"].reshape(S, 2 * F)
    ref = ocfg.rx_batch(t["iq"], t["taps"], t["meta"], S, 2 * F)
    got_bytes = [[] for _ in range(S)]
    dev_bytes = [[] for _ in range(S)]
    for half in range(2):
        sl = slice(half * F, (half + 1) * F)
        a = np.ascontiguousarray(iq[:, sl]).reshape(S * F, fs)
        b = np.ascontiguousarray(taps[:, sl]).reshape(S * F, 64)
        m = np.ascontiguousarray(meta[:, sl]).reshape(S * F)
        g = rx.rx(a, b, m, S, F)
        tm = torch.from_numpy(m.view(np.uint8).reshape(S * F, -1).copy()).cuda()
        d = rx_d.rx(torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda(), tm, S, F)
        rx_d.sync()
        d_nb, d_by = d["nbytes"].cpu().numpy(), d["bytes"].cpu().numpy()
        assert np.array_equal(g["nbytes"], d_nb)
        assert np.array_equal(g["ntbs"], d["ntbs"].cpu().numpy())
        assert np.array_equal(g["res"].view(np.uint8), d["res"].cpu().numpy())
        for s in range(S):
            got_bytes[s].append(g["bytes"][s, :g["nbytes"][s]])
            dev_bytes[s].append(d_by[s, :d_nb[s]])
    for s in range(S):
        assert np.array_equal(np.concatenate(got_bytes[s]), ref["bytes"][s]), s
        assert np.array_equal(np.concatenate(dev_bytes[s]), ref["bytes"][s]), s
    rx.close()
    rx_d.close()
