"""Regression tests for defects found by review."""
import numpy as np
import pytest

import pyoracle as orc
from common import ALIST_300, make_oracle_cfg, make_receiver

pytestmark = pytest.mark.gpu


def test_no_fec_host_pipelined_results_wait_for_the_side_stream():
    """fec_enabled = 0, DTL_MEM_HOST, >= 8 streams and >= 16 MB of input: rx_host_pipelined copies `res` out per chunk.  The
    feedback fields are written by stream_state_kernel on the side stream; the no-FEC branch has to join it first."""
    ocfg = make_oracle_cfg()
    rx = make_receiver()
    S, F = 16, 80                                      # 16 x 80 x 1840 x 8 B = 18.8 MB -> chunked path
    assert S * F * rx.frame_samples * 8 >= 16 << 20
    rng = np.random.default_rng(2)
    cpf = rng.integers(1, 5, size=S * F).astype(np.int32)
    t = ocfg.tx_batch(S, F, seed=77, cnst=1, snr_db=26.0, cnst_per_frame=cpf)
    ref = ocfg.rx_batch(t["iq"], t["taps"], t["meta"], S, F, want_llr=False, want_syms=False)
    for rep in range(3):                               # a race does not show every time
        rx.reset(S)
        got = rx.rx(t["iq"], t["taps"], t["meta"], S, F)
        for k in ("fb_cnst", "fb_fec", "emitted", "lost_frames"):
            assert np.array_equal(got["res"][k], ref["res"][k]), (k, rep)
        assert np.allclose(got["res"]["snr_db"], ref["res"]["snr_db"], rtol=1e-4)
        assert np.array_equal(got["frame_crc"], ref["frame_crc"])
        for s in range(S):
            assert np.array_equal(got["bytes"][s, :got["nbytes"][s]], ref["bytes"][s])
    assert len(set(ref["res"]["fb_cnst"])) > 1, "feedback never moved: vacuous"
    rx.close()


@pytest.mark.parametrize("fec", [False, True])
def test_mcs_stays_at_zero_when_the_estimate_is_below_every_threshold(fec):
    """SNR estimate <= 0 dB at MCS 0 for more than decision_th frames (a link in a deep fade): the reference only asserts
    current_mcs_id > 0 (ofdm_adaptive_feedback_decision.cc:70); here the MCS stays at 0 (no out-of-bounds table reads), and
    recovers when the SNR comes back."""
    al = (ALIST_300,) if fec else ()
    ocfg = make_oracle_cfg(alists=al)
    rx = make_receiver(alists=al)
    S, F = 3, 10
    bad = ocfg.tx_batch(S, F, seed=5, cnst=orc.BPSK, fec_id=1 if fec else 0, snr_db=-4.0)
    good = ocfg.tx_batch(S, F, seed=6, cnst=orc.BPSK, fec_id=1 if fec else 0, snr_db=20.0)
    for k in ("frame_no",):
        good["meta"][k] = (good["meta"][k] + F) % 4096
    iq = np.concatenate([bad["iq"].reshape(S, F, -1), good["iq"].reshape(S, F, -1)], axis=1).reshape(2 * S * F, -1)
    taps = np.concatenate([bad["taps"].reshape(S, F, -1), good["taps"].reshape(S, F, -1)], axis=1).reshape(2 * S * F, -1)
    meta = np.concatenate([bad["meta"].reshape(S, F), good["meta"].reshape(S, F)], axis=1).reshape(-1)
    if fec:                                            # tag bookkeeping of the second half starts a new TB sequence
        meta["tb_no"][np.arange(2 * S * F) % (2 * F) >= F] += 100
    ref = ocfg.rx_batch(iq, taps, meta, S, 2 * F)
    got = rx.rx(iq, taps, meta, S, 2 * F)
    assert (ref["res"]["snr_db"].reshape(S, 2 * F)[:, :F] < 0).all()
    for k in ("fb_cnst", "fb_fec", "lost_frames"):
        assert np.array_equal(got["res"][k], ref["res"][k]), k
    fb = got["res"]["fb_cnst"].reshape(S, 2 * F)
    assert (fb[:, :F] == orc.BPSK).all() and (fb[:, -1] == orc.QPSK).all()
    rx.close()


def test_one_frame_per_call_120_consecutive_frames_match_the_oracle():
    """The latency mode of a drop-in shim: S = 1, F = 1, DTL_MEM_HOST, one tagged frame per call.  120
    consecutive frames of one stream, SNR low enough that BP iterates and some TBs fail: per-frame results, TB records and
    payload equal the oracle's run over the same 120 frames (per-stream state -- TB reassembly across frames, feedback
    counters, frame numbers -- lives in the handle between the calls)."""
    import pyoracle as orc
    from common import ALIST_300, make_oracle_cfg
    ocfg = make_oracle_cfg(alists=(ALIST_300,), initial_mcs=1)
    rx = make_receiver(alists=(ALIST_300,), initial_mcs=1)
    F = 120
    t = ocfg.tx_batch(1, F, seed=123, cnst=orc.QPSK, fec_id=1, snr_db=16.5)
    freq = ocfg.rx_demod(t["iq"])
    ref = ocfg.rx_batch(freq, t["taps"], t["meta"], 1, F, freq_domain=True)
    fq = freq.reshape(F, ocfg.frame_len, 64)
    rx.reset(1)
    got_bytes, got_tbs = [], []
    for j in range(F):
        g = rx.rx(np.ascontiguousarray(fq[j:j + 1]), np.ascontiguousarray(t["taps"][j:j + 1]), t["meta"][j:j + 1].copy(), 1, 1,
                  freq_domain=True)
        for k in ("fb_cnst", "fb_fec", "emitted", "lost_frames"):
            assert g["res"][k][0] == ref["res"][k][j], (k, j)
        assert np.isclose(g["res"]["snr_db"][0], ref["res"]["snr_db"][j], rtol=1e-9)
        got_bytes.append(g["bytes"][0, :g["nbytes"][0]].copy())
        tb = g["tbs"][0, :g["ntbs"][0]].copy()
        tb["frame_index"] += j                                   # frame index within the call -> within the run
        got_tbs.append(tb)
    tbs = np.concatenate(got_tbs)
    r = ref["tbs"][0]
    assert r["avg_it"].max() > 0 and (r["crc_ok"] == 0).any() and (r["crc_ok"] == 1).any(), "operating point not challenging"
    for k in ("tb_no", "crc_ok", "nbytes", "avg_it"):
        assert np.array_equal(tbs[k], r[k]), k
    assert np.array_equal(np.concatenate(got_bytes), ref["bytes"][0])
    rx.close()


def test_rx_iq_from_host_memory_in_chunks_equals_the_device_path():
    """dtl_b200_rx_iq with DTL_MEM_HOST and a big batch goes through the chunked pipeline (IQ alone is uploaded, every chunk's
    header branch makes its taps and tags on the device): results equal the one-shot device-buffer path and the oracle."""
    import torch
    ocfg = make_oracle_cfg(alists=(ALIST_300,), initial_mcs=1)
    rx = make_receiver(alists=(ALIST_300,), initial_mcs=1)
    S, F = 24, 48                                      # 24 x 48 x 1920 x 8 B = 17.7 MB -> chunked path (six chunks of four streams)
    assert S * F * rx.frame_samples * 8 >= 16 << 20
    t = ocfg.tx_batch(S, F, seed=91, cnst=orc.QPSK, fec_id=1, snr_db=17.0)
    iq = np.ascontiguousarray(t["iq"])
    rx.reset(S)
    host = rx.rx_iq(iq, S, F)
    rx.reset(S)
    dev = rx.rx_iq(torch.from_numpy(iq.view(np.float32)).to("cuda:0").view(torch.complex64), S, F)
    rx.sync()
    cpu = lambda v: v.cpu().numpy() if hasattr(v, "cpu") else v
    for k in ("res", "meta", "hres", "nbytes", "ntbs", "tbs", "frame_crc"):
        assert cpu(dev[k]).tobytes() == np.asarray(host[k]).tobytes(), k
    nb = np.asarray(host["nbytes"])
    db = cpu(dev["bytes"]).reshape(S, -1)
    ref_h = ocfg.rx_headers(iq, S, F)
    ref = ocfg.rx_batch(iq, ref_h["taps"], ref_h["meta"], S, F)
    tbs = host["tbs"]
    assert any((tbs[s, :host["ntbs"][s]]["crc_ok"] == 0).any() for s in range(S)) and nb.sum() > 0, "operating point not challenging"
    for s in range(S):
        assert np.array_equal(host["bytes"][s, :nb[s]], db[s, :nb[s]])
        assert np.array_equal(host["bytes"][s, :nb[s]], ref["bytes"][s])
    rx.close()
