"""The oracle's arithmetic CONTRACT against the reference's LITERAL expressions, and the CUDA path against the literal mode.

The CUDA equaliser is bit-exact against the oracle's contract mode (oracle/src/contract.hpp: every fp32 operation an
explicit IEEE single operation, division as multiplication by a reciprocal, constellation-point reciprocals tabulated).
The reference itself divides `std::complex<float>` values (lib/dtl/ofdm_adaptive_equalizer.cc:241-256) -- the oracle keeps
those expressions as `literal` mode (oracle/src/equalizer.cc:65-68,78-81; `Config.set_literal`, `FrameEqualizer(literal=True)`).
This file states what holds between the two, on the inputs of BASELINE.json's five configurations:

  bit-exact   hard decisions (port 0), feedback MCS, emitted symbols, lost-frame counters, CRC flags, TB records, payload bytes
  bounded     equalised symbols (port 1) 1e-6, channel taps 1e-6, SNR tag [dB] 1e-6, noise tag 5e-6 (all relative)

(measured worst cases over these inputs: 3.7e-7 / 2.0e-7 / 2.5e-7 / 1.8e-6).  These -- not "bit-exact taps" -- are the
honest bars against the reference's own expressions; they sit far inside north_star's 1e-4.
"""
import numpy as np
import pytest

import pyoracle as po
from common import ALIST_300, make_oracle_cfg, make_receiver

ALL = (1, 2, 3, 4, 5, 6)
# (name, oracle/product configuration, constellation, SNR dB, S, F): configs[0..4] of BASELINE.json
CASES = [
    ("config1_nofec_qpsk", dict(), 2, 14.0, 4, 8),
    ("config1_nofec_8psk", dict(), 3, 20.0, 4, 8),
    ("config2_fec_qpsk", dict(alists=(ALIST_300,), initial_mcs=1, lut="config_fec"), 2, 17.0, 4, 8),
    ("config2_fec_16qam", dict(alists=(ALIST_300,), initial_mcs=1, lut="config_fec"), 4, 22.0, 4, 8),
    ("config3_bpsk", dict(cnsts=ALL, alists=(ALIST_300,)), 1, 12.0, 4, 8),
    ("config3_64qam", dict(cnsts=ALL, alists=(ALIST_300,)), 5, 30.0, 4, 8),
    ("config3_256qam", dict(cnsts=ALL, alists=(ALIST_300,)), 6, 34.0, 4, 8),
    ("config4_16qam_ldpc", dict(alists=(ALIST_300,), initial_mcs=3), 4, 22.0, 6, 8),
    ("config5_fft1024_256qam", dict(fft_len=1024, frame_len=5, cnsts=ALL, alists=(ALIST_300,)), 6, 36.0, 2, 4),
]
TOL = dict(soft=1e-6, taps=1e-6, snr_db=1e-6, noise=5e-6)


def _rel(a, b, floor=1e-3):
    a, b = np.asarray(a), np.asarray(b)
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), floor))) if a.size else 0.0


def _batch(kw, cnst, snr, S, F, seed):
    ocfg = make_oracle_cfg(**kw)
    fec = bool(kw.get("alists"))
    t = ocfg.tx_batch(S, F, seed=seed, cnst=cnst, fec_id=1 if fec else 0, snr_db=snr)
    return ocfg, t, fec


def _compare_exact(a, b, S, fec):
    """a, b: outputs of rx_batch-shaped calls (lists per stream); every bit-exact item of the module docstring"""
    for k in ("fb_cnst", "fb_fec", "emitted", "lost_frames"):
        assert np.array_equal(a["res"][k], b["res"][k]), k
    for s in range(S):
        assert np.array_equal(a["bytes"][s], b["bytes"][s]), "payload bytes, stream %d" % s
        if fec:
            for k in ("tb_no", "crc_ok", "nbytes", "frame_index"):
                assert np.array_equal(a["tbs"][s][k], b["tbs"][s][k]), "TB records: " + k
            # iteration counts: exact on the blocks that decode; a block that never converges wanders for as long as the
            # last bits of its LLRs say (the two modes' LLRs differ by ~1e-7 relative)
            ok = b["tbs"][s]["crc_ok"] == 1
            assert np.array_equal(a["tbs"][s]["avg_it"][ok], b["tbs"][s]["avg_it"][ok]), "avg_it of the decoded TBs"
    if not fec:
        assert np.array_equal(a["frame_crc"], b["frame_crc"])


@pytest.mark.parametrize("name,kw,cnst,snr,S,F", CASES, ids=[c[0] for c in CASES])
def test_contract_mode_vs_reference_literal_mode(name, kw, cnst, snr, S, F):
    """CPU: the two oracle modes on the same frequency-domain frames, block level and whole path."""
    ocfg, t, fec = _batch(kw, cnst, snr, S, F, seed=900 + cnst)
    freq = ocfg.rx_demod(t["iq"])
    worst = dict(soft=0.0, taps=0.0, snr_db=0.0, noise=0.0)
    for s in range(S):
        con, lit = po.FrameEqualizer(ocfg, literal=False), po.FrameEqualizer(ocfg, literal=True)
        for j in range(F):
            f = s * F + j
            na, ha, sa, ra, ta = con.work(freq[f], t["taps"][f], t["meta"][f])
            nb, hb, sb, rb, tb = lit.work(freq[f], t["taps"][f], t["meta"][f])
            assert na == nb and np.array_equal(ha, hb), "hard decisions differ between the modes (frame %d)" % f
            assert (ra["fb_cnst"], ra["fb_fec"], ra["lost_frames"]) == (rb["fb_cnst"], rb["fb_fec"], rb["lost_frames"])
            worst["soft"] = max(worst["soft"], _rel(sa, sb))
            worst["taps"] = max(worst["taps"], _rel(ta, tb))
            worst["snr_db"] = max(worst["snr_db"], _rel(ra["snr_db"], rb["snr_db"]))
            worst["noise"] = max(worst["noise"], _rel(ra["noise"], rb["noise"], floor=1e-12))
    for k, v in worst.items():
        assert v <= TOL[k], "%s: %g > %g" % (k, v, TOL[k])
    # whole path: decoded payload, CRC flags, TB records, feedback
    a = ocfg.rx_batch(t["iq"], t["taps"], t["meta"], S, F)
    ocfg.set_literal(True)
    b = ocfg.rx_batch(t["iq"], t["taps"], t["meta"], S, F)
    ocfg.set_literal(False)
    _compare_exact(a, b, S, fec)
    assert sum(x.size for x in b["bytes"]) > 0


@pytest.mark.gpu
@pytest.mark.parametrize("name,kw,cnst,snr,S,F", CASES, ids=[c[0] for c in CASES])
def test_cuda_path_vs_reference_literal_mode(name, kw, cnst, snr, S, F):
    """GPU twin: libdtl_b200 against the oracle's LITERAL mode (the reference's own expressions)."""
    ocfg, t, fec = _batch(kw, cnst, snr, S, F, seed=900 + cnst)
    ocfg.set_literal(True)
    rx = make_receiver(**kw)
    N, L = rx.fft_len, rx.frame_length
    freq = ocfg.rx_demod(t["iq"]).reshape(S * F, L, N)            # same frequency-domain input on both sides
    ref = ocfg.rx_batch(freq.reshape(S * F, -1), t["taps"], t["meta"], S, F, freq_domain=True)
    got = rx.equalize(freq, t["taps"], t["meta"], S, F, want=("hard", "soft", "taps", "res"))
    assert np.array_equal(got["hard"].reshape(S * F, -1), ref["hard"]), "hard decisions (port 0)"
    for k in ("fb_cnst", "fb_fec", "emitted", "lost_frames"):
        assert np.array_equal(got["res"][k], ref["res"][k]), k
    mask = np.abs(ref["soft"]) > 0                                  # data carriers only (port 1 is written there only)
    assert _rel(got["soft"].reshape(S * F, -1)[mask], ref["soft"][mask]) <= TOL["soft"]
    assert _rel(got["res"]["snr_db"], ref["res"]["snr_db"]) <= TOL["snr_db"]
    assert _rel(got["res"]["noise"], ref["res"]["noise"], floor=1e-12) <= TOL["noise"]
    # taps after the frame against the literal frame equaliser, frame by frame
    worst = 0.0
    for s in range(S):
        lit = po.FrameEqualizer(ocfg, literal=True)
        for j in range(F):
            f = s * F + j
            worst = max(worst, _rel(got["taps"][f], lit.work(freq[f].reshape(-1), t["taps"][f], t["meta"][f])[4]))
    assert worst <= TOL["taps"]
    # whole path from the same frequency-domain frames: bytes, CRC flags, TB records
    rx.reset(S)                                                     # a fresh set of receivers, as the oracle's call
    full = rx.rx(freq, t["taps"], t["meta"], S, F, freq_domain=True)
    g = dict(res=full["res"], frame_crc=full["frame_crc"],
             bytes=[full["bytes"][s, :full["nbytes"][s]] for s in range(S)],
             tbs=[full["tbs"][s, :full["ntbs"][s]] for s in range(S)])
    _compare_exact(g, ref, S, fec)      # iteration counts on identical LLRs are pinned by test_ldpc_decode_matches_awgn_bp
    rx.close()
