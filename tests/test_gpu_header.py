"""Header branch (python/dtl/ofdm_receiver.py:100-150): channel estimation from the
sync words, header equaliser, header parser -- the stream tags of a frame from its IQ alone -- and the receive path
that starts from IQ only.  Oracle: oracle/src/header.cc.  Bar: every parsed field, the detected carrier offset and the
CRC flags bit-exact; channel taps 1e-4 relative (the FFT upstream rounds differently); decoded payload bit-exact."""
import numpy as np
import pytest

import pyoracle as orc
from common import ALIST_300, ALIST_100_23, make_oracle_cfg, make_receiver

pytestmark = pytest.mark.gpu

ALL = (1, 2, 3, 4, 5, 6)
FIELDS = ("constellation", "carrier_offset", "frame_no", "payload_len", "packet_len", "fec_id", "tb_no", "tb_offset",
          "tb_payload", "reserved")


def _compare_headers(got, ref, taps_rtol=1e-4):
    for k in FIELDS:
        assert np.array_equal(got["meta"][k], ref["meta"][k]), k
    assert np.array_equal(got["hres"]["crc_ok"], ref["ok"])
    assert np.array_equal(got["hres"]["feedback_cnst"], ref["fb"][:, 0])
    assert np.array_equal(got["hres"]["feedback_fec"], ref["fb"][:, 1])
    good = ref["ok"] == 1
    a, b = got["taps"][good], ref["taps"][good]
    assert np.array_equal(a == 0, b == 0), "unused carriers must carry zero taps"
    nz = b != 0
    assert np.all(np.abs(a[nz] - b[nz]) <= taps_rtol * np.abs(b[nz]))


@pytest.mark.parametrize("fec", [False, True])
def test_header_branch_vs_oracle(fec):
    al = (ALIST_300,) if fec else ()
    ocfg = make_oracle_cfg(cnsts=ALL, alists=al)
    rx = make_receiver(cnsts=ALL, alists=al)
    S, F = 12, 9
    cpf = np.repeat((np.arange(S) % 6 + 1).astype(np.int32), F)
    t = ocfg.tx_batch(S, F, seed=77 + fec, cnst=1, fec_id=1 if fec else 0, snr_db=18.0, cnst_per_frame=cpf)
    ref = ocfg.rx_headers(t["iq"], S, F)
    got = rx.rx_headers(t["iq"], S, F)
    assert ref["ok"].all()
    for k in FIELDS:                                    # at 18 dB the parsed tags are the transmitted ones
        assert np.array_equal(ref["meta"][k], t["meta"][k]), k
    _compare_headers(got, ref)
    rx.close()


@pytest.mark.parametrize("g", [-4, -2, 2, 4])
def test_header_branch_integer_carrier_offset(g):
    """An integer carrier offset of g bins (even, as ofdm_chanest_vcvc assumes) is detected, undone by the header
    equaliser, handed to the payload branch as the "ofdm_sync_carr_offset" tag, and the payload still decodes."""
    ocfg = make_oracle_cfg(cnsts=ALL, alists=(ALIST_300,))
    rx = make_receiver(cnsts=ALL, alists=(ALIST_300,))
    S, F = 4, 8
    t = ocfg.tx_batch(S, F, seed=500 + g, cnst=orc.QPSK, fec_id=1, snr_db=24.0)
    fs = ocfg.frame_samples
    rot = np.exp(2j * np.pi * g * np.arange(fs) / 64.0)
    iq = (t["iq"].reshape(S * F, fs) * rot[None, :]).astype(np.complex64)
    ref = ocfg.rx_headers(iq, S, F)
    got = rx.rx_headers(iq, S, F)
    assert ref["ok"].all() and np.all(ref["meta"]["carrier_offset"] == g)
    _compare_headers(got, ref)
    full = rx.rx_iq(iq, S, F)
    oref = ocfg.rx_batch(iq, ref["taps"], ref["meta"], S, F)
    for s in range(S):
        gb = full["bytes"][s, :full["nbytes"][s]]
        assert np.array_equal(gb, oref["bytes"][s]) and gb.size > 0
        assert np.array_equal(gb, t["payload"][s][:gb.size])
    rx.close()


@pytest.mark.parametrize("g", [-2, 4])
def test_header_phase_when_the_header_is_longer_than_the_payload(g):
    """frame_len = 1 with FEC: two header symbols, one payload symbol.  The header symbols' carrier-offset phases come from the
    header's own table (the payload table of such a frame has one symbol entry; its next entry is the channel-state rotation)."""
    ocfg = make_oracle_cfg(frame_len=1, cnsts=ALL, alists=(ALIST_100_23,))
    rx = make_receiver(frame_len=1, cnsts=ALL, alists=(ALIST_100_23,))
    assert ocfg.header_syms == 2 and ocfg.frame_len == 1
    S, F = 3, 10
    t = ocfg.tx_batch(S, F, seed=900 + g, cnst=5, fec_id=1, snr_db=36.0)          # 64QAM: one 48-carrier symbol holds a TB of the n = 100 code
    fs = ocfg.frame_samples
    rot = np.exp(2j * np.pi * g * np.arange(fs) / 64.0)
    iq = (t["iq"].reshape(S * F, fs) * rot[None, :]).astype(np.complex64)
    ref = ocfg.rx_headers(iq, S, F)
    got = rx.rx_headers(iq, S, F)
    assert ref["ok"].all() and np.all(ref["meta"]["carrier_offset"] == g)
    _compare_headers(got, ref)
    rx.close()


def test_header_crc_failures_at_low_snr():
    """Around 7 dB a fifth of the 96-bit BPSK headers fail the CRC-16: the same frames fail on both sides."""
    ocfg = make_oracle_cfg(cnsts=ALL, alists=(ALIST_300,))
    rx = make_receiver(cnsts=ALL, alists=(ALIST_300,))
    S, F = 16, 16
    t = ocfg.tx_batch(S, F, seed=10, cnst=orc.BPSK, fec_id=1, snr_db=7.0)
    ref = ocfg.rx_headers(t["iq"], S, F)
    assert 0.3 < ref["ok"].mean() < 0.98
    _compare_headers(rx.rx_headers(t["iq"], S, F), ref)
    rx.close()


def test_dropped_frames_take_no_part_in_the_payload_branch():
    """Headers of chosen frames are destroyed (first header symbol negated) on an otherwise clean link: those frames are
    dropped -- no tags, no counters, no TB input -- and everything else matches the oracle run on the same tags."""
    ocfg = make_oracle_cfg(cnsts=ALL, alists=(ALIST_300,), initial_mcs=1)
    rx = make_receiver(cnsts=ALL, alists=(ALIST_300,), initial_mcs=1)
    S, F = 8, 14
    t = ocfg.tx_batch(S, F, seed=321, cnst=orc.QPSK, fec_id=1, snr_db=26.0)
    fs, sym = ocfg.frame_samples, 64 + 16
    iq = t["iq"].reshape(S * F, fs).copy()
    kill = np.zeros(S * F, dtype=bool)
    kill[np.random.default_rng(5).choice(S * F, size=17, replace=False)] = True
    iq[kill, 2 * sym:3 * sym] *= -1
    ref = ocfg.rx_headers(iq, S, F)
    assert np.array_equal(ref["ok"] == 0, kill)
    full = rx.rx_iq(iq, S, F)
    oref = ocfg.rx_batch(iq, ref["taps"], ref["meta"], S, F)
    assert np.array_equal(full["hres"]["crc_ok"], ref["ok"])
    live = ~kill
    for k in ("fb_cnst", "fb_fec", "emitted", "lost_frames"):
        assert np.array_equal(full["res"][k][live], oref["res"][k][live]), k
    assert np.all(full["res"]["emitted"][kill] == 0)
    assert full["res"]["lost_frames"].max() > 0                     # the gaps show up in the frame counter
    n_bytes = 0
    for s in range(S):
        gb = full["bytes"][s, :full["nbytes"][s]]
        assert np.array_equal(gb, oref["bytes"][s])
        n_bytes += gb.size
        g, r = full["tbs"][s, :full["ntbs"][s]], oref["tbs"][s]
        assert g.size == r.size
        for k in ("tb_no", "crc_ok", "nbytes", "frame_index"):
            assert np.array_equal(g[k], r[k]), k
    assert n_bytes > 0
    rx.close()


@pytest.mark.parametrize("fec", [False, True])
def test_rx_from_iq_alone(fec):
    """dtl_b200_rx_iq == header branch + payload branch of the oracle, and what comes out is what was sent."""
    al = (ALIST_300,) if fec else ()
    ocfg = make_oracle_cfg(cnsts=ALL, alists=al, initial_mcs=1)
    rx = make_receiver(cnsts=ALL, alists=al, initial_mcs=1)
    S, F = 10, 10
    cpf = np.repeat((np.arange(S) % 4 + 1).astype(np.int32), F)
    t = ocfg.tx_batch(S, F, seed=901 + fec, cnst=1, fec_id=1 if fec else 0, snr_db=30.0, cnst_per_frame=cpf)
    ref = ocfg.rx_headers(t["iq"], S, F)
    oref = ocfg.rx_batch(t["iq"], ref["taps"], ref["meta"], S, F)
    got = rx.rx_iq(t["iq"], S, F)
    for k in FIELDS:
        assert np.array_equal(got["meta"][k], ref["meta"][k]), k
    for k in ("fb_cnst", "fb_fec", "emitted", "lost_frames"):
        assert np.array_equal(got["res"][k], oref["res"][k]), k
    assert np.allclose(got["res"]["snr_db"], oref["res"]["snr_db"], rtol=1e-4)
    total = 0
    for s in range(S):
        gb = got["bytes"][s, :got["nbytes"][s]]
        assert np.array_equal(gb, oref["bytes"][s])
        assert np.array_equal(gb, t["payload"][s][:gb.size])
        total += gb.size
    assert total > 0
    rx.close()


def test_rx_from_iq_scaled_phy_1024():
    """Config 5's PHY (1024-point FFT, 96 candidate carrier offsets, 1536 header bits) through the header branch."""
    kw = dict(fft_len=1024, frame_len=5, cnsts=ALL, alists=(ALIST_300,), initial_mcs=5)
    ocfg = make_oracle_cfg(**kw)
    rx = make_receiver(**kw)
    S, F = 3, 5
    t = ocfg.tx_batch(S, F, seed=5100, cnst=orc.QAM256, fec_id=1, snr_db=40.0)
    ref = ocfg.rx_headers(t["iq"], S, F)
    got = rx.rx_headers(t["iq"], S, F)
    assert ref["ok"].all()
    _compare_headers(got, ref)
    full = rx.rx_iq(t["iq"], S, F)
    oref = ocfg.rx_batch(t["iq"], ref["taps"], ref["meta"], S, F)
    for s in range(S):
        gb = full["bytes"][s, :full["nbytes"][s]]
        assert np.array_equal(gb, oref["bytes"][s]) and gb.size > 0
    rx.close()
