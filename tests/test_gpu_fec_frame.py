"""TX FEC framer on the GPU (lib/dtl/ofdm_adaptive_fec_frame_bvb_impl.cc:343-432,
lib/dtl/tb_encoder.cc:32-99, lib/dtl/ldpc_enc.cc:58-68).  Oracle: oracle/src/txrx.cc tx_framer fed with the same bytes.
Bar: chunks and every tag bit-exact, bytes consumed and TBs started equal; and what the framer sends comes back
through the receive path (header formatter -> modulator -> channel -> dtl_b200_rx_iq)."""
import numpy as np
import pytest

import pyoracle as orc
from common import ALIST_300, make_oracle_cfg, make_receiver

pytestmark = pytest.mark.gpu

ALL = (1, 2, 3, 4, 5, 6)
FIELDS = ("constellation", "frame_no", "payload_len", "packet_len", "fec_id", "tb_no", "tb_offset", "tb_payload")


def _compare(got, ref):
    for k in FIELDS:
        assert np.array_equal(got["meta"][k], ref["meta"][k]), k
    assert np.array_equal(got["consumed"], ref["consumed"])
    assert np.array_equal(got["tbs_started"], ref["started"])
    assert np.array_equal(got["chunks"], ref["chunks"])


@pytest.mark.parametrize("cnst", ALL)
def test_fec_frame_vs_oracle_fixed_constellation(cnst):
    ocfg = make_oracle_cfg(cnsts=ALL, alists=(ALIST_300,))
    rx = make_receiver(cnsts=ALL, alists=(ALIST_300,))
    S, F = 7, 11
    rng = np.random.default_rng(40 + cnst)
    data = rng.integers(0, 256, size=(S, F * 1200), dtype=np.uint8)
    c = np.full(S * F, cnst, dtype=np.int32)
    f = np.ones(S * F, dtype=np.int32)
    ref = ocfg.frame_batch(data, c, f, S, F)
    got = rx.fec_frame(data, c, f, S, F)
    assert ref["started"].min() > 0
    _compare(got, ref)
    rx.close()


def test_fec_frame_constellation_changes_between_frames_and_input_runs_out():
    """The MCS changes from frame to frame (TB sized by the frame that starts it, continued at the new bits per symbol);
    streams whose input runs out stop starting TBs and send what is in flight."""
    ocfg = make_oracle_cfg(cnsts=ALL, alists=(ALIST_300,))
    rx = make_receiver(cnsts=ALL, alists=(ALIST_300,))
    S, F = 9, 16
    rng = np.random.default_rng(77)
    data = rng.integers(0, 256, size=(S, 6000), dtype=np.uint8)
    c = rng.integers(1, 7, size=S * F).astype(np.int32)
    f = np.ones(S * F, dtype=np.int32)
    in_len = rng.integers(0, 6000, size=S).astype(np.int32)
    in_len[0] = 0
    ref = ocfg.frame_batch(data, c, f, S, F, in_len=in_len)
    got = rx.fec_frame(data, c, f, S, F, in_len=in_len)
    assert ref["started"][0] == 0 and ref["started"].max() > 3
    _compare(got, ref)
    rx.close()


def test_fec_frame_tb_in_flight_continues_in_the_next_call():
    ocfg = make_oracle_cfg(cnsts=ALL, alists=(ALIST_300,))
    rx = make_receiver(cnsts=ALL, alists=(ALIST_300,))
    S, F = 5, 6
    rng = np.random.default_rng(78)
    state = None
    for call in range(4):
        data = rng.integers(0, 256, size=(S, F * 800), dtype=np.uint8)
        c = rng.integers(1, 7, size=S * F).astype(np.int32)
        f = np.ones(S * F, dtype=np.int32)
        ref = ocfg.frame_batch(data, c, f, S, F, state=state)
        state = ref["state"]
        _compare(rx.fec_frame(data, c, f, S, F), ref)
    rx.close()


@pytest.mark.parametrize("cnst,snr", [(2, 24.0), (4, 32.0)])
def test_fec_frame_roundtrip_through_the_receive_path(cnst, snr):
    """framer -> header formatter -> modulator -> flat channel + noise -> dtl_b200_rx_iq: the bytes of every TB completed
    inside the call come back, in order."""
    import torch
    from dtl_b200 import synth
    rx = make_receiver(cnsts=ALL, alists=(ALIST_300,), initial_mcs=cnst - 1)
    S, F = 64, 24
    rng = np.random.default_rng(79)
    data = rng.integers(0, 256, size=(S, F * 400), dtype=np.uint8)
    c = np.full(S * F, cnst, dtype=np.int32)
    f = np.ones(S * F, dtype=np.int32)
    fr = rx.fec_frame(data, c, f, S, F)
    hdr = synth.format_headers(fr["meta"], rx.header_len * len(rx.config.occupied_carriers[0]), True)
    iq = rx.tx(hdr, fr["chunks"], fr["meta"])
    g = np.random.default_rng(80)
    h = (1 + 1j) * np.exp(2j * np.pi * g.random(S))
    sd = np.sqrt(rx.fft_len * 2.0 / 10.0 ** (snr / 10.0)) / np.sqrt(2.0)
    iq = iq.reshape(S, -1) * h[:, None]
    iq = (iq + sd * (g.standard_normal(iq.shape) + 1j * g.standard_normal(iq.shape))).astype(np.complex64)
    got = rx.rx_iq(iq.reshape(S * F, -1), S, F)
    n_ok = 0
    for s in range(S):
        tbs = got["tbs"][s, :got["ntbs"][s]]
        ok = tbs["crc_ok"].astype(bool)
        assert ok.size > 0
        pm = int(tbs["nbytes"].max())
        want = data[s, :ok.size * pm].reshape(ok.size, pm)[ok].reshape(-1)
        assert np.array_equal(got["bytes"][s, :got["nbytes"][s]], want)
        n_ok += int(ok.sum())
    assert n_ok >= 0.9 * got["ntbs"].sum()
    rx.close()
