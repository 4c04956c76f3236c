"""The decoder BASELINE.json's north_star names -- layered normalised min-sum, fp32, one codeword per warp, warp-vote exit
(DTL_LDPC_LAYERED, gr-dtl_b200/csrc/bp_layered.cuh) -- through the whole transport-block chain, beside the reference-arithmetic
decoder (DTL_LDPC_REFERENCE) on the same input.

What can be asserted about a decoder with a different schedule and arithmetic:
  * every transport block that passes its CRC-32 carries exactly the transmitted bytes, in order (no false accepts: the CRC
    and the zero-syndrome exit are the same on both decoders);
  * the set of passing TBs contains reference mode's set minus a small measured fraction (stated per case below; on the
    operating points tested the layered decoder passes MORE blocks than the reference one, whose LLR-domain arithmetic treats
    the demapper's sigma-scaled LLRs as true LLRs);
  * the frame-level outputs (feedback decisions, SNR tags, lost-frame counters) do not depend on the decoder at all.
"""
import numpy as np
import pytest

import dtl_b200
from common import ALIST_300, ALIST_100_23, ALIST_100_27, make_receiver
from test_gpu_configs import _check_roundtrip, _np

pytestmark = pytest.mark.gpu

ALL = (1, 2, 3, 4, 5, 6)
# (name, receiver configuration, constellation, SNR grid [dB], S, F, largest tolerated loss of reference-mode passing TBs)
CASES = [
    ("config2_qpsk", dict(alists=(ALIST_300,), initial_mcs=1, lut="config_fec"), 2, (16.0, 17.0, 19.0), 128, 24, 0.02),
    ("config2_16qam", dict(alists=(ALIST_300,), initial_mcs=3, lut="config_fec"), 4, (20.0, 22.0, 25.0), 128, 24, 0.02),
    ("config3_bpsk", dict(cnsts=ALL, alists=(ALIST_300,)), 1, (9.0, 11.0, 14.0), 96, 16, 0.02),
    ("config3_8psk", dict(cnsts=ALL, alists=(ALIST_300,), initial_mcs=2), 3, (16.0, 18.0, 22.0), 96, 16, 0.02),
    ("config3_64qam", dict(cnsts=ALL, alists=(ALIST_300,), initial_mcs=4), 5, (30.0, 33.0, 38.0), 96, 16, 0.02),
    ("config4_16qam_shard", dict(alists=(ALIST_300,), initial_mcs=3), 4, (22.0,), 512, 32, 0.02),
    ("config5_fft1024_256qam", dict(fft_len=1024, frame_len=5, cnsts=ALL, alists=(ALIST_300,), initial_mcs=5), 6, (33.0, 36.0, 40.0, 46.0), 48, 16, 0.02),
]


@pytest.mark.parametrize("name,kw,cnst,snrs,S,F,max_loss", CASES, ids=[c[0] for c in CASES])
def test_layered_through_the_tb_chain(name, kw, cnst, snrs, S, F, max_loss):
    from dtl_b200 import synth, tb_result_dtype
    rx_ref = make_receiver(ldpc_mode=dtl_b200.LDPC_REFERENCE, **kw)
    rx_lay = make_receiver(ldpc_mode=dtl_b200.LDPC_LAYERED, **kw)
    seen_fail = False
    for i, snr in enumerate(snrs):
        b = synth.make_batch(rx_ref, S, F, cnst, fec_id=1, seed=9000 + 17 * cnst + i, snr_db=snr)
        rx_ref.reset(S)
        rx_lay.reset(S)
        g_ref = rx_ref.rx(b["iq"], b["taps"], b["meta"], S, F)
        g_lay = rx_lay.rx(b["iq"], b["taps"], b["meta"], S, F)
        rx_ref.sync()
        rx_lay.sync()
        _check_roundtrip(g_lay, b, S, True, min_ok=0.0)                 # every CRC-passing TB carries the transmitted bytes
        _check_roundtrip(g_ref, b, S, True, min_ok=0.0)
        t_ref = _np(g_ref["tbs"]).view(tb_result_dtype).reshape(S, -1)
        t_lay = _np(g_lay["tbs"]).view(tb_result_dtype).reshape(S, -1)
        n = _np(g_ref["ntbs"])
        assert np.array_equal(n, _np(g_lay["ntbs"]))
        ok_ref = np.concatenate([t_ref[s, :n[s]]["crc_ok"] for s in range(S)]).astype(bool)
        ok_lay = np.concatenate([t_lay[s, :n[s]]["crc_ok"] for s in range(S)]).astype(bool)
        lost = int((ok_ref & ~ok_lay).sum())
        assert lost <= max_loss * max(1, int(ok_ref.sum())), \
            "%s @ %g dB: layered fails %d of the %d TBs reference mode decodes (%d vs %d pass of %d)" % (
                name, snr, lost, ok_ref.sum(), ok_lay.sum(), ok_ref.sum(), ok_ref.size)
        seen_fail |= bool((~ok_lay).any())
        r_ref, r_lay = _np(g_ref["res"]).view(dtl_b200.frame_result_dtype), _np(g_lay["res"]).view(dtl_b200.frame_result_dtype)
        assert r_ref.tobytes() == r_lay.tobytes(), "frame-level results must not depend on the decoder"
    if len(snrs) > 1:
        assert seen_fail, "SNR grid too comfortable: the decoder was never challenged"
    rx_ref.close()
    rx_lay.close()


def test_layered_two_codes_plain_decoder():
    """fec_dec::decode batches in layered mode on the two n=100 codes (row weights 5 and 4, one static pad column each)"""
    import pyoracle as po
    from common import make_oracle_cfg
    ocfg = make_oracle_cfg(frame_len=10, alists=(ALIST_100_23, ALIST_100_27))
    rx = make_receiver(frame_len=10, alists=(ALIST_100_23, ALIST_100_27), ldpc_mode=dtl_b200.LDPC_LAYERED)
    rng = np.random.default_rng(4)
    for fec in (1, 2):
        info = ocfg.code_info(fec)
        data = rng.integers(0, 2, (512, info["K"]), dtype=np.uint8)
        cw = ocfg.ldpc_encode(fec, data)
        sig = 1.1
        llr = (2.0 * ((2.0 * cw - 1.0) + sig * rng.standard_normal(cw.shape)) / sig ** 2).astype(np.float32)
        ref_bits, _ = ocfg.ldpc_decode(fec, llr)
        bits, nit = rx.ldpc_decode(fec, llr)
        ok_ref, ok = (ref_bits == data).all(axis=1), (bits == data).all(axis=1)
        assert ok_ref.sum() > 200 and ok.sum() >= ok_ref.sum() - 5 and nit.max() > 0
        # a word the decoder reports as converged (nit < max) is a codeword: its systematic part re-encodes to a zero syndrome
        conv = nit < 15
        recw = ocfg.ldpc_encode(fec, bits[conv])
        assert all(ocfg.ldpc_syndrome(fec, c) == 0 for c in recw[:64])
    rx.close()
