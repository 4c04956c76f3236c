"""The product's synthetic-traffic generator (gr-dtl_b200/dtl_b200/synth.py: TX framing restated for many streams)
against the oracle's TX framer and, on the GPU, through both receive implementations."""
import numpy as np
import pytest

import pyoracle as orc
from common import ALIST_100_27, ALIST_300, make_oracle_cfg, make_receiver


@pytest.mark.parametrize("cnst", [1, 2, 3, 4, 5, 6])
@pytest.mark.parametrize("alist,n,k", [(ALIST_300, 300, 152), (ALIST_100_27, 100, 27)])
def test_fec_plan_tags_match_oracle_framer(cnst, alist, n, k):
    from dtl_b200 import synth
    ocfg = make_oracle_cfg(alists=(alist,), initial_mcs=0, cnsts=(1, 2, 3, 4, 5, 6))
    F = 40
    t = ocfg.tx_batch(1, F, seed=3, cnst=cnst, fec_id=1, want_iq=False, want_taps=False)
    p = synth.fec_plan(F, ocfg.capacity, orc.bps(cnst), n, k, 1)
    m = p.meta.copy()
    m["constellation"] = cnst
    for f in m.dtype.names:
        assert np.array_equal(m[f], t["meta"][f]), f
    assert sum(1 for x in p.tb_done_frame if x >= 0) == int(t["payload_tbs"][0])


def test_bits_to_chunks_is_lsb_first_repack():
    from dtl_b200 import synth
    rng = np.random.default_rng(1)
    for bps in (1, 2, 3, 4, 6, 8):
        bits = rng.integers(0, 2, size=(3, 100), dtype=np.uint8)
        got = synth._bits_to_chunks(bits, bps)
        for r in range(3):
            assert np.array_equal(got[r], orc.repack(bits[r], 1, bps))


@pytest.mark.gpu
@pytest.mark.parametrize("cnst,snr,fec", [(2, 18.0, True), (4, 24.0, True), (3, 30.0, False), (6, 40.0, True)])
def test_synth_round_trip_gpu_and_oracle(cnst, snr, fec):
    """synth frames -> GPU modulator -> channel -> (a) the CUDA receive path, (b) the oracle: both return exactly the
    generated payload (and therefore agree with each other)."""
    from dtl_b200 import synth
    cn = (1, 2, 3, 4, 5, 6)
    alists = (ALIST_300,) if fec else ()
    rx = make_receiver(alists=alists, cnsts=cn, initial_mcs=cn.index(cnst))
    ocfg = make_oracle_cfg(alists=alists, cnsts=cn, initial_mcs=cn.index(cnst))
    S, F = 6, 9
    b = synth.make_batch(rx, S, F, cnst, fec_id=1 if fec else 0, seed=11, snr_db=snr, device=False)
    got = rx.rx(b["iq"], b["taps"], b["meta"], S, F)
    ref = ocfg.rx_batch(b["iq"], b["taps"], b["meta"], S, F)
    for s in range(S):
        gb = got["bytes"][s, :got["nbytes"][s]]
        assert np.array_equal(gb, ref["bytes"][s])
        if fec:      # CRC-passing TBs carry exactly what was generated
            tb = got["tbs"][s, :got["ntbs"][s]]
            assert len(tb) == b["n_done"]
            want = b["payload"][s].reshape(-1, b["tb_bytes"])[tb["crc_ok"].astype(bool)].reshape(-1)
            assert tb["crc_ok"].sum() >= len(tb) // 2
        else:
            want = b["payload"][s]
        assert np.array_equal(gb, want), "stream %d" % s
    assert np.array_equal(got["res"]["fb_cnst"], ref["res"]["fb_cnst"])
    rx.close()


@pytest.mark.parametrize("cnst", [1, 2, 3, 4, 5, 6])
def test_oracle_framer_with_caller_bytes_matches_fec_plan_and_is_continuous(cnst):
    """The oracle framer fed with the caller's bytes (the reference for dtl_b200_fec_frame, tests/test_gpu_fec_frame.py):
    its tags equal the product's framing plan, it consumes exactly the TB payloads it started, and F frames in one call
    equal the same frames in two calls (the TB in flight continues)."""
    from dtl_b200 import synth
    ocfg = make_oracle_cfg(alists=(ALIST_300,), initial_mcs=0, cnsts=(1, 2, 3, 4, 5, 6))
    S, F = 3, 12
    rng = np.random.default_rng(100 + cnst)
    data = rng.integers(0, 256, size=(S, F * 1200), dtype=np.uint8)
    c = np.full(S * F, cnst, dtype=np.int32)
    f = np.ones(S * F, dtype=np.int32)
    one = ocfg.frame_batch(data, c, f, S, F)
    p = synth.fec_plan(F, ocfg.capacity, orc.bps(cnst), 300, 152, 1)
    m = p.meta.copy()
    m["constellation"] = cnst
    for s in range(S):
        for name in m.dtype.names:
            assert np.array_equal(one["meta"][name].reshape(S, F)[s], m[name]), name
    assert np.all(one["started"] == len(p.tb_bits))
    assert np.all(one["consumed"] == sum(p.tb_payload_bytes))
    state, pos, parts = None, np.zeros(S, dtype=np.int64), []
    for h in range(2):
        d = np.stack([data[s, pos[s]:pos[s] + 8000] for s in range(S)])
        r = ocfg.frame_batch(d, c.reshape(S, F)[:, h * 6:(h + 1) * 6].reshape(-1), np.ones(S * 6, dtype=np.int32), S, 6, state=state)
        state = r["state"]
        pos += r["consumed"]
        parts.append(r["chunks"].reshape(S, 6, -1))
    assert np.array_equal(np.concatenate(parts, axis=1).reshape(S * F, -1), one["chunks"])
    assert np.array_equal(pos, one["consumed"])
