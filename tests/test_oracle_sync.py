"""The oracle's synchronisation front end on its own (oracle/src/sync.cc; CPU): what the GPU parity test in test_gpu_sync.py
leans on.  gr-digital's ofdm_sync_sc_cfb is not under the reference tree (system GNU Radio), so
these are property tests of the restatement -- the known answers of the Schmidl & Cox estimator on a synthetic stream --
plus the literal scenarios of lib/dtl/ofdm_adaptive_frame_detect_bb_impl.cc:64-156 for the trigger regulariser."""
import numpy as np

import pyoracle as orc
from common import ALIST_300, make_oracle_cfg


def _stream(ocfg, F, seed, eps, lead, drop=(), snr=26.0):
    t = ocfg.tx_batch(1, F, seed=seed, cnst=orc.QPSK, fec_id=1, snr_db=snr)
    rng = np.random.default_rng(seed)
    noise = lambda n: (0.02 * (rng.standard_normal(n) + 1j * rng.standard_normal(n))).astype(np.complex64)
    parts = [noise(lead)]
    for j in range(F):
        parts.append(noise(ocfg.frame_samples) if j in drop else t["iq"][j])
    parts.append(noise(3000))
    x = np.concatenate(parts)
    x = (x * np.exp(2j * np.pi * eps * np.arange(x.size) / ocfg.fft_len)).astype(np.complex64)
    return t, x


def _run(ocfg, x, chunk):
    sy = orc.StreamSync(ocfg, 0.95)
    out = dict(starts=[], frames=[], raw=[], fixed=[])
    for c0 in range(0, x.size, chunk):
        r = sy.push(x[c0:c0 + chunk])
        out["starts"] += [int(v) for v in r["starts"]]
        out["frames"] += [f for f in r["frames"]]
        out["raw"] += [int(v) for v in r["raw"]]
        out["fixed"] += [int(v) for v in r["fixed"]]
    return out


def test_trigger_position_cfo_removal_and_payload():
    """the plateau's middle = the middle of the first symbol's cyclic prefix, delayed by N + cp; arg(P) = pi * eps, so the
    mixer leaves no fractional CFO: headers parse and the payload decodes from the frames as cut"""
    ocfg = make_oracle_cfg(alists=(ALIST_300,), initial_mcs=1)
    F, lead, eps = 8, 777, 0.23
    t, x = _stream(ocfg, F, 5, eps, lead)
    D, fs = ocfg.fft_len + ocfg.cp_len, ocfg.frame_samples
    r = _run(ocfg, x, 5000)
    assert len(r["starts"]) >= F
    off = np.array(r["starts"][:F]) - D - (lead + fs * np.arange(F))
    assert (np.abs(off + ocfg.cp_len // 2) <= 3).all(), off
    fr = np.array(r["frames"][:F])
    h = ocfg.rx_headers(fr, 1, F)
    assert h["ok"].all()
    ref = ocfg.rx_batch(fr, h["taps"], h["meta"], 1, F)
    got = ref["bytes"][0]
    assert got.size > 0 and np.array_equal(got, t["payload"][0][:got.size])
    assert ref["tbs"][0]["crc_ok"].all()


def test_chunking_changes_nothing_but_the_regulariser_windows():
    """raw triggers and the samples of a frame cut at the same start do not depend on how the stream was chunked (the
    moving sums are summed in an order fixed by the global sample index); fix_sync's window IS the chunk, as in the reference,
    so only its insertions may differ -- with every frame present there are none before the stream's end"""
    ocfg = make_oracle_cfg(alists=(ALIST_300,), initial_mcs=1)
    F = 7
    _, x = _stream(ocfg, F, 11, -0.18, 333)
    a, b, c = _run(ocfg, x, 4096), _run(ocfg, x, 10007), _run(ocfg, x, x.size)
    assert a["raw"][:F] == b["raw"][:F] == c["raw"][:F]
    assert a["starts"][:F] == b["starts"][:F] == c["starts"][:F]
    for j in range(F):
        assert np.array_equal(a["frames"][j], b["frames"][j]) and np.array_equal(a["frames"][j], c["frames"][j])


def test_regulariser_inserts_the_missing_trigger_once_in_sync():
    """frame_detect_bb_impl.cc:123-137: in sync (more than three triggers at the frame period), a frame without a trigger gets
    one inserted at last + frame_samples; its header fails and the frame after it is found again by the detector"""
    ocfg = make_oracle_cfg(alists=(ALIST_300,), initial_mcs=1)
    F, lead, lost = 12, 500, 7
    _, x = _stream(ocfg, F, 3, 0.1, lead, drop={lost})
    fs = ocfg.frame_samples
    r = _run(ocfg, x, 9000)
    fixed, raw = np.array(r["fixed"]), np.array(r["raw"])
    assert raw.size >= F - 1 and fixed.size >= F
    inserted = sorted(set(fixed.tolist()) - set(raw.tolist()))
    assert inserted and inserted[0] == fixed[lost - 1] + fs
    assert (np.diff(fixed[:F]) == fs).all()
    h = ocfg.rx_headers(np.array(r["frames"][:F]), 1, F)
    assert not h["ok"][lost] and h["ok"][:lost].all() and h["ok"][lost + 1:F].all()


def test_not_in_sync_a_lost_frame_is_just_lost():
    """before d_in_sync (fewer than four regular triggers) nothing is inserted: frame_detect_bb_impl.cc:88-97"""
    ocfg = make_oracle_cfg(alists=(ALIST_300,), initial_mcs=1)
    F = 6
    _, x = _stream(ocfg, F, 8, 0.05, 600, drop={2})
    r = _run(ocfg, x, x.size)
    fixed = np.array(r["fixed"])
    fs = ocfg.frame_samples
    assert fixed.size >= F - 1 and fixed[2] - fixed[1] == 2 * fs
