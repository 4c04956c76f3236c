"""Shared helpers of the test-suite: builds the SAME stream configuration for the oracle (test infrastructure,
oracle/pyoracle.py) and for the product (gr-dtl_b200/dtl_b200)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
ALIST_300 = os.path.join(GOLDEN, "n_0300_k_0152_gap_03.alist")
ALIST_100_23 = os.path.join(GOLDEN, "n_0100_k_0023_gap_10.alist")
ALIST_100_27 = os.path.join(GOLDEN, "n_0100_k_0027_gap_04.alist")

DBL_MIN = sys.float_info.min


def mcs_table(cnsts, fec, lut="config"):
    """(snr threshold, (constellation, fec id)) rows.  lut "config": the thresholds of examples/config.json:3-6 (13 / 18 / 23 dB;
    every test that does not say otherwise, FEC or not); lut "config_fec": examples/config_fec.json:4-7 (11 / 16 / 21 dB, what
    BASELINE.json configs[1] ships with).  The 64-/256-QAM rows (+5 dB each) are this repo's extension."""
    th = {1: DBL_MIN, 2: 13, 3: 18, 4: 23, 5: 28, 6: 33} if lut == "config" else {1: DBL_MIN, 2: 11, 3: 16, 4: 21, 5: 26, 6: 31}
    rows = [(th[c], (c, 1 if fec else 0)) for c in sorted(cnsts)]
    rows[0] = (DBL_MIN, rows[0][1])
    return rows


def make_oracle_cfg(fft_len=64, frame_len=20, alists=(), cnsts=(1, 2, 3, 4), initial_mcs=0, cp_len=None, lut="config"):
    import pyoracle as orc
    return orc.Config(fft_len=fft_len, cp_len=cp_len, frame_len=frame_len, alists=list(alists),
                      mcs=mcs_table(cnsts, bool(alists), lut), initial_mcs=initial_mcs, constellations=tuple(cnsts))


def make_product_cfg(fft_len=64, frame_len=20, alists=(), cnsts=(1, 2, 3, 4), initial_mcs=0, cp_len=None, lut="config"):
    import dtl_b200
    kw = {}
    if fft_len != 64:
        occ, pil = dtl_b200.scaled_phy_carriers(fft_len)
        kw = dict(occupied_carriers=occ, pilot_carriers=pil)
    names = ["fec_%d" % (i + 1) for i in range(len(alists))]
    mcs = tuple((snr, (dtl_b200.constellation_type_t(c), names[f - 1] if f else "no_fec"))
                for snr, (c, f) in mcs_table(cnsts, bool(alists), lut))
    return dtl_b200.ofdm_adaptive_rx_config(fft_len=fft_len, cp_len=cp_len if cp_len is not None else fft_len // 4,
                                            frame_length=frame_len, fec=bool(alists),
                                            fec_codes=tuple(zip(names, alists)), mcs=mcs, initial_mcs_id=initial_mcs, **kw)


def make_receiver(**kw):
    import dtl_b200
    extra = {k: kw.pop(k) for k in ("ldpc_mode", "ldpc_sigma2", "ldpc_max_iterations", "device") if k in kw}
    return dtl_b200.ofdm_receiver(make_product_cfg(**kw), **extra)


def assert_llr_close(got, ref, rtol=1e-4, sat=60.0):
    """LLR comparison with the stated tolerance: |got-ref| <= rtol*max(1,|ref|) wherever |ref| < sat; beyond
    that both sides are in expf's subnormal/underflow range (sums of a few ulps, +-inf, or inf-inf = NaN), where
    only sign and saturation are compared."""
    got = np.asarray(got, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    assert got.shape == ref.shape
    fin = np.isfinite(ref) & (np.abs(ref) < sat)
    err = np.abs(got[fin] - ref[fin])
    tol = rtol * np.maximum(1.0, np.abs(ref[fin]))
    assert np.all(err <= tol), "LLR mismatch: max err %g (tol %g), %d of %d" % (err.max(), tol.min(), (err > tol).sum(), err.size)
    big = ~fin & ~np.isnan(ref)
    g = got[big]
    assert np.all(np.isnan(g) | ((np.sign(g) == np.sign(ref[big])) & (np.abs(g) > 0.5 * sat))), "saturated LLR sign mismatch"
