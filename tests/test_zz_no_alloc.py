"""dtl_b200_equalize_noalloc (include/dtl_b200.h): the equaliser pass for callers that cannot call cudaMalloc -- its scratch is a
static device array inside the library.  No counterpart in the reference; checked against the existing entry point
dtl_b200_equalize on the same device inputs, two calls in a row (the second with other data and fewer frames, so stale scratch
contents of the first would show), plus the size limit and the no-resize rule.  Collected last on purpose (file name)."""
import numpy as np
import pytest

import pyoracle as orc
from common import ALIST_300, assert_llr_close, make_oracle_cfg, make_receiver

pytestmark = pytest.mark.gpu

MAX_FLOATS = 524288            # DTL_B200_NOALLOC_MAX_PILOT_FLOATS


def _case(ocfg, S, F, seed, cnsts, snr):
    """frequency-domain frames of S streams, the constellation changing from stream to stream"""
    parts = [ocfg.tx_batch(1, F, seed=seed + s, cnst=cnsts[s % len(cnsts)], fec_id=1, snr_db=snr[s % len(snr)]) for s in range(S)]
    iq = np.concatenate([p["iq"] for p in parts])
    taps = np.concatenate([p["taps"] for p in parts])
    meta = np.concatenate([p["meta"] for p in parts])
    return ocfg.rx_demod(iq).reshape(S * F, ocfg.frame_len, ocfg.fft_len), taps, meta


def test_equalize_noalloc_equals_equalize_two_calls_in_a_row():
    import torch
    import dtl_b200
    ocfg = make_oracle_cfg(cnsts=(1, 2, 3, 4, 5, 6), alists=(ALIST_300,), initial_mcs=1)
    rx = make_receiver(cnsts=(1, 2, 3, 4, 5, 6), alists=(ALIST_300,), initial_mcs=1)
    dev = torch.device("cuda:0")
    up = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).to(dev)
    for S, F, seed, cn, snr in ((6, 10, 800, (4, 2, 6, 1, 3, 5), (22.0, 15.0, 36.0, 9.0, 19.0, 30.0)),
                                (3, 7, 900, (3, 4, 1), (17.0, 20.0, 6.0))):
        freq, taps, meta = _case(ocfg, S, F, seed, cn, snr)
        meta = meta.copy()
        meta["frame_no"][F + 2] += 3                      # a gap in one stream: lost_frames / feedback state must agree too
        d_freq, d_taps, d_meta = up(freq), up(taps), up(meta)
        n = S * F
        # the existing entry point on the same device buffers
        rx.reset(S)
        ref = rx.equalize(d_freq, d_taps, d_meta, S, F, want=("res", "llr"))
        rx.sync()
        # the new one: every buffer exists before the call
        out = dict(res=rx._new(dtl_b200.MEM_DEVICE, (n,), dtl_b200.frame_result_dtype),
                   llr=rx._new(dtl_b200.MEM_DEVICE, (n, rx.llr_stride), np.float32),
                   n_llr=rx._new(dtl_b200.MEM_DEVICE, (n,), np.int32))
        out["llr"].fill_(float("nan"))
        rx.reset(S)
        rx.sync()
        before, free_before = torch.cuda.memory_allocated(dev), torch.cuda.mem_get_info(dev)[0]
        rx.equalize_noalloc(d_freq, d_taps, d_meta, S, F, out)
        rx.sync()
        # nothing allocated through torch, and the driver's free memory did not shrink (a cudaMalloc in the library would show)
        assert torch.cuda.memory_allocated(dev) == before and torch.cuda.mem_get_info(dev)[0] >= free_before
        to_np = lambda t, dt: t.cpu().numpy().view(dt).reshape(-1)
        r0, r1 = to_np(ref["res"], dtl_b200.frame_result_dtype), to_np(out["res"], dtl_b200.frame_result_dtype)
        for k in ("fb_cnst", "fb_fec", "emitted", "lost_frames"):
            assert np.array_equal(r0[k], r1[k]), k
        assert r1["lost_frames"].max() >= 3
        assert np.allclose(r1["snr_db"], r0["snr_db"], rtol=1e-9, atol=0) and np.allclose(r1["noise"], r0["noise"], rtol=1e-9, atol=0)
        n0, n1 = to_np(ref["n_llr"], np.int32), to_np(out["n_llr"], np.int32)
        assert np.array_equal(n0, n1) and n1.min() > 0
        l0 = ref["llr"].cpu().numpy().reshape(n, -1)
        l1 = out["llr"].cpu().numpy().reshape(n, -1)
        for f in range(n):
            assert_llr_close(l1[f, :n1[f]], l0[f, :n0[f]])       # the tolerance of test_gpu_parity.py (1e-4 relative)
    rx.close()


def test_equalize_noalloc_rejects_what_it_cannot_do_without_allocating():
    import torch
    import dtl_b200
    rx = make_receiver(alists=(ALIST_300,), initial_mcs=1)
    dev = torch.device("cuda:0")
    npil = 80                                                    # 4 pilot carriers x 20 symbols
    F = 8
    S = MAX_FLOATS // (npil * F) + 1                             # one stream beyond the static scratch
    n = S * F
    z = lambda nbytes: torch.zeros(nbytes, dtype=torch.uint8, device=dev)
    d_freq, d_taps = z(n * 20 * 64 * 8), z(n * 64 * 8)
    d_meta = z(n * dtl_b200.frame_meta_dtype.itemsize)
    out = dict(res=rx._new(dtl_b200.MEM_DEVICE, (n,), dtl_b200.frame_result_dtype))
    rx.reset(S)
    with pytest.raises(dtl_b200.DtlError) as ei:
        rx.equalize_noalloc(d_freq, d_taps, d_meta, S, F, out)
    assert ei.value.status == dtl_b200._capi.ERR_CAPACITY and "static" in str(ei.value)
    # per-stream state sized for another stream count: the entry point refuses instead of resizing (which would allocate)
    with pytest.raises(dtl_b200.DtlError) as ei:
        rx.equalize_noalloc(d_freq, d_taps, d_meta, S - 1, F, out)
    assert ei.value.status == dtl_b200._capi.ERR_INVALID_ARG and "dtl_b200_reset" in str(ei.value)
    # and one stream fewer fits
    rx.reset(S - 1)
    rx.equalize_noalloc(d_freq, d_taps, d_meta, S - 1, F, out)      # all-zero tags: every frame rejected on the device (error word)
    with pytest.raises(dtl_b200.DtlError):
        rx.sync()
    rx.close()
