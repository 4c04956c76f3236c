"""The reference's own FEC round-trip scenario on the GPU (python/dtl/qa_ofdm_adaptive_fec.py:60-61,71-171) and a stream
whose FEC id switches between the two codes (lib/dtl/tb_decoder.cc:92-133).

  dtl_b200_fec_frame -> dtl_b200_map -> dtl_b200_soft_demap (no noise tag: sigma 0.001) -> dtl_b200_fec_decode

with BOTH n=100 codes (n_0100_k_0023_gap_10, n_0100_k_0027_gap_04) in ONE decoder list, fec id 1 and 2, frame_len 3 and 10
x 48 carriers, BPSK / QPSK / 8PSK / 16QAM, int(k * 69.69) input items of value 0/1.  See tests/test_oracle_golden.py
(test_fec_chain_qa_scenario) for what is restated and what differs from the qa file (byte input, whole-TB framing).
Covers on the product side: a handle with n_codes = 2, the mixed-code branch of ldpc_screen_kernel, a two-code
ldpc_bp_kernel launch, tb_plan_kernel's hand-over of `info_fec` when the FEC id changes, +-inf LLRs, the n=100 geometry.
"""
import numpy as np
import pytest

import pyoracle as po
from common import ALIST_100_23, ALIST_100_27, make_oracle_cfg, make_receiver
from test_oracle_golden import qa_fec_geometry, qa_fec_input

pytestmark = pytest.mark.gpu

CODES = (ALIST_100_23, ALIST_100_27)
FIELDS = ("constellation", "frame_no", "payload_len", "packet_len", "fec_id", "tb_no", "tb_offset", "tb_payload")


@pytest.mark.parametrize("frame_len", [3, 10])
def test_qa_fec_round_trip_all_constellations_both_codes(frame_len):
    """eight streams in one call: (BPSK, QPSK, 8PSK, 16QAM) x (fec 1, fec 2); every stream gets the qa's input length"""
    ocfg = make_oracle_cfg(frame_len=frame_len, alists=CODES)
    rx = make_receiver(frame_len=frame_len, alists=CODES)
    assert rx.frame_capacity == frame_len * 48
    combos = [(c, f) for c in (1, 2, 3, 4) for f in (1, 2)]
    S = len(combos)
    datas = [qa_fec_input(ocfg, f, 7 * c + f) for c, f in combos]
    geo = [qa_fec_geometry(ocfg, c, f) for c, f in combos]
    ntb = [d.size // g[0] for d, g in zip(datas, geo)]
    F = max((n * (g[1] + 8) + g[2] - 1) // g[2] + 2 for n, g in zip(ntb, geo))
    in_cap = max(d.size for d in datas)
    data = np.zeros((S, in_cap), dtype=np.uint8)
    in_len = np.zeros(S, dtype=np.int32)
    for s, d in enumerate(datas):
        data[s, :d.size] = d
        in_len[s] = d.size
    cn = np.repeat(np.array([c for c, _ in combos], dtype=np.int32), F)
    fe = np.repeat(np.array([f for _, f in combos], dtype=np.int32), F)
    ref = ocfg.frame_batch(data, cn, fe, S, F, in_len=in_len)
    fr = rx.fec_frame(data, cn, fe, S, F, in_len=in_len)
    for k in FIELDS:
        assert np.array_equal(fr["meta"][k], ref["meta"][k]), k
    assert np.array_equal(fr["chunks"], ref["chunks"]) and np.array_equal(fr["consumed"], ref["consumed"])
    syms = rx.chunks_to_symbols(fr["chunks"], fr["meta"])                       # ofdm_adaptive_chunks_to_symbols_bc
    llr = rx.soft_demap(syms, fr["meta"], noise=None)                            # soft_cf, no noise tag: sigma = 0.001 (:89)
    for s, (c, _) in enumerate(combos):
        m = fr["meta"][s * F]
        # saturated: +-inf where the demapper sums exponentials (8PSK, 16QAM: expf underflows as in the reference), a finite
        # |LLR| ~ 4000 of the same sign where BPSK / QPSK use the linear form of the same ratio
        assert np.all(np.abs(llr[s * F, :m["packet_len"] * po.bps(c)]) > 1000.0), "LLRs must saturate without a noise tag"
    # decoded in several calls of <= 512 frames per stream (a call releases at most 1024 TBs per stream; the TB in flight
    # at the end of a call is continued by the next one)
    out, recs = [[] for _ in range(S)], [[] for _ in range(S)]
    llr3, meta2 = llr.reshape(S, F, -1), fr["meta"].reshape(S, F)
    for j0 in range(0, F, 512):
        j1 = min(F, j0 + 512)
        got = rx.fec_decode(np.ascontiguousarray(llr3[:, j0:j1]).reshape(S * (j1 - j0), -1),
                            np.ascontiguousarray(meta2[:, j0:j1]).reshape(-1), S, j1 - j0, bytes_cap=in_cap + 64, tb_cap=1030)
        for s in range(S):
            out[s].append(got["bytes"][s, :got["nbytes"][s]].copy())
            recs[s].append(got["tbs"][s, :got["ntbs"][s]].copy())
    for s, ((c, f), d, g, n) in enumerate(zip(combos, datas, geo, ntb)):
        assert fr["consumed"][s] == n * g[0] and d.size - fr["consumed"][s] < g[0]
        tbs = np.concatenate(recs[s])
        assert tbs.size == n and tbs["crc_ok"].all() and not tbs["avg_it"].any(), (c, f)
        assert np.array_equal(np.concatenate(out[s]), d[:n * g[0]]), "qa :140-141, cnst %d fec %d" % (c, f)
    rx.close()


def _noisy_chain(ocfg, rx, S, F, cn, fe, noise_var, seed):
    """framer on both sides, then symbols + AWGN with a noise tag; returns (meta, oracle LLRs [S*F, stride], GPU LLRs)"""
    rng = np.random.default_rng(seed)
    data = rng.integers(0, 256, size=(S, F * 64), dtype=np.uint8)
    ref = ocfg.frame_batch(data, cn, fe, S, F)
    fr = rx.fec_frame(data, cn, fe, S, F)
    for k in FIELDS:
        assert np.array_equal(fr["meta"][k], ref["meta"][k]), k
    assert np.array_equal(fr["chunks"], ref["chunks"])
    meta, cap, mb = fr["meta"], rx.frame_capacity, rx.max_bps
    syms = rx.chunks_to_symbols(fr["chunks"], meta)
    sd = np.sqrt(noise_var / 2.0)
    syms = (syms + sd * (rng.standard_normal(syms.shape) + 1j * rng.standard_normal(syms.shape))).astype(np.complex64)
    noise = np.full(S * F, noise_var, dtype=np.float64)
    g_llr = rx.soft_demap(syms, meta, noise=noise)
    o_llr = np.zeros((S * F, cap * mb), dtype=np.float32)
    for f in range(S * F):
        c = int(meta["constellation"][f])
        v = po.soft_dec(c, syms[f], float(np.sqrt(noise_var)))      # soft_cf_impl.cc:100-101: npwr := sqrt(noise tag)
        o_llr[f, :v.size] = v
    return data, meta, o_llr, g_llr


@pytest.mark.parametrize("cnst,noise_var", [(2, 0.12), (2, 0.2), (4, 0.12)])
def test_fec_id_switches_inside_a_stream_two_code_bp(cnst, noise_var):
    """The FEC id (the MCS feedback) changes from frame to frame between the two n=100 codes while transport blocks are in
    flight, with noise so that belief propagation runs: TB records (tb_no, crc_ok, nbytes, avg_it, frame_index) and payload
    bytes bit-exact against the oracle's tb_decoder on identical LLRs; the LLRs themselves within 1e-4."""
    from common import assert_llr_close
    frame_len = 10
    ocfg = make_oracle_cfg(frame_len=frame_len, alists=CODES)
    rx = make_receiver(frame_len=frame_len, alists=CODES)
    S, F = 12, 40
    rng = np.random.default_rng(5 + cnst)
    fe = rng.integers(1, 3, size=S * F).astype(np.int32)
    fe[:F] = 1
    fe[F:2 * F] = 2                                             # two streams that never switch, ten that do
    cn = np.full(S * F, cnst, dtype=np.int32)
    data, meta, o_llr, g_llr = _noisy_chain(ocfg, rx, S, F, cn, fe, noise_var, seed=11 + cnst)
    assert len(set(meta["fec_id"][2 * F:3 * F])) == 2, "the framer never switched codes: vacuous test"
    b = po.bps(cnst)
    for f in range(S * F):
        n = int(meta["packet_len"][f]) * b
        assert_llr_close(g_llr[f, :n], o_llr[f, :n])
    got = rx.fec_decode(o_llr[:, :rx.llr_stride].copy(), meta, S, F)
    n_it = n_ok = n_all = 0
    for s in range(S):
        dec = po.FecDecoder(ocfg)
        by, recs = [], []
        for j in range(F):
            o, t = dec.frame(o_llr[s * F + j], meta[s * F + j])
            by.append(o)
            recs.append(t)
        recs = np.concatenate(recs)
        g = got["tbs"][s, :got["ntbs"][s]]
        assert np.array_equal(g, recs), "TB records differ from the oracle (stream %d)" % s
        assert np.array_equal(got["bytes"][s, :got["nbytes"][s]], np.concatenate(by))
        n_it += int(recs["avg_it"].sum()); n_ok += int(recs["crc_ok"].sum()); n_all += recs.size
    assert n_it > 0, "noise too low: belief propagation never iterated"
    assert 0 < n_ok, "noise too high: no transport block decoded"
    rx.close()


def test_two_code_handle_plain_decoder_matches_oracle():
    """fec_dec::decode batches of both codes from one handle (make_ldpc_decoders with two alists)"""
    ocfg = make_oracle_cfg(frame_len=10, alists=CODES)
    rx = make_receiver(frame_len=10, alists=CODES)
    rng = np.random.default_rng(3)
    for fec in (1, 2):
        info = ocfg.code_info(fec)
        data = rng.integers(0, 2, (256, info["K"]), dtype=np.uint8)
        cw = ocfg.ldpc_encode(fec, data)
        assert np.array_equal(rx.ldpc_encode(fec, data), cw)
        llr = ((2.0 * cw - 1.0) * 2.0 + 1.6 * rng.standard_normal(cw.shape)).astype(np.float32)
        llr[:8] = np.where(cw[:8] == 1, np.inf, -np.inf)          # saturated input (the qa scenario's LLRs)
        ob, oi = ocfg.ldpc_decode(fec, llr)
        gb, gi = rx.ldpc_decode(fec, llr)
        assert np.array_equal(gb, ob) and np.array_equal(gi, oi) and oi.max() > 0
    rx.close()


def test_fec_decode_device_llrs_at_any_alignment_and_row_stride():
    """dtl_b200_fec_decode on DEVICE LLRs whose base is not 16-byte aligned and whose row stride is not a multiple of four
    floats: the decoder fetches a queued codeword's run with ONE bulk copy of its 16-byte-aligned superset where that stays
    inside the row and through the gathered block otherwise -- TB records and bytes equal the host-buffer call's (which equal
    the oracle's, test above), for every lead-in 0..3."""
    import torch
    frame_len = 10
    ocfg = make_oracle_cfg(frame_len=frame_len, alists=CODES)
    rx = make_receiver(frame_len=frame_len, alists=CODES)
    S, F = 12, 40
    rng = np.random.default_rng(9)
    fe = rng.integers(1, 3, size=S * F).astype(np.int32)
    cn = np.full(S * F, 2, dtype=np.int32)
    data, meta, o_llr, g_llr = _noisy_chain(ocfg, rx, S, F, cn, fe, 0.12, seed=21)
    llr = o_llr[:, :rx.llr_stride].copy()
    rx.reset(S)
    ref = rx.fec_decode(llr, meta, S, F)
    assert sum(int(ref["tbs"][s, :ref["ntbs"][s]]["avg_it"].sum()) for s in range(S)) > 0, "belief propagation never iterated"
    tm = torch.from_numpy(meta.view(np.uint8).reshape(S * F, -1).copy()).cuda()
    for front, extra in ((0, 0), (1, 1), (2, 3), (3, 2)):
        w = rx.llr_stride + extra
        buf = torch.full((front + S * F * w + 8,), float("nan"), dtype=torch.float32, device="cuda")
        view = buf[front:front + S * F * w].view(S * F, w)
        view[:, :rx.llr_stride] = torch.from_numpy(llr).cuda()
        assert view.data_ptr() % 16 == (4 * front) % 16 and view.is_contiguous()
        rx.reset(S)
        got = rx.fec_decode(view, tm, S, F)
        rx.sync()
        nt = got["ntbs"].cpu().numpy()
        nb = got["nbytes"].cpu().numpy()
        assert np.array_equal(nt, ref["ntbs"]) and np.array_equal(nb, ref["nbytes"]), (front, extra)
        from dtl_b200._capi import tb_result_dtype
        tb = got["tbs"].cpu().numpy().view(tb_result_dtype).reshape(S, -1)
        by = got["bytes"].cpu().numpy().reshape(S, -1)
        for s in range(S):
            assert tb[s, :nt[s]].tobytes() == ref["tbs"][s, :nt[s]].tobytes(), (front, extra, s)
            assert np.array_equal(by[s, :nb[s]], ref["bytes"][s, :nb[s]]), (front, extra, s)
    rx.close()
