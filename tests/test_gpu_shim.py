"""The GNU Radio block shims (gr-dtl_b200/gr_shim/*.cc: the reference's own make() signatures, io signatures, stream tags and
message ports, bodies calling libdtl_b200.so) driven ONE FRAME PER work() CALL by gr_shim/test/shim_driver.cc, which plays the
scheduler with an in-tree stand-in of gr::block / tag_t / pmt (GNU Radio is not installed here).  Oracle: the same frames of
one stream through oracle.rx_batch (frame equaliser -> serialiser -> soft demapper -> TB decoder)."""
import os
import struct
import subprocess

import numpy as np
import pytest

import pyoracle as orc
from common import ALIST_300, ALIST_100_23, ROOT, assert_llr_close, make_oracle_cfg

pytestmark = pytest.mark.gpu
SHIM = os.path.join(ROOT, "gr-dtl_b200", "gr_shim")
DRIVER = os.path.join(SHIM, "_build", "shim_driver")


def write_blob(path, recs):
    with open(path, "wb") as f:
        for k, v in recs.items():
            b = v if isinstance(v, bytes) else np.ascontiguousarray(v).tobytes()
            f.write(struct.pack("<I", len(k)) + k.encode() + struct.pack("<Q", len(b)) + b)


def read_blob(path):
    out, b, o = {}, open(path, "rb").read(), 0
    while o < len(b):
        nl, = struct.unpack_from("<I", b, o); o += 4
        k = b[o:o + nl].decode(); o += nl
        nb, = struct.unpack_from("<Q", b, o); o += 8
        out[k] = b[o:o + nb]; o += nb
    return out


def scenario(ocfg, F, freq, taps, meta, fec, cw_llr=None, cw_fec=None):
    i32 = lambda x: np.asarray(x, dtype=np.int32)
    npil = len(ocfg.pilots)
    recs = {
        "cfg": i32([ocfg.fft_len, ocfg.cp_len, ocfg.frame_len, ocfg.header_syms, F, int(fec), ocfg.initial_mcs, 5]),
        "hysteresis": np.array([1.0]), "cnsts": i32(ocfg.constellations),
        "occ_len": i32([len(ocfg.occupied)]), "occ": i32(ocfg.occupied),
        "pc_len": i32([npil] * ocfg.n_pilot_carrier_sets), "pc": i32(ocfg.pilots * ocfg.n_pilot_carrier_sets),
        "ps_len": i32([npil] * ocfg.pilot_symbols.shape[0]), "ps": ocfg.pilot_symbols.astype(np.complex64),
        "mcs_snr": np.array([m[0] for m in ocfg.mcs], dtype=np.float64),
        "mcs_cnst": i32([m[1][0] for m in ocfg.mcs]), "mcs_fec": i32([m[1][1] for m in ocfg.mcs]),
        "alists": "\n".join(ocfg.alists).encode(),
        "freq": freq.astype(np.complex64), "taps": taps.astype(np.complex64),
        "meta": np.stack([meta[k] for k in ("constellation", "carrier_offset", "frame_no", "payload_len", "packet_len", "fec_id",
                                            "tb_no", "tb_offset", "tb_payload", "reserved")], axis=1).astype(np.int32),
    }
    if cw_llr is not None:
        recs["cw_llr"] = np.concatenate([c.astype(np.float32) for c in cw_llr])
        recs["cw_fec"] = i32(cw_fec)
    return recs


def run_driver(tmp_path, recs):
    assert os.path.exists(DRIVER), "gr_shim is not built: python __graft_entry__.py"
    a, b = str(tmp_path / "scenario.bin"), str(tmp_path / "result.bin")
    write_blob(a, recs)
    p = subprocess.run([DRIVER, a, b], capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stderr
    return read_blob(b)


@pytest.mark.parametrize("cnst,snr", [(2, 17.0), (4, 21.0)])
def test_payload_branch_through_the_blocks_frame_by_frame(tmp_path, cnst, snr):
    ocfg = make_oracle_cfg(alists=(ALIST_300,), initial_mcs=1)
    F, N, L = 24, 64, ocfg.frame_len
    t = ocfg.tx_batch(1, F, seed=70 + cnst, cnst=cnst, fec_id=1, snr_db=snr)
    freq = ocfg.rx_demod(t["iq"])
    ref = ocfg.rx_batch(freq, t["taps"], t["meta"], 1, F, freq_domain=True)
    # two codewords straight through fec_dec::decode as well
    rng = np.random.default_rng(5)
    info = ocfg.code_info(1)
    data = rng.integers(0, 2, (6, info["K"]), dtype=np.uint8)
    cw = ocfg.ldpc_encode(1, data)
    llr = (2.0 * ((2.0 * cw - 1.0) + 0.9 * rng.standard_normal(cw.shape)) / 0.81).astype(np.float32)
    got = run_driver(tmp_path, scenario(ocfg, F, freq, t["taps"], t["meta"], True, list(llr), [1] * 6))
    assert got["error"] == b""
    f32, i32 = (lambda k: np.frombuffer(got[k], dtype=np.float32)), (lambda k: np.frombuffer(got[k], dtype=np.int32))
    hard = np.frombuffer(got["hard"], dtype=np.complex64).reshape(F, L, N)
    rh = ref["hard"].reshape(F, L, N)
    active = np.abs(rh).sum(axis=(0, 1)) > 0
    assert np.array_equal(np.ascontiguousarray(hard[:, :, active]).view(np.uint32), np.ascontiguousarray(rh[:, :, active]).view(np.uint32))
    assert np.allclose(np.frombuffer(got["snr"]), ref["res"]["snr_db"], rtol=1e-6)
    assert np.allclose(np.frombuffer(got["noise"]), ref["res"]["noise"], rtol=1e-6)
    assert np.array_equal(i32("fb").reshape(F, 2), np.stack([ref["res"]["fb_cnst"], ref["res"]["fb_fec"]], axis=1))
    assert np.array_equal(i32("emitted"), ref["res"]["emitted"])
    assert np.array_equal(i32("n_llr"), ref["n_llr"])
    o = 0
    for f in range(F):
        n = int(ref["n_llr"][f])
        assert_llr_close(f32("llr")[o:o + n], ref["llr"][f, :n])
        o += n
    # the decoder block emits unpacked bits, LSB first, one len tag per CRC-passing transport block
    tbs = ref["tbs"][0]
    assert tbs["crc_ok"].sum() > 0 and tbs["avg_it"].max() > 0
    bits = np.frombuffer(got["bits"], dtype=np.uint8)
    assert np.array_equal(np.packbits(bits, bitorder="little"), ref["bytes"][0])
    assert np.array_equal(i32("bits_len"), 8 * tbs["nbytes"][tbs["crc_ok"] == 1])
    tb = i32("tb").reshape(-1, 4)
    assert np.array_equal(tb[:, 0], tbs["tb_no"]) and np.array_equal(tb[:, 3], tbs["avg_it"])
    assert tb[-1, 1] == tbs["crc_ok"].sum() and tb[-1, 2] == (tbs["crc_ok"] == 0).sum()
    # fec_dec::decode / fec_enc::encode objects
    rb, rn = ocfg.ldpc_decode(1, llr)
    assert np.array_equal(np.frombuffer(got["cw_bits"], dtype=np.uint8).reshape(6, -1), rb)
    assert np.array_equal(i32("cw_nit"), rn)
    assert np.array_equal(np.frombuffer(got["cw_wire"], dtype=np.uint8).reshape(6, -1), ocfg.ldpc_encode(1, rb))
    assert list(i32("code_nk")[:2]) == [info["N"], info["K"]]


def test_missing_constellation_tag_throws_what_the_reference_throws(tmp_path):
    ocfg = make_oracle_cfg()
    t = ocfg.tx_batch(1, 2, seed=3, cnst=orc.QPSK, snr_db=20.0)
    meta = t["meta"].copy()
    meta["constellation"][1] = 0
    got = run_driver(tmp_path, scenario(ocfg, 2, ocfg.rx_demod(t["iq"]), t["taps"], meta, False))
    assert got["error"] == b"Missing constellation tag."          # frame_equalizer_vcvc_impl.cc:150-152
    assert np.frombuffer(got["emitted"], dtype=np.int32).size == 1


def test_chunks_to_symbols_and_decoder_cb_blocks(tmp_path):
    """python/dtl/qa_ofdm_adaptive_chunks_to_symbols_bc.py:39-63 through the shims: [0,1,2,3,3,2,1,0] in QPSK comes back, the
    symbols in between are the constellation's points; the same for every constellation; a stream without the constellation
    tag throws "Constellation not found" (chunks_to_symbols_bc_impl.cc:69-71)."""
    ocfg = make_oracle_cfg(cnsts=(1, 2, 3, 4, 5, 6))
    t = ocfg.tx_batch(1, 1, seed=1, cnst=2, snr_db=30.0)
    rng = np.random.default_rng(12)
    bps = {1: 1, 2: 2, 3: 3, 4: 4, 5: 6, 6: 8}
    cn = [2, 1, 3, 4, 5, 6]
    chunks = np.stack([np.array([0, 1, 2, 3, 3, 2, 1, 0], dtype=np.uint8)] +
                      [rng.integers(0, 1 << bps[c], 8, dtype=np.uint8) for c in cn[1:]])
    recs = scenario(ocfg, 0, np.zeros((0, 20, 64), np.complex64), np.zeros((0, 64), np.complex64), t["meta"][:0], False)
    recs["md_chunks"] = chunks
    recs["md_cnst"] = np.asarray(cn, dtype=np.int32)
    got = run_driver(tmp_path, recs)
    assert got["md_error"] == b""
    back = np.frombuffer(got["md_out"], dtype=np.uint8).reshape(len(cn), 8)
    assert np.array_equal(back, chunks) and list(back[0]) == [0, 1, 2, 3, 3, 2, 1, 0]
    syms = np.frombuffer(got["md_syms"], dtype=np.complex64).reshape(len(cn), 8)
    for k, c in enumerate(cn):
        assert np.array_equal(syms[k].view(np.uint32), orc.map_chunks(c, chunks[k]).astype(np.complex64).view(np.uint32)), c
    recs["md_cnst"] = np.asarray([0], dtype=np.int32)
    recs["md_chunks"] = chunks[:1]
    assert run_driver(tmp_path, recs)["md_error"] == b"Constellation not found"
