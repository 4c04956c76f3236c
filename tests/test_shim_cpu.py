"""CPU-side checks of the GNU Radio block shims (gr-dtl_b200/gr_shim): the library builds against the in-tree stand-in,
exports the reference's factory functions under the reference's names, and fails loudly without a CUDA device."""
import os
import subprocess

import numpy as np

from common import ROOT, make_oracle_cfg

SHIM = os.path.join(ROOT, "gr-dtl_b200", "gr_shim")


def _build():
    subprocess.check_call(["make", "-C", SHIM], stdout=subprocess.DEVNULL)


def test_shim_exports_the_reference_factories():
    _build()
    out = subprocess.run(["nm", "-D", "--demangle", "--defined-only", os.path.join(SHIM, "_build", "libgnuradio-dtl-b200.so")],
                         capture_output=True, text=True, check=True).stdout
    for sym in ("gr::dtl::ofdm_adaptive_frame_equalizer_vcvc::make(",           # ofdm_adaptive_frame_equalizer_vcvc.h:30-36
                "gr::dtl::ofdm_adaptive_payload_equalizer::make(",              # ofdm_adaptive_equalizer.h:106-117
                "gr::dtl::ofdm_adaptive_feedback_decision::make(",              # ofdm_adaptive_feedback_decision.h:49-52
                "gr::dtl::ofdm_adaptive_constellation_soft_cf::make(",          # ofdm_adaptive_constellation_soft_cf.h:28
                "gr::dtl::ofdm_adaptive_fec_decoder::make(",                    # ofdm_adaptive_fec_decoder.h:27
                "gr::dtl::ofdm_adaptive_constellation_decoder_cb::make(",       # ofdm_adaptive_constellation_decoder_cb.h:34-35
                "gr::dtl::ofdm_adaptive_chunks_to_symbols_bc::make(",           # ofdm_adaptive_chunks_to_symbols_bc.h:32-35
                "gr::dtl::make_ldpc_decoders(", "gr::dtl::make_ldpc_encoders("):    # fec.h:33-35
        assert sym in out, sym
    # the only undefined dtl_b200_* symbols are C ABI entry points declared in include/dtl_b200.h
    und = subprocess.run(["nm", "-D", "--undefined-only", os.path.join(SHIM, "_build", "libgnuradio-dtl-b200.so")],
                         capture_output=True, text=True, check=True).stdout
    used = sorted({l.split()[-1] for l in und.splitlines() if "dtl_b200_" in l})
    header = open(os.path.join(ROOT, "include", "dtl_b200.h")).read()
    assert used and all(u + "(" in header for u in used), used
    assert {"dtl_b200_equalize", "dtl_b200_soft_demap", "dtl_b200_hard_demap", "dtl_b200_map", "dtl_b200_fec_decode", "dtl_b200_ldpc_decode",
            "dtl_b200_ldpc_encode"} <= set(used)


def test_shim_sources_do_not_touch_the_oracle_or_the_reference():
    for f in os.listdir(SHIM):
        if f.endswith((".cc", ".h")):
            txt = open(os.path.join(SHIM, f)).read()
            assert "oracle" not in txt and "#include \"/root/reference" not in txt, f


def test_driver_fails_loudly_without_a_gpu(tmp_path):
    import torch
    if torch.cuda.is_available():
        return
    _build()
    from test_gpu_shim import scenario, write_blob
    ocfg = make_oracle_cfg()
    t = ocfg.tx_batch(1, 1, seed=1, cnst=2, snr_db=20.0)
    a, b = str(tmp_path / "s.bin"), str(tmp_path / "r.bin")
    write_blob(a, scenario(ocfg, 1, ocfg.rx_demod(t["iq"]), t["taps"], t["meta"], False))
    p = subprocess.run([os.path.join(SHIM, "_build", "shim_driver"), a, b], capture_output=True, text=True)
    # the first work() call creates the handle: no CUDA device -> the block throws, nothing is computed on the CPU
    got = open(b, "rb").read() if os.path.exists(b) else b""
    assert p.returncode != 0 or b"no CUDA device" in got, (p.returncode, p.stderr)
