"""CPU-side checks of the drop-in boundary: the library loads, exports every symbol include/dtl_b200.h declares,
its PODs have the layout the ctypes mirror assumes, and it refuses to run without a CUDA device."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

from common import ROOT, make_product_cfg

HEADER = os.path.join(ROOT, "include", "dtl_b200.h")


def _declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(dtl_b200_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    import dtl_b200
    from dtl_b200 import _capi
    L = dtl_b200.lib()
    declared = _declared_symbols()
    assert len(declared) >= 25
    for name in declared:
        assert hasattr(L, name), name
    assert sorted(_capi.SYMBOLS) == declared
    assert L.dtl_b200_abi_version() == 1


def test_pod_layout_matches_header(tmp_path):
    from dtl_b200 import _capi
    prog = tmp_path / "sz.c"
    prog.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "dtl_b200.h"\n'
                    'int main(void){printf("%zu %zu %zu %zu %zu %zu %zu\\n", sizeof(dtl_config), sizeof(dtl_frame_meta),'
                    'sizeof(dtl_frame_result), sizeof(dtl_tb_result), offsetof(dtl_config, mcs_snr),'
                    'offsetof(dtl_config, ldpc_sigma2), offsetof(dtl_config, device));return 0;}\n')
    exe = tmp_path / "sz"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), str(prog), "-o", str(exe)])
    got = [int(x) for x in subprocess.check_output([str(exe)]).split()]
    want = [C.sizeof(_capi.dtl_config), _capi.frame_meta_dtype.itemsize, _capi.frame_result_dtype.itemsize,
            _capi.tb_result_dtype.itemsize, _capi.dtl_config.mcs_snr.offset, _capi.dtl_config.ldpc_sigma2.offset,
            _capi.dtl_config.device.offset]
    assert got == want


def test_no_cpu_fallback():
    """Without a CUDA device the product refuses to construct (DTL_ERR_NO_DEVICE); with one it constructs."""
    import dtl_b200
    try:
        import torch
        have_gpu = torch.cuda.is_available()
    except Exception:
        have_gpu = False
    if have_gpu:
        rx = dtl_b200.ofdm_receiver(make_product_cfg())
        assert rx.frame_capacity == 960
        rx.close()
    else:
        with pytest.raises(dtl_b200.DtlError) as ei:
            dtl_b200.ofdm_receiver(make_product_cfg())
        assert ei.value.status == dtl_b200._capi.ERR_NO_DEVICE


def test_product_never_touches_the_oracle():
    """Nothing under gr-dtl_b200/ may import, include or link oracle/ (test infrastructure only)."""
    bad = []
    for base, _, files in os.walk(os.path.join(ROOT, "gr-dtl_b200")):
        if os.path.basename(base) in ("build", "lib", "__pycache__"):
            continue
        for f in files:
            if f.endswith((".py", ".cc", ".cu", ".cuh", ".hpp", ".h", "Makefile")):
                txt = open(os.path.join(base, f), errors="ignore").read()
                if re.search(r"pyoracle|liboracle|oracle/src|import oracle|#include \"oracle", txt):
                    bad.append(os.path.join(base, f))
    assert not bad, bad


def test_config_mirror_defaults():
    """python/dtl/qa_ofdm_adaptive_config.py: defaults (fft_len == 64) + helpers of python/dtl/ofdm_adaptive.py"""
    import dtl_b200
    cfg = dtl_b200.ofdm_adaptive_rx_config()
    assert cfg.fft_len == 64 and cfg.cp_len == 16 and cfg.frame_length == 20
    assert len(cfg.pilot_symbols) == 127 and cfg.pilot_symbols[4] == (-1, -1, -1, 1)
    assert dtl_b200.frame_capacity(20, cfg.occupied_carriers) == 960
    assert dtl_b200.max_bps([1, 2, 3, 4]) == 4
    assert len(cfg.sync_word1) == 64 and np.count_nonzero(cfg.sync_word2) == 52
