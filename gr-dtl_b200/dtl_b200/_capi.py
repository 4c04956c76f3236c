"""ctypes binding of libdtl_b200.so (include/dtl_b200.h).

The library is the product: there is no Python/NumPy/torch implementation of the path behind this
module, and importing it without the built library raises.  Build with `python __graft_entry__.py`
(or `make -C gr-dtl_b200/csrc`).
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("DTL_B200_LIB") or os.path.join(os.path.dirname(_HERE), "lib", "libdtl_b200.so")

OK, ERR_INVALID_ARG, ERR_MISSING_TAG, ERR_CUDA, ERR_NO_DEVICE, ERR_UNSUPPORTED, ERR_CAPACITY, ERR_IO = range(8)
MEM_HOST, MEM_DEVICE = 0, 1
LDPC_REFERENCE, LDPC_LAYERED = 0, 1
STAGES = ("fft", "eq_demap", "stream_state", "tb_plan", "ldpc", "tb_finish", "tx", "demap", "header", "sync")

frame_meta_dtype = np.dtype([
    ("constellation", "<i4"), ("carrier_offset", "<i4"), ("frame_no", "<i4"), ("payload_len", "<i4"),
    ("packet_len", "<i4"), ("fec_id", "<i4"), ("tb_no", "<i4"), ("tb_offset", "<i4"), ("tb_payload", "<i4"),
    ("reserved", "<i4")])
frame_result_dtype = np.dtype([
    ("snr_db", "<f8"), ("noise", "<f8"), ("fb_cnst", "<i4"), ("fb_fec", "<i4"), ("emitted", "<i4"),
    ("lost_frames", "<i4")])
tb_result_dtype = np.dtype([("tb_no", "<i4"), ("crc_ok", "<i4"), ("nbytes", "<i4"), ("avg_it", "<i4"),
                            ("frame_index", "<i4")])
header_result_dtype = np.dtype([("crc_ok", "<i4"), ("feedback_cnst", "<i4"), ("feedback_fec", "<i4"), ("reserved", "<i4")])

_i32p, _f32p, _f64p = C.POINTER(C.c_int32), C.POINTER(C.c_float), C.POINTER(C.c_double)


class dtl_config(C.Structure):
    _fields_ = [
        ("fft_len", C.c_int32), ("cp_len", C.c_int32), ("frame_len", C.c_int32), ("header_syms", C.c_int32),
        ("n_occupied_sets", C.c_int32), ("occupied_set_len", _i32p), ("occupied_carriers", _i32p),
        ("n_pilot_carrier_sets", C.c_int32), ("pilot_carrier_set_len", _i32p), ("pilot_carriers", _i32p),
        ("n_pilot_symbol_sets", C.c_int32), ("pilot_symbol_set_len", _i32p), ("pilot_symbols", _f32p),
        ("symbols_skipped", C.c_int32), ("alpha", C.c_float), ("input_is_shifted", C.c_int32),
        ("n_mcs", C.c_int32), ("mcs_snr", _f64p), ("mcs_cnst", _i32p), ("mcs_fec", _i32p),
        ("initial_mcs", C.c_int32), ("hysteresis", C.c_double), ("decision_th", C.c_int32),
        ("n_codes", C.c_int32), ("alist_paths", C.POINTER(C.c_char_p)),
        ("fec_enabled", C.c_int32), ("max_bps", C.c_int32), ("ldpc_max_iterations", C.c_int32),
        ("ldpc_mode", C.c_int32), ("ldpc_sigma2", C.c_double),
        ("sync_word1", _f32p), ("sync_word2", _f32p),
        ("device", C.c_int32), ("sync_threshold", C.c_float), ("reserved", C.c_int32 * 6)]


# every symbol include/dtl_b200.h declares (tests/test_capi_cpu.py checks the header and the built library against this list)
SYMBOLS = (
    "dtl_b200_abi_version", "dtl_b200_create", "dtl_b200_destroy", "dtl_b200_last_error", "dtl_b200_reset",
    "dtl_b200_sync", "dtl_b200_cuda_stream", "dtl_b200_frame_capacity", "dtl_b200_frame_samples",
    "dtl_b200_llr_stride", "dtl_b200_bytes_stride", "dtl_b200_code_info", "dtl_b200_launch_count",
    "dtl_b200_profile", "dtl_b200_stage_times",
    "dtl_b200_host_alloc", "dtl_b200_host_free", "dtl_b200_demod", "dtl_b200_fft", "dtl_b200_equalize", "dtl_b200_equalize_noalloc",
    "dtl_b200_soft_demap", "dtl_b200_hard_demap", "dtl_b200_map", "dtl_b200_fec_decode", "dtl_b200_ldpc_decode", "dtl_b200_ldpc_encode",
    "dtl_b200_rx", "dtl_b200_tx", "dtl_b200_rx_headers", "dtl_b200_rx_iq", "dtl_b200_fec_frame", "dtl_b200_rx_stream")

_lib = None


class DtlError(RuntimeError):
    def __init__(self, status, msg):
        super().__init__("libdtl_b200 status %d: %s" % (status, msg))
        self.status = status


def lib():
    """Load libdtl_b200.so; raises if it has not been built (there is no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError("libdtl_b200.so is not built (%s): run `python __graft_entry__.py`; "
                              "this package has no CPU/PyTorch fallback" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        vp, i32, i64 = C.c_void_p, C.c_int32, C.c_int64
        L.dtl_b200_abi_version.restype = C.c_int
        L.dtl_b200_create.argtypes = [C.POINTER(dtl_config), C.POINTER(vp)]
        L.dtl_b200_destroy.argtypes = [vp]
        L.dtl_b200_destroy.restype = None
        L.dtl_b200_last_error.argtypes = [vp]
        L.dtl_b200_last_error.restype = C.c_char_p
        L.dtl_b200_reset.argtypes = [vp, i32]
        L.dtl_b200_sync.argtypes = [vp]
        L.dtl_b200_cuda_stream.argtypes = [vp]
        L.dtl_b200_cuda_stream.restype = vp
        for f in ("frame_capacity", "frame_samples", "llr_stride", "bytes_stride"):
            getattr(L, "dtl_b200_" + f).argtypes = [vp]
            getattr(L, "dtl_b200_" + f).restype = i32
        L.dtl_b200_code_info.argtypes = [vp, i32, _i32p, _i32p, vp]
        L.dtl_b200_launch_count.argtypes = [vp]
        L.dtl_b200_launch_count.restype = C.c_uint64
        L.dtl_b200_profile.argtypes = [vp, C.c_int]
        L.dtl_b200_stage_times.argtypes = [vp, vp, vp, i32, C.c_int]
        L.dtl_b200_host_alloc.argtypes = [C.POINTER(vp), C.c_size_t]
        L.dtl_b200_host_free.argtypes = [vp]
        L.dtl_b200_host_free.restype = None
        L.dtl_b200_demod.argtypes = [vp, vp, i64, vp, C.c_int]
        L.dtl_b200_fft.argtypes = [vp, vp, i64, C.c_int, vp, C.c_int]
        L.dtl_b200_equalize.argtypes = [vp, vp, vp, vp, i32, i32, vp, vp, vp, vp, vp, vp, vp, vp, vp, C.c_int]
        L.dtl_b200_equalize_noalloc.argtypes = [vp, vp, vp, vp, i32, i32, vp, vp, vp]
        L.dtl_b200_soft_demap.argtypes = [vp, vp, i32, vp, vp, i64, vp, i32, C.c_int]
        L.dtl_b200_hard_demap.argtypes = [vp, vp, i32, vp, i64, vp, C.c_int]
        L.dtl_b200_map.argtypes = [vp, vp, i32, vp, i64, vp, C.c_int]
        L.dtl_b200_fec_decode.argtypes = [vp, vp, i32, vp, i32, i32, vp, i32, vp, vp, i32, vp, C.c_int]
        L.dtl_b200_ldpc_decode.argtypes = [vp, i32, vp, i64, vp, vp, C.c_int]
        L.dtl_b200_ldpc_encode.argtypes = [vp, i32, vp, i64, vp, C.c_int]
        L.dtl_b200_rx.argtypes = [vp, C.c_int, vp, vp, vp, i32, i32, vp, vp, i32, vp, vp, i32, vp, vp, C.c_int]
        L.dtl_b200_tx.argtypes = [vp, vp, vp, vp, i64, vp, C.c_int]
        L.dtl_b200_fec_frame.argtypes = [vp, vp, i32, vp, vp, vp, i32, i32, vp, vp, vp, vp, C.c_int]
        L.dtl_b200_rx_headers.argtypes = [vp, vp, i32, i32, vp, vp, vp, C.c_int]
        L.dtl_b200_rx_iq.argtypes = [vp, vp, i32, i32, vp, vp, vp, vp, i32, vp, vp, i32, vp, vp, C.c_int]
        L.dtl_b200_rx_stream.argtypes = [vp, vp, i32, i64, i32, vp, vp, vp, vp, vp, vp, vp, i32, vp, vp, i32, vp, vp, C.c_int]
        _lib = L
    return _lib
