"""Batched mirror of python/dtl/ofdm_receiver.py (payload branch, :152-239) and of the stage-5 half of
python/dtl/ofdm_transmitter.py (:95-98,168-189) on top of libdtl_b200.so.

The reference wires one GNU Radio block per stage and the scheduler hands each block one tagged frame per
work() call; here one object owns a libdtl_b200 handle and every method takes S streams x F frames at once.
Stream tags travel as `frame_meta_dtype` records, tags/messages come back as `frame_result_dtype` /
`tb_result_dtype` records.  Arrays may be NumPy (host; the call returns when outputs are valid) or torch CUDA
tensors (device; the call only enqueues on the handle's stream -- wrap it in `with rx.stream_context():`).
"""
import ctypes as C

import numpy as np

from . import _capi
from ._capi import (MEM_DEVICE, MEM_HOST, DtlError, frame_meta_dtype, frame_result_dtype, tb_result_dtype)
from .config import get_bits_per_symbol, frame_capacity, max_bps


def _is_torch(a):
    return type(a).__module__.startswith("torch")


class _Arg:
    """pointer + memory kind of a NumPy array / torch tensor (None -> NULL)"""

    def __init__(self, a, dtype=None):
        self.keep = a
        if a is None:
            self.ptr, self.mem = None, None
        elif _is_torch(a):
            if not a.is_contiguous():
                raise ValueError("tensor must be contiguous")
            self.ptr, self.mem = C.c_void_p(a.data_ptr()), (MEM_DEVICE if a.is_cuda else MEM_HOST)
        else:
            if dtype is not None and a.dtype != dtype:
                raise ValueError("expected dtype %s, got %s" % (dtype, a.dtype))
            if not a.flags["C_CONTIGUOUS"]:
                raise ValueError("array must be C-contiguous")
            self.ptr, self.mem = C.c_void_p(a.ctypes.data), MEM_HOST


def _mem_of(*args):
    kinds = {a.mem for a in args if a.mem is not None}
    if len(kinds) != 1:
        raise ValueError("all buffers of one call must live in the same memory space")
    return kinds.pop()


class ofdm_receiver:
    """__init__ mirrors ofdm_receiver.__init__(config) (python/dtl/ofdm_receiver.py:21-57)."""

    def __init__(self, config, device=0, ldpc_mode=_capi.LDPC_REFERENCE, ldpc_sigma2=2.0, ldpc_max_iterations=15,
                 alist_dir=""):
        import os
        self.config = config
        self.fft_len, self.cp_len, self.frame_length = config.fft_len, config.cp_len, config.frame_length
        mc_schemes = list(zip(*config.mcs))[1]
        cnsts = list(zip(*mc_schemes))[0]
        self.constellations = sorted(set(int(c) for c in cnsts))
        self.fec = len(config.fec_codes)
        self.codes_alist, self.codes_id = [], {}
        if self.fec:
            self.codes_alist = [os.path.join(alist_dir, p) if alist_dir and not os.path.isabs(p) else p
                                for p in list(zip(*config.fec_codes))[1]]
            self.codes_id = {name: i + 1 for (i, name) in enumerate(list(zip(*config.fec_codes))[0])}
        self.mcs = [(snr_th, (int(cnst), self.codes_id.get(code_name, 0))) for (snr_th, (cnst, code_name)) in config.mcs]
        self.initial_mcs_id = config.initial_mcs_id
        if [self.fft_len, self.fft_len] != [len(config.sync_word1), len(config.sync_word2)]:
            raise ValueError("Length of sync sequence(s) must be FFT length.")
        header_len = 2 if self.fec else 1
        self.header_len = header_len
        self.max_bps = max_bps(self.constellations)
        self._lib = _capi.lib()
        # ---- dtl_config
        cfg = _capi.dtl_config()
        keep = []

        def i32(v):
            a = np.ascontiguousarray(v, dtype=np.int32)
            keep.append(a)
            return a.ctypes.data_as(C.POINTER(C.c_int32))

        def f32(v):
            a = np.ascontiguousarray(v).astype(np.complex64).view(np.float32)
            keep.append(a)
            return a.ctypes.data_as(C.POINTER(C.c_float))

        def f64(v):
            a = np.ascontiguousarray(v, dtype=np.float64)
            keep.append(a)
            return a.ctypes.data_as(C.POINTER(C.c_double))

        cfg.fft_len, cfg.cp_len, cfg.frame_len, cfg.header_syms = self.fft_len, self.cp_len, self.frame_length, header_len
        occ, pc, ps = config.occupied_carriers, config.pilot_carriers, config.pilot_symbols
        cfg.n_occupied_sets = len(occ)
        cfg.occupied_set_len = i32([len(s) for s in occ])
        cfg.occupied_carriers = i32([c for s in occ for c in s])
        cfg.n_pilot_carrier_sets = len(pc)
        cfg.pilot_carrier_set_len = i32([len(s) for s in pc] or [0])
        cfg.pilot_carriers = i32([c for s in pc for c in s] or [0])
        cfg.n_pilot_symbol_sets = len(ps)
        cfg.pilot_symbol_set_len = i32([len(s) for s in ps] or [0])
        cfg.pilot_symbols = f32([c for s in ps for c in s] or [0])
        cfg.symbols_skipped, cfg.alpha, cfg.input_is_shifted = header_len, 0.1, 1   # ofdm_receiver.py:162-163
        cfg.n_mcs = len(self.mcs)
        cfg.mcs_snr = f64([m[0] for m in self.mcs])
        cfg.mcs_cnst = i32([m[1][0] for m in self.mcs])
        cfg.mcs_fec = i32([m[1][1] for m in self.mcs])
        cfg.initial_mcs, cfg.hysteresis, cfg.decision_th = self.initial_mcs_id, 1.0, 5   # ofdm_receiver.py:167
        cfg.n_codes = len(self.codes_alist)
        paths = (C.c_char_p * max(1, len(self.codes_alist)))(*[p.encode() for p in self.codes_alist])
        keep.append(paths)
        cfg.alist_paths = C.cast(paths, C.POINTER(C.c_char_p))
        cfg.fec_enabled, cfg.max_bps = int(bool(self.fec)), self.max_bps
        cfg.ldpc_max_iterations, cfg.ldpc_mode, cfg.ldpc_sigma2 = ldpc_max_iterations, ldpc_mode, ldpc_sigma2
        cfg.sync_word1, cfg.sync_word2 = f32(config.sync_word1), f32(config.sync_word2)
        cfg.device = device
        cfg.sync_threshold = float(getattr(config, "sync_threshold", 0.95))
        self._keep = keep
        h = C.c_void_p()
        st = self._lib.dtl_b200_create(C.byref(cfg), C.byref(h))
        if st != _capi.OK:
            raise DtlError(st, (self._lib.dtl_b200_last_error(None) or b"").decode())
        self._h = h
        self._pinned = []
        self.device = device
        self.frame_capacity = self._lib.dtl_b200_frame_capacity(h)
        assert self.frame_capacity == frame_capacity(self.frame_length, occ)
        self.frame_samples = self._lib.dtl_b200_frame_samples(h)
        self.llr_stride = self._lib.dtl_b200_llr_stride(h)
        self.bytes_stride = self._lib.dtl_b200_bytes_stride(h)

    # ---- plumbing -------------------------------------------------------------------------------------
    def close(self):
        if getattr(self, "_h", None):
            self._lib.dtl_b200_destroy(self._h)
            self._h = None
            for p in self._pinned:
                self._lib.dtl_b200_host_free(p)
            self._pinned = []

    def __del__(self):
        try:
            import sys
            if sys is None or sys.is_finalizing():      # interpreter teardown: the CUDA runtime may already be gone
                return
            self.close()
        except Exception:
            pass

    def _ck(self, st):
        if st != _capi.OK:
            raise DtlError(st, (self._lib.dtl_b200_last_error(self._h) or b"").decode())

    def reset(self, n_streams):
        self._ck(self._lib.dtl_b200_reset(self._h, int(n_streams)))

    def sync(self):
        self._ck(self._lib.dtl_b200_sync(self._h))

    def profile(self, enable=True):
        """bracket every stage's kernels with CUDA events on the handle's stream (dtl_b200_profile)"""
        self._ck(self._lib.dtl_b200_profile(self._h, int(enable)))

    def stage_times(self, reset=True):
        """{stage: (milliseconds, kernel launches)} accumulated since the last reset (dtl_b200_stage_times)"""
        n = len(_capi.STAGES)
        ms, cnt = np.zeros(n, dtype=np.float64), np.zeros(n, dtype=np.int64)
        self._ck(self._lib.dtl_b200_stage_times(self._h, ms.ctypes.data_as(C.c_void_p), cnt.ctypes.data_as(C.c_void_p), n,
                                                int(reset)))
        return {k: (float(ms[i]), int(cnt[i])) for i, k in enumerate(_capi.STAGES)}

    def host_alloc(self, shape, dtype):
        """page-locked NumPy array (dtl_b200_host_alloc); freed with the receiver"""
        dtype = np.dtype(dtype)
        nbytes = int(np.prod(shape)) * dtype.itemsize
        p = C.c_void_p()
        st = self._lib.dtl_b200_host_alloc(C.byref(p), max(nbytes, 1))
        if st != _capi.OK:
            raise DtlError(st, "dtl_b200_host_alloc failed")
        self._pinned.append(p)
        buf = (C.c_uint8 * max(nbytes, 1)).from_address(p.value)
        return np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)

    @property
    def launch_count(self):
        return int(self._lib.dtl_b200_launch_count(self._h))

    @property
    def cuda_stream(self):
        return int(self._lib.dtl_b200_cuda_stream(self._h) or 0)

    def torch_stream(self):
        import torch
        return torch.cuda.ExternalStream(self.cuda_stream, device=self.device)

    def stream_context(self):
        import torch
        return torch.cuda.stream(self.torch_stream())

    def code_info(self, fec_id):
        n, k = C.c_int32(0), C.c_int32(0)
        perm = np.zeros(65536, dtype=np.int32)
        self._ck(self._lib.dtl_b200_code_info(self._h, fec_id, C.byref(n), C.byref(k), perm.ctypes.data_as(C.c_void_p)))
        return dict(N=n.value, K=k.value, permute=perm[:n.value].copy())

    def _new(self, like_mem, shape, dtype):
        if like_mem == MEM_HOST:
            return np.zeros(shape, dtype=dtype)
        import torch
        if isinstance(dtype, np.dtype) and dtype.fields is not None:
            n = int(np.prod(shape)) * dtype.itemsize
            t = torch.zeros(n, dtype=torch.uint8, device="cuda:%d" % self.device)
            torch.cuda.current_stream(t.device).synchronize()
            return t
        tdt = {np.dtype(np.float32): torch.float32, np.dtype(np.complex64): torch.complex64, np.dtype(np.uint8): torch.uint8,
               np.dtype(np.int32): torch.int32, np.dtype(np.float64): torch.float64, np.dtype(np.int64): torch.int64}[np.dtype(dtype)]
        t = torch.zeros(shape, dtype=tdt, device="cuda:%d" % self.device)
        # the fill runs on torch's current stream, the library works on its own: without this a large fill can land
        # after (on top of) what the library's kernels wrote
        torch.cuda.current_stream(t.device).synchronize()
        return t

    # ---- stage 1: header_payload_demux (CP strip) + fft_vcc (ofdm_receiver.py:79-86,153) -----------------
    def demod(self, iq, out=None):
        a = _Arg(iq)
        n = iq.numel() // self.frame_samples if _is_torch(iq) else iq.size // self.frame_samples
        if out is None:
            out = self._new(a.mem, (n, self.frame_length, self.fft_len), np.complex64)
        self._ck(self._lib.dtl_b200_demod(self._h, a.ptr, n, _Arg(out).ptr, a.mem))
        return out

    def fft(self, x, forward=True):
        """fft.fft_vcc(fft_len, forward, (), True)"""
        a = _Arg(x)
        n = (x.numel() if _is_torch(x) else x.size) // self.fft_len
        out = self._new(a.mem, (n, self.fft_len), np.complex64)
        self._ck(self._lib.dtl_b200_fft(self._h, a.ptr, n, int(forward), _Arg(out).ptr, a.mem))
        return out

    # ---- stages 2+3: dtl.ofdm_adaptive_frame_equalizer_vcvc + serialiser + demapper (:165-217) --------------
    def equalize(self, freq, taps, meta, n_streams, frames_per_stream, want=("hard", "soft", "taps", "res", "llr")):
        """Returns a dict with the requested outputs: hard/soft [S*F, L, N] (ports 0/1 of the frame equaliser),
        taps (channel state after the frame), res (tags + feedback message), llr/n_llr (soft_cf output, FEC
        chain), bytes/nbytes/crc_ok (constellation_decoder_cb + frame_pack_bb output, no-FEC chain)."""
        S, F = int(n_streams), int(frames_per_stream)
        n, L, N = S * F, self.frame_length, self.fft_len
        a_f, a_t, a_m = _Arg(freq), _Arg(taps), _Arg(meta)
        mem = _mem_of(a_f, a_t, a_m)
        o = {}
        if "hard" in want:
            o["hard"] = self._new(mem, (n, L, N), np.complex64)
        if "soft" in want:
            o["soft"] = self._new(mem, (n, L, N), np.complex64)
        if "taps" in want:
            o["taps"] = self._new(mem, (n, N), np.complex64)
        if "res" in want:
            o["res"] = self._new(mem, (n,), frame_result_dtype)
        if "llr" in want:
            o["llr"] = self._new(mem, (n, self.llr_stride), np.float32)
            o["n_llr"] = self._new(mem, (n,), np.int32)
        if "bytes" in want:
            o["bytes"] = self._new(mem, (n, self.bytes_stride), np.uint8)
            o["nbytes"] = self._new(mem, (n,), np.int32)
            o["crc_ok"] = self._new(mem, (n,), np.int32)
        g = lambda k: _Arg(o.get(k)).ptr
        self._ck(self._lib.dtl_b200_equalize(self._h, a_f.ptr, a_t.ptr, a_m.ptr, S, F, g("hard"), g("soft"), g("taps"),
                                             g("res"), g("llr"), g("n_llr"), g("bytes"), g("nbytes"), g("crc_ok"), mem))
        return o

    def equalize_noalloc(self, freq, taps, meta, n_streams, frames_per_stream, out):
        """dtl_b200_equalize_noalloc: the lean equaliser pass for callers that must not allocate.  Device tensors only, and
        `out` must carry the preallocated device buffers res [S*F] (+ optionally llr [S*F, llr_stride], n_llr [S*F]); nothing
        is allocated here or in the library.  Call reset(n_streams) first; sync() before reading."""
        a_f, a_t, a_m = _Arg(freq), _Arg(taps), _Arg(meta)
        if _mem_of(a_f, a_t, a_m) != MEM_DEVICE or "res" not in out or any(not _is_torch(v) for v in out.values()):
            raise ValueError("equalize_noalloc takes device buffers only, and a preallocated out['res']")
        g = lambda k: _Arg(out.get(k)).ptr
        self._ck(self._lib.dtl_b200_equalize_noalloc(self._h, a_f.ptr, a_t.ptr, a_m.ptr, int(n_streams), int(frames_per_stream),
                                                     g("res"), g("llr"), g("n_llr")))
        return out

    # ---- stand-alone blocks ---------------------------------------------------------------------------------
    def soft_demap(self, syms, meta, noise=None):
        """dtl.ofdm_adaptive_constellation_soft_cf(constellations, len_key): syms [n_frames, stride]"""
        a_s, a_m, a_n = _Arg(syms), _Arg(meta), _Arg(noise)
        mem = _mem_of(a_s, a_m, a_n)
        n, stride = syms.shape[0], syms.shape[1]
        out = self._new(mem, (n, stride * self.max_bps), np.float32)
        self._ck(self._lib.dtl_b200_soft_demap(self._h, a_s.ptr, stride, a_m.ptr, a_n.ptr, n, _Arg(out).ptr,
                                               stride * self.max_bps, mem))
        return out

    def hard_demap(self, syms, meta):
        """dtl.ofdm_adaptive_constellation_decoder_cb(constellations, tsb_key)"""
        a_s, a_m = _Arg(syms), _Arg(meta)
        mem = _mem_of(a_s, a_m)
        out = self._new(mem, tuple(syms.shape), np.uint8)
        self._ck(self._lib.dtl_b200_hard_demap(self._h, a_s.ptr, syms.shape[1], a_m.ptr, syms.shape[0], _Arg(out).ptr, mem))
        return out

    def chunks_to_symbols(self, chunks, meta):
        """dtl.ofdm_adaptive_chunks_to_symbols_bc(constellations, tsb_key)"""
        a_c, a_m = _Arg(chunks), _Arg(meta)
        mem = _mem_of(a_c, a_m)
        out = self._new(mem, tuple(chunks.shape), np.complex64)
        self._ck(self._lib.dtl_b200_map(self._h, a_c.ptr, chunks.shape[1], a_m.ptr, chunks.shape[0], _Arg(out).ptr, mem))
        return out

    # ---- stage 4: dtl.ofdm_adaptive_fec_decoder + dtl.ofdm_adaptive_fec_pack_bb (:196-202) ------------------------
    def fec_decode(self, llr, meta, n_streams, frames_per_stream, bytes_cap=None, tb_cap=None):
        S, F = int(n_streams), int(frames_per_stream)
        a_l, a_m = _Arg(llr), _Arg(meta)
        mem = _mem_of(a_l, a_m)
        bytes_cap = bytes_cap or (self.bytes_stride + 8) * (F + 1)
        tb_cap = tb_cap or 2 * F + 1
        o = dict(bytes=self._new(mem, (S, bytes_cap), np.uint8), nbytes=self._new(mem, (S,), np.int32),
                 tbs=self._new(mem, (S, tb_cap), tb_result_dtype), ntbs=self._new(mem, (S,), np.int32))
        stride = llr.shape[-1]
        self._ck(self._lib.dtl_b200_fec_decode(self._h, a_l.ptr, stride, a_m.ptr, S, F, _Arg(o["bytes"]).ptr, bytes_cap,
                                               _Arg(o["nbytes"]).ptr, _Arg(o["tbs"]).ptr, tb_cap, _Arg(o["ntbs"]).ptr, mem))
        return o

    def ldpc_decode(self, fec_id, llr):
        """fec_dec::decode for a batch: llr [count, n] in wire order -> (bits [count, k], nit [count])"""
        a = _Arg(llr)
        count = llr.shape[0]
        info = self.code_info(fec_id)
        bits = self._new(a.mem, (count, info["K"]), np.uint8)
        nit = self._new(a.mem, (count,), np.int32)
        self._ck(self._lib.dtl_b200_ldpc_decode(self._h, fec_id, a.ptr, count, _Arg(bits).ptr, _Arg(nit).ptr, a.mem))
        return bits, nit

    def ldpc_encode(self, fec_id, data):
        """ldpc_enc::encode for a batch: data bits [count, k] -> wire bits [count, n] = [parity | data]"""
        a = _Arg(data)
        count = data.shape[0]
        info = self.code_info(fec_id)
        out = self._new(a.mem, (count, info["N"]), np.uint8)
        self._ck(self._lib.dtl_b200_ldpc_encode(self._h, fec_id, a.ptr, count, _Arg(out).ptr, a.mem))
        return out

    # ---- header branch (python/dtl/ofdm_receiver.py:100-150) ----------------------------------------------------
    def rx_headers(self, iq, n_streams, frames_per_stream):
        """chanest + header equaliser + header parser: the stream tags of every frame from its IQ alone.
        Returns meta [S*F], taps [S*F, N] (as the payload equaliser receives them) and hres [S*F]."""
        from ._capi import header_result_dtype
        S, F = int(n_streams), int(frames_per_stream)
        a = _Arg(iq)
        o = dict(meta=self._new(a.mem, (S * F,), frame_meta_dtype), taps=self._new(a.mem, (S * F, self.fft_len), np.complex64),
                 hres=self._new(a.mem, (S * F,), header_result_dtype))
        self._ck(self._lib.dtl_b200_rx_headers(self._h, a.ptr, S, F, _Arg(o["meta"]).ptr, _Arg(o["taps"]).ptr,
                                               _Arg(o["hres"]).ptr, a.mem))
        return o

    def rx_iq(self, iq, n_streams, frames_per_stream, out=None, bytes_cap=None, tb_cap=None):
        """The receive path from IQ alone: header branch, then the payload branch on the tags it produced."""
        from ._capi import header_result_dtype
        S, F = int(n_streams), int(frames_per_stream)
        a = _Arg(iq)
        mem = a.mem
        if out is None:
            bytes_cap = bytes_cap or (self.bytes_stride + 8) * (F + 1)
            tb_cap = tb_cap or 2 * F + 1
            out = dict(res=self._new(mem, (S * F,), frame_result_dtype), bytes=self._new(mem, (S, bytes_cap), np.uint8),
                       nbytes=self._new(mem, (S,), np.int32), tbs=self._new(mem, (S, tb_cap), tb_result_dtype),
                       ntbs=self._new(mem, (S,), np.int32), frame_crc=self._new(mem, (S * F,), np.int32),
                       meta=self._new(mem, (S * F,), frame_meta_dtype), hres=self._new(mem, (S * F,), header_result_dtype))
        o = out
        bytes_cap = o["bytes"].shape[1]
        tb_cap = o["tbs"].shape[1] if not _is_torch(o["tbs"]) else o["tbs"].numel() // (S * tb_result_dtype.itemsize)
        g = lambda k: _Arg(o.get(k)).ptr
        self._ck(self._lib.dtl_b200_rx_iq(self._h, a.ptr, S, F, g("meta"), g("hres"), g("res"), g("bytes"), bytes_cap,
                                          g("nbytes"), g("tbs"), tb_cap, g("ntbs"), g("frame_crc"), mem))
        return o

    def rx_stream(self, iq, n_streams, frames_cap, bytes_cap=None, tb_cap=None, want_frames=False, out=None):
        """The receive path from CONTINUOUS IQ (python/dtl/ofdm_receiver.py:70-97 in front of the header and payload branches):
        Schmidl & Cox synchroniser, fine-CFO mixer, trigger regulariser, then rx_iq on the frames found.  iq: [S, T] the next T
        samples of every stream (state is kept between calls).  Returns the rx_iq outputs for S x frames_cap frame slots plus
        n_frames [S] and frame_start [S * frames_cap] (-1 = empty slot)."""
        from ._capi import header_result_dtype
        S, F = int(n_streams), int(frames_cap)
        a = _Arg(iq)
        mem = a.mem
        T = (iq.numel() if _is_torch(iq) else iq.size) // S
        if out is not None:                                # preallocated buffers (same keys): no allocation in a timed loop
            o = out
            bytes_cap = o["bytes"].shape[1]
            tb_cap = o["tbs"].shape[1] if not _is_torch(o["tbs"]) else o["tbs"].numel() // (S * tb_result_dtype.itemsize)
        else:
            bytes_cap = bytes_cap or (self.bytes_stride + 8) * (F + 1)
            tb_cap = tb_cap or 2 * F + 1
            o = dict(res=self._new(mem, (S * F,), frame_result_dtype), bytes=self._new(mem, (S, bytes_cap), np.uint8),
                     nbytes=self._new(mem, (S,), np.int32), tbs=self._new(mem, (S, tb_cap), tb_result_dtype),
                     ntbs=self._new(mem, (S,), np.int32), frame_crc=self._new(mem, (S * F,), np.int32),
                     meta=self._new(mem, (S * F,), frame_meta_dtype), hres=self._new(mem, (S * F,), header_result_dtype),
                     n_frames=self._new(mem, (S,), np.int32), frame_start=self._new(mem, (S * F,), np.int64))
            if want_frames:
                o["frames"] = self._new(mem, (S * F, self.frame_samples), np.complex64)
        g = lambda k: _Arg(o.get(k)).ptr
        self._ck(self._lib.dtl_b200_rx_stream(self._h, a.ptr, S, T, F, g("n_frames"), g("frame_start"), g("frames"), g("meta"),
                                              g("hres"), g("res"), g("bytes"), bytes_cap, g("nbytes"), g("tbs"), tb_cap,
                                              g("ntbs"), g("frame_crc"), mem))
        return o

    # ---- the whole payload branch ------------------------------------------------------------------------------
    def rx(self, inp, taps, meta, n_streams, frames_per_stream, freq_domain=False, out=None, bytes_cap=None, tb_cap=None):
        """IQ (or frequency-domain symbols) -> decoded payload bytes per stream.  `out` may carry preallocated
        buffers (same keys as the returned dict) to keep allocations out of a timed loop."""
        S, F = int(n_streams), int(frames_per_stream)
        a_i, a_t, a_m = _Arg(inp), _Arg(taps), _Arg(meta)
        mem = _mem_of(a_i, a_t, a_m)
        if out is None:
            bytes_cap = bytes_cap or (self.bytes_stride + 8) * (F + 1)
            tb_cap = tb_cap or 2 * F + 1
            out = dict(res=self._new(mem, (S * F,), frame_result_dtype), bytes=self._new(mem, (S, bytes_cap), np.uint8),
                       nbytes=self._new(mem, (S,), np.int32), tbs=self._new(mem, (S, tb_cap), tb_result_dtype),
                       ntbs=self._new(mem, (S,), np.int32), frame_crc=self._new(mem, (S * F,), np.int32))
        o = out
        bytes_cap = o["bytes"].shape[1]
        tb_cap = o["tbs"].shape[1] if not _is_torch(o["tbs"]) else o["tbs"].numel() // (S * tb_result_dtype.itemsize)
        g = lambda k: _Arg(o.get(k)).ptr
        self._ck(self._lib.dtl_b200_rx(self._h, int(freq_domain), a_i.ptr, a_t.ptr, a_m.ptr, S, F, g("res"), g("bytes"),
                                       bytes_cap, g("nbytes"), g("tbs"), tb_cap, g("ntbs"), g("frame_crc"), mem))
        return o

    # ---- stage 5: chunks_to_symbols + carrier allocator + IFFT + cyclic prefixer (ofdm_transmitter.py:95-98,168-189)
    def fec_frame(self, data, cnst, fec_id, S, F, in_len=None):
        """TX FEC framer (dtl.ofdm_adaptive_fec_frame_bvb, lib/dtl/ofdm_adaptive_fec_frame_bvb_impl.cc:343-432): user bytes
        u8 [S, in_cap], per-frame constellation and FEC id int32 [S*F] -> dict(chunks u8 [S*F, frame_capacity],
        meta [S*F], consumed int32 [S], tbs_started int32 [S]).  The TB in flight continues in the next call."""
        a_d, a_c, a_f = _Arg(data), _Arg(cnst), _Arg(fec_id)
        a_l = _Arg(in_len) if in_len is not None else None
        mem = _mem_of(a_d, a_c, a_f) if a_l is None else _mem_of(a_d, a_c, a_f, a_l)
        in_cap = int(data.shape[-1]) if len(data.shape) > 1 else int(data.shape[0]) // max(S, 1)
        out = dict(chunks=self._new(mem, (S * F, self.frame_capacity), np.uint8),
                   meta=self._new(mem, (S * F,), _capi.frame_meta_dtype),
                   consumed=self._new(mem, (S,), np.int32), tbs_started=self._new(mem, (S,), np.int32))
        self._ck(self._lib.dtl_b200_fec_frame(self._h, a_d.ptr, in_cap, a_l.ptr if a_l is not None else None, a_c.ptr, a_f.ptr,
                                              S, F, _Arg(out["chunks"]).ptr, _Arg(out["meta"]).ptr,
                                              _Arg(out["consumed"]).ptr, _Arg(out["tbs_started"]).ptr, mem))
        return out

    def tx(self, hdr_bits, chunks, meta):
        a_h, a_c, a_m = _Arg(hdr_bits), _Arg(chunks), _Arg(meta)
        mem = _mem_of(a_h, a_c, a_m)
        n = chunks.shape[0]
        out = self._new(mem, (n, self.frame_samples), np.complex64)
        self._ck(self._lib.dtl_b200_tx(self._h, a_h.ptr, a_c.ptr, a_m.ptr, n, _Arg(out).ptr, mem))
        return out
