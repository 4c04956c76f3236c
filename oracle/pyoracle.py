"""ctypes binding of the CPU oracle (oracle/_build/liboracle.so).

TEST INFRASTRUCTURE ONLY: imported by tests/, by __graft_entry__.smoke() and by bench.py's
cpu_baseline / --impl reference legs -- never by the product package (gr-dtl_b200/).
Parity status of the oracle itself: see oracle/src/oracle.hpp.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "liboracle.so")

# constellation_type_t (include/gnuradio/dtl/ofdm_adaptive_utils.h:22-28) + extension 5, 6
UNKNOWN, BPSK, QPSK, PSK8, QAM16, QAM64, QAM256 = range(7)

frame_meta_dtype = np.dtype([
    ("constellation", "<i4"), ("carrier_offset", "<i4"), ("frame_no", "<i4"), ("payload_len", "<i4"),
    ("packet_len", "<i4"), ("fec_id", "<i4"), ("tb_no", "<i4"), ("tb_offset", "<i4"), ("tb_payload", "<i4"),
    ("reserved", "<i4")])
frame_result_dtype = np.dtype([
    ("snr_db", "<f8"), ("noise", "<f8"), ("fb_cnst", "<i4"), ("fb_fec", "<i4"), ("emitted", "<i4"),
    ("lost_frames", "<i4")])
tb_result_dtype = np.dtype([("tb_no", "<i4"), ("crc_ok", "<i4"), ("nbytes", "<i4"), ("avg_it", "<i4"),
                            ("frame_index", "<i4")])


def build(force=False):
    if force or not os.path.exists(_LIB_PATH):
        subprocess.check_call(["make", "-C", _HERE], stdout=subprocess.DEVNULL)
    return _LIB_PATH


def build_native():
    """-O3 -march=native build on THIS machine (bench.py's CPU arm: the timed baseline is compiled for the host it runs
    on); falls back to the portable library when the compiler is missing.  Call before lib()."""
    global _LIB_PATH
    native = os.path.join(_HERE, "_build", "liboracle_native.so")
    try:
        subprocess.check_call(["make", "-C", _HERE, "native"], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        if _lib is None:
            _LIB_PATH = native
        return "-O3 -march=native (built on this host)"
    except Exception:
        return "-O3 -march=x86-64-v3 (portable build; native build failed on this host)"


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        L.orc_cfg_create.restype = C.c_void_p
        L.orc_fb_create.restype = C.c_void_p
        L.orc_feq_create.restype = C.c_void_p
        L.orc_fdec_create.restype = C.c_void_p
        L.orc_hdr_create.restype = C.c_void_p
        L.orc_crc.restype = C.c_uint64
        L.orc_min_dist2.restype = C.c_double
        L.orc_rx_batch.restype = C.c_double
        _lib = L
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _c64_as_f32(a):
    a = np.ascontiguousarray(a, dtype=np.complex64)
    return a


# ---- primitives ---------------------------------------------------------------------------------
def bps(cnst):
    return lib().orc_bps(int(cnst))


def points(cnst):
    out = np.zeros(256, dtype=np.complex64)
    n = lib().orc_points(int(cnst), _p(out))
    return out[:n].copy()


def min_dist2(cnst):
    return lib().orc_min_dist2(int(cnst))


def decide(cnst, sym):
    sym = _c64_as_f32(sym)
    out = np.zeros(sym.size, dtype=np.uint8)
    lib().orc_decide(int(cnst), _p(sym), sym.size, _p(out))
    return out


def map_chunks(cnst, chunks):
    chunks = np.ascontiguousarray(chunks, dtype=np.uint8)
    out = np.zeros(chunks.size, dtype=np.complex64)
    lib().orc_map(int(cnst), _p(chunks), chunks.size, _p(out))
    return out


def soft_dec(cnst, sym, npwr):
    sym = _c64_as_f32(sym)
    out = np.zeros(sym.size * bps(cnst), dtype=np.float32)
    lib().orc_soft_dec(int(cnst), _p(sym), sym.size, C.c_float(npwr), _p(out))
    return out


def crc(nbits, poly, init, xorout, in_refl, out_refl, data):
    data = np.ascontiguousarray(data, dtype=np.uint8)
    return int(lib().orc_crc(C.c_uint(nbits), C.c_uint64(poly), C.c_uint64(init), C.c_uint64(xorout),
                             int(in_refl), int(out_refl), _p(data), C.c_size_t(data.size)))


def crc32(data):
    return crc(32, 0x04C11DB7, 0xFFFFFFFF, 0xFFFFFFFF, True, True, data)


def repack(data, bits_in, bits_out, lsb_first=True):
    data = np.ascontiguousarray(data, dtype=np.uint8)
    out = np.zeros(data.size * bits_in // bits_out + 2, dtype=np.uint8)
    n = lib().orc_repack(_p(data), C.c_size_t(data.size), bits_in, bits_out, int(lsb_first), _p(out))
    return out[:n].copy()


def frame_pack(cnst, chunks):
    chunks = np.ascontiguousarray(chunks, dtype=np.uint8)
    out = np.zeros(chunks.size + 8, dtype=np.uint8)
    ok = C.c_int(0)
    n = lib().orc_frame_pack(int(cnst), _p(chunks), chunks.size, _p(out), C.byref(ok))
    return out[:n].copy(), bool(ok.value)


class FeedbackDecision:
    def __init__(self, hyst, th, lut, initial=0):
        snr = np.array([x[0] for x in lut], dtype=np.float64)
        cn = np.array([x[1][0] for x in lut], dtype=np.int32)
        fe = np.array([x[1][1] for x in lut], dtype=np.int32)
        self.h = C.c_void_p(lib().orc_fb_create(C.c_double(hyst), th, len(lut), _p(snr), _p(cn), _p(fe), initial))

    def get_feedback(self, snr):
        c, f = C.c_int(0), C.c_int(0)
        lib().orc_fb_get(self.h, C.c_double(snr), C.byref(c), C.byref(f))
        return c.value, f.value

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_fb_destroy(self.h)
            self.h = None


def snr_simple(x):
    x = _c64_as_f32(x)
    a, b, c = C.c_double(0), C.c_double(0), C.c_double(0)
    lib().orc_snr_simple(_p(x), x.size, C.byref(a), C.byref(b), C.byref(c))
    return a.value, b.value, c.value


def fft_shifted(x, n, forward=True):
    x = _c64_as_f32(x)
    out = np.zeros_like(x)
    lib().orc_fft(_p(x), _p(out), n, int(forward), x.size // n)
    return out


# ---- configuration ---------------------------------------------------------------------------------
# python/dtl/ofdm_adaptive_config.py:21-30
PILOT_SCRAMBLE_SEQ = (
    1, 1, 1, 1, -1, -1, -1, 1, -1, -1, -1, -1, 1, 1, -1, 1, -1, -1, 1, 1, -1, 1, 1, -1, 1, 1, 1, 1, 1, 1, -1, 1,
    1, 1, -1, 1, 1, -1, -1, 1, 1, 1, -1, 1, -1, -1, -1, 1, -1, 1, -1, -1, 1, -1, -1, 1, 1, 1, 1, 1, -1, -1, 1, 1,
    -1, -1, 1, -1, 1, -1, 1, 1, -1, -1, -1, 1, 1, -1, -1, -1, -1, 1, -1, -1, 1, -1, 1, 1, 1, 1, -1, 1, -1, 1, -1, 1,
    -1, -1, -1, -1, -1, 1, -1, 1, 1, -1, 1, -1, 1, 1, 1, -1, -1, 1, -1, -1, -1, 1, 1, 1, -1, -1, -1, -1, -1, -1, -1)


def default_carriers(fft_len=64):
    """64: python/dtl/ofdm_adaptive_config.py:17-20.  1024: this repo's scaled-PHY extension
    (SURVEY.md 8d config 5): carriers -416..416 without DC, every 13th a pilot."""
    if fft_len == 64:
        occ = (list(range(-26, -21)) + list(range(-20, -7)) + list(range(-6, 0)) + list(range(1, 7)) +
               list(range(8, 21)) + list(range(22, 27)))
        pil = [-21, -7, 7, 21]
        return occ, pil
    if fft_len == 1024:
        allc = [c for c in range(-416, 417) if c != 0]
        pil = [c for c in allc if c % 13 == 0]
        occ = [c for c in allc if c % 13 != 0]
        return occ, pil
    raise ValueError(fft_len)


def sync_words(fft_len, occupied, pilots):
    """gr-digital ofdm_txrx._make_sync_word1/2 (RECALLED): numpy legacy RandomState seed 42."""
    active = set()
    for c in list(occupied) + list(pilots):
        active.add(c + fft_len if c < 0 else c)
    rs = np.random.RandomState(42)
    bpsk = {0: np.sqrt(2), 1: -np.sqrt(2)}
    sw1 = [bpsk[rs.randint(2)] if (x in active and x % 2) else 0 for x in range(fft_len)]
    rs = np.random.RandomState(42)
    bpsk2 = {0: 1, 1: -1}
    sw2 = [bpsk2[rs.randint(2)] if x in active else 0 for x in range(fft_len)]
    sw2[0] = 0
    return (np.fft.fftshift(np.array(sw1)).astype(np.complex64),
            np.fft.fftshift(np.array(sw2)).astype(np.complex64))


class Config:
    """Stream configuration (ofdm_adaptive_config defaults + ofdm_receiver wiring)."""

    def __init__(self, fft_len=64, cp_len=None, frame_len=20, occupied=None, pilots=None, pilot_symbols=None,
                 n_pilot_carrier_sets=2, alists=(), mcs=None, initial_mcs=0, constellations=(1, 2, 3, 4), fec=None,
                 pilot_carrier_sets=None, pilot_symbol_sets=None):
        if cp_len is None:
            cp_len = fft_len // 4
        if occupied is None:
            occupied, pilots = default_carriers(fft_len)
        self.fft_len, self.cp_len, self.frame_len = fft_len, cp_len, frame_len
        self.occupied, self.pilots = list(occupied), list(pilots)
        npil = len(self.pilots)
        if pilot_symbols is None:
            # config.py:31-32 generalised: pattern (x, x, x, -x) repeated over the pilot carriers
            pat = np.array([1, 1, 1, -1] * ((npil + 3) // 4))[:npil]
            pilot_symbols = np.array([[x * p for p in pat] for x in PILOT_SCRAMBLE_SEQ], dtype=np.complex64)
        self.pilot_symbols = np.ascontiguousarray(pilot_symbols, dtype=np.complex64).reshape(-1, max(npil, 1))
        self.n_pilot_carrier_sets = n_pilot_carrier_sets if npil else 0
        occ = np.array(self.occupied, dtype=np.int32)
        pil = np.array(self.pilots * max(self.n_pilot_carrier_sets, 1), dtype=np.int32)
        self.h = C.c_void_p(lib().orc_cfg_create(fft_len, cp_len, frame_len, occ.size, _p(occ),
                                                 self.n_pilot_carrier_sets, npil, _p(pil),
                                                 self.pilot_symbols.shape[0] if npil else 0, _p(self.pilot_symbols)))
        if pilot_carrier_sets is not None:
            cl = np.array([len(x) for x in pilot_carrier_sets], dtype=np.int32)
            cf = np.array([c for x in pilot_carrier_sets for c in x], dtype=np.int32)
            sl = np.array([len(x) for x in pilot_symbol_sets], dtype=np.int32)
            sf = np.array([c for x in pilot_symbol_sets for c in x], dtype=np.complex64)
            lib().orc_cfg_set_pilots(self.h, cl.size, _p(cl), _p(cf), sl.size, _p(sl), _p(sf))
        self.alists = list(alists)
        for a in self.alists:
            if lib().orc_cfg_add_code(self.h, a.encode()) < 0:
                raise RuntimeError("cannot load alist " + a)
        self.fec = bool(self.alists) if fec is None else bool(fec)
        lib().orc_cfg_set_fec(self.h, int(self.fec))
        self.constellations = list(constellations)
        cn = np.array(self.constellations, dtype=np.int32)
        lib().orc_cfg_set_constellations(self.h, cn.size, _p(cn))
        if mcs is None:
            # python/dtl/ofdm_adaptive_config.py:40-41 (no FEC)
            mcs = [(2.2250738585072014e-308, (BPSK, 0)), (13, (QPSK, 0)), (18, (PSK8, 0)), (23, (QAM16, 0))]
        self.mcs, self.initial_mcs = list(mcs), initial_mcs
        snr = np.array([m[0] for m in mcs], dtype=np.float64)
        mc = np.array([m[1][0] for m in mcs], dtype=np.int32)
        mf = np.array([m[1][1] for m in mcs], dtype=np.int32)
        lib().orc_cfg_set_mcs(self.h, len(mcs), _p(snr), _p(mc), _p(mf), initial_mcs)
        self.sync_word1, self.sync_word2 = sync_words(fft_len, self.occupied, self.pilots)
        lib().orc_cfg_set_sync(self.h, _p(self.sync_word1), _p(self.sync_word2))
        self.capacity = lib().orc_cfg_capacity(self.h)
        self.frame_samples = lib().orc_cfg_frame_samples(self.h)
        self.header_syms = lib().orc_cfg_header_syms(self.h)
        self.max_bps = max(bps(c) for c in self.constellations)

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_cfg_destroy(self.h)
            self.h = None

    def set_literal(self, on=True):
        """rx_batch's payload equaliser evaluates the reference's own std::complex<float> expressions
        (lib/dtl/ofdm_adaptive_equalizer.cc:241-256) instead of the arithmetic contract of src/contract.hpp"""
        lib().orc_cfg_set_literal(self.h, int(on))
        return self

    # ---- LDPC ----
    def code_info(self, fec_id):
        N, K, M = C.c_int(0), C.c_int(0), C.c_int(0)
        perm = np.zeros(4096, dtype=np.int32)
        edges = lib().orc_code_info(self.h, fec_id, C.byref(N), C.byref(K), C.byref(M), _p(perm))
        return dict(N=N.value, K=K.value, M=M.value, edges=edges, permute=perm[:N.value].copy())

    def ldpc_encode(self, fec_id, data):
        info = self.code_info(fec_id)
        data = np.ascontiguousarray(data, dtype=np.uint8).reshape(-1, info["K"])
        out = np.zeros((data.shape[0], info["N"]), dtype=np.uint8)
        lib().orc_ldpc_encode(self.h, fec_id, _p(data), data.shape[0], _p(out))
        return out

    def ldpc_decode(self, fec_id, llr):
        info = self.code_info(fec_id)
        llr = np.ascontiguousarray(llr, dtype=np.float32).reshape(-1, info["N"])
        out = np.zeros((llr.shape[0], info["K"]), dtype=np.uint8)
        nit = np.zeros(llr.shape[0], dtype=np.int32)
        lib().orc_ldpc_decode(self.h, fec_id, _p(llr), llr.shape[0], _p(out), _p(nit))
        return out, nit

    def ldpc_syndrome(self, fec_id, wire):
        wire = np.ascontiguousarray(wire, dtype=np.uint8)
        return lib().orc_ldpc_syndrome(self.h, fec_id, _p(wire))

    # ---- batch TX / RX ----
    def tx_batch(self, S, F, seed, cnst, fec_id=0, snr_db=30.0, nthreads=8, cnst_per_frame=None,
                 want_iq=True, want_taps=True):
        cap, N = self.capacity, self.fft_len
        chunks = np.zeros((S * F, cap), dtype=np.uint8)
        meta = np.zeros(S * F, dtype=frame_meta_dtype)
        iq = np.zeros((S * F, self.frame_samples), dtype=np.complex64) if want_iq else None
        taps = np.zeros((S * F, N), dtype=np.complex64) if want_taps else None
        pcap = (cap * 8 // 8 + 8) * F
        payload = np.zeros((S, pcap), dtype=np.uint8)
        plen = np.zeros(S, dtype=np.int32)
        ptbs = np.zeros(S, dtype=np.int32)
        cpf = None if cnst_per_frame is None else np.ascontiguousarray(cnst_per_frame, dtype=np.int32)
        rc = lib().orc_tx_batch(self.h, S, F, C.c_uint64(seed), int(cnst), int(fec_id), C.c_double(snr_db), nthreads,
                                _p(cpf), _p(chunks), _p(meta), _p(iq), _p(taps), _p(payload), pcap, _p(plen), _p(ptbs))
        if rc != 0:
            raise RuntimeError("orc_tx_batch failed")
        return dict(chunks=chunks, meta=meta, iq=iq, taps=taps, payload=[payload[s, :plen[s]].copy() for s in range(S)],
                    payload_tbs=ptbs)

    def rx_batch(self, inp, taps, meta, S, F, freq_domain=False, nthreads=8, want_llr=True, want_syms=True):
        cap, N, L = self.capacity, self.fft_len, self.frame_len
        inp = _c64_as_f32(inp)
        taps = None if taps is None else _c64_as_f32(taps)
        meta = np.ascontiguousarray(meta, dtype=frame_meta_dtype)
        res = np.zeros(S * F, dtype=frame_result_dtype)
        stride = cap * 8
        llr = np.zeros((S * F, stride), dtype=np.float32) if want_llr else None
        n_llr = np.zeros(S * F, dtype=np.int32)
        hard = np.zeros((S * F, L * N), dtype=np.complex64) if want_syms else None
        soft = np.zeros((S * F, L * N), dtype=np.complex64) if want_syms else None
        bcap = (cap + 8) * F
        by = np.zeros((S, bcap), dtype=np.uint8)
        nby = np.zeros(S, dtype=np.int32)
        tcap = 2 * F + 2
        tbs = np.zeros((S, tcap), dtype=tb_result_dtype)
        ntbs = np.zeros(S, dtype=np.int32)
        fcrc = np.zeros(S * F, dtype=np.int32)
        secs = lib().orc_rx_batch(self.h, int(freq_domain), _p(inp), _p(taps), _p(meta), S, F, nthreads, _p(res),
                                  _p(llr), stride, _p(n_llr), _p(hard), _p(soft), _p(by), bcap, _p(nby), _p(tbs), tcap,
                                  _p(ntbs), _p(fcrc))
        return dict(res=res, llr=llr, n_llr=n_llr, hard=hard, soft=soft,
                    bytes=[by[s, :nby[s]].copy() for s in range(S)],
                    tbs=[tbs[s, :ntbs[s]].copy() for s in range(S)], frame_crc=fcrc, seconds=secs)

    def rx_headers(self, iq, S, F, nthreads=8):
        """Header branch of the receiver (oracle/src/header.cc): the tags the payload branch receives, from IQ alone."""
        iq = _c64_as_f32(iq)
        meta = np.zeros(S * F, dtype=frame_meta_dtype)
        taps = np.zeros((S * F, self.fft_len), dtype=np.complex64)
        ok = np.zeros(S * F, dtype=np.int32)
        fb = np.zeros((S * F, 2), dtype=np.int32)
        if lib().orc_rx_headers(self.h, _p(iq), S, F, nthreads, _p(meta), _p(taps), _p(ok), _p(fb)) != 0:
            raise RuntimeError("orc_rx_headers failed (sync words not configured)")
        return dict(meta=meta, taps=taps, ok=ok, fb=fb)

    def frame_batch(self, data, cnst, fec_id, S, F, in_len=None, state=None):
        """TX FEC framer alone (oracle/src/txrx.cc tx_framer fed with the caller's bytes): data u8 [S, in_cap], cnst /
        fec_id int32 [S*F] -> dict(chunks u8 [S*F, capacity], meta [S*F], consumed [S], started [S]).  `state`: a list
        returned by an earlier call (the TB in flight continues), or None."""
        data = np.ascontiguousarray(data, dtype=np.uint8).reshape(S, -1)
        cnst = np.ascontiguousarray(cnst, dtype=np.int32).reshape(S * F)
        fec_id = np.ascontiguousarray(fec_id, dtype=np.int32).reshape(S * F)
        chunks = np.zeros((S * F, self.capacity), dtype=np.uint8)
        meta = np.zeros(S * F, dtype=frame_meta_dtype)
        consumed = np.zeros(S, dtype=np.int32)
        started = np.zeros(S, dtype=np.int32)
        st = state if state is not None else (C.c_void_p * S)()
        il = None if in_len is None else np.ascontiguousarray(in_len, dtype=np.int32)
        f = lib().orc_frame_batch
        f.restype = C.c_int
        f.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p,
                      C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        if f(self.h, S, F, _p(cnst), _p(fec_id), _p(data), data.shape[1], None if il is None else _p(il), _p(chunks), _p(meta),
             _p(consumed), _p(started), C.cast(st, C.c_void_p)) != 0:
            raise RuntimeError("orc_frame_batch failed")
        return dict(chunks=chunks, meta=meta, consumed=consumed, started=started, state=st)

    def rx_demod(self, iq):
        iq = _c64_as_f32(iq).reshape(-1, self.frame_samples)
        out = np.zeros((iq.shape[0], self.frame_len * self.fft_len), dtype=np.complex64)
        lib().orc_rx_demod(self.h, _p(iq), iq.shape[0], _p(out))
        return out

    def serialize(self, frame, packet_len):
        frame = _c64_as_f32(frame)
        out = np.zeros(self.capacity, dtype=np.complex64)
        n = lib().orc_serialize(self.h, _p(frame), self.frame_len, int(packet_len), _p(out))
        return out[:n].copy()

    def tx_modulate(self, hdr_syms, payload_syms):
        out = np.zeros(self.frame_samples, dtype=np.complex64)
        lib().orc_tx_modulate_frame(self.h, _p(_c64_as_f32(hdr_syms)), _p(_c64_as_f32(payload_syms)), _p(out))
        return out


class StreamSync:
    """Synchronisation front end of one continuous IQ stream (oracle/src/sync.cc): push(chunk) -> frames cut at the
    regularised Schmidl & Cox triggers, fine CFO removed."""

    def __init__(self, cfg, threshold=0.95):
        self.cfg = cfg
        f = lib().orc_sync_create
        f.restype = C.c_void_p
        self.h = C.c_void_p(f(cfg.h, C.c_float(threshold)))

    def push(self, x, frame_cap=64, trig_cap=256):
        x = _c64_as_f32(x)
        fs = self.cfg.frame_samples
        frames = np.zeros((frame_cap, fs), dtype=np.complex64)
        starts = np.zeros(frame_cap, dtype=np.int64)
        raw, fixed = np.zeros(trig_cap, dtype=np.int64), np.zeros(trig_cap, dtype=np.int64)
        nr, nf = C.c_int(0), C.c_int(0)
        n = lib().orc_sync_push(self.h, _p(x), x.size, _p(frames), _p(starts), frame_cap, _p(raw), C.byref(nr), _p(fixed),
                                C.byref(nf), trig_cap)
        if n > frame_cap:
            raise RuntimeError("frame_cap too small")
        return dict(frames=frames[:n].copy(), starts=starts[:n].copy(), raw=raw[:nr.value].copy(), fixed=fixed[:nf.value].copy())

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_sync_destroy(self.h)
            self.h = None


class FrameEqualizer:
    def __init__(self, cfg, literal=False, symbols_skipped=None, alpha=0.1):
        self.cfg = cfg
        if symbols_skipped is None:
            symbols_skipped = cfg.header_syms
        self.h = C.c_void_p(lib().orc_feq_create(cfg.h, int(literal), symbols_skipped, C.c_float(alpha)))

    def work(self, frame, taps, meta, n_sym=None):
        N = self.cfg.fft_len
        frame = _c64_as_f32(frame)
        n_sym = frame.size // N if n_sym is None else n_sym
        hard = np.zeros(n_sym * N, dtype=np.complex64)
        soft = np.zeros(n_sym * N, dtype=np.complex64)
        res = np.zeros(1, dtype=frame_result_dtype)
        tout = np.zeros(N, dtype=np.complex64)
        m = np.ascontiguousarray(meta, dtype=frame_meta_dtype).reshape(1)
        tp = None if taps is None else _c64_as_f32(taps)
        r = lib().orc_feq_work(self.h, _p(frame), n_sym, _p(tp), _p(m), _p(hard), _p(soft), _p(res), _p(tout))
        return r, hard, soft, res[0], tout

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_feq_destroy(self.h)
            self.h = None


class FecDecoder:
    def __init__(self, cfg):
        self.cfg = cfg
        self.h = C.c_void_p(lib().orc_fdec_create(cfg.h))
        self.idx = 0

    def frame(self, llr, meta):
        llr = np.ascontiguousarray(llr, dtype=np.float32)
        out = np.zeros(4 * self.cfg.capacity + 64, dtype=np.uint8)
        tbs = np.zeros(4, dtype=tb_result_dtype)
        nt = C.c_int(0)
        m = np.ascontiguousarray(meta, dtype=frame_meta_dtype).reshape(1)
        n = lib().orc_fdec_frame(self.h, _p(llr), _p(m), self.idx, _p(out), out.size, _p(tbs), 4, C.byref(nt))
        self.idx += 1
        return out[:n].copy(), tbs[:nt.value].copy()

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_fdec_destroy(self.h)
            self.h = None


class PacketHeader:
    def __init__(self, header_len_bits, payload_syms, has_fec, max_cnst=QAM16):
        self.n = header_len_bits
        self.h = C.c_void_p(lib().orc_hdr_create(header_len_bits, payload_syms, int(has_fec), max_cnst))

    def format(self, meta, fb_cnst=0, fb_fec=0):
        out = np.zeros(self.n, dtype=np.uint8)
        m = np.ascontiguousarray(meta, dtype=frame_meta_dtype).reshape(1)
        lib().orc_hdr_format(self.h, _p(m), fb_cnst, fb_fec, _p(out))
        return out

    def parse(self, bits):
        bits = np.ascontiguousarray(bits, dtype=np.uint8)
        m = np.zeros(1, dtype=frame_meta_dtype)
        a, b = C.c_int(0), C.c_int(0)
        ok = lib().orc_hdr_parse(self.h, _p(bits), _p(m), C.byref(a), C.byref(b))
        return bool(ok), m[0], a.value, b.value

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_hdr_destroy(self.h)
            self.h = None
