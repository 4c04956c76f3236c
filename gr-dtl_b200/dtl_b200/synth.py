"""Synthetic loopback traffic for the receive path: the TX framing the reference does on the host
(CRC-32 append, transport-block split with shortening, bit -> chunk repack, per-frame tags) restated for
S streams at once, with the LDPC systematic encode, the mapper/IFFT/CP modulator and the channel running on
the GPU through libdtl_b200.

  lib/dtl/ofdm_adaptive_fec_frame_bvb_impl.cc:343-432   TB construction, CRC-32, tags (FEC framer)
  lib/dtl/tb_encoder.cc:32-74,88-99                     codeword split with shortening, buf_out repack
  lib/dtl/ldpc_enc.cc:58-68                             systematic encode (GPU: dtl_b200_ldpc_encode)
  lib/dtl/ofdm_adaptive_frame_bb_impl.cc:236-262        no-FEC frame: payload + CRC-32, full frames
  lib/dtl/fec_utils.cc:104-121                          compute_tb_len, align_bits_to_bytes
  python/dtl/ofdm_transmitter.py:95-98,168-189          mapper + allocator + IFFT + CP (GPU: dtl_b200_tx)

All streams of a batch use the same constellation / code, so the framing STRUCTURE (which TB bits go to which
frame, every tag value) is computed once in plain Python and applied to all streams with array operations.
The reference's framers are paced by wall clock and the no-FEC one pads with std::random_device bytes
(ofdm_adaptive_frame_bb_impl.cc:355-364); here frames are always full and seeded.
"""
import zlib

import numpy as np

from ._capi import frame_meta_dtype
from .config import get_bits_per_symbol


def compute_tb_len(cw_len, frame_len_bits):
    """lib/dtl/fec_utils.cc:104-112: codewords per transport block"""
    return 1 + frame_len_bits // cw_len if frame_len_bits > cw_len else 1


def _align(bits):
    """lib/dtl/fec_utils.cc:114-121"""
    return (bits + 7) // 8


class _Plan:
    """Framing structure of F consecutive frames of one stream (identical for every stream of the batch)."""

    def __init__(self):
        self.tb_payload_bytes = []      # user bytes of TB t (without CRC)
        self.tb_ncws = []
        self.tb_bits = []               # encoded length of TB t in bits
        self.segs = []                  # (frame, first chunk, tb, first TB bit, n bits)
        self.meta = None
        self.tb_done_frame = []         # frame in which the last bit of TB t went out (-1: still in flight)


def fec_plan(F, cap, bps, n, k, fec_id=1):
    """Replays the FEC framer's per-frame decisions (fec_frame_bvb_impl.cc:343-432); tests/test_synth.py pins the tags
    against the test-suite's CPU framer and the round trip through the receive path."""
    ncheck = n - k
    p = _Plan()
    meta = np.zeros(F, dtype=frame_meta_dtype)
    cur = -1                            # TB in flight
    sent = 0                            # bits of it already sent
    syms = lambda nbytes: (nbytes * 8 + bps - 1) // bps

    def new_tb():
        frame_bytes = cap * bps // 8
        ncws = compute_tb_len(n, frame_bytes * 8)
        payload_max = (ncws * k) // 8 - 4
        if payload_max <= 0:
            raise ValueError("Misconfiguration: TB should be able to carry at least 1 user data bit")
        p.tb_payload_bytes.append(payload_max)
        p.tb_ncws.append(ncws)
        p.tb_bits.append(ncws * ncheck + (payload_max + 4) * 8)
        p.tb_done_frame.append(-1)
        return len(p.tb_bits) - 1

    for j in range(F):
        used = payload = offset = 0
        started = False
        if cur >= 0 and sent < p.tb_bits[cur]:
            remaining = p.tb_bits[cur] - sent
            out_bytes = min(_align(remaining), (cap - used) * bps // 8)
            nb = min(out_bytes * 8, remaining)
            p.segs.append((j, used, cur, sent, nb))
            sent += nb
            used += syms(out_bytes)
            payload += out_bytes
            if sent == p.tb_bits[cur]:
                p.tb_done_frame[cur] = j
        if cur < 0 or sent == p.tb_bits[cur]:
            avail = (cap - used) * bps // 8
            if avail > 0:
                offset = payload
                cur, sent, started = new_tb(), 0, True
                remaining = p.tb_bits[cur]
                out_bytes = min(_align(remaining), avail)
                extra = bps - (offset * 8) % bps if (offset * 8) % bps else 0
                while out_bytes > 1 and extra and (out_bytes * 8) % bps and (out_bytes * 8) % bps <= extra:
                    out_bytes -= 1
                nb = min(out_bytes * 8, remaining)
                p.segs.append((j, used, cur, 0, nb))
                sent += nb
                used += syms(out_bytes)
                payload += out_bytes
                if sent == p.tb_bits[cur]:
                    p.tb_done_frame[cur] = j
        m = meta[j]
        m["constellation"] = 0          # filled by the caller
        m["frame_no"] = j & 0xFFF
        m["payload_len"] = payload
        m["packet_len"] = syms(payload)
        m["fec_id"] = fec_id
        m["tb_no"] = len(p.tb_bits) & 0xFF
        m["tb_offset"] = offset if started else 0
        m["tb_payload"] = ((p.tb_payload_bytes[cur] + 4) * 8) & 0xFFFF
    p.meta = meta
    return p


def _crc32_rows(rows):
    """CRC-32 of every row (crc_util::append_crc, lib/dtl/crc_util.cc:31-39: zlib polynomial, little-endian)"""
    out = np.empty((rows.shape[0], 4), dtype=np.uint8)
    for i in range(rows.shape[0]):
        out[i] = np.frombuffer(np.uint32(zlib.crc32(rows[i].tobytes())).tobytes(), dtype=np.uint8)
    return out


def _bits_to_chunks(bits, bps):
    """repack(1 -> bps) LSB-first (lib/testbed/repack.cc:31-69); a partial last chunk is zero-padded"""
    n = bits.shape[-1]
    nsym = (n + bps - 1) // bps
    if nsym * bps != n:
        pad = np.zeros(bits.shape[:-1] + (nsym * bps - n,), dtype=np.uint8)
        bits = np.concatenate([bits, pad], axis=-1)
    b = bits.reshape(bits.shape[:-1] + (nsym, bps)).astype(np.uint16)
    w = (1 << np.arange(bps, dtype=np.uint16))
    return (b * w).sum(axis=-1).astype(np.uint8)


def format_headers(meta, nbits, has_fec, feedback_cnst=0, feedback_fec=0):
    """ofdm_adaptive_packet_header::header_formatter (lib/dtl/ofdm_adaptive_packet_header.cc:141-200, no scrambling)
    for an array of frames: payload length 12 bits, frame number 12, constellation 4, feedback constellation 4,
    [TB number 12, FEC feedback 4, TB offset 12, FEC id 4, TB payload 16,] CRC-16 -- every field LSB first, one bit
    per item.  The CRC (0x1021, init 0xFFFF, output reflected) runs over the bits packed MSB first the way
    pack_crc does (:99-111, including its `i*len + j < header_len` bound)."""
    n = meta.shape[0]
    bits = np.zeros((n, nbits), dtype=np.uint8)
    k = 0

    def put(val, nb):
        nonlocal k
        v = np.asarray(val, dtype=np.int64)
        for i in range(nb):
            if k < nbits:
                bits[:, k] = (v >> i) & 1
            k += 1

    put(meta["payload_len"] & 0xFFF, 12)
    put(meta["frame_no"] & 0xFFF, 12)
    put(meta["constellation"] & 0xF, 4)
    put(np.full(n, feedback_cnst & 0xF), 4)
    if has_fec:
        put(meta["tb_no"], 12)
        put(np.full(n, feedback_fec), 4)
        put(meta["tb_offset"], 12)
        put(meta["fec_id"], 4)
        put(meta["tb_payload"], 16)
    ln = (nbits - 16) // 8 + (1 if (nbits - 16) % 8 else 0)
    crc = np.full(n, 0xFFFF, dtype=np.int64)
    for i in range(ln):
        byte = np.zeros(n, dtype=np.int64)
        for j in range(8):
            if i * ln + j < nbits:
                byte = ((byte << 1) | bits[:, i * 8 + j]) & 0xFF
        crc ^= byte << 8
        for _ in range(8):
            crc = np.where(crc & 0x8000, (crc << 1) ^ 0x1021, crc << 1) & 0xFFFF
    refl = np.zeros(n, dtype=np.int64)
    for i in range(16):
        refl |= ((crc >> i) & 1) << (15 - i)
    put(refl, 16)
    return bits


def make_frames(rx, S, F, cnst, fec_id=0, seed=1):
    """Frames of S streams x F frames: dict(chunks u8 [S*F, cap], meta [S*F], hdr_bits u8 [S*F, H*D],
    payload = list of S byte arrays (user bytes of every TB / frame completed inside the F frames))."""
    cnst = int(cnst)
    bps, cap = get_bits_per_symbol(cnst), rx.frame_capacity
    rng = np.random.Generator(np.random.PCG64(seed))
    chunks = np.zeros((S, F, cap), dtype=np.uint8)
    meta = np.zeros((S, F), dtype=frame_meta_dtype)
    if not fec_id:
        in_bytes = cap * bps // 8 - 4
        data = rng.integers(0, 256, size=(S * F, in_bytes), dtype=np.uint8)
        full = np.concatenate([data, _crc32_rows(data)], axis=1)
        bits = np.unpackbits(full, axis=1, bitorder="little")
        ch = _bits_to_chunks(bits, bps)
        chunks.reshape(S * F, cap)[:, :ch.shape[1]] = ch
        meta["constellation"] = cnst
        meta["frame_no"] = (np.arange(F) & 0xFFF)[None, :]
        meta["payload_len"] = in_bytes + 4
        meta["packet_len"] = ch.shape[1]
        payload = [data[s * F:(s + 1) * F].reshape(-1) for s in range(S)]
        tb_bytes, n_done = in_bytes, F
    else:
        info = rx.code_info(fec_id)
        n, k = info["N"], info["K"]
        ncheck = n - k
        plan = fec_plan(F, cap, bps, n, k, fec_id)
        T, pm, ncws = len(plan.tb_bits), plan.tb_payload_bytes[0], plan.tb_ncws[0]
        data = rng.integers(0, 256, size=(S * T, pm), dtype=np.uint8)
        full = np.concatenate([data, _crc32_rows(data)], axis=1)
        bits = np.unpackbits(full, axis=1, bitorder="little")          # [S*T, len]
        length = bits.shape[1]
        cw_in = np.zeros((S * T, ncws, k), dtype=np.uint8)
        k_new, pos, left = [], 0, length                                # tb_encoder.cc:44-52
        for i in range(ncws):
            kn = left // (ncws - i) + (1 if left % (ncws - i) else 0)
            cw_in[:, i, :kn] = bits[:, pos:pos + kn]
            k_new.append(kn)
            pos += kn
            left -= kn
        wire = rx.ldpc_encode(fec_id, cw_in.reshape(-1, k)).reshape(S * T, ncws, n)
        tb = np.concatenate([wire[:, i, :ncheck + k_new[i]] for i in range(ncws)], axis=1).reshape(S, T, -1)
        assert tb.shape[2] == plan.tb_bits[0]
        for (j, c0, t, b0, nb) in plan.segs:
            ch = _bits_to_chunks(tb[:, t, b0:b0 + nb], bps)
            chunks[:, j, c0:c0 + ch.shape[1]] = ch
        meta[:] = plan.meta[None, :]
        meta["constellation"] = cnst
        done = [t for t in range(T) if plan.tb_done_frame[t] >= 0]
        d3 = data.reshape(S, T, pm)
        payload = [d3[s, done].reshape(-1) for s in range(S)]
        tb_bytes, n_done = pm, len(done)
    hdr = format_headers(meta.reshape(S * F), rx.header_len * len(rx.config.occupied_carriers[0]), bool(fec_id))
    return dict(chunks=chunks.reshape(S * F, cap), meta=meta.reshape(S * F), hdr_bits=hdr, payload=payload,
                tb_bytes=tb_bytes, n_done=n_done)


def make_batch(rx, S, F, cnst, fec_id=0, seed=1, snr_db=30.0, device=True):
    """Frames -> IQ through the GPU modulator (dtl_b200_tx) and a flat channel h_s = (1+1j) e^{j phi_s} with AWGN at
    `snr_db`; taps = N h_s + estimation noise stand in for digital.ofdm_chanest_vcvc (python/dtl/ofdm_receiver.py:102-103),
    which is not on the payload path.  Returns torch CUDA tensors (device=True) or NumPy arrays."""
    import torch
    fr = make_frames(rx, S, F, cnst, fec_id, seed)
    dev = torch.device("cuda:%d" % rx.device)
    N, fs = rx.fft_len, rx.frame_samples
    t_chunks = torch.from_numpy(fr["chunks"]).to(dev)
    t_hdr = torch.from_numpy(fr["hdr_bits"]).to(dev)
    t_meta = torch.from_numpy(fr["meta"].view(np.uint8).reshape(S * F, -1).copy()).to(dev)
    torch.cuda.synchronize(dev)
    iq = rx.tx(t_hdr, t_chunks, t_meta)
    rx.sync()
    g = torch.Generator(device=dev)
    g.manual_seed(int(seed) * 7919 + 17)
    phi = 2 * np.pi * torch.rand(S, generator=g, device=dev, dtype=torch.float64)
    h = (torch.polar(torch.ones_like(phi), phi) * complex(1, 1)).to(torch.complex64)           # [S]
    sigma_t = float(np.sqrt(N * 2.0 / 10.0 ** (snr_db / 10.0)))
    sd = sigma_t / np.sqrt(2.0)
    iq = iq.view(S, F * fs) * h[:, None]
    iq = torch.view_as_real(iq)
    iq += sd * torch.randn(iq.shape, generator=g, device=dev, dtype=torch.float32)
    iq = torch.view_as_complex(iq).reshape(S * F, fs).contiguous()
    taps = (N * h)[:, None, None].expand(S, F, N).contiguous()
    taps = torch.view_as_real(taps)
    taps += float(sd * np.sqrt(N)) * torch.randn(taps.shape, generator=g, device=dev, dtype=torch.float32)
    taps = torch.view_as_complex(taps).reshape(S * F, N).contiguous()
    torch.cuda.synchronize(dev)
    out = dict(iq=iq, taps=taps, meta=t_meta, meta_np=fr["meta"], payload=fr["payload"], chunks=fr["chunks"],
               tb_bytes=fr["tb_bytes"], n_done=fr["n_done"])
    if not device:
        out.update(iq=iq.cpu().numpy(), taps=taps.cpu().numpy(), meta=fr["meta"])
    return out
