"""Mirror of python/dtl/ofdm_adaptive_config.py:10-89 (same field names and defaults) plus the helpers of
python/dtl/ofdm_adaptive.py:2-8.  `digital.ofdm_txrx._make_sync_word1/2` are restated here because GNU Radio is
not a dependency of this package."""
import dataclasses as dc
import enum
import sys
import typing as t

import numpy as np


class constellation_type_t(enum.IntEnum):
    """include/gnuradio/dtl/ofdm_adaptive_utils.h:22-28; QAM64/QAM256 are this package's extension."""
    UNKNOWN = 0
    BPSK = 1
    QPSK = 2
    PSK8 = 3
    QAM16 = 4
    QAM64 = 5
    QAM256 = 6


_BPS = {0: 0, 1: 1, 2: 2, 3: 3, 4: 4, 5: 6, 6: 8}


def get_bits_per_symbol(cnst):
    """lib/dtl/ofdm_adaptive_utils.cc:64-70"""
    return _BPS[int(cnst)]


def frame_capacity(n_ofdm_syms, ocupied_carriers):
    """python/dtl/ofdm_adaptive.py:2-6"""
    return sum(len(ocupied_carriers[i % len(ocupied_carriers)]) for i in range(n_ofdm_syms))


def max_bps(cnsts):
    """Bits per symbol of the densest allowed constellation.  (python/dtl/ofdm_adaptive.py:8-9 takes the max of
    the enum VALUES, which equals bits/symbol only up to 16QAM; with the QAM64/256 extension the bit count is
    what tb_decoder's capacity needs.)"""
    return max(get_bits_per_symbol(c) for c in cnsts)


def _make_sync_word1(fft_len, occupied_carriers, pilot_carriers):
    active = _active(fft_len, occupied_carriers, pilot_carriers)
    rs = np.random.RandomState(42)
    bpsk = {0: np.sqrt(2), 1: -np.sqrt(2)}
    sw1 = [bpsk[rs.randint(2)] if (x in active and x % 2) else 0 for x in range(fft_len)]
    return np.fft.fftshift(np.array(sw1)).astype(np.complex64)


def _make_sync_word2(fft_len, occupied_carriers, pilot_carriers):
    active = _active(fft_len, occupied_carriers, pilot_carriers)
    rs = np.random.RandomState(42)
    bpsk = {0: 1, 1: -1}
    sw2 = [bpsk[rs.randint(2)] if x in active else 0 for x in range(fft_len)]
    sw2[0] = 0
    return np.fft.fftshift(np.array(sw2)).astype(np.complex64)


def _active(fft_len, occupied_carriers, pilot_carriers):
    active = set()
    for sets in (occupied_carriers, pilot_carriers):
        for s in sets:
            for c in s:
                active.add(c + fft_len if c < 0 else c)
    return active


_OCC64 = (list(range(-26, -21)) + list(range(-20, -7)) + list(range(-6, 0)) + list(range(1, 7)) +
          list(range(8, 21)) + list(range(22, 27)),)
_PIL64 = ((-21, -7, 7, 21,), (-21, -7, 7, 21,))
_SCRAMBLE = (
    1, 1, 1, 1, -1, -1, -1, 1, -1, -1, -1, -1, 1, 1, -1, 1, -1, -1, 1, 1, -1, 1, 1, -1, 1, 1, 1, 1, 1, 1, -1, 1,
    1, 1, -1, 1, 1, -1, -1, 1, 1, 1, -1, 1, -1, -1, -1, 1, -1, 1, -1, -1, 1, -1, -1, 1, 1, 1, 1, 1, -1, -1, 1, 1,
    -1, -1, 1, -1, 1, -1, 1, 1, -1, -1, -1, 1, 1, -1, -1, -1, -1, 1, -1, -1, 1, -1, 1, 1, 1, 1, -1, 1, -1, 1, -1, 1,
    -1, -1, -1, -1, -1, 1, -1, 1, 1, -1, 1, -1, 1, 1, 1, -1, -1, 1, -1, -1, -1, 1, 1, 1, -1, -1, -1, -1, -1, -1, -1)


@dc.dataclass
class ofdm_adaptive_config:
    fft_len: int = 64
    cp_len: int = 16
    frame_length_tag_key: str = "frame_length"
    packet_length_tag_key: str = "packet_length"
    frame_no_tag_key: str = "frame_no"
    occupied_carriers: t.Tuple[t.List[int]] = _OCC64
    pilot_carriers: t.Tuple[t.Tuple[int]] = _PIL64
    pilot_sym_scramble_seq: t.Tuple[int] = _SCRAMBLE
    pilot_symbols: t.Optional[tuple] = None
    sync_word1: t.Optional[t.Any] = None
    sync_word2: t.Optional[t.Any] = None
    rolloff: int = 0
    scramble_bits: bool = False
    frame_length: int = 20
    frame_store_folder: str = "/tmp"
    fec: bool = False
    fec_codes: t.Tuple = ()          # (("name", "file.alist"), ...)
    mcs: t.Tuple = ((sys.float_info.min, (constellation_type_t.BPSK, "no_fec")),
                    (13, (constellation_type_t.QPSK, "no_fec")),
                    (18, (constellation_type_t.PSK8, "no_fec")),
                    (23, (constellation_type_t.QAM16, "no_fec")),)
    initial_mcs_id: int = 0

    def __post_init__(self):
        if self.pilot_symbols is None:
            npil = len(self.pilot_carriers[0]) if self.pilot_carriers else 0
            pat = ([1, 1, 1, -1] * ((npil + 3) // 4))[:npil]
            self.pilot_symbols = tuple(tuple(x * p for p in pat) for x in self.pilot_sym_scramble_seq)
        if self.sync_word1 is None:
            self.sync_word1 = _make_sync_word1(self.fft_len, self.occupied_carriers, self.pilot_carriers)
        if self.sync_word2 is None:
            self.sync_word2 = _make_sync_word2(self.fft_len, self.occupied_carriers, self.pilot_carriers)


@dc.dataclass
class ofdm_adaptive_rx_config(ofdm_adaptive_config):
    sync_threshold: float = 0.95
    use_sync_correct: bool = True


@dc.dataclass
class ofdm_adaptive_tx_config(ofdm_adaptive_config):
    max_empty_frames: int = -1
    sample_rate: int = 700000


def scaled_phy_carriers(fft_len=1024):
    """Carrier plan of the scaled PHY (BASELINE.json config 5; not in the reference): carriers -416..416
    without DC, every 13th a pilot."""
    allc = [c for c in range(-416, 417) if c != 0]
    pil = tuple(c for c in allc if c % 13 == 0)
    occ = [c for c in allc if c % 13 != 0]
    return (occ,), (pil, pil)
