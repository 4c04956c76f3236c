"""dtl_b200 -- gr-dtl's adaptive-OFDM receive hot path on B200 (libdtl_b200.so + this thin mirror of the
reference's Python layer).  Importing the package loads the CUDA library; there is no fallback."""
from ._capi import (LDPC_LAYERED, LDPC_REFERENCE, MEM_DEVICE, MEM_HOST, DtlError, frame_meta_dtype, frame_result_dtype,
                    tb_result_dtype, header_result_dtype, lib, LIB_PATH)
from .config import (constellation_type_t, frame_capacity, get_bits_per_symbol, max_bps, ofdm_adaptive_config,
                     ofdm_adaptive_rx_config, ofdm_adaptive_tx_config, scaled_phy_carriers)
from .ofdm_receiver import ofdm_receiver

lib()   # fail loudly at import time when the CUDA library is missing
