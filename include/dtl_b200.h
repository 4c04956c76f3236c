/* dtl_b200.h -- C ABI of libdtl_b200.so: gr-dtl's adaptive-OFDM receive hot path (and the matching
 * TX map/IFFT path) as hand-written CUDA kernels for sm_100a.
 *
 * This is the drop-in boundary: the GNU Radio block shims (gr-dtl_b200/gr_shim/, same block names,
 * io signatures, stream tags and make() parameters as the reference) and the Python mirror
 * (gr-dtl_b200/dtl_b200/) call ONLY these entry points.  Plain pointers and sizes, no C++/torch/GNU
 * Radio types, no exceptions across the boundary: every call returns a dtl_status.
 *
 * Batching.  The reference's scheduler hands each block one tagged frame per work() call; here every
 * call takes `n_streams x frames_per_stream` frames at once.  Frames are laid out stream-major,
 * f = s * frames_per_stream + j; frames of one stream are consecutive in time.  The stream tags the
 * reference reads on the first item of a frame travel as one `dtl_frame_meta` per frame, and the tags /
 * messages it emits come back as one `dtl_frame_result` per frame (+ `dtl_tb_result` per transport block).
 *
 * Memory.  Every data pointer of a call is a HOST pointer when `mem == DTL_MEM_HOST` (the call stages
 * through pinned buffers, copies in, runs, copies out and returns when the outputs are valid) or a
 * DEVICE pointer of the handle's device when `mem == DTL_MEM_DEVICE` (the call only enqueues kernels
 * on the handle's stream; use dtl_b200_sync before reading).  Output pointers may be NULL = not wanted.
 *
 * Threading: a handle is not thread-safe (GNU Radio never re-enters work() of one block instance);
 * distinct handles are independent and own one CUDA stream each.
 *
 * Each entry point cites the reference interface (file:line under mihaipstef/gr-dtl) it replaces.
 */
#ifndef DTL_B200_H
#define DTL_B200_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DTL_B200_ABI_VERSION 1

typedef enum dtl_status {
    DTL_OK = 0,
    DTL_ERR_INVALID_ARG = 1,    /* std::invalid_argument in the reference (unknown constellation, ...) */
    DTL_ERR_MISSING_TAG = 2,    /* "Missing constellation tag" / "Tags missing" / "FEC tags missing" */
    DTL_ERR_CUDA = 3,           /* CUDA runtime failure; dtl_b200_last_error() has the text */
    DTL_ERR_NO_DEVICE = 4,      /* no CUDA device: there is NO CPU fallback */
    DTL_ERR_UNSUPPORTED = 5,    /* configuration outside what the kernels cover */
    DTL_ERR_CAPACITY = 6,       /* an output buffer / internal table was too small */
    DTL_ERR_IO = 7              /* alist file unreadable / malformed */
} dtl_status;

typedef enum dtl_mem { DTL_MEM_HOST = 0, DTL_MEM_DEVICE = 1 } dtl_mem;

/* constellation_type_t, include/gnuradio/dtl/ofdm_adaptive_utils.h:22-28; 5 and 6 are extensions that
 * still fit the 4-bit header field (lib/dtl/ofdm_adaptive_packet_header.cc:157,173). */
enum { DTL_CNST_UNKNOWN = 0, DTL_CNST_BPSK = 1, DTL_CNST_QPSK = 2, DTL_CNST_PSK8 = 3, DTL_CNST_QAM16 = 4,
       DTL_CNST_QAM64 = 5, DTL_CNST_QAM256 = 6, DTL_CNST_MAX = 6 };

/* POD mirror of the stream tags on the first item of a frame
 * (lib/dtl/ofdm_adaptive_utils.cc:23-39 key names; read at
 *  lib/dtl/ofdm_adaptive_frame_equalizer_vcvc_impl.cc:114-150,
 *  lib/dtl/ofdm_adaptive_constellation_soft_cf_impl.cc:84-111, lib/dtl/fec_utils.cc:60-102). */
typedef struct dtl_frame_meta {
    int32_t constellation;   /* "constellation_tag_key" */
    int32_t carrier_offset;  /* "ofdm_sync_carr_offset" */
    int32_t frame_no;        /* "frame_no" (mod 4096) */
    int32_t payload_len;     /* "payload_length_key", bytes */
    int32_t packet_len;      /* "packet_length": constellation symbols = ceil(payload*8/bps) */
    int32_t fec_id;          /* "fec_key" (1-based; 0 = none) */
    int32_t tb_no;           /* "fec_tb_key" */
    int32_t tb_offset;       /* "fec_offset_key", bytes */
    int32_t tb_payload;      /* "fec_tb_payload_key", bits */
    int32_t reserved;
} dtl_frame_meta;

/* Tags / feedback message the frame equaliser emits per frame
 * (lib/dtl/ofdm_adaptive_frame_equalizer_vcvc_impl.cc:196-251). */
typedef struct dtl_frame_result {
    double snr_db;           /* "estimated_snr_tag_key" */
    double noise;            /* "noise_tag_key" (linear) */
    int32_t fb_cnst;         /* "feedback_constellation_key" */
    int32_t fb_fec;          /* "fec_feedback_key" */
    int32_t emitted;         /* OFDM symbols emitted (0 when payload_len == 0, :218-254) */
    int32_t lost_frames;     /* running lost-frame count of the stream (:124-138) */
} dtl_frame_result;

/* One decoded transport block (monitor message of lib/dtl/ofdm_adaptive_fec_decoder_impl.cc:184-196). */
typedef struct dtl_tb_result {
    int32_t tb_no, crc_ok, nbytes, avg_it, frame_index;
} dtl_tb_result;

/* LDPC decoder flavour.  REFERENCE = gr-fec awgn_bp restated: flooding sum-product in the likelihood-
 * ratio domain, double precision, <= max_iterations, early exit on zero syndrome (lib/dtl/ldpc_dec.cc:55-67).
 * LAYERED = layered normalised min-sum, fp32, one codeword per warp (throughput mode; decodes to the
 * same codeword whenever both converge, iteration counts differ). */
enum { DTL_LDPC_REFERENCE = 0, DTL_LDPC_LAYERED = 1 };

/* Everything the reference passes through make() arguments of the hot-path blocks
 * (python/dtl/ofdm_receiver.py:59-246, python/dtl/ofdm_adaptive_config.py:10-89). */
typedef struct dtl_config {
    int32_t fft_len, cp_len;
    int32_t frame_len;                  /* payload OFDM symbols per frame (config.py:39) */
    int32_t header_syms;                /* 1 without FEC, 2 with (ofdm_receiver.py:66-68) */
    /* carrier plan: flattened variable-length sets, carrier indices relative to DC (negative allowed) */
    int32_t n_occupied_sets; const int32_t* occupied_set_len; const int32_t* occupied_carriers;
    int32_t n_pilot_carrier_sets; const int32_t* pilot_carrier_set_len; const int32_t* pilot_carriers;
    int32_t n_pilot_symbol_sets; const int32_t* pilot_symbol_set_len; const float* pilot_symbols; /* re,im */
    /* ofdm_adaptive_payload_equalizer::make (include/gnuradio/dtl/ofdm_adaptive_equalizer.h:106-117) */
    int32_t symbols_skipped; float alpha; int32_t input_is_shifted;
    /* ofdm_adaptive_feedback_decision(hysterisis, decision_th, lut, initial) (ofdm_receiver.py:167) */
    int32_t n_mcs; const double* mcs_snr; const int32_t* mcs_cnst; const int32_t* mcs_fec;
    int32_t initial_mcs; double hysteresis; int32_t decision_th;
    /* make_ldpc_decoders(alist_fnames) (lib/dtl/ldpc_dec.cc:23-31): fec id i+1 <- alist_paths[i] */
    int32_t n_codes; const char* const* alist_paths;
    int32_t fec_enabled;                /* FEC receive chain (soft demap + decoder) vs hard demap + frame_pack */
    int32_t max_bps;                    /* max bits/symbol over the allowed constellations */
    int32_t ldpc_max_iterations;        /* 15 (ldpc_dec.cc:27) */
    int32_t ldpc_mode;                  /* DTL_LDPC_* */
    double ldpc_sigma2;                 /* awgn_bp sigma^2; the reference never sets it -- 2.0 = "LLR in" */
    const float* sync_word1; const float* sync_word2;   /* [fft_len] re,im shifted layout; TX only, may be NULL */
    int32_t device;                     /* CUDA device ordinal */
    float sync_threshold;               /* digital.ofdm_sync_sc_cfb threshold (config.py:56; <= 0: 0.95) -- dtl_b200_rx_stream only */
    int32_t reserved[6];
} dtl_config;

typedef struct dtl_b200 dtl_b200;

/* ---- lifecycle --------------------------------------------------------------------------------------- */
int dtl_b200_abi_version(void);
/* Replaces the constructors of ofdm_adaptive_payload_equalizer, ofdm_adaptive_frame_equalizer_vcvc,
 * ofdm_adaptive_constellation_soft_cf, ofdm_adaptive_fec_decoder and make_ldpc_decoders.  Fails with
 * DTL_ERR_NO_DEVICE when no CUDA device is present. */
dtl_status dtl_b200_create(const dtl_config* cfg, dtl_b200** out);
void dtl_b200_destroy(dtl_b200* h);
const char* dtl_b200_last_error(const dtl_b200* h);   /* h may be NULL: error of the last failed create */
/* Sizes per-stream state (feedback decision, frame counters, TB reassembly carry) for n_streams and
 * resets it -- the state a fresh flowgraph of n_streams receivers would start with. */
dtl_status dtl_b200_reset(dtl_b200* h, int32_t n_streams);
dtl_status dtl_b200_sync(dtl_b200* h);
void* dtl_b200_cuda_stream(dtl_b200* h);              /* cudaStream_t of the handle */
/* derived quantities: frame_capacity (python/dtl/ofdm_adaptive.py:2-6), samples per frame incl. CP,
 * sync and header symbols, LLR stride of one frame (floats), hard-byte stride of one frame */
int32_t dtl_b200_frame_capacity(const dtl_b200* h);
int32_t dtl_b200_frame_samples(const dtl_b200* h);
int32_t dtl_b200_llr_stride(const dtl_b200* h);
int32_t dtl_b200_bytes_stride(const dtl_b200* h);
/* cldpc facts of code `fec_id` (1-based): n, k and (optionally) the column permutation. */
dtl_status dtl_b200_code_info(const dtl_b200* h, int32_t fec_id, int32_t* n, int32_t* k, int32_t* permute);
/* number of kernels this handle has launched since create (bench.py's gpu_launches) */
uint64_t dtl_b200_launch_count(const dtl_b200* h);
/* Per-stage device time.  While enabled the handle brackets every stage's kernels with CUDA events on its own
 * stream; dtl_b200_stage_times waits for the stream and returns, per stage, the accumulated milliseconds and
 * the number of kernel launches since the last reset (ms / launches may be NULL; n <= DTL_STAGE_COUNT). */
enum { DTL_STAGE_FFT = 0, DTL_STAGE_EQ = 1, DTL_STAGE_STATE = 2, DTL_STAGE_TBPLAN = 3, DTL_STAGE_LDPC = 4,
       DTL_STAGE_TBFINISH = 5, DTL_STAGE_TX = 6, DTL_STAGE_DEMAP = 7, DTL_STAGE_HEADER = 8, DTL_STAGE_SYNC = 9, DTL_STAGE_COUNT = 10 };
dtl_status dtl_b200_profile(dtl_b200* h, int enable);
dtl_status dtl_b200_stage_times(dtl_b200* h, double* ms, int64_t* launches, int32_t n, int reset);
/* Page-locked host memory for DTL_MEM_HOST buffers (what a GNU Radio custom buffer would hand to work()):
 * copies from/to such memory overlap with the kernels of the call; pageable memory works too, slower. */
dtl_status dtl_b200_host_alloc(void** out, size_t bytes);
void dtl_b200_host_free(void* p);

/* ---- stage 1: CP strip + FFT ---------------------------------------------------------------------------
 * digital.header_payload_demux(output_symbols=True) + fft.fft_vcc(fft_len, True, (), True) on the payload
 * symbols (python/dtl/ofdm_receiver.py:79-86,153).  iq: c64 [n_frames][frame_samples];
 * freq: c64 [n_frames][frame_len][fft_len], DC at index fft_len/2, unnormalised. */
dtl_status dtl_b200_demod(dtl_b200* h, const float* iq, int64_t n_frames, float* freq, int mem);
/* fft.fft_vcc(fft_len, forward, (), shift=True) on n_vec contiguous vectors
 * (python/dtl/ofdm_receiver.py:101,153; python/dtl/ofdm_transmitter.py:176-181). */
dtl_status dtl_b200_fft(dtl_b200* h, const float* in, int64_t n_vec, int forward, float* out, int mem);

/* ---- stages 2+3: frame equaliser (+ serialiser + demapper) ------------------------------------------------
 * ofdm_adaptive_frame_equalizer_vcvc_impl::work (lib/dtl/ofdm_adaptive_frame_equalizer_vcvc_impl.cc:99-255)
 * -> ofdm_adaptive_equalizer_base::frame_equalize (lib/dtl/ofdm_adaptive_equalizer.cc:217-268)
 * -> digital.ofdm_serializer_vcc (python/dtl/ofdm_receiver.py:175-180)
 * -> ofdm_adaptive_constellation_soft_cf_impl::general_work (lib/dtl/ofdm_adaptive_constellation_soft_cf_impl.cc:68-156)
 *    or, without FEC, ofdm_adaptive_constellation_decoder_cb_impl::work (..._decoder_cb_impl.cc:69-93) +
 *    ofdm_adaptive_frame_pack_bb_impl::work (lib/dtl/ofdm_adaptive_frame_pack_bb_impl.cc:73-123),
 * fused into one pass over the frame.
 *   freq      c64 [S*F][frame_len][fft_len]   (in)
 *   taps      c64 [S*F][fft_len]              (in)  "ofdm_sync_chan_taps"
 *   meta          [S*F]                        (in)
 *   out_hard  c64 [S*F][frame_len][fft_len]   (out, optional) port 0
 *   out_soft  c64 [S*F][frame_len][fft_len]   (out, optional) port 1 (data carriers only, as the reference)
 *   taps_out  c64 [S*F][fft_len]              (out, optional) channel state after the frame (:186-193)
 *   res           [S*F]                        (out, optional)
 *   llr       f32 [S*F][llr_stride]           (out, optional) LSB-first LLRs, + = bit 1; n_llr[S*F]
 *   bytes     u8  [S*F][bytes_stride]         (out, optional) no-FEC payload bytes (CRC stripped),
 *   nbytes, crc_ok i32 [S*F]                   (out, optional) per frame
 */
dtl_status dtl_b200_equalize(dtl_b200* h, const float* freq, const float* taps, const dtl_frame_meta* meta,
                             int32_t n_streams, int32_t frames_per_stream,
                             float* out_hard, float* out_soft, float* taps_out, dtl_frame_result* res,
                             float* llr, int32_t* n_llr, uint8_t* bytes, int32_t* nbytes, int32_t* crc_ok, int mem);

/* ---- the equaliser pass without any allocation ----------------------------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call cudaMalloc (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same arithmetic and the same kernels as
 * dtl_b200_equalize for the FEC chain's outputs -- res (tags + feedback), llr / n_llr (nullable) -- but
 *   * every pointer is a DEVICE pointer of the handle's device; the call only enqueues on the handle's stream
 *     (dtl_b200_sync before reading) and calls no cudaMalloc / cudaFree or allocating library routine;
 *   * the per-frame scratch (|equalised pilot| magnitudes between the equaliser pass and the SNR-tag pass) is a static device
 *     array inside the library: DTL_B200_NOALLOC_MAX_PILOT_FLOATS floats = 2 MiB that every GPU which loads libdtl_b200.so holds
 *     until exit.  n_streams * frames_per_stream * (pilot carriers per frame) beyond it -> DTL_ERR_CAPACITY
 *     (64-carrier PHY, 20 symbols: 80 pilots per frame = 6553 frames per call);
 *   * per-stream state must already be sized: dtl_b200_reset(h, n_streams) before the first call, else DTL_ERR_INVALID_ARG;
 *     carrier plans the lean kernel does not cover -> DTL_ERR_UNSUPPORTED.
 * All calls on one device share the scratch array, which keeps the last call's contents; every slice that is read is written
 * first by the same call, so nothing carries over -- but calls that OVERLAP on the GPU (two handles / streams of one device)
 * race on it: serialise them. */
#define DTL_B200_NOALLOC_MAX_PILOT_FLOATS 524288
dtl_status dtl_b200_equalize_noalloc(dtl_b200* h, const float* freq, const float* taps, const dtl_frame_meta* meta,
                                     int32_t n_streams, int32_t frames_per_stream,
                                     dtl_frame_result* res, float* llr, int32_t* n_llr);

/* ---- stage 3 stand-alone blocks -----------------------------------------------------------------------------
 * ofdm_adaptive_constellation_soft_cf: syms c64 [n_frames][stride_syms], per frame meta.constellation,
 * meta.packet_len symbols, noise[n_frames] = "noise_tag_key" (NULL or <0 entry: sigma = 0.001, :89);
 * llr f32 [n_frames][stride_syms*bps_max]. */
dtl_status dtl_b200_soft_demap(dtl_b200* h, const float* syms, int32_t stride_syms, const dtl_frame_meta* meta,
                               const double* noise, int64_t n_frames, float* llr, int32_t llr_stride, int mem);
/* ofdm_adaptive_constellation_decoder_cb::work: decision_maker per symbol. chunks u8 same shape as syms. */
dtl_status dtl_b200_hard_demap(dtl_b200* h, const float* syms, int32_t stride_syms, const dtl_frame_meta* meta,
                               int64_t n_frames, uint8_t* chunks, int mem);
/* ofdm_adaptive_chunks_to_symbols_bc_impl::work (lib/dtl/ofdm_adaptive_chunks_to_symbols_bc_impl.cc:59-81). */
dtl_status dtl_b200_map(dtl_b200* h, const uint8_t* chunks, int32_t stride, const dtl_frame_meta* meta,
                        int64_t n_frames, float* syms, int mem);

/* ---- stage 4: FEC decoder -----------------------------------------------------------------------------------
 * ofdm_adaptive_fec_decoder_impl::general_work (lib/dtl/ofdm_adaptive_fec_decoder_impl.cc:80-217) +
 * tb_decoder (lib/dtl/tb_decoder.cc:32-195) + ldpc_dec::decode (lib/dtl/ldpc_dec.cc:62-71) +
 * ofdm_adaptive_fec_pack_bb (lib/dtl/ofdm_adaptive_fec_pack_bb_impl.cc:48-69).
 *   llr    f32 [S*F][llr_stride] (in), meta [S*F] (in)
 *   bytes  u8 [S][bytes_cap]: payload bytes of the CRC-passing TBs of each stream, in order; nbytes[S]
 *   tbs       [S][tb_cap], ntbs[S]: every TB released in this call
 */
dtl_status dtl_b200_fec_decode(dtl_b200* h, const float* llr, int32_t llr_stride, const dtl_frame_meta* meta,
                               int32_t n_streams, int32_t frames_per_stream,
                               uint8_t* bytes, int32_t bytes_cap, int32_t* nbytes,
                               dtl_tb_result* tbs, int32_t tb_cap, int32_t* ntbs, int mem);
/* fec_dec::decode(const float* in, int* nit, unsigned char* out) (include/gnuradio/dtl/fec.h:25-31),
 * `count` codewords per call: llr f32 [count][n] wire order, bits u8 [count][k], nit i32 [count]. */
dtl_status dtl_b200_ldpc_decode(dtl_b200* h, int32_t fec_id, const float* llr, int64_t count,
                                uint8_t* bits, int32_t* nit, int mem);

/* fec_enc::encode(const unsigned char* in_data, int len, unsigned char* out_data) (include/gnuradio/dtl/fec.h:17-23; ldpc_enc::encode lib/dtl/ldpc_enc.cc:58-68)
 * for `count` blocks: data u8 [count][k] one bit per byte -> wire u8 [count][n] = [parity (n-k) | data (k)]. */
dtl_status dtl_b200_ldpc_encode(dtl_b200* h, int32_t fec_id, const uint8_t* data, int64_t count, uint8_t* wire, int mem);

/* ---- the whole receive path (stages 1-4 back to back, intermediates stay in HBM) -----------------------------
 * in_kind 0: iq c64 [S*F][frame_samples]; 1: freq c64 [S*F][frame_len][fft_len] (stage 1 skipped).
 * With FEC: bytes/nbytes/tbs/ntbs as dtl_b200_fec_decode.  Without: bytes [S][bytes_cap] = concatenated
 * frame payloads, frame_crc[S*F]. */
dtl_status dtl_b200_rx(dtl_b200* h, int in_kind, const float* in, const float* taps, const dtl_frame_meta* meta,
                       int32_t n_streams, int32_t frames_per_stream, dtl_frame_result* res,
                       uint8_t* bytes, int32_t bytes_cap, int32_t* nbytes,
                       dtl_tb_result* tbs, int32_t tb_cap, int32_t* ntbs, int32_t* frame_crc, int mem);

/* ---- stage 5: TX map + carrier allocation + IFFT + cyclic prefix ------------------------------------------------
 * ofdm_adaptive_chunks_to_symbols_bc + digital.ofdm_carrier_allocator_cvc + fft.fft_vcc(N, False, (), True) +
 * digital.ofdm_cyclic_prefixer (python/dtl/ofdm_transmitter.py:95-98,168-189).
 *   hdr_bits u8 [n_frames][header_syms*D] (BPSK header chunks), chunks u8 [n_frames][frame_capacity],
 *   iq c64 [n_frames][frame_samples] */
dtl_status dtl_b200_tx(dtl_b200* h, const uint8_t* hdr_bits, const uint8_t* chunks, const dtl_frame_meta* meta,
                       int64_t n_frames, float* iq, int mem);

/* ---- TX FEC framer ----------------------------------------------------------------------
 * dtl.ofdm_adaptive_fec_frame_bvb (lib/dtl/ofdm_adaptive_fec_frame_bvb_impl.cc:343-432) with tb_encoder
 * (lib/dtl/tb_encoder.cc:32-74,88-99), crc_util::append_crc (lib/dtl/crc_util.cc:31-39) and ldpc_enc::encode
 * (lib/dtl/ldpc_enc.cc:58-68): user bytes -> frames of constellation chunks + the tags the header formatter and the
 * modulator (dtl_b200_tx) consume.  S streams x F consecutive frames per call; the TB in flight at the end of the call is
 * continued by the next call (per-stream state, cleared by dtl_b200_reset).  The reference block is paced by wall clock and
 * by what its input holds: here a TB is started whenever the frame has room and the stream's input still holds a whole TB
 * payload (never on less).
 *   data  u8  [S][in_cap] (in), in_len [S] valid bytes per stream (NULL: in_cap)
 *   cnst, fec_id [S*F] (in): constellation and 1-based FEC id of each frame (the MCS the feedback selected)
 *   chunks u8 [S*F][frame_capacity] (out), meta [S*F] (out: constellation, frame_no, payload_len, packet_len, fec_id,
 *   tb_no, tb_offset, tb_payload), consumed [S] (out, optional: input bytes used), tbs_started [S] (out, optional) */
dtl_status dtl_b200_fec_frame(dtl_b200* h, const uint8_t* data, int32_t in_cap, const int32_t* in_len,
                              const int32_t* cnst, const int32_t* fec_id, int32_t n_streams, int32_t frames_per_stream,
                              uint8_t* chunks, dtl_frame_meta* meta, int32_t* consumed, int32_t* tbs_started, int mem);

/* ---- header branch: the tags of a frame from its IQ alone ------------------------
 * python/dtl/ofdm_receiver.py:100-150: fft_vcc on the two sync symbols and the header symbols,
 * digital.ofdm_chanest_vcvc(sync_word1, sync_word2, header_len) (integer carrier offset + channel taps),
 * digital.ofdm_frame_equalizer_vcvc around dtl.ofdm_adaptive_header_equalizer (BPSK, symbols_skipped 0, alpha 0.1,
 * propagate_channel_state: lib/dtl/ofdm_adaptive_equalizer.cc:161-174 -> frame_equalize :217-268),
 * digital.ofdm_serializer_vcc, constellation_decoder_cb(BPSK) and packet_headerparser_b on
 * ofdm_adaptive_packet_header::header_parser (lib/dtl/ofdm_adaptive_packet_header.cc:232-312).
 *   iq    c64 [S*F][frame_samples] (in)
 *   meta      [S*F] (out): the stream tags the payload branch receives; a header whose CRC-16 fails is dropped by
 *             header_payload_demux: meta.reserved = 1, every other field but carrier_offset 0
 *   taps  c64 [S*F][fft_len] (out): "ofdm_sync_chan_taps" as the payload equaliser receives it (the header equaliser's
 *             channel state after the header symbols)
 *   hres      [S*F] (out, nullable): header CRC flag and the feedback fields the peer wrote into the header
 * The parser's fallback constellation (ofdm_adaptive_packet_header.cc:70,287-289) is per-stream state (dtl_b200_reset). */
typedef struct dtl_header_result {
    int32_t crc_ok;          /* 1: header parsed; 0: CRC-16 mismatch, frame dropped */
    int32_t feedback_cnst;   /* "feedback_constellation_key" */
    int32_t feedback_fec;    /* "fec_feedback_key" */
    int32_t reserved;
} dtl_header_result;
dtl_status dtl_b200_rx_headers(dtl_b200* h, const float* iq, int32_t n_streams, int32_t frames_per_stream,
                               dtl_frame_meta* meta, float* taps, dtl_header_result* hres, int mem);

/* Header branch + dtl_b200_rx(in_kind 0) back to back: the receive path from IQ alone.  Dropped frames (header CRC)
 * take no part in the payload branch (no tags, no counters, no bytes), as in the flowgraph.  meta_out / hres nullable. */
dtl_status dtl_b200_rx_iq(dtl_b200* h, const float* iq, int32_t n_streams, int32_t frames_per_stream,
                          dtl_frame_meta* meta_out, dtl_header_result* hres, dtl_frame_result* res,
                          uint8_t* bytes, int32_t bytes_cap, int32_t* nbytes,
                          dtl_tb_result* tbs, int32_t tb_cap, int32_t* ntbs, int32_t* frame_crc, int mem);

/* ---- synchronisation front end + the receive path: continuous IQ in, payload bytes out ------
 * python/dtl/ofdm_receiver.py:70-97: digital.ofdm_sync_sc_cfb(fft_len, cp_len, threshold) (Schmidl & Cox timing metric, plateau
 * detector, fine frequency estimate), blocks.delay(fft_len + cp_len) + analog.frequency_modulator_fc(-2 / fft_len) +
 * blocks.multiply_cc (fine CFO removed), dtl.ofdm_adaptive_frame_detect_bb (lib/dtl/ofdm_adaptive_frame_detect_bb_impl.cc:64-156,
 * trigger regulariser) and the trigger input of digital.header_payload_demux; then dtl_b200_rx_iq on the frames found.
 *   iq          c64 [S][samples_per_stream] (in): the next samples of S continuous streams.  Detector, oscillator, regulariser
 *               and the samples a later call still needs are per-stream state (cleared by dtl_b200_reset); one call = one
 *               work() call of the regulariser (its output depends on the chunking, as in the reference).
 *   frames_cap  F: frame slots per stream in every per-frame output; frames completed by this call fill slots 0..n_frames[s]-1,
 *               the other slots are reported as dropped frames (meta.reserved = 1).  More than F frames: DTL_ERR_CAPACITY.
 *   n_frames [S] (out), frame_start i64 [S*F] (out, nullable): position of each frame's first sample in the stream, -1 = empty
 *   frames_out  c64 [S*F][frame_samples] (out, nullable): the frames as header_payload_demux receives them
 * The first call after dtl_b200_reset fixes the largest samples_per_stream later calls may use. */
dtl_status dtl_b200_rx_stream(dtl_b200* h, const float* iq, int32_t n_streams, int64_t samples_per_stream, int32_t frames_cap,
                              int32_t* n_frames, int64_t* frame_start, float* frames_out,
                              dtl_frame_meta* meta_out, dtl_header_result* hres, dtl_frame_result* res,
                              uint8_t* bytes, int32_t bytes_cap, int32_t* nbytes,
                              dtl_tb_result* tbs, int32_t tb_cap, int32_t* ntbs, int32_t* frame_crc, int mem);

#ifdef __cplusplus
}
#endif
#endif /* DTL_B200_H */
