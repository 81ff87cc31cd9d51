/* lora_phy_cuda.h -- C ABI of the B200 (sm_100a) batched LoRa PHY.
 *
 * This is the drop-in boundary for the reference's PHY hot path.  The reference has no
 * FFI of its own: its boundary is the C++ free-function API in
 * include/lora_phy/phy.hpp.  Each entry point below names the reference interface it
 * replaces (file:line relative to the reference tree); include/lora_phy/phy.hpp in THIS
 * repository re-implements those C++ signatures as thin inline wrappers over this ABI.
 *
 * Conventions
 *   - plain pointers and sizes only; no CUDA or torch types in any signature
 *     (`stream` is a cudaStream_t passed as void*, NULL = the context's own stream);
 *   - `d_` pointers are DEVICE pointers, `h_` pointers are HOST pointers;
 *   - IQ is interleaved float32 I,Q (std::complex<float> layout), packets are rows of a
 *     [n_packets][packet_stride] array, a packet holds whole symbols of (1<<sf)*osr samples;
 *   - every function returns 0 / a count on success or a NEGATIVE errno, exactly like the
 *     reference (phy.hpp:100-141): -EINVAL bad argument, -ERANGE capacity / <2 symbols /
 *     missing scratch, -ENOMEM missing buffer, plus -EIO for a CUDA failure
 *     (text via lora_phy_cuda_last_error);
 *   - there is NO CPU fallback: without a CUDA device create() fails with -ENODEV;
 *   - threading: a context owns the staging buffers, scratch and stream its calls use; every entry point holds
 *     the context's mutex from its first use of them to its last (the *_host variants: through the copy-out), so
 *     calls from several threads on ONE context are safe and serialise.  For parallelism use one context per
 *     host thread, like one workspace per thread in the reference (phy.hpp:74-78); contexts on the same device
 *     are independent.  Device work of one context is ordered on its stream, and the device-pointer entry
 *     points return without waiting for it.
 */
#ifndef LORA_PHY_CUDA_H
#define LORA_PHY_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LORA_PHY_CUDA_ABI_VERSION 1

typedef struct lora_phy_cuda_ctx lora_phy_cuda_ctx;

/* lora_params (phy.hpp:53-60) + the call-site choices the reference makes per function. */
typedef struct lora_phy_cuda_cfg {
    uint32_t sf;        /* spreading factor 2..12 (N = 1<<sf); phy.hpp:54 */
    uint32_t bw_hz;     /* 125000 / 250000 / 500000; phy.hpp:40-44 */
    uint32_t osr;       /* oversampling ratio >= 1; phy.hpp:57 */
    uint32_t window;    /* 0 = none, 1 = Hann; phy.hpp:31-34 */
    uint32_t flags;     /* LORA_PHY_CUDA_F_* */
    uint32_t sync_word; /* modulator only; phy.hpp:59 */
    float    amplitude; /* modulator only; clamped to [-1,1] like LoRaMod.cpp:18 */
    uint32_t reserved;
} lora_phy_cuda_cfg;

/* demodulate_batch flags */
#define LORA_PHY_CUDA_F_DECHIRP     0x1u /* high-level demodulate(): multiply by the on-chip conjugate
                                            downchirp first and estimate offsets on the raw symbols
                                            without index tie-break (phy.cpp:182-243, :116) */
#define LORA_PHY_CUDA_F_NO_SCRATCH  0x2u /* model a lora_demod_workspace without scratch: packets whose
                                            max |I|,|Q| exceeds 1 get status -ERANGE (LoRaDemod.cpp:70-72) */

/* Split one demodulate_batch call into its two launches (bench.py times the kernels separately
 * with CUDA events; a normal call sets neither):                                            */
#define LORA_PHY_CUDA_F_ONLY_PREPASS 0x10u /* run only the per-packet prepass (max-abs, estimator, sync) */
#define LORA_PHY_CUDA_F_ONLY_SYMBOLS 0x20u /* run only the data-symbol kernel, reusing the packet state
                                              the previous prepass of the same batch left in the context */

#define LORA_PHY_CUDA_F_EXACT         0x100u /* A/B: demodulate every window with the operation-for-operation replay of the
                                               reference's float arithmetic (bit-exact spectra), instead of the default
                                               margin-certified fast path that falls back to that replay per window only
                                               when the argmax is not provably the reference's.
                                               Symbols, sync word, status, cfo and time_offset are identical either way. */
#define LORA_PHY_CUDA_F_TWO_KERNEL   0x80u /* A/B: force the prepass + symbols kernel pair (demod.cu: the route of osr > 1,
                                              SF2-6 and estimate-only calls) instead of the default */

/* ---- context -------------------------------------------------------------------------- */
/* replaces lora_demod_init / init (LoRaDemod.cpp:11-33, phy.cpp:27-52): builds, ON THE DEVICE,
 * the kissfft twiddle tables, Hann windows and downchirp tables for SF 2..12.            */
int  lora_phy_cuda_create(lora_phy_cuda_ctx** out, int device);
void lora_phy_cuda_destroy(lora_phy_cuda_ctx* ctx);              /* lora_demod_free, LoRaDemod.cpp:35 */
const char* lora_phy_cuda_last_error(const lora_phy_cuda_ctx* ctx);
int  lora_phy_cuda_abi_version(void);
int  lora_phy_cuda_device(const lora_phy_cuda_ctx* ctx);
/* Number of kernels this context has launched so far (bench.py's gpu_launches). */
uint64_t lora_phy_cuda_launch_count(const lora_phy_cuda_ctx* ctx);
/* Make every later call of this context enqueue on `stream` (cudaStream_t as void*). */
int  lora_phy_cuda_set_stream(lora_phy_cuda_ctx* ctx, void* stream);
int  lora_phy_cuda_synchronize(lora_phy_cuda_ctx* ctx);
/* Counters of the default (margin-certified) demodulator, accumulated since the context was created; waits for the
 * context's stream.  out[0] windows walked by the throughput transform, out[1] windows replayed through the exact
 * chain because the proof failed (or the packet was exact throughout), out[2] packets whose two estimator symbols
 * were redone after a max-abs misprediction, out[3] packets walked a second time.  No reference counterpart.       */
int  lora_phy_cuda_cert_stats(lora_phy_cuda_ctx* ctx, uint64_t out[4]);
/* The same counters and the later ones, n <= 8 of them: out[4] windows settled by the pruned transform (one bin class,
 * accepted by the Parseval test) without running the full one; out[5] packets the streaming first pass (SF10-12) handed
 * to the certified kernel with every symbol still open; out[6..7] reserved (zero).  No reference counterpart.            */
int  lora_phy_cuda_cert_stats_ex(lora_phy_cuda_ctx* ctx, uint64_t* out, size_t n);

/* ---- demodulator: K1 + K2a ------------------------------------------------------------- */
/* replaces lora_demodulate (LoRaDemod.cpp:50-197) -- and, with F_DECHIRP, demodulate
 * (phy.cpp:182-243) -- for n_packets independent packets.
 *   d_iq            [n_packets][packet_stride] complex64
 *   d_iq            must be 8-byte aligned (complex64); it need NOT be 16-byte aligned and nothing outside
 *                   [d_iq, d_iq + n_packets*packet_stride) is ever read
 *   d_sample_counts NULL (every packet has samples_per_packet samples) or [n_packets] ragged lengths; a count above
 *                   samples_per_packet is clamped to it on the device (the *_host variant returns -EINVAL instead);
 *                   a trailing partial symbol is ignored (LoRaDemod.cpp:57)
 *   d_symbols       [n_packets][symbol_stride] uint16: the data symbols (sync symbols stripped when
 *                   the packet has >= 2 symbols, LoRaDemod.cpp:166-175)
 *   d_sync          [n_packets] recovered sync word (LoRaDemod.cpp:178-193), may be NULL
 *   d_cfo, d_time_offset  [n_packets] lora_metrics (phy.hpp:67-71), may be NULL
 *   d_status        [n_packets] symbols written, or -ERANGE, may be NULL
 * Asynchronous on the context's stream.                                                     */
int lora_phy_cuda_demodulate_batch(lora_phy_cuda_ctx* ctx, const lora_phy_cuda_cfg* cfg,
                                   const void* d_iq, size_t packet_stride,
                                   const uint32_t* d_sample_counts, size_t samples_per_packet,
                                   size_t n_packets,
                                   uint16_t* d_symbols, size_t symbol_stride,
                                   uint8_t* d_sync, float* d_cfo, float* d_time_offset,
                                   int32_t* d_status);

/* Same call with HOST buffers: stages through pinned memory and device scratch owned by the
 * context, copies results back and synchronises.  This is what the C++ wrappers use.       */
int lora_phy_cuda_demodulate_batch_host(lora_phy_cuda_ctx* ctx, const lora_phy_cuda_cfg* cfg,
                                        const void* h_iq, size_t packet_stride,
                                        const uint32_t* h_sample_counts, size_t samples_per_packet,
                                        size_t n_packets,
                                        uint16_t* h_symbols, size_t symbol_stride,
                                        uint8_t* h_sync, float* h_cfo, float* h_time_offset,
                                        int32_t* h_status);

/* replaces estimate_offsets (phy.cpp:81-148): every whole symbol of each packet contributes.
 * Outputs [n_packets] cfo / time_offset; packets with no whole symbol leave them untouched. */
int lora_phy_cuda_estimate_offsets_batch(lora_phy_cuda_ctx* ctx, const lora_phy_cuda_cfg* cfg,
                                         const void* d_iq, size_t packet_stride,
                                         const uint32_t* d_sample_counts, size_t samples_per_packet,
                                         size_t n_packets, float* d_cfo, float* d_time_offset);

/* replaces compensate_offsets (phy.cpp:150-180), in place, per packet. */
int lora_phy_cuda_compensate_offsets_batch(lora_phy_cuda_ctx* ctx, const lora_phy_cuda_cfg* cfg,
                                           void* d_iq, size_t packet_stride,
                                           size_t samples_per_packet, size_t n_packets,
                                           const float* d_cfo, const float* d_time_offset);

/* HOST-buffer variants of the two calls above (stage, run, copy back, synchronise). */
int lora_phy_cuda_estimate_offsets_batch_host(lora_phy_cuda_ctx* ctx, const lora_phy_cuda_cfg* cfg,
                                              const void* h_iq, size_t packet_stride,
                                              size_t samples_per_packet, size_t n_packets,
                                              float* h_cfo, float* h_time_offset);
int lora_phy_cuda_compensate_offsets_batch_host(lora_phy_cuda_ctx* ctx, const lora_phy_cuda_cfg* cfg,
                                                void* h_iq, size_t packet_stride,
                                                size_t samples_per_packet, size_t n_packets,
                                                const float* h_cfo, const float* h_time_offset);

/* One N-point window per row through the detector (LoRaDetector.hpp:39-74): spectrum (may be
 * NULL), first-max index, power, powerAvg and fIndex.  Used by the parity tests.           */
int lora_phy_cuda_detect_batch(lora_phy_cuda_ctx* ctx, uint32_t sf, const void* d_windows,
                               size_t n_windows, void* d_spectrum, uint32_t* d_index,
                               float* d_power, float* d_power_avg, float* d_findex);

/* ---- modulator: K3 ---------------------------------------------------------------------- */
/* replaces lora_modulate / modulate (LoRaMod.cpp:8-43, phy.cpp:68-79).
 *   d_symbols [n_packets][symbol_stride] uint16, symbols_per_packet used from each row
 *   d_iq      [n_packets][packet_stride] complex64, (symbols_per_packet+2)*N*osr samples written
 * Returns samples per packet, or -ERANGE if packet_stride is too small (checked BEFORE launch,
 * unlike phy.cpp:74-77 which reports it after writing).                                      */
long lora_phy_cuda_modulate_batch(lora_phy_cuda_ctx* ctx, const lora_phy_cuda_cfg* cfg,
                                  const uint16_t* d_symbols, size_t symbol_stride,
                                  size_t symbols_per_packet, size_t n_packets,
                                  void* d_iq, size_t packet_stride);
/* Allocation-free variant of lora_phy_cuda_modulate_batch.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same arguments, same output bit for bit; the only difference is where the modulator keeps its
 * per-chirp start phases: a static device array inside the library (4 MiB on every GPU that loads it, held until
 * exit) instead of the context's growable buffer.  The call performs no cudaMalloc / cudaFree / allocating library
 * routine.
 *   limit   n_packets * (symbols_per_packet + 2) <= LORA_PHY_CUDA_STATIC_MAX_CHIRPS, else -E2BIG with
 *           lora_phy_cuda_last_error() naming the limit (nothing is launched).
 *   sharing every call on a device -- from any context or thread -- uses the SAME array: calls whose kernels may
 *           overlap on the GPU (different streams, different contexts) race on it and must be ordered by the
 *           caller.  The array keeps the previous call's contents; the modulator overwrites every entry it reads,
 *           so nothing has to be cleared between calls.                                                          */
#define LORA_PHY_CUDA_STATIC_MAX_CHIRPS 1048576u
long lora_phy_cuda_modulate_batch_static(lora_phy_cuda_ctx* ctx, const lora_phy_cuda_cfg* cfg,
                                         const uint16_t* d_symbols, size_t symbol_stride,
                                         size_t symbols_per_packet, size_t n_packets,
                                         void* d_iq, size_t packet_stride);
long lora_phy_cuda_modulate_batch_host(lora_phy_cuda_ctx* ctx, const lora_phy_cuda_cfg* cfg,
                                       const uint16_t* h_symbols, size_t symbol_stride,
                                       size_t symbols_per_packet, size_t n_packets,
                                       void* h_iq, size_t packet_stride);

/* Upstream transmitter frame format (external/LoRa-SDR/LoRaMod.cpp:109-238 with the chirp generator of
 * external/LoRa-SDR/ChirpGenerator.hpp:22-47): per frame 10 up-chirps, sync word0 / word1 chirps at
 * (nibble*8) bins, 2 1/4 down-chirps, one up-chirp per symbol, padding_symbols symbols of zeros; the phase
 * runs through the whole frame.  cfg supplies sf, osr (power of two), amplitude (used as given, upstream
 * default 0.3) and sync_word; bw_hz / window / flags are ignored.  This is the stream
 * lora_phy_cuda_detect_stream consumes.  Returns samples per frame = lora_phy_cuda_frame_samples(...).   */
long lora_phy_cuda_frame_samples(uint32_t sf, uint32_t osr, size_t n_symbols, uint32_t padding_symbols);
long lora_phy_cuda_modulate_frames_batch(lora_phy_cuda_ctx* ctx, const lora_phy_cuda_cfg* cfg,
                                         uint32_t padding_symbols, const uint16_t* d_symbols,
                                         size_t symbol_stride, size_t symbols_per_frame, size_t n_frames,
                                         void* d_iq, size_t frame_stride);

/* Multiply every symbol window by the conjugate downchirp of (sf, osr, bw) in place: the
 * caller-side dechirp every reference test performs before lora_demodulate
 * (tests/e2e_chain_test.cpp:83-92, tests/performance_test.cpp:117-122).                      */
int lora_phy_cuda_dechirp_batch(lora_phy_cuda_ctx* ctx, const lora_phy_cuda_cfg* cfg,
                                void* d_iq, size_t packet_stride, size_t samples_per_packet,
                                size_t n_packets);

/* ---- bit pipeline: K4 (shipped coder) ----------------------------------------------------- */
/* replaces lora_encode / encode (LoRaEncoder.cpp:6-18, phy.cpp:58-66): 2 symbols per byte.   */
long lora_phy_cuda_encode_batch(lora_phy_cuda_ctx* ctx, const uint8_t* d_payload,
                                size_t payload_stride, size_t payload_len, size_t n_packets,
                                uint16_t* d_symbols, size_t symbol_stride);
/* replaces lora_decode / decode (LoRaDecoder.cpp:7-21, phy.cpp:245-261): returns bytes per
 * packet or -EINVAL for an odd symbol count; d_crc_ok[n_packets] (may be NULL) gets the
 * phy.cpp:252-259 check (bytes [2,n-2) vs little-endian CRC in the last two bytes).
 * d_symbol_counts NULL or [n_packets] int32 (e.g. the status array of demodulate_batch;
 * negative or odd counts yield 0 bytes and crc_ok = 0).                                       */
long lora_phy_cuda_decode_batch(lora_phy_cuda_ctx* ctx, const uint16_t* d_symbols,
                                size_t symbol_stride, const int32_t* d_symbol_counts,
                                size_t symbols_per_packet, size_t n_packets,
                                uint8_t* d_payload, size_t payload_stride, uint8_t* d_crc_ok);
long lora_phy_cuda_encode_batch_host(lora_phy_cuda_ctx* ctx, const uint8_t* h_payload,
                                     size_t payload_stride, size_t payload_len, size_t n_packets,
                                     uint16_t* h_symbols, size_t symbol_stride);
long lora_phy_cuda_decode_batch_host(lora_phy_cuda_ctx* ctx, const uint16_t* h_symbols,
                                     size_t symbol_stride, const int32_t* h_symbol_counts,
                                     size_t symbols_per_packet, size_t n_packets,
                                     uint8_t* h_payload, size_t payload_stride, uint8_t* h_crc_ok);

/* ---- streaming front end: K2 -------------------------------------------------------------------- */
/* Sliding-window preamble / sync-word detector with CFO / STO tracking on CONTINUOUS IQ, n_streams
 * independent streams of one spreading factor: the upstream receiver the reference vendors
 * (external/LoRa-SDR/LoRaDemod.cpp:145-327; frame format of LoRaMod.cpp:141-215: 10 up-chirps, two sync
 * chirps at nibble*8, 2 1/4 down-chirps, data).  Per stream up to max_frames frames are reported, in
 * order of completion:
 *   d_frame_start [n_streams][max_frames]  sample index of the window that matched sync word0 (after STO tracking)
 *   d_frame_err   [n_streams][max_frames]  averaged down-chirp bin error (_freqError: the integer CFO estimate)
 *   d_frame_nsym  [n_streams][max_frames]  symbols in the frame (<= mtu; ends at mtu or when snr < thresh_db)
 *   d_symbols     [n_streams][max_frames][mtu] int16
 *   d_n_frames    [n_streams] frames completed (may exceed max_frames: the excess is counted, not stored)
 *   d_end_pos     [n_streams] (may be NULL) first sample the receiver did not consume
 * d_sample_counts NULL or [n_streams] ragged lengths.  thresh_db is upstream's squelch (default -30).     */
int lora_phy_cuda_detect_stream(lora_phy_cuda_ctx* ctx, uint32_t sf, uint32_t sync_word, float thresh_db,
                                uint32_t mtu, const void* d_iq, size_t stream_stride,
                                const uint32_t* d_sample_counts, size_t n_samples, size_t n_streams,
                                size_t max_frames, uint32_t* d_frame_start, int32_t* d_frame_err,
                                int32_t* d_frame_nsym, int16_t* d_symbols, uint32_t* d_n_frames,
                                uint32_t* d_end_pos);

/* Mixed batch (BASELINE.json configs[3]): stream s runs at spreading factor h_sf[s] (HOST array, 7..12); rows of
 * one SF are walked by one launch each, results land in the rows of the same output arrays.  Bandwidth does not
 * enter the receiver's arithmetic at one sample per chip (it only names the sample rate); oversampled captures
 * are decimated by the caller, as upstream's LoRaDemod expects.  Everything else as lora_phy_cuda_detect_stream. */
int lora_phy_cuda_detect_streams_mixed(lora_phy_cuda_ctx* ctx, const uint8_t* h_sf, uint32_t sync_word, float thresh_db,
                                       uint32_t mtu, const void* d_iq, size_t stream_stride,
                                       const uint32_t* d_sample_counts, size_t n_samples, size_t n_streams,
                                       size_t max_frames, uint32_t* d_frame_start, int32_t* d_frame_err,
                                       int32_t* d_frame_nsym, int16_t* d_symbols, uint32_t* d_n_frames,
                                       uint32_t* d_end_pos);

/* cfg 4 of BASELINE.json: continuous streams that differ in spreading factor, bandwidth and oversampling, optionally
 * pre-compensated.  Per stream: h_sf[s] in 7..12; h_bw_scale[s] in {1, 2, 4} (BW125 / 250 / 500 with the REFERENCE's
 * semantics -- bw_scale multiplies chirp slope and symbol offset, ChirpGenerator.hpp:28-30, LoRaMod.cpp:24-25,36-37 --
 * so dechirp tables, sync-word targets and the bin->time conversions of the state machine scale with it and the
 * symbols come out as (bw_scale*s) mod N); h_osr[s] in 1..8 samples per chip (windows of N*osr samples, phase 0 of
 * every chip as LoRaDemod.cpp:94-99 takes it, slides in chips).  NULL h_bw_scale / h_osr = 1.  One launch per distinct
 * (sf, bw_scale, osr); the launches run side by side on streams the context owns (they start after the work already
 * queued on the context's stream and the context's stream waits for all of them: to the caller the call is ordered
 * like any other).  d_cfo / d_time_offset: NULL, or [n_streams] offsets applied to each stream exactly as
 * compensate_offsets (phy.cpp:150-180: rotate by exp(-j*2*pi*cfo*n/(N*osr)), shift by round(time_offset) with zero
 * fill) INSIDE the window load -- no compensated copy of the stream is made.  frame_start / end_pos are in samples.
 * With bw_scale = osr = 1 and no offsets this is lora_phy_cuda_detect_streams_mixed.                                  */
int lora_phy_cuda_detect_streams_ex(lora_phy_cuda_ctx* ctx, const uint8_t* h_sf, const uint8_t* h_bw_scale,
                                    const uint8_t* h_osr, uint32_t sync_word, float thresh_db, uint32_t mtu,
                                    const void* d_iq, size_t stream_stride, const uint32_t* d_sample_counts,
                                    size_t n_samples, size_t n_streams, size_t max_frames,
                                    const float* d_cfo, const float* d_time_offset,
                                    uint32_t* d_frame_start, int32_t* d_frame_err, int32_t* d_frame_nsym,
                                    int16_t* d_symbols, uint32_t* d_n_frames, uint32_t* d_end_pos);

/* ---- bit pipeline: K4 (full upstream coding chain) -------------------------------------------- */
/* The chain the reference vendors under external/LoRa-SDR and its README's LoRaParams{cr, explicit_header,
 * crc_enabled} refer to (LoRaEncoder.cpp:161-233, LoRaDecoder.cpp:196-397): CRC-16, 3-byte explicit
 * header, Hamming / parity FEC per coding rate, whitening, diagonal interleaver, gray mapping.           */
typedef struct lora_phy_cuda_chain_cfg {
    uint32_t sf;              /* 7..12 */
    uint32_t ppm;             /* bits per symbol actually used; 0 = sf (LoRaEncoder.cpp:165) */
    uint32_t cr;              /* redundancy bits 0..4 = coding rate 4/4, 4/5, 4/6, 4/7, 4/8 */
    uint32_t explicit_header; /* LoRaEncoder.cpp:182 / LoRaDecoder.cpp:284 */
    uint32_t crc;             /* encoder: append the CRC-16 and flag it in the header */
    uint32_t whitening;       /* encoder: table whitening on (upstream default) */
    uint32_t error_check;     /* decoder: drop packets with any non-zero FEC syndrome (_errorCheck) */
    uint32_t crc_check;       /* decoder: drop packets whose CRC does not match (_crcc) */
    uint32_t out_header;      /* decoder: keep the 3 header bytes and the CRC in the output (_hdr) */
    uint32_t data_length;     /* decoder, implicit header: payload bytes (_dataLength) */
} lora_phy_cuda_chain_cfg;

/* Symbols one packet of payload_len bytes encodes to (LoRaEncoder.cpp:171-172), or -EINVAL. */
long lora_phy_cuda_chain_symbols(const lora_phy_cuda_chain_cfg* cc, size_t payload_len);
/* d_payload [n_packets][payload_stride] uint8 -> d_symbols [n_packets][symbol_stride] uint16; returns the
 * symbols per packet, -ERANGE if symbol_stride is smaller, -EINVAL on a bad configuration.          */
long lora_phy_cuda_encode_chain_batch(lora_phy_cuda_ctx* ctx, const lora_phy_cuda_chain_cfg* cc,
                                      const uint8_t* d_payload, size_t payload_stride, size_t payload_len,
                                      size_t n_packets, uint16_t* d_symbols, size_t symbol_stride);
/* d_symbols [n_packets][symbol_stride] (n_symbols used per packet, or d_symbol_counts[p] when given, e.g.
 * the status array of demodulate_batch) -> d_payload [n_packets][payload_stride].
 * d_len[p] = bytes written, -1 = dropped by the decoder (LoRaDecoder.cpp drop()), -ERANGE = payload_stride
 * too small.  d_flags[p] (may be NULL): bit0 a codeword had a non-zero syndrome, bit1 uncorrectable
 * H(8,4) word, bit2 the packet carries a CRC, bit3 the CRC matched.  Returns 0 or a negative errno.   */
int lora_phy_cuda_decode_chain_batch(lora_phy_cuda_ctx* ctx, const lora_phy_cuda_chain_cfg* cc,
                                     const uint16_t* d_symbols, size_t symbol_stride,
                                     const int32_t* d_symbol_counts, size_t n_symbols, size_t n_packets,
                                     uint8_t* d_payload, size_t payload_stride, int32_t* d_len, uint8_t* d_flags);

/* ---- whole-modem pipelines (the README's LoRaPHY::modulate / LoRaPHY::demodulate, batched) ------ */
/* tx: payload bytes -> coding chain (cc) -> lora_modulate (cfg: sf, bw, osr, amplitude, sync_word), i.e.
 * encode_chain_batch + modulate_batch without leaving the device.  Returns samples per packet
 * (= (lora_phy_cuda_chain_symbols(cc, payload_len) + 2) * N * osr) or a negative errno.
 * rx: raw IQ -> caller-side dechirp + lora_demodulate (the receive flow of tests/e2e_chain_test.cpp:83-101;
 * the reference's high-level demodulate() does not round-trip its own modulator) -> decode
 * chain.  The device variant dechirps d_iq IN PLACE.  d_len / d_flags as in decode_chain_batch; d_flags,
 * d_sync, d_cfo, d_time_offset may be NULL.                                                              */
long lora_phy_cuda_tx_batch(lora_phy_cuda_ctx* ctx, const lora_phy_cuda_cfg* cfg, const lora_phy_cuda_chain_cfg* cc,
                            const uint8_t* d_payload, size_t payload_stride, size_t payload_len, size_t n_packets,
                            void* d_iq, size_t packet_stride);
int lora_phy_cuda_rx_batch(lora_phy_cuda_ctx* ctx, const lora_phy_cuda_cfg* cfg, const lora_phy_cuda_chain_cfg* cc,
                           void* d_iq, size_t packet_stride, size_t samples_per_packet, size_t n_packets,
                           uint8_t* d_payload, size_t payload_stride, int32_t* d_len, uint8_t* d_flags,
                           uint8_t* d_sync, float* d_cfo, float* d_time_offset);
long lora_phy_cuda_tx_batch_host(lora_phy_cuda_ctx* ctx, const lora_phy_cuda_cfg* cfg, const lora_phy_cuda_chain_cfg* cc,
                                 const uint8_t* h_payload, size_t payload_stride, size_t payload_len, size_t n_packets,
                                 void* h_iq, size_t packet_stride);
int lora_phy_cuda_rx_batch_host(lora_phy_cuda_ctx* ctx, const lora_phy_cuda_cfg* cfg, const lora_phy_cuda_chain_cfg* cc,
                                const void* h_iq, size_t packet_stride, size_t samples_per_packet, size_t n_packets,
                                uint8_t* h_payload, size_t payload_stride, int32_t* h_len, uint8_t* h_flags,
                                uint8_t* h_sync, float* h_cfo, float* h_time_offset);

/* ---- AWGN Monte-Carlo driver (BASELINE.json configs[4]) ---------------------------------------- */
/* For packets first_packet .. first_packet + n_packets - 1 of sweep cell cell_id, entirely on the device:
 * draw payload_len payload bytes (Philox4x32-10 keyed by seed; counter = byte group, packet, cell) ->
 * upstream coding chain (cc) -> lora_modulate (LoRaMod.cpp:8-43; SF cc->sf, BW125, osr 1, amplitude 1,
 * sync 0x12) -> add complex Gaussian noise of standard deviation noise_std per I/Q component (Box-Muller
 * on the same Philox stream; SNR = -10 log10(2 noise_std^2)) and dechirp -> lora_demodulate
 * (LoRaDemod.cpp:50-197) -> decode chain -> compare with what was sent.  The reference ships no AWGN
 * code; the channel is defined in csrc/montecarlo.cu and restated in oracle/lora_mc_oracle.c so that
 * any packet can be regenerated and replayed on the host.  Packet ranges are independent: shard a
 * sweep over GPUs by giving each its own [first_packet, n_packets).
 *   d_counts[8] uint64, ACCUMULATED (zero it first): packets, packet errors, packets the decoder dropped
 *               or mis-sized, bit errors, bits compared, symbol errors, symbols compared, sync-word errors
 *   d_packet_ok [n_packets] (may be NULL); d_rx_symbols / d_rx_payload / d_rx_len (may be NULL):
 *               demodulated symbols, decoded bytes (payload_len per packet) and decoder length per packet
 * Work is chunked so that the IQ scratch stays below 4 GiB (LORA_PHY_MC_SCRATCH_MB).                */
int lora_phy_cuda_awgn_roundtrip(lora_phy_cuda_ctx* ctx, const lora_phy_cuda_chain_cfg* cc, float noise_std,
                                 uint64_t seed, uint32_t cell_id, uint64_t first_packet, size_t n_packets,
                                 size_t payload_len, uint64_t* d_counts, uint8_t* d_packet_ok,
                                 uint16_t* d_rx_symbols, size_t rx_symbol_stride, uint8_t* d_rx_payload,
                                 size_t rx_payload_stride, int32_t* d_rx_len);

/* ---- tables (parity tests read these back) -------------------------------------------------- */
/* which: 0 = kissfft twiddles (kissfft.hh:24-29), 1 = Hann window (LoRaDemod.cpp:18-24),
 * 2 = downchirp for cfg bw (phy.cpp:205-208).  Copies N entries to host memory.                */
int lora_phy_cuda_copy_table(lora_phy_cuda_ctx* ctx, uint32_t sf, uint32_t bw_hz, int which,
                             void* h_out);

#ifdef __cplusplus
}
#endif
#endif /* LORA_PHY_CUDA_H */
