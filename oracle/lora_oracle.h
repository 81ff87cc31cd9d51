/* oracle/lora_oracle.h -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Plain-C CPU restatement of the reference's LoRa PHY hot path.  Every function
 * cites the reference file:line whose arithmetic it restates.  Parity status:
 * PINNED -- tests/test_oracle.py checks it (a) bit-for-bit against
 * oracle/_ref/libref_lora_phy.so (the reference's own sources compiled unchanged)
 * on seeded random inputs, and (b) against the reference's known-answer vectors
 * (roundtrip_test.cpp:30, whitening_test.cpp:30-49, equal_power_iq.b64,
 * sync_word_iq.b64, no_alloc_test.cpp:35).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may load this library.  The product (liblora_phy_cuda.so) never
 * links or calls it.
 */
#ifndef LORA_ORACLE_H
#define LORA_ORACLE_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct { float re, im; } ocplx;

#define ORACLE_MAX_N 4096

/* kissfft.hh:24-29 */
void oracle_twiddles(unsigned sf, ocplx* tw);
/* kissfft.hh:100-185 (forward, radix 4 then 2) */
void oracle_fft(unsigned sf, const ocplx* in, ocplx* out);
/* LoRaDetector.hpp:39-74 */
size_t oracle_detect(unsigned sf, const ocplx* in, ocplx* spectrum,
                     float* power, float* power_avg, float* findex);
/* LoRaDemod.cpp:11-33 window part */
void oracle_window(unsigned sf, int hann, float* w);
/* LoRaDemod.cpp:50-197.  with_scratch==0 models a workspace without scratch (-ERANGE path). */
long oracle_lora_demodulate(unsigned sf, int window, int with_scratch,
                            const ocplx* samples, size_t sample_count, unsigned osr,
                            uint16_t* out_symbols, uint8_t* out_sync,
                            float* out_cfo, float* out_time_offset);
/* ChirpGenerator.hpp:24-51 */
int oracle_gen_chirp(ocplx* out, int N, int osr, int NN, float f0, int down, float ampl,
                     float* phase, float bw_scale);
/* LoRaMod.cpp:8-43 */
size_t oracle_lora_modulate(const uint16_t* symbols, size_t n, ocplx* out, unsigned sf,
                            unsigned osr, unsigned bw_hz, float amplitude, uint8_t sync);
/* LoRaEncoder.cpp:6-18, LoRaDecoder.cpp:7-21 */
size_t oracle_lora_encode(const uint8_t* bytes, size_t n, uint16_t* out);
long oracle_lora_decode(const uint16_t* symbols, size_t n, uint8_t* out);

/* phy.cpp:81-148 (high-level estimator: no index tie-break, all symbols given) */
void oracle_estimate_offsets(unsigned sf, unsigned osr, int window, const ocplx* samples,
                             size_t sample_count, float* cfo, float* time_offset);
/* phy.cpp:150-180 */
void oracle_compensate_offsets(unsigned sf, unsigned osr, float cfo, float time_offset,
                               ocplx* samples, size_t sample_count);
/* phy.cpp:182-243; returns count or -errno; *sync_out = recovered sync word */
long oracle_demodulate(unsigned sf, unsigned bw_hz, unsigned osr, int window,
                       const ocplx* iq, size_t sample_count, uint16_t* symbols,
                       size_t symbol_cap, uint8_t* sync_out, float* cfo, float* time_offset);
/* phy.cpp:245-261 */
long oracle_decode(const uint16_t* symbols, size_t n, uint8_t* payload, size_t cap, int* crc_ok);

/* LoRaCodes.hpp helpers */
uint8_t  oracle_header_checksum(const uint8_t* h);                 /* :43-67  */
uint16_t oracle_data_checksum(const uint8_t* d, int n);            /* :69-105 */
void oracle_whiten_sx1232(uint8_t* b, uint16_t n);                 /* :111-137 */
void oracle_whiten_table(uint8_t* b, uint16_t n, int ofs, int rdd);/* :147-167 */
void oracle_whiten_lfsr(uint8_t* b, uint16_t n, int ofs, size_t rdd);/* :176-189 */
uint16_t oracle_binary_to_gray16(uint16_t v);                      /* :201-207 */
uint16_t oracle_gray_to_binary16(uint16_t v);                      /* :212-222 */
uint8_t oracle_enc_h84(uint8_t v);                                 /* :229-242 */
uint8_t oracle_dec_h84(uint8_t v, int* err, int* bad);             /* :250-281 */
uint8_t oracle_enc_h74(uint8_t v);                                 /* :287-299 */
uint8_t oracle_dec_h74(uint8_t v, int* err);                       /* :306-334 */
uint8_t oracle_enc_p54(uint8_t v);                                 /* :347-351 */
uint8_t oracle_chk_p54(uint8_t v, int* err);                       /* :340-345 */
uint8_t oracle_enc_p64(uint8_t v);                                 /* :367-371 */
uint8_t oracle_chk_p64(uint8_t v, int* err);                       /* :357-365 */
void oracle_interleave(const uint8_t* cw, size_t ncw, uint16_t* sym, size_t ppm, size_t rdd);   /* :376-393 */
void oracle_deinterleave(const uint16_t* sym, size_t nsym, uint8_t* cw, size_t ppm, size_t rdd);/* :396-412 */

/* Upstream coding chain (external/LoRa-SDR/LoRaEncoder.cpp:161-233, LoRaDecoder.cpp:196-397), restated in
 * lora_chain_oracle.c.  ppm == 0 means ppm = sf. */
typedef struct {
    unsigned sf, ppm, rdd;          /* rdd 0..4 = coding rate 4/4 .. 4/8 */
    int explicit_hdr, crc, whitening;
    int error_check, crc_check, out_hdr;   /* decoder: _errorCheck, _crcc, _hdr */
    unsigned data_length;           /* decoder, implicit-header mode */
} oracle_chain_cfg;
size_t oracle_chain_num_symbols(const oracle_chain_cfg* c, size_t payload_len);
long oracle_encode_chain(const oracle_chain_cfg* c, const uint8_t* payload, size_t payload_len, uint16_t* symbols);
long oracle_decode_chain(const oracle_chain_cfg* c, const uint16_t* symbols, size_t n_symbols, uint8_t* out,
                         size_t cap, int* flags);

/* Upstream streaming receiver (external/LoRa-SDR/LoRaDemod.cpp:68-117,145-327), restated in
 * lora_stream_oracle.c: tables, per-window dechirp + fine tune + detect, frame-sync state machine. */
void oracle_stream_tables(unsigned sf, ocplx* up, ocplx* down, ocplx* fine);
long oracle_stream_demod(unsigned sf, unsigned sync, float thresh, unsigned mtu, const ocplx* iq, size_t n,
                         size_t max_frames, uint32_t* frame_start, int32_t* frame_err, int32_t* frame_nsym,
                         int16_t* symbols, uint32_t* end_pos);
void oracle_stream_tables_bw(unsigned sf, unsigned bw_k, ocplx* up, ocplx* down, ocplx* fine);
long oracle_stream_demod_ex(unsigned sf, unsigned osr, unsigned bw_k, unsigned sync, float thresh, unsigned mtu, const ocplx* iq,
                            size_t n, size_t max_frames, uint32_t* frame_start, int32_t* frame_err, int32_t* frame_nsym,
                            int16_t* symbols, uint32_t* end_pos);
size_t oracle_frame_modulate_bw(const uint16_t* symbols, size_t n, ocplx* out, unsigned sf, unsigned ovs, unsigned sync,
                                float ampl, unsigned padding, unsigned bw_k);
/* Upstream transmitter frame format (external/LoRa-SDR/LoRaMod.cpp:109-238, ChirpGenerator.hpp:22-47). */
size_t oracle_frame_samples(unsigned sf, unsigned ovs, size_t n_symbols, unsigned padding);
size_t oracle_frame_modulate(const uint16_t* symbols, size_t n, ocplx* out, unsigned sf, unsigned ovs, unsigned sync,
                             float ampl, unsigned padding);

/* AWGN Monte-Carlo spot check (lora_mc_oracle.c): Philox4x32-10, payload / noise draws, one packet round trip. */
void oracle_philox4x32(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]);
void oracle_mc_payload(uint64_t seed, uint32_t cell, uint64_t packet, size_t len, uint8_t* out);
void oracle_mc_noise_sample(uint64_t seed, uint32_t cell, uint64_t packet, uint32_t i, float std, float* re, float* im);
long oracle_awgn_roundtrip(const oracle_chain_cfg* cc, float noise_std, uint64_t seed, uint32_t cell, uint64_t packet,
                           size_t payload_len, uint8_t* tx_payload, uint16_t* tx_sym, uint16_t* rx_sym,
                           uint8_t* rx_payload, size_t rx_cap, int* flags);

/* Timed CPU arm when oracle/_ref is unavailable: one packet per thread. Returns seconds. */
double oracle_demod_batch_mt(unsigned sf, int window, unsigned osr, const ocplx* iq,
                             size_t n_packets, size_t samples_per_packet,
                             uint16_t* out_symbols, size_t symbol_stride, uint8_t* out_sync,
                             int n_threads);

#ifdef __cplusplus
}
#endif
#endif
