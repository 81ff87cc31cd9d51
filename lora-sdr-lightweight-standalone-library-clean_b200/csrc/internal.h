// internal.h -- launcher prototypes shared by the .cu translation units (not part of the ABI).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include "../../include/lora_phy_cuda.h"
#include "fft_exact.cuh"
#include "mc_noise.cuh"

namespace lk {

using fx::cf;
struct DemodArgs;

struct DetectArgs {
    const cf* windows; uint32_t n_windows;
    const cf* tw; const cf* ct; float power_scale;
    cf* spectrum; uint32_t* index; float* power; float* power_avg; float* findex;
};

// float(20*log10(double(N))) for N = 1<<sf (LoRaDetector.hpp:29), bit patterns checked against
// the oracle by tests/test_tables.py
__host__ __device__ inline float power_scale_of(int sf)
{
    const uint32_t bits[13] = { 0x0u, 0x40c0a8c1u, 0x4140a8c1u, 0x41907e91u, 0x41c0a8c1u, 0x41f0d2f1u,
                                0x42107e91u, 0x422893a9u, 0x4240a8c1u, 0x4258bdd9u, 0x4270d2f1u,
                                0x42847405u, 0x42907e91u };
    union { uint32_t u; float f; } v; v.u = bits[sf]; return v.f;
}

cudaError_t launch_demod(int sf, const DemodArgs& a, int sm_count, cudaStream_t stream, bool prepass,
                         bool symbols, unsigned* launches);
cudaError_t launch_detect(int sf, const DetectArgs& a, int sm_count, cudaStream_t stream);
// demod_cert.cu: pipeline per packet, throughput FFT with a per-window proof that its argmax is the reference's, exact
// replay of the windows that cannot be proven (SF 7..12, osr == 1); F_EXACT = the same kernel, every window replayed
bool cert_supported(int sf, const DemodArgs& a);
cudaError_t launch_demod_cert(int sf, const DemodArgs& a, int sm_count, cudaStream_t stream);

// tables.cu
cudaError_t launch_build_tables(int sf, cf* tw, cf* ct, int ct_size, float* hann, cudaStream_t stream);
cudaError_t launch_build_chirp(int sf, int osr, float bw_scale, cf* out /* N*osr */, cudaStream_t stream);
int compact_size_of(int sf);

// demod_lean.cu: streaming pruned first pass (SF10-12); see api.cu for the three-launch route
bool lean_supported(int sf, const DemodArgs& a);
cudaError_t launch_demod_lean(int sf, const DemodArgs& a, int sm_count, cudaStream_t stream);

// modcodes.cu
struct ModArgs {
    const uint16_t* symbols; size_t symbol_stride; uint32_t symbols_per_packet; uint32_t n_packets;
    cf* iq; size_t packet_stride; int sf; uint32_t osr; float bw_scale; float amplitude; uint32_t sync;
    float* phase_scratch;   // [n_packets][chirps] start phase of every chirp
    McChannel channel;      // Monte-Carlo driver: AWGN + dechirp fused into the store (channel.on), else untouched
    int frame;              // 0: shipped framing (2 sync chirps + data); 1: upstream frame format (LoRaMod.cpp:109-238)
    uint32_t padding;       // frame format: trailing zero symbols
};
// chirps per packet: shipped framing n+2; upstream frame 10 up + 2 sync + 2 down + quarter down + n data
__host__ __device__ inline uint32_t mod_chirps(const ModArgs& a) { return a.symbols_per_packet + (a.frame ? 15u : 2u); }
cudaError_t launch_modulate(const ModArgs& a, int sm_count, cudaStream_t stream, unsigned* launches);
// allocation-free variant's scratch: a static device array of this many chirp phases (modcodes.cu); the device address
// of it on the CURRENT device (cudaGetSymbolAddress, cached per device)
constexpr size_t MOD_STATIC_MAX_CHIRPS = LORA_PHY_CUDA_STATIC_MAX_CHIRPS;
cudaError_t mod_static_scratch(float** out);
cudaError_t launch_dechirp(cf* iq, size_t packet_stride, uint32_t samples_per_packet, uint32_t n_packets,
                           const cf* table, uint32_t table_len, cudaStream_t stream);
cudaError_t launch_encode(const uint8_t* payload, size_t payload_stride, uint32_t payload_len, uint32_t n_packets,
                          uint16_t* symbols, size_t symbol_stride, cudaStream_t stream);
cudaError_t launch_decode(const uint16_t* symbols, size_t symbol_stride, const int32_t* counts,
                          uint32_t symbols_per_packet, uint32_t n_packets, uint8_t* payload,
                          size_t payload_stride, uint8_t* crc_ok, cudaStream_t stream);
cudaError_t launch_compensate(const cf* in, cf* out, size_t packet_stride, uint32_t samples_per_packet,
                              uint32_t n_packets, int sf, uint32_t osr, const float* cfo,
                              const float* time_offset, cudaStream_t stream);

// montecarlo.cu: AWGN Monte-Carlo driver kernels (payload draw, channel + dechirp, scoring)
struct McArgs {
    unsigned long long seed, first_packet; uint32_t cell; uint32_t n_packets;
    uint8_t* payload; size_t payload_stride; uint32_t payload_len;
    cf* iq; size_t packet_stride; uint32_t spp; const cf* chirp; uint32_t chirp_len; float noise_std;
    const uint16_t* tx_sym; const uint16_t* rx_sym; size_t sym_stride; uint32_t n_sym;
    const uint8_t* rx_payload; size_t rx_stride; const int32_t* rx_len; const int32_t* status; const uint8_t* sync;
    uint32_t sync_word; unsigned long long* counts; uint8_t* packet_ok;
};
cudaError_t launch_mc_payload(const McArgs& a, int sm_count, cudaStream_t stream);
cudaError_t launch_awgn_dechirp(const McArgs& a, int sm_count, cudaStream_t stream);
cudaError_t launch_mc_score(const McArgs& a, cudaStream_t stream);

// stream_sync.cu: sliding-window preamble / sync-word detector on continuous IQ, one thread group per stream
struct StreamSyncArgs {
    const cf* iq; size_t stream_stride; const uint32_t* counts; uint32_t n_samples; uint32_t n_streams;
    const cf* up; const cf* down; const cf* fine;          // LoRaDemod.cpp:96-117 tables (N, N, N*128 entries)
    const cf* tw; const cf* ct; float power_scale;
    uint32_t sync; float thresh; uint32_t mtu; uint32_t max_frames;
    uint32_t osr, bw_k;                                     // samples per chip (windows are N*osr samples); bandwidth scale 1 / 2 / 4
    uint32_t sync0, sync1;                                  // (value + 4) / 8 targets of the two sync-word chirps: nibble * bw_k mod N/8
    const float* comp_cfo; const float* comp_to;            // NULL, or [n_streams] compensate_offsets fused into the window load
    const uint32_t* ids;                                    // NULL, or [n_streams] rows this launch walks (mixed-SF batches)
    int force_serial;                                       // tests: take the literal serial loops (LORA_PHY_STREAM_SERIAL=1)
    uint32_t* frame_start; int32_t* frame_err; int32_t* frame_nsym; int16_t* symbols;   // [n_streams][max_frames]([mtu])
    uint32_t* n_frames; uint32_t* end_pos;                 // [n_streams]
};
cudaError_t launch_stream_sync(int sf, const StreamSyncArgs& a, cudaStream_t stream);

// chain.cu: upstream coding chain (CRC / header / FEC / whitening / interleaver / gray), one warp per packet
struct ChainArgs {
    unsigned sf, ppm, rdd;
    int explicit_hdr, crc, whitening, error_check, crc_check, out_hdr;
    unsigned data_length;
};
cudaError_t launch_chain_encode(const ChainArgs& a, const uint8_t* payload, size_t payload_stride, uint32_t len,
                                uint32_t n_packets, uint16_t* symbols, size_t symbol_stride, int sm_count,
                                cudaStream_t stream);
cudaError_t launch_chain_decode(const ChainArgs& a, const uint16_t* symbols, size_t symbol_stride,
                                const int32_t* counts, uint32_t n_symbols, uint32_t n_packets, const uint8_t* lfsr,
                                uint8_t* payload, size_t payload_stride, int32_t* out_len, uint8_t* out_flags,
                                int sm_count, cudaStream_t stream);
void chain_lfsr_table(uint8_t* out);
int chain_lfsr_table_size();
int chain_max_symbols();

} // namespace lk
