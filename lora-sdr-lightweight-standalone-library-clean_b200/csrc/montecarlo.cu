// montecarlo.cu -- AWGN Monte-Carlo driver kernels (BASELINE.json configs[4]; SURVEY.md 8d cfg 5).
// The reference ships no AWGN code; the channel is a counter RNG + Box-Muller defined identically here and
// in oracle/lora_mc_oracle.c so that any packet of a sweep can be regenerated and re-run on the host:
//   Philox4x32-10, counter = (index, packet_lo, packet_hi ^ cell << 8, domain), key = 64-bit seed
//   u1, u2 = ((x >> 9) + 0.5) * 2^-23;  r = sqrtf(-2 ln10 * log10f(u1));  n = std * r * (cosf, sinf)(2 pi u2)
// with log10f / cosf / sinf the glibc restatements of glibc_math.cuh (bit-identical to the host's libm).
//   mc_payload_kernel    payload bytes of every packet
//   awgn_dechirp_kernel  iq = (iq + noise) * downchirp in one pass over the modulated batch (the caller-side
//                        dechirp of tests/e2e_chain_test.cpp:83-92 fused with the channel)
//   mc_score_kernel      compares decoded bytes / demodulated symbols with what was sent, accumulates counters
#include "internal.h"
#include "glibc_math.cuh"
#include "mc_noise.cuh"

namespace lk {

namespace {

__device__ __forceinline__ uint4 mc_block(const McArgs& a, unsigned long long packet, uint32_t index, uint32_t domain)
{
    return mc_philox(a.seed, a.cell, packet, index, domain);
}

__global__ void __launch_bounds__(256) mc_payload_kernel(const McArgs a)
{
    const unsigned long long groups = (a.payload_len + 15) / 16, total = (unsigned long long)a.n_packets * groups;
    for (unsigned long long k = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; k < total;
         k += (unsigned long long)gridDim.x * blockDim.x) {
        const uint32_t p = (uint32_t)(k / groups), g = (uint32_t)(k % groups);
        const uint4 x = mc_block(a, a.first_packet + p, g, MC_DOMAIN_PAYLOAD);
        const uint32_t w[4] = { x.x, x.y, x.z, x.w };
        uint8_t* out = a.payload + (size_t)p * a.payload_stride + 16 * g;
        const uint32_t n = min(16u, a.payload_len - 16 * g);
        for (uint32_t b = 0; b < n; ++b) out[b] = (uint8_t)(w[b >> 2] >> (8 * (b & 3)));
    }
}

__global__ void __launch_bounds__(256) awgn_dechirp_kernel(const McArgs a)
{
    const unsigned long long total = (unsigned long long)a.n_packets * a.spp;
    const uint32_t mask = a.chirp_len - 1;                                  // N (osr = 1): a power of two
    for (unsigned long long k = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; k < total;
         k += (unsigned long long)gridDim.x * blockDim.x) {
        const uint32_t p = (uint32_t)(k / a.spp), i = (uint32_t)(k % a.spp);
        McChannel ch;
        ch.on = 1; ch.seed = a.seed; ch.first_packet = a.first_packet; ch.cell = a.cell; ch.noise_std = a.noise_std;
        ch.chirp = a.chirp; ch.chirp_mask = mask;
        cf* q = a.iq + (size_t)p * a.packet_stride + i;
        *q = mc_channel_sample(ch, a.first_packet + p, i, *q);
    }
}

__device__ __forceinline__ unsigned long long warp_sum(unsigned long long v)
{
#pragma unroll
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// counts: 0 packets, 1 packet errors, 2 packets the decoder dropped or mis-sized, 3 bit errors, 4 bits compared,
//         5 symbol errors, 6 symbols compared, 7 sync-word errors
__global__ void __launch_bounds__(128) mc_score_kernel(const McArgs a)
{
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long c[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };
    if (p < a.n_packets) {
        const uint8_t* tx = a.payload + (size_t)p * a.payload_stride;
        const uint8_t* rx = a.rx_payload + (size_t)p * a.rx_stride;
        const int len = a.rx_len[p];
        const uint32_t cmp = len < 0 ? 0u : min((uint32_t)len, a.payload_len);
        uint32_t biterr = 0;
        for (uint32_t b = 0; b < cmp; ++b) biterr += __popc((uint32_t)(tx[b] ^ rx[b]));
        const bool ok = len == (int)a.payload_len && biterr == 0;
        const uint16_t* ts = a.tx_sym + (size_t)p * a.sym_stride;
        const uint16_t* rs = a.rx_sym + (size_t)p * a.sym_stride;
        const int got = a.status[p];
        const uint32_t ns = got < 0 ? 0u : min((uint32_t)got, a.n_sym);
        uint32_t symerr = 0;
        for (uint32_t s = 0; s < ns; ++s) symerr += ts[s] != rs[s];
        c[0] = 1; c[1] = !ok; c[2] = len != (int)a.payload_len; c[3] = biterr; c[4] = 8ull * cmp;
        c[5] = symerr; c[6] = ns; c[7] = a.sync[p] != (uint8_t)a.sync_word;
        if (a.packet_ok) a.packet_ok[p] = ok;
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const unsigned long long s = warp_sum(c[k]);
        if ((threadIdx.x & 31) == 0 && s) atomicAdd(a.counts + k, s);
    }
}

unsigned grid_for(unsigned long long items, unsigned block, int sm_count)
{
    const unsigned long long need = (items + block - 1) / block, cap = (unsigned long long)sm_count * 16;
    return (unsigned)(need < cap ? (need ? need : 1) : cap);
}

} // namespace

cudaError_t launch_mc_payload(const McArgs& a, int sm_count, cudaStream_t stream)
{
    if (!a.n_packets || !a.payload_len) return cudaSuccess;
    mc_payload_kernel<<<grid_for((unsigned long long)a.n_packets * ((a.payload_len + 15) / 16), 256, sm_count), 256, 0, stream>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_awgn_dechirp(const McArgs& a, int sm_count, cudaStream_t stream)
{
    if (!a.n_packets || !a.spp) return cudaSuccess;
    awgn_dechirp_kernel<<<grid_for((unsigned long long)a.n_packets * a.spp, 256, sm_count), 256, 0, stream>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_mc_score(const McArgs& a, cudaStream_t stream)
{
    if (!a.n_packets) return cudaSuccess;
    mc_score_kernel<<<(a.n_packets + 127) / 128, 128, 0, stream>>>(a);
    return cudaGetLastError();
}

} // namespace lk
