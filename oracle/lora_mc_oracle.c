/* oracle/lora_mc_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU side of the AWGN Monte-Carlo spot check (BASELINE.json configs[4], SURVEY.md 8d cfg 5).  The reference
 * ships no AWGN code (its awgn vectors and tests/awgn_sweep.py are absent), so the channel is defined here
 * and in csrc/montecarlo.cu by the same published pieces:
 *   - Philox4x32-10 counter RNG (Salmon et al., SC'11; Random123 known answers checked in tests/test_mc_oracle.py)
 *     counter = (index, packet_lo, packet_hi ^ cell << 8 ..., domain), key = 64-bit seed
 *   - Box-Muller on two 23-bit uniforms, written with the libm calls the rest of the path already replays
 *     bit for bit (log10f, sqrtf, cosf/sinf), so host and device produce identical float noise
 * and the signal path around it is the reference's own: upstream coding chain -> lora_modulate
 * (LoRaMod.cpp:8-43) -> + noise -> caller-side dechirp (tests/e2e_chain_test.cpp:83-92) -> lora_demodulate
 * (LoRaDemod.cpp:50-197) -> upstream decode chain.  PARITY STATUS: the signal path is pinned as in
 * lora_oracle.h; the channel has no reference counterpart (CUDA == this file is the contract).
 */
#define _USE_MATH_DEFINES
#define _DEFAULT_SOURCE
#include "lora_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

void oracle_philox4x32(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4])
{
    uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3], k0 = key[0], k1 = key[1];
    for (int r = 0; r < 10; ++r) {
        const uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
        const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t)p1;
        const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

enum { MC_DOMAIN_NOISE = 0x4e4f4953, MC_DOMAIN_PAYLOAD = 0x50415931 };

static void mc_block(uint64_t seed, uint32_t cell, uint64_t packet, uint32_t index, uint32_t domain, uint32_t out[4])
{
    const uint32_t ctr[4] = { index, (uint32_t)packet, (uint32_t)(packet >> 32) ^ (cell << 8), domain };
    const uint32_t key[2] = { (uint32_t)seed, (uint32_t)(seed >> 32) };
    oracle_philox4x32(ctr, key, out);
}

/* payload byte b of a packet: byte (b & 15) of Philox block b >> 4, little-endian words */
void oracle_mc_payload(uint64_t seed, uint32_t cell, uint64_t packet, size_t len, uint8_t* out)
{
    uint32_t x[4] = { 0, 0, 0, 0 };
    for (size_t b = 0; b < len; ++b) {
        if ((b & 15) == 0) mc_block(seed, cell, packet, (uint32_t)(b >> 4), MC_DOMAIN_PAYLOAD, x);
        out[b] = (uint8_t)(x[(b >> 2) & 3] >> (8 * (b & 3)));
    }
}

/* complex noise sample i of a packet: u1, u2 = ((x >> 9) + 0.5) * 2^-23 in (0, 1);
 * r = sqrtf(-2 ln10 * log10f(u1)); theta = 2 pi u2; n = std * r * (cosf, sinf)(theta) */
void oracle_mc_noise_sample(uint64_t seed, uint32_t cell, uint64_t packet, uint32_t i, float std, float* re, float* im)
{
    uint32_t x[4];
    mc_block(seed, cell, packet, i, MC_DOMAIN_NOISE, x);
    const float u1 = ((float)(x[0] >> 9) + 0.5f) * 0x1p-23f;
    const float u2 = ((float)(x[1] >> 9) + 0.5f) * 0x1p-23f;
    const float r = sqrtf(-4.605170185988092f * log10f(u1));
    const float th = 6.283185307179586f * u2;
    *re = std * (r * cosf(th));
    *im = std * (r * sinf(th));
}

/* One packet of the sweep.  Outputs (each may be NULL): the payload drawn, the symbols sent, the symbols
 * demodulated, the decoded bytes and decoder flags.  Returns the decoder's length (-1 = dropped). */
long oracle_awgn_roundtrip(const oracle_chain_cfg* cc, float noise_std, uint64_t seed, uint32_t cell, uint64_t packet,
                           size_t payload_len, uint8_t* tx_payload, uint16_t* tx_sym, uint16_t* rx_sym,
                           uint8_t* rx_payload, size_t rx_cap, int* flags)
{
    const unsigned sf = cc->sf;
    const size_t N = (size_t)1 << sf;
    const size_t nsym = oracle_chain_num_symbols(cc, payload_len);
    uint8_t* pay = (uint8_t*)malloc(payload_len ? payload_len : 1);
    uint16_t* sym = (uint16_t*)malloc(sizeof(uint16_t) * (nsym + 2));
    uint16_t* got = (uint16_t*)malloc(sizeof(uint16_t) * (nsym + 2));
    ocplx* iq = (ocplx*)malloc(sizeof(ocplx) * (nsym + 2) * N);
    ocplx* down = (ocplx*)malloc(sizeof(ocplx) * N);
    oracle_mc_payload(seed, cell, packet, payload_len, pay);
    oracle_encode_chain(cc, pay, payload_len, sym);
    const size_t total = oracle_lora_modulate(sym, nsym, iq, sf, 1, 125000, 1.0f, 0x12);
    float ph = 0.0f;
    oracle_gen_chirp(down, (int)N, 1, (int)N, 0.0f, 1, 1.0f, &ph, 1.0f);
    for (size_t i = 0; i < total; ++i) {
        float nr, ni;
        oracle_mc_noise_sample(seed, cell, packet, (uint32_t)i, noise_std, &nr, &ni);
        const float yr = iq[i].re + nr, yi = iq[i].im + ni;
        const ocplx d = down[i % N];
        iq[i].re = yr * d.re - yi * d.im;
        iq[i].im = yr * d.im + yi * d.re;
    }
    uint8_t sync = 0;
    float cfo = 0, to = 0;
    const long n = oracle_lora_demodulate(sf, 0, 1, iq, total, 1, got, &sync, &cfo, &to);
    int fl = 0;
    long r = -1;
    if (n >= 0) r = oracle_decode_chain(cc, got, (size_t)n, rx_payload, rx_cap, &fl);
    if (tx_payload) memcpy(tx_payload, pay, payload_len);
    if (tx_sym) memcpy(tx_sym, sym, sizeof(uint16_t) * nsym);
    if (rx_sym && n > 0) memcpy(rx_sym, got, sizeof(uint16_t) * (size_t)n);
    if (flags) *flags = fl;
    free(pay); free(sym); free(got); free(iq); free(down);
    return r;
}
