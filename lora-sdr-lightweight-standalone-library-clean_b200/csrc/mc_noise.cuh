// mc_noise.cuh -- the Monte-Carlo channel's noise sample, shared by the fused synthesis kernel (modcodes.cu) and the
// stand-alone channel kernel (montecarlo.cu); defined identically in oracle/lora_mc_oracle.c:
//   Philox4x32-10, counter = (index, packet_lo, packet_hi ^ cell << 8, domain), key = 64-bit seed
//   u1, u2 = ((x >> 9) + 0.5) * 2^-23;  r = sqrtf(-2 ln10 * log10f(u1));  n = std * r * (cosf, sinf)(2 pi u2)
// with log10f / cosf / sinf the glibc restatements of glibc_math.cuh (bit-identical to the host's libm).
#pragma once
#include <stdint.h>
#include "fft_exact.cuh"
#include "glibc_math.cuh"

namespace lk {

constexpr uint32_t MC_DOMAIN_NOISE = 0x4e4f4953u, MC_DOMAIN_PAYLOAD = 0x50415931u;

struct McChannel {                      // AWGN + caller-side dechirp applied where the modulator stores a sample
    int on;
    unsigned long long seed, first_packet;
    uint32_t cell;
    float noise_std;
    const fx::cf* chirp;                // conjugate downchirp, chirp_len entries (osr == 1: N, a power of two)
    uint32_t chirp_mask;
};

__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint32_t k0, uint32_t k1)
{
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
        c = make_uint4(hi1 ^ c.y ^ k0, lo1, hi0 ^ c.w ^ k1, lo0);
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    return c;
}

__device__ __forceinline__ uint4 mc_philox(unsigned long long seed, uint32_t cell, unsigned long long packet, uint32_t index,
                                           uint32_t domain)
{
    return philox4x32_10(make_uint4(index, (uint32_t)packet, (uint32_t)(packet >> 32) ^ (cell << 8), domain),
                         (uint32_t)seed, (uint32_t)(seed >> 32));
}

// (v + noise_i) * downchirp[i & mask] for sample i of `packet`
__device__ __forceinline__ fx::cf mc_channel_sample(const McChannel& ch, unsigned long long packet, uint32_t i, fx::cf v)
{
    const uint4 x = mc_philox(ch.seed, ch.cell, packet, i, MC_DOMAIN_NOISE);
    const float u1 = fx::fmul(fx::fadd((float)(x.x >> 9), 0.5f), 0x1p-23f);
    const float u2 = fx::fmul(fx::fadd((float)(x.y >> 9), 0.5f), 0x1p-23f);
    const float r = __fsqrt_rn(fx::fmul(-4.605170185988092f, gm::log10f_g(u1)));
    float sn, cs;
    gm::sincosf_g(fx::fmul(6.283185307179586f, u2), &sn, &cs);
    const fx::cf y = fx::mk(fx::fadd(v.x, fx::fmul(ch.noise_std, fx::fmul(r, cs))),
                            fx::fadd(v.y, fx::fmul(ch.noise_std, fx::fmul(r, sn))));
    const float2 t = __ldg(reinterpret_cast<const float2*>(ch.chirp + (i & ch.chirp_mask)));
    return fx::cmul(y, fx::mk(t.x, t.y));
}

} // namespace lk
