// fft_prune.cuh -- the pruned transform of the certified demodulator: ONE residue class of bins instead of all N.
//
// A dechirped LoRa symbol is one tone.  After pass 0 of the decimation-in-frequency plan (a 16-point DFT over the
// samples i = tid + T*c a thread loads anyway) the tone's energy sits in ONE of the thread's 16 outputs, the same
// k0 = bin mod 16 on every thread, and the bins of that class are a T-point DFT over the threads:
//     X[k0 + 16 k'] = sum_tid  W_N^{tid k0} V_tid[k0]  W_T^{tid k'},      T = N / 16.
// So the walk computes pass 0, takes k0 from the window's first thread, and transforms ONE value per thread:
//     T <= 32   radix-2 decimation in frequency across the lanes with shfl.xor -- no shared memory, no barrier;
//     T  > 32   (R = T/32 warps per window) the values meet in a T-entry buffer, warp ka forms its output of the
//               R-point DFT across warps, c_l = W_T^{l ka} sum_w u_{l+32w} W_R^{w ka}, then the 32-lane transform:
//               lane l of warp ka ends with bin k' = ka + R * bitrev5(l).
// That is N/16 bins for 1/16 of the work of passes 1..  The certificate does not need the other bins: tier 1 accepts
// a window iff the top bin found holds more than 9/16 of N*E, and by Parseval no bin of the exact transform -- computed
// here or not -- can then come near it.  It also makes ANY choice of k0 safe: a wrong class
// cannot pass the test.  Windows that fail (noise, weak signal) take the full transform of fft_fast.cuh.
//
// Host build: the index maps below are replayed by tests/cpu/test_fft_fast_emulation.cpp.
#pragma once
#include "fft_fast.cuh"

namespace fp {

using ff::cf;

// Tier 1 of the certificate (demod_cert.cu, demod_lean.cu): the top bin found is accepted iff q1 > TIER1 * N * E.
// Parseval on the exact transform Z of the true inputs: every other bin has |Z_k|^2 <= S^2 - |Z_top|^2, S^2 = N E.
// With y = sqrt(q1) / S >= 0.75, our error K_o S and the reference's K_r S (K_o + K_r = K < 2^-7, the kernels check it):
// z = |Z_top| / S >= 0.742, the reference computes |X^_top| >= (z - K_r) S >= 0.734 S and every other
// |X^_k| <= (sqrt(1 - z^2) + K_r) S <= 0.679 S: a strict, unique maximum with a margin of 0.05 S.  (The first version
// of the test asked for 3/4 of the energy; 9/16 keeps windows down to ~1 dB per-sample SNR and carrier corrections
// up to 0.4 bin on the cheap path.)
constexpr float TIER1 = 0.5625f;

template <int SF>
struct Prune {
    static constexpr int N = 1 << SF, T = N / 16;
    static constexpr int W = T < 32 ? T : 32;                 // lanes of one lane transform
    static constexpr int R = T / W;                           // warps per window (1 for SF7-9)
    static constexpr int LW = W == 8 ? 3 : W == 16 ? 4 : 5;   // log2 W
    static constexpr int STAGES_TW = LW - 1;                  // stages with a non-trivial twiddle (the last has none)

    FF_M static constexpr int bitrev(int x)
    {
        int r = 0;
        for (int b = 0; b < LW; ++b) if ((x >> b) & 1) r |= 1 << (LW - 1 - b);
        return r;
    }
    // stage s = 0 .. LW-1 pairs lane l with l ^ h, h = W >> (s + 1); the upper lane of a pair keeps
    // (lower - upper) * W_{2h}^{l mod h}, the lower one the sum.  Exponent of W_W, 0 for lower lanes:
    FF_M static constexpr int half(int s) { return W >> (s + 1); }
    FF_M static constexpr int stage_exp(int s, int lane)
    {
        const int h = half(s);
        return (lane & h) ? (lane & (h - 1)) * (W / (2 * h)) : 0;
    }
    // bin held after the lane transform by lane l (index inside its window's W lanes) of warp ka, class k0
    FF_M static constexpr int bin(int k0, int ka, int l) { return k0 + 16 * (ka + R * bitrev(l)); }
    // exponent of W_T for the cross-warp stage's twiddle on lane l of warp ka, and of W_R for term w of its sum
    FF_M static constexpr int cross_exp(int ka, int l) { return (l * ka) & (T - 1); }
    FF_M static constexpr int sum_exp(int ka, int w) { return (w * ka) & (R - 1); }
};

#if defined(__CUDACC__)
// One stage of the lane transform.  tw = this lane's factor for the stage ((1, 0) on lower lanes).
__device__ __forceinline__ cf lane_stage(cf v, int h, bool upper, cf tw, bool last)
{
    const float ox = __shfl_xor_sync(0xffffffffu, v.x, h), oy = __shfl_xor_sync(0xffffffffu, v.y, h);
    // upper: other - own, lower: other + own (sign flip of the own value: one logic op per component)
    const unsigned flip = upper ? 0x80000000u : 0u;
    const cf own = fx::mk(__uint_as_float(__float_as_uint(v.x) ^ flip), __uint_as_float(__float_as_uint(v.y) ^ flip));
    const cf r = ff::add(fx::mk(ox, oy), own);
    return last ? r : ff::fcmul(r, tw);
}

// W-lane forward DFT of one value per lane; lane l (inside its group of W lanes) ends with bin bitrev(l).
// tws: shared-memory table [stage][lane & (W - 1)] of the stage factors.
template <int SF>
__device__ __forceinline__ cf lane_dft(cf v, int l, const cf* __restrict__ tws)
{
    using P = Prune<SF>;
#pragma unroll
    for (int s = 0; s < P::LW; ++s) {
        const int h = P::half(s);
        const bool last = s == P::LW - 1;
        const cf tw = last ? fx::mk(1.0f, 0.0f) : tws[s * P::W + l];
        v = lane_stage(v, h, (l & h) != 0, tw, last);
    }
    return v;
}
#endif

} // namespace fp
