// fft_fast.cuh -- the throughput FFT of the margin-certified demodulator path (demod_cert.cu).
//
// fft_exact.cuh replays KISS FFT operation for operation so that spectra are bit-identical to the
// reference (kissfft.hh:106-185); that costs the un-fused radix-4 arithmetic and an exchange layout
// dictated by kissfft's digit order.  This transform is free of that contract: its spectrum only has
// to be ACCURATE (north_star: |X|^2 within 1e-4 relative; measured ~1e-6), because the kernel proves
// per window that the argmax of the reference's own float spectrum is the bin found here, and replays
// the exact chain when it cannot ("certification").
//
//   * N = 2^SF points on T = N/16 threads, 16 complex values per thread, decimation in frequency:
//     pass 0 is a radix-16 butterfly over the samples a thread loads anyway (i = tid + T*c), later
//     passes are radix 16 / 8 / 4 over the remaining index fields (SF7 [16,8], SF8 [16,16],
//     SF9 [16,8,4], SF10 [16,8,8], SF11 [16,16,8], SF12 [16,16,16]);
//   * fused multiply-add everywhere, W8 scalings folded into the next butterfly's FMAs;
//   * pass-0 twiddles (W_N^{tid*k}, 15 per thread) live in registers for the kernel's lifetime, the
//     twiddles of the middle pass come from a 32..256-entry shared-memory table;
//   * exchanges go through ONE buffer with a LINEAR address map: every index bit ("slot") has its own
//     stride, so an element's address is base(thread) + constant(register) -- no per-element address
//     arithmetic -- and the strides are chosen such that the 16 lanes of every 64-bit access phase hit
//     16 different bank pairs (tests/cpu/test_fft_fast_emulation.cpp checks injectivity, conflicts and
//     the spectrum against a double-precision DFT on the host).
//
// Index bookkeeping: the SF index bits of an element are "slots" SF-1..0.  Field p (the bits pass p
// transforms) occupies slots [lo(p), lo(p)+width(p)); before pass p they hold time bits, after it the
// bits of the frequency digit k_p, in place.  bin = sum_p k_p << (width(0)+..+width(p-1)).
#pragma once
#include <stdint.h>
#include "fft_exact.cuh"

#if defined(__CUDACC__)
#define FF_HD __host__ __device__ __forceinline__
#define FF_M __host__ __device__ __forceinline__
#else
#include <cmath>
#define FF_HD static inline
#define FF_M inline
#endif

namespace ff {

using fx::cf;
using fx::mk;

FF_HD float fma_(float a, float b, float c)
{
#if defined(__CUDA_ARCH__)
    return __fmaf_rn(a, b, c);
#else
    return fmaf(a, b, c);
#endif
}
FF_HD float mul_(float a, float b) { return fx::fmul(a, b); }
// complex add / sub: ONE packed f32x2 instruction on the device (the (re, im) pair sits in an aligned register pair)
FF_HD cf add(cf a, cf b) { return fx::cadd(a, b); }
FF_HD cf sub(cf a, cf b) { return fx::csub(a, b); }
// a + s * b and a - s * b with a real scalar s: one packed fused multiply-add each
FF_HD cf axpy(cf a, float s, cf b)
{
#if defined(__CUDA_ARCH__)
    unsigned long long r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(fx::pk2(mk(s, s))), "l"(fx::pk2(b)), "l"(fx::pk2(a)));
    return fx::upk2(r);
#else
    return mk(fma_(s, b.x, a.x), fma_(s, b.y, a.y));
#endif
}
// a * w: two multiplies, two fused multiply-adds
FF_HD cf fcmul(cf a, cf w) { return mk(fma_(-a.y, w.y, mul_(a.x, w.x)), fma_(a.y, w.x, mul_(a.x, w.y))); }
// acc + (z.re^2, z.im^2): one packed fused multiply-add
FF_HD cf sqacc(cf acc, cf z)
{
#if defined(__CUDA_ARCH__)
    unsigned long long r;
    asm("fma.rn.f32x2 %0, %1, %1, %2;" : "=l"(r) : "l"(fx::pk2(z)), "l"(fx::pk2(acc)));
    return fx::upk2(r);
#else
    return mk(fma_(z.x, z.x, acc.x), fma_(z.y, z.y, acc.y));
#endif
}
// |z|^2
FF_HD float fnorm(cf a) { return fma_(a.y, a.y, mul_(a.x, a.x)); }

// ---- butterflies: forward DFT (kernel e^{-2 pi j n k / r}), natural order in, natural order out ----------
FF_HD void dft2(cf& a0, cf& a1)
{
    const cf t = sub(a0, a1);
    a0 = add(a0, a1);
    a1 = t;
}

FF_HD void dft4(cf& a0, cf& a1, cf& a2, cf& a3)
{
    const cf t0 = add(a0, a2), t1 = sub(a0, a2), t2 = add(a1, a3), t3 = sub(a1, a3);
    a0 = add(t0, t2);
    a2 = sub(t0, t2);
    a1 = mk(fx::fadd(t1.x, t3.y), fx::fsub(t1.y, t3.x));      // t1 - j t3
    a3 = mk(fx::fsub(t1.x, t3.y), fx::fadd(t1.y, t3.x));      // t1 + j t3
}

FF_HD void dft8(cf& a0, cf& a1, cf& a2, cf& a3, cf& a4, cf& a5, cf& a6, cf& a7)
{
    const float S = 0.70710678118654752440f;
    cf s0 = add(a0, a4), s1 = add(a1, a5), s2 = add(a2, a6), s3 = add(a3, a7);
    const cf d0 = sub(a0, a4), d1 = sub(a1, a5), d2 = sub(a2, a6), d3 = sub(a3, a7);
    dft4(s0, s1, s2, s3);                                       // X0 X2 X4 X6
    // odd half: d_e * W8^e.  W8^1 = S(1 - j), W8^2 = -j, W8^3 = S(-1 - j); the factor S rides on the FMAs below
    const cf p1 = mk(fx::fadd(d1.x, d1.y), fx::fsub(d1.y, d1.x));
    const cf p3 = mk(fx::fsub(d3.y, d3.x), fx::fsub(-d3.x, d3.y));
    const cf r2 = mk(d2.y, -d2.x);
    const cf t0 = add(d0, r2), t1 = sub(d0, r2);
    const cf u = add(p1, p3), v = sub(p1, p3);                  // t2 = S u, t3 = S v
    a0 = s0; a2 = s1; a4 = s2; a6 = s3;
    a1 = axpy(t0, S, u);                                        // t0 + t2
    a5 = axpy(t0, -S, u);                                       // t0 - t2
    a3 = mk(fma_(S, v.y, t1.x), fma_(-S, v.x, t1.y));           // t1 - j t3
    a7 = mk(fma_(-S, v.y, t1.x), fma_(S, v.x, t1.y));           // t1 + j t3
}

FF_HD void dft16(cf (&a)[16])
{
    const float C1 = 0.92387953251128675613f, S1 = 0.38268343236508977173f, S = 0.70710678118654752440f;
    cf s[8], d[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) { s[e] = add(a[e], a[e + 8]); d[e] = sub(a[e], a[e + 8]); }
    // d_e * W16^e, W16^e = (cos(2 pi e/16), -sin(2 pi e/16))
    d[1] = fcmul(d[1], mk(C1, -S1));
    d[2] = mk(mul_(S, fx::fadd(d[2].x, d[2].y)), mul_(S, fx::fsub(d[2].y, d[2].x)));
    d[3] = fcmul(d[3], mk(S1, -C1));
    d[4] = mk(d[4].y, -d[4].x);
    d[5] = fcmul(d[5], mk(-S1, -C1));
    d[6] = mk(mul_(S, fx::fsub(d[6].y, d[6].x)), mul_(-S, fx::fadd(d[6].x, d[6].y)));
    d[7] = fcmul(d[7], mk(-C1, -S1));
    dft8(s[0], s[1], s[2], s[3], s[4], s[5], s[6], s[7]);
    dft8(d[0], d[1], d[2], d[3], d[4], d[5], d[6], d[7]);
#pragma unroll
    for (int q = 0; q < 8; ++q) { a[2 * q] = s[q]; a[2 * q + 1] = d[q]; }
}

// ---- plan: radices, register / thread slots of every pass, exchange strides -----------------------------------
template <int SF>
struct Plan {
    static_assert(SF >= 7 && SF <= 12, "fast plan covers SF7..SF12");
    static constexpr int N = 1 << SF;
    static constexpr int R = 16;
    static constexpr int T = N / 16;
    static constexpr int LOG_T = SF - 4;
    static constexpr int NP = SF <= 8 ? 2 : 3;

    // bits transformed by pass p
    FF_M static constexpr int width(int p)
    {
        if (p == 0) return 4;
        if (SF == 7) return 3;
        if (SF == 8) return 4;
        if (SF == 9) return p == 1 ? 3 : 2;
        if (SF == 10) return 3;
        if (SF == 11) return p == 1 ? 4 : 3;
        return 4;
    }
    // lowest slot of field p
    FF_M static constexpr int lo(int p)
    {
        int s = SF;
        for (int q = 0; q <= p; ++q) s -= width(q);
        return s;
    }
    // slot carried by bit b of the register index during pass p: bits [0, width) are the field's own slots (the
    // digit), the rest select the independent set and are slots of frequency digits already computed
    FF_M static constexpr int reg_slot(int p, int b)
    {
        if (b < width(p)) return lo(p) + b;
        if (SF == 9 && p == 2) return b == 2 ? 4 : 8;            // sets: k1 bit 2, k0 bit 3
        if (SF == 10 && p == 2) return 5;                        // k1 bit 2
        if (SF == 11 && p == 2) return 6;                        // k1 bit 3
        return SF - 1;                                            // k0 bit 3 (SF7 p1, SF9 p1, SF10 p1)
    }
    FF_M static constexpr bool slot_in_regs(int p, int s)
    {
        for (int b = 0; b < 4; ++b) if (reg_slot(p, b) == s) return true;
        return false;
    }
    // slot carried by bit b of the thread index during pass p: every slot not in the registers, ascending
    FF_M static constexpr int th_slot(int p, int b)
    {
        int n = 0;
        for (int s = 0; s < SF; ++s) {
            if (slot_in_regs(p, s)) continue;
            if (n == b) return s;
            ++n;
        }
        return -1;
    }
    // Exchange x (between pass x and x + 1): address stride of every slot, in complex entries.  The low four thread
    // slots of the writing pass and of the reading pass each get strides that are distinct sums modulo 16
    // (bank pairs of a 64-bit access phase); every stride exceeds the sum of the smaller ones (injective map).
    FF_M static constexpr int stride(int x, int s)
    {
        if (SF == 7) { const int t[7] = { 1, 2, 4, 9, 18, 36, 72 }; return t[s]; }
        if (SF == 8) { const int t[8] = { 1, 2, 4, 8, 17, 34, 68, 136 }; return t[s]; }
        if (SF == 9) {
            const int t0[9] = { 1, 2, 4, 8, 16, 36, 72, 144, 288 };
            const int t1[9] = { 1, 2, 17, 34, 68, 4, 8, 136, 272 };
            return x == 0 ? t0[s] : t1[s];
        }
        if (SF == 10) {
            const int t0[10] = { 1, 2, 4, 8, 16, 32, 72, 144, 288, 576 };
            const int t1[10] = { 1, 2, 4, 17, 34, 68, 8, 148, 288, 576 };
            return x == 0 ? t0[s] : t1[s];
        }
        if (SF == 11) {
            const int t0[11] = { 1, 2, 4, 8, 16, 32, 64, 136, 272, 544, 1088 };
            const int t1[11] = { 1, 2, 4, 17, 34, 68, 136, 8, 272, 544, 1088 };
            return x == 0 ? t0[s] : t1[s];
        }
        {
            const int t0[12] = { 1, 2, 4, 8, 16, 32, 64, 128, 256, 512, 1024, 2048 };
            const int t1[12] = { 1, 2, 4, 8, 17, 34, 68, 136, 272, 544, 1088, 2176 };
            return x == 0 ? t0[s] : t1[s];
        }
    }
    FF_M static constexpr int footprint(int x)
    {
        int n = 1;
        for (int s = 0; s < SF; ++s) n += stride(x, s);
        return n;
    }
    // entries one window needs in its stage slot: the raw samples (+1 for an odd-aligned fetch) and every exchange;
    // a multiple of 16 plus 8 for SF7 so that the two windows a half-warp touches differ by 8 bank pairs
    FF_M static constexpr int pitch()
    {
        int n = N + 2;
        for (int x = 0; x + 1 < NP; ++x) if (footprint(x) > n) n = footprint(x);
        n = (n + 15) & ~15;
        return SF == 7 ? n + 8 : n;
    }

    // ---- per-register (compile-time) and per-thread (start-up) parts of the linear address maps ---------------
    FF_M static constexpr int wr_off(int x, int e)
    {
        int a = 0;
        for (int b = 0; b < 4; ++b) if ((e >> b) & 1) a += stride(x, reg_slot(x, b));
        return a;
    }
    FF_M static constexpr int rd_off(int x, int e)
    {
        int a = 0;
        for (int b = 0; b < 4; ++b) if ((e >> b) & 1) a += stride(x, reg_slot(x + 1, b));
        return a;
    }
    FF_M static int wr_base(int x, int tid)
    {
        int a = 0;
        for (int b = 0; b < LOG_T; ++b) if ((tid >> b) & 1) a += stride(x, th_slot(x, b));
        return a;
    }
    FF_M static int rd_base(int x, int tid)
    {
        int a = 0;
        for (int b = 0; b < LOG_T; ++b) if ((tid >> b) & 1) a += stride(x, th_slot(x + 1, b));
        return a;
    }
    // weight of a slot in the output bin: field p contributes k_p << (width(0) + .. + width(p-1))
    FF_M static constexpr int bin_weight(int s)
    {
        int sh = 0;
        for (int p = 0; p < NP; ++p) {
            if (s >= lo(p) && s < lo(p) + width(p)) return 1 << (sh + s - lo(p));
            sh += width(p);
        }
        return 0;
    }
    FF_M static constexpr int bin_off(int e)
    {
        int a = 0;
        for (int b = 0; b < 4; ++b) if ((e >> b) & 1) a += bin_weight(reg_slot(NP - 1, b));
        return a;
    }
    FF_M static int bin_base(int tid)
    {
        int a = 0;
        for (int b = 0; b < LOG_T; ++b) if ((tid >> b) & 1) a += bin_weight(th_slot(NP - 1, b));
        return a;
    }
    // Twiddle after pass p < NP-1: W_M^{k * i_rest}, M = 2^(lo(p)+width(p)), k = the pass's frequency digit (register
    // bits [0, width)), i_rest = the element's remaining time index = its slots below lo(p).  In every plan above those
    // slots are thread slots, so i_rest is a per-thread constant.
    FF_M static int i_rest(int p, int tid)
    {
        int a = 0;
        for (int b = 0; b < LOG_T; ++b) {
            const int s = th_slot(p, b);
            if (s < lo(p) && ((tid >> b) & 1)) a += 1 << s;
        }
        return a;
    }
    // middle-pass twiddle table (3-pass plans): entry [k << lo(1) | i_rest] = W_M1^{k * i_rest}
    static constexpr int TW1_SIZE = NP == 3 ? (1 << (lo(1) + width(1))) : 0;
    // exponent of W_N for the pass-0 twiddle of register k on thread tid (lo(0) + 4 == SF: i_rest(0) == tid)
    FF_M static int tw0_exp(int tid, int k) { return (k * tid) & (N - 1); }

    // ---- the passes ---------------------------------------------------------------------------------------------
    FF_M static void butterflies(int p, cf (&v)[16])
    {
        const int w = width(p);
        if (w == 4) {
            dft16(v);
        } else if (w == 3) {
            dft8(v[0], v[1], v[2], v[3], v[4], v[5], v[6], v[7]);
            dft8(v[8], v[9], v[10], v[11], v[12], v[13], v[14], v[15]);
        } else {
#pragma unroll
            for (int s = 0; s < 4; ++s) dft4(v[4 * s], v[4 * s + 1], v[4 * s + 2], v[4 * s + 3]);
        }
    }
};

} // namespace ff
