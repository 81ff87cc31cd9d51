// SPDX-License-Identifier: LGPL-2.1-or-later
// Attribution: the algorithms, polynomial coefficients and tables below restate functions of the GNU C Library 2.39
// (sysdeps/ieee754/flt-32/{s_sinf,s_cosf,s_sincosf,sincosf.h,e_logf,e_log10f,s_atanf,e_atan2f}.c, sysdeps/ieee754/dbl-64/
// e_hypotf-style hypotf), Copyright (C) Free Software Foundation, Inc., distributed under the GNU Lesser General
// Public License v2.1 or later; atanf/atan2f/log10f descend from fdlibm, Copyright (C) 1993 Sun Microsystems, Inc.
// ("Permission to use, copy, modify, and distribute this software is freely granted, provided that this notice is
// preserved").  This file is a derived work of that material and carries the same licence; it is third-party-derived,
// not reference-derived (the reference only CALLS libm).
//
// glibc_math.cuh -- device restatements of the glibc 2.39 libm entry points the
// reference's hot path resolves to (SURVEY.md section 1, L0 "Below" column):
//
//   std::cos/std::sin(float)  -> cosf / sinf     (LoRaDemod.cpp:156-157, phy.cpp:221-222,
//                                                 ChirpGenerator.hpp:39,47 via std::polar,
//                                                 kissfft.hh:28 via cexpf)
//   std::abs(complex<float>)  -> hypotf          (LoRaDetector.hpp:66-67)
//   std::arg(complex<float>)  -> atan2f          (LoRaDemod.cpp:115, phy.cpp:125)
//   std::log10(float)         -> log10f          (LoRaDetector.hpp:63-64)
//
// Why restate instead of calling CUDA's libdevice: the reference rotates every sample
// by (cosf(ph), sinf(ph)) before the FFT and takes argmax of |X|^2 with a strict
// first-max tie-break.  Demodulated symbols are required to be bit-exact, so the
// rotated samples must be the same floats the reference produces.  glibc's sinf/cosf
// evaluate fixed double-precision polynomials (error 0.56 ulp, NOT correctly rounded),
// so only the same polynomial reproduces them; CUDA's sinf (<= 2 ulp) does not.  B200
// executes FP64 at half the FP32 rate, which makes the double-precision evaluation
// affordable per sample ("K1 instruction budget").
//
// Everything is __host__ __device__ so tests/cpu/test_glibc_math.cpp can compare each
// function with the real glibc over hundreds of millions of arguments on the CPU.
// Compile with -fmad=false (nvcc) / -ffp-contract=off (g++): float-path functions
// (atanf, atan2f, log10f) rely on one rounding per operation.
#pragma once

#include <stdint.h>
#include <math.h>

#if defined(__CUDACC__)
#define GM_HD __host__ __device__ __forceinline__
#else
#define GM_HD static inline
#endif

namespace gm {

GM_HD uint32_t f2u(float f) {
#if defined(__CUDA_ARCH__)
    return __float_as_uint(f);
#else
    union { float f; uint32_t u; } v; v.f = f; return v.u;
#endif
}
GM_HD float u2f(uint32_t u) {
#if defined(__CUDA_ARCH__)
    return __uint_as_float(u);
#else
    union { float f; uint32_t u; } v; v.u = u; return v.f;
#endif
}
GM_HD double ll2d_bits(uint64_t u) {
#if defined(__CUDA_ARCH__)
    return __longlong_as_double((long long)u);
#else
    union { double d; uint64_t u; } v; v.u = u; return v.d;
#endif
}

// ------------------------------------------------------------------------------------
// sinf / cosf : glibc sysdeps/ieee754/flt-32/{s_sinf.c,s_cosf.c,s_sincosf.c,sincosf.h}
//
// x86-64 glibc resolves sinf/cosf/sincosf through an ifunc to the *_fma build of those
// files (sysdeps/x86_64/fpu/multiarch) on every CPU with FMA+AVX2 -- i.e. on every host a
// B200 is attached to.  That build contracts each `a + b*c` of sincosf.h into one fused
// multiply-add.  The restatement below fuses exactly those expressions (and no others), and
// tests/cpu/test_glibc_math.cpp shows 0 mismatches against the running glibc over ALL 2^32
// float arguments for sinf, cosf and sincosf.
// ------------------------------------------------------------------------------------
GM_HD double dfma(double a, double b, double c) {
#if defined(__CUDA_ARCH__)
    return __fma_rn(a, b, c);
#else
    return fma(a, b, c);
#endif
}
GM_HD double dmul(double a, double b) {
#if defined(__CUDA_ARCH__)
    return __dmul_rn(a, b);
#else
    return a * b;
#endif
}

GM_HD double sc_hpi_inv() { return 0x1.45F306DC9C883p+23; }   // 2/pi * 2^24
GM_HD double sc_hpi()     { return 0x1.921FB54442D18p0; }      // pi/2

// sincosf.h polynomials on the reduced argument (|x| <= pi/4), __sincosf_table[0].
// Table [1] negates every cosine coefficient, which negates the result exactly; the
// quadrant sign multiplies x by +-1, which negates the (odd) sine result exactly.
GM_HD double sc_sin_poly(double x, double x2)
{
    const double s1 = -0x1.555545995a603p-3, s2 = 0x1.1107605230bc4p-7, s3 = -0x1.994eb3774cf24p-13;
    const double x3 = dmul(x, x2);
    const double t1 = dfma(x2, s3, s2);
    const double x7 = dmul(x3, x2);
    const double s = dfma(x3, s1, x);
    return dfma(x7, t1, s);
}
GM_HD double sc_cos_poly(double x2)
{
    const double c0 = 0x1p0, c1 = -0x1.ffffffd0c621cp-2, c2 = 0x1.55553e1068f19p-5,
                 c3 = -0x1.6c087e89a359dp-10, c4 = 0x1.99343027bf8c3p-16;
    const double x4 = dmul(x2, x2);
    const double t2 = dfma(x2, c4, c3);
    const double t1 = dfma(x2, c1, c0);
    const double x6 = dmul(x4, x2);
    const double c = dfma(x4, c2, t1);
    return dfma(x6, t2, c);
}

// __inv_pio4[]: consecutive 32-bit windows (8-bit steps) of 4/pi's binary expansion.  Tables indexed at run time live at
// namespace scope (device: global memory through the read-only cache): as function-local arrays they were rebuilt
// in LOCAL memory by every call -- two dozen stores per sample in the Monte-Carlo channel.
#define GM_INV_PIO4_INIT { \
        0xa2u, 0xa2f9u, 0xa2f983u, 0xa2f9836eu, 0xf9836e4eu, 0x836e4e44u, 0x6e4e4415u, 0x4e441529u, \
        0x441529fcu, 0x1529fc27u, 0x29fc2757u, 0xfc2757d1u, 0x2757d1f5u, 0x57d1f534u, 0xd1f534ddu, \
        0xf534ddc0u, 0x34ddc0dbu, 0xddc0db62u, 0xc0db6295u, 0xdb629599u, 0x6295993cu, 0x95993c43u, \
        0x993c4390u, 0x3c439041u }
#if defined(__CUDACC__)
static __device__ const uint32_t gm_inv_pio4_d[24] = GM_INV_PIO4_INIT;
#endif
static const uint32_t gm_inv_pio4_h[24] = GM_INV_PIO4_INIT;
GM_HD uint32_t sc_inv_pio4(int i)
{
#if defined(__CUDA_ARCH__)
    return __ldg(&gm_inv_pio4_d[i]);
#else
    return gm_inv_pio4_h[i];
#endif
}

GM_HD double sc_reduce_large(uint32_t xi, int* np)
{
    const int a = (xi >> 26) & 15;
    const int shift = (xi >> 23) & 7;
    uint64_t n, res0, res1, res2;
    xi = (xi & 0xffffff) | 0x800000;
    xi <<= shift;
    res0 = (uint32_t)(xi * sc_inv_pio4(a));           // 32-bit product promoted, as in glibc
    res1 = (uint64_t)xi * sc_inv_pio4(a + 4);
    res2 = (uint64_t)xi * sc_inv_pio4(a + 8);
    res0 = (res2 >> 32) | (res0 << 32);
    res0 += res1;
    n = (res0 + (1ULL << 61)) >> 62;
    res0 -= n << 62;
    double x = (double)(int64_t)res0;
    *np = (int)n;
    return dmul(x, 0x1.921FB54442D18p-62);              // pi63
}

// Quadrant bookkeeping (sincosf.h sincosf_poly + the callers' sign set-up): with Ps / Pc the
// sine / cosine polynomial on the reduced argument, n the quadrant from the reduction and
// m = n + (sign of y, large-argument path only):  S = Ps * sign[m&3], C = Pc * (m&2 ? -1 : 1),
// and (sin, cos) = n odd ? (C, S) : (S, C).
GM_HD void sc_assemble(float ps, float pc, unsigned n, unsigned m, float* sn, float* cs)
{
    const uint32_t S = f2u(ps) ^ (((m + 1u) & 2u) << 30);
    const uint32_t Cc = f2u(pc) ^ ((m & 2u) << 30);
    *sn = u2f((n & 1u) ? Cc : S);
    *cs = u2f((n & 1u) ? S : Cc);
}

// Core of the |y| < 120 path: reduction + both polynomials + quadrant assembly.  For
// |y| < pi/4 the reduction yields n = 0 and x unchanged, so glibc's separate small-argument
// branch computes the same thing.  NOT valid for |y| < 2^-12 (glibc returns (y, 1.0f) there).
//
// glibc's reduce_fast computes n = ((int32_t)(x * hpi_inv) + 0x800000) >> 24 (truncate, bias, floor)
// and x - n * hpi.  Here the quadrant comes from the magic-number trick instead: adding 1.5 * 2^76
// rounds q = x * hpi_inv to the nearest multiple of 2^24 (ulp of that binade), leaves n in the low
// mantissa bits and n * 2^24 after subtracting the constant again -- two DADDs instead of
// F2I + IADD + SHF + I2F, and no trip through the conversion pipe.  Round-to-nearest-even differs
// from glibc's truncate-and-bias only for q within 1 of a half-way point; tests/cpu/test_glibc_math.cpp
// walks ALL floats with 2^-12 <= |y| < 120 (both signs): sin and cos are bit-identical to glibc for
// every one of them (0 mismatches), so the substitution is exact on the domain this core is used on.
GM_HD uint32_t d_lo32(double d) {
#if defined(__CUDA_ARCH__)
    return (uint32_t)__double2loint(d);
#else
    union { double d; uint64_t u; } v; v.d = d; return (uint32_t)v.u;
#endif
}
GM_HD double dadd(double a, double b) {
#if defined(__CUDA_ARCH__)
    return __dadd_rn(a, b);
#else
    return a + b;
#endif
}
// a ^ (b & 0x80000000): one LOP3 on the device
GM_HD uint32_t xor_signbit_of(uint32_t a, uint32_t b) {
#if defined(__CUDA_ARCH__)
    uint32_t d;
    asm("lop3.b32 %0, %1, %2, 0x80000000, 0x78;" : "=r"(d) : "r"(a), "r"(b));
    return d;
#else
    return a ^ (b & 0x80000000u);
#endif
}

GM_HD void sincosf_core(float y, float* sn, float* cs)
{
    const double x = (double)y;
    const double q = dmul(x, sc_hpi_inv());
    const double magic = 0x1.8p76;
    const double t = dadd(q, magic);
    const uint32_t n = d_lo32(t);                         // quadrant (mod 2^32)
    const double nd = dadd(t, -magic);                    // n * 2^24, exact
    const double xr = dfma(-nd, 0x1.921FB54442D18p-24, x); // x - n * pi/2, one rounding like glibc's FMA build
    const double x2 = dmul(xr, xr);
    const float ps = (float)sc_sin_poly(xr, x2);
    const float pc = (float)sc_cos_poly(x2);
    // sc_assemble with m == n: sin = odd ? pc : ps with sign bit1(n); cos = odd ? ps : pc with sign bit1(n + 1)
    const uint32_t t30 = n << 30;
    const bool odd = (n & 1u) != 0;
    const uint32_t a = f2u(odd ? pc : ps), b = f2u(odd ? ps : pc);
    *sn = u2f(xor_signbit_of(a, t30));
    *cs = u2f(xor_signbit_of(b, t30 + 0x40000000u));
}

// |y| < 120, any magnitude: the core plus glibc's tiny-argument result.
GM_HD void sincosf_mid(float y, float* sn, float* cs)
{
    float s, c;
    sincosf_core(y, &s, &c);
    if ((f2u(y) & 0x7fffffffu) < 0x39800000u) { s = y; c = 1.0f; }   // |y| < 2^-12 (n == 0 there)
    *sn = s; *cs = c;
}

GM_HD void sincosf_g(float y, float* sn, float* cs)
{
    const uint32_t iy = f2u(y), ay = iy & 0x7fffffffu;
    if (ay < 0x42f00000u) { sincosf_mid(y, sn, cs); return; }       // |y| < 120
    if (ay >= 0x7f800000u) { *sn = *cs = y - y; return; }           // inf / NaN
    int n;
    const double xr = sc_reduce_large(iy, &n);
    const double x2 = dmul(xr, xr);
    sc_assemble((float)sc_sin_poly(xr, x2), (float)sc_cos_poly(x2), (unsigned)n, (unsigned)n + (iy >> 31), sn, cs);
}

GM_HD float sinf_g(float y) { float s, c; sincosf_g(y, &s, &c); return s; }
GM_HD float cosf_g(float y) { float s, c; sincosf_g(y, &s, &c); return c; }

// ------------------------------------------------------------------------------------
// hypotf : glibc sysdeps/ieee754/flt-32/e_hypotf.c (2.35+): double-precision sqrt
// ------------------------------------------------------------------------------------
GM_HD float hypotf_g(float x, float y)
{
    const uint32_t ax = f2u(x) & 0x7fffffffu, ay = f2u(y) & 0x7fffffffu;
    if (ax == 0x7f800000u || ay == 0x7f800000u) return u2f(0x7f800000u);  // inf wins over NaN
    if (ax > 0x7f800000u || ay > 0x7f800000u) return x + y;
#if defined(__CUDA_ARCH__)
    return (float)__dsqrt_rn((double)x * (double)x + (double)y * (double)y);
#else
    return (float)sqrt((double)x * (double)x + (double)y * (double)y);
#endif
}

// ------------------------------------------------------------------------------------
// atanf / atan2f : glibc sysdeps/ieee754/flt-32/{s_atanf.c,e_atan2f.c} (fdlibm, float ops)
// ------------------------------------------------------------------------------------
GM_HD float atanf_g(float x)
{
    const float hi[4] = { 4.6364760399e-01f, 7.8539812565e-01f, 9.8279368877e-01f, 1.5707962513e+00f };
    const float lo[4] = { 5.0121582440e-09f, 3.7748947079e-08f, 3.4473217170e-08f, 7.5497894159e-08f };
    // (indexed through selects below, never through memory)
    const float aT[11] = { 3.3333334327e-01f, -2.0000000298e-01f, 1.4285714924e-01f, -1.1111110449e-01f,
                           9.0908870101e-02f, -7.6918758452e-02f, 6.6610731184e-02f, -5.8335702866e-02f,
                           4.9768779427e-02f, -3.6531571299e-02f, 1.6285819933e-02f };
    const int32_t hx = (int32_t)f2u(x);
    const int32_t ix = hx & 0x7fffffff;
    int id;
    if (ix >= 0x4c000000) {                    // |x| >= 2^25
        if (ix > 0x7f800000) return x + x;
        return hx > 0 ? hi[3] + lo[3] : -hi[3] - lo[3];
    }
    if (ix < 0x3ee00000) {                     // |x| < 0.4375
        if (ix < 0x31000000) return x;         // |x| < 2^-29
        id = -1;
    } else {
        x = fabsf(x);
        if (ix < 0x3f980000) {
            if (ix < 0x3f300000) { id = 0; x = (2.0f * x - 1.0f) / (2.0f + x); }
            else                 { id = 1; x = (x - 1.0f) / (x + 1.0f); }
        } else {
            if (ix < 0x401c0000) { id = 2; x = (x - 1.5f) / (1.0f + 1.5f * x); }
            else                 { id = 3; x = -1.0f / x; }
        }
    }
    float z = x * x;
    float w = z * z;
    float s1 = z * (aT[0] + w * (aT[2] + w * (aT[4] + w * (aT[6] + w * (aT[8] + w * aT[10])))));
    float s2 = w * (aT[1] + w * (aT[3] + w * (aT[5] + w * (aT[7] + w * aT[9]))));
    if (id < 0) return x - x * (s1 + s2);
    const float hid = id == 0 ? hi[0] : id == 1 ? hi[1] : id == 2 ? hi[2] : hi[3];
    const float lod = id == 0 ? lo[0] : id == 1 ? lo[1] : id == 2 ? lo[2] : lo[3];
    z = hid - ((x * (s1 + s2) - lod) - x);
    return hx < 0 ? -z : z;
}

GM_HD float atan2f_g(float y, float x)
{
    const float tiny = 1.0e-30f, pi_o_4 = 7.8539818525e-01f, pi_o_2 = 1.5707963705e+00f,
                pi = 3.1415927410e+00f, pi_lo = -8.7422776573e-08f;
    const int32_t hx = (int32_t)f2u(x), hy = (int32_t)f2u(y);
    const int32_t ix = hx & 0x7fffffff, iy = hy & 0x7fffffff;
    if (ix > 0x7f800000 || iy > 0x7f800000) return x + y;
    if (hx == 0x3f800000) return atanf_g(y);
    const int m = ((hy >> 31) & 1) | ((hx >> 30) & 2);
    if (iy == 0) {
        switch (m) {
        case 0: case 1: return y;
        case 2: return pi + tiny;
        default: return -pi - tiny;
        }
    }
    if (ix == 0) return hy < 0 ? -pi_o_2 - tiny : pi_o_2 + tiny;
    if (ix == 0x7f800000) {
        if (iy == 0x7f800000) {
            switch (m) {
            case 0: return pi_o_4 + tiny;
            case 1: return -pi_o_4 - tiny;
            case 2: return 3.0f * pi_o_4 + tiny;
            default: return -3.0f * pi_o_4 - tiny;
            }
        } else {
            switch (m) {
            case 0: return 0.0f;
            case 1: return -0.0f;
            case 2: return pi + tiny;
            default: return -pi - tiny;
            }
        }
    }
    if (iy == 0x7f800000) return hy < 0 ? -pi_o_2 - tiny : pi_o_2 + tiny;
    const int k = (iy - ix) >> 23;
    float z;
    if (k > 60) z = pi_o_2 + 0.5f * pi_lo;
    else if (hx < 0 && k < -60) z = 0.0f;
    else z = atanf_g(fabsf(y / x));
    switch (m) {
    case 0: return z;
    case 1: return u2f(f2u(z) ^ 0x80000000u);
    case 2: return pi - (z - pi_lo);
    default: return (z - pi_lo) - pi;
    }
}

// ------------------------------------------------------------------------------------
// logf / log10f : glibc sysdeps/ieee754/flt-32/{e_logf.c,e_log10f.c}
// ------------------------------------------------------------------------------------
#define GM_LOGF_TAB_INIT { \
        0x1.661ec79f8f3bep+0, -0x1.57bf7808caadep-2, 0x1.571ed4aaf883dp+0, -0x1.2bef0a7c06ddbp-2, \
        0x1.49539f0f010bp+0, -0x1.01eae7f513a67p-2, 0x1.3c995b0b80385p+0, -0x1.b31d8a68224e9p-3, \
        0x1.30d190c8864a5p+0, -0x1.6574f0ac07758p-3, 0x1.25e227b0b8eap+0, -0x1.1aa2bc79c81p-3, \
        0x1.1bb4a4a1a343fp+0, -0x1.a4e76ce8c0e5ep-4, 0x1.12358f08ae5bap+0, -0x1.1973c5a611cccp-4, \
        0x1.0953f419900a7p+0, -0x1.252f438e10c1ep-5, 0x1p+0, 0x0p+0, \
        0x1.e608cfd9a47acp-1, 0x1.aa5aa5df25984p-5, 0x1.ca4b31f026aap-1, 0x1.c5e53aa362eb4p-4, \
        0x1.b2036576afce6p-1, 0x1.526e57720db08p-3, 0x1.9c2d163a1aa2dp-1, 0x1.bc2860d22477p-3, \
        0x1.886e6037841edp-1, 0x1.1058bc8a07ee1p-2, 0x1.767dcf5534862p-1, 0x1.4043057b6ee09p-2 }
#if defined(__CUDACC__)
static __device__ const double gm_logf_tab_d[32] = GM_LOGF_TAB_INIT;
#endif
static const double gm_logf_tab_h[32] = GM_LOGF_TAB_INIT;
GM_HD void logf_tab(int i, double* invc, double* logc)
{
#if defined(__CUDA_ARCH__)
    *invc = __ldg(&gm_logf_tab_d[2 * i]); *logc = __ldg(&gm_logf_tab_d[2 * i + 1]);
#else
    *invc = gm_logf_tab_h[2 * i]; *logc = gm_logf_tab_h[2 * i + 1];
#endif
}

GM_HD float logf_g(float x)
{
    uint32_t ix = f2u(x);
    if (ix == 0x3f800000u) return 0.0f;
    if (ix - 0x00800000u >= 0x7f800000u - 0x00800000u) {
        if (ix * 2 == 0) return -u2f(0x7f800000u);                     // log(+-0) = -inf
        if (ix == 0x7f800000u) return x;
        if ((ix & 0x80000000u) || ix * 2 >= 0xff000000u) return (x - x) / (x - x); // NaN
        ix = f2u(x * 0x1p23f);
        ix -= 23u << 23;
    }
    const uint32_t tmp = ix - 0x3f330000u;
    const int i = (int)((tmp >> 19) % 16);
    const int k = (int32_t)tmp >> 23;
    const uint32_t iz = ix - (tmp & (0x1ffu << 23));
    double invc, logc;
    logf_tab(i, &invc, &logc);
    const double z = (double)u2f(iz);
#if defined(GM_LOGF_NOFMA)
    const double r = z * invc - 1;
    const double y0 = logc + (double)k * 0x1.62e42fefa39efp-1;
    const double r2 = r * r;
    double y = 0x1.5575b0be00b6ap-2 * r + -0x1.ffffef20a4123p-2;
    y = -0x1.00ea348b88334p-2 * r2 + y;
    y = y * r2 + (y0 + r);
#else   // e_logf.c as built for the x86-64 FMA ifunc variant: every a*b+c fused
    const double r = dfma(z, invc, -1.0);
    const double y0 = dfma((double)k, 0x1.62e42fefa39efp-1, logc);
    const double r2 = dmul(r, r);
    double y = dfma(0x1.5575b0be00b6ap-2, r, -0x1.ffffef20a4123p-2);
    y = dfma(-0x1.00ea348b88334p-2, r2, y);
    y = dfma(y, r2, y0 + r);
#endif
    return (float)y;
}

GM_HD float log10f_g(float x)
{
    const float two25 = 3.3554432000e+07f, ivln10 = 4.3429449201e-01f,
                log10_2hi = 3.0102920532e-01f, log10_2lo = 7.9034151668e-07f;
    int32_t hx = (int32_t)f2u(x);
    int32_t k = 0;
    if (hx < 0x00800000) {
        if ((hx & 0x7fffffff) == 0) return -two25 / fabsf(x);   // -inf
        if (hx < 0) return (x - x) / (x - x);                   // NaN
        k -= 25; x *= two25;
        hx = (int32_t)f2u(x);
    }
    if (hx >= 0x7f800000) return x + x;
    k += (hx >> 23) - 127;
    const int32_t i = (int32_t)(((uint32_t)k & 0x80000000u) >> 31);
    hx = (hx & 0x007fffff) | ((0x7f - i) << 23);
    const float y = (float)(k + i);
    x = u2f((uint32_t)hx);
    const float z = y * log10_2lo + ivln10 * logf_g(x);
    return z + y * log10_2hi;
}

} // namespace gm
