// fft_soa.cuh -- struct-of-arrays packed butterflies for the certified demodulator's throughput transform.
//
// fft_fast.cuh keeps a complex value as an (re, im) register pair: complex add / sub are one packed f32x2
// instruction, but everything that mixes re and im -- the +-j rotations inside the butterflies, the W8 terms,
// every complex multiply, |z|^2 -- stays scalar.  Here a packed register pair holds the SAME component of TWO
// complex values that go through identical arithmetic (lane 0 / lane 1), so every operation of a butterfly,
// twiddle multiply or modulus is one packed instruction for two points:
//   pass 0 (radix 16, decimation in frequency): after the first stage the sums s_e and the twiddled differences
//     d_e feed two independent 8-point transforms -> lanes (s_e, d_e), outputs (X[2q], X[2q+1]);
//   the pass-0 twiddle multiplies pair q by (W^{tid*2q}, W^{tid*(2q+1)}): one packed complex multiply;
//   the exchange moves PAIRS (two 64-bit planes, re and im), so the reading pass gets its lanes -- the two
//   independent 8-point sets of a [16, 8] plan, k0 even / k0 odd -- without any transposition;
//   pass 1: one packed 8-point transform; |Y|^2 of two bins per multiply-add pair.
// Used for SF7 (plan [16, 8]); the scalar / AoS-packed code of fft_fast.cuh serves the other spreading factors.
// Host build: P2 is two floats (tests/cpu/test_fft_fast_emulation.cpp replays the pass and its exchange).
#pragma once
#include <string.h>
#include "fft_fast.cuh"

namespace fs {

#if defined(__CUDA_ARCH__)
struct P2 { unsigned long long v; };
__device__ __forceinline__ P2 mk2(float lo, float hi) { P2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r.v) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ float lo(P2 a) { float x, y; asm("mov.b64 {%0, %1}, %2;" : "=f"(x), "=f"(y) : "l"(a.v)); (void)y; return x; }
__device__ __forceinline__ float hi(P2 a) { float x, y; asm("mov.b64 {%0, %1}, %2;" : "=f"(x), "=f"(y) : "l"(a.v)); (void)x; return y; }
__device__ __forceinline__ P2 add(P2 a, P2 b) { P2 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v)); return r; }
__device__ __forceinline__ P2 sub(P2 a, P2 b) { P2 r; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v)); return r; }
__device__ __forceinline__ P2 mul(P2 a, P2 b) { P2 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v)); return r; }
__device__ __forceinline__ P2 fma(P2 a, P2 b, P2 c) { P2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r.v) : "l"(a.v), "l"(b.v), "l"(c.v)); return r; }
__device__ __forceinline__ unsigned long long bits(P2 a) { return a.v; }
__device__ __forceinline__ P2 from_bits(unsigned long long b) { P2 r; r.v = b; return r; }
#else
struct P2 { float l, h; };
static inline P2 mk2(float lo, float hi) { P2 r; r.l = lo; r.h = hi; return r; }
static inline float lo(P2 a) { return a.l; }
static inline float hi(P2 a) { return a.h; }
static inline P2 add(P2 a, P2 b) { return mk2(a.l + b.l, a.h + b.h); }
static inline P2 sub(P2 a, P2 b) { return mk2(a.l - b.l, a.h - b.h); }
static inline P2 mul(P2 a, P2 b) { return mk2(a.l * b.l, a.h * b.h); }
static inline P2 fma(P2 a, P2 b, P2 c) { return mk2(fmaf(a.l, b.l, c.l), fmaf(a.h, b.h, c.h)); }
static inline unsigned long long bits(P2 a) { unsigned long long b; memcpy(&b, &a, 8); return b; }
static inline P2 from_bits(unsigned long long b) { P2 r; memcpy(&r, &b, 8); return r; }
#endif

// forward 4-point DFT on (re, im) component registers, natural order in and out
FF_HD void dft4(P2& r0, P2& i0, P2& r1, P2& i1, P2& r2, P2& i2, P2& r3, P2& i3)
{
    const P2 t0r = add(r0, r2), t0i = add(i0, i2), t1r = sub(r0, r2), t1i = sub(i0, i2);
    const P2 t2r = add(r1, r3), t2i = add(i1, i3), t3r = sub(r1, r3), t3i = sub(i1, i3);
    r0 = add(t0r, t2r); i0 = add(t0i, t2i);
    r2 = sub(t0r, t2r); i2 = sub(t0i, t2i);
    r1 = add(t1r, t3i); i1 = sub(t1i, t3r);             // t1 - j t3
    r3 = sub(t1r, t3i); i3 = add(t1i, t3r);             // t1 + j t3
}

// forward 8-point DFT, natural order in and out; S = (1/sqrt2, 1/sqrt2), NS = -S
FF_HD void dft8(P2 (&r)[8], P2 (&i)[8], P2 S, P2 NS)
{
    P2 s0r = add(r[0], r[4]), s0i = add(i[0], i[4]), s1r = add(r[1], r[5]), s1i = add(i[1], i[5]);
    P2 s2r = add(r[2], r[6]), s2i = add(i[2], i[6]), s3r = add(r[3], r[7]), s3i = add(i[3], i[7]);
    const P2 d0r = sub(r[0], r[4]), d0i = sub(i[0], i[4]), d1r = sub(r[1], r[5]), d1i = sub(i[1], i[5]);
    const P2 d2r = sub(r[2], r[6]), d2i = sub(i[2], i[6]), d3r = sub(r[3], r[7]), d3i = sub(i[3], i[7]);
    dft4(s0r, s0i, s1r, s1i, s2r, s2i, s3r, s3i);         // X0 X2 X4 X6
    // odd half: d_e * W8^e with W8^1 = S(1 - j), W8^2 = -j, W8^3 = S(-1 - j); S rides on the FMAs
    const P2 p1r = add(d1r, d1i), p1i = sub(d1i, d1r);
    const P2 p3r = sub(d3i, d3r), q3 = add(d3r, d3i);     // p3 = (p3r, -q3)
    const P2 t0r = add(d0r, d2i), t0i = sub(d0i, d2r);    // d0 - j d2
    const P2 t1r = sub(d0r, d2i), t1i = add(d0i, d2r);    // d0 + j d2
    const P2 ur = add(p1r, p3r), ui = sub(p1i, q3);       // p1 + p3
    const P2 vr = sub(p1r, p3r), vi = add(p1i, q3);       // p1 - p3
    r[0] = s0r; i[0] = s0i; r[2] = s1r; i[2] = s1i; r[4] = s2r; i[4] = s2i; r[6] = s3r; i[6] = s3i;
    r[1] = fma(S, ur, t0r);  i[1] = fma(S, ui, t0i);      // t0 + S u
    r[5] = fma(NS, ur, t0r); i[5] = fma(NS, ui, t0i);     // t0 - S u
    r[3] = fma(S, vi, t1r);  i[3] = fma(NS, vr, t1i);     // t1 - j S v
    r[7] = fma(NS, vi, t1r); i[7] = fma(S, vr, t1i);      // t1 + j S v
}

// Pass 0 of the [16, 8] plan on 16 complex values y[c] (sample tid + T*c, already rotated): first radix-2 stage and the
// W16 twiddles in scalar arithmetic (they cross the lanes), written straight into the lane layout (s_e, d_e); then ONE
// packed 8-point transform.  Output pair q = (X[2q], X[2q+1]).
FF_HD void pass0_r16(const fx::cf (&y)[16], P2 (&r)[8], P2 (&i)[8], P2 S, P2 NS)
{
    using fx::fadd; using fx::fsub; using ff::mul_; using ff::fma_;
    const float C1 = 0.92387953251128675613f, S1 = 0.38268343236508977173f, Q = 0.70710678118654752440f;
    float dr[8], di[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) { dr[e] = fsub(y[e].x, y[e + 8].x); di[e] = fsub(y[e].y, y[e + 8].y); }
    // d_e * W16^e, W16^e = (cos(2 pi e/16), -sin(2 pi e/16)): (a + jb)(c - js) = (ac + bs) + j(bc - as)
    float tr[8], ti[8];
    tr[0] = dr[0]; ti[0] = di[0];
    tr[1] = fma_(di[1], S1, mul_(dr[1], C1));   ti[1] = fma_(-dr[1], S1, mul_(di[1], C1));
    tr[2] = mul_(Q, fadd(dr[2], di[2]));        ti[2] = mul_(Q, fsub(di[2], dr[2]));
    tr[3] = fma_(di[3], C1, mul_(dr[3], S1));   ti[3] = fma_(-dr[3], C1, mul_(di[3], S1));
    tr[4] = di[4];                              ti[4] = -dr[4];
    tr[5] = fma_(di[5], C1, mul_(dr[5], -S1));  ti[5] = fma_(-dr[5], C1, mul_(di[5], -S1));
    tr[6] = mul_(Q, fsub(di[6], dr[6]));        ti[6] = mul_(-Q, fadd(dr[6], di[6]));
    tr[7] = fma_(di[7], S1, mul_(dr[7], -C1));  ti[7] = fma_(-dr[7], S1, mul_(di[7], -C1));
#pragma unroll
    for (int e = 0; e < 8; ++e) {
        r[e] = mk2(fadd(y[e].x, y[e + 8].x), tr[e]);
        i[e] = mk2(fadd(y[e].y, y[e + 8].y), ti[e]);
    }
    dft8(r, i, S, NS);
}

// (r, i) *= (wr, wi): packed complex multiply.  The product difference is written as mul - mul: ptxas contracts it into
// one FFMA2 with a negated operand (it does so even under -fmad=false, which is why the exact path never feeds a packed
// multiply into a packed add -- here the fusion is wanted and accuracy is all that matters).
FF_HD void cmul(P2& r, P2& i, P2 wr, P2 wi)
{
    const P2 nr = sub(mul(r, wr), mul(i, wi));
    i = fma(r, wi, mul(i, wr));
    r = nr;
}

// |z|^2 of two bins
FF_HD P2 norm2(P2 r, P2 i) { return fma(i, i, mul(r, r)); }

// Exchange of the [16, 8] plan at SF7 in PAIR entries (8 bytes: one component of two values): the writing thread t
// (= the low index digit i1) holds pairs q = 0..7, the reading thread t' (= q) needs the pairs of every t -- an 8 x 8
// transpose with row pitch 9: conflict-free for the 8 threads of a window, and the neighbouring window of the same
// half-warp sits 152 entries (8 mod 16) away.  Two planes: re at [0, 72), im at [72, 144) of the window's slot.
constexpr int SF7_PLANE = 72;
FF_HD constexpr int sf7_wr(int t, int q) { return t + 9 * q; }
FF_HD constexpr int sf7_rd(int tq, int r) { return 9 * tq + r; }
// bin held by lane `lane` of register pair r on reading thread tq after pass 1: k0 + 16 k1, k0 = 2 tq + lane, k1 = r
FF_HD constexpr int sf7_bin(int tq, int r, int lane) { return 2 * tq + lane + 16 * r; }

} // namespace fs
