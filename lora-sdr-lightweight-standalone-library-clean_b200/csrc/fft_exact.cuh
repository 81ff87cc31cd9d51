// fft_exact.cuh -- register/shared-memory FFT whose every floating-point operation is the
// one KISS FFT performs for the same transform (reference: include/lora_phy/kissfft.hh
// :81-97 stage plan, :106-143 decimation-in-time recursion, :155-162 kf_bfly2, :164-185
// kf_bfly4).  IEEE add/sub/mul without contraction are deterministic, so reproducing the
// operand order reproduces the reference spectrum bit for bit -- which is what makes the
// strict first-max argmax (LoRaDetector.hpp:46-58) agree on noise-only and tied inputs.
//
// This is a B200 design, not a port of the recursion:
//   * one transform is spread over T = N/16 threads, each holding 16 complex values in
//     registers; two radix-4 stages (or the odd radix-2 + radix-4 head) are fused into one
//     register "pass", so a 4096-point transform needs only 2 shared-memory exchanges;
//   * pass 0 consumes samples in the order the global loads deliver them (thread t holds
//     samples t, t+T, t+2T, ...: coalesced) -- the digit reversal of the recursion's leaf
//     copy is folded into the first exchange's store addresses;
//   * twiddles for the fused passes come from compact per-pass tables (consecutive lanes
//     read consecutive entries) gathered from the exact kissfft table.
//
// Everything is __host__ __device__: tests/cpu/test_fft_emulation.cpp runs the same code
// with an emulated thread loop against the oracle and counts shared-memory bank conflicts.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define FX_HD __host__ __device__ __forceinline__
#define FX_MHD __host__ __device__ __forceinline__
#else
#define FX_HD static inline
#define FX_MHD inline
#endif

namespace fx {

struct alignas(8) cf { float x, y; };

// one rounding per operation, never contracted into FMA
FX_HD float fmul(float a, float b) {
#if defined(__CUDA_ARCH__)
    return __fmul_rn(a, b);
#else
    return a * b;
#endif
}
FX_HD float fadd(float a, float b) {
#if defined(__CUDA_ARCH__)
    return __fadd_rn(a, b);
#else
    return a + b;
#endif
}
FX_HD float fsub(float a, float b) {
#if defined(__CUDA_ARCH__)
    return __fsub_rn(a, b);
#else
    return a - b;
#endif
}
FX_HD cf mk(float x, float y) { cf r; r.x = x; r.y = y; return r; }

// sm_100a packed FP32: one FADD2 / FMUL2 instruction performs the two IEEE round-to-nearest
// operations of a complex add / scale (same results as two scalar ops, half the issue slots).
// Rule kept throughout: a packed MULTIPLY never feeds a packed ADD -- ptxas 12.9 contracts
// mul.rn.f32x2 + add.rn.f32x2 into FFMA2 even with --fmad=false, which would round once
// instead of twice.  Products are therefore combined with SCALAR adds (never contracted), and
// the Makefile's sass-check target fails the build if any FFMA2 shows up in the library.
#if defined(__CUDA_ARCH__)
__device__ __forceinline__ unsigned long long pk2(cf a)
{
    unsigned long long r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a.x), "f"(a.y));
    return r;
}
__device__ __forceinline__ cf upk2(unsigned long long v)
{
    cf r;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(r.x), "=f"(r.y) : "l"(v));
    return r;
}
#endif
FX_HD cf cadd(cf a, cf b)
{
#if defined(__CUDA_ARCH__)
    unsigned long long r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(pk2(a)), "l"(pk2(b)));
    return upk2(r);
#else
    return mk(fadd(a.x, b.x), fadd(a.y, b.y));
#endif
}
FX_HD cf csub(cf a, cf b)
{
#if defined(__CUDA_ARCH__)
    unsigned long long r;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(pk2(a)), "l"(pk2(b)));
    return upk2(r);
#else
    return mk(fsub(a.x, b.x), fsub(a.y, b.y));
#endif
}
// (a.x*s, a.y*s); on the device the result must feed a multiply or a scalar add, never cadd/csub
FX_HD cf cscale2(cf a, float s)
{
#if defined(__CUDA_ARCH__)
    unsigned long long r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(pk2(a)), "l"(pk2(mk(s, s))));
    return upk2(r);
#else
    return mk(fmul(a.x, s), fmul(a.y, s));
#endif
}
// std::complex<float> operator* for finite operands: (ac - bd, ad + bc).  Device: four products
// by two FMUL2, combined by two scalar adds (bc + ad == ad + bc bit for bit).
FX_HD cf cmul(cf a, cf b)
{
    const cf p1 = cscale2(a, b.x);      // (ac, bc)
    const cf p2 = cscale2(a, b.y);      // (ad, bd)
    return mk(fsub(p1.x, p2.y), fadd(p1.y, p2.x));
}
// |z|^2 = re*re + im*im (LoRaDetector.hpp:48)
FX_HD float cnorm(cf a)
{
#if defined(__CUDA_ARCH__)
    unsigned long long r;
    asm("mul.rn.f32x2 %0, %1, %1;" : "=l"(r) : "l"(pk2(a)));
    const cf p = upk2(r);
    return fadd(p.x, p.y);
#else
    return fadd(fmul(a.x, a.x), fmul(a.y, a.y));
#endif
}
// scalar scale: safe in front of cadd/csub (see the rule above)
FX_HD cf cscale(cf a, float s) { return mk(fmul(a.x, s), fmul(a.y, s)); }

// kf_bfly4 (kissfft.hh:164-185), forward transform, one k.  f0..f3 = F[k], F[k+m], F[k+2m], F[k+3m].
// TW=false: the three twiddles are tw[0] == (1, 0) and the products are skipped.  That only
// changes the sign of exact zeros (and NaN propagation from infinities), neither of which can
// alter |X|^2 comparisons; the estimator build (TW products kept) needs the zero signs for atan2f.
template <bool TW>
FX_HD void bfly4(cf& f0, cf& f1, cf& f2, cf& f3, cf t1, cf t2, cf t3)
{
    cf s0 = TW ? cmul(f1, t1) : f1;
    cf s1 = TW ? cmul(f2, t2) : f2;
    cf s2 = TW ? cmul(f3, t3) : f3;
    cf s5 = csub(f0, s1);
    f0 = cadd(f0, s1);
    cf s3 = cadd(s0, s2);
    cf s4 = csub(s0, s2);
    s4 = mk(s4.y, -s4.x);
    f2 = csub(f0, s3);
    f0 = cadd(f0, s3);
    f1 = cadd(s5, s4);
    f3 = csub(s5, s4);
}
// kf_bfly2 (kissfft.hh:155-162)
template <bool TW>
FX_HD void bfly2(cf& f0, cf& f1, cf t)
{
    cf p = TW ? cmul(f1, t) : f1;
    f1 = csub(f0, p);
    f0 = cadd(f0, p);
}

template <int SF>
struct Plan {
    static constexpr int N = 1 << SF;
    static constexpr int R = N >= 16 ? 16 : N;          // complex values per thread
    static constexpr int LOG_R = N >= 16 ? 4 : SF;
    static constexpr int T = N / R;                     // threads per transform
    static constexpr int LOG_T = SF - LOG_R;
    static constexpr int NP = (SF + 3) / 4;             // register passes
    static constexpr int W0 = SF - 4 * (NP - 1);        // position bits resolved by pass 0
    static constexpr int SETS0 = R >> W0;               // independent butterfly sets in pass 0
    static constexpr int STAGES = (SF + 1) / 2;         // kissfft stages: radix 4 ..., trailing 2 if SF odd
    static constexpr bool ODD = (SF & 1) != 0;

    // entries of the compact twiddle table (passes >= 1): 15 * m_a each
    FX_MHD static constexpr int compact_size() {
        int n = 0;
        for (int p = 1; p < NP; ++p) n += 15 << (W0 + 4 * (p - 1));
        return n > 0 ? n : 1;
    }
    FX_MHD static constexpr int compact_offset(int pass) {
        int n = 0;
        for (int p = 1; p < pass; ++p) n += 15 << (W0 + 4 * (p - 1));
        return n;
    }

    // input sample index -> position in kissfft's working array after the leaf copy
    // (kissfft.hh:112-131): digit j of the input index (radix p_j, weight prod_{l<j} p_l)
    // lands at weight m_j = N / prod_{l<=j} p_l.
    FX_MHD static int pos_from_input(int i)
    {
        int o = 0;
#pragma unroll
        for (int j = 0; j < STAGES; ++j) {
            const bool last2 = ODD && j == STAGES - 1;
            const int q = last2 ? ((i >> (2 * j)) & 1) : ((i >> (2 * j)) & 3);
            const int shift = last2 ? 0 : (SF - 2 * (j + 1));
            o |= q << shift;
        }
        return o;
    }

    // position held by register e of thread tid during pass p >= 1
    FX_MHD static int pos_in_pass(int p, int tid, int e)
    {
        const int lo = W0 + 4 * (p - 1);
        return ((tid >> lo) << (lo + 4)) | (e << lo) | (tid & ((1 << lo) - 1));
    }

    // Shared-memory placement of position o inside one transform's exchange buffer.  The
    // XOR folds the top four position bits (the ones that differ between the lanes of a
    // half-warp when pass 0 scatters its digit-reversed results) into the low four bits so
    // that 64-bit stores/loads of all passes are bank-conflict free.
    FX_MHD static int phys(int o)
    {
        // SF10: the pass-1 exchange varies position bits {0,1} and {6,7} across a half-warp, and the single fold left
        // it 4-way conflicted (3.8 G conflict cycles per cfg-sized step); a second field makes every exchange
        // conflict-free (exhaustive search over two-field XOR swizzles: tools/fft_swizzle_search.py).  SF9 / SF11 have
        // conflict-free two-field swizzles too, but their 2-way conflicts cost less than the extra address
        // arithmetic in these issue-bound kernels (measured: SF9 33.7 -> 36.8 ms with it), so they keep one field.
        if constexpr (SF == 10) return o ^ ((o >> 4) & 15) ^ ((o >> 6) & 15);
        else if constexpr (SF >= 8) return o ^ ((o >> (SF - 4)) & 15);
        else if constexpr (SF == 7) return o ^ ((o >> 4) & 7);
        else return o;
    }

    // ---- pass 0: v[c] = sample (tid + T*c); resolves position bits [0, W0) ------------
    template <bool FULL>
    FX_MHD static void pass0(cf (&v)[R], const cf* __restrict__ tw)
    {
        // FULL keeps the reference's multiply by tw[0] (whose imaginary part is -0.0f)
        const cf one = FULL ? tw[0] : mk(1.0f, 0.0f);
        if constexpr (W0 == 1) {
#pragma unroll
            for (int s = 0; s < SETS0; ++s) bfly2<FULL>(v[s], v[s + SETS0], one);
        } else if constexpr (W0 == 2) {
#pragma unroll
            for (int s = 0; s < SETS0; ++s)
                bfly4<FULL>(v[s], v[s + SETS0], v[s + 2 * SETS0], v[s + 3 * SETS0], one, one, one);
        } else if constexpr (W0 == 3) {
            // c = c0 + SETS0*a + 4*SETS0*b ; radix-2 over b (m = 1), then radix-4 over a (m = 2, k = b)
#pragma unroll
            for (int s = 0; s < SETS0; ++s)
#pragma unroll
                for (int a = 0; a < 4; ++a) bfly2<FULL>(v[s + SETS0 * a], v[s + SETS0 * a + 4 * SETS0], one);
            const cf t1 = tw[N / 8], t2 = tw[2 * (N / 8)], t3 = tw[3 * (N / 8)];
#pragma unroll
            for (int s = 0; s < SETS0; ++s) {
                bfly4<FULL>(v[s], v[s + SETS0], v[s + 2 * SETS0], v[s + 3 * SETS0], one, one, one);
                bfly4<true>(v[s + 4 * SETS0], v[s + 5 * SETS0], v[s + 6 * SETS0], v[s + 7 * SETS0], t1, t2, t3);
            }
        } else {
            // W0 == 4: c = a + 4b ; radix-4 over b (m = 1, k = 0), then radix-4 over a (m = 4, k = b)
#pragma unroll
            for (int a = 0; a < 4; ++a) bfly4<FULL>(v[a], v[a + 4], v[a + 8], v[a + 12], one, one, one);
            bfly4<FULL>(v[0], v[1], v[2], v[3], one, one, one);
#pragma unroll
            for (int b = 1; b < 4; ++b) {
                const cf t1 = tw[b * (N / 16)], t2 = tw[2 * b * (N / 16)], t3 = tw[3 * b * (N / 16)];
                bfly4<true>(v[4 * b], v[4 * b + 1], v[4 * b + 2], v[4 * b + 3], t1, t2, t3);
            }
        }
    }

    // position of register c after pass 0
    FX_MHD static int pos_after_pass0(int tid, int c) { return pos_from_input(tid + T * c); }

    // spectrum bin held by register r of thread tid once all passes are done
    FX_MHD static int bin_of_reg(int tid, int r)
    {
        return NP == 1 ? pos_after_pass0(tid, r) : pos_in_pass(NP - 1, tid, r);
    }

    // ---- pass p >= 1: two fused radix-4 stages, v[e], e = c1 + 4*c2 ----------------------
    // ct = compact twiddle table of this SF (see build_compact()).
    FX_MHD static void pass(int p, cf (&v)[R], int tid, const cf* __restrict__ ct)
    {
        const int lo = W0 + 4 * (p - 1);
        const int ma = 1 << lo;
        const int k = tid & (ma - 1);
        const cf* A = ct + compact_offset(p);
        const cf* B = A + 3 * ma;
        const cf a1 = A[k], a2 = A[ma + k], a3 = A[2 * ma + k];
#pragma unroll
        for (int c2 = 0; c2 < 4; ++c2)
            bfly4<true>(v[4 * c2], v[4 * c2 + 1], v[4 * c2 + 2], v[4 * c2 + 3], a1, a2, a3);
#pragma unroll
        for (int c1 = 0; c1 < 4; ++c1) {
            const int kk = k + ma * c1;
            const cf b1 = B[kk], b2 = B[4 * ma + kk], b3 = B[8 * ma + kk];
            bfly4<true>(v[c1], v[c1 + 4], v[c1 + 8], v[c1 + 12], b1, b2, b3);
        }
    }

    // Fill the compact table from the exact kissfft table tw[N] (host or device, any thread split).
    // entry j of [0, compact_size())
    FX_MHD static cf compact_entry(int j, const cf* tw)
    {
        for (int p = 1; p < NP; ++p) {
            const int lo = W0 + 4 * (p - 1);
            const int ma = 1 << lo;
            const int off = compact_offset(p);
            if (j < off + 15 * ma) {
                int r = j - off;
                if (r < 3 * ma) {                          // A[c1-1][k] = tw[k * fs_a * c1]
                    const int c1 = r / ma + 1, k = r % ma;
                    return tw[k * (N >> (lo + 2)) * c1];
                }
                r -= 3 * ma;                               // B[c2-1][k'] = tw[k' * fs_b * c2]
                const int c2 = r / (4 * ma) + 1, kk = r % (4 * ma);
                return tw[kk * (N >> (lo + 4)) * c2];
            }
        }
        return mk(1.0f, 0.0f);
    }
};

} // namespace fx
