// stream_common.cuh -- pieces shared by the streaming kernels (demod_cert.cu, demod_lean.cu): the exact transform on a
// shared-memory exchange buffer, the reference's per-sample chain, window fetches with cp.async.bulk.
#pragma once
#include "demod_device.cuh"
#include "stream_ptx.cuh"

namespace lk {

#ifndef STREAM_CTAS_PER_SM
#define STREAM_CTAS_PER_SM 3      // 24 warps per SM at <= 85 registers; 3 x ~72 KB of stages fit the 227 KB of shared memory
#endif

struct __align__(16) WinDesc {       // written by the window's leader at prefetch time
    float scale, rate, start;
    uint32_t flags;                   // WD_*
    unsigned long long out_off;       // index into a.symbols (data windows) / packet id (scan windows)
    uint32_t aux0, aux1;              // unit-type specific
};
enum { WD_ACTIVE = 1, WD_ODD = 2, WD_SLOW = 4, WD_DIRECT = 8, WD_DEFERRED = 16 };

template <int SF>
struct StreamCfg {
    using P = fx::Plan<SF>;
    static constexpr int T = P::T, N = P::N;
    static constexpr int DOM_THREADS = T <= 32 ? 32 : CTA_THREADS;   // threads sharing one pipeline
    static constexpr int DOMS = CTA_THREADS / DOM_THREADS;           // pipelines per CTA
    static constexpr int GD = DOM_THREADS / T;                       // windows per pipeline step
    static constexpr int PITCH = N + 8;     // +1 entry for the odd-aligned tail sample, +8 keeps 64-byte
                                            // alignment and shifts neighbouring SF7 windows by 16 banks
    static constexpr int SLOT = GD * PITCH; // cf entries per stage
    static constexpr int STAGES = 2;
    static constexpr size_t slot_bytes = (size_t)DOMS * STAGES * SLOT * sizeof(cf);
    static constexpr size_t desc_bytes = (size_t)DOMS * STAGES * GD * sizeof(WinDesc);
    static constexpr size_t bar_bytes = (size_t)DOMS * STAGES * sizeof(uint64_t);
    static constexpr size_t bytes = slot_bytes + desc_bytes + bar_bytes + 8 * sizeof(Peak);
};

template <int SF>
__device__ __forceinline__ void domain_sync()
{
    if (StreamCfg<SF>::T <= 32) __syncwarp(); else __syncthreads();
}

// exact-order FFT with every exchange done in the window's own slot (N entries at `ex`)
template <int SF, bool FULL>
__device__ __forceinline__ void fft_inplace(cf (&v)[fx::Plan<SF>::R], int tid, const cf* __restrict__ tw,
                                            const cf* __restrict__ ct, cf* ex)
{
    using P = fx::Plan<SF>;
    P::template pass0<FULL>(v, tw);
#pragma unroll
    for (int p = 1; p < P::NP; ++p) {
        group_sync<P::T>();           // raw samples (p == 1) / previous exchange (p > 1) fully consumed; the slot is
                                      // the window's own, so only its thread group meets (named barrier per group)
#pragma unroll
        for (int r = 0; r < P::R; ++r) {
            const int o = (p == 1) ? P::pos_after_pass0(tid, r) : P::pos_in_pass(p - 1, tid, r);
            ex[P::phys(o)] = v[r];
        }
        group_sync<P::T>();
#pragma unroll
        for (int r = 0; r < P::R; ++r) v[r] = ex[P::phys(P::pos_in_pass(p, tid, r))];
        P::pass(p, v, tid, ct);
    }
}

// rare path: some |phase| of the window is below 2^-12 or reaches 120 rad (glibc's other branches)
static __device__ __noinline__ float2 sincos_slow(float ph)
{
    float sn, cs;
    gm::sincosf_g(ph, &sn, &cs);
    return make_float2(cs, sn);
}

// Window classification for the fast sincos core: phases run monotonically from `start` (i = 0)
// to `last` (i = N-1); the core needs 2^-12 <= |ph| < 120 for every sample.
__device__ __forceinline__ bool phases_need_slow_path(float start, float last)
{
    const float lo = fminf(fabsf(start), fabsf(last)), hi = fmaxf(fabsf(start), fabsf(last));
    const bool same_sign = (__float_as_uint(start) ^ __float_as_uint(last)) < 0x80000000u;
    return !(same_sign && lo >= 0x1p-11f && hi < 119.0f);
}

// Per-sample chain of one data / sync window held in shared memory (LoRaDemod.cpp:154-163,
// phy.cpp:220-231): normalise -> [downchirp] -> CFO rotation -> [Hann], into pass-0 register order.
template <int SF, bool EXTRA, bool TRACK = false>
__device__ __forceinline__ void chain_window(cf (&v)[fx::Plan<SF>::R], const cf* raw /* window + tid */, int tid,
                                             float scale, float rate, float start, bool slow,
                                             const cf* __restrict__ down, const float* __restrict__ w,
                                             float* track = nullptr)
{
    using P = fx::Plan<SF>;
    constexpr int T = P::T;
    const float fi0 = (float)tid;
    if (!slow) {
        float m = TRACK ? *track : 0.0f;
#pragma unroll
        for (int c = 0; c < P::R; ++c) {
            const cf r = raw[T * c];
            // running max(|re|,|im|) of the raw samples (LoRaDemod.cpp:61-68) for finite data: one FMNMX3.
            // NaN samples are ignored here; the caller detects them through the poisoned spectrum.
            if (TRACK) m = fmaxf(fmaxf(m, fabsf(r.x)), fabsf(r.y));
            cf x = fx::cscale2(r, scale);                                            // LoRaDemod.cpp:75-77
            if (EXTRA && down) x = fx::cmul(x, ldg_cf(down + tid + T * c));          // phy.cpp:224
            const float ph = fx::fadd(start, fx::fmul(rate, fx::fadd(fi0, (float)(T * c))));
            float sn, cs;
#if defined(LORA_ABLATE_SINCOS)
            // measurement-only build (tools/ablate.sh, never the product): a 3-instruction stand-in for the exact
            // sincosf, to time the chain without it
            sn = ph; cs = fx::fsub(1.0f, fx::fmul(ph, ph));
#else
            gm::sincosf_core(ph, &sn, &cs);                                          // LoRaDemod.cpp:155-159
#endif
            x = fx::cmul(x, fx::mk(cs, sn));
            if (EXTRA && w) x = fx::cscale(x, __ldg(w + tid + T * c));               // LoRaDemod.cpp:160-161
            v[c] = x;
        }
        if (TRACK) *track = m;
    } else {
        float m = TRACK ? *track : 0.0f;
#pragma unroll
        for (int c = 0; c < P::R; ++c) {
            const cf r = raw[T * c];
            if (TRACK) m = fmaxf(fmaxf(m, fabsf(r.x)), fabsf(r.y));
            cf x = fx::cscale(r, scale);
            if (EXTRA && down) x = fx::cmul(x, ldg_cf(down + tid + T * c));
            const float ph = fx::fadd(start, fx::fmul(rate, fx::fadd(fi0, (float)(T * c))));
            const float2 q = sincos_slow(ph);
            x = fx::cmul(x, fx::mk(q.x, q.y));
            if (EXTRA && w) x = fx::cscale(x, __ldg(w + tid + T * c));
            v[c] = x;
        }
        if (TRACK) *track = m;
    }
}

// window base of symbol s with the guarded time-offset shift (LoRaDemod.cpp:143-150), osr == 1
__device__ __forceinline__ uint32_t shifted_base(uint32_t s, uint32_t N, int t_off, uint32_t count)
{
    uint32_t base = s * N;
    if (t_off > 0) {
        if ((unsigned long long)base + (unsigned)t_off + N <= count) base += (uint32_t)t_off;
    } else if (t_off < 0) {
        const uint32_t off = (uint32_t)(-(long long)t_off);
        if (off <= base) base -= off;
    }
    return base;
}

// start phase of symbol s (LoRaDemod.cpp:152-153), osr == 1
__device__ __forceinline__ float start_phase(float rate, uint32_t s, uint32_t N, int t_off)
{
    return fx::fmul(rate, fx::fadd(__uint2float_rn(s * N), __int2float_rn(t_off)));
}

__device__ __forceinline__ bool window_is_odd(const cf* src)      // starts 8 bytes after a 16-byte boundary
{
    return ((reinterpret_cast<uintptr_t>(src) >> 3) & 1u) != 0;
}

// Leader-side fetch of one full window into its stage slot.  An even window (16-byte aligned) is one bulk copy
// to win[0..N).  An odd window (8 bytes past a 16-byte boundary) lands one entry later -- sample i at win[i + 1],
// readers start at win + 1: its aligned interior (samples 1..N-2) is the bulk copy, the two end samples are plain
// loads, so nothing outside the window is ever read (a batch may start 8 bytes into an allocation).  The caller
// publishes its descriptor BEFORE calling: the mbarrier arrive releases it to the waiters.
template <int N>
__device__ __forceinline__ void fetch_window(cf* win, const cf* src, uint64_t* bar)
{
    const uintptr_t addr = reinterpret_cast<uintptr_t>(src);
    const unsigned odd = (unsigned)(addr >> 3) & 1u;
    if (odd) {
        win[1] = ldg_cf(src);
        win[N] = ldg_cf(src + (N - 1));
        ptx::mbar_arrive_expect_tx(bar, (unsigned)((N - 2) * sizeof(cf)));
        ptx::bulk_g2s(win + 2, src + 1, (unsigned)((N - 2) * sizeof(cf)), bar);
    } else {
        ptx::mbar_arrive_expect_tx(bar, (unsigned)(N * sizeof(cf)));
        ptx::bulk_g2s(win, src, (unsigned)(N * sizeof(cf)), bar);
    }
}

// fetch_window on 32-bit shared addresses (window slot, mbarrier)
template <int N>
__device__ __forceinline__ void fetch_window_sa(uint32_t win, const cf* src, uint32_t bar)
{
    const uintptr_t addr = reinterpret_cast<uintptr_t>(src);
    const unsigned odd = (unsigned)(addr >> 3) & 1u;
    if (odd) {
        const cf a = ldg_cf(src), b = ldg_cf(src + (N - 1));
        ptx::sts_cf_sa(win + 8u, a.x, a.y);
        ptx::sts_cf_sa(win + 8u * N, b.x, b.y);
        ptx::mbar_arrive_expect_tx_sa(bar, (unsigned)((N - 2) * sizeof(cf)));
        ptx::bulk_g2s_sa(win + 16u, src + 1, (unsigned)((N - 2) * sizeof(cf)), bar);
    } else {
        ptx::mbar_arrive_expect_tx_sa(bar, (unsigned)(N * sizeof(cf)));
        ptx::bulk_g2s_sa(win, src, (unsigned)(N * sizeof(cf)), bar);
    }
}

} // namespace lk
