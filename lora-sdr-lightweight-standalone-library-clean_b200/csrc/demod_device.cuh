// demod_device.cuh -- device building blocks of the demodulator kernels (K1, K2a).
//
// One "group" of T = N/16 threads owns one N-point window at a time: it loads the window
// straight from global memory into registers (coalesced), applies the reference's per-sample
// chain (normalise -> [downchirp] -> CFO rotation -> [Hann]), runs the exact-order FFT of
// fft_exact.cuh and reduces |X|^2 to the first-max bin.  Reference loops restated here:
//   LoRaDemod.cpp:142-176 / phy.cpp:204-238  (per-symbol data loop)
//   LoRaDemod.cpp:86-124 / phy.cpp:97-134    (estimator inner loops)
//   LoRaDetector.hpp:39-74                   (detect)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "fft_exact.cuh"
#include "glibc_math.cuh"

namespace lk {

using fx::cf;

constexpr int CTA_THREADS = 256;

struct PacketState {   // written by the prepass kernel, read by the symbol kernel
    float scale;       // 1 / max_amp when the packet is normalised (LoRaDemod.cpp:74)
    float rate;        // -2*pi*cfo/N (LoRaDemod.cpp:139)
    int   t_off;       // round(time_offset) (LoRaDemod.cpp:138)
    int   flags;       // PS_*
};
enum { PS_NORM = 1, PS_SKIP = 2 };

enum { MODE_LEGACY = 0, MODE_HIGHLEVEL = 1, MODE_EST_ONLY = 2 };

struct DemodArgs {
    const cf* iq;
    size_t packet_stride;
    const uint32_t* sample_counts;
    uint32_t samples_per_packet;
    uint32_t n_packets;
    uint32_t osr;
    int mode;
    int no_scratch;
    uint32_t est_uniform;         // uniform trip count of the estimator symbol loop
    uint32_t data_symbols_max;    // symbol-kernel items per packet
    unsigned long long dsm_magic; // floor(2^64 / data_symbols_max) + 1: item / dsm == umul64hi(item, magic) for item < 2^32
    const float* window;          // nullptr = none
    const cf* down;               // nullptr unless MODE_HIGHLEVEL
    const cf* tw;
    const cf* ct;
    float power_scale;
    PacketState* state;
    uint16_t* symbols;
    size_t symbol_stride;
    uint8_t* sync;
    float* cfo;
    float* time_offset;
    int32_t* status;
    int no_prune;                 // certified kernel: the copy without the pruned transform (noisy batches, api.cu)
    int unit_scale;               // prepass: assume no 1/max normalisation is needed (no max-abs scan); the lean pass checks
    int skip_sync;                // prepass: leave the two sync symbols to a later pass
    uint8_t* flagged;             // lean pass (demod_lean.cu): packets it hands to the certified kernel
    const uint8_t* only_flagged;  // certified kernel: skip packets whose byte is 0 (nullptr = every packet); 2 = the lean pass
                                  // settled every symbol but the packet's max-abs (pmax) exceeds 1: redo the estimator with the
                                  // true scale, and keep the lean symbols if rate / t_off stay within the certified tolerance
    float* pmax;                  // lean pass: max-abs of the packets it marks 2
    unsigned* n_flagged;          // lean pass: number of packets it flagged (the certified pass returns at once on 0)
    int force_exact;              // LORA_PHY_CUDA_F_EXACT: the certified kernel replays the exact chain for every window
    unsigned long long* cert_stats;   // [8] windows walked, windows replayed exactly, second estimates, re-walked packets, windows settled by the pruned transform
};

__device__ __forceinline__ cf ldg_cf(const cf* p)
{
    const float2 t = __ldg(reinterpret_cast<const float2*>(p));
    return fx::mk(t.x, t.y);
}

template <int T>
__device__ __forceinline__ void group_sync()
{
    // a group's windows, exchange slots and reduction cells are its own: only its T threads meet.
    // Sub-warp groups share a warp barrier; a group of several warps gets its own named barrier
    // (ids 1..CTA_THREADS/T; id 0 stays __syncthreads), so the groups of a CTA do not wait on each other.
    if (T <= 32) __syncwarp();
    else if (T >= CTA_THREADS) __syncthreads();
    else asm volatile("bar.sync %0, %1;" :: "r"(1 + (int)threadIdx.x / T), "n"(T) : "memory");
}

// ---- per-sample chain + load into pass-0 register order ---------------------------------------
template <int SF>
__device__ __forceinline__ void load_window(cf (&v)[fx::Plan<SF>::R], const cf* __restrict__ src, int tid,
                                            unsigned osr, bool active, bool norm, float scale,
                                            const cf* __restrict__ down, bool rotate, float start, float rate,
                                            const float* __restrict__ w)
{
    using P = fx::Plan<SF>;
#pragma unroll
    for (int c = 0; c < P::R; ++c) {
        const int i = tid + P::T * c;
        cf x = fx::mk(0.0f, 0.0f);
        if (active) {
            x = ldg_cf(src + (size_t)i * osr);
            if (norm) x = fx::cscale(x, scale);                       // LoRaDemod.cpp:75-77
            if (down) x = fx::cmul(x, ldg_cf(down + i));              // phy.cpp:224
            if (rotate) {                                             // LoRaDemod.cpp:155-159
                const float ph = fx::fadd(start, fx::fmul(rate, (float)i));
                float sn, cs;
                gm::sincosf_g(ph, &sn, &cs);
                x = fx::cmul(x, fx::mk(cs, sn));
            }
            if (w) x = fx::cscale(x, __ldg(w + i));                   // LoRaDemod.cpp:160-161
        }
        v[c] = x;
    }
}

// ---- exact-order FFT of one group --------------------------------------------------------------
// buf0/buf1: the group's two N-entry exchange buffers; flip alternates between them so that a
// buffer is only rewritten after a barrier that follows its last read.
template <int SF, bool FULL>
__device__ __forceinline__ void fft_group(cf (&v)[fx::Plan<SF>::R], int tid, const cf* __restrict__ tw,
                                          const cf* __restrict__ ct, cf* buf0, cf* buf1, int& flip)
{
    using P = fx::Plan<SF>;
    P::template pass0<FULL>(v, tw);
#pragma unroll
    for (int p = 1; p < P::NP; ++p) {
        cf* dst = flip ? buf1 : buf0;
        flip ^= 1;
#pragma unroll
        for (int r = 0; r < P::R; ++r) {
            const int o = (p == 1) ? P::pos_after_pass0(tid, r) : P::pos_in_pass(p - 1, tid, r);
            dst[P::phys(o)] = v[r];
        }
        group_sync<P::T>();
#pragma unroll
        for (int r = 0; r < P::R; ++r) v[r] = dst[P::phys(P::pos_in_pass(p, tid, r))];
        P::pass(p, v, tid, ct);
    }
}

// ---- |X|^2 first-max (LoRaDetector.hpp:46-58) ---------------------------------------------------
struct Peak { float val; int idx; };

__device__ __forceinline__ Peak peak_merge(Peak a, Peak b)
{
    // larger power wins; on equal power the lower bin wins (strict '>' in an ascending scan)
    return (b.val > a.val || (b.val == a.val && b.idx < a.idx)) ? b : a;
}

template <int SF>
__device__ __forceinline__ Peak local_peak(const cf (&v)[fx::Plan<SF>::R], int tid)
{
    using P = fx::Plan<SF>;
    Peak pk; pk.val = 0.0f; pk.idx = 0;
#pragma unroll
    for (int r = 0; r < P::R; ++r) {
        const float m = fx::cnorm(v[r]);
        const int bin = P::bin_of_reg(tid, r);
        if (P::NP > 1) {                       // registers hold ascending bins
            if (m > pk.val) { pk.val = m; pk.idx = bin; }
        } else {
            if (m > pk.val || (m == pk.val && bin < pk.idx)) { pk.val = m; pk.idx = bin; }
        }
    }
    return pk;
}

// All threads of the group return the group's peak.  red: CTA_THREADS/32 Peak slots in smem.
template <int SF>
__device__ __forceinline__ Peak group_peak(Peak pk, Peak* red)
{
    using P = fx::Plan<SF>;
    constexpr int T = P::T;
    constexpr int W = T < 32 ? T : 32;
#pragma unroll
    for (int off = W / 2; off > 0; off >>= 1) {
        Peak o;
        o.val = __shfl_xor_sync(0xffffffffu, pk.val, off);
        o.idx = __shfl_xor_sync(0xffffffffu, pk.idx, off);
        pk = peak_merge(pk, o);
    }
    if (T > 32) {
        const int warp = threadIdx.x >> 5;
        if ((threadIdx.x & 31) == 0) red[warp] = pk;
        group_sync<T>();
        const int w0 = (threadIdx.x / T) * (T / 32);
        pk = red[w0];
#pragma unroll
        for (int k = 1; k < T / 32; ++k) pk = peak_merge(pk, red[w0 + k]);
    }
    return pk;
}

template <int T>
__device__ __forceinline__ float group_max(float m, float* redf)
{
    constexpr int W = T < 32 ? T : 32;
#pragma unroll
    for (int off = W / 2; off > 0; off >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, off));
    if (T > 32) {
        const int warp = threadIdx.x >> 5;
        if ((threadIdx.x & 31) == 0) redf[warp] = m;
        group_sync<T>();
        const int w0 = (threadIdx.x / T) * (T / 32);
        m = redf[w0];
#pragma unroll
        for (int k = 1; k < T / 32; ++k) m = fmaxf(m, redf[w0 + k]);
        group_sync<T>();             // the group's own redf cells may be rewritten
    }
    return m;
}

// value held by the group's thread 0, delivered to every thread of the group
template <int T>
__device__ __forceinline__ float group_bcast(float x, int tid, float* slot /* one per group, smem */)
{
    if (T <= 32) {
        const int lane = threadIdx.x & 31;
        return __shfl_sync(0xffffffffu, x, lane - tid);
    }
    __syncthreads();
    if (tid == 0) *slot = x;
    __syncthreads();
    return *slot;
}

// ---- one data / sync symbol: LoRaDemod.cpp:142-176, phy.cpp:204-238 -----------------------------
template <int SF, bool FULL>
__device__ __forceinline__ int demod_symbol(const DemodArgs& a, const cf* __restrict__ pkt, uint32_t count,
                                            uint32_t s, bool active, const PacketState& st, int tid,
                                            cf* buf0, cf* buf1, int& flip, Peak* red)
{
    using P = fx::Plan<SF>;
    const uint32_t step = (uint32_t)P::N * a.osr;
    size_t base = (size_t)s * step;
    const int t_off = st.t_off;
    if (t_off > 0) {
        if (base + (size_t)t_off + step <= count) base += (size_t)t_off;
    } else if (t_off < 0) {
        const size_t off = (size_t)(-(long long)t_off);
        if (off <= base) base -= off;
    }
    const float start = fx::fmul(st.rate, fx::fadd(__ull2float_rn((unsigned long long)s * P::N),
                                                   __fdiv_rn(__int2float_rn(t_off), __uint2float_rn(a.osr))));
    const bool rotate = st.rate != 0.0f;   // cos(0)=1, sin(0)=0: the product is the identity
    cf v[P::R];
    load_window<SF>(v, pkt + base, tid, a.osr, active, (st.flags & PS_NORM) != 0, st.scale, a.down,
                    rotate, start, st.rate, a.window);
    fft_group<SF, FULL>(v, tid, a.tw, a.ct, buf0, buf1, flip);
    const Peak pk = group_peak<SF>(local_peak<SF>(v, tid), red);
    return pk.idx;
}

} // namespace lk
