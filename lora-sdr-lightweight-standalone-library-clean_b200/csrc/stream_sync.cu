// stream_sync.cu -- K2, the sliding-window preamble / sync-word detector on continuous IQ:
// external/LoRa-SDR/LoRaDemod.cpp:145-327 (work loop and frame-sync state machine) over the tables of
// :96-117, with the detector of LoRaDetector.hpp:39-74.
//
// A stream is inherently sequential (how far the window slides depends on the bin it just detected), so
// the parallelism is streams x the T = N/16 threads of one window: one thread group owns one stream and
// walks it window by window -- dechirp with the current chirp table and the 128-step fine-tune table,
// exact-order FFT (fft_exact.cuh, full twiddles: the spectrum feeds fIndex), first-max, power / powerAvg /
// fIndex, then the state machine on the group's thread 0:
//   FRAMESYNC   preamble tracking: slide by N - value, accumulate fIndex into the fine frequency error;
//               on (prev+4)/8 == 0 and sync word0 matched, look one window ahead for sync word1
//   DOWNCHIRP0/1 signed bin error of the two down-chirps, averaged (the CFO estimate reported per frame)
//   QUARTERCHIRP consume N/4 + err/2 (the STO correction), fold err/2 into the fine frequency error
//   DATASYMBOLS emit symbols until mtu or squelch (snr < threshold), report the frame
// Two pieces of the reference are serial by construction, and their order is observable in the results: the fine-tune
// index recurrence (an int decremented by a float and truncated, per sample, :160-163) and the detector's
// double-precision running total (LoRaDetector.hpp:45-52).  Both run in parallel under a checked proof (sync_window)
// and fall back to the literal loop on thread 0 when it fails; LORA_PHY_STREAM_SERIAL=1 forces the literal loops.
//
// Streams do not wait for each other: every barrier and vote of the walk spans one DOMAIN only -- the warp for
// windows of up to 32 threads (SF7-9; the 4 / 2 / 1 streams of a warp step together), the stream's own thread
// group (named barrier) above.  A stream that looks ahead for sync word1, or falls back to a serial loop, delays the
// streams of its warp, not the 32 streams of its CTA.  The next window always starts inside the 3N chips after the
// current window's start, so each step asks L2 for the lines up to 4N chips ahead (prefetch.global.L2): the window
// load that follows a slide decision finds its samples on chip.
#include <type_traits>
#include "demod_device.cuh"
#include "internal.h"

namespace lk {

namespace {

constexpr int FINE_STEPS = 128;
// Resident threads per SM the register budget is cut for.  Multi-warp windows (SF10-12) gain 3-6 % from 768 (80
// registers, a few spilled words, 3 x 64 KB of exchange buffers); sub-warp windows are latency-bound walks of few warps,
// where the tighter budget costs 25 % (same-box A/B, 4 GB batches) -- they keep 128 registers.
template <int SF>
constexpr int sync_threads_per_sm() { return fx::Plan<SF>::T <= 32 ? 512 : 768; }
enum { ST_FRAMESYNC = 0, ST_DOWNCHIRP0, ST_DOWNCHIRP1, ST_QUARTERCHIRP, ST_DATASYMBOLS };

struct GroupCtl {            // written by the group's thread 0, read by the whole group
    uint32_t pos;
    int live, down, need2, ft;
    float fine_err;
    int ok, ft2;
};

// CTA size: multi-warp windows use the CTA_THREADS-wide group helpers; sub-warp windows need no CTA-wide anything,
// so they run in 128-thread CTAs (twice the CTAs for the same streams: a few thousand SF7 streams reach every SM)
template <int SF>
constexpr int sync_cta() { return fx::Plan<SF>::T <= 32 ? 128 : CTA_THREADS; }

template <int SF>
struct SyncSmem {
    using P = fx::Plan<SF>;
    static constexpr int CTA = sync_cta<SF>();
    static constexpr int G = CTA / P::T;
    static constexpr size_t buf_bytes = 2 * (size_t)G * P::N * sizeof(cf);
    static constexpr size_t red_off = buf_bytes;                       // 8 Peaks, then 8 doubles
    static constexpr size_t ctl_off = red_off + 8 * sizeof(Peak) + 8 * sizeof(double);
    static constexpr size_t bytes = ctl_off + (size_t)G * sizeof(GroupCtl);
};

struct DetectOut { int value; float snr, findex; };

// one step of the fine-tune index recurrence (:160-163): int -= float * size_t, truncated, wrapped once
__device__ __forceinline__ int ft_step(int idx, float d, int wrap)
{
    int r = (int)fx::fsub((float)idx, d);
    if (r < 0) r += wrap;
    else if (r >= wrap) r -= wrap;
    return r;
}

// barrier / vote over one domain: the warp (T <= 32) or the stream's own thread group.  For T > 32 the predicate must
// already be uniform over the group (it is read from the group's control block after a barrier); the barrier keeps the
// control block from being rewritten before every thread has read it.
template <int T>
__device__ __forceinline__ bool dom_any(bool p)
{
    if (T <= 32) return __any_sync(0xffffffffu, p) != 0;
    group_sync<T>();
    return p;
}

// sum of the group's doubles (any order: the caller proves the order does not matter)
template <int T>
__device__ __forceinline__ double group_dsum(double x, double* redd)
{
    constexpr int W = T < 32 ? T : 32;
#pragma unroll
    for (int off = W / 2; off > 0; off >>= 1) x += __shfl_xor_sync(0xffffffffu, x, off);
    if (T > 32) {
        const int warp = threadIdx.x >> 5;
        if ((threadIdx.x & 31) == 0) redd[warp] = x;
        group_sync<T>();
        const int w0 = (threadIdx.x / T) * (T / 32);
        x = redd[w0];
#pragma unroll
        for (int k = 1; k < T / 32; ++k) x += redd[w0 + k];
    }
    return x;
}

// Sample m of the stream as the detector sees it.  Without compensation: iq[m].  With per-stream (cfo, time_offset)
// (compensate_offsets, phy.cpp:150-180, fused into the load -- no temporary copy of the stream): the reference first
// rotates sample n by exp(j*rate*n), rate = -2*pi*cfo/(N*osr), then shifts the buffer by round(time_offset) with zero fill,
// so position m holds the rotated sample m - offset, or zero when that falls outside [0, n).
struct StreamView {
    const cf* iq; uint32_t n; float rate; int offset; bool comp;
};
// The compensated sample is a call, not inline code: 64 inlined copies of the double-precision sincosf (16 samples x
// two windows x two load loops) spread the walk's hot path over 150 KB of instructions, and a lone warp per scheduler
// -- all a small group of streams gives the GPU -- then waits on instruction fetches (measured on the cfg-4 groups:
// 3x per window against the plain kernel, without any compensation requested).
__device__ __noinline__ cf stream_sample_comp(const cf* __restrict__ iq, uint32_t n, float rate, int offset, uint32_t m)
{
    const long long k = (long long)m - offset;
    // |offset| >= n: the reference leaves the buffer unshifted (phy.cpp:166,171)
    const long long idx = ((offset > 0 && (uint32_t)offset < n) || (offset < 0 && (uint32_t)(-offset) < n)) ? k : (long long)m;
    if (idx < 0 || idx >= (long long)n) return fx::mk(0.0f, 0.0f);
    const cf x = ldg_cf(iq + idx);
    float sn, cs;
    gm::sincosf_g(fx::fmul(rate, (float)(uint32_t)idx), &sn, &cs);
    return fx::cmul(x, fx::mk(cs, sn));
}
// one window at wpos: dechirp + fine tune (:157-166), FFT, detect (LoRaDetector.hpp:39-74).
// thread 0 of the group returns the detector outputs and the fine-tune index advanced past the window.
//
// The two pieces of the reference that are serial by construction are run in parallel with a proof
// obligation and fall back to the literal serial loop when the proof fails:
//  * fine-tune index: inside one float binade the recurrence is a constant integer step k, so the
//    sequence is idx0 - i*k (mod N*128).  Every thread checks ft_step(c_i) == c_{i+1} for its own
//    samples; if all agree the arithmetic progression IS the recurrence (induction from idx0).  Where the
//    check fails the walk restarts from the literal next value (pieces, below).
//  * detector total: the ascending double sum differs from any other summation order by at most
//    2(N-1)u * total.  If float(total - max) is the same at both ends of that interval (the map is
//    monotone) the order cannot matter and the tree sum is used.
template <int SF, bool EX>
__device__ __forceinline__ DetectOut sync_window(const StreamSyncArgs& a, const StreamView& view, uint32_t wpos, bool live,
                                                 const cf* __restrict__ chirp, int idx0, int* ft_out, float fine_err,
                                                 int tid, cf* buf0, cf* buf1, int& flip, Peak* red,
                                                 double* redd, volatile int* ok_flag)
{
    using P = fx::Plan<SF>;
    constexpr int T = P::T, N = P::N;
    constexpr int wrap = N * FINE_STEPS;
    const bool constant = fine_err == 0.0f;
    const float d = fx::fmul(fine_err, (float)FINE_STEPS);
    // k: the integer step of the first sample, before wrapping
    const int k = idx0 - (int)fx::fsub((float)idx0, d);
    // With |d| <= wrap / 2 the one wrap of ft_step is a reduction mod wrap (a power of two), and the next term of the
    // progression is c_i - k mod wrap: the check per sample is ((int)(float(c_i) - d) - c_i + k) mod wrap == 0.
    // A larger |d| (never seen: it is 128 x a fraction of a bin plus at most N/4 bins) takes the literal loop.
    //
    // Where the check fails the recurrence has left the progression -- in practice once per wrap of the index (the
    // truncation rounds the other way below zero) or at a change of float binade.  Everything up to the first failing
    // sample i* is still the literal sequence, so the walk restarts there: c_{i*+1} = ft_step(c_{i*}) computed literally, a
    // new step k from it, the remaining samples checked against the new progression -- up to SEGMENTS pieces, then
    // the literal loop.  (A carrier offset of a quarter bin wraps the index every eighth window; at SF12 the serial
    // loop costs ten windows' time.)
    constexpr int SEGMENTS = 4;
    constexpr unsigned MASK = (unsigned)(wrap - 1);
    bool serial = a.force_serial || !(fabsf(d) <= (float)(wrap / 2));      // uniform over the group
    const bool check = !constant && live && !serial;
    const unsigned c0 = (unsigned)idx0 - (unsigned)tid * (unsigned)k, cstep = (unsigned)T * (unsigned)k;   // c_i, i = tid + T c
    int bad = 0;
    if (check) {
#pragma unroll
        for (int c = 0; c < P::R; ++c) {
            const int ci = (int)((c0 - (unsigned)c * cstep) & MASK);
            bad |= ((int)fx::fsub((float)ci, d) - ci + k) & (wrap - 1);
        }
    }
    int* const flag = const_cast<int*>(ok_flag);            // the group's cell: first failing sample, N when none
    // first failing sample among this thread's, at or after j0, against the progression (e0 at j0, step kk)
    auto first_fail = [&](int j0, unsigned e0, int kk) {
        int mb = N;
#pragma unroll 1
        for (int c = P::R - 1; c >= 0; --c) {
            const int i = tid + T * c;
            if (i < j0) break;
            const int ci = (int)((e0 - (unsigned)(i - j0) * (unsigned)kk) & MASK);
            if (((int)fx::fsub((float)ci, d) - ci + kk) & (wrap - 1)) mb = i;
        }
        return mb;
    };
    auto group_min = [&](int mine) {
        if (tid == 0) *flag = N;
        group_sync<T>();
        if (mine < N) atomicMin(flag, mine);
        group_sync<T>();
        return *ok_flag;
    };
    auto vote = [&](bool p) { return T > 32 ? p : (__any_sync(0xffffffffu, p) != 0); };   // T > 32: p is uniform over the group
    int first_bad = group_min(bad ? first_fail(0, (unsigned)idx0, k) : N);
    // the literal indices live in the exchange buffer the transform writes SECOND (after its first barrier, which every
    // thread reaches with its samples loaded), so they need no memory of their own
    int* ftidx = reinterpret_cast<int*>(flip ? buf0 : buf1);
    int ft_end = (int)(((unsigned)idx0 - (unsigned)N * (unsigned)k) & MASK);   // d == 0: k == 0
    bool pieces = false;                                                     // indices come from the table
    if (vote(check && first_bad < N)) {                                      // cold: a warp's groups go through it together
        int j0 = 0, kk = k;
        unsigned e0 = (unsigned)idx0;
#pragma unroll 1
        for (int seg = 1;; ++seg) {
            const bool more = check && !serial && first_bad < N;
            if (!vote(more)) break;
            group_sync<T>();                                 // everyone has read the cell before it is reset

            if (more && seg >= SEGMENTS) serial = true;
            const bool go = more && !serial;
            if (go) {
#pragma unroll 1
                for (int c = 0; c < P::R; ++c) {                             // the finished stretch j0 .. first_bad
                    const int i = tid + T * c;
                    if (i >= j0 && i <= first_bad) ftidx[i] = (int)((e0 - (unsigned)(i - j0) * (unsigned)kk) & MASK);
                }
                const int cb = (int)((e0 - (unsigned)(first_bad - j0) * (unsigned)kk) & MASK);
                e0 = (unsigned)ft_step(cb, d, wrap);
                j0 = first_bad + 1;
                kk = (int)e0 - (int)fx::fsub((float)(int)e0, d);
                pieces = true;
            }
            const int fb = group_min(go ? first_fail(j0, e0, kk) : N);
            first_bad = go ? fb : N;
        }
        if (pieces && !serial) {                                             // the last piece, and where the index ends
#pragma unroll 1
            for (int c = 0; c < P::R; ++c) {
                const int i = tid + T * c;
                if (i >= j0) ftidx[i] = (int)((e0 - (unsigned)(i - j0) * (unsigned)kk) & MASK);
            }
            ft_end = (int)((e0 - (unsigned)(N - j0) * (unsigned)kk) & MASK);
        }
    }
    // a warp may hold several groups: they enter the fallback together (a multi-warp group is alone in its domain)
    if (vote(serial)) {                                      // literal recurrence on thread 0
        if (serial && tid == 0) {
            int idx = idx0;
            for (int i = 0; i < N; ++i) {
                ftidx[i] = idx;
                idx = ft_step(idx, d, wrap);
            }
            ft_end = idx;
        }
        group_sync<T>();
    }
    const bool progression = !serial && !pieces;
    if (tid == 0 && live) *ft_out = ft_end;
    cf v[P::R];
    // two copies of the load loop, so that the usual one has no shared-memory read between its global loads (a select on
    // `progression` inside ONE loop keeps the compiler from batching the 3 x R loads: measured 2x slower at SF7-9)
    // (the compensated variant likewise: a per-sample `if (comp)` around a call splits the loop into 16 basic blocks)
    auto load = [&](auto from_table, auto comp) {
        constexpr bool TABLE = decltype(from_table)::value, COMP = decltype(comp)::value;
        if constexpr (EX && !COMP && T <= 32) {              // (multi-warp windows have warps to overlap with; there the plain loop is faster)
            // Strided window (osr > 1 / bandwidth scale): the sample addresses are runtime products, not immediates, and the
            // compiler then serialises the loop -- three loads, their arithmetic, the next three loads: 16 memory round
            // trips per window on a walk that has nothing else to hide them behind.  Two explicit batches of eight.
            const cf* src = view.iq + wpos + (size_t)tid * a.osr;
            const size_t step = (size_t)T * a.osr;
#pragma unroll
            for (int h = 0; h < P::R; h += 8) {
                cf raw[8], chv[8], fnv[8];
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    const int c = h + q, i = tid + T * c;
                    raw[q] = chv[q] = fnv[q] = fx::mk(0.0f, 0.0f);
                    if (live) {
                        const int ci = TABLE ? ftidx[i] : (int)((c0 - (unsigned)c * cstep) & MASK);
                        raw[q] = ldg_cf(src + (size_t)c * step);
                        chv[q] = ldg_cf(chirp + i);
                        fnv[q] = ldg_cf(a.fine + ci);
                    }
                }
#pragma unroll
                for (int q = 0; q < 8; ++q)
                    v[h + q] = live ? fx::cmul(fx::cmul(raw[q], chv[q]), fnv[q]) : fx::mk(0.0f, 0.0f);
            }
            return;
        }
#pragma unroll
        for (int c = 0; c < P::R; ++c) {
            const int i = tid + T * c;
            cf x = fx::mk(0.0f, 0.0f);
            if (live) {
                // (pointer arithmetic, not a 32-bit sample index: `iq + wpos + i` folds i into the load's immediate offset;
                //  `iq + (wpos + i)` in uint32 does not and measured 2x slower on the sub-warp walks)
                const cf raw = COMP ? stream_sample_comp(view.iq, view.n, view.rate, view.offset, wpos + (uint32_t)i * a.osr)
                             : EX   ? ldg_cf(view.iq + wpos + (size_t)i * a.osr)
                                    : ldg_cf(view.iq + wpos + i);
                x = fx::cmul(raw, ldg_cf(chirp + i));
                const int ci = TABLE ? ftidx[i] : (int)((c0 - (unsigned)c * cstep) & MASK);
                x = fx::cmul(x, ldg_cf(a.fine + ci));
            }
            v[c] = x;
        }
    };
    if (EX && view.comp) {
        if (progression) load(std::false_type{}, std::true_type{}); else load(std::true_type{}, std::true_type{});
    } else {
        if (progression) load(std::false_type{}, std::false_type{}); else load(std::true_type{}, std::false_type{});
    }
    fft_group<SF, true>(v, tid, a.tw, a.ct, buf0, buf1, flip);
    const Peak pk = group_peak<SF>(local_peak<SF>(v, tid), red);
    double part = 0.0;
#pragma unroll
    for (int r = 0; r < P::R; ++r) part += (double)fx::cnorm(v[r]);
    const double tree = group_dsum<T>(part, redd);
    cf* spec = flip ? buf1 : buf0;
    flip ^= 1;
#pragma unroll
    for (int r = 0; r < P::R; ++r) spec[P::bin_of_reg(tid, r)] = v[r];
    group_sync<T>();
    DetectOut o; o.value = 0; o.snr = 0.0f; o.findex = 0.0f;
    float rest = 0.0f;
    if (tid == 0 && live) {
        const double eps = tree * (4.0 * (double)N * 0x1p-53);
        rest = (float)((tree - eps) - (double)pk.val);
        if (a.force_serial || !(rest == (float)((tree + eps) - (double)pk.val))) {
            double total = 0.0;                              // LoRaDetector.hpp:45-52, ascending
            for (int i = 0; i < N; ++i) total += (double)fx::cnorm(spec[i]);
            rest = (float)(total - (double)pk.val);
        }
    }
    // The detector's two hypotf and two log10f (LoRaDetector.hpp:60-72) run on threads 0 and 1 side by side -- same code on
    // two lanes, half the serial tail: thread 0 takes the left neighbour and the fundamental, thread 1 the right neighbour
    // and the noise.  (T >= 8, so both sit in one warp; every thread of the warp executes the shuffles.)
    const int lane = threadIdx.x & 31;
    const float rest1 = __shfl_sync(0xffffffffu, rest, (lane - (tid & 31)) & 31);
    float hyp = 0.0f, lg = 0.0f;
    if (tid < 2 && live) {
        const int nb = tid == 0 ? (pk.idx > 0 ? pk.idx - 1 : N - 1) : (pk.idx < N - 1 ? pk.idx + 1 : 0);
        const cf c = spec[nb];
        hyp = gm::hypotf_g(c.x, c.y);
        lg = gm::log10f_g(__fsqrt_rn(tid == 0 ? pk.val : rest1));
    }
    const float right = __shfl_sync(0xffffffffu, hyp, (lane + 1) & 31);
    const float lg_noise = __shfl_sync(0xffffffffu, lg, (lane + 1) & 31);
    if (tid == 0 && live) {
        const float fundamental = __fsqrt_rn(pk.val);
        const float left = hyp;
        const double demon = (2.0 * (double)fundamental) - (double)right - (double)left;
        const float power = fx::fsub(fx::fmul(20.0f, lg), a.power_scale);
        const float power_avg = fx::fsub(fx::fmul(20.0f, lg_noise), a.power_scale);
        o.value = pk.idx;
        o.snr = fx::fsub(power, power_avg);
        o.findex = (demon == 0.0) ? 0.0f : (float)(0.5 * (double)fx::fsub(right, left) / demon);
    }
    group_sync<T>();                                         // spectrum / ftidx free for the next window
    return o;
}

template <int SF, bool EX>
__global__ void __launch_bounds__(sync_cta<SF>(), sync_threads_per_sm<SF>() / sync_cta<SF>()) stream_sync_kernel(const StreamSyncArgs a)
{
    using P = fx::Plan<SF>;
    using S = SyncSmem<SF>;
    constexpr int T = P::T, N = P::N, G = S::G;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int g = threadIdx.x / T, tid = threadIdx.x % T;
    cf* base = reinterpret_cast<cf*>(smem_raw);
    cf* buf0 = base + (size_t)g * N;
    cf* buf1 = base + (size_t)G * N + (size_t)g * N;
    Peak* red = reinterpret_cast<Peak*>(smem_raw + S::red_off);
    double* redd = reinterpret_cast<double*>(smem_raw + S::red_off + 8 * sizeof(Peak));
    GroupCtl& ctl = reinterpret_cast<GroupCtl*>(smem_raw + S::ctl_off)[g];
    int flip = 0;

    const uint32_t slot = blockIdx.x * G + g;
    const bool have = slot < a.n_streams;
    const uint32_t sid = have ? (a.ids ? a.ids[slot] : slot) : 0;
    const cf* iq = a.iq + (size_t)(have ? sid : 0) * a.stream_stride;
    const uint32_t n = have ? (a.counts ? min(a.counts[sid], a.n_samples) : a.n_samples) : 0;
    const uint32_t osr = EX ? a.osr : 1u, bwk = EX ? a.bw_k : 1u;
    StreamView view;
    view.iq = iq; view.n = n; view.comp = have && a.comp_cfo != nullptr; view.rate = 0.0f; view.offset = 0;
    if (view.comp) {                                                          // phy.cpp:158-165
        const float PI_F = 3.14159265358979323846f;
        view.rate = __fdiv_rn(fx::fmul(-2.0f * PI_F, a.comp_cfo[sid]), fx::fmul((float)N, (float)osr));
        view.offset = (int)roundf(a.comp_to[sid]);
    }

    // state machine registers (thread 0 of the group)
    int state = ST_FRAMESYNC, prev_value = 0, freq_error = 0, fine_idx = 0, down = 0;
    float fine_err = 0.0f;
    uint32_t pos = 0, sym_count = 0, frames = 0, cur_start = 0;
    uint32_t pf_end = 0;                                                       // samples below this were asked for already

    for (;;) {
        if (tid == 0) {
            ctl.pos = pos;
            ctl.live = have && (unsigned long long)pos + 2ull * N * osr <= n;  // LoRaDemod.cpp:148
            ctl.down = down;
            ctl.ft = fine_idx;
            ctl.fine_err = fine_err;
            ctl.need2 = 0;
        }
        group_sync<T>();
        const bool live = ctl.live != 0;
        if (!dom_any<T>(live)) break;
        const uint32_t wpos = ctl.pos;
        const float ferr = ctl.fine_err;
        if (live) {                                                            // L2 prefetch of the samples the next windows can touch
            const unsigned long long span = (unsigned long long)N * osr;
            unsigned long long lo = wpos + span, hi = wpos + 4ull * span;
            if (lo < pf_end) lo = pf_end;
            if (hi > n) hi = n;
            if (hi > lo) {
                const long long shift = view.comp ? (long long)view.offset : 0ll;  // stream_sample reads m - offset
                for (unsigned long long m = (lo & ~15ull) + 16ull * tid; m < hi; m += 16ull * T) {
                    const long long idx = (long long)m - shift;
                    if (idx >= 0 && idx < (long long)n) asm volatile("prefetch.global.L2 [%0];" ::"l"(iq + idx));
                }
                pf_end = (uint32_t)hi;
            }
        }
        const cf* chirp = ctl.down ? a.down : a.up;
        // ONE copy of the window code for the window itself (w = 0) and the look-ahead for sync word1 (w = 1, :188-205;
        // it works on a copy of the fine-tune index: _fineTuneIndex itself is not advanced)
        DetectOut d; d.value = 0; d.snr = 0.0f; d.findex = 0.0f;
        bool squelched = false, syncd = false, match0 = false, need2 = false;
        int value1 = 0;
#pragma unroll 1
        for (int w = 0; w < 2; ++w) {
            int ftv = 0;
            const DetectOut o = sync_window<SF, EX>(a, view, wpos + (w ? (uint32_t)N * osr : 0u), live && (w == 0 || need2), chirp,
                                                    w ? ctl.ft2 : ctl.ft, &ftv, ferr, tid, buf0, buf1, flip, red, redd, &ctl.ok);
            if (w) {
                if (tid == 0 && need2) { value1 = o.value; d.findex = o.findex; }
                break;
            }
            d = o;
            if (tid == 0 && live) {
                fine_idx = ftv;
                squelched = d.snr < a.thresh;
                if (state == ST_FRAMESYNC) {                                   // :179-185
                    syncd = !squelched && (prev_value + 4) / 8 == 0;
                    match0 = (unsigned)(d.value + 4) / 8 == a.sync0;
                    if (syncd && match0) { ctl.need2 = 1; ctl.ft2 = ftv; }
                }
            }
            group_sync<T>();
            need2 = ctl.need2 != 0;
            if (!dom_any<T>(need2)) break;
        }
        if (tid == 0 && live) {
            uint32_t total = 0;
            switch (state) {
            case ST_FRAMESYNC: {
                const bool match1 = need2 && (unsigned)(value1 + 4) / 8 == a.sync1;
                if (syncd && match0 && match1) { total = 2 * N; state = ST_DOWNCHIRP0; down = 1; cur_start = wpos; }
                else if (!squelched) { total = (uint32_t)(N - d.value / (int)bwk); fine_err = fx::fadd(fine_err, d.findex); }   // bw_k = 1: upstream
                else { total = N; fine_err = 0.0f; fine_idx = 0; }
            } break;
            case ST_DOWNCHIRP0: {                                              // :240-250
                state = ST_DOWNCHIRP1; total = N;
                freq_error = d.value > N / 2 ? d.value - N : d.value;
            } break;
            case ST_DOWNCHIRP1: {                                              // :253-270
                state = ST_QUARTERCHIRP; total = N; down = 0;
                const int error = d.value > N / 2 ? d.value - N : d.value;
                freq_error = (freq_error + error) / 2;
            } break;
            case ST_QUARTERCHIRP: {                                            // :273-283
                state = ST_DATASYMBOLS;
                total = (uint32_t)(N / 4 + (freq_error / 2) / (int)bwk);
                fine_err = fx::fadd(fine_err, (float)(freq_error / 2));
                sym_count = 0;
            } break;
            default: {                                                         // ST_DATASYMBOLS, :286-310
                total = N;
                if (frames < a.max_frames)
                    a.symbols[((size_t)sid * a.max_frames + frames) * a.mtu + sym_count] = (int16_t)d.value;
                ++sym_count;
                if (sym_count >= a.mtu || squelched) {
                    if (frames < a.max_frames) {
                        const size_t f = (size_t)sid * a.max_frames + frames;
                        a.frame_start[f] = cur_start;
                        a.frame_err[f] = freq_error;
                        a.frame_nsym[f] = (int32_t)sym_count;
                    }
                    ++frames;
                    fine_err = 0.0f;
                    state = ST_FRAMESYNC;
                }
            } break;
            }
            pos = wpos + total * osr;
            prev_value = (int)(short)d.value;
        }
    }
    if (tid == 0 && have) {
        if (a.n_frames) a.n_frames[sid] = frames;
        if (a.end_pos) a.end_pos[sid] = pos;
    }
}

template <int SF>
cudaError_t launch_sync_sf(const StreamSyncArgs& a, cudaStream_t stream)
{
    using S = SyncSmem<SF>;
    const bool ex = a.osr != 1 || a.bw_k != 1 || a.comp_cfo != nullptr;
    const unsigned grid = (a.n_streams + S::G - 1) / S::G;
    if (ex) {
        cudaFuncSetAttribute(stream_sync_kernel<SF, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)S::bytes);
        stream_sync_kernel<SF, true><<<grid, S::CTA, S::bytes, stream>>>(a);
    } else {
        cudaFuncSetAttribute(stream_sync_kernel<SF, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)S::bytes);
        stream_sync_kernel<SF, false><<<grid, S::CTA, S::bytes, stream>>>(a);
    }
    return cudaGetLastError();
}

} // namespace

cudaError_t launch_stream_sync(int sf, const StreamSyncArgs& a, cudaStream_t stream)
{
    switch (sf) {
    case 7: return launch_sync_sf<7>(a, stream);
    case 8: return launch_sync_sf<8>(a, stream);
    case 9: return launch_sync_sf<9>(a, stream);
    case 10: return launch_sync_sf<10>(a, stream);
    case 11: return launch_sync_sf<11>(a, stream);
    case 12: return launch_sync_sf<12>(a, stream);
    default: return cudaErrorInvalidValue;
    }
}

} // namespace lk
