// demod_lean.cu -- the default demodulator for pre-dechirped IQ (lora_demodulate, LoRaDemod.cpp:50-197; SF7-12, osr 1,
// no downchirp / window tables): ONE WARP PER PACKET, the packet streamed once in 4 KB steps, every window settled by the
// pruned transform of fft_prune.cuh and the tier-1 Parseval test of demod_cert.cu.  What it cannot settle it FLAGS for
// demod_cert_kernel, which api.cu launches right after over the flagged packets only.
//
// Why a kernel of its own: in demod_cert.cu a window is held in registers (16 samples per thread: the layout its full
// transform needs) -- 2..8 warps, four to six barriers and a 9-35 KB slot per window at SF10-12, 128 registers and a
// slow path the hot loop pays for at every SF.  The pruned transform needs none of that.  With f_c = F_t G^c (DESIGN.md,
// "geometric variant") the class-k0 output of pass 0 is a POLYNOMIAL in the raw samples,
//     V_t[k0] = F_t * sum_c r_{t + T c} h^c,   h = G W16^k0,  T = N / 16,
//   SF7-9   T <= 32: a 4 KB step is GD = 512 / N whole windows side by side (4 / 2 / 1), 16 samples per thread, four
//           halving levels, then the T-point transform across the lanes by shfl.xor;
//   SF10-12 T = 32 R: a window is R = N / 512 steps; lane l keeps R complex accumulators (t = l + 32 j), each step adds
//           16 / R terms to each (halving levels inside the step, one multiply-add with h^(c0) from the table's powers of
//           two to join).  At the end of the window the lane holds the R inputs of fft_prune.cuh's cross-warp stage: an
//           R-point DFT in registers, the W_T twiddles, R lane transforms, top bin.
// No shared-memory exchange, no barrier, no window ever resident; 2 x 4 KB of ring per warp, one cp.async.bulk per step.
//
// Per round of packets:
//   estimator phase (the ring is idle and doubles as the exact transform's exchange buffer): the exact two-symbol
//     estimator (LoRaDemod.cpp:80-139: exact-order FFT of fft_exact.cuh, first-max, fIndex, atan2f) -> cfo, time_offset,
//     rate, t_off bit for bit.  A 2^SF-point exact transform needs T threads: teams of T / 32 warps at SF10-12 (packet
//     after packet of the team's members), the warp itself at SF7-9 (SF7: four windows side by side = two jobs per
//     pass, so a round is two packets per warp).  The 1/max normalisation (LoRaDemod.cpp:61-78) is PREDICTED from the
//     warp's previous packet; the stream tracks the true max-abs, and a miss is re-estimated with the true scale in the
//     next estimator phase: the symbols (scale-invariant decisions) stand iff t_off is unchanged and the rate moved by
//     no more than the tolerance the certificate budgets, else the packet is flagged.
//   stream phase: every window as above.  A time offset (t_off != 0: every window but one shifted, LoRaDemod.cpp:143-150)
//     is streamed as one contiguous run from the 16-byte boundary below its first sample, the unshifted window comes
//     from global memory.  The first refused window (noise, weak signal, NaN / Inf), an irregular packet (partial
//     symbol, 8-byte alignment, non-finite max) flags the packet and abandons it.
// Symbols, sync word, status, cfo, time_offset are bit-exact either way: accepted windows carry demod_cert.cu's proof
// (same constants, same budget: DESIGN.md section 5), everything else is that kernel.
#include <atomic>
#include <type_traits>
#include "stream_common.cuh"
#include "fft_fast.cuh"
#include "fft_prune.cuh"
#include "internal.h"

namespace lk {

#ifndef LEAN_CTA_NARROW
#define LEAN_CTA_NARROW 256                   // SF7-9 (a window is at most one warp): threads per CTA, any multiple of 32
#endif
#ifndef LEAN_CTAS_NARROW
#define LEAN_CTAS_NARROW 2                    // ... and resident CTAs per SM for SF8 / SF9 (occupancy A/B knobs; measured on B200:
                                              // SF7 5729 -> 6146 GB/s with 3 x 256 threads (80 registers), SF8 6886 -> 6124)
#endif
#ifndef LEAN_CTAS_SF7
#define LEAN_CTAS_SF7 3
#endif
#ifndef LEAN_RETRY
#define LEAN_RETRY 1                          // SF7-9: a refused window is tried again with the two neighbouring bin classes (A/B knob)
#endif
#ifndef LEAN_PAIR_SF7
#define LEAN_PAIR_SF7 1                       // SF7: one exact estimator pass serves two jobs (A/B knob)
#endif
#ifndef LEAN_CTA
#define LEAN_CTA 256
#endif
#ifndef LEAN_CTAS_PER_SM
#define LEAN_CTAS_PER_SM 2
#endif
#ifndef LEAN_STAGES
#define LEAN_STAGES 2
#endif

constexpr float LEAN_U = 5.9604644775390625e-08f;
constexpr float LEAN_C0 = 384.0f;                 // CERT_C0 of demod_cert.cu (same budget)
constexpr int LEAN_CHUNK = 512;                   // samples per step: 4 KB
constexpr int LEAN_SLOT = LEAN_CHUNK + 2;         // entries per ring stage: a stream that starts 8 bytes into a 16-byte unit is
                                                  // copied from the boundary below it (one extra sample at either end)

template <int SF>
struct LeanCfg {
    static constexpr bool WIDE = SF >= 10;                                // a window spans R > 1 chunks (else GD windows per chunk)
    static constexpr int N = 1 << SF, T = N / 16, R = WIDE ? N / LEAN_CHUNK : 1;   // R = T / 32 accumulators per lane
    static constexpr int GD = WIDE ? 1 : LEAN_CHUNK / N;                  // windows per chunk (SF7: 4, SF8: 2, SF9: 1)
    static constexpr int JOBS = (LEAN_PAIR_SF7 && SF == 7) ? 2 : 1;       // packets per warp and round: SF7's estimator holds four
                                                                          // windows side by side = the two symbols of TWO packets
    static constexpr int CP = 16 / R;                                     // terms per accumulator and chunk
    static constexpr int LV = CP == 8 ? 3 : CP == 4 ? 2 : 1;              // halving levels inside a chunk
    static constexpr int CTA = WIDE ? LEAN_CTA : LEAN_CTA_NARROW, CTAS = WIDE ? LEAN_CTAS_PER_SM : (SF == 7 ? LEAN_CTAS_SF7 : LEAN_CTAS_NARROW);
    static constexpr int WARPS = CTA / 32;
    static constexpr size_t ring_bytes = (size_t)WARPS * LEAN_STAGES * LEAN_SLOT * sizeof(cf);
    static constexpr size_t bar_off = ring_bytes;
    static constexpr size_t tws_off = (bar_off + (size_t)WARPS * LEAN_STAGES * 8 + 15) & ~(size_t)15;   // [4][32] lane-stage factors
    static constexpr size_t twl_off = tws_off + 4 * 32 * sizeof(cf);                                   // W_N^e, e < 512
    static constexpr int TWL = WIDE ? 512 : N;                                                         // (narrow: the whole circle)
    static constexpr size_t tw32_off = twl_off + (size_t)TWL * sizeof(cf);                             // W_N^(32 e), e < 128
    static constexpr size_t twx_off = tw32_off + (WIDE ? 128 : 0) * sizeof(cf);                        // W_T^(l ka), [ka][l]
    static constexpr size_t w16_off = twx_off + (size_t)R * 32 * sizeof(cf);                           // W16^j
    static constexpr size_t hp_off = w16_off + 16 * sizeof(cf);                                        // per warp: [k0][4] powers of h
    static constexpr size_t g32_off = hp_off + (size_t)WARPS * 64 * sizeof(cf);                        // per warp: G32^j, j < 8
    static constexpr size_t cell_off = g32_off + (size_t)WARPS * 8 * sizeof(cf);                       // per warp: LeanTeamCell (48 B)
    static constexpr size_t est_off = cell_off + (size_t)WARPS * JOBS * 48;                            // per team: LeanEst[4] + LeanRes[2]
    static constexpr size_t red_off = est_off + (size_t)WARPS * 128;                                   // Peak[8] of the exact estimator
    static constexpr size_t bytes = red_off + 8 * 8;
};

__device__ __forceinline__ cf lean_cma(cf acc, cf a, cf w)            // acc + a * w: four fused multiply-adds
{
    return fx::mk(ff::fma_(-a.y, w.y, ff::fma_(a.x, w.x, acc.x)), ff::fma_(a.y, w.x, ff::fma_(a.x, w.y, acc.y)));
}

struct LeanTeamCell {       // one per warp: what the team's estimator phase hands to / takes from the warp's stream
    float rate; int t_off; int ok; float used_max;          // estimate for the packet about to be streamed
    int vpid; float vmax; float vrate; int vtoff;           // request: redo packet vpid's estimate with max-abs vmax
    float pred;                                             // max-abs the next estimate should assume (the warp's last packet's)
    int pad[3];
};
struct LeanEst { int idx; float fi; float phase; int pad; };
struct LeanRes { float cfo, toff, rate; int t_off; };       // a paired pass leaves each job's result here

template <int SF>
__global__ void __launch_bounds__(LeanCfg<SF>::CTA, LeanCfg<SF>::CTAS) demod_lean_kernel(const DemodArgs a)
{
    using C = LeanCfg<SF>;
    using PR = fp::Prune<SF>;
    using P = fx::Plan<SF>;
    constexpr int N = C::N, T = C::T, R = C::R, LV = C::LV, S = LEAN_STAGES, TW = T > 32 ? T / 32 : 1;
    constexpr bool WIDE = C::WIDE;
    constexpr int GD = C::GD, GDE = T < 32 ? 32 / T : 1;       // windows per chunk; windows the estimator runs side by side
    static_assert(!WIDE || C::CTA == CTA_THREADS, "the exact estimator's group helpers assume CTA_THREADS for multi-warp windows");
    static_assert((size_t)TW * S * LEAN_SLOT >= (size_t)N * (GDE < 2 ? 1 : 2), "the team's ring doubles as the exact transform's exchange buffer");
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int warp = threadIdx.x >> 5, l = threadIdx.x & 31;
    // team = the threads of one estimator transform's owner set: T / 32 warps (SF10-12) or one warp (SF7-9, where a warp
    // holds 32 / T windows side by side); tid = thread inside a window, grp = window group inside the warp
    const int team = WIDE ? threadIdx.x / T : warp, tid = threadIdx.x % T, wt = warp % TW, grp = WIDE ? 0 : l / T;
    const bool lead = WIDE ? tid == 0 : l == 0;                // the team's first thread
    cf* ring = reinterpret_cast<cf*>(smem_raw) + (size_t)warp * S * LEAN_SLOT;
    cf* ex = reinterpret_cast<cf*>(smem_raw) + (size_t)(team * TW) * S * LEAN_SLOT + (size_t)grp * N;   // N entries: exchange, then spectrum
    uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw + C::bar_off) + warp * S;
    cf* tws = reinterpret_cast<cf*>(smem_raw + C::tws_off);
    cf* twl = reinterpret_cast<cf*>(smem_raw + C::twl_off);
    cf* tw32 = reinterpret_cast<cf*>(smem_raw + C::tw32_off);
    cf* twx = reinterpret_cast<cf*>(smem_raw + C::twx_off);
    cf* w16 = reinterpret_cast<cf*>(smem_raw + C::w16_off);
    cf* hp = reinterpret_cast<cf*>(smem_raw + C::hp_off) + warp * 64;
    cf* g32 = reinterpret_cast<cf*>(smem_raw + C::g32_off) + warp * 8;
    constexpr int JOBS = C::JOBS;
    LeanTeamCell* cells = reinterpret_cast<LeanTeamCell*>(smem_raw + C::cell_off) + team * TW * JOBS;
    LeanEst* est = reinterpret_cast<LeanEst*>(smem_raw + C::est_off + (size_t)(WIDE ? team * TW : warp) * 128);
    LeanRes* res = reinterpret_cast<LeanRes*>(est + 4);
    Peak* red = reinterpret_cast<Peak*>(smem_raw + C::red_off);

    if (l == 0) {
#pragma unroll
        for (int s = 0; s < S; ++s) ptx::mbar_init(&full[s], 1);
        ptx::fence_mbar_init();
    }
    for (int j = threadIdx.x; j < PR::STAGES_TW * PR::W; j += C::CTA) {
        double sn, cs;
        sincospi(-2.0 * (double)PR::stage_exp(j / PR::W, j % PR::W) / (double)PR::W, &sn, &cs);
        tws[j] = fx::mk((float)cs, (float)sn);
    }
    for (int j = threadIdx.x; j < C::TWL; j += C::CTA) {
        double sn, cs;
        sincospi(-2.0 * (double)(j & (N - 1)) / (double)N, &sn, &cs);
        twl[j] = fx::mk((float)cs, (float)sn);
    }
    for (int j = threadIdx.x; j < (WIDE ? 128 : 0); j += C::CTA) {
        double sn, cs;
        sincospi(-2.0 * (double)((32 * j) & (N - 1)) / (double)N, &sn, &cs);
        tw32[j] = fx::mk((float)cs, (float)sn);
    }
    for (int j = threadIdx.x; j < R * 32; j += C::CTA) {
        double sn, cs;
        sincospi(-2.0 * (double)PR::cross_exp(j / 32, j % 32) / (double)T, &sn, &cs);
        twx[j] = fx::mk((float)cs, (float)sn);
    }
    for (int j = threadIdx.x; j < 16; j += C::CTA) {
        double sn, cs;
        sincospi(-2.0 * (double)j / 16.0, &sn, &cs);
        w16[j] = fx::mk((float)cs, (float)sn);
    }
    __syncthreads();

    const uint32_t ring0 = ptx::smem_u32(ring), bar0 = ptx::smem_u32(full);
    int stage = 0;
    unsigned phase_bits = 0;
    unsigned st_pruned = 0, st_flagged = 0, st_est2 = 0;
    const uint32_t n_warps = gridDim.x * C::WARPS;
    const uint32_t gw0 = blockIdx.x * C::WARPS + (uint32_t)(team * TW);            // the team's first warp, grid-wide
    float pred_max = 0.0f;                                     // max-abs of this warp's previous packet
    int vpid[JOBS]; float vmax[JOBS], vrate[JOBS]; int vtoff[JOBS];                // this warp's pending re-estimates
#pragma unroll
    for (int j = 0; j < JOBS; ++j) { vpid[j] = -1; vmax[j] = 0.0f; vrate[j] = 0.0f; vtoff[j] = 0; }

    // The exact two-symbol estimator of packet `pid` with the samples scaled by 1/use_max when use_max > 1
    // (LoRaDemod.cpp:61-139), by the whole team; thread 0 of the team gets cfo / time_offset / rate / t_off.
    auto estimate = [&](uint32_t pid, float use_max, float& cfo, float& toff, float& rate, int& t_off) {
        const cf* pkt = a.iq + (size_t)pid * a.packet_stride;
        const bool norm = use_max > 1.0f;
        const float scale = norm ? __fdiv_rn(1.0f, use_max) : 1.0f;
#pragma unroll 1
        for (int s0 = 0; s0 < 2; s0 += GDE) {
            const int s = s0 + grp;
            const bool live = s < 2;
            cf v[P::R];
#pragma unroll
            for (int c = 0; c < P::R; ++c) {
                cf x = fx::mk(0.0f, 0.0f);
                if (live) {
                    x = ldg_cf(pkt + (size_t)s * N + tid + T * c);
                    if (norm) x = fx::cscale(x, scale);                                // LoRaDemod.cpp:75-77
                }
                v[c] = x;
            }
            fft_inplace<SF, true>(v, tid, a.tw, a.ct, ex);
            const Peak pk = group_peak<SF>(local_peak<SF>(v, tid), red);
            group_sync<T>();                                   // last exchange fully read: the buffer becomes the spectrum
#pragma unroll
            for (int r = 0; r < P::R; ++r) ex[P::bin_of_reg(tid, r)] = v[r];
            group_sync<T>();
            if (tid == 0 && live) {
                LeanEst e; e.idx = 0; e.fi = 0.0f; e.phase = 0.0f;
                cf best_bin = fx::mk(0.0f, 0.0f);
                if (pk.val > 0.0f) {                           // candidate kept iff its power beats -1e30 dB: max > 0
                    const float fundamental = __fsqrt_rn(pk.val);
                    const cf lft = ex[pk.idx > 0 ? pk.idx - 1 : N - 1];
                    const cf rgt = ex[pk.idx < N - 1 ? pk.idx + 1 : 0];
                    const float left = gm::hypotf_g(lft.x, lft.y), right = gm::hypotf_g(rgt.x, rgt.y);
                    const double demon = (2.0 * (double)fundamental) - (double)right - (double)left;
                    e.fi = (demon == 0.0) ? 0.0f : (float)(0.5 * (double)fx::fsub(right, left) / demon);
                    e.idx = pk.idx;
                    best_bin = ex[pk.idx];
                }
                e.phase = gm::atan2f_g(best_bin.y, best_bin.x);
                est[s] = e;
            }
            group_sync<T>();
        }
        if (lead) {                                            // LoRaDemod.cpp:126-139
            const float PI_F = 3.14159265358979323846f;
            const LeanEst e0 = est[0], e1 = est[1];
            float sum_index = fx::fadd(0.0f, fx::fadd((float)e0.idx, e0.fi));
            sum_index = fx::fadd(sum_index, fx::fadd((float)e1.idx, e1.fi));
            float dd = fx::fsub(e1.phase, e0.phase);
            while (dd > PI_F) dd = fx::fsub(dd, 2.0f * PI_F);
            while (dd < -PI_F) dd = fx::fadd(dd, 2.0f * PI_F);
            const float phase_diff = fx::fadd(0.0f, dd);
            const float avg_index = __fdiv_rn(sum_index, 2.0f);
            const float cfo_coarse = __fdiv_rn(avg_index, (float)N);
            const float cfo_fine = __fdiv_rn(__fdiv_rn(phase_diff, 1.0f), fx::fmul(2.0f * PI_F, (float)N));
            cfo = fx::fadd(cfo_coarse, cfo_fine);
            const float frac = fx::fsub(avg_index, floorf(fx::fadd(avg_index, 0.5f)));
            const float avg_t = __fdiv_rn(0.0f, 2.0f);
            toff = fx::fsub(avg_t, fx::fmul(frac, (float)N));
            t_off = (int)roundf(toff);
            rate = __fdiv_rn(fx::fmul(-2.0f * PI_F, cfo), (float)N);
        }
    };

    // SF7 (T = 8: four windows side by side): ONE exact pass for two jobs -- groups 0, 1 hold the two symbols of job 0,
    // groups 2, 3 those of job 1; each job's first thread leaves cfo / time_offset / rate / t_off in res[job].
    auto estimate_pair = [&](uint32_t pid0, float max0, bool on0, uint32_t pid1, float max1, bool on1) {
        const int job = grp >> 1, s = grp & 1;
        const bool live = job ? on1 : on0;
        const cf* pkt = a.iq + (size_t)(job ? pid1 : pid0) * a.packet_stride;
        const float use_max = job ? max1 : max0;
        const bool norm = use_max > 1.0f;
        const float scale = norm ? __fdiv_rn(1.0f, use_max) : 1.0f;
        cf v[P::R];
#pragma unroll
        for (int c = 0; c < P::R; ++c) {
            cf x = fx::mk(0.0f, 0.0f);
            if (live) {
                x = ldg_cf(pkt + (size_t)s * N + tid + T * c);
                if (norm) x = fx::cscale(x, scale);                                    // LoRaDemod.cpp:75-77
            }
            v[c] = x;
        }
        fft_inplace<SF, true>(v, tid, a.tw, a.ct, ex);
        const Peak pk = group_peak<SF>(local_peak<SF>(v, tid), red);
        group_sync<T>();
#pragma unroll
        for (int r = 0; r < P::R; ++r) ex[P::bin_of_reg(tid, r)] = v[r];
        group_sync<T>();
        if (tid == 0 && live) {
            LeanEst e; e.idx = 0; e.fi = 0.0f; e.phase = 0.0f; e.pad = 0;
            cf best_bin = fx::mk(0.0f, 0.0f);
            if (pk.val > 0.0f) {
                const float fundamental = __fsqrt_rn(pk.val);
                const cf lft = ex[pk.idx > 0 ? pk.idx - 1 : N - 1];
                const cf rgt = ex[pk.idx < N - 1 ? pk.idx + 1 : 0];
                const float left = gm::hypotf_g(lft.x, lft.y), right = gm::hypotf_g(rgt.x, rgt.y);
                const double demon = (2.0 * (double)fundamental) - (double)right - (double)left;
                e.fi = (demon == 0.0) ? 0.0f : (float)(0.5 * (double)fx::fsub(right, left) / demon);
                e.idx = pk.idx;
                best_bin = ex[pk.idx];
            }
            e.phase = gm::atan2f_g(best_bin.y, best_bin.x);
            est[grp] = e;
        }
        group_sync<T>();
        if (tid == 0 && s == 0 && live) {                      // LoRaDemod.cpp:126-139, one thread per job
            const float PI_F = 3.14159265358979323846f;
            const LeanEst e0 = est[2 * job], e1 = est[2 * job + 1];
            float sum_index = fx::fadd(0.0f, fx::fadd((float)e0.idx, e0.fi));
            sum_index = fx::fadd(sum_index, fx::fadd((float)e1.idx, e1.fi));
            float dd = fx::fsub(e1.phase, e0.phase);
            while (dd > PI_F) dd = fx::fsub(dd, 2.0f * PI_F);
            while (dd < -PI_F) dd = fx::fadd(dd, 2.0f * PI_F);
            const float phase_diff = fx::fadd(0.0f, dd);
            const float avg_index = __fdiv_rn(sum_index, 2.0f);
            const float cfo_coarse = __fdiv_rn(avg_index, (float)N);
            const float cfo_fine = __fdiv_rn(__fdiv_rn(phase_diff, 1.0f), fx::fmul(2.0f * PI_F, (float)N));
            LeanRes r;
            r.cfo = fx::fadd(cfo_coarse, cfo_fine);
            const float frac = fx::fsub(avg_index, floorf(fx::fadd(avg_index, 0.5f)));
            const float avg_t = __fdiv_rn(0.0f, 2.0f);
            r.toff = fx::fsub(avg_t, fx::fmul(frac, (float)N));
            r.t_off = (int)roundf(r.toff);
            r.rate = __fdiv_rn(fx::fmul(-2.0f * PI_F, r.cfo), (float)N);
            res[job] = r;
        }
        group_sync<T>();
    };

#pragma unroll 1
    for (uint32_t round = 0;; ++round) {
        // ---- estimator phase (the team's ring is idle): pending re-estimates, then the packets about to be streamed ------
        if (l == 0) {
#pragma unroll
            for (int j = 0; j < JOBS; ++j) {
                LeanTeamCell& c = cells[wt * JOBS + j];
                c.vpid = vpid[j]; c.vmax = vmax[j]; c.vrate = vrate[j]; c.vtoff = vtoff[j]; c.pred = pred_max;
            }
        }
        group_sync<T>();
        bool any = false;
        if constexpr (JOBS == 2) {
            // SF7: two re-estimates in one exact pass (when any is pending), the two new packets in another
            const int v0 = cells[0].vpid, v1 = cells[1].vpid;
            if (v0 >= 0 || v1 >= 0) {
                any = true;
                estimate_pair((uint32_t)max(v0, 0), cells[0].vmax, v0 >= 0, (uint32_t)max(v1, 0), cells[1].vmax, v1 >= 0);
                if (lead) {
#pragma unroll
                    for (int q = 0; q < 2; ++q) {
                        const LeanTeamCell& c = cells[q];
                        if (c.vpid < 0) continue;
                        const LeanRes r = res[q];
                        // the streamed symbols stand iff the window bases and (within the certified tolerance) the rate do
                        const bool same = r.t_off == c.vtoff && fabsf(fx::fsub(r.rate, c.vrate)) * (float)N <= 64.0f * LEAN_U;
                        if (same) {
                            if (a.cfo) a.cfo[c.vpid] = r.cfo;
                            if (a.time_offset) a.time_offset[c.vpid] = r.toff;
                        } else {
                            a.flagged[c.vpid] = 1; ++st_flagged;
                        }
                        ++st_est2;
                    }
                }
                group_sync<T>();
            }
            uint32_t np[2] = { 0u, 0u };
            bool non[2] = { false, false };
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const unsigned long long pm = (unsigned long long)gw0 + ((unsigned long long)round * 2 + (unsigned)j) * n_warps;
                int ok = -1;
                if (pm < a.n_packets) {
                    const uint32_t pid = (uint32_t)pm;
                    const cf* pkt = a.iq + (size_t)pid * a.packet_stride;
                    const uint32_t count = a.sample_counts ? min(a.sample_counts[pid], a.samples_per_packet) : a.samples_per_packet;
                    const uint32_t total = count / (uint32_t)N;
                    const bool shape = total >= 2 && count == total * (uint32_t)N && (reinterpret_cast<uintptr_t>(pkt) & 15) == 0 &&
                                       cells[j].pred < 3.0e38f;
                    ok = shape ? 1 : 0;
                    np[j] = pid; non[j] = shape;
                    any = true;
                }
                if (lead) cells[j].ok = ok;
            }
            if (non[0] || non[1]) {
                estimate_pair(np[0], cells[0].pred, non[0], np[1], cells[1].pred, non[1]);
                if (lead) {
#pragma unroll
                    for (int q = 0; q < 2; ++q) {
                        if (!non[q]) continue;
                        const LeanRes r = res[q];
                        LeanTeamCell& c = cells[q];
                        if (a.cfo) a.cfo[np[q]] = r.cfo;
                        if (a.time_offset) a.time_offset[np[q]] = r.toff;
                        c.rate = r.rate; c.t_off = r.t_off; c.used_max = c.pred;
                        c.pad[2] = fabsf(r.toff) > 0.02f ? 1 : 0;
                    }
                }
                group_sync<T>();
            }
        } else
#pragma unroll 1
        for (int m = 0; m < TW; ++m) {
            const int vp = cells[m].vpid;
            if (vp >= 0) {
                any = true;
                float cfo = 0.0f, toff = 0.0f, rate = 0.0f; int t_off = 0;
                estimate((uint32_t)vp, cells[m].vmax, cfo, toff, rate, t_off);
                if (lead) {
                    // the streamed symbols stand iff the window bases and (within the certified tolerance) the rate do
                    const bool same = t_off == cells[m].vtoff && fabsf(fx::fsub(rate, cells[m].vrate)) * (float)N <= 64.0f * LEAN_U;
                    if (same) {
                        if (a.cfo) a.cfo[vp] = cfo;
                        if (a.time_offset) a.time_offset[vp] = toff;
                    } else {
                        a.flagged[vp] = 1; ++st_flagged;
                    }
                    ++st_est2;
                }
            }
            const unsigned long long pm = (unsigned long long)gw0 + (unsigned)m + (unsigned long long)round * n_warps;
            int ok = 0;
            if (pm < a.n_packets) {
                any = true;
                const uint32_t pid = (uint32_t)pm;
                const cf* pkt = a.iq + (size_t)pid * a.packet_stride;
                const uint32_t count = a.sample_counts ? min(a.sample_counts[pid], a.samples_per_packet) : a.samples_per_packet;
                const uint32_t total = count / (uint32_t)N;
                // regular packets only; everything else is the certified kernel's
                const bool shape = total >= 2 && count == total * (uint32_t)N && (reinterpret_cast<uintptr_t>(pkt) & 15) == 0 &&
                                   cells[m].pred < 3.0e38f;
                if (shape) {
                    float cfo = 0.0f, toff = 0.0f, rate = 0.0f; int t_off = 0;
                    estimate(pid, cells[m].pred, cfo, toff, rate, t_off);
                    if (lead) {
                        if (a.cfo) a.cfo[pid] = cfo;
                        if (a.time_offset) a.time_offset[pid] = toff;
                        cells[m].rate = rate; cells[m].t_off = t_off; cells[m].used_max = cells[m].pred;
                        cells[m].pad[2] = fabsf(toff) > 0.02f ? 1 : 0;         // sync peaks off the grid: not a clean packet
                    }
                    ok = 1;
                }
                if (lead) cells[m].ok = ok;
            } else if (lead) {
                cells[m].ok = -1;                              // no packet this round
            }
        }
        ptx::fence_proxy_async();                              // exchange-buffer stores precede the bulk copies into the ring
        group_sync<T>();
        if (!any) break;
#pragma unroll
        for (int j = 0; j < JOBS; ++j) vpid[j] = -1;

        // ---- stream phase: every warp its own packet(s) --------------------------------------------------------------------
#pragma unroll 1
        for (int job = 0; job < JOBS; ++job) {
        const unsigned long long pme = (unsigned long long)blockIdx.x * C::WARPS + (unsigned)warp +
                                       ((unsigned long long)round * JOBS + (unsigned)job) * n_warps;
        LeanTeamCell* mycell = &cells[wt * JOBS + job];
        const LeanTeamCell mine = *mycell;
        if (mine.ok >= 0 && pme < a.n_packets) {
            const uint32_t pid = (uint32_t)pme;
            const cf* pkt = a.iq + (size_t)pid * a.packet_stride;
            const uint32_t count = a.sample_counts ? min(a.sample_counts[pid], a.samples_per_packet) : a.samples_per_packet;
            const uint32_t total = count / (uint32_t)N;
            const float rate = mine.rate;
            const float start_max = fabsf(rate) * ((float)total * (float)N) * 1.01f;
            const float Kc = LEAN_U * (LEAN_C0 + start_max + 2.0f * fabsf(rate) * (float)N);
            // tier 1 available (demod_cert.cu)
            bool bad = !(mine.ok == 1 && mine.t_off > -N && mine.t_off < N && Kc < 0.0078125f);
            const float tier1 = fp::TIER1 * (float)N;
            float mmax = 0.0f;
            int sw0 = 0, sw1 = 0;
            if constexpr (WIDE) {
            if (!bad) {
                // per packet: F_l = exp(j rate l); G^(1,2,4,8) (G = exp(j rate T)) and G32^j (exp(j rate 32 j)) in double;
                // hp[k0][m] = G^(2^m) W16^(k0 2^m)
                cf geo_f;
                {
                    float sn, cs;
                    __sincosf(fx::fmul(rate, (float)l), &sn, &cs);     // |rate l| < 2 pi / 128
                    geo_f = fx::mk(cs, sn);
                    double ds = 0.0, dc = 1.0;
                    if (l < 4) sincos((double)rate * (double)(T << l), &ds, &dc);
                    else if (l < 12) sincos((double)rate * (double)(32 * (l - 4)), &ds, &dc);
                    const float gx = (float)dc, gy = (float)ds;
                    if (l >= 4 && l < 12) g32[l - 4] = fx::mk(gx, gy);
#pragma unroll
                    for (int h2 = 0; h2 < 2; ++h2) {
                        const int e = l + 32 * h2, k = e >> 2, mi = e & 3;
                        const cf g = fx::mk(__shfl_sync(0xffffffffu, gx, mi), __shfl_sync(0xffffffffu, gy, mi));
                        hp[e] = ff::fcmul(g, w16[(k << mi) & 15]);
                    }
                    __syncwarp();
                }
                const cf geo_g1 = hp[0];

                // the stream: 4 KB chunks, contiguous.  With a time offset (LoRaDemod.cpp:143-150, count == total N here) every
                // window but the last (t_off > 0) or the first (t_off < 0) is shifted: those are one contiguous run through
                // the ring (copied from the 16-byte boundary below an odd first sample), the unshifted window is read from
                // global memory by the same code afterwards.
                // Two copies of the stream: clean packets (no time offset, sync peaks on the grid) run code in which the offset,
                // the odd-alignment entry, the extra unshifted window and the retry are compile-time absent; the others the
                // general one.
                auto run = [&](auto shifted_tag) {
                constexpr bool SH = decltype(shifted_tag)::value;
                const int toff = SH ? mine.t_off : 0;
                const uint32_t run0 = toff > 0 ? (uint32_t)toff : (toff < 0 ? (uint32_t)(N + toff) : 0u);
                const uint32_t n_run = toff ? total - 1 : total;
                const uint32_t sym0 = toff < 0 ? 1u : 0u;
                const uint32_t odd = run0 & 1u;
                const cf* src0 = pkt + (run0 - odd);
                const uint32_t n_chunks = n_run * (uint32_t)R;
                auto issue = [&](uint32_t g, int stg) {                // lane 0
                    const uint32_t bytes = LEAN_CHUNK * sizeof(cf) + odd * 16u;
                    ptx::mbar_arrive_expect_tx_sa(bar0 + 8u * (uint32_t)stg, bytes);
                    ptx::bulk_g2s_sa(ring0 + (uint32_t)stg * (LEAN_SLOT * sizeof(cf)), src0 + (size_t)g * LEAN_CHUNK, bytes,
                                     bar0 + 8u * (uint32_t)stg);
                };
                if (l == 0) {
                    int stg = stage;
                    for (uint32_t g = 0; g < (uint32_t)(S - 1) && g < n_chunks; ++g) { issue(g, stg); stg = stg + 1 == S ? 0 : stg + 1; }
                }
                uint32_t g = 0;
                const uint32_t n_win = n_run + (toff ? 1u : 0u);
#pragma unroll 1
                for (uint32_t w = 0; w < n_win; ++w) {
                    const bool direct = SH && w >= n_run;              // the unshifted window, from global memory
                    const uint32_t s = direct ? (toff > 0 ? total - 1 : 0u) : sym0 + w;
                    cf acc[R];
                    cf e0 = fx::mk(0.0f, 0.0f), e1 = fx::mk(0.0f, 0.0f);
                    cf h1 = fx::mk(1.0f, 0.0f), h2 = h1, h4 = h1, h8 = h1;
                    int k0 = 0;
#pragma unroll
                    for (int q = 0; q < R; ++q) {
                        cf v[16];
                        if (!direct) {
                            __syncwarp();                              // every lane has finished the slot that is refilled now
                            if (l == 0 && g + (uint32_t)(S - 1) < n_chunks) {
                                int stg = stage + (S - 1);
                                if (stg >= S) stg -= S;
                                issue(g + (uint32_t)(S - 1), stg);
                            }
                            ptx::mbar_wait_sa(bar0 + 8u * (uint32_t)stage, (phase_bits >> stage) & 1u);
                            phase_bits ^= 1u << stage;
                            const cf* raw = ring + (size_t)stage * LEAN_SLOT + odd + l;
#pragma unroll
                            for (int m = 0; m < 16; ++m) v[m] = raw[32 * m];
                            stage = stage + 1 == S ? 0 : stage + 1;
                            ++g;
                        } else {
                            const cf* gsrc = pkt + (size_t)s * N + (size_t)q * LEAN_CHUNK + l;
#pragma unroll
                            for (int m = 0; m < 16; ++m) v[m] = ldg_cf(gsrc + 32 * m);
                        }
                        float mw = 0.0f;
#pragma unroll
                        for (int m = 0; m < 16; ++m) {
                            const cf r = v[m];
                            mw = fmaxf(fmaxf(mw, fabsf(r.x)), fabsf(r.y));                     // LoRaDemod.cpp:61-68
                            if (m & 1) e1 = ff::sqacc(e1, r); else e0 = ff::sqacc(e0, r);
                        }
                        mmax = fmaxf(mmax, mw);
                        if (q == 0) {
                            // bin class from the phase step between samples T apart (m and m + R), four pairs per lane
                            cf z = fx::mk(0.0f, 0.0f);
#pragma unroll
                            for (int m = 0; m < 4; ++m) {
                                const cf p1 = v[m + R], p0 = v[m];
                                z.x = ff::fma_(p1.x, p0.x, z.x); z.x = ff::fma_(p1.y, p0.y, z.x);
                                z.y = ff::fma_(p1.y, p0.x, z.y); z.y = ff::fma_(-p1.x, p0.y, z.y);
                            }
#pragma unroll
                            for (int off = 16; off > 0; off >>= 1) {
                                z.x += __shfl_xor_sync(0xffffffffu, z.x, off);
                                z.y += __shfl_xor_sync(0xffffffffu, z.y, off);
                            }
                            z = ff::fcmul(z, geo_g1);
                            const float ax = fabsf(z.x), ay = fabsf(z.y);
                            const float mn = fminf(ax, ay), mx = fmaxf(ax, ay);
                            int qq = (mn > 0.19891237f * mx ? 1 : 0) + (mn > 0.66817864f * mx ? 1 : 0);
                            if (ay > ax) qq = 4 - qq;
                            if (z.x < 0.0f) qq = 8 - qq;
                            if (z.y < 0.0f) qq = 16 - qq;
                            k0 = qq & 15;
                            const float4 hA = *reinterpret_cast<const float4*>(&hp[4 * k0]);
                            const float4 hB = *reinterpret_cast<const float4*>(&hp[4 * k0 + 2]);
                            h1 = fx::mk(hA.x, hA.y); h2 = fx::mk(hA.z, hA.w); h4 = fx::mk(hB.x, hB.y); h8 = fx::mk(hB.z, hB.w);
                        }
                        // inside the chunk: accumulator j gets sum_{c' < CP} v[j + R c'] h^c' by halving
                        if (LV >= 3) {
#pragma unroll
                            for (int i = 0; i < 4 * R; ++i) v[i] = lean_cma(v[i], v[i + 4 * R], h4);
                        }
                        if (LV >= 2) {
#pragma unroll
                            for (int i = 0; i < 2 * R; ++i) v[i] = lean_cma(v[i], v[i + 2 * R], h2);
                        }
#pragma unroll
                        for (int i = 0; i < R; ++i) v[i] = lean_cma(v[i], v[i + R], h1);
                        if (q == 0) {
#pragma unroll
                            for (int j = 0; j < R; ++j) acc[j] = v[j];
                        } else {
                            // h^(q 16 / R) from the table's powers of two (at most two products of accurate factors: a running
                            // product would carry seven roundings into the window's last chunk)
                            const int e = q * (16 / R);                    // 2 .. 14, compile-time
                            cf hq = fx::mk(1.0f, 0.0f);
                            bool first = true;
                            if (e & 2) { hq = h2; first = false; }
                            if (e & 4) { hq = first ? h4 : ff::fcmul(hq, h4); first = false; }
                            if (e & 8) { hq = first ? h8 : ff::fcmul(hq, h8); first = false; }
#pragma unroll
                            for (int j = 0; j < R; ++j) acc[j] = lean_cma(acc[j], v[j], hq);
                        }
                    }
                    // end of the window: thread t = l + 32 j holds acc[j]; twiddle exp(j t theta), theta = rate - 2 pi k0 / N
                    const cf el = ff::fcmul(geo_f, twl[l * k0]);                               // F_l W_N^(l k0)
#pragma unroll
                    for (int j = 0; j < R; ++j) {
                        cf u = ff::fcmul(acc[j], el);
                        if (j > 0) u = ff::fcmul(u, ff::fcmul(g32[j], tw32[j * k0]));          // G32^j W_N^(32 j k0)
                        acc[j] = u;
                    }
                    // R-point DFT across the window's warps-that-were (natural order), cross twiddle, R lane transforms
                    if (R == 2) ff::dft2(acc[0], acc[1]);
                    if (R == 4) ff::dft4(acc[0], acc[1], acc[2], acc[3]);
                    if (R == 8) ff::dft8(acc[0], acc[1], acc[2], acc[3], acc[4], acc[5], acc[6], acc[7]);
#pragma unroll
                    for (int ka = 1; ka < R; ++ka) acc[ka] = ff::fcmul(acc[ka], twx[ka * 32 + l]);
#pragma unroll
                    for (int sgi = 0; sgi < 5; ++sgi) {
                        const int h = 16 >> sgi;
                        const bool last = sgi == 4;
                        const cf tw = last ? fx::mk(1.0f, 0.0f) : tws[sgi * 32 + l];
#pragma unroll
                        for (int ka = 0; ka < R; ++ka) acc[ka] = fp::lane_stage(acc[ka], h, (l & h) != 0, tw, last);
                    }
                    unsigned qb = 0u;
                    int kab = 0;
#pragma unroll
                    for (int ka = 0; ka < R; ++ka) {
                        const unsigned qv = __float_as_uint(ff::fnorm(acc[ka]));              // >= +0, or NaN (sorts above everything)
                        if (qv > qb) { qb = qv; kab = ka; }
                    }
                    unsigned qmax = qb;
                    float en = ((e0.x + e0.y) + (e1.x + e1.y)) * 1.000002f;
#pragma unroll
                    for (int off = 16; off > 0; off >>= 1) {
                        qmax = max(qmax, __shfl_xor_sync(0xffffffffu, qmax, off));
                        en += __shfl_xor_sync(0xffffffffu, en, off);
                    }
                    const float q1 = __uint_as_float(qmax);
                    const unsigned ball = __ballot_sync(0xffffffffu, qb == qmax);
                    const int src = __ffs((int)ball) - 1;
                    const int idx = __shfl_sync(0xffffffffu, k0 + 16 * (kab + R * (int)(__brev((unsigned)l) >> 27)), src);
#pragma unroll
                    for (int off = 16; off > 0; off >>= 1) mmax = fmaxf(mmax, __shfl_xor_sync(0xffffffffu, mmax, off));
                    // accepted iff the top bin holds 9/16 of N E (tier 1 of demod_cert.cu).  The decision does not depend on
                    // the scale the estimate assumed (positive, finite; the rate it changes is re-checked); NaN samples are
                    // invisible to fmaxf but poison E, an infinite max wipes the packet in the reference.  E is taken on the
                    // raw samples: an upper bound of the scaled energy (scale <= 1) is all the test needs... but q1 is raw
                    // too, so both sides carry the same factor.
                    const bool cert = (q1 > tier1 * en) && (en < 1.0e30f) && (q1 >= 8.3e-25f) && (mmax < 3.0e38f);
                    if (!cert) { bad = true; break; }
                    if (s == 0) sw0 = idx;
                    else if (s == 1) sw1 = idx;
                    else if (l == 0) a.symbols[(size_t)pid * a.symbol_stride + (s - 2)] = (uint16_t)idx;
                }
                // samples no window covered (the max-abs is over the whole packet, LoRaDemod.cpp:61-68): the first t_off of
                // the packet, or its last |t_off|
                if constexpr (SH) if (toff != 0 && !bad) {
                    const uint32_t a0 = (uint32_t)(toff > 0 ? toff : -toff);
                    const cf* e = pkt + (toff > 0 ? 0u : count - a0);
                    for (uint32_t i = (uint32_t)l; i < a0; i += 32u) {
                        const cf r = ldg_cf(e + i);
                        mmax = fmaxf(fmaxf(mmax, fabsf(r.x)), fabsf(r.y));
                    }
#pragma unroll
                    for (int off = 16; off > 0; off >>= 1) mmax = fmaxf(mmax, __shfl_xor_sync(0xffffffffu, mmax, off));
                    if (!(mmax < 3.0e38f)) bad = true;
                }
                if (bad) {
                    // abandon the packet: consume the copies still in flight so that the ring's phases stay consistent
                    const uint32_t end = min(n_chunks, g + (uint32_t)(S - 1));
                    for (; g < end; ++g) {
                        ptx::mbar_wait_sa(bar0 + 8u * (uint32_t)stage, (phase_bits >> stage) & 1u);
                        phase_bits ^= 1u << stage;
                        stage = stage + 1 == S ? 0 : stage + 1;
                    }
                    __syncwarp();
                }
                };
                if (mine.t_off != 0 || mine.pad[2] != 0) run(std::true_type{}); else run(std::false_type{});
            }
            } else {
            if (!bad) {
                // SF7-9: GD = 512 / N whole windows per 4 KB step, T = N / 16 <= 32 lanes per window (demod_cert.cu's
                // geometric pruned step, with one bulk copy per step and nothing else in the kernel)
                const int gq = l / T;
                cf geo_f;
                {
                    float sn, cs;
                    __sincosf(fx::fmul(rate, (float)tid), &sn, &cs);   // |rate tid| < 2 pi / 16
                    geo_f = fx::mk(cs, sn);
                    double ds = 0.0, dc = 1.0;
                    if (l < 4) sincos((double)rate * (double)(T << l), &ds, &dc);
                    const float gx = (float)dc, gy = (float)ds;
#pragma unroll
                    for (int h2 = 0; h2 < 2; ++h2) {
                        const int e = l + 32 * h2, k = e >> 2, mi = e & 3;
                        const cf g = fx::mk(__shfl_sync(0xffffffffu, gx, mi), __shfl_sync(0xffffffffu, gy, mi));
                        hp[e] = ff::fcmul(g, w16[(k << mi) & 15]);
                    }
                    __syncwarp();
                }
                const cf geo_g1 = hp[0];
                // Window bases (LoRaDemod.cpp:143-150, count == total N here): with t_off > 0 every window but the last
                // starts t_off later; with t_off < 0 every window but the first starts |t_off| earlier.  The shifted windows
                // are one contiguous run and go through the ring (from the 16-byte boundary below its first sample when
                // that one is odd); the one unshifted window comes straight from global memory afterwards.
                // Two copies of the stream: clean packets (no time offset, sync peaks on the grid) run code in which the offset,
                // the odd-alignment entry, the extra unshifted window and the retry are compile-time absent; the others the
                // general one.
                auto run = [&](auto shifted_tag) {
                constexpr bool SH = decltype(shifted_tag)::value;
                const int toff = SH ? mine.t_off : 0;
                const uint32_t run0 = toff > 0 ? (uint32_t)toff : (toff < 0 ? (uint32_t)(N + toff) : 0u);    // first sample of the run
                const uint32_t n_run = toff ? total - 1 : total;                                          // windows in it
                const uint32_t sym0 = toff < 0 ? 1u : 0u;                                                  // symbol of its first window
                const uint32_t odd = run0 & 1u;
                const cf* src0 = pkt + (run0 - odd);
                const uint32_t n_chunks = (n_run + (uint32_t)GD - 1) / (uint32_t)GD;
                auto issue = [&](uint32_t g, int stg) {                // lane 0: the step's windows are contiguous
                    const uint32_t left = n_run - g * (uint32_t)GD;
                    const uint32_t bytes = (left < (uint32_t)GD ? left : (uint32_t)GD) * (uint32_t)(N * sizeof(cf)) + odd * 16u;
                    ptx::mbar_arrive_expect_tx_sa(bar0 + 8u * (uint32_t)stg, bytes);
                    ptx::bulk_g2s_sa(ring0 + (uint32_t)stg * (LEAN_SLOT * sizeof(cf)), src0 + (size_t)g * LEAN_CHUNK, bytes,
                                     bar0 + 8u * (uint32_t)stg);
                };
                if (l == 0) {
                    int stg = stage;
                    for (uint32_t g = 0; g < (uint32_t)(S - 1) && g < n_chunks; ++g) { issue(g, stg); stg = stg + 1 == S ? 0 : stg + 1; }
                }
                uint32_t g = 0;
                // one more step after the run when t_off != 0: the unshifted window (group 0 of the warp), from global memory
                const uint32_t n_steps = n_chunks + (toff ? 1u : 0u);
#pragma unroll 1
                for (uint32_t step = 0; step < n_steps && !bad; ++step) {
                    const bool direct = SH && step >= n_chunks;
                    cf v[16];
                    uint32_t s;
                    bool active;
                    if (!direct) {
                        __syncwarp();                                  // every lane has finished the slot that is refilled now
                        if (l == 0 && g + (uint32_t)(S - 1) < n_chunks) {
                            int stg = stage + (S - 1);
                            if (stg >= S) stg -= S;
                            issue(g + (uint32_t)(S - 1), stg);
                        }
                        ptx::mbar_wait_sa(bar0 + 8u * (uint32_t)stage, (phase_bits >> stage) & 1u);
                        phase_bits ^= 1u << stage;
                        const uint32_t w = g * (uint32_t)GD + (uint32_t)gq;
                        s = sym0 + w;
                        active = w < n_run;
                        const cf* raw = ring + (size_t)stage * LEAN_SLOT + odd + gq * N + tid;
#pragma unroll
                        for (int c = 0; c < 16; ++c) v[c] = raw[T * c];
                        stage = stage + 1 == S ? 0 : stage + 1;
                        ++g;
                    } else {
                        s = toff > 0 ? total - 1 : 0u;
                        active = gq == 0;
                        const cf* gsrc = pkt + (size_t)s * N + tid;
#pragma unroll
                        for (int c = 0; c < 16; ++c) v[c] = active ? ldg_cf(gsrc + T * c) : fx::mk(0.0f, 0.0f);
                    }
                    float mw = 0.0f;
                    cf e0 = fx::mk(0.0f, 0.0f), e1 = fx::mk(0.0f, 0.0f);
#pragma unroll
                    for (int c = 0; c < 16; ++c) {
                        const cf r = v[c];
                        mw = fmaxf(fmaxf(mw, fabsf(r.x)), fabsf(r.y));                         // LoRaDemod.cpp:61-68
                        if (c & 1) e1 = ff::sqacc(e1, r); else e0 = ff::sqacc(e0, r);
                    }
                    if (active) mmax = fmaxf(mmax, mw);
                    float en = ((e0.x + e0.y) + (e1.x + e1.y)) * 1.000002f;
                    // bin class from the phase step between neighbouring samples of a thread (T apart), four pairs per thread
                    cf z = fx::mk(0.0f, 0.0f);
#pragma unroll
                    for (int c = 0; c < 16; c += 4) {
                        z.x = ff::fma_(v[c + 1].x, v[c].x, z.x); z.x = ff::fma_(v[c + 1].y, v[c].y, z.x);
                        z.y = ff::fma_(v[c + 1].y, v[c].x, z.y); z.y = ff::fma_(-v[c + 1].x, v[c].y, z.y);
                    }
#pragma unroll
                    for (int off = T / 2; off > 0; off >>= 1) {
                        z.x += __shfl_xor_sync(0xffffffffu, z.x, off);
                        z.y += __shfl_xor_sync(0xffffffffu, z.y, off);
                    }
                    z = ff::fcmul(z, geo_g1);
                    int k0;
                    {
                        const float ax = fabsf(z.x), ay = fabsf(z.y);
                        const float mn = fminf(ax, ay), mx = fmaxf(ax, ay);
                        int qq = (mn > 0.19891237f * mx ? 1 : 0) + (mn > 0.66817864f * mx ? 1 : 0);
                        if (ay > ax) qq = 4 - qq;
                        if (z.x < 0.0f) qq = 8 - qq;
                        if (z.y < 0.0f) qq = 16 - qq;
                        k0 = qq & 15;
                    }
                    // The class is judged from 4 x T sample pairs; in noise it can land one sector off.  Any class that passes
                    // the test is right (a wrong one cannot hold 9/16 of the energy), so a refused window is tried again with
                    // the two neighbouring classes before the packet is given up -- from the samples still in the slot.
                    float en_w = en;
#pragma unroll
                    for (int off = T / 2; off > 0; off >>= 1) en_w += __shfl_xor_sync(0xffffffffu, en_w, off);
                    bool settled = !active;
                    auto settle = [&](int kk) {                        // the pruned transform of class kk on v (consumed)
                        const float4 hA = *reinterpret_cast<const float4*>(&hp[4 * kk]);
                        const float4 hB = *reinterpret_cast<const float4*>(&hp[4 * kk + 2]);
                        const cf h1 = fx::mk(hA.x, hA.y), h2 = fx::mk(hA.z, hA.w), h4 = fx::mk(hB.x, hB.y), h8 = fx::mk(hB.z, hB.w);
#pragma unroll
                        for (int c = 0; c < 8; ++c) v[c] = lean_cma(v[c], v[c + 8], h8);
#pragma unroll
                        for (int c = 0; c < 4; ++c) v[c] = lean_cma(v[c], v[c + 4], h4);
#pragma unroll
                        for (int c = 0; c < 2; ++c) v[c] = lean_cma(v[c], v[c + 2], h2);
                        const cf pv = lean_cma(v[0], v[1], h1);
                        const cf cval = ff::fcmul(ff::fcmul(pv, twl[(tid * kk) & (N - 1)]), geo_f);
                        const cf yv = fp::lane_dft<SF>(cval, tid, tws);
                        const unsigned qb = __float_as_uint(ff::fnorm(yv));           // >= +0, or NaN (sorts above everything)
                        unsigned qmax = qb;
#pragma unroll
                        for (int off = T / 2; off > 0; off >>= 1) qmax = max(qmax, __shfl_xor_sync(0xffffffffu, qmax, off));
                        const float q1 = __uint_as_float(qmax);
                        const bool cert = (q1 > tier1 * en_w) && (en_w < 1.0e30f) && (q1 >= 8.3e-25f);
                        if (!settled && cert) {
                            if (qb == qmax) {                              // the lane that holds the top bin (unique when certified)
                                const int idx = kk + 16 * (int)(__brev((unsigned)tid) >> (32 - PR::LW));
                                if (s < 2) mycell->pad[s] = idx;
                                else a.symbols[(size_t)pid * a.symbol_stride + (s - 2)] = (uint16_t)idx;
                            }
                            settled = true;
                        }
                    };
                    settle(k0);
                    // The class is judged from 4 x T sample pairs; in noise it can land one sector off.  Any class that passes
                    // the test is right (a wrong one cannot hold 9/16 of the energy), so a refused window is tried again with
                    // the two neighbouring classes before the packet is given up -- from the samples still in the slot.
                    if (SH && LEAN_RETRY && __builtin_expect(__any_sync(0xffffffffu, !settled) != 0, 0)) {
                        const int prev = stage == 0 ? S - 1 : stage - 1;   // the slot this step consumed (not refilled before the next step)
                        const cf* vsrc = direct ? pkt + (size_t)s * N + tid : ring + (size_t)prev * LEAN_SLOT + odd + gq * N + tid;
#pragma unroll 1
                        for (int att = 1; att < 3; ++att) {
#pragma unroll
                            for (int c = 0; c < 16; ++c)
                                v[c] = (direct && !active) ? fx::mk(0.0f, 0.0f) : (direct ? ldg_cf(vsrc + T * c) : vsrc[T * c]);
                            settle((k0 + (att == 1 ? 1 : 15)) & 15);
                            if (!__any_sync(0xffffffffu, !settled)) break;
                        }
                    }
                    if (__any_sync(0xffffffffu, !settled)) bad = true;
                }
                // samples no window covered (the max-abs is over the whole packet, LoRaDemod.cpp:61-68): the first t_off of
                // the packet, or its last |t_off|
                if constexpr (SH) if (toff != 0 && !bad) {
                    const uint32_t a0 = (uint32_t)(toff > 0 ? toff : -toff);
                    const cf* e = pkt + (toff > 0 ? 0u : count - a0);
                    for (uint32_t i = (uint32_t)l; i < a0; i += 32u) {
                        const cf r = ldg_cf(e + i);
                        mmax = fmaxf(fmaxf(mmax, fabsf(r.x)), fabsf(r.y));
                    }
                }
#pragma unroll
                for (int off = 16; off > 0; off >>= 1) mmax = fmaxf(mmax, __shfl_xor_sync(0xffffffffu, mmax, off));
                if (!(mmax < 3.0e38f)) bad = true;                         // an infinite max wipes the packet in the reference
                __syncwarp();
                sw0 = mycell->pad[0]; sw1 = mycell->pad[1];
                if (bad) {
                    // abandon the packet: consume the copies still in flight so that the ring's phases stay consistent
                    const uint32_t end = min(n_chunks, g + (uint32_t)(S - 1));
                    for (; g < end; ++g) {
                        ptx::mbar_wait_sa(bar0 + 8u * (uint32_t)stage, (phase_bits >> stage) & 1u);
                        phase_bits ^= 1u << stage;
                        stage = stage + 1 == S ? 0 : stage + 1;
                    }
                    __syncwarp();
                }
                };
                if (mine.t_off != 0 || mine.pad[2] != 0) run(std::true_type{}); else run(std::false_type{});
            }
            }
            if (bad) {
                if (l == 0) { a.flagged[pid] = 1; ++st_flagged; }
            } else {
                if (l == 0) {
                    st_pruned += total;
                    if (a.sync) a.sync[pid] = (uint8_t)((((sw0 >> (SF - 4)) & 0xf) << 4) | ((sw1 >> (SF - 4)) & 0xf));
                    if (a.status) a.status[pid] = (int)(total - 2);
                }
                // the estimate assumed max-abs used_max (LoRaDemod.cpp:61-78): right iff the packet's own max says the same
                const bool hit = (mmax > 1.0f) ? (mmax == mine.used_max) : !(mine.used_max > 1.0f);
                if (!hit) { vpid[job] = (int)pid; vmax[job] = mmax; vrate[job] = rate; vtoff[job] = mine.t_off; }
                pred_max = mmax;
            }
        }
        }                                                      // jobs of the round
    }
    if (a.cert_stats && l == 0) {
        if (st_pruned) { atomicAdd(a.cert_stats + 0, (unsigned long long)st_pruned); atomicAdd(a.cert_stats + 4, (unsigned long long)st_pruned); }
        if (st_flagged) atomicAdd(a.cert_stats + 5, (unsigned long long)st_flagged);
    }
    if (a.n_flagged && st_flagged) atomicAdd(a.n_flagged, st_flagged);
    if (a.cert_stats && lead && st_est2) atomicAdd(a.cert_stats + 2, (unsigned long long)st_est2);
}

template <int SF>
static cudaError_t launch_lean_sf(const DemodArgs& a, int sm_count, cudaStream_t stream)
{
    using C = LeanCfg<SF>;
    static std::atomic<bool> attr_done[64];
    int dev = 0;
    cudaGetDevice(&dev);
    if (!attr_done[dev & 63].load(std::memory_order_acquire)) {
        cudaFuncSetAttribute(demod_lean_kernel<SF>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::bytes);
        attr_done[dev & 63].store(true, std::memory_order_release);
    }
    const unsigned max_ctas = (unsigned)sm_count * C::CTAS;               // persistent grid
    const unsigned need = (a.n_packets + C::WARPS - 1) / C::WARPS;
    demod_lean_kernel<SF><<<need < max_ctas ? need : max_ctas, C::CTA, C::bytes, stream>>>(a);
    return cudaGetLastError();
}

bool lean_supported(int sf, const DemodArgs& a)
{
    return sf >= 7 && sf <= 12 && a.osr == 1 && a.mode == MODE_LEGACY && !a.window && !a.down && !a.no_scratch &&
           !a.force_exact && (unsigned long long)a.samples_per_packet < 0xf0000000ull;
}

cudaError_t launch_demod_lean(int sf, const DemodArgs& a, int sm_count, cudaStream_t stream)
{
    switch (sf) {
    case 7: return launch_lean_sf<7>(a, sm_count, stream);
    case 8: return launch_lean_sf<8>(a, sm_count, stream);
    case 9: return launch_lean_sf<9>(a, sm_count, stream);
    case 10: return launch_lean_sf<10>(a, sm_count, stream);
    case 11: return launch_lean_sf<11>(a, sm_count, stream);
    case 12: return launch_lean_sf<12>(a, sm_count, stream);
    default: return cudaErrorInvalidValue;
    }
}

} // namespace lk
