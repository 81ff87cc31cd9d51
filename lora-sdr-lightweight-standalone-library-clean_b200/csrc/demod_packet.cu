// demod_packet.cu -- the whole demodulator in ONE launch, one CTA per packet (SF 7..12, osr == 1):
// lora_demodulate (LoRaDemod.cpp:50-197) and demodulate (phy.cpp:182-243) for a batch.
//
// The reference walks a packet three times: max-abs scan of every sample (:61-68), the
// two-symbol CFO/STO estimator (:80-136), then the per-symbol loop (:142-176) whose inputs
// (1/max scale, rate, t_off) come from the first two.  Here a persistent CTA owns one packet at
// a time and runs the three walks back to back as three phases:
//
//   A  scan      all 256 threads stream the packet with 128-bit loads -> max(|re|,|im|) -> scale
//   B  estimate  symbols 0 and 1 in parallel on two thread groups (full-twiddle exact FFT, peak,
//                fIndex, atan2f phase), combined in symbol order by thread 0 -> cfo, time_offset,
//                rate, t_off
//   C  symbols   every symbol window of the packet (sync symbols included) through the streaming
//                pipelines of demod_stream.cu: bulk asynchronous copy global -> shared one unit
//                ahead, normalise -> [downchirp] -> CFO rotation -> [Hann], exact-order FFT in
//                place, first-max argmax -> uint16
//
// Why this shape: the per-sample chain is issue-bound (glibc-exact sincosf in FP64), the scan is
// HBM-bound.  Three CTAs are resident per SM and drift out of step, so one CTA's scan streams
// from HBM while its neighbours' symbol phases fill the issue slots -- the scan costs no time of
// its own.  Phase A leaves the packet in L2 (a CTA's packet x 444 resident CTAs is 30 MB at SF7 /
// 66 symbols, far below the 126 MB L2), so phases B and C are served by L2 and HBM sees every IQ
// byte once; packets too long for that (SF12: 4.3 MB each) read HBM twice, which the issue-bound
// phase C still hides.
#include <atomic>
#include "stream_common.cuh"
#include <cstdlib>
#include "internal.h"

namespace lk {

struct EstSym { int idx; float fi; float phase; uint32_t t; };

struct PktShared {
    float redf[8];
    EstSym est[2];
    PacketState st;
    int sw[2];
    int status;
    uint32_t run;
};

template <int SF>
struct PacketCfg : StreamCfg<SF> {
    using B = StreamCfg<SF>;
    static constexpr size_t sh_off = (B::bytes + 15) & ~(size_t)15;
    static constexpr size_t bytes = sh_off + sizeof(PktShared);
};

// max(m, |a|, |b|) that returns NaN when any input is NaN (sm_100 three-input max; the abs fold into the operands)
__device__ __forceinline__ float max3_nan_abs(float m, float a, float b)
{
    float d;
    asm("{\n\t.reg .f32 aa, bb;\n\tabs.f32 aa, %2;\n\tabs.f32 bb, %3;\n\tmax.NaN.f32 %0, %1, aa, bb;\n\t}" : "=f"(d) : "f"(m), "f"(a), "f"(b));
    return d;
}

__device__ __forceinline__ float amp_of(float re, float im)      // LoRaDemod.cpp:63-66
{
    const float r = fabsf(re), i = fabsf(im);
    return (r < i) ? i : r;
}

template <int SF, bool EXTRA>
__global__ void __launch_bounds__(CTA_THREADS, STREAM_CTAS_PER_SM) demod_packet_kernel(const DemodArgs a)
{
    using C = PacketCfg<SF>;
    using P = fx::Plan<SF>;
    constexpr int T = C::T, N = C::N, GD = C::GD;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int dom = threadIdx.x / C::DOM_THREADS;
    const int dt = threadIdx.x % C::DOM_THREADS;
    const int gq = dt / T, tid = dt % T;
    const int g = threadIdx.x / T;                 // CTA-wide group number (phase B)

    cf* slots = reinterpret_cast<cf*>(smem_raw) + (size_t)dom * C::STAGES * C::SLOT;
    WinDesc* descs = reinterpret_cast<WinDesc*>(smem_raw + C::slot_bytes) + (size_t)dom * C::STAGES * GD;
    uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw + C::slot_bytes + C::desc_bytes) + dom * C::STAGES;
    Peak* red = reinterpret_cast<Peak*>(smem_raw + C::slot_bytes + C::desc_bytes + C::bar_bytes);
    PktShared& sh = *reinterpret_cast<PktShared*>(smem_raw + C::sh_off);

    if (dt == 0) {
#pragma unroll
        for (int s = 0; s < C::STAGES; ++s) ptx::mbar_init(&full[s], GD);
        ptx::fence_mbar_init();
    }
    __syncthreads();

    int stage = 0;
    unsigned phase_bits = 0;                       // bit s = parity the next wait on stage s expects

    for (uint32_t pid = blockIdx.x; pid < a.n_packets; pid += gridDim.x) {
        const cf* pkt = a.iq + (size_t)pid * a.packet_stride;
        const uint32_t count = a.sample_counts ? min(a.sample_counts[pid], a.samples_per_packet) : a.samples_per_packet;
        const uint32_t total = count / (uint32_t)N;
        const bool have_sync = total >= 2;

        // ---- phase A: max-abs scan, LoRaDemod.cpp:61-78 (legacy path only) ----------------------
        float scale = 1.0f;
        int flags = 0, status = 0;
        if (a.mode == MODE_LEGACY) {
            float m = 0.0f;
            if ((reinterpret_cast<uintptr_t>(pkt) & 15) == 0) {
                // Fast scan: one NaN-propagating 3-input max per sample (|re|, |im| are operand modifiers).  The
                // reference ignores a sample whose real part is NaN and takes |re| when only the imaginary part is
                // (std::max / '>' semantics, LoRaDemod.cpp:63-66); a NaN anywhere poisons m here, and the CTA then
                // rescans with the literal comparison chain below.
                const float4* p4 = reinterpret_cast<const float4*>(pkt);
                const uint32_t pairs = count >> 1;
#pragma unroll 8
                for (uint32_t j = threadIdx.x; j < pairs; j += CTA_THREADS) {
                    const float4 q = __ldg(p4 + j);
                    m = max3_nan_abs(m, q.x, q.y);
                    m = max3_nan_abs(m, q.z, q.w);
                }
                if ((count & 1) && threadIdx.x == 0) {
                    const cf x = ldg_cf(pkt + count - 1);
                    m = max3_nan_abs(m, x.x, x.y);
                }
                if (__syncthreads_or(m != m)) {
                    m = 0.0f;
#pragma unroll 4
                    for (uint32_t j = threadIdx.x; j < count; j += CTA_THREADS) {
                        const cf x = ldg_cf(pkt + j);
                        m = fmaxf(m, amp_of(x.x, x.y));
                    }
                }
            } else {
#pragma unroll 4
                for (uint32_t j = threadIdx.x; j < count; j += CTA_THREADS) {
                    const cf x = ldg_cf(pkt + j);
                    m = fmaxf(m, amp_of(x.x, x.y));
                }
            }
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, off));
            if ((threadIdx.x & 31) == 0) sh.redf[threadIdx.x >> 5] = m;
            __syncthreads();
            m = sh.redf[0];
#pragma unroll
            for (int k = 1; k < CTA_THREADS / 32; ++k) m = fmaxf(m, sh.redf[k]);
            if (m > 1.0f) {
                if (a.no_scratch) { flags |= PS_SKIP; status = -34; /* -ERANGE, LoRaDemod.cpp:70-72 */ }
                else { flags |= PS_NORM; scale = __fdiv_rn(1.0f, m); }
            }
        } else {
            // phy.cpp:190-194 (argument checks; per packet when lengths are ragged)
            if (count % (uint32_t)N != 0) { flags |= PS_SKIP; status = -22; }
            else if (total < 2) { flags |= PS_SKIP; status = -34; }
        }
        const bool run = !(flags & PS_SKIP);
        const bool norm = (flags & PS_NORM) != 0;
        const uint32_t n_est = run ? (total < 2 ? total : 2u) : 0u;

        // ---- phase B: estimator, LoRaDemod.cpp:80-136 / phy.cpp:93-147 (osr == 1) ------------------
        // symbol s is handled by thread group s; groups own their window slots of stage 0 / 1 as
        // exchange buffers (the ring is idle: no bulk copy is in flight between packets)
        const bool est_here = (T > 32) ? (n_est > 0) : (n_est > 0 && (int)(threadIdx.x / 32) * 32 / T < 2);
        if (est_here) {
            cf* buf0 = slots + (size_t)gq * C::PITCH;
            cf* buf1 = slots + (size_t)C::SLOT + (size_t)gq * C::PITCH;
            int flip = 0;
            constexpr int G = CTA_THREADS / T;
            for (uint32_t s0 = 0; s0 < n_est; s0 += G) {       // G == 1 (SF12): two rounds
                const uint32_t s = s0 + g;
                const bool live = s < n_est;
                cf v[P::R];
                load_window<SF>(v, pkt + (size_t)(live ? s : 0) * N, tid, 1, live, norm, scale, nullptr, false, 0.0f, 0.0f,
                                a.window);
                fft_group<SF, true>(v, tid, a.tw, a.ct, buf0, buf1, flip);
                const Peak pk = group_peak<SF>(local_peak<SF>(v, tid), red);
                cf* spec = flip ? buf1 : buf0;
                flip ^= 1;
#pragma unroll
                for (int r = 0; r < P::R; ++r) spec[P::bin_of_reg(tid, r)] = v[r];
                group_sync<T>();
                if (tid == 0 && live) {
                    const float fundamental = __fsqrt_rn(pk.val);
                    const float p = fx::fsub(fx::fmul(20.0f, gm::log10f_g(fundamental)), a.power_scale);
                    EstSym e; e.idx = 0; e.fi = 0.0f; e.phase = 0.0f; e.t = 0;
                    cf best_bin = fx::mk(0.0f, 0.0f);
                    if (p > -1e30f) {                              // LoRaDemod.cpp:101-108 with a single phase t = 0
                        const cf l = spec[pk.idx > 0 ? pk.idx - 1 : N - 1];
                        const cf r = spec[pk.idx < N - 1 ? pk.idx + 1 : 0];
                        const float left = gm::hypotf_g(l.x, l.y), right = gm::hypotf_g(r.x, r.y);
                        const double demon = (2.0 * (double)fundamental) - (double)right - (double)left;
                        e.fi = (demon == 0.0) ? 0.0f : (float)(0.5 * (double)fx::fsub(right, left) / demon);
                        e.idx = pk.idx;
                        best_bin = spec[pk.idx];
                    }
                    e.phase = gm::atan2f_g(best_bin.y, best_bin.x);
                    sh.est[s] = e;
                }
                group_sync<T>();                                   // spectrum read before the buffer is reused
            }
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            PacketState st; st.scale = scale; st.rate = 0.0f; st.t_off = 0; st.flags = flags;
            if (run) {
                const float PI_F = 3.14159265358979323846f;
                float sum_index = 0.0f, phase_diff = 0.0f, prev_phase = 0.0f;
                unsigned sum_t = 0;
                for (uint32_t s = 0; s < n_est; ++s) {
                    const EstSym e = sh.est[s];
                    sum_t += e.t;
                    sum_index = fx::fadd(sum_index, fx::fadd((float)e.idx, e.fi));
                    if (s > 0) {
                        float d = fx::fsub(e.phase, prev_phase);
                        while (d > PI_F) d = fx::fsub(d, 2.0f * PI_F);
                        while (d < -PI_F) d = fx::fadd(d, 2.0f * PI_F);
                        phase_diff = fx::fadd(phase_diff, d);
                    }
                    prev_phase = e.phase;
                }
                const float avg_index = __fdiv_rn(sum_index, (float)n_est);
                const float cfo_coarse = __fdiv_rn(avg_index, (float)N);
                float cfo_fine = 0.0f;
                if (n_est > 1)
                    cfo_fine = __fdiv_rn(__fdiv_rn(phase_diff, (float)(n_est - 1)), fx::fmul(2.0f * PI_F, (float)N));
                const float cfo = fx::fadd(cfo_coarse, cfo_fine);
                const float frac = fx::fsub(avg_index, floorf(fx::fadd(avg_index, 0.5f)));
                const float avg_t = __fdiv_rn((float)sum_t, (float)n_est);
                const float toff = fx::fsub(avg_t, fx::fmul(fx::fmul(frac, (float)N), 1.0f));
                if (a.cfo) a.cfo[pid] = cfo;
                if (a.time_offset) a.time_offset[pid] = toff;
                st.t_off = (int)roundf(toff);
                st.rate = __fdiv_rn(fx::fmul(-2.0f * PI_F, cfo), (float)N);
                status = have_sync ? (int)(total - 2) : (int)total;
            }
            sh.st = st;
            sh.sw[0] = 0; sh.sw[1] = 0;
            sh.status = status;
            sh.run = run ? 1u : 0u;
        }
        ptx::fence_proxy_async();      // phase-B generic-proxy writes to the ring precede phase-C bulk copies
        __syncthreads();

        // ---- phase C: every symbol of the packet, LoRaDemod.cpp:142-176 / phy.cpp:204-238 -----------
        if (run) {
            const PacketState st = sh.st;
            const uint32_t n_units = (total + GD - 1) / GD;

            auto prefetch = [&](uint32_t u, int stg) {
                if (tid != 0) return;
                WinDesc d;
                d.scale = st.scale; d.rate = st.rate; d.start = 0.0f; d.flags = 0; d.out_off = 0; d.aux0 = 0; d.aux1 = 0;
                const uint32_t s = u * GD + gq;
                cf* win = slots + (size_t)stg * C::SLOT + (size_t)gq * C::PITCH;
                if (s < total) {
                    const uint32_t base = shifted_base(s, (uint32_t)N, st.t_off, count);   // LoRaDemod.cpp:143-150
                    const cf* src = pkt + base;
                    d.start = start_phase(st.rate, s, (uint32_t)N, st.t_off);
                    d.flags = WD_ACTIVE;
                    if (phases_need_slow_path(d.start, fx::fadd(d.start, fx::fmul(st.rate, (float)(N - 1))))) d.flags |= WD_SLOW;
                    if (window_is_odd(src)) d.flags |= WD_ODD;
                    d.aux0 = s;
                    descs[stg * GD + gq] = d;
                    fetch_window<N>(win, src, &full[stg]);
                } else {
                    descs[stg * GD + gq] = d;
                    ptx::mbar_arrive(&full[stg]);
                }
            };

            uint32_t u = (uint32_t)dom;
            if (u < n_units) prefetch(u, stage);
            while (u < n_units) {
                const uint32_t u_next = u + C::DOMS;
                if (u_next < n_units) prefetch(u_next, stage ^ 1);

                ptx::mbar_wait(&full[stage], (phase_bits >> stage) & 1u);
                phase_bits ^= 1u << stage;

                const WinDesc d = descs[stage * GD + gq];
                cf* win = slots + (size_t)stage * C::SLOT + (size_t)gq * C::PITCH;
                const bool active = (d.flags & WD_ACTIVE) != 0;
                cf v[P::R];
                // an inactive window (past the packet's last symbol) runs the same code on whatever its slot
                // holds: the result is discarded, and no register zeroing / branch sits in the hot path
                chain_window<SF, EXTRA>(v, win + ((d.flags & WD_ODD) ? 1 : 0) + tid, tid, d.scale, d.rate, d.start,
                                        (d.flags & WD_SLOW) != 0, a.down, a.window);
                fft_inplace<SF, false>(v, tid, a.tw, a.ct, win);
                const Peak pk = group_peak<SF>(local_peak<SF>(v, tid), red);
                if (tid == 0 && active) {
                    const uint32_t s = d.aux0;
                    if (have_sync && s < 2) sh.sw[s] = pk.idx;
                    else a.symbols[(size_t)pid * a.symbol_stride + (have_sync ? s - 2 : s)] = (uint16_t)pk.idx;
                }
                ptx::fence_proxy_async();     // our generic-proxy stores to the slot precede the next bulk copy into it
                group_sync<T>();              // the slot is this group's own: groups stay coupled only through the stage mbarrier
                u = u_next;
                stage ^= 1;
            }
        }
        __syncthreads();

        // ---- bookkeeping: LoRaDemod.cpp:166-196 ----------------------------------------------------------
        if (threadIdx.x == 0) {
            if (a.sync) {
                const unsigned shift = SF > 4 ? SF - 4 : 0;
                a.sync[pid] = (run && have_sync) ? (uint8_t)((((sh.sw[0] >> shift) & 0xf) << 4) | ((sh.sw[1] >> shift) & 0xf)) : 0;
            }
            if (a.status) a.status[pid] = sh.status;
        }
        __syncthreads();                       // sh and the ring are free for the next packet
    }
}

// ---------------------------------------------------------------------------------------------
// SF 7..11: one PIPELINE per packet (a warp for SF7-9, the window's 2 / 4 warps for SF10 / 11), single pass,
// speculative scale.
//
// A window needs T = N/16 <= 32 threads, so a warp is a complete pipeline (demod_stream.cu) and can
// own a packet without any CTA-level barrier.  The reference's first walk (max-abs, LoRaDemod.cpp
// :61-68) only matters when some |re| or |im| exceeds 1: the warp therefore demodulates the packet
// assuming it does not (scale = 1, which multiplies exactly), tracks the running max of the samples
// it streams anyway, and only if the max turns out to exceed 1 walks the packet a second time with
// the now known 1/max (clean unit-amplitude IQ: about 1 % of SF7 packets, whose second walk is
// served by L2).  HBM sees every IQ byte of a packet once; the estimator's serial epilogue stalls
// one warp of 24 per SM, not a CTA.
//   E units  the first min(total, 2) symbols, raw: full-twiddle exact FFT in place, peak, fIndex,
//            phase -> cfo, time_offset, rate, t_off (LoRaDemod.cpp:80-136 / phy.cpp:93-147)
//   C units  every symbol of the packet, sync symbols included (LoRaDemod.cpp:142-176)
// ---------------------------------------------------------------------------------------------
struct WarpShared { EstSym est[2]; int sw[2]; float bc_rate; int bc_toff; };

// A pipeline (the threads that own one packet at a time) is a warp for SF7-9 (T <= 32: GD = 32/T windows per
// step) and the T = 64 / 128 threads of one window for SF10 / 11 (GD = 1, 4 / 2 pipelines per CTA, each with its
// own named barrier, ring and mbarriers), so packets never wait on each other at CTA level.
template <int SF>
struct WarpCfg {
    using P = fx::Plan<SF>;
    static constexpr int T = P::T, N = P::N;
    static constexpr int DOM_THREADS = T <= 32 ? 32 : T;
    static constexpr int DOMS = CTA_THREADS / DOM_THREADS;
    static constexpr int GD = DOM_THREADS / T;
    static constexpr int PITCH = N + 8;
    static constexpr int SLOT = GD * PITCH;
    static constexpr int STAGES = 2;
    static constexpr size_t slot_bytes = (size_t)DOMS * STAGES * SLOT * sizeof(cf);
    static constexpr size_t desc_bytes = (size_t)DOMS * STAGES * GD * sizeof(WinDesc);
    static constexpr size_t bar_bytes = (size_t)DOMS * STAGES * sizeof(uint64_t);
    static constexpr size_t red_bytes = 8 * sizeof(Peak) + 8 * sizeof(float);
    static constexpr size_t ws_off = (slot_bytes + desc_bytes + bar_bytes + red_bytes + 15) & ~(size_t)15;
    static constexpr size_t bytes = ws_off + (size_t)DOMS * sizeof(WarpShared);
};

// ---- pipeline-wide primitives: warp intrinsics for DT == 32, the pipeline's named barrier otherwise ----------
template <int DT>
__device__ __forceinline__ void pipe_sync()
{
    if (DT <= 32) __syncwarp(); else group_sync<DT>();
}

template <int DT>
__device__ __forceinline__ bool pipe_any(bool p)
{
    if (DT <= 32) return __any_sync(0xffffffffu, p) != 0;
    if (DT >= CTA_THREADS) return __syncthreads_or(p) != 0;
    int r;
    asm volatile("{\n\t.reg .pred q, r;\n\tsetp.ne.b32 q, %1, 0;\n\tbarrier.cta.red.or.pred r, %2, %3, q;\n\tselp.b32 %0, 1, 0, r;\n\t}"
                 : "=r"(r) : "r"((int)p), "r"(1 + (int)threadIdx.x / DT), "n"(DT) : "memory");
    return r != 0;
}

__device__ __forceinline__ float warp_max(float m)
{
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, off));
    return m;
}

template <int DT>
__device__ __forceinline__ float pipe_max(float m, float* redf)
{
    if (DT <= 32) return warp_max(m);
    return group_max<DT>(m, redf);
}

// max over [0, count) of max(|re|,|im|) with one pipeline: 128-bit loads, eight in flight per thread
template <int DT>
__device__ __forceinline__ float pipe_scan_max(const cf* __restrict__ pkt, uint32_t count, int lane, float* redf)
{
    float m = 0.0f;
    uint32_t head = 0;
    if ((reinterpret_cast<uintptr_t>(pkt) & 15) != 0 && count > 0) {        // odd start: peel one sample
        if (lane == 0) { const cf x = ldg_cf(pkt); m = fmaxf(m, amp_of(x.x, x.y)); }
        head = 1;
    }
    const float4* p4 = reinterpret_cast<const float4*>(pkt + head);
    const uint32_t pairs = (count - head) >> 1;
#pragma unroll 8
    for (uint32_t j = lane; j < pairs; j += DT) {
        const float4 q = __ldg(p4 + j);
        m = fmaxf(m, amp_of(q.x, q.y));
        m = fmaxf(m, amp_of(q.z, q.w));
    }
    if (((count - head) & 1) && lane == 0) { const cf x = ldg_cf(pkt + count - 1); m = fmaxf(m, amp_of(x.x, x.y)); }
    return pipe_max<DT>(m, redf);
}

template <int SF, bool EXTRA>
__global__ void __launch_bounds__(CTA_THREADS, STREAM_CTAS_PER_SM) demod_warp_kernel(const DemodArgs a)
{
    using C = WarpCfg<SF>;
    using P = fx::Plan<SF>;
    constexpr int T = C::T, N = C::N, GD = C::GD;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    constexpr int DT = C::DOM_THREADS;
    const int dom = threadIdx.x / DT;              // pipeline index inside the CTA
    const int lane = threadIdx.x % DT;             // thread index inside the pipeline
    const int gq = lane / T, tid = lane % T;

    cf* slots = reinterpret_cast<cf*>(smem_raw) + (size_t)dom * C::STAGES * C::SLOT;
    WinDesc* descs = reinterpret_cast<WinDesc*>(smem_raw + C::slot_bytes) + (size_t)dom * C::STAGES * GD;
    uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw + C::slot_bytes + C::desc_bytes) + dom * C::STAGES;
    Peak* red = reinterpret_cast<Peak*>(smem_raw + C::slot_bytes + C::desc_bytes + C::bar_bytes);   // unused for T <= 32
    float* redf = reinterpret_cast<float*>(red + 8);
    WarpShared& ws = reinterpret_cast<WarpShared*>(smem_raw + C::ws_off)[dom];

    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < C::STAGES; ++s) ptx::mbar_init(&full[s], GD);
        ptx::fence_mbar_init();
    }
    pipe_sync<DT>();

    int stage = 0;
    unsigned phase_bits = 0;
    const uint32_t n_warps = gridDim.x * C::DOMS;
    constexpr int DISTRUST_PACKETS = 4;
    float pred_max = 0.0f;                         // max-abs of the last packet this warp saw
    int distrust = 0;                              // packets left to scan first after a misprediction

    for (uint32_t pid = blockIdx.x * C::DOMS + dom; pid < a.n_packets; pid += n_warps) {
        const cf* pkt = a.iq + (size_t)pid * a.packet_stride;
        const uint32_t count = a.sample_counts ? min(a.sample_counts[pid], a.samples_per_packet) : a.samples_per_packet;
        const uint32_t total = count / (uint32_t)N;
        const bool have_sync = total >= 2;
        const bool legacy = a.mode == MODE_LEGACY;

        int flags = 0, status = 0;
        if (!legacy) {                                     // phy.cpp:190-194
            if (count % (uint32_t)N != 0) { flags |= PS_SKIP; status = -22; }
            else if (total < 2) { flags |= PS_SKIP; status = -34; }
        }
        float cfo = 0.0f, toff = 0.0f;
        bool have_metrics = false;

        // Which max-abs to demodulate with (LoRaDemod.cpp:61-78).  Trusting: the previous packet's max
        // (batches are usually homogeneous: clean unit-amplitude IQ has the same max in every packet),
        // verified against the max tracked during the walk.  Distrusting (after a miss, or when a
        // packet above 1 must be refused): scan first with plain loads, then walk once.
        float use_max = pred_max;
        bool known = !legacy;
        if (legacy && (distrust > 0 || a.no_scratch)) {
            use_max = pipe_scan_max<DT>(pkt, count, lane, redf);
            known = true;
            if (distrust > 0) distrust = (use_max == pred_max) ? distrust - 1 : DISTRUST_PACKETS;
            pred_max = use_max;
        }

        for (int attempt = 0; attempt < 2 && !(flags & PS_SKIP); ++attempt) {
            float scale = 1.0f;
            flags &= ~PS_NORM;
            if (legacy && use_max > 1.0f) {
                if (a.no_scratch) { flags |= PS_SKIP; status = -34; have_metrics = false; break; }   // LoRaDemod.cpp:70-72
                flags |= PS_NORM;
                scale = __fdiv_rn(1.0f, use_max);
            }
            const bool norm = (flags & PS_NORM) != 0;
            const uint32_t n_est = total < 2 ? total : 2u;

            // ---- E units ---------------------------------------------------------------------------
            for (uint32_t s0 = 0; s0 < n_est; s0 += GD) {
                const uint32_t s = s0 + gq;
                const bool live = s < n_est;
                cf* win = slots + (size_t)stage * C::SLOT + (size_t)gq * C::PITCH;
                if (tid == 0) {
                    WinDesc d;
                    d.scale = scale; d.rate = 0.0f; d.start = 0.0f; d.flags = 0; d.out_off = 0; d.aux0 = s; d.aux1 = 0;
                    if (live) {
                        const cf* src = pkt + (size_t)s * N;
                        d.flags = WD_ACTIVE | (window_is_odd(src) ? WD_ODD : 0);
                        descs[stage * GD + gq] = d;
                        fetch_window<N>(win, src, &full[stage]);
                    } else {
                        descs[stage * GD + gq] = d;
                        ptx::mbar_arrive(&full[stage]);
                    }
                }
                ptx::mbar_wait(&full[stage], (phase_bits >> stage) & 1u);
                phase_bits ^= 1u << stage;
                const WinDesc d = descs[stage * GD + gq];
                cf v[P::R];
                const cf* raw = win + ((d.flags & WD_ODD) ? 1 : 0) + tid;
#pragma unroll
                for (int c = 0; c < P::R; ++c) {
                    cf x = fx::mk(0.0f, 0.0f);
                    if (live) {
                        x = raw[T * c];
                        if (norm) x = fx::cscale(x, scale);                                  // LoRaDemod.cpp:75-77
                        if (EXTRA && a.window) x = fx::cscale(x, __ldg(a.window + tid + T * c));
                    }
                    v[c] = x;
                }
                fft_inplace<SF, true>(v, tid, a.tw, a.ct, win);
                const Peak pk = group_peak<SF>(local_peak<SF>(v, tid), red);
                pipe_sync<DT>();                              // last exchange fully read: the slot becomes the spectrum
#pragma unroll
                for (int r = 0; r < P::R; ++r) win[P::bin_of_reg(tid, r)] = v[r];
                pipe_sync<DT>();
                if (tid == 0 && live) {
                    EstSym e; e.idx = 0; e.fi = 0.0f; e.phase = 0.0f; e.t = 0;
                    cf best_bin = fx::mk(0.0f, 0.0f);
                    // one sample phase (osr == 1): the candidate is kept iff its power beats -1e30 dB, i.e.
                    // iff 20*log10f(sqrtf(max)) is not -inf: max > 0 (LoRaDemod.cpp:96-108, LoRaDetector.hpp:60-64)
                    if (pk.val > 0.0f) {
                        const float fundamental = __fsqrt_rn(pk.val);
                        const cf l = win[pk.idx > 0 ? pk.idx - 1 : N - 1];
                        const cf r = win[pk.idx < N - 1 ? pk.idx + 1 : 0];
                        const float left = gm::hypotf_g(l.x, l.y), right = gm::hypotf_g(r.x, r.y);
                        const double demon = (2.0 * (double)fundamental) - (double)right - (double)left;
                        e.fi = (demon == 0.0) ? 0.0f : (float)(0.5 * (double)fx::fsub(right, left) / demon);
                        e.idx = pk.idx;
                        best_bin = win[pk.idx];
                    }
                    e.phase = gm::atan2f_g(best_bin.y, best_bin.x);
                    ws.est[s] = e;
                }
                ptx::fence_proxy_async();
                pipe_sync<DT>();
                stage ^= 1;
            }

            // ---- combine (lane 0), LoRaDemod.cpp:126-139 ---------------------------------------------
            float rate = 0.0f;
            int t_off = 0;
            if (lane == 0) {
                const float PI_F = 3.14159265358979323846f;
                float sum_index = 0.0f, phase_diff = 0.0f, prev_phase = 0.0f;
                for (uint32_t s = 0; s < n_est; ++s) {
                    const EstSym e = ws.est[s];
                    sum_index = fx::fadd(sum_index, fx::fadd((float)e.idx, e.fi));
                    if (s > 0) {
                        float dd = fx::fsub(e.phase, prev_phase);
                        while (dd > PI_F) dd = fx::fsub(dd, 2.0f * PI_F);
                        while (dd < -PI_F) dd = fx::fadd(dd, 2.0f * PI_F);
                        phase_diff = fx::fadd(phase_diff, dd);
                    }
                    prev_phase = e.phase;
                }
                const float avg_index = __fdiv_rn(sum_index, (float)n_est);
                const float cfo_coarse = __fdiv_rn(avg_index, (float)N);
                float cfo_fine = 0.0f;
                if (n_est > 1)
                    cfo_fine = __fdiv_rn(__fdiv_rn(phase_diff, (float)(n_est - 1)), fx::fmul(2.0f * PI_F, (float)N));
                cfo = fx::fadd(cfo_coarse, cfo_fine);
                const float frac = fx::fsub(avg_index, floorf(fx::fadd(avg_index, 0.5f)));
                const float avg_t = __fdiv_rn(0.0f, (float)n_est);
                toff = fx::fsub(avg_t, fx::fmul(frac, (float)N));
                t_off = (int)roundf(toff);
                rate = __fdiv_rn(fx::fmul(-2.0f * PI_F, cfo), (float)N);
                ws.sw[0] = 0; ws.sw[1] = 0;
            }
            if (DT <= 32) {
                rate = __shfl_sync(0xffffffffu, rate, 0);
                t_off = __shfl_sync(0xffffffffu, t_off, 0);
            } else {
                if (lane == 0) { ws.bc_rate = rate; ws.bc_toff = t_off; }
                pipe_sync<DT>();
                rate = ws.bc_rate; t_off = ws.bc_toff;
            }
            have_metrics = true;

            // ---- C units --------------------------------------------------------------------------
            // No descriptors in shared memory: every lane derives its window's base / start phase / path from
            // (rate, t_off, s) in registers; the window's leader only issues the bulk copy one unit ahead.
            const uint32_t n_units = (total + GD - 1) / GD;
            float m = 0.0f;
            bool suspicious = false;                       // a window whose spectrum had no positive bin (NaN / all-zero input)
            const bool odd_pkt = window_is_odd(pkt);       // N is even: every window of the packet shares the parity unless shifted
            auto prefetch = [&](uint32_t u, int stg) {
                if (tid != 0) return;
                const uint32_t s = u * GD + gq;
                if (s < total) {
                    const uint32_t base = shifted_base(s, (uint32_t)N, t_off, count);       // LoRaDemod.cpp:143-150
                    fetch_window<N>(slots + (size_t)stg * C::SLOT + (size_t)gq * C::PITCH, pkt + base, &full[stg]);
                } else {
                    ptx::mbar_arrive(&full[stg]);
                }
            };
            if (n_units) prefetch(0, stage);
            for (uint32_t u = 0; u < n_units; ++u) {
                if (u + 1 < n_units) prefetch(u + 1, stage ^ 1);
                const uint32_t s = u * GD + gq;
                const bool active = s < total;
                const uint32_t base = shifted_base(s, (uint32_t)N, t_off, count);
                const bool odd = odd_pkt != ((base & 1u) != 0);
                const float start = start_phase(rate, s, (uint32_t)N, t_off);           // LoRaDemod.cpp:152-153
                const bool slow = phases_need_slow_path(start, fx::fadd(start, fx::fmul(rate, (float)(N - 1))));
                ptx::mbar_wait(&full[stage], (phase_bits >> stage) & 1u);
                phase_bits ^= 1u << stage;
                cf* win = slots + (size_t)stage * C::SLOT + (size_t)gq * C::PITCH;
                cf v[P::R];
                const cf* raw = win + (odd ? 1 : 0) + tid;
                // an inactive window (past the packet's last symbol) runs the same code on whatever its slot
                // holds: the result is discarded, and no register zeroing / branch sits in the hot path
                float mw = 0.0f;
                // one copy of the chain in the loop (the running max costs one FMNMX3 per sample and is simply unused when
                // the max is already known): a second, untracked copy made the loop body 10 KB longer, and the SM
                // instruction cache -- 24 warps walk this loop out of phase -- is what the issue rate is short of
                chain_window<SF, EXTRA, true>(v, raw, tid, scale, rate, start, slow, a.down, a.window, &mw);
                if (active) m = fmaxf(m, mw);
#if !defined(LORA_ABLATE_FFT)                      // measurement-only variant (tools/ablate.sh): the chain without the FFT
                fft_inplace<SF, false>(v, tid, a.tw, a.ct, win);
#endif
                const Peak pk = group_peak<SF>(local_peak<SF>(v, tid), red);
                if (tid == 0 && active) {
                    if (!(pk.val > 0.0f)) suspicious = true;
                    if (have_sync && s < 2) ws.sw[s] = pk.idx;
                    else a.symbols[(size_t)pid * a.symbol_stride + (have_sync ? s - 2 : s)] = (uint16_t)pk.idx;
                }
                ptx::fence_proxy_async();
                pipe_sync<DT>();
                stage ^= 1;
            }

            // ---- verify the trusted max against the samples actually walked ------------------------------
            if (known) break;
            if (t_off != 0 || count != total * (uint32_t)N || pipe_any<DT>(suspicious)) {
                // shifted windows / a trailing partial symbol leave samples the units did not stream, and NaN samples
                // are invisible to the tracked max: scan the packet with the reference's own comparison
                m = pipe_scan_max<DT>(pkt, count, lane, redf);
            } else {
                m = pipe_max<DT>(m, redf);
            }
            const bool ok = (m > 1.0f) ? (m == use_max) : !(use_max > 1.0f);
            pred_max = m;
            if (ok) break;
            use_max = m;                                   // second walk with the packet's own max
            known = true;
            distrust = DISTRUST_PACKETS;
        }

        // ---- bookkeeping: LoRaDemod.cpp:166-196 ----------------------------------------------------------
        pipe_sync<DT>();
        if (lane == 0) {
            const bool run = !(flags & PS_SKIP);
            if (run) status = have_sync ? (int)(total - 2) : (int)total;
            if (have_metrics) {
                if (a.cfo) a.cfo[pid] = cfo;
                if (a.time_offset) a.time_offset[pid] = toff;
            }
            if (a.sync) {
                const unsigned shift = SF > 4 ? SF - 4 : 0;
                a.sync[pid] = (run && have_sync) ? (uint8_t)((((ws.sw[0] >> shift) & 0xf) << 4) | ((ws.sw[1] >> shift) & 0xf)) : 0;
            }
            if (a.status) a.status[pid] = status;
        }
        pipe_sync<DT>();
    }
}

// measured (A/B in one run, 1.965 GHz): SF10 44.4 -> 39.7 ms, SF11 42.0 -> 40.8 ms with a pipeline per packet; SF12's
// pipeline is the whole CTA and cfg 3's packets alternate between two max-abs classes, so it keeps the scan-first kernel
constexpr int PIPE_MAX_SF_DEFAULT = 11;

template <int SF>
static cudaError_t launch_warp_sf(const DemodArgs& a, int sm_count, cudaStream_t stream)
{
    using C = WarpCfg<SF>;
    const bool extra = a.down != nullptr || a.window != nullptr;
    static std::atomic<bool> attr_done[64];           // per device; the attribute set is idempotent, so a racing first launch is harmless
    int dev = 0;
    cudaGetDevice(&dev);
    if (!attr_done[dev & 63].load(std::memory_order_acquire)) {
        cudaFuncSetAttribute(demod_warp_kernel<SF, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::bytes);
        cudaFuncSetAttribute(demod_warp_kernel<SF, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::bytes);
        attr_done[dev & 63].store(true, std::memory_order_release);
    }
    const unsigned max_ctas = (unsigned)sm_count * STREAM_CTAS_PER_SM;    // persistent grid: every CTA resident
    const unsigned need = (a.n_packets + C::DOMS - 1) / C::DOMS;
    const unsigned grid = need < max_ctas ? need : max_ctas;
    if (extra) demod_warp_kernel<SF, true><<<grid, CTA_THREADS, C::bytes, stream>>>(a);
    else demod_warp_kernel<SF, false><<<grid, CTA_THREADS, C::bytes, stream>>>(a);
    return cudaGetLastError();
}

template <int SF>
static cudaError_t launch_packet_sf(const DemodArgs& a, int sm_count, cudaStream_t stream)
{
    using C = PacketCfg<SF>;
    const bool extra = a.down != nullptr || a.window != nullptr;
    static std::atomic<bool> attr_done[64];           // per device; the attribute set is idempotent, so a racing first launch is harmless
    int dev = 0;
    cudaGetDevice(&dev);
    if (!attr_done[dev & 63].load(std::memory_order_acquire)) {
        cudaFuncSetAttribute(demod_packet_kernel<SF, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::bytes);
        cudaFuncSetAttribute(demod_packet_kernel<SF, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::bytes);
        attr_done[dev & 63].store(true, std::memory_order_release);
    }
    const unsigned max_ctas = (unsigned)sm_count * STREAM_CTAS_PER_SM;    // persistent grid: every CTA resident
    const unsigned grid = a.n_packets < max_ctas ? a.n_packets : max_ctas;
    if (extra) demod_packet_kernel<SF, true><<<grid, CTA_THREADS, C::bytes, stream>>>(a);
    else demod_packet_kernel<SF, false><<<grid, CTA_THREADS, C::bytes, stream>>>(a);
    return cudaGetLastError();
}

bool packet_supported(int sf, const DemodArgs& a)
{
    return sf >= 7 && sf <= 12 && a.osr == 1 && (a.mode == MODE_LEGACY || a.mode == MODE_HIGHLEVEL) &&
           (unsigned long long)a.samples_per_packet < 0xf0000000ull;
}

cudaError_t launch_demod_packet(int sf, const DemodArgs& a, int sm_count, cudaStream_t stream)
{
    // pipeline-per-packet kernel (a warp for SF7-9, the window's 2 / 4 warps for SF10 / 11) up to PIPE_MAX_SF,
    // CTA-per-packet kernel above; LORA_PHY_PIPE_MAX_SF (9..12) overrides the split for A/B measurements
    static int pipe_max_sf = [] { const char* e = getenv("LORA_PHY_PIPE_MAX_SF"); const int v = e ? atoi(e) : 0;
                                  return (v >= 9 && v <= 12) ? v : PIPE_MAX_SF_DEFAULT; }();
    switch (sf) {
    case 7: return launch_warp_sf<7>(a, sm_count, stream);
    case 8: return launch_warp_sf<8>(a, sm_count, stream);
    case 9: return launch_warp_sf<9>(a, sm_count, stream);
    case 10: return pipe_max_sf >= 10 ? launch_warp_sf<10>(a, sm_count, stream) : launch_packet_sf<10>(a, sm_count, stream);
    case 11: return pipe_max_sf >= 11 ? launch_warp_sf<11>(a, sm_count, stream) : launch_packet_sf<11>(a, sm_count, stream);
    case 12: return pipe_max_sf >= 12 ? launch_warp_sf<12>(a, sm_count, stream) : launch_packet_sf<12>(a, sm_count, stream);
    default: return cudaErrorInvalidValue;
    }
}

} // namespace lk
