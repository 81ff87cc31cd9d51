// demod_fused.cu -- the whole demodulator in ONE persistent launch (SF >= 7, osr == 1):
// lora_demodulate (LoRaDemod.cpp:50-197) and demodulate (phy.cpp:182-243) for a batch.
//
// The reference walks a packet three times: max-abs scan of every sample (:61-68), the
// two-symbol CFO/STO estimator (:80-136), then the per-symbol loop (:142-176) whose inputs
// (1/max scale, rate, t_off) come from the first two.  Here the three become three kinds of
// work units, numbered in one global order and dealt round-robin to the pipelines of a persistent grid:
//
//   SCAN  unit  GD windows of one packet group: max(|re|,|im|) -> atomicMax + arrival count
//   EST   unit  one packet group: wait for its scans, scale, estimator FFTs on symbols 0/1,
//               cfo / time_offset / rate / t_off, the two sync symbols, status; publish
//               PacketState and raise the group's ready flag
//   DATA  unit  GD data-symbol windows: rotate + FFT + first-max argmax -> uint16 symbol
//
// Ticket block b holds [SCAN units of group b][EST unit of group b-LE][DATA units of group b-LD]
// with 0 < LE < LD, so every unit only depends on units with SMALLER numbers.  Each pipeline walks its
// units in increasing order and only blocks on a dependency while it holds no older unfinished unit, so
// the globally oldest unfinished unit can always run; the grid is launched cooperatively (all CTAs
// resident), which makes that argument complete: no deadlock.
// The look-ahead (LD groups) is sized so that a packet's windows are still in the 126 MB L2
// when its EST and DATA units read them again: HBM sees every IQ byte once.
//
// Windows travel HBM/L2 -> shared memory by 1-D bulk asynchronous copies (cp.async.bulk +
// mbarrier complete_tx), issued one unit ahead by each window's leader thread.
#include "stream_common.cuh"
#include "internal.h"

namespace lk {

struct EstRes { int idx; float fi; float phase; int pad; };

struct EShared {                       // per pipeline scratch of the EST unit (up to 2 packets, 4 windows)
    uint32_t pid[2], count[2], ok[2], n_est[2];
    float scale[2];
    int flags[2], status[2];
    EstRes est[4];
    int sw[4];
    PacketState pst[2];
    uint32_t bcast;                    // ticket broadcast slot (CTA-wide pipelines)
    uint32_t pad[3];
};

template <int SF>
struct FusedCfg : StreamCfg<SF> {
    using B = StreamCfg<SF>;
    static constexpr int EP = B::GD >= 2 ? B::GD / 2 : 1;          // packets per group (= per EST unit)
    static constexpr size_t es_bytes = (size_t)B::DOMS * sizeof(EShared);
    static constexpr size_t bytes = B::bytes + es_bytes + 8 * sizeof(float);
};

enum { UNIT_NONE = 0, UNIT_SCAN = 1, UNIT_EST = 2, UNIT_DATA = 3 };

struct Unit { int type; uint32_t group; uint32_t idx; };

__device__ __forceinline__ uint32_t ld_acquire_u32(const uint32_t* p)
{
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_u32(uint32_t* p, uint32_t v)
{
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// compute unit number -> unit.  Block b = [EST unit of group b][ub DATA units of group b - data_lag];
// groups outside [0, n_groups) are no-ops.
__device__ __forceinline__ Unit decode_compute(const FusedArgs& f, uint32_t t)
{
    Unit u; u.type = UNIT_NONE; u.group = 0; u.idx = 0;
    if (t >= f.total_units) return u;
    const uint32_t blk = (uint32_t)__umul64hi((unsigned long long)t, f.block_magic);   // t / block_units
    const uint32_t w = t - blk * f.block_units;
    long long g;
    int type;
    if (w == 0) { type = UNIT_EST; g = (long long)blk; u.idx = 0; }
    else { type = UNIT_DATA; g = (long long)blk - f.data_lag; u.idx = w - 1; }
    if (g >= 0 && g < (long long)f.n_groups) { u.type = type; u.group = (uint32_t)g; }
    return u;
}
// scan unit number -> unit (group = t / ua)
__device__ __forceinline__ Unit decode_scan(const FusedArgs& f, uint32_t t)
{
    Unit u; u.type = UNIT_NONE; u.group = 0; u.idx = 0;
    if (t >= f.scan_units) return u;
    u.group = (uint32_t)__umul64hi((unsigned long long)t, f.ua_magic);
    u.idx = t - u.group * f.ua;
    u.type = UNIT_SCAN;
    return u;
}

template <int SF, bool EXTRA>
__global__ void __launch_bounds__(CTA_THREADS, STREAM_CTAS_PER_SM) demod_fused_kernel(const FusedArgs f)
{
    using C = FusedCfg<SF>;
    using P = fx::Plan<SF>;
    constexpr int T = C::T, N = C::N, GD = C::GD, EP = C::EP;
    constexpr unsigned FULLMASK = 0xffffffffu;
    const DemodArgs& a = f.d;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int dom = threadIdx.x / C::DOM_THREADS;
    const int dt = threadIdx.x % C::DOM_THREADS;
    const int gq = dt / T, tid = dt % T;

    cf* slots = reinterpret_cast<cf*>(smem_raw) + (size_t)dom * C::STAGES * C::SLOT;
    WinDesc* descs = reinterpret_cast<WinDesc*>(smem_raw + C::slot_bytes) + (size_t)dom * C::STAGES * GD;
    uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw + C::slot_bytes + C::desc_bytes) + dom * C::STAGES;
    Peak* red = reinterpret_cast<Peak*>(smem_raw + C::slot_bytes + C::desc_bytes + C::bar_bytes);
    EShared* es = reinterpret_cast<EShared*>(smem_raw + C::B::bytes) + dom;
    float* redf = reinterpret_cast<float*>(smem_raw + C::B::bytes + C::es_bytes);

    if (dt == 0) {
#pragma unroll
        for (int s = 0; s < C::STAGES; ++s) ptx::mbar_init(&full[s], GD);
        ptx::fence_mbar_init();
    }
    __syncthreads();

    unsigned phase_bits = 0;                     // bit s = parity the next wait on stage s expects
    auto wait_stage = [&](int stage) {
        ptx::mbar_wait(&full[stage], (phase_bits >> stage) & 1u);
        phase_bits ^= 1u << stage;
    };
    auto slot_of = [&](int stage) -> cf* { return slots + (size_t)stage * C::SLOT + (size_t)gq * C::PITCH; };

    // ---- DATA window: descriptor + fetch (leader only).  blocking == false: if the packet's state is
    //      not published yet the window is marked WD_DEFERRED and fetched later by fetch_deferred(). ----
    auto data_prefetch = [&](const Unit& u, int stage, bool blocking) {
        WinDesc d;
        d.scale = 1.0f; d.rate = 0.0f; d.start = 0.0f; d.flags = 0; d.out_off = 0; d.aux0 = 0; d.aux1 = 0;
        const uint32_t item = u.idx * GD + gq;
        const uint32_t e = (uint32_t)__umul64hi((unsigned long long)item, a.dsm_magic);      // item / dsm
        const uint32_t k = item - e * a.data_symbols_max;
        const uint32_t pid = u.group * EP + e;
        const cf* src = nullptr;
        if (e < (uint32_t)EP && pid < a.n_packets) {
            const uint32_t count = a.sample_counts ? min(a.sample_counts[pid], a.samples_per_packet) : a.samples_per_packet;
            const uint32_t total = count / (uint32_t)N;
            const uint32_t s = total >= 2 ? k + 2 : k;
            if (s < total) {
                bool ready = ld_acquire_u32(f.eready + pid) != 0;
                if (!ready && blocking) {
                    while (ld_acquire_u32(f.eready + pid) == 0) __nanosleep(64);
                    ready = true;
                }
                if (!ready) {
                    d.flags = WD_DEFERRED;
                } else {
                    const int4 sraw = __ldcg(reinterpret_cast<const int4*>(a.state + pid));
                    const float scale = __int_as_float(sraw.x), rate = __int_as_float(sraw.y);
                    const int t_off = sraw.z, sflags = sraw.w;
                    if (!(sflags & PS_SKIP)) {
                        src = a.iq + (size_t)pid * a.packet_stride + shifted_base(s, (uint32_t)N, t_off, count);
                        d.scale = scale;
                        d.rate = rate;
                        d.start = start_phase(rate, s, (uint32_t)N, t_off);
                        d.flags = WD_ACTIVE;
                        if (phases_need_slow_path(d.start, fx::fadd(d.start, fx::fmul(rate, (float)(N - 1))))) d.flags |= WD_SLOW;
                        if (window_is_odd(src)) d.flags |= WD_ODD;
                        d.out_off = (unsigned long long)pid * a.symbol_stride + k;
                    }
                }
            }
        }
        descs[stage * GD + gq] = d;
        if (src) fetch_window<N>(slot_of(stage), src, &full[stage]);
        else ptx::mbar_arrive(&full[stage]);
    };

    // ---- SCAN window: descriptor + fetch (leader only) ----------------------------------------------
    auto scan_prefetch = [&](const Unit& u, int stage) {
        WinDesc d;
        d.scale = 1.0f; d.rate = 0.0f; d.start = 0.0f; d.flags = 0; d.out_off = 0; d.aux0 = 0; d.aux1 = 0;
        const uint32_t item = u.idx * GD + gq;
        const uint32_t e = (uint32_t)__umul64hi((unsigned long long)item, f.awin_magic);     // item / awin
        const uint32_t j = item - e * f.awin;
        const uint32_t pid = u.group * EP + e;
        const cf* src = nullptr;
        if (e < (uint32_t)EP && pid < a.n_packets) {
            const uint32_t count = a.sample_counts ? min(a.sample_counts[pid], a.samples_per_packet) : a.samples_per_packet;
            const unsigned long long off = (unsigned long long)j * N;
            if (off < count) {
                const uint32_t len = (uint32_t)min((unsigned long long)N, count - off);
                d.flags = WD_ACTIVE;
                d.out_off = pid;
                d.aux0 = len;
                d.aux1 = (uint32_t)off;
                const cf* p = a.iq + (size_t)pid * a.packet_stride + off;
                if (len == (uint32_t)N) { src = p; if (window_is_odd(src)) d.flags |= WD_ODD; }
                else d.flags |= WD_DIRECT;                              // trailing partial window: plain loads
            }
        }
        descs[stage * GD + gq] = d;
        if (src) fetch_window<N>(slot_of(stage), src, &full[stage]);
        else ptx::mbar_arrive(&full[stage]);
    };

    auto prefetch = [&](const Unit& u, int stage) {
        if (tid != 0) return;
        if (u.type == UNIT_DATA) data_prefetch(u, stage, false);
        else if (u.type == UNIT_SCAN) scan_prefetch(u, stage);
        else {                                   // EST fetches its windows itself; NONE has nothing to fetch
            WinDesc d;
            d.scale = 1.0f; d.rate = 0.0f; d.start = 0.0f; d.flags = 0; d.out_off = 0; d.aux0 = 0; d.aux1 = 0;
            descs[stage * GD + gq] = d;
            ptx::mbar_arrive(&full[stage]);
        }
    };

    // ---- SCAN unit: LoRaDemod.cpp:61-68 ---------------------------------------------------------------
    auto run_scan = [&](const Unit& u, int stage) {
        // throttle: stay at most scan_lead groups ahead of the estimator, so that what is scanned now is still
        // in L2 when the compute pipelines come for it (waits on a SMALLER group: cannot deadlock)
        if (dt == 0 && u.group >= f.scan_lead) {
            const uint32_t pid = (u.group - f.scan_lead) * EP;
            while (ld_acquire_u32(f.eready + pid) == 0) __nanosleep(256);
        }
        domain_sync<SF>();
        const WinDesc d = descs[stage * GD + gq];
        float m = 0.0f;
        if (d.flags & WD_ACTIVE) {
            if (d.flags & WD_DIRECT) {
                const cf* p = a.iq + (size_t)d.out_off * a.packet_stride + d.aux1;
                for (uint32_t i = tid; i < d.aux0; i += T) {
                    const cf x = ldg_cf(p + i);
                    const float r = fabsf(x.x), im = fabsf(x.y);
                    m = fmaxf(m, (r < im) ? im : r);
                }
            } else {
                const cf* raw = slot_of(stage) + ((d.flags & WD_ODD) ? 1 : 0) + tid;
#pragma unroll
                for (int c = 0; c < P::R; ++c) {
                    const cf x = raw[T * c];
                    const float r = fabsf(x.x), im = fabsf(x.y);
                    m = fmaxf(m, (r < im) ? im : r);
                }
            }
        }
        m = group_max<T>(m, redf);
        if (tid == 0 && (d.flags & WD_ACTIVE)) {
            atomicMax(f.pmax + d.out_off, __float_as_uint(m));      // m >= 0: unsigned order == float order
            __threadfence();
            atomicAdd(f.acount + d.out_off, 1u);
        }
    };

    // ---- DATA unit: LoRaDemod.cpp:142-176 / phy.cpp:204-238 + LoRaDetector.hpp:39-58 -------------------
    auto run_data = [&](const Unit& u, int stage) {
        WinDesc d = descs[stage * GD + gq];
        const bool deferred = (d.flags & WD_DEFERRED) != 0;
        const bool any_deferred = (T <= 32) ? (__any_sync(FULLMASK, deferred) != 0) : (__syncthreads_or(deferred) != 0);
        if (any_deferred) {                      // this pipeline holds no older ticket now: blocking is safe
            if (tid == 0) {
                if (deferred) data_prefetch(u, stage, true);
                else ptx::mbar_arrive(&full[stage]);
            }
            wait_stage(stage);
            d = descs[stage * GD + gq];
        }
        cf* win = slot_of(stage);
        const bool active = (d.flags & WD_ACTIVE) != 0;
        cf v[P::R];
        if (active) {
            chain_window<SF, EXTRA>(v, win + ((d.flags & WD_ODD) ? 1 : 0) + tid, tid, d.scale, d.rate, d.start,
                                    (d.flags & WD_SLOW) != 0, a.down, a.window);
        } else {
#pragma unroll
            for (int c = 0; c < P::R; ++c) v[c] = fx::mk(0.0f, 0.0f);
        }
        fft_inplace<SF, false>(v, tid, a.tw, a.ct, win);
        const Peak pk = group_peak<SF>(local_peak<SF>(v, tid), red);
        if (tid == 0 && active) a.symbols[d.out_off] = (uint16_t)pk.idx;
    };

    // ---- EST unit: LoRaDemod.cpp:69-136 + :166-196 (legacy) / phy.cpp:190-197 + :81-148 + :232-241 -----
    // Windows of the unit: w = 0 .. 2*EP-1, packet e = w / 2, symbol = w % 2.  With GD >= 2 all of them run
    // side by side (group gq owns window gq); with GD == 1 the single group walks them one after another.
    auto run_est = [&](const Unit& u, int stage) {
        cf* win = slot_of(stage);
        const float PI_F = 3.14159265358979323846f;

        // (1) packet headers: wait for the scans, derive scale / early status
        if (tid == 0 && (GD == 1 || (gq & 1) == 0)) {
            const int e = GD == 1 ? 0 : gq >> 1;
            const uint32_t pid = u.group * EP + e;
            const bool ok = pid < a.n_packets;
            uint32_t count = 0;
            float scale = 1.0f;
            int flags = 0, status = 0;
            if (ok) {
                count = a.sample_counts ? min(a.sample_counts[pid], a.samples_per_packet) : a.samples_per_packet;
                if (a.mode == MODE_LEGACY) {
                    const uint32_t expected = (uint32_t)(((unsigned long long)count + N - 1) / N);
                    while (ld_acquire_u32(f.acount + pid) < expected) __nanosleep(64);
                    const float m = __uint_as_float(__ldcg(f.pmax + pid));
                    if (m > 1.0f) {
                        if (a.no_scratch) { flags |= PS_SKIP; status = -34; }            // -ERANGE, LoRaDemod.cpp:70-72
                        else { flags |= PS_NORM; scale = __fdiv_rn(1.0f, m); }           // LoRaDemod.cpp:74
                    }
                } else {                                                                  // phy.cpp:190-194
                    if (count % (uint32_t)N != 0) { flags |= PS_SKIP; status = -22; }
                    else if (count / (uint32_t)N < 2) { flags |= PS_SKIP; status = -34; }
                }
            }
            const uint32_t total = count / (uint32_t)N;
            es->pid[e] = pid; es->count[e] = count; es->ok[e] = ok ? 1u : 0u;
            es->scale[e] = scale; es->flags[e] = flags; es->status[e] = status;
            es->n_est[e] = (ok && !(flags & PS_SKIP)) ? (total < 2 ? total : 2u) : 0u;
        }
        domain_sync<SF>();

        // (2) estimator windows: raw symbols 0 / 1, scaled, [Hann], no rotation, full-precision FFT
        for (int w0 = 0; w0 < 2 * EP; w0 += GD) {
            const int w = w0 + gq, e = w >> 1, sym = w & 1;
            const bool live = (uint32_t)sym < es->n_est[e];
            const cf* src = a.iq + (size_t)es->pid[e] * a.packet_stride + (size_t)sym * N;
            const bool odd = live && window_is_odd(src);
            if (tid == 0) {
                if (live) fetch_window<N>(win, src, &full[stage]);
                else ptx::mbar_arrive(&full[stage]);
            }
            wait_stage(stage);
            cf v[P::R];
            const bool norm = (es->flags[e] & PS_NORM) != 0;
            const float scale = es->scale[e];
#pragma unroll
            for (int c = 0; c < P::R; ++c) {
                cf x = fx::mk(0.0f, 0.0f);
                if (live) {
                    x = win[(odd ? 1 : 0) + tid + T * c];
                    if (norm) x = fx::cscale(x, scale);                                   // LoRaDemod.cpp:75-77
                    if (a.window) x = fx::cscale(x, __ldg(a.window + tid + T * c));       // LoRaDemod.cpp:96-97
                }
                v[c] = x;
            }
            fft_inplace<SF, true>(v, tid, a.tw, a.ct, win);
            const Peak pk = group_peak<SF>(local_peak<SF>(v, tid), red);
            domain_sync<SF>();                   // last exchange fully read: the slot becomes the spectrum
#pragma unroll
            for (int r = 0; r < P::R; ++r) win[P::bin_of_reg(tid, r)] = v[r];
            domain_sync<SF>();
            if (tid == 0) {
                EstRes res; res.idx = 0; res.fi = 0.0f; res.phase = 0.0f; res.pad = 0;
                if (live) {                       // LoRaDetector.hpp:60-71, LoRaDemod.cpp:101-115
                    const float fundamental = __fsqrt_rn(pk.val);
                    const float p = fx::fsub(fx::fmul(20.0f, gm::log10f_g(fundamental)), a.power_scale);
                    cf best_bin = fx::mk(0.0f, 0.0f);
                    if (p > -1e30f) {             // osr == 1: a single candidate per symbol
                        const cf l = win[pk.idx > 0 ? pk.idx - 1 : N - 1];
                        const cf r = win[pk.idx < N - 1 ? pk.idx + 1 : 0];
                        const float left = gm::hypotf_g(l.x, l.y), right = gm::hypotf_g(r.x, r.y);
                        const double demon = (2.0 * (double)fundamental) - (double)right - (double)left;
                        res.fi = (demon == 0.0) ? 0.0f : (float)(0.5 * (double)fx::fsub(right, left) / demon);
                        res.idx = pk.idx;
                        best_bin = win[pk.idx];
                    }
                    res.phase = gm::atan2f_g(best_bin.y, best_bin.x);
                }
                es->est[w] = res;
            }
            ptx::fence_proxy_async();            // spectrum stores precede the next bulk copy into the slot
            domain_sync<SF>();
        }

        // (3) offsets: LoRaDemod.cpp:126-139 / phy.cpp:136-147
        if (tid == 0 && (GD == 1 || (gq & 1) == 0)) {
            const int e = GD == 1 ? 0 : gq >> 1;
            PacketState st; st.scale = es->scale[e]; st.rate = 0.0f; st.t_off = 0; st.flags = es->flags[e];
            const uint32_t n_est = es->n_est[e];
            if (es->ok[e] && !(st.flags & PS_SKIP)) {     // n_est == 0 (empty packet) yields 0/0 = NaN like the reference
                float sum_index = 0.0f, phase_diff = 0.0f;
                for (uint32_t s = 0; s < n_est; ++s) {
                    const EstRes r = es->est[2 * e + s];
                    sum_index = fx::fadd(sum_index, fx::fadd((float)r.idx, r.fi));
                    if (s > 0) {
                        float d = fx::fsub(r.phase, es->est[2 * e + s - 1].phase);
                        while (d > PI_F) d = fx::fsub(d, 2.0f * PI_F);
                        while (d < -PI_F) d = fx::fadd(d, 2.0f * PI_F);
                        phase_diff = fx::fadd(phase_diff, d);
                    }
                }
                const float avg_index = __fdiv_rn(sum_index, (float)n_est);
                const float cfo_coarse = __fdiv_rn(avg_index, (float)N);
                float cfo_fine = 0.0f;
                if (n_est > 1)
                    cfo_fine = __fdiv_rn(__fdiv_rn(phase_diff, (float)(n_est - 1)), fx::fmul(2.0f * PI_F, (float)N));
                const float cfo = fx::fadd(cfo_coarse, cfo_fine);
                const float frac = fx::fsub(avg_index, floorf(fx::fadd(avg_index, 0.5f)));
                const float avg_t = __fdiv_rn(0.0f, (float)n_est);                        // osr == 1: best_t is always 0
                const float toff = fx::fsub(avg_t, fx::fmul(fx::fmul(frac, (float)N), 1.0f));
                const uint32_t pid = es->pid[e];
                if (a.cfo) a.cfo[pid] = cfo;
                if (a.time_offset) a.time_offset[pid] = toff;
                st.t_off = (int)roundf(toff);
                st.rate = __fdiv_rn(fx::fmul(-2.0f * PI_F, cfo), (float)N);
            }
            es->pst[e] = st;
        }
        domain_sync<SF>();

        // (4) the two sync symbols through the data-symbol chain: LoRaDemod.cpp:142-175
        for (int w0 = 0; w0 < 2 * EP; w0 += GD) {
            const int w = w0 + gq, e = w >> 1, sym = w & 1;
            const PacketState st = es->pst[e];
            const uint32_t count = es->count[e];
            const bool active = es->ok[e] && !(st.flags & PS_SKIP) && count / (uint32_t)N >= 2;
            const cf* src = a.iq + (size_t)es->pid[e] * a.packet_stride + shifted_base((uint32_t)sym, (uint32_t)N, st.t_off, count);
            const bool odd = active && window_is_odd(src);
            if (tid == 0) {
                if (active) fetch_window<N>(win, src, &full[stage]);
                else ptx::mbar_arrive(&full[stage]);
            }
            wait_stage(stage);
            cf v[P::R];
            if (active) {
                const float start = start_phase(st.rate, (uint32_t)sym, (uint32_t)N, st.t_off);
                const bool slow = phases_need_slow_path(start, fx::fadd(start, fx::fmul(st.rate, (float)(N - 1))));
                chain_window<SF, EXTRA>(v, win + (odd ? 1 : 0) + tid, tid, st.scale, st.rate, start, slow, a.down, a.window);
            } else {
#pragma unroll
                for (int c = 0; c < P::R; ++c) v[c] = fx::mk(0.0f, 0.0f);
            }
            fft_inplace<SF, false>(v, tid, a.tw, a.ct, win);
            const Peak pk = group_peak<SF>(local_peak<SF>(v, tid), red);
            if (tid == 0) es->sw[w] = active ? pk.idx : 0;
            ptx::fence_proxy_async();
            domain_sync<SF>();
        }

        // (5) publish: sync word, status, PacketState, then the ready flag (LoRaDemod.cpp:178-196)
        if (tid == 0 && (GD == 1 || (gq & 1) == 0)) {
            const int e = GD == 1 ? 0 : gq >> 1;
            if (es->ok[e]) {
                const uint32_t pid = es->pid[e];
                const PacketState st = es->pst[e];
                const uint32_t total = es->count[e] / (uint32_t)N;
                const bool run = !(st.flags & PS_SKIP), have_sync = total >= 2;
                a.state[pid] = st;
                if (a.sync) {
                    const unsigned shift = SF > 4 ? SF - 4 : 0;
                    a.sync[pid] = (run && have_sync) ? (uint8_t)((((es->sw[2 * e] >> shift) & 0xf) << 4) | ((es->sw[2 * e + 1] >> shift) & 0xf)) : 0;
                }
                if (a.status) a.status[pid] = run ? (have_sync ? (int)(total - 2) : (int)total) : es->status[e];
                __threadfence();
                st_release_u32(f.eready + pid, 1u);
            }
        }
    };

    // ---- main loop.  The last scan_ctas CTAs are scan pipelines (memory-latency bound, nearly free in issue
    //      slots), the others compute pipelines.  Units of each kind are dealt round-robin to the pipelines of
    //      that kind; every pipeline walks its units in increasing order, prefetching one ahead. ----
    const bool scanner = blockIdx.x >= gridDim.x - f.scan_ctas;
    const uint32_t n_pipes = (scanner ? f.scan_ctas : gridDim.x - f.scan_ctas) * C::DOMS;
    const uint32_t limit = scanner ? f.scan_units : f.total_units;
    uint32_t t_cur = (scanner ? blockIdx.x - (gridDim.x - f.scan_ctas) : blockIdx.x) * C::DOMS + dom;
    Unit cur = scanner ? decode_scan(f, t_cur) : decode_compute(f, t_cur);
    int stage = 0;
    if (t_cur < limit) prefetch(cur, 0);

    while (t_cur < limit) {
        const uint32_t t_nxt = t_cur + n_pipes;          // limit + n_pipes < 2^32 (fused_plan)
        const Unit nxt = scanner ? decode_scan(f, t_nxt) : decode_compute(f, t_nxt);
        if (t_nxt < limit) prefetch(nxt, stage ^ 1);

        wait_stage(stage);
        if (cur.type == UNIT_DATA) run_data(cur, stage);
        else if (cur.type == UNIT_SCAN) run_scan(cur, stage);
        else if (cur.type == UNIT_EST) run_est(cur, stage);

        ptx::fence_proxy_async();     // our generic-proxy stores to the slot precede the next bulk copy into it
        domain_sync<SF>();
        t_cur = t_nxt;
        cur = nxt;
        stage ^= 1;
    }
}

template <int SF>
static cudaError_t launch_fused_sf(const FusedArgs& f, int sm_count, cudaStream_t stream)
{
    using C = FusedCfg<SF>;
    const bool extra = f.d.down != nullptr || f.d.window != nullptr;
    static bool attr_done[64] = { false };
    int dev = 0;
    cudaGetDevice(&dev);
    if (!attr_done[dev & 63]) {
        cudaFuncSetAttribute(demod_fused_kernel<SF, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::bytes);
        cudaFuncSetAttribute(demod_fused_kernel<SF, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::bytes);
        attr_done[dev & 63] = true;
    }
    // every CTA must be resident (units wait on units of other pipelines): size the grid from the
    // occupancy query and launch cooperatively, which fails instead of deadlocking if it cannot fit
    int per_sm = 0;
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, extra ? (const void*)demod_fused_kernel<SF, true>
                                                                                 : (const void*)demod_fused_kernel<SF, false>,
                                                                  CTA_THREADS, C::bytes);
    if (e != cudaSuccess) return e;
    if (per_sm < 1) return cudaErrorLaunchOutOfResources;
    const unsigned max_ctas = (unsigned)sm_count * (unsigned)per_sm;
    const unsigned long long need_c = ((unsigned long long)f.total_units + C::DOMS - 1) / C::DOMS;
    FusedArgs g = f;
    unsigned grid;
    if (f.awin == 0) {                      // high-level mode: nothing to scan
        g.scan_ctas = 0;
        grid = need_c < max_ctas ? (unsigned)(need_c ? need_c : 1) : max_ctas;
    } else {                                // one CTA in three scans (one per SM at 3 resident CTAs per SM)
        unsigned compute = max_ctas - max_ctas / 3;
        if (need_c < compute) compute = (unsigned)(need_c ? need_c : 1);
        unsigned scan = (compute + 1) / 2;
        const unsigned long long need_s = ((unsigned long long)f.scan_units + C::DOMS - 1) / C::DOMS;
        if (need_s < scan) scan = (unsigned)(need_s ? need_s : 1);
        if (compute + scan > max_ctas) scan = max_ctas - compute;
        if (scan < 1) return cudaErrorLaunchOutOfResources;
        g.scan_ctas = scan;
        grid = compute + scan;
    }
    void* args[] = { &g };
    return cudaLaunchCooperativeKernel(extra ? (const void*)demod_fused_kernel<SF, true> : (const void*)demod_fused_kernel<SF, false>,
                                       dim3(grid), dim3(CTA_THREADS), args, C::bytes, stream);
}

// Geometry of the ticket schedule for (sf, batch); false if this batch must take the two-kernel path.
bool fused_plan(int sf, const DemodArgs& a, uint32_t lookahead_bytes, FusedArgs* f)
{
    if (sf < 7 || sf > 12 || a.osr != 1 || a.mode == MODE_EST_ONLY) return false;
    if ((unsigned long long)a.samples_per_packet >= 0xf0000000ull) return false;
    const uint32_t N = 1u << sf;
    const uint32_t T = N / 16, dom_threads = T <= 32 ? 32 : CTA_THREADS, GD = dom_threads / T;
    const uint32_t EP = GD >= 2 ? GD / 2 : 1;
    f->d = a;
    f->n_groups = (a.n_packets + EP - 1) / EP;
    f->awin = a.mode == MODE_LEGACY ? (a.samples_per_packet + N - 1) / N : 0;
    f->ua = (EP * f->awin + GD - 1) / GD;
    f->ub = (EP * a.data_symbols_max + GD - 1) / GD;
    f->block_units = 1 + f->ub;
    // EST -> DATA distance: an EST unit is 2..4 window steps with exposed fetches; its DATA units must be dealt
    // well after it finished.  Scan lead: how far the scanners may run ahead of the estimator (L2 residency).
    const unsigned long long group_bytes = (unsigned long long)EP * a.samples_per_packet * 8ull + 1ull;
    unsigned long long ld = lookahead_bytes / group_bytes;
    if (ld < 8) ld = 8;
    if (ld > 8192) ld = 8192;
    f->data_lag = (uint32_t)ld;
    f->scan_lead = (uint32_t)(ld / 4 > 4 ? ld / 4 : 4);
    f->scan_ctas = 0;
    const unsigned long long total = ((unsigned long long)f->n_groups + f->data_lag) * f->block_units;
    const unsigned long long scans = (unsigned long long)f->n_groups * f->ua;
    if (total >= 0xf0000000ull || scans >= 0xf0000000ull) return false;
    f->total_units = (uint32_t)total;
    f->scan_units = (uint32_t)scans;
    f->block_magic = (~0ull / f->block_units) + 1ull;
    f->awin_magic = f->awin ? (~0ull / f->awin) + 1ull : 0ull;
    f->ua_magic = f->ua ? (~0ull / f->ua) + 1ull : 0ull;
    return true;
}

cudaError_t launch_demod_fused(int sf, const FusedArgs& f, int sm_count, cudaStream_t stream)
{
    switch (sf) {
    case 7: return launch_fused_sf<7>(f, sm_count, stream);
    case 8: return launch_fused_sf<8>(f, sm_count, stream);
    case 9: return launch_fused_sf<9>(f, sm_count, stream);
    case 10: return launch_fused_sf<10>(f, sm_count, stream);
    case 11: return launch_fused_sf<11>(f, sm_count, stream);
    case 12: return launch_fused_sf<12>(f, sm_count, stream);
    default: return cudaErrorInvalidValue;
    }
}

} // namespace lk
