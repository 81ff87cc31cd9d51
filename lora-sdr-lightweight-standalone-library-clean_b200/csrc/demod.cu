// demod.cu -- demodulator kernels for sm_100a.
//
//   prepass_kernel  (K2a + sync symbols)  one group per packet: max-abs scan and scale
//                   (LoRaDemod.cpp:61-78), two-symbol CFO/STO estimator (LoRaDemod.cpp:80-136,
//                   phy.cpp:81-148), then the two sync symbols -> sync word, status.
//   symbols_kernel  (K1)  one group per data symbol: per-sample chain + exact-order FFT +
//                   first-max argmax (LoRaDemod.cpp:142-176, LoRaDetector.hpp:39-58).
//   detect_kernel   LoRaDetector::detect on caller-supplied windows (parity tests).
//
// Compiled with -fmad=false: every float operation rounds once, like the reference binary.
#include <atomic>
#include "demod_device.cuh"
#include "internal.h"

namespace lk {

template <int SF>
struct Smem {
    using P = fx::Plan<SF>;
    static constexpr int G = CTA_THREADS / P::T;          // groups (windows in flight) per CTA
    static constexpr int BUF = G * P::N;                  // cf entries per exchange buffer
    static constexpr size_t bytes = 2 * (size_t)BUF * sizeof(cf) + 8 * sizeof(Peak) + 8 * sizeof(float) +
                                    (size_t)G * sizeof(float);
};

template <int SF>
__device__ __forceinline__ void carve(unsigned char* raw, int g, cf*& buf0, cf*& buf1, Peak*& red,
                                      float*& redf, float*& slot)
{
    using S = Smem<SF>;
    cf* base = reinterpret_cast<cf*>(raw);
    buf0 = base + g * S::P::N;
    buf1 = base + S::BUF + g * S::P::N;
    red = reinterpret_cast<Peak*>(base + 2 * S::BUF);
    redf = reinterpret_cast<float*>(red + 8);
    slot = redf + 8 + g;
}

// ---------------------------------------------------------------------------------------------
// prepass: one group per packet
// ---------------------------------------------------------------------------------------------
template <int SF>
__global__ void __launch_bounds__(CTA_THREADS, 2) prepass_kernel(const DemodArgs a)
{
    using P = fx::Plan<SF>;
    constexpr int T = P::T, G = Smem<SF>::G, N = P::N;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int g = threadIdx.x / T, tid = threadIdx.x % T;
    cf *buf0, *buf1; Peak* red; float *redf, *slot;
    carve<SF>(smem_raw, g, buf0, buf1, red, redf, slot);
    int flip = 0;
    const uint32_t step = (uint32_t)N * a.osr;

    for (uint32_t pid0 = blockIdx.x * G; pid0 < a.n_packets; pid0 += gridDim.x * G) {
        const uint32_t pid = pid0 + g;
        const bool active = pid < a.n_packets;
        const cf* pkt = a.iq + (size_t)(active ? pid : 0) * a.packet_stride;
        const uint32_t count = active ? (a.sample_counts ? min(a.sample_counts[pid], a.samples_per_packet) : a.samples_per_packet) : 0;
        const uint32_t total = count / step;

        // -- 1. max-abs scan (legacy path only) : LoRaDemod.cpp:61-68 --
        PacketState st; st.scale = 1.0f; st.rate = 0.0f; st.t_off = 0; st.flags = 0;
        int status = 0;
        if (a.mode == MODE_LEGACY && !a.unit_scale) {
            float m = 0.0f;
            const bool vec = ((reinterpret_cast<uintptr_t>(pkt) & 15) == 0);
            if (vec) {
                const float4* p4 = reinterpret_cast<const float4*>(pkt);
                const uint32_t pairs = count >> 1;
                for (uint32_t j = tid; j < pairs; j += T) {
                    const float4 q = __ldg(p4 + j);
                    float r0 = fabsf(q.x), i0 = fabsf(q.y), r1 = fabsf(q.z), i1 = fabsf(q.w);
                    m = fmaxf(m, (r0 < i0) ? i0 : r0);
                    m = fmaxf(m, (r1 < i1) ? i1 : r1);
                }
                if ((count & 1) && tid == 0) {
                    const cf x = ldg_cf(pkt + count - 1);
                    float r = fabsf(x.x), im = fabsf(x.y);
                    m = fmaxf(m, (r < im) ? im : r);
                }
            } else {
                for (uint32_t j = tid; j < count; j += T) {
                    const cf x = ldg_cf(pkt + j);
                    float r = fabsf(x.x), im = fabsf(x.y);
                    m = fmaxf(m, (r < im) ? im : r);
                }
            }
            m = group_max<T>(m, redf);
            if (m > 1.0f) {
                if (a.no_scratch) { st.flags |= PS_SKIP; status = -34; /* -ERANGE, LoRaDemod.cpp:70-72 */ }
                else { st.flags |= PS_NORM; st.scale = __fdiv_rn(1.0f, m); }
            }
        } else if (a.mode == MODE_HIGHLEVEL) {
            // phy.cpp:190-194 (argument checks; per packet when lengths are ragged)
            if (count % step != 0) { st.flags |= PS_SKIP; status = -22; }
            else if (total < 2) { st.flags |= PS_SKIP; status = -34; }
        }
        const bool norm = (st.flags & PS_NORM) != 0;
        const bool run = active && !(st.flags & PS_SKIP);

        // -- 2. estimator : LoRaDemod.cpp:80-136 / phy.cpp:93-147 --
        uint32_t n_est = total;
        if (a.mode != MODE_EST_ONLY && n_est > 2) n_est = 2;
        if (!run) n_est = 0;
        float sum_index = 0.0f, phase_diff = 0.0f, prev_phase = 0.0f;
        bool have_prev = false;
        unsigned sum_t = 0;
        for (uint32_t s = 0; s < a.est_uniform; ++s) {
            const bool live = s < n_est;
            float best_p = -1e30f, best_fi = 0.0f;
            int best_idx = 0;
            unsigned best_t = 0;
            cf best_bin = fx::mk(0.0f, 0.0f);
            for (unsigned t = 0; t < a.osr; ++t) {
                cf v[P::R];
                load_window<SF>(v, pkt + (size_t)s * step + t, tid, a.osr, live, norm, st.scale, nullptr,
                                false, 0.0f, 0.0f, a.window);
                fft_group<SF, true>(v, tid, a.tw, a.ct, buf0, buf1, flip);
                const Peak pk = group_peak<SF>(local_peak<SF>(v, tid), red);
                // spectrum to shared memory: thread 0 needs X[idx] and its two neighbours
                cf* spec = flip ? buf1 : buf0;
                flip ^= 1;
#pragma unroll
                for (int r = 0; r < P::R; ++r) spec[P::bin_of_reg(tid, r)] = v[r];
                group_sync<T>();
                if (tid == 0 && live) {
                    const float fundamental = __fsqrt_rn(pk.val);
                    const float p = fx::fsub(fx::fmul(20.0f, gm::log10f_g(fundamental)), a.power_scale);
                    const cf l = spec[pk.idx > 0 ? pk.idx - 1 : N - 1];
                    const cf r = spec[pk.idx < N - 1 ? pk.idx + 1 : 0];
                    const float left = gm::hypotf_g(l.x, l.y), right = gm::hypotf_g(r.x, r.y);
                    const double demon = (2.0 * (double)fundamental) - (double)right - (double)left;
                    const float fi = (demon == 0.0) ? 0.0f : (float)(0.5 * (double)fx::fsub(right, left) / demon);
                    const bool tie = (a.mode == MODE_LEGACY) && p == best_p && pk.idx < best_idx; // LoRaDemod.cpp:102 vs phy.cpp:116
                    if (p > best_p || tie) {
                        best_p = p; best_idx = pk.idx; best_fi = fi; best_t = t; best_bin = spec[pk.idx];
                    }
                }
            }
            if (tid == 0 && live) {
                sum_t += best_t;
                sum_index = fx::fadd(sum_index, fx::fadd((float)best_idx, best_fi));
                const float phase = gm::atan2f_g(best_bin.y, best_bin.x);
                if (have_prev) {
                    const float PI_F = 3.14159265358979323846f;
                    float d = fx::fsub(phase, prev_phase);
                    while (d > PI_F) d = fx::fsub(d, 2.0f * PI_F);
                    while (d < -PI_F) d = fx::fadd(d, 2.0f * PI_F);
                    phase_diff = fx::fadd(phase_diff, d);
                }
                prev_phase = phase;
                have_prev = true;
            }
        }
        float cfo = 0.0f, toff = 0.0f;
        if (tid == 0 && run && (a.mode != MODE_EST_ONLY || n_est > 0)) {
            const float PI_F = 3.14159265358979323846f;
            const float avg_index = __fdiv_rn(sum_index, (float)n_est);
            const float cfo_coarse = __fdiv_rn(avg_index, (float)N);
            float cfo_fine = 0.0f;
            if (n_est > 1)
                cfo_fine = __fdiv_rn(__fdiv_rn(phase_diff, (float)(n_est - 1)), fx::fmul(2.0f * PI_F, (float)N));
            cfo = fx::fadd(cfo_coarse, cfo_fine);
            const float frac = fx::fsub(avg_index, floorf(fx::fadd(avg_index, 0.5f)));
            const float avg_t = __fdiv_rn((float)sum_t, (float)n_est);
            toff = fx::fsub(avg_t, fx::fmul(fx::fmul(frac, (float)N), (float)a.osr));
            if (a.cfo) a.cfo[pid] = cfo;
            if (a.time_offset) a.time_offset[pid] = toff;
            st.t_off = (int)roundf(toff);
            st.rate = __fdiv_rn(fx::fmul(-2.0f * PI_F, cfo), (float)N);
        }
        if (a.mode == MODE_EST_ONLY) continue;

        st.rate = group_bcast<T>(st.rate, tid, slot);
        st.t_off = __float_as_int(group_bcast<T>(__int_as_float(st.t_off), tid, slot));

        // -- 3. sync symbols and bookkeeping : LoRaDemod.cpp:166-196 --
        const bool have_sync = total >= 2;
        int sw[2] = { 0, 0 };
        if (!a.skip_sync) {
#pragma unroll
            for (uint32_t s = 0; s < 2; ++s)
                sw[s] = demod_symbol<SF, false>(a, pkt, count, s, run && have_sync, st, tid, buf0, buf1, flip, red);
        }
        if (tid == 0 && active) {
            if (run) {
                if (a.state) a.state[pid] = st;
                if (a.sync) {
                    const unsigned shift = SF > 4 ? SF - 4 : 0;
                    a.sync[pid] = have_sync ? (uint8_t)((((sw[0] >> shift) & 0xf) << 4) | ((sw[1] >> shift) & 0xf)) : 0;
                }
                status = have_sync ? (int)(total - 2) : (int)total;
            } else {
                if (a.state) a.state[pid] = st;
                if (a.sync) a.sync[pid] = 0;
            }
            if (a.status) a.status[pid] = status;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// symbols: one group per (packet, data symbol)
// ---------------------------------------------------------------------------------------------
template <int SF>
__global__ void __launch_bounds__(CTA_THREADS, 2) symbols_kernel(const DemodArgs a)
{
    using P = fx::Plan<SF>;
    constexpr int T = P::T, G = Smem<SF>::G, N = P::N;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int g = threadIdx.x / T, tid = threadIdx.x % T;
    cf *buf0, *buf1; Peak* red; float *redf, *slot;
    carve<SF>(smem_raw, g, buf0, buf1, red, redf, slot);
    int flip = 0;
    const uint32_t step = (uint32_t)N * a.osr;
    const unsigned long long n_items = (unsigned long long)a.n_packets * a.data_symbols_max;

    for (unsigned long long item0 = (unsigned long long)blockIdx.x * G; item0 < n_items;
         item0 += (unsigned long long)gridDim.x * G) {
        const unsigned long long item = item0 + g;
        bool active = item < n_items;
        const uint32_t pid = active ? (uint32_t)(item / a.data_symbols_max) : 0;
        const uint32_t k = active ? (uint32_t)(item % a.data_symbols_max) : 0;   // output slot
        const uint32_t count = a.sample_counts ? min(a.sample_counts[pid], a.samples_per_packet) : a.samples_per_packet;
        const uint32_t total = count / step;
        const bool have_sync = total >= 2;
        const uint32_t s = have_sync ? k + 2 : k;
        PacketState st = a.state[pid];
        active = active && s < total && !(st.flags & PS_SKIP);
        const cf* pkt = a.iq + (size_t)pid * a.packet_stride;
        const int idx = demod_symbol<SF, false>(a, pkt, count, s, active, st, tid, buf0, buf1, flip, red);
        if (tid == 0 && active) a.symbols[(size_t)pid * a.symbol_stride + k] = (uint16_t)idx;
    }
}

// ---------------------------------------------------------------------------------------------
// detect: LoRaDetector.hpp:39-74 on raw windows
// ---------------------------------------------------------------------------------------------
template <int SF>
__global__ void __launch_bounds__(CTA_THREADS, 2) detect_kernel(const DetectArgs a)
{
    using P = fx::Plan<SF>;
    constexpr int T = P::T, G = Smem<SF>::G, N = P::N;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int g = threadIdx.x / T, tid = threadIdx.x % T;
    cf *buf0, *buf1; Peak* red; float *redf, *slot;
    carve<SF>(smem_raw, g, buf0, buf1, red, redf, slot);
    int flip = 0;
    for (uint32_t w0 = blockIdx.x * G; w0 < a.n_windows; w0 += gridDim.x * G) {
        const uint32_t wi = w0 + g;
        const bool active = wi < a.n_windows;
        cf v[P::R];
        load_window<SF>(v, a.windows + (size_t)(active ? wi : 0) * N, tid, 1, active, false, 1.0f, nullptr,
                        false, 0.0f, 0.0f, nullptr);
        fft_group<SF, true>(v, tid, a.tw, a.ct, buf0, buf1, flip);
        const Peak pk = group_peak<SF>(local_peak<SF>(v, tid), red);
        cf* spec = flip ? buf1 : buf0;
        flip ^= 1;
#pragma unroll
        for (int r = 0; r < P::R; ++r) {
            const int bin = P::bin_of_reg(tid, r);
            spec[bin] = v[r];
            if (active && a.spectrum) a.spectrum[(size_t)wi * N + bin] = v[r];
        }
        group_sync<T>();
        if (tid == 0 && active) {
            double total = 0.0;                                  // LoRaDetector.hpp:45-52, ascending
            for (int i = 0; i < N; ++i)
                total += (double)fx::fadd(fx::fmul(spec[i].x, spec[i].x), fx::fmul(spec[i].y, spec[i].y));
            const float noise = __fsqrt_rn((float)(total - (double)pk.val));
            const float fundamental = __fsqrt_rn(pk.val);
            const cf l = spec[pk.idx > 0 ? pk.idx - 1 : N - 1];
            const cf r = spec[pk.idx < N - 1 ? pk.idx + 1 : 0];
            const float left = gm::hypotf_g(l.x, l.y), right = gm::hypotf_g(r.x, r.y);
            const double demon = (2.0 * (double)fundamental) - (double)right - (double)left;
            if (a.index) a.index[wi] = (uint32_t)pk.idx;
            if (a.power) a.power[wi] = fx::fsub(fx::fmul(20.0f, gm::log10f_g(fundamental)), a.power_scale);
            if (a.power_avg) a.power_avg[wi] = fx::fsub(fx::fmul(20.0f, gm::log10f_g(noise)), a.power_scale);
            if (a.findex) a.findex[wi] = (demon == 0.0) ? 0.0f : (float)(0.5 * (double)fx::fsub(right, left) / demon);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// launchers
// ---------------------------------------------------------------------------------------------
template <int SF>
static cudaError_t launch_demod_sf(const DemodArgs& a, int sm_count, cudaStream_t stream, bool prepass,
                                   bool symbols, unsigned* launches)
{
    using S = Smem<SF>;
    static std::atomic<bool> attr_done[64];           // per device; the attribute set is idempotent, so a racing first launch is harmless
    int dev = 0;
    cudaGetDevice(&dev);
    if (!attr_done[dev & 63].load(std::memory_order_acquire)) {
        cudaFuncSetAttribute(prepass_kernel<SF>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)S::bytes);
        cudaFuncSetAttribute(symbols_kernel<SF>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)S::bytes);
        attr_done[dev & 63].store(true, std::memory_order_release);
    }
    const unsigned max_ctas = (unsigned)sm_count * 64u;
    if (prepass && a.n_packets) {
        unsigned need = (a.n_packets + S::G - 1) / S::G;
        unsigned grid = need < max_ctas ? need : max_ctas;
        prepass_kernel<SF><<<grid, CTA_THREADS, S::bytes, stream>>>(a);
        ++*launches;
    }
    const unsigned long long items = (unsigned long long)a.n_packets * a.data_symbols_max;
    if (symbols && items) {
        unsigned long long need = (items + S::G - 1) / S::G;
        unsigned grid = need < max_ctas ? (unsigned)need : max_ctas;
        symbols_kernel<SF><<<grid, CTA_THREADS, S::bytes, stream>>>(a);
        ++*launches;
    }
    return cudaGetLastError();
}

cudaError_t launch_demod(int sf, const DemodArgs& a, int sm_count, cudaStream_t stream, bool prepass,
                         bool symbols, unsigned* launches)
{
    switch (sf) {
    case 2: return launch_demod_sf<2>(a, sm_count, stream, prepass, symbols, launches);
    case 3: return launch_demod_sf<3>(a, sm_count, stream, prepass, symbols, launches);
    case 4: return launch_demod_sf<4>(a, sm_count, stream, prepass, symbols, launches);
    case 5: return launch_demod_sf<5>(a, sm_count, stream, prepass, symbols, launches);
    case 6: return launch_demod_sf<6>(a, sm_count, stream, prepass, symbols, launches);
    case 7: return launch_demod_sf<7>(a, sm_count, stream, prepass, symbols, launches);
    case 8: return launch_demod_sf<8>(a, sm_count, stream, prepass, symbols, launches);
    case 9: return launch_demod_sf<9>(a, sm_count, stream, prepass, symbols, launches);
    case 10: return launch_demod_sf<10>(a, sm_count, stream, prepass, symbols, launches);
    case 11: return launch_demod_sf<11>(a, sm_count, stream, prepass, symbols, launches);
    case 12: return launch_demod_sf<12>(a, sm_count, stream, prepass, symbols, launches);
    default: return cudaErrorInvalidValue;
    }
}

template <int SF>
static cudaError_t launch_detect_sf(const DetectArgs& a, int sm_count, cudaStream_t stream)
{
    using S = Smem<SF>;
    cudaFuncSetAttribute(detect_kernel<SF>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)S::bytes);
    unsigned need = (a.n_windows + S::G - 1) / S::G;
    unsigned max_ctas = (unsigned)sm_count * 16u;
    detect_kernel<SF><<<need < max_ctas ? need : max_ctas, CTA_THREADS, S::bytes, stream>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_detect(int sf, const DetectArgs& a, int sm_count, cudaStream_t stream)
{
    switch (sf) {
    case 2: return launch_detect_sf<2>(a, sm_count, stream);
    case 3: return launch_detect_sf<3>(a, sm_count, stream);
    case 4: return launch_detect_sf<4>(a, sm_count, stream);
    case 5: return launch_detect_sf<5>(a, sm_count, stream);
    case 6: return launch_detect_sf<6>(a, sm_count, stream);
    case 7: return launch_detect_sf<7>(a, sm_count, stream);
    case 8: return launch_detect_sf<8>(a, sm_count, stream);
    case 9: return launch_detect_sf<9>(a, sm_count, stream);
    case 10: return launch_detect_sf<10>(a, sm_count, stream);
    case 11: return launch_detect_sf<11>(a, sm_count, stream);
    case 12: return launch_detect_sf<12>(a, sm_count, stream);
    default: return cudaErrorInvalidValue;
    }
}

} // namespace lk
