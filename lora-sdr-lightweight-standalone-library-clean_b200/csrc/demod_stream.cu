// demod_stream.cu -- K1, streaming form (SF >= 7, osr == 1): the data-symbol kernel.
//
// Reference loop restated: LoRaDemod.cpp:142-176 / phy.cpp:204-238 (per-symbol chain) and
// LoRaDetector.hpp:39-58 (FFT + first-max argmax), for every data symbol of every packet.
//
// Shape of the kernel (B200):
//   * persistent grid, 2 CTAs of 256 threads per SM; a CTA is split into independent
//     "pipelines": one warp each for SF <= 9 (a window needs T = N/16 <= 32 threads), the whole
//     CTA for SF >= 10.  A pipeline owns two shared-memory stages and two mbarriers.
//   * every symbol window (N complex64, contiguous in HBM) is fetched by ONE bulk asynchronous
//     copy (cp.async.bulk global->shared, mbarrier complete_tx) issued by the window's leader
//     thread one step ahead of use, so HBM latency is hidden behind the previous step's math
//     and no thread ever waits on a global load for samples;
//   * threads read their 16 samples from shared memory, apply normalise -> [downchirp] ->
//     CFO rotation (glibc-exact sincosf in FP64 FMA) -> [Hann], run the exact-order FFT of
//     fft_exact.cuh with the exchanges done IN PLACE in the window's own stage slot, reduce
//     |X|^2 to the first-max bin and store one uint16;
//   * IQ is read from HBM exactly once by this kernel (the packet prepass reads it once more;
//     see DESIGN.md for the fused schedule).
#include <atomic>
#include "stream_common.cuh"
#include "internal.h"

namespace lk {

template <int SF, bool EXTRA>
__global__ void __launch_bounds__(CTA_THREADS, STREAM_CTAS_PER_SM) symbols_stream_kernel(const DemodArgs a)
{
    using C = StreamCfg<SF>;
    using P = fx::Plan<SF>;
    constexpr int T = C::T, N = C::N, GD = C::GD;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int dom = threadIdx.x / C::DOM_THREADS;
    const int dt = threadIdx.x % C::DOM_THREADS;
    const int gq = dt / T, tid = dt % T;

    cf* slots = reinterpret_cast<cf*>(smem_raw) + (size_t)dom * C::STAGES * C::SLOT;
    WinDesc* descs = reinterpret_cast<WinDesc*>(smem_raw + C::slot_bytes) + (size_t)dom * C::STAGES * GD;
    uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw + C::slot_bytes + C::desc_bytes) + dom * C::STAGES;
    Peak* red = reinterpret_cast<Peak*>(smem_raw + C::slot_bytes + C::desc_bytes + C::bar_bytes);

    if (dt == 0) {
#pragma unroll
        for (int s = 0; s < C::STAGES; ++s) ptx::mbar_init(&full[s], GD);
        ptx::fence_mbar_init();
    }
    __syncthreads();

    // the launcher guarantees n_items < 2^32 (else the generic kernel runs)
    const uint32_t n_items = a.n_packets * a.data_symbols_max;
    const uint32_t n_units = (n_items + GD - 1) / GD;
    const uint32_t stride = gridDim.x * C::DOMS;

    // leader of window gq: describe item, publish the descriptor, start the bulk copy
    auto prefetch = [&](uint32_t u, int stage) {
        if (tid != 0) return;
        WinDesc d;
        d.scale = 1.0f; d.rate = 0.0f; d.start = 0.0f; d.flags = 0; d.out_off = 0; d.aux0 = 0; d.aux1 = 0;
        const uint32_t item = u * GD + gq;
        const cf* src = nullptr;
        if (item < n_items) {
            const uint32_t pid = (uint32_t)__umul64hi((unsigned long long)item, a.dsm_magic);   // item / dsm
            const uint32_t k = item - pid * a.data_symbols_max;
            const uint32_t count = a.sample_counts ? min(a.sample_counts[pid], a.samples_per_packet) : a.samples_per_packet;
            const uint32_t total = count / (uint32_t)N;
            const uint32_t s = total >= 2 ? k + 2 : k;
            const int4 sraw = __ldg(reinterpret_cast<const int4*>(a.state + pid));
            PacketState st;
            st.scale = __int_as_float(sraw.x); st.rate = __int_as_float(sraw.y); st.t_off = sraw.z; st.flags = sraw.w;
            if (s < total && !(st.flags & PS_SKIP)) {
                const uint32_t base = shifted_base(s, (uint32_t)N, st.t_off, count);   // LoRaDemod.cpp:143-150
                src = a.iq + (size_t)pid * a.packet_stride + base;
                d.scale = st.scale;
                d.rate = st.rate;
                d.start = start_phase(st.rate, s, (uint32_t)N, st.t_off);
                d.flags = WD_ACTIVE;
                if (phases_need_slow_path(d.start, fx::fadd(d.start, fx::fmul(st.rate, (float)(N - 1))))) d.flags |= WD_SLOW;
                d.out_off = (unsigned long long)pid * a.symbol_stride + k;
            }
        }
        cf* win = slots + (size_t)stage * C::SLOT + (size_t)gq * C::PITCH;
        if (src) {
            if (window_is_odd(src)) d.flags |= WD_ODD;
            descs[stage * GD + gq] = d;
            fetch_window<N>(win, src, &full[stage]);
        } else {
            descs[stage * GD + gq] = d;
            ptx::mbar_arrive(&full[stage]);
        }
    };

    uint32_t u = blockIdx.x * C::DOMS + dom;
    int stage = 0;
    unsigned phase_bits = 0;                     // bit s = parity the next wait on stage s expects
    if (u < n_units) prefetch(u, 0);

    while (u < n_units) {
        const uint32_t u_next = u + stride;        // n_units + stride < 2^32 is checked by the launcher
        if (u_next < n_units) prefetch(u_next, stage ^ 1);

        ptx::mbar_wait(&full[stage], (phase_bits >> stage) & 1u);
        phase_bits ^= 1u << stage;

        const WinDesc d = descs[stage * GD + gq];
        cf* win = slots + (size_t)stage * C::SLOT + (size_t)gq * C::PITCH;
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

        ptx::fence_proxy_async();     // our generic-proxy stores to the slot precede the next bulk copy into it
        domain_sync<SF>();
        u = u_next;
        stage ^= 1;
    }
}

template <int SF>
static cudaError_t launch_stream_sf(const DemodArgs& a, int sm_count, cudaStream_t stream)
{
    using C = StreamCfg<SF>;
    const bool extra = a.down != nullptr || a.window != nullptr;
    static std::atomic<bool> attr_done[64];           // per device; the attribute set is idempotent, so a racing first launch is harmless
    int dev = 0;
    cudaGetDevice(&dev);
    if (!attr_done[dev & 63].load(std::memory_order_acquire)) {
        cudaFuncSetAttribute(symbols_stream_kernel<SF, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::bytes);
        cudaFuncSetAttribute(symbols_stream_kernel<SF, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::bytes);
        attr_done[dev & 63].store(true, std::memory_order_release);
    }
    const unsigned long long items = (unsigned long long)a.n_packets * a.data_symbols_max;
    const unsigned long long units = (items + C::GD - 1) / C::GD;
    unsigned long long need = (units + C::DOMS - 1) / C::DOMS;
    const unsigned max_ctas = (unsigned)sm_count * STREAM_CTAS_PER_SM;   // persistent grid: every CTA resident
    const unsigned grid = need < max_ctas ? (unsigned)need : max_ctas;
    if (extra) symbols_stream_kernel<SF, true><<<grid, CTA_THREADS, C::bytes, stream>>>(a);
    else symbols_stream_kernel<SF, false><<<grid, CTA_THREADS, C::bytes, stream>>>(a);
    return cudaGetLastError();
}

bool stream_supported(int sf, const DemodArgs& a)
{
    const unsigned long long items = (unsigned long long)a.n_packets * a.data_symbols_max;
    return sf >= 7 && sf <= 12 && a.osr == 1 && items < 0xf0000000ull &&
           (unsigned long long)a.samples_per_packet < 0xf0000000ull;
}

cudaError_t launch_symbols_stream(int sf, const DemodArgs& a, int sm_count, cudaStream_t stream)
{
    switch (sf) {
    case 7: return launch_stream_sf<7>(a, sm_count, stream);
    case 8: return launch_stream_sf<8>(a, sm_count, stream);
    case 9: return launch_stream_sf<9>(a, sm_count, stream);
    case 10: return launch_stream_sf<10>(a, sm_count, stream);
    case 11: return launch_stream_sf<11>(a, sm_count, stream);
    case 12: return launch_stream_sf<12>(a, sm_count, stream);
    default: return cudaErrorInvalidValue;
    }
}

} // namespace lk
