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
#include "demod_device.cuh"
#include "internal.h"
#include "stream_ptx.cuh"

namespace lk {

struct __align__(16) WinDesc {       // written by the window's leader at prefetch time
    float scale, rate, start;
    uint32_t flags;                   // WD_*
    unsigned long long out_off;       // index into a.symbols
    unsigned long long pad;
};
enum { WD_ACTIVE = 1, WD_ODD = 2, WD_SLOW = 4 };

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
template <int SF>
__device__ __forceinline__ void fft_inplace(cf (&v)[fx::Plan<SF>::R], int tid, const cf* __restrict__ tw,
                                            const cf* __restrict__ ct, cf* ex)
{
    using P = fx::Plan<SF>;
    P::template pass0<false>(v, tw);
#pragma unroll
    for (int p = 1; p < P::NP; ++p) {
        domain_sync<SF>();            // raw samples (p == 1) / previous exchange (p > 1) fully consumed
#pragma unroll
        for (int r = 0; r < P::R; ++r) {
            const int o = (p == 1) ? P::pos_after_pass0(tid, r) : P::pos_in_pass(p - 1, tid, r);
            ex[P::phys(o)] = v[r];
        }
        domain_sync<SF>();
#pragma unroll
        for (int r = 0; r < P::R; ++r) v[r] = ex[P::phys(P::pos_in_pass(p, tid, r))];
        P::pass(p, v, tid, ct);
    }
}

// rare path: some |phase| of the window is below 2^-12 or reaches 120 rad (glibc's other branches)
__device__ __noinline__ float2 sincos_slow(float ph)
{
    float sn, cs;
    gm::sincosf_g(ph, &sn, &cs);
    return make_float2(cs, sn);
}

template <int SF, bool EXTRA>
__global__ void __launch_bounds__(CTA_THREADS, 2) symbols_stream_kernel(const DemodArgs a)
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
        d.scale = 1.0f; d.rate = 0.0f; d.start = 0.0f; d.flags = 0; d.out_off = 0; d.pad = 0;
        const uint32_t item = u * GD + gq;
        const cf* src = nullptr;
        if (item < n_items) {
            const uint32_t pid = (uint32_t)__umul64hi((unsigned long long)item, a.dsm_magic);   // item / dsm
            const uint32_t k = item - pid * a.data_symbols_max;
            const uint32_t count = a.sample_counts ? a.sample_counts[pid] : a.samples_per_packet;
            const uint32_t total = count / (uint32_t)N;
            const uint32_t s = total >= 2 ? k + 2 : k;
            const int4 sraw = __ldg(reinterpret_cast<const int4*>(a.state + pid));
            PacketState st;
            st.scale = __int_as_float(sraw.x); st.rate = __int_as_float(sraw.y); st.t_off = sraw.z; st.flags = sraw.w;
            if (s < total && !(st.flags & PS_SKIP)) {
                uint32_t base = s * (uint32_t)N;                   // LoRaDemod.cpp:143-150
                if (st.t_off > 0) {
                    if ((unsigned long long)base + (unsigned)st.t_off + N <= count) base += (uint32_t)st.t_off;
                } else if (st.t_off < 0) {
                    const uint32_t off = (uint32_t)(-(long long)st.t_off);
                    if (off <= base) base -= off;
                }
                src = a.iq + (size_t)pid * a.packet_stride + base;
                d.scale = st.scale;
                d.rate = st.rate;
                d.start = fx::fmul(st.rate, fx::fadd(__uint2float_rn(s * (uint32_t)N), __int2float_rn(st.t_off)));
                // phases run monotonically from `start` (i = 0) to `last` (i = N-1): the fast sincos
                // core needs 2^-12 <= |ph| < 120 for every sample of the window
                const float last = fx::fadd(d.start, fx::fmul(st.rate, (float)(N - 1)));
                const float lo = fminf(fabsf(d.start), fabsf(last)), hi = fmaxf(fabsf(d.start), fabsf(last));
                const bool same_sign = (__float_as_uint(d.start) ^ __float_as_uint(last)) < 0x80000000u;
                d.flags = WD_ACTIVE;
                if (!(same_sign && lo >= 0x1p-11f && hi < 119.0f)) d.flags |= WD_SLOW;
                d.out_off = (unsigned long long)pid * a.symbol_stride + k;
            }
        }
        cf* win = slots + (size_t)stage * C::SLOT + (size_t)gq * C::PITCH;
        if (src) {
            const uintptr_t addr = reinterpret_cast<uintptr_t>(src);
            const unsigned odd = (unsigned)(addr >> 3) & 1u;      // window starts 8 bytes after a 16-byte boundary
            if (odd) {
                d.flags |= WD_ODD;
                win[N] = ldg_cf(src + (N - 1));                   // the one sample the aligned copy misses
            }
            descs[stage * GD + gq] = d;
            ptx::mbar_arrive_expect_tx(&full[stage], (unsigned)(N * sizeof(cf)));
            ptx::bulk_g2s(win, reinterpret_cast<const void*>(addr - 8u * odd), (unsigned)(N * sizeof(cf)), &full[stage]);
        } else {
            descs[stage * GD + gq] = d;
            ptx::mbar_arrive(&full[stage]);
        }
    };

    uint32_t u = blockIdx.x * C::DOMS + dom;
    int stage = 0;
    unsigned phase_bits = 0;                     // bit s = parity the next wait on stage s expects
    const float fi0 = (float)tid;
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
            const cf* raw = win + ((d.flags & WD_ODD) ? 1 : 0) + tid;
            if (!(d.flags & WD_SLOW)) {
#pragma unroll
                for (int c = 0; c < P::R; ++c) {
                    cf x = fx::cscale2(raw[T * c], d.scale);                                 // LoRaDemod.cpp:75-77
                    if (EXTRA && a.down) x = fx::cmul(x, ldg_cf(a.down + tid + T * c));      // phy.cpp:224
                    const float ph = fx::fadd(d.start, fx::fmul(d.rate, fx::fadd(fi0, (float)(T * c))));
                    float sn, cs;
                    gm::sincosf_core(ph, &sn, &cs);                                          // LoRaDemod.cpp:155-159
                    x = fx::cmul(x, fx::mk(cs, sn));
                    if (EXTRA && a.window) x = fx::cscale(x, __ldg(a.window + tid + T * c)); // LoRaDemod.cpp:160-161
                    v[c] = x;
                }
            } else {
#pragma unroll
                for (int c = 0; c < P::R; ++c) {
                    cf x = fx::cscale(raw[T * c], d.scale);
                    if (EXTRA && a.down) x = fx::cmul(x, ldg_cf(a.down + tid + T * c));
                    const float ph = fx::fadd(d.start, fx::fmul(d.rate, fx::fadd(fi0, (float)(T * c))));
                    const float2 r = sincos_slow(ph);
                    x = fx::cmul(x, fx::mk(r.x, r.y));
                    if (EXTRA && a.window) x = fx::cscale(x, __ldg(a.window + tid + T * c));
                    v[c] = x;
                }
            }
        } else {
#pragma unroll
            for (int c = 0; c < P::R; ++c) v[c] = fx::mk(0.0f, 0.0f);
        }

        fft_inplace<SF>(v, tid, a.tw, a.ct, win);
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
    static bool attr_done[64] = { false };
    int dev = 0;
    cudaGetDevice(&dev);
    if (!attr_done[dev & 63]) {
        cudaFuncSetAttribute(symbols_stream_kernel<SF, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::bytes);
        cudaFuncSetAttribute(symbols_stream_kernel<SF, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::bytes);
        attr_done[dev & 63] = true;
    }
    const unsigned long long items = (unsigned long long)a.n_packets * a.data_symbols_max;
    const unsigned long long units = (items + C::GD - 1) / C::GD;
    unsigned long long need = (units + C::DOMS - 1) / C::DOMS;
    const unsigned max_ctas = (unsigned)sm_count * 2u;           // persistent: 2 resident CTAs per SM
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
