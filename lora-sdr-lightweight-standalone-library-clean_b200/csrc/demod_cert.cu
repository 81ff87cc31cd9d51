// demod_cert.cu -- the default demodulator (SF 7..12, osr == 1): lora_demodulate (LoRaDemod.cpp:50-197) and
// demodulate (phy.cpp:182-243) for a batch, ONE launch, one pipeline per packet, HBM read once -- with the
// per-window work done by a THROUGHPUT transform whose argmax is PROVEN to be the reference's, window by window.
//
// Why: replaying the reference's float operations (glibc-exact sincosf per sample in FP64, un-fused kissfft
// butterflies; round 1's kernels) costs 92 thread-instructions per sample and leaves the kernel issue-bound at a
// third of the HBM rate.  north_star asks for bit-exact SYMBOLS (and 1e-4 on |X|^2), not bit-exact spectra, so:
//
//   E units (exact)   symbols 0,1 raw through the exact-order FFT: cfo, time_offset, rate, t_off bit for bit
//                     (LoRaDemod.cpp:80-139) -- round 1's code.
//   C units (fast)    every symbol: samples * f_c, where f_c = scale * [down_i] * exp(j*rate*i) * [w_i] is computed ONCE
//                     PER PACKET per thread (|X|^2 does not depend on the per-symbol start phase, LoRaDemod.cpp:152-159),
//                     then ff::Plan (fft_fast.cuh: FMA radix-16/8/4, 1-2 exchanges), |Y|^2, and the top bin.
//   certification     the reference's computed spectrum X^ and ours Y both approximate the same ideal DFT Z; with
//                     S = sqrt(N * sum|u_i|^2) >= sum|u_i| every bin obeys | |X^_k| - |Y_k| | <= K*S, K = u*(C0 +
//                     |start| + 2|rate|N), u = 2^-24 (DESIGN.md derives C0 from kissfft's stage count, its float twiddle
//                     table, the libm bounds and our own rounding).  The window is accepted iff the top |Y|^2 beats the
//                     second by more than that margin (tier 1: the top bin holds > 9/16 of N*E, by Parseval nothing else
//                     can come near; tier 2: the true second maximum).  Then X^ has the same strict, unique maximum.
//   fallback          windows that fail -- ties, noise-only windows with close bins, NaN / Inf / denormal input -- are
//                     replayed through round 1's exact chain (stream_common.cuh) at the end of the packet, same launch.
//                     Symbols, sync word, status, cfo and time_offset are therefore bit-exact BY CONSTRUCTION.
//
// The 1/max normalisation (LoRaDemod.cpp:61-78) no longer forces a second walk: the certified decision is invariant to
// the (positive, finite) scale, so a packet is walked once with the previous packet's max folded into f_c while the
// true max is tracked; on a mismatch only the two E units are redone with the true scale, and the walk stands if the
// new rate / t_off agree with the ones it used to within the tolerance its margin already budgeted.
#include <atomic>
#include <cstdlib>
#include "stream_common.cuh"
#include "fft_fast.cuh"
#include "fft_soa.cuh"
#include "fft_prune.cuh"
#include "internal.h"

namespace lk {

#ifndef CERT_CTAS_PER_SM
#define CERT_CTAS_PER_SM 2
#endif

#ifndef CERT_CTAS_SMALL
#define CERT_CTAS_SMALL CERT_CTAS_PER_SM      // resident CTAs per SM for SF7-9 (with CERT_CTA_SMALL: the occupancy A/B knob)
#endif
constexpr int cert_ctas(int sf) { return sf <= 9 ? CERT_CTAS_SMALL : CERT_CTAS_PER_SM; }

#ifndef CERT_CTA_SMALL
#define CERT_CTA_SMALL 256                    // threads per CTA for SF7-9 (a window is at most one warp there: any multiple of 32)
#endif

#ifndef CERT_TW0_SMEM_MAX_SF
#define CERT_TW0_SMEM_MAX_SF 0                // pass-0 twiddles in shared memory up to this SF (A/B knob; registers measured faster)
#endif

#ifndef CERT_SOA_SF7
#define CERT_SOA_SF7 1                        // SF7: struct-of-arrays packed transform (fft_soa.cuh); 0 = the AoS code of fft_fast.cuh (A/B)
#endif

#ifndef CERT_SOA_TW_SMEM
#define CERT_SOA_TW_SMEM 0                    // SF7 SoA path: pass-0 twiddle pairs from shared memory instead of 32 registers (A/B knob)
#endif

#ifndef CERT_EARLY_REFILL
#define CERT_EARLY_REFILL 0                   // warp pipelines: refill at the top of the step (needs an end-of-step __syncwarp), A/B knob
#endif

#ifndef CERT_PRUNE
#define CERT_PRUNE 1                          // pruned transform (fft_prune.cuh) first, full transform for windows that fail tier 1; 0 = A/B
#endif

#ifndef CERT_GEOM
#define CERT_GEOM 1                           // single-warp windows without downchirp / window tables: the pruned transform works on
                                              // the RAW samples (rotation folded into a per-packet geometric ratio); 0 = A/B
#endif

#ifndef CERT_GEOM_MAX_T
#define CERT_GEOM_MAX_T 256                   // widest window (threads) that takes the geometric variant (A/B knob)
#endif

#ifndef CERT_ONE_LEADER
#define CERT_ONE_LEADER 0                     // 1: lane 0 issues the bulk copies of all the warp's windows in a row -- measured SLOWER (SF7 4845 -> 4051 GB/s):
                                              // the diverged leaders compute their addresses in parallel and only the copy itself goes through the elect loop
#endif

#ifndef CERT_L2_AHEAD
#define CERT_L2_AHEAD 0                       // steps ahead of the walk that are pulled into L2 (0 = off); see the walk loop
#endif

constexpr int CERT_CAP = 512;                 // symbols per packet the fallback bitmap / symbol staging cover (255-byte payload
                                              // in the shipped framing); longer packets take the exact chain throughout
// ---- certification constants (units of u = 2^-24), DESIGN.md section 5 ---------------------------------------------
constexpr float CERT_U = 5.9604644775390625e-08f;
constexpr float CERT_C0 = 384.0f;             // 160 reference (6 radix-4 stages x 19u + input chain + libm + |X|^2) + 96 ours
                                              // + 16 key quantisation + 64 rate tolerance + slack
constexpr float CERT_RATE_TOL = 64.0f * CERT_U;   // |rate' - rate| * N a finished walk tolerates (budgeted in C0)

struct EstSymC { int idx; float fi; float phase; };

struct CertShared {                           // one per pipeline
    EstSymC est[2];
    int sw[2];
    float bc_rate; int bc_toff;
    int k0;                                   // pruned transform: bin class chosen by the window's first thread
    int fail_seq;                             //   bumped when a pruned window fails tier 1 (multi-warp windows; read one step later)
    int tcnt;                                 //   warps of the window that have published their top bin
    unsigned tq[8]; int tbin[8]; float te[8]; //   per warp: top |Y|^2 bits, its bin, energy of the warp's samples
    int tk0[8];                               //   per warp: the bin class it transformed (all must agree)
    alignas(16) cf hp[64];                             //   geometric variant: [k0][m] = (G w_k0)^(2^m), G = exp(j rate T), w = W16 (per packet)
    unsigned fb[CERT_CAP / 32];               // windows that failed certification
    uint16_t sym[CERT_CAP];                   // data symbols of the packet, flushed in one coalesced run
};

template <int SF, int CTA>
struct CertCfg {
    using F = ff::Plan<SF>;
    static constexpr int T = F::T, N = F::N;
    static constexpr int DT = T <= 32 ? 32 : T;            // threads of one pipeline (one packet at a time)
    static constexpr int DOMS = CTA / DT;
    static constexpr int GD = DT / T;                      // windows per pipeline step
    static constexpr int PITCH = F::pitch();
    static constexpr int SLOT = GD * PITCH;
    static constexpr size_t stage_bytes = (size_t)DOMS * SLOT * sizeof(cf);
    static constexpr size_t red_bytes = 256;               // Peak[8] | RedCell[8] | float[8] (multi-warp windows only)
    // pass-0 twiddles W_N^{tid*k}: a [16][T] table in shared memory for SF7-9 (15 conflict-free 64-bit loads per window
    // instead of 30 registers for the kernel's lifetime -- the loop is register-bound there), registers for SF10-12
    static constexpr bool TW0_SMEM = SF <= CERT_TW0_SMEM_MAX_SF;
    static constexpr size_t tw0_bytes = (TW0_SMEM || (SF == 7 && CERT_SOA_SF7 && CERT_SOA_TW_SMEM)) ? (size_t)N * sizeof(cf) : 0;
    // pruned transform: stage factors of the lane transform [4][32], W_R^j [8], W_N^j (SF7 only: its pass-0 twiddles live
    // in the struct-of-arrays layout), and for multi-warp windows the T-entry meeting buffer of every pipeline
    using PR = fp::Prune<SF>;
    static constexpr size_t pr_tws = 0, pr_wr = 4 * 32 * sizeof(cf), pr_twn = pr_wr + 8 * sizeof(cf);
    static constexpr size_t pr_w16 = pr_twn + (SF <= 9 ? (size_t)N * sizeof(cf) : 0);       // W16^j, 16 entries
    static constexpr size_t pr_ub = pr_w16 + 16 * sizeof(cf);
    static constexpr size_t pr_bytes = CERT_PRUNE ? pr_ub + (PR::R > 1 ? (size_t)DOMS * 2 * T * sizeof(cf) : 0) : 0;   // two buffers: steps alternate
    static constexpr size_t fixed_bytes = (size_t)DOMS * sizeof(CertShared) + red_bytes + (size_t)F::TW1_SIZE * sizeof(cf) + tw0_bytes + pr_bytes + 256;
    // as many ring stages as fit CERT_CTAS_PER_SM CTAs into the 227 KB of an SM, at most 3
    static constexpr int fit = (int)(((size_t)227 * 1024 / cert_ctas(SF) - 1024 - fixed_bytes) / stage_bytes);
    static_assert(fit >= 2, "the ring needs two stages");
    static constexpr int STAGES = fit >= 3 ? 3 : 2;
    static constexpr size_t slot_bytes = stage_bytes * STAGES;
    static constexpr size_t bar_off = slot_bytes;
    static constexpr size_t red_off = bar_off + (size_t)DOMS * STAGES * sizeof(uint64_t);
    static constexpr size_t tw1_off = (red_off + red_bytes + 15) & ~(size_t)15;
    static constexpr size_t tw0_off = (tw1_off + (size_t)F::TW1_SIZE * sizeof(cf) + 15) & ~(size_t)15;
    static constexpr size_t ws_off = (tw0_off + tw0_bytes + 15) & ~(size_t)15;
    static constexpr size_t pr_off = (ws_off + (size_t)DOMS * sizeof(CertShared) + 15) & ~(size_t)15;
    static constexpr size_t bytes = pr_off + pr_bytes;
};

struct RedCell { unsigned key; float e; unsigned m1; unsigned m2; };      // per warp, multi-warp windows only

__device__ __forceinline__ unsigned umax3(unsigned a, unsigned b, unsigned c) { return max(max(a, b), c); }

// |Y|^2 with the register index in the low four mantissa bits: (bits & mask) | e as ONE three-input logic op (the mask
// lives in a register: two immediates would cost two instructions)
__device__ __forceinline__ unsigned key_of(float q, unsigned mask, int e)
{
    unsigned k;
    asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(k) : "r"(__float_as_uint(q)), "r"(mask), "r"((unsigned)e));
    return k;
}

template <int DT>
__device__ __forceinline__ void csync()
{
    if (DT <= 32) __syncwarp(); else group_sync<DT>();
}

template <int DT>
__device__ __forceinline__ bool cany(bool p)
{
    if (DT <= 32) return __any_sync(0xffffffffu, p) != 0;
    if (DT >= CTA_THREADS) return __syncthreads_or(p) != 0;
    int r;
    asm volatile("{\n\t.reg .pred q, r;\n\tsetp.ne.b32 q, %1, 0;\n\tbarrier.cta.red.or.pred r, %2, %3, q;\n\tselp.b32 %0, 1, 0, r;\n\t}"
                 : "=r"(r) : "r"((int)p), "r"(1 + (int)threadIdx.x / DT), "n"(DT) : "memory");
    return r != 0;
}

// max over [0, count) of max(|re|,|im|) with the reference's comparison chain (LoRaDemod.cpp:61-68), one pipeline
template <int DT>
__device__ __forceinline__ float cert_scan_max(const cf* __restrict__ pkt, uint32_t count, int lane, float* redf)
{
    float m = 0.0f;
    uint32_t head = 0;
    if ((reinterpret_cast<uintptr_t>(pkt) & 15) != 0 && count > 0) {
        if (lane == 0) { const cf x = ldg_cf(pkt); const float r = fabsf(x.x), i = fabsf(x.y); m = fmaxf(m, (r < i) ? i : r); }
        head = 1;
    }
    const float4* p4 = reinterpret_cast<const float4*>(pkt + head);
    const uint32_t pairs = (count - head) >> 1;
#pragma unroll 8
    for (uint32_t j = lane; j < pairs; j += DT) {
        const float4 q = __ldg(p4 + j);
        const float r0 = fabsf(q.x), i0 = fabsf(q.y), r1 = fabsf(q.z), i1 = fabsf(q.w);
        m = fmaxf(m, (r0 < i0) ? i0 : r0);
        m = fmaxf(m, (r1 < i1) ? i1 : r1);
    }
    if (((count - head) & 1) && lane == 0) {
        const cf x = ldg_cf(pkt + count - 1);
        const float r = fabsf(x.x), i = fabsf(x.y);
        m = fmaxf(m, (r < i) ? i : r);
    }
    if (DT <= 32) {
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, off));
        return m;
    }
    return group_max<DT>(m, redf);
}

// PR_ON = false: the copy without the pruned transform in its walk, for batches the lean route has given up on (noise:
// no window passes tier 1 on one bin class, so the attempt is never made -- but its code sits in the walk loop, and the
// walk then waits on instruction fetches: `no_instruction` 2.0 stalled warps per issue against 0.6 without it).
template <int SF, bool EXTRA, int CTA, bool PR_ON = true>
__global__ void __launch_bounds__(CTA, cert_ctas(SF)) demod_cert_kernel(const DemodArgs a)
{
    using C = CertCfg<SF, CTA>;
    using P = fx::Plan<SF>;
    using F = ff::Plan<SF>;
    constexpr int T = C::T, N = C::N, GD = C::GD, DT = C::DT, STAGES = C::STAGES, NP = F::NP;
    static_assert(T <= 32 || CTA == CTA_THREADS, "multi-warp windows use the CTA_THREADS-wide helpers");
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int dom = threadIdx.x / DT;
    const int lane = threadIdx.x % DT;
    const int gq = lane / T, tid = lane % T;

    cf* slots = reinterpret_cast<cf*>(smem_raw) + (size_t)dom * STAGES * C::SLOT;
    uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw + C::bar_off) + dom * STAGES;
    Peak* red = reinterpret_cast<Peak*>(smem_raw + C::red_off);                   // [8] exact units (group_peak)
    RedCell* redc = reinterpret_cast<RedCell*>(smem_raw + C::red_off + 64);       // [8] fast units, multi-warp windows
    float* redf = reinterpret_cast<float*>(smem_raw + C::red_off + 192);          // [8] group_max
    cf* tw1 = reinterpret_cast<cf*>(smem_raw + C::tw1_off);
    CertShared& ws = reinterpret_cast<CertShared*>(smem_raw + C::ws_off)[dom];

    if (a.only_flagged && a.n_flagged && *a.n_flagged == 0u) return;     // lean route: nothing was handed over
    // ---- start-up: barriers, twiddles, address bases ------------------------------------------------------------
    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; ++s) ptx::mbar_init(&full[s], GD);
        ptx::fence_mbar_init();
    }
    if (NP == 3) {
        constexpr int LO1 = F::lo(1), M1 = 1 << (F::lo(1) + F::width(1));
        for (int j = threadIdx.x; j < F::TW1_SIZE; j += CTA) {
            const int k = j >> LO1, ir = j & ((1 << LO1) - 1);
            double sn, cs;
            sincospi(-2.0 * (double)((k * ir) & (M1 - 1)) / (double)M1, &sn, &cs);
            tw1[j] = fx::mk((float)cs, (float)sn);
        }
    }
    cf tw0[16];                                    // W_N^{tid * k}: correctly rounded; registers (SF10-12) or shared memory
    cf* tw0s = reinterpret_cast<cf*>(smem_raw + C::tw0_off);
    if (C::TW0_SMEM) {
        for (int j = threadIdx.x; j < N; j += CTA) {             // entry [k * T + t]
            double sn, cs;
            sincospi(-2.0 * (double)F::tw0_exp(j % T, j / T) / (double)N, &sn, &cs);
            tw0s[j] = fx::mk((float)cs, (float)sn);
        }
    } else {
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            double sn, cs;
            sincospi(-2.0 * (double)F::tw0_exp(tid, k) / (double)N, &sn, &cs);
            tw0[k] = fx::mk((float)cs, (float)sn);
        }
    }
    constexpr bool SOA = SF == 7 && CERT_SOA_SF7 != 0;
    fs::P2 twr[8], twi[8];                         // SOA: pass-0 twiddle pairs (W^{tid*2q}, W^{tid*(2q+1)})
    constexpr bool SOA_TWS = SOA && CERT_SOA_TW_SMEM != 0;
    unsigned long long* twp = reinterpret_cast<unsigned long long*>(smem_raw + C::tw0_off);   // [plane][q][t] pairs
    if (SOA_TWS) {
        for (int j = threadIdx.x; j < 64; j += CTA) {
            const int q = j >> 3, t = j & 7;
            double s0, c0, s1, c1;
            sincospi(-2.0 * (double)((t * 2 * q) & (N - 1)) / (double)N, &s0, &c0);
            sincospi(-2.0 * (double)((t * (2 * q + 1)) & (N - 1)) / (double)N, &s1, &c1);
            twp[j] = fs::bits(fs::mk2((float)c0, (float)c1));
            twp[64 + j] = fs::bits(fs::mk2((float)s0, (float)s1));
        }
    }
    if (SOA && !SOA_TWS) {
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            double s0, c0, s1, c1;
            sincospi(-2.0 * (double)((tid * 2 * q) & (N - 1)) / (double)N, &s0, &c0);
            sincospi(-2.0 * (double)((tid * (2 * q + 1)) & (N - 1)) / (double)N, &s1, &c1);
            twr[q] = fs::mk2((float)c0, (float)c1);
            twi[q] = fs::mk2((float)s0, (float)s1);
        }
    }
    // pruned transform (fft_prune.cuh): tables and this thread's cross-warp twiddle W_T^{lane * warp}
    using PR = fp::Prune<SF>;
    constexpr bool PRUNE = CERT_PRUNE != 0 && PR_ON;
    cf* pr_tws = reinterpret_cast<cf*>(smem_raw + C::pr_off + C::pr_tws);
    cf* pr_wr = reinterpret_cast<cf*>(smem_raw + C::pr_off + C::pr_wr);
    cf* pr_twn = reinterpret_cast<cf*>(smem_raw + C::pr_off + C::pr_twn);
    cf* pr_w16 = reinterpret_cast<cf*>(smem_raw + C::pr_off + C::pr_w16);
    cf* pr_ub = reinterpret_cast<cf*>(smem_raw + C::pr_off + C::pr_ub) + (PR::R > 1 ? (size_t)dom * 2 * T : 0);
    cf pr_twc = fx::mk(1.0f, 0.0f);
    if (PRUNE) {
        for (int j = threadIdx.x; j < PR::STAGES_TW * PR::W; j += CTA) {
            double sn, cs;
            sincospi(-2.0 * (double)PR::stage_exp(j / PR::W, j % PR::W) / (double)PR::W, &sn, &cs);
            pr_tws[j] = fx::mk((float)cs, (float)sn);
        }
        for (int j = threadIdx.x; j < 8; j += CTA) {
            double sn, cs;
            sincospi(-2.0 * (double)(j & (PR::R - 1)) / (double)PR::R, &sn, &cs);
            pr_wr[j] = fx::mk((float)cs, (float)sn);
        }
        for (int j = threadIdx.x; j < 16; j += CTA) {
            double sn, cs;
            sincospi(-2.0 * (double)j / 16.0, &sn, &cs);
            pr_w16[j] = fx::mk((float)cs, (float)sn);
        }
        if (SF <= 9)
            for (int j = threadIdx.x; j < N; j += CTA) {
                double sn, cs;
                sincospi(-2.0 * (double)j / (double)N, &sn, &cs);
                pr_twn[j] = fx::mk((float)cs, (float)sn);
            }
        if (PR::R > 1) {
            double sn, cs;
            sincospi(-2.0 * (double)PR::cross_exp(tid >> 5, tid & 31) / (double)T, &sn, &cs);
            pr_twc = fx::mk((float)cs, (float)sn);
        }
        if (lane == 0) { ws.k0 = 0; ws.fail_seq = 0; ws.tcnt = 0; }
    }
    const fs::P2 soa_s = fs::mk2(0.70710678118654752440f, 0.70710678118654752440f);
    const fs::P2 soa_ns = fs::mk2(-0.70710678118654752440f, -0.70710678118654752440f);
    const int wr0 = F::wr_base(0, tid), rd0 = F::rd_base(0, tid);
    const int wr1 = NP == 3 ? F::wr_base(1, tid) : 0, rd1 = NP == 3 ? F::rd_base(1, tid) : 0;
    const int ir1 = NP == 3 ? F::i_rest(1, tid) : 0;
    const int bin0 = F::bin_base(tid);
    unsigned keymask = 0xfffffff0u;
    asm volatile("" : "+r"(keymask));              // opaque to constant folding: stays in a register for key_of
    __syncthreads();

    int stage = 0;
    unsigned phase_bits = 0;
    const uint32_t n_pipes = gridDim.x * C::DOMS;
    float pred_max = 0.0f;                         // max-abs of the last packet this pipeline saw
    unsigned st_windows = 0, st_fallback = 0, st_est2 = 0, st_rewalk = 0;      // per-thread statistics, summed at the end
    unsigned st_pruned = 0;
    // pruned-transform policy, uniform over the pipeline: after a window fails tier 1 the next `skip` steps go straight
    // to the full transform; the back-off doubles on every failed retry (noisy traffic settles on the full transform
    // with a vanishing number of retries) and clears on success
    unsigned pr_skip = 0, pr_backoff = 0;
    int pr_seen = 0;

    for (uint32_t pid = blockIdx.x * C::DOMS + dom; pid < a.n_packets; pid += n_pipes) {
        const int lean_flag = a.only_flagged ? (int)a.only_flagged[pid] : 1;
        if (lean_flag == 0) continue;                                   // third pass of the lean route: flagged packets only
        const cf* pkt = a.iq + (size_t)pid * a.packet_stride;
        const uint32_t count = a.sample_counts ? min(a.sample_counts[pid], a.samples_per_packet) : a.samples_per_packet;
        const uint32_t total = count / (uint32_t)N;
        const bool have_sync = total >= 2;
        const bool legacy = a.mode == MODE_LEGACY;
        const uint32_t n_units = (total + GD - 1) / GD;
        const uint32_t n_data = have_sync ? total - 2 : total;
        uint16_t* out_sym = a.symbols + (size_t)pid * a.symbol_stride;

        int flags = 0, status = 0;
        if (!legacy) {                                     // phy.cpp:190-194
            if (count % (uint32_t)N != 0) { flags |= PS_SKIP; status = -22; }
            else if (total < 2) { flags |= PS_SKIP; status = -34; }
        }
        float cfo = 0.0f, toff = 0.0f;
        bool have_metrics = false;
        bool all_exact = total > (uint32_t)CERT_CAP || a.force_exact != 0;
        if (lane < CERT_CAP / 32) ws.fb[lane] = 0u;
        if (lane == 0) { ws.sw[0] = 0; ws.sw[1] = 0; }

        // which max-abs to walk with (LoRaDemod.cpp:61-78): the previous packet's, verified against the samples the walk
        // streams anyway.  A miss costs a second pair of exact estimator symbols and nothing else (measured: scanning
        // first after a miss, as the round-1 kernel did, is slower at every SF -- SF12 24.4 -> 31.3 ms -- because clean
        // batches alternate between max-abs classes <= 1 and 1 + 2^-23 and the scan re-reads HBM for long packets).
        // Scanned first: a packet that may have to be refused (no scratch), one that is exact throughout, and any packet
        // after a non-finite max.
        float use_max = pred_max;
        bool known = !legacy;
        if (legacy && (a.no_scratch || all_exact || !(pred_max < 3.0e38f))) {
            use_max = cert_scan_max<DT>(pkt, count, lane, redf);
            known = true;
            pred_max = use_max;
        }

        float rate = 0.0f, scale = 1.0f, rate_walk = 0.0f;
        int t_off = 0, t_off_walk = 0;
        bool walked = false;
        const bool odd_pkt = window_is_odd(pkt);
        // Lean route, flag 2: the streaming pass has settled every symbol with the rate of a unit-scale estimate and found
        // the packet's max-abs above 1.  Its decisions are scale-invariant (as a finished walk's are, see below), so this
        // pass only repeats the two exact estimator symbols with the true scale -- cfo / time_offset bit for bit -- and
        // keeps the lean symbols when t_off is unchanged and the rate moved by no more than the budgeted tolerance.
        bool verify_only = lean_flag == 2 && legacy && !all_exact;
        if (verify_only) {
            const PacketState ps = a.state[pid];
            walked = true; rate_walk = ps.rate; t_off_walk = ps.t_off;
            use_max = a.pmax[pid]; known = true;
        }

#pragma unroll 1
        for (int attempt = 0; attempt < 3 && !(flags & PS_SKIP); ++attempt) {
            scale = 1.0f;
            flags &= ~PS_NORM;
            if (legacy && use_max > 1.0f) {
                if (a.no_scratch) { flags |= PS_SKIP; status = -34; have_metrics = false; break; }   // LoRaDemod.cpp:70-72
                flags |= PS_NORM;
                scale = __fdiv_rn(1.0f, use_max);
            }
            const bool norm = (flags & PS_NORM) != 0;
            const uint32_t n_est = total < 2 ? total : 2u;

            // ---- E units (exact), LoRaDemod.cpp:80-124 / phy.cpp:93-134 ---------------------------------------------
            for (uint32_t s0 = 0; s0 < n_est; s0 += GD) {
                const uint32_t s = s0 + gq;
                const bool live = s < n_est;
                cf* win = slots + (size_t)stage * C::SLOT + (size_t)gq * C::PITCH;
                if (tid == 0) {
                    if (live) fetch_window<N>(win, pkt + (size_t)s * N, &full[stage]);
                    else ptx::mbar_arrive(&full[stage]);
                }
                ptx::mbar_wait(&full[stage], (phase_bits >> stage) & 1u);
                phase_bits ^= 1u << stage;
                cf v[P::R];
                const cf* raw = win + (odd_pkt ? 1 : 0) + tid;          // s * N is even: the window shares the packet's parity
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
                csync<DT>();                                  // last exchange fully read: the slot becomes the spectrum
#pragma unroll
                for (int r = 0; r < P::R; ++r) win[P::bin_of_reg(tid, r)] = v[r];
                csync<DT>();
                if (tid == 0 && live) {
                    EstSymC e; e.idx = 0; e.fi = 0.0f; e.phase = 0.0f;
                    cf best_bin = fx::mk(0.0f, 0.0f);
                    if (pk.val > 0.0f) {                       // candidate kept iff its power beats -1e30 dB: max > 0
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
                csync<DT>();
                stage = stage + 1 == STAGES ? 0 : stage + 1;
            }

            // ---- combine (lane 0), LoRaDemod.cpp:126-139 -----------------------------------------------------------
            if (lane == 0) {
                const float PI_F = 3.14159265358979323846f;
                float sum_index = 0.0f, phase_diff = 0.0f, prev_phase = 0.0f;
                for (uint32_t s = 0; s < n_est; ++s) {
                    const EstSymC e = ws.est[s];
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
            }
            if (DT <= 32) {
                rate = __shfl_sync(0xffffffffu, rate, 0);
                t_off = __shfl_sync(0xffffffffu, t_off, 0);
            } else {
                if (lane == 0) { ws.bc_rate = rate; ws.bc_toff = t_off; }
                csync<DT>();
                rate = ws.bc_rate; t_off = ws.bc_toff;
            }
            have_metrics = true;

            if (!(use_max < 3.0e38f)) all_exact = true;        // 1/inf == 0 wipes the packet: nothing is scale-invariant there
            if (all_exact) { verify_only = false; break; }
            if (walked) {
                // second estimate after a max-abs misprediction: the finished walk is scale-invariant, so it stands when
                // it used the same window bases and a rate within the tolerance its certification margin budgeted
                if (lane == 0) ++st_est2;
                const bool same = t_off == t_off_walk && fabsf(fx::fsub(rate, rate_walk)) * (float)N <= CERT_RATE_TOL;
                if (same) break;
                verify_only = false;                           // lean symbols do not stand: this pass walks the packet
                if (lane == 0) ++st_rewalk;
                if (lane < CERT_CAP / 32) ws.fb[lane] = 0u;
                csync<DT>();
            }

            // ---- per-packet factors f_c = scale * [down_i] * exp(j*rate*i) * [w_i], i = tid + T*c ------------------
            constexpr bool GEOM = PRUNE && CERT_GEOM != 0 && !EXTRA && T <= CERT_GEOM_MAX_T;
            cf f[16];
            bool f_ready = false;
            auto make_f = [&]() {
#pragma unroll
                for (int c = 0; c < 16; ++c) {
                    const int i = tid + T * c;
                    float ph = fx::fmul(rate, (float)i);
                    ph = __fmaf_rn(-6.283185307179586f, rintf(ph * 0.15915494309189535f), ph);      // into [-pi, pi]
                    float sn, cs;
                    __sincosf(ph, &sn, &cs);
                    cf r = fx::mk(cs * scale, sn * scale);
                    if (EXTRA && a.down) r = ff::fcmul(ldg_cf(a.down + i), r);
                    if (EXTRA && a.window) { const float w = __ldg(a.window + i); r = fx::mk(r.x * w, r.y * w); }
                    f[c] = r;
                }
                f_ready = true;
            };
            // Geometric variant: f_c = F_tid * G^c with F_tid = scale exp(j rate tid), G = exp(j rate T), so the bin class k0
            // of the pruned transform is sum_c r_c h^c, h = G W16^k0 -- a polynomial in the RAW samples.  Per packet: F_tid
            // (one sin/cos.approx, <= 9u), G^(1,2,4,8) in double on four lanes (rate T m is exact: T m is a power of two),
            // and the table hp[k0][m] = G^(2^m) W16^(k0 2^m) (one float complex product: <= 3.5u each).
            cf geo_f = fx::mk(scale, 0.0f), geo_g1 = fx::mk(1.0f, 0.0f);
            if (GEOM) {
                float ph = fx::fmul(rate, (float)tid);
                float sn, cs;
                __sincosf(ph, &sn, &cs);                                   // |rate tid| < 2 pi / 16
                geo_f = fx::mk(cs * scale, sn * scale);
                if (lane < 32) {                                           // the pipeline's first warp
                    double ds = 0.0, dc = 1.0;
                    if (lane < 4) sincos((double)rate * (double)(T << lane), &ds, &dc);
                    const float gx = (float)dc, gy = (float)ds;
#pragma unroll
                    for (int h2 = 0; h2 < 2; ++h2) {
                        const int e = lane + 32 * h2, k = e >> 2, mi = e & 3;      // entry [k][mi]
                        const cf g = fx::mk(__shfl_sync(0xffffffffu, gx, mi), __shfl_sync(0xffffffffu, gy, mi));
                        ws.hp[e] = ff::fcmul(g, pr_w16[(k << mi) & 15]);
                    }
                }
                csync<DT>();
                geo_g1 = ws.hp[0];                                         // G itself (W16^0 = 1)
            } else {
                make_f();
            }
            // certification constant: K = u * (C0 + |start|max + 2|rate|N); accepted iff d^2 > 16 K^2 (N E) q1
            const float start_max = fabsf(rate) * ((float)total * (float)N + fabsf((float)t_off)) * 1.01f;
            const float Kc = CERT_U * (CERT_C0 + start_max + 2.0f * fabsf(rate) * (float)N);
            const float c2 = 16.32f * Kc * Kc * (float)N;
            const float tier1 = (Kc < 0.0078125f) ? fp::TIER1 * (float)N : __int_as_float(0x7f800000);

            // ---- C units: the walk ------------------------------------------------------------------------------
            // Loop bookkeeping is kept out of the issue slots: the window slot, its mbarrier and the prefetch target are
            // cursors advanced by a constant (shared-memory addresses as 32-bit integers), a packet without time shift
            // (t_off == 0, the usual case) has window s at sample s*N with the packet's own parity.
            float m = 0.0f;
            bool suspicious = false;
            const bool fast_pkt = Kc < 0.0078125f;             // tier 1 available at all (pruned windows have no tier 2)
            const bool regular = t_off == 0;
            const uint32_t slot0 = ptx::smem_u32(slots) + (uint32_t)(gq * C::PITCH * sizeof(cf));
            const uint32_t bar0 = ptx::smem_u32(full);
            constexpr uint32_t SLOT_B = C::SLOT * sizeof(cf);
            auto prefetch = [&](uint32_t sp, int stg) {              // leader only: window of symbol sp into stage stg
                if (sp < total) {
                    const uint32_t base = regular ? sp * (uint32_t)N : shifted_base(sp, (uint32_t)N, t_off, count);   // LoRaDemod.cpp:143-150
                    fetch_window_sa<N>(slot0 + (uint32_t)stg * SLOT_B, pkt + base, bar0 + 8u * (uint32_t)stg);
                } else {
                    ptx::mbar_arrive_sa(bar0 + 8u * (uint32_t)stg);
                }
            };
            constexpr bool ONE_LEADER = CERT_ONE_LEADER != 0 && GD > 1;
            const uint32_t slotb = ptx::smem_u32(slots);
            auto prefetch_step = [&](uint32_t s_first, int stg) {    // lane 0 only: the GD windows of one step
#pragma unroll
                for (int g = 0; g < GD; ++g) {
                    const uint32_t sp = s_first + (uint32_t)g;
                    if (sp < total) {
                        const uint32_t base = regular ? sp * (uint32_t)N : shifted_base(sp, (uint32_t)N, t_off, count);
                        fetch_window_sa<N>(slotb + (uint32_t)(g * C::PITCH * sizeof(cf)) + (uint32_t)stg * SLOT_B, pkt + base,
                                           bar0 + 8u * (uint32_t)stg);
                    } else {
                        ptx::mbar_arrive_sa(bar0 + 8u * (uint32_t)stg);
                    }
                }
            };
            if (ONE_LEADER ? lane == 0 : tid == 0) {
                int stg = stage;
                for (uint32_t u = 0; u < (uint32_t)(STAGES - 1) && u < n_units; ++u) {
                    if (ONE_LEADER) prefetch_step(u * GD, stg); else prefetch(u * GD + gq, stg);
                    stg = stg + 1 == STAGES ? 0 : stg + 1;
                }
            }
            uint32_t s = (uint32_t)gq;                              // this window's symbol; += GD per step
            int roff = (odd_pkt ? 1 : 0) + tid;
#pragma unroll 1
            for (uint32_t u = 0; u < n_units; ++u, s += GD) {
                // The slot refilled during this step is the one the PREVIOUS step consumed.  Its refill is issued after
                // this step's first barrier: a thread that has reached that barrier has finished the previous step (its
                // last generic-proxy store there is followed by fence.proxy.async), so no end-of-step barrier is needed.
                constexpr bool EARLY = CERT_EARLY_REFILL != 0 && DT <= 32;
                auto refill_now = [&]() {
                    if (ONE_LEADER ? lane == 0 : tid == 0) {
                        int stg = stage + (STAGES - 1);
                        if (stg >= STAGES) stg -= STAGES;
                        if (u + (STAGES - 1) < n_units) {
                            if (ONE_LEADER) prefetch_step(s + (STAGES - 1) * GD, stg);       // lane 0: s is the step's first symbol
                            else prefetch(s + (STAGES - 1) * GD, stg);
                        }
                    }
                };
                auto refill = [&]() { if (!EARLY) refill_now(); };
                // The ring holds one or two steps per pipeline -- 64 KB in flight per SM with 16 warps, short of what the HBM
                // latency-bandwidth product asks for (~100 KB per SM).  One lane therefore pulls the windows CERT_L2_AHEAD steps
                // ahead (this packet's, or the first ones of the pipeline's next packet) into L2 with a bulk prefetch: no
                // shared memory, nothing to wait for, and the ring's own copies then complete at L2 latency.
                if (CERT_L2_AHEAD > 0 && lane == 0) {
                    const uint32_t v = u + CERT_L2_AHEAD;
                    const cf* src = nullptr;
                    if (v < n_units) src = pkt + (size_t)v * GD * N;
                    else if ((unsigned long long)pid + n_pipes < a.n_packets && v - n_units < n_units)
                        src = pkt + (size_t)n_pipes * a.packet_stride + (size_t)(v - n_units) * GD * N;
                    if (src) ptx::bulk_prefetch_l2(reinterpret_cast<const void*>(reinterpret_cast<uintptr_t>(src) & ~(uintptr_t)15),
                                                   (unsigned)(GD * N * sizeof(cf)));
                }
                if (EARLY) refill_now();
                const bool active = s < total;
                if (!regular) {
                    const uint32_t base = shifted_base(s, (uint32_t)N, t_off, count);
                    roff = ((odd_pkt != ((base & 1u) != 0)) ? 1 : 0) + tid;
                }
                ptx::mbar_wait_sa(bar0 + 8u * (uint32_t)stage, (phase_bits >> stage) & 1u);
                phase_bits ^= 1u << stage;
                cf* win = slots + (size_t)stage * C::SLOT + (size_t)gq * C::PITCH;
                const cf* raw = win + roff;

                const bool fast = PRUNE && pr_skip == 0 && fast_pkt;
                bool pr_failed = false;
                bool refilled = false;
                auto refill_once = [&]() { if (!refilled) { refill(); refilled = true; } };
                cf v[16];
                float mw = 0.0f;
                cf e0 = fx::mk(0.0f, 0.0f), e1 = fx::mk(0.0f, 0.0f);                   // (sum re^2, sum im^2), two chains
                float en;
                auto load_rotated = [&]() {
                    // rotated samples: the full transform and the pruned transform of the variants with tables
                    if (GEOM && !f_ready) make_f();
                    e0 = fx::mk(0.0f, 0.0f); e1 = fx::mk(0.0f, 0.0f);
#pragma unroll
                    for (int c = 0; c < 16; ++c) {
                        const cf r = raw[T * c];
                        mw = fmaxf(fmaxf(mw, fabsf(r.x)), fabsf(r.y));                 // running max-abs (LoRaDemod.cpp:61-68)
                        const cf y = ff::fcmul(r, f[c]);
                        if (c & 1) e1 = ff::sqacc(e1, y); else e0 = ff::sqacc(e0, y);  // one packed FMA per sample
                        v[c] = y;
                    }
                    en = (e0.x + e0.y) + (e1.x + e1.y);
                };
                // the full transform runs when the pruned one is not tried -- or, for single-warp windows, when it has just
                // failed: the raw samples are still in the slot (the pruned transform never writes there)
                bool full = !(PRUNE && fast);
                if (PRUNE && fast && !GEOM) {
                    load_rotated();
                } else if (PRUNE && fast) {
#pragma unroll
                    for (int c = 0; c < 16; ++c) {
                        const cf r = raw[T * c];
                        mw = fmaxf(fmaxf(mw, fabsf(r.x)), fabsf(r.y));
                        if (c & 1) e1 = ff::sqacc(e1, r); else e0 = ff::sqacc(e0, r);
                        v[c] = r;
                    }
                    // energy of the rotated samples: |f_c| = scale up to the rounding of sin^2 + cos^2 (a hair of head-room)
                    en = ((e0.x + e0.y) + (e1.x + e1.y)) * (scale * scale) * 1.000002f;
                }
                if (PRUNE && fast) {
                    if (active) m = fmaxf(m, mw);
                    // ---- pruned transform (fft_prune.cuh): the tone's bin class k0, then ONE value per thread -----------------
                    int k0 = 0;
                    cf cval;
                    if (GEOM) {
                        // Geometric variant, on the raw samples.  Bin class: the phase step between neighbouring samples of a
                        // thread is 2 pi (bin mod 16) / 16 once the rotation's own step G is added; four pairs per thread,
                        // summed over the warp.  (Every warp of a multi-warp window decides for itself -- a tone gives them
                        // all the same answer -- and the choices are compared before anything is accepted.)
                        if (T <= 32 || (tid >> 5) == 0) {              // multi-warp windows: the first warp decides
                            cf z = fx::mk(0.0f, 0.0f);
#pragma unroll
                            for (int c = 0; c < 16; c += 4) {
                                z.x = ff::fma_(v[c + 1].x, v[c].x, z.x); z.x = ff::fma_(v[c + 1].y, v[c].y, z.x);
                                z.y = ff::fma_(v[c + 1].y, v[c].x, z.y); z.y = ff::fma_(-v[c + 1].x, v[c].y, z.y);
                            }
#pragma unroll
                            for (int off = PR::W / 2; off > 0; off >>= 1) {
                                z.x += __shfl_xor_sync(0xffffffffu, z.x, off);
                                z.y += __shfl_xor_sync(0xffffffffu, z.y, off);
                            }
                            z = ff::fcmul(z, geo_g1);
                            // nearest multiple of 22.5 degrees, by octant folding (an inexact answer only costs a failed test)
                            const float ax = fabsf(z.x), ay = fabsf(z.y);
                            const float mn = fminf(ax, ay), mx = fmaxf(ax, ay);
                            int q = (mn > 0.19891237f * mx ? 1 : 0) + (mn > 0.66817864f * mx ? 1 : 0);
                            if (ay > ax) q = 4 - q;
                            if (z.x < 0.0f) q = 8 - q;
                            if (z.y < 0.0f) q = 16 - q;
                            k0 = q & 15;
                            if (T > 32 && tid == 0) ws.k0 = k0;
                        }
                        if (T <= 32) {
                            __syncwarp();
                        } else {
                            group_sync<T>();                   // k0 visible; every thread of the window has its raw samples
                            k0 = ws.k0;
                        }
                        refill_once();
                        // sum_c r_c h^c, h = G W16^k0, by halving: four levels of complex multiply-adds with h^8, h^4, h^2, h
                        const float4 hA = *reinterpret_cast<const float4*>(&ws.hp[4 * k0]);          // h, h^2
                        const float4 hB = *reinterpret_cast<const float4*>(&ws.hp[4 * k0 + 2]);      // h^4, h^8
                        const cf h1 = fx::mk(hA.x, hA.y), h2 = fx::mk(hA.z, hA.w), h4 = fx::mk(hB.x, hB.y), h8 = fx::mk(hB.z, hB.w);
                        auto cma = [](cf acc, cf a, cf w) {               // acc + a * w: four fused multiply-adds
                            return fx::mk(ff::fma_(-a.y, w.y, ff::fma_(a.x, w.x, acc.x)), ff::fma_(a.y, w.x, ff::fma_(a.x, w.y, acc.y)));
                        };
#pragma unroll
                        for (int c = 0; c < 8; ++c) v[c] = cma(v[c], v[c + 8], h8);
#pragma unroll
                        for (int c = 0; c < 4; ++c) v[c] = cma(v[c], v[c + 4], h4);
#pragma unroll
                        for (int c = 0; c < 2; ++c) v[c] = cma(v[c], v[c + 2], h2);
                        const cf pv = cma(v[0], v[1], h1);
                        cf tw = fx::mk(1.0f, 0.0f);
                        if (T <= 32) {
                            tw = pr_twn[(tid * k0) & (N - 1)];
                        } else {
                            switch (k0) {                                  // uniform over the warp
#define CERT_PICK(K) case K: tw = tw0[K]; break;
                                CERT_PICK(1) CERT_PICK(2) CERT_PICK(3) CERT_PICK(4) CERT_PICK(5) CERT_PICK(6) CERT_PICK(7) CERT_PICK(8)
                                CERT_PICK(9) CERT_PICK(10) CERT_PICK(11) CERT_PICK(12) CERT_PICK(13) CERT_PICK(14) CERT_PICK(15)
#undef CERT_PICK
                                default: break;
                            }
                        }
                        cval = ff::fcmul(ff::fcmul(pv, tw), geo_f);
                    } else {
                        // Variants with downchirp / window tables: pass 0 on the rotated samples, class from the window's first
                        // thread (its 16 outputs ranked by |re| + |im|), that output times its pass-0 twiddle
                        ff::dft16(v);
                        if (T <= 32 || (tid >> 5) == 0) {
                            unsigned kb = 0u;
#pragma unroll
                            for (int e = 0; e < 16; e += 2)
                                kb = umax3(kb, key_of(fabsf(v[e].x) + fabsf(v[e].y), keymask, e),
                                           key_of(fabsf(v[e + 1].x) + fabsf(v[e + 1].y), keymask, e + 1));
                            if (T <= 32) k0 = __shfl_sync(0xffffffffu, (int)(kb & 15u), (int)(threadIdx.x & 31u) & ~(T - 1));
                            else if (tid == 0) ws.k0 = (int)(kb & 15u);
                        }
                        if (T > 32) {
                            group_sync<T>();                   // k0 visible; every thread of the window has its raw samples
                            k0 = ws.k0;
                        } else {
                            __syncwarp();
                        }
                        refill_once();
                        cf sel = v[0], tw = fx::mk(1.0f, 0.0f);
                        constexpr bool TW_TABLE = T < 32;      // several windows per warp: k0 differs between lanes -> no branch
                        if (T < 32) {
                            // binary select tree on the bits of k0 (30 selects), twiddle from the W_N table
                            const bool b0 = (k0 & 1) != 0, b1 = (k0 & 2) != 0, b2 = (k0 & 4) != 0, b3 = (k0 & 8) != 0;
                            cf t8[8], t4[4], t2[2];
#pragma unroll
                            for (int j = 0; j < 8; ++j) t8[j] = fx::mk(b0 ? v[2 * j + 1].x : v[2 * j].x, b0 ? v[2 * j + 1].y : v[2 * j].y);
#pragma unroll
                            for (int j = 0; j < 4; ++j) t4[j] = fx::mk(b1 ? t8[2 * j + 1].x : t8[2 * j].x, b1 ? t8[2 * j + 1].y : t8[2 * j].y);
#pragma unroll
                            for (int j = 0; j < 2; ++j) t2[j] = fx::mk(b2 ? t4[2 * j + 1].x : t4[2 * j].x, b2 ? t4[2 * j + 1].y : t4[2 * j].y);
                            sel = fx::mk(b3 ? t2[1].x : t2[0].x, b3 ? t2[1].y : t2[0].y);
                        } else
                        switch (k0) {
#define CERT_PICK(K) case K: sel = v[K]; tw = tw0[K]; break;
                            CERT_PICK(1) CERT_PICK(2) CERT_PICK(3) CERT_PICK(4) CERT_PICK(5) CERT_PICK(6) CERT_PICK(7) CERT_PICK(8)
                            CERT_PICK(9) CERT_PICK(10) CERT_PICK(11) CERT_PICK(12) CERT_PICK(13) CERT_PICK(14) CERT_PICK(15)
#undef CERT_PICK
                            default: break;
                        }
                        if (TW_TABLE) tw = pr_twn[(tid * k0) & (N - 1)];
                        cval = ff::fcmul(sel, tw);
                    }
                    if (PR::R > 1) {
                        // T > 32: warp ka forms its output of the R-point transform across the window's warps.  Two meeting
                        // buffers, alternating: a warp may start the next step while another still reads this one.
                        cf* ub = pr_ub + (u & 1u) * T;
                        ub[tid] = cval;
                        group_sync<T>();
                        refill_once();
                        const int fsq = ws.fail_seq;           // verdict of the previous pruned step, published before this barrier
                        if (fsq != pr_seen) { pr_seen = fsq; pr_failed = true; }
                        const int ka = tid >> 5, l5 = tid & 31;
                        cf acc = ub[l5];
#pragma unroll
                        for (int w = 1; w < PR::R; ++w) {
                            const cf t = ff::fcmul(ub[l5 + 32 * w], pr_wr[(w * ka) & (PR::R - 1)]);
                            acc = ff::add(acc, t);
                        }
                        cval = ff::fcmul(acc, pr_twc);
                    }
                    const int lw = tid & (PR::W - 1);
                    const cf yv = fp::lane_dft<SF>(cval, lw, pr_tws);
                    const unsigned qb = __float_as_uint(ff::fnorm(yv));       // >= +0, or NaN (sorts above everything)
                    unsigned qmax = qb;
#pragma unroll
                    for (int off = PR::W / 2; off > 0; off >>= 1) {
                        qmax = max(qmax, __shfl_xor_sync(0xffffffffu, qmax, off));
                        en += __shfl_xor_sync(0xffffffffu, en, off);
                    }
                    if (PR::R == 1) {
                        const float q1 = __uint_as_float(qmax);
                        const bool cert = (q1 > tier1 * en) && (en < 1.0e30f) && (q1 >= 8.3e-25f);
                        if (active) {
                            if (!cert && !(en < 1.0e30f)) suspicious = true;
                            if (qb == qmax && cert) {                          // the lane that holds the top bin (unique when certified)
                                const int idx = k0 + 16 * (int)(__brev((unsigned)lw) >> (32 - PR::LW));
                                if (have_sync && s < 2) ws.sw[s] = idx;
                                else ws.sym[have_sync ? s - 2 : s] = (uint16_t)idx;
                            }
                        }
                        pr_failed = __any_sync(0xffffffffu, active && !cert) != 0;
                        if (pr_failed) full = true;                            // same step, full transform (tier 1 / tier 2)
                        else if (tid == 0 && active) ++st_pruned;
                    } else {
                        // every warp publishes its top bin; the last one to arrive decides for the window
                        const unsigned ball = __ballot_sync(0xffffffffu, qb == qmax);
                        if ((threadIdx.x & 31u) == 0u) {
                            const int ka = tid >> 5;
                            const int lt = __ffs((int)ball) - 1;
                            ws.tq[ka] = qmax;
                            ws.tbin[ka] = k0 + 16 * (ka + PR::R * (int)(__brev((unsigned)lt) >> (32 - PR::LW)));
                            ws.te[ka] = en;
                            ws.tk0[ka] = k0;
                            __threadfence_block();
                            const int ticket = atomicAdd(&ws.tcnt, 1);
                            if (ticket == PR::R - 1) {
                                __threadfence_block();
                                ws.tcnt = 0;
                                unsigned bq = 0u; int bb = 0; float et = 0.0f;
                                bool agree = true;                 // every warp transformed the same bin class
#pragma unroll
                                for (int w = 0; w < PR::R; ++w) {
                                    const unsigned qw = ws.tq[w];
                                    et += ws.te[w];
                                    agree = agree && ws.tk0[w] == k0;
                                    if (qw > bq) { bq = qw; bb = ws.tbin[w]; }
                                }
                                const float q1 = __uint_as_float(bq);
                                const bool cert = agree && (q1 > tier1 * et) && (et < 1.0e30f) && (q1 >= 8.3e-25f);
                                if (active) {
                                    if (cert) {
                                        if (have_sync && s < 2) ws.sw[s] = bb;
                                        else ws.sym[have_sync ? s - 2 : s] = (uint16_t)bb;
                                        ++st_pruned;
                                    } else {
                                        if (!(et < 1.0e30f)) suspicious = true;
                                        atomicOr(&ws.fb[s >> 5], 1u << (s & 31));
                                        ws.fail_seq = ws.fail_seq + 1;
                                    }
                                }
                            }
                        }
                    }
                }
                if (full) {
                load_rotated();
                if (active) m = fmaxf(m, mw);
                unsigned key[16];
#if defined(CERT_ABLATE)
                // measurement-only build (tools/ablate.sh, never the product): the walk without its transform -- what the
                // streaming skeleton (ring, rotation, max / energy tracking, keys, reductions, bookkeeping) costs on its own
                if constexpr (true) {
                    group_sync<T>();
                    refill_once();
#pragma unroll
                    for (int e = 0; e < 16; ++e) key[e] = key_of(ff::fnorm(v[e]), keymask, e);
                } else if constexpr (SOA) {
#else
                if constexpr (SOA) {
#endif
                    // struct-of-arrays transform (fft_soa.cuh): every butterfly / twiddle / modulus op is packed over two points
                    fs::P2 pr[8], pi[8];
                    fs::pass0_r16(v, pr, pi, soa_s, soa_ns);
#pragma unroll
                    for (int q = 0; q < 8; ++q) {
                        if (SOA_TWS) fs::cmul(pr[q], pi[q], fs::from_bits(twp[8 * q + tid]), fs::from_bits(twp[64 + 8 * q + tid]));
                        else fs::cmul(pr[q], pi[q], twr[q], twi[q]);
                    }
                    unsigned long long* ex = reinterpret_cast<unsigned long long*>(win);
                    group_sync<T>();                           // every thread of the window has its raw samples
                    refill_once();
#pragma unroll
                    for (int q = 0; q < 8; ++q) {
                        ex[fs::sf7_wr(0, q) + tid] = fs::bits(pr[q]);
                        ex[fs::SF7_PLANE + fs::sf7_wr(0, q) + tid] = fs::bits(pi[q]);
                    }
                    group_sync<T>();
#pragma unroll
                    for (int r = 0; r < 8; ++r) {
                        pr[r] = fs::from_bits(ex[9 * tid + r]);
                        pi[r] = fs::from_bits(ex[fs::SF7_PLANE + 9 * tid + r]);
                    }
                    fs::dft8(pr, pi, soa_s, soa_ns);
#pragma unroll
                    for (int r = 0; r < 8; ++r) {
                        const fs::P2 q2 = fs::norm2(pr[r], pi[r]);
                        key[r] = key_of(fs::lo(q2), keymask, r);          // bin 2 tid + 16 r
                        key[r + 8] = key_of(fs::hi(q2), keymask, r + 8);  // bin 2 tid + 1 + 16 r
                    }
                } else {
                // pass 0 in registers, twiddle, exchange(s) in the window's own slot
                ff::dft16(v);
#pragma unroll
                for (int k = 1; k < 16; ++k) v[k] = ff::fcmul(v[k], C::TW0_SMEM ? tw0s[k * T + tid] : tw0[k]);
                // exchange 0 in the window's own slot.  With natural strides (SF12) a thread overwrites exactly the entries it
                // read its own samples from, unless the window sits one entry later (odd alignment): no barrier needed first.
                constexpr bool OWN = F::stride(0, SF - 1) == (1 << (SF - 1));
                const bool pre_sync = !OWN || roff != tid;
                if (pre_sync) { group_sync<T>(); refill_once(); }  // every thread of the window has its raw samples
#pragma unroll
                for (int e = 0; e < 16; ++e) win[wr0 + F::wr_off(0, e)] = v[e];
                group_sync<T>();
                if (!pre_sync) refill_once();
#pragma unroll
                for (int e = 0; e < 16; ++e) v[e] = win[rd0 + F::rd_off(0, e)];
                F::butterflies(1, v);
                if (NP == 3) {
                    constexpr int W1 = F::width(1), LO1 = F::lo(1);
#pragma unroll
                    for (int k = 1; k < (1 << W1); ++k) {
                        const cf t = tw1[(k << LO1) + ir1];
#pragma unroll
                        for (int e = k; e < 16; e += (1 << W1)) v[e] = ff::fcmul(v[e], t);
                    }
                    group_sync<T>();
#pragma unroll
                    for (int e = 0; e < 16; ++e) win[wr1 + F::wr_off(1, e)] = v[e];
                    group_sync<T>();
#pragma unroll
                    for (int e = 0; e < 16; ++e) v[e] = win[rd1 + F::rd_off(1, e)];
                    F::butterflies(2, v);
                }

                // |Y|^2 keyed with the register index in the low four mantissa bits; top key of the window
#pragma unroll
                for (int e = 0; e < 16; ++e) key[e] = key_of(ff::fnorm(v[e]), keymask, e);
                }
                unsigned kloc = 0u;
#pragma unroll
                for (int e = 0; e < 16; e += 2) kloc = umax3(kloc, key[e], key[e + 1]);
                unsigned kmax = kloc;
                constexpr int W = T < 32 ? T : 32;
#pragma unroll
                for (int off = W / 2; off > 0; off >>= 1) {
                    kmax = max(kmax, __shfl_xor_sync(0xffffffffu, kmax, off));
                    en += __shfl_xor_sync(0xffffffffu, en, off);
                }
                if (T > 32) {
                    const int warp = threadIdx.x >> 5;
                    if ((threadIdx.x & 31) == 0) { redc[warp].key = kmax; redc[warp].e = en; }
                    group_sync<T>();
                    const int w0 = (threadIdx.x / T) * (T / 32);
                    kmax = redc[w0].key; en = redc[w0].e;
#pragma unroll
                    for (int k = 1; k < T / 32; ++k) { kmax = max(kmax, redc[w0 + k].key); en += redc[w0 + k].e; }
                }
                const float q1 = __uint_as_float(kmax & ~15u);
                // tier 1: the top bin holds more than 9/16 of N*E (Parseval: every other bin is below half its modulus)
                bool cert = (q1 > tier1 * en) && (en < 1.0e30f) && (q1 >= 8.3e-25f);
#if defined(CERT_ABLATE)
                cert = true;
#endif
                // (a window that is the whole pipeline needs no vote: cert is uniform across it)
                if (GD == 1 ? (!cert && active) : cany<DT>(!cert && active)) {
                    // tier 2: the true second maximum of the window.  Keys inside a thread are distinct (register index in
                    // the low bits); (m1, m2) pairs merge as sorted pairs, so a key held by two threads shows up as m2 == m1
                    unsigned m1 = kloc, m2 = 0u;
#pragma unroll
                    for (int e = 0; e < 16; ++e) m2 = max(m2, key[e] == kloc ? 0u : key[e]);
#pragma unroll
                    for (int off = W / 2; off > 0; off >>= 1) {
                        const unsigned o1 = __shfl_xor_sync(0xffffffffu, m1, off), o2 = __shfl_xor_sync(0xffffffffu, m2, off);
                        m2 = umax3(min(m1, o1), m2, o2);
                        m1 = max(m1, o1);
                    }
                    if (T > 32) {
                        const int warp = threadIdx.x >> 5;
                        if ((threadIdx.x & 31) == 0) { redc[warp].m1 = m1; redc[warp].m2 = m2; }
                        group_sync<T>();
                        const int w0 = (threadIdx.x / T) * (T / 32);
                        m1 = redc[w0].m1; m2 = redc[w0].m2;
#pragma unroll
                        for (int k = 1; k < T / 32; ++k) {
                            const unsigned o1 = redc[w0 + k].m1, o2 = redc[w0 + k].m2;
                            m2 = umax3(min(m1, o1), m2, o2);
                            m1 = max(m1, o1);
                        }
                    }
                    const float q2 = __uint_as_float(m2 & ~15u);
                    const float d = q1 - q2 * 1.000004f;
                    const bool t2 = (d > 0.0f) && (d * d > c2 * en * q1) && (en < 1.0e30f) && (q1 >= 8.3e-25f) && (q1 < 1.0e30f);
                    cert = cert || t2;
                }
                if (active) {
                    if (!cert && !(en < 1.0e30f)) suspicious = true;       // NaN / Inf in the window: invisible to the tracked max
                    if (kloc == kmax) {                                    // the lane that holds the top bin
                        if (cert) {
                            const unsigned e = kmax & 15u;
                            const int idx = SOA ? fs::sf7_bin(tid, (int)(e & 7u), (int)(e >> 3))
                                                : bin0 + ((e & 1u) ? F::bin_off(1) : 0) + ((e & 2u) ? F::bin_off(2) : 0) +
                                                  ((e & 4u) ? F::bin_off(4) : 0) + ((e & 8u) ? F::bin_off(8) : 0);
                            if (have_sync && s < 2) ws.sw[s] = idx;
                            else ws.sym[have_sync ? s - 2 : s] = (uint16_t)idx;
                        } else {
                            atomicOr(&ws.fb[s >> 5], 1u << (s & 31));
                        }
                    }
                }
                if (pr_skip) --pr_skip;
                }                                              // full transform
                if (PRUNE && fast) {
                    if (pr_failed) { pr_backoff = min(2u * pr_backoff + 16u, 4096u); pr_skip = pr_backoff; }
                    else pr_backoff = 0u;
                }
                ptx::fence_proxy_async();                      // this thread's stores to the slot precede its next refill
                if (EARLY) csync<DT>();
                stage = stage + 1 == STAGES ? 0 : stage + 1;
            }
            csync<DT>();                                       // the last step's slot and staging cells are settled
            if (lane == 0) st_windows += total;
            walked = true;
            rate_walk = rate;
            t_off_walk = t_off;

            // ---- verify the trusted max against the samples actually walked -----------------------------------------
            if (known) break;
            if (cany<DT>(suspicious)) {
                m = cert_scan_max<DT>(pkt, count, lane, redf);       // NaN somewhere: the reference's own comparison chain
            } else {
                if (DT <= 32) {
#pragma unroll
                    for (int off = 16; off > 0; off >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, off));
                } else {
                    m = group_max<DT>(m, redf);
                }
                // Shifted windows (LoRaDemod.cpp:143-150) and a partial last symbol leave a few samples the walk never
                // read: with t_off > 0 the windows cover [t_off, total N) (the last ones stay unshifted), with t_off < 0
                // [0, total N - |t_off|).  Only those ends are scanned (a superset is harmless for a maximum) -- a full
                // second pass over the packet doubled the HBM traffic of every noisy packet.
                const uint32_t whole = total * (uint32_t)N;
                if (t_off > 0) {
                    const uint32_t n0 = min((uint32_t)t_off, count);
                    m = fmaxf(m, cert_scan_max<DT>(pkt, n0, lane, redf));
                    if (count > whole) m = fmaxf(m, cert_scan_max<DT>(pkt + whole, count - whole, lane, redf));
                } else if (t_off < 0) {
                    const uint32_t a0 = (uint32_t)(-t_off), lo = whole > a0 ? whole - a0 : 0u;
                    m = fmaxf(m, cert_scan_max<DT>(pkt + lo, count - lo, lane, redf));
                } else if (count != whole) {
                    m = fmaxf(m, cert_scan_max<DT>(pkt + whole, count - whole, lane, redf));
                }
            }
            const bool ok = (m > 1.0f) ? (m == use_max) : !(use_max > 1.0f);
            pred_max = m;
            known = true;
            if (ok) break;
            use_max = m;                                   // redo the two E units with the packet's own max
        }

        // ---- exact units: windows that failed certification (or the whole packet), LoRaDemod.cpp:142-176 ------------
        csync<DT>();
        const bool run = !(flags & PS_SKIP);
        bool any_fb = all_exact;
        if (!all_exact) {
            bool mine = false;
            if (lane < CERT_CAP / 32) mine = ws.fb[lane] != 0u;
            any_fb = cany<DT>(mine);
        }
        if (run && any_fb) {
            uint32_t cursor = 0;
            while (true) {
                int mine_s = -1;
                bool found = false;
#pragma unroll 1
                for (int g = 0; g < GD; ++g) {
                    uint32_t s = cursor;
                    if (!all_exact) {
                        while (s < total && s < (uint32_t)CERT_CAP) {
                            const unsigned w = ws.fb[s >> 5] >> (s & 31);
                            if (w) { s += __ffs(w) - 1; break; }
                            s = (s | 31u) + 1u;
                        }
                        if (s >= (uint32_t)CERT_CAP) s = total;
                    }
                    if (s < total) { if (g == gq) mine_s = (int)s; cursor = s + 1; found = true; }
                    else cursor = total;
                }
                if (!found) break;
                const bool active = mine_s >= 0;
                const uint32_t s = active ? (uint32_t)mine_s : 0u;
                const uint32_t base = shifted_base(s, (uint32_t)N, t_off, count);
                cf* win = slots + (size_t)stage * C::SLOT + (size_t)gq * C::PITCH;
                if (tid == 0) {
                    if (active) fetch_window<N>(win, pkt + base, &full[stage]);
                    else ptx::mbar_arrive(&full[stage]);
                }
                const bool odd = odd_pkt != ((base & 1u) != 0);
                const float start = start_phase(rate, s, (uint32_t)N, t_off);           // LoRaDemod.cpp:152-153
                const bool slow = phases_need_slow_path(start, fx::fadd(start, fx::fmul(rate, (float)(N - 1))));
                ptx::mbar_wait(&full[stage], (phase_bits >> stage) & 1u);
                phase_bits ^= 1u << stage;
                cf v[P::R];
                chain_window<SF, EXTRA, false>(v, win + (odd ? 1 : 0) + tid, tid, scale, rate, start, slow, a.down, a.window);
                fft_inplace<SF, false>(v, tid, a.tw, a.ct, win);
                const Peak pk = group_peak<SF>(local_peak<SF>(v, tid), red);
                if (tid == 0 && active) {
                    const uint32_t ds = have_sync ? s - 2 : s;
                    if (have_sync && s < 2) ws.sw[s] = pk.idx;
                    else if (ds < (uint32_t)CERT_CAP) ws.sym[ds] = (uint16_t)pk.idx;
                    else out_sym[ds] = (uint16_t)pk.idx;                 // beyond the staging buffer (exact-throughout packets only)
                    if (all_exact) ++st_windows;
                    ++st_fallback;
                }
                ptx::fence_proxy_async();
                csync<DT>();
                stage = stage + 1 == STAGES ? 0 : stage + 1;
            }
        }

        // ---- bookkeeping: LoRaDemod.cpp:166-196; the packet's symbols leave in one coalesced run ----------------------
        csync<DT>();
        if (run && !verify_only) {
            const uint32_t n_stage = n_data < (uint32_t)CERT_CAP ? n_data : (uint32_t)CERT_CAP;
            for (uint32_t i = lane; i < n_stage; i += DT) out_sym[i] = ws.sym[i];
        }
        if (lane == 0) {
            if (run) status = have_sync ? (int)(total - 2) : (int)total;
            if (have_metrics) {
                if (a.cfo) a.cfo[pid] = cfo;
                if (a.time_offset) a.time_offset[pid] = toff;
            }
            if (a.sync && !verify_only) {
                const unsigned shift = SF > 4 ? SF - 4 : 0;
                a.sync[pid] = (run && have_sync) ? (uint8_t)((((ws.sw[0] >> shift) & 0xf) << 4) | ((ws.sw[1] >> shift) & 0xf)) : 0;
            }
            if (a.status) a.status[pid] = status;
        }
        csync<DT>();
    }

    if (a.cert_stats) {
        if (st_windows) atomicAdd(a.cert_stats + 0, (unsigned long long)st_windows);
        if (st_fallback) atomicAdd(a.cert_stats + 1, (unsigned long long)st_fallback);
        if (lane == 0 && st_est2) atomicAdd(a.cert_stats + 2, (unsigned long long)st_est2);
        if (lane == 0 && st_rewalk) atomicAdd(a.cert_stats + 3, (unsigned long long)st_rewalk);
        if (st_pruned) atomicAdd(a.cert_stats + 4, (unsigned long long)st_pruned);
    }
}

template <int SF, int CTA>
static cudaError_t launch_cert_sf(const DemodArgs& a, int sm_count, cudaStream_t stream)
{
    using C = CertCfg<SF, CTA>;
    const bool extra = a.down != nullptr || a.window != nullptr;
    static std::atomic<bool> attr_done[64];
    int dev = 0;
    cudaGetDevice(&dev);
    if (!attr_done[dev & 63].load(std::memory_order_acquire)) {
        cudaFuncSetAttribute(demod_cert_kernel<SF, false, CTA>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::bytes);
        cudaFuncSetAttribute(demod_cert_kernel<SF, true, CTA>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::bytes);
        cudaFuncSetAttribute(demod_cert_kernel<SF, false, CTA, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::bytes);
        attr_done[dev & 63].store(true, std::memory_order_release);
    }
    const unsigned max_ctas = (unsigned)sm_count * cert_ctas(SF);         // persistent grid: every CTA resident
    const unsigned need = (a.n_packets + C::DOMS - 1) / C::DOMS;
    const unsigned grid = need < max_ctas ? need : max_ctas;
    if (extra) demod_cert_kernel<SF, true, CTA><<<grid, CTA, C::bytes, stream>>>(a);
    else if (a.no_prune) demod_cert_kernel<SF, false, CTA, false><<<grid, CTA, C::bytes, stream>>>(a);
    else demod_cert_kernel<SF, false, CTA><<<grid, CTA, C::bytes, stream>>>(a);
    return cudaGetLastError();
}

bool cert_supported(int sf, const DemodArgs& a)
{
    return sf >= 7 && sf <= 12 && a.osr == 1 && (a.mode == MODE_LEGACY || a.mode == MODE_HIGHLEVEL) &&
           (unsigned long long)a.samples_per_packet < 0xf0000000ull;
}

cudaError_t launch_demod_cert(int sf, const DemodArgs& a, int sm_count, cudaStream_t stream)
{
    switch (sf) {
    case 7: return launch_cert_sf<7, CERT_CTA_SMALL>(a, sm_count, stream);
    case 8: return launch_cert_sf<8, CERT_CTA_SMALL>(a, sm_count, stream);
    case 9: return launch_cert_sf<9, CERT_CTA_SMALL>(a, sm_count, stream);
    case 10: return launch_cert_sf<10, CTA_THREADS>(a, sm_count, stream);
    case 11: return launch_cert_sf<11, CTA_THREADS>(a, sm_count, stream);
    case 12: return launch_cert_sf<12, CTA_THREADS>(a, sm_count, stream);
    default: return cudaErrorInvalidValue;
    }
}

} // namespace lk
