// api.cu -- the C ABI of include/lora_phy_cuda.h: context, plan tables, argument checks with
// the reference's errno contract, launches, and the host-buffer staging variants.
#include "../../include/lora_phy_cuda.h"
#include "demod_device.cuh"
#include "internal.h"

#include <algorithm>
#include <cerrno>
#include <cmath>
#include <complex>
#include <vector>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <new>
#include <string>
#include <tuple>

using lk::cf;

namespace {

struct SfTables { cf* tw = nullptr; cf* ct = nullptr; float* hann = nullptr; };

struct DevBuf {
    void* p = nullptr; size_t cap = 0;
    cudaError_t reserve(size_t bytes)
    {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        size_t want = bytes + bytes / 8 + 256;
        cudaError_t e = cudaMalloc(&p, want);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};

struct PinBuf {
    void* p = nullptr; size_t cap = 0;
    cudaError_t reserve(size_t bytes)
    {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFreeHost(p);
        p = nullptr; cap = 0;
        size_t want = bytes + bytes / 8 + 256;
        cudaError_t e = cudaMallocHost(&p, want);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
};

} // namespace

struct lora_phy_cuda_ctx {
    int device = 0;
    int sm_count = 148;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    std::string err;
    uint64_t launches = 0;
    SfTables tab[13];
    std::map<std::tuple<int, int, int>, cf*> chirps;   // (sf, bw_hz, osr) -> conjugate downchirp, N*osr entries
    DevBuf state, phases, tmp;                          // PacketState[], modulator phases, compensate scratch
    DevBuf flag, pmax;                                  // lean route: one byte / one max-abs per packet (handed to the certified kernel)
    // lean route, adaptive: the share of packets the streaming pass handed over in the previous call comes back through a
    // pinned word (asynchronously); when more than a third were handed over (noisy traffic: measured break-even) the next
    // calls skip that pass and retry it every 8th call
    unsigned* h_flagged = nullptr;
    cudaEvent_t ev_flagged = nullptr;
    size_t flagged_of = 0;                              // packets of the call the pending word belongs to
    bool flagged_pending = false;
    int lean_skip = 0;
    cf* sync_tab[3][13] = {};                            // stream detector tables per (bandwidth scale, sf): up | down | fine
    DevBuf txrx[3];                                      // tx/rx pipelines: symbols, status, host-variant results
    DevBuf mc[8];                                        // Monte-Carlo driver: payload, tx symbols, IQ, rx symbols, status+sync, decoded, len, flags
    std::mutex mc_mu;
    lk::McChannel mc_channel = {};                      // set by awgn_roundtrip around its modulate call (fused channel)
    size_t mc_scratch_bytes = (size_t)4 << 30;           // IQ bytes per Monte-Carlo chunk (LORA_PHY_MC_SCRATCH_MB)
    DevBuf lfsr;                                         // dewhitening byte table of the coding chain (chain.cu)
    unsigned long long* cert_stats = nullptr;           // [8] counters of the certified demodulator (demod_cert.cu)
    DevBuf d_in, d_out, d_aux;                          // host-variant staging on the device
    PinBuf h_in, h_out;
    // host-variant pipeline (demodulate_batch_host): the batch is cut into chunks, chunk k+1 crosses PCIe on copy_stream
    // into the other half of d_in while chunk k is demodulated on `stream`
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t ev_h2d[2] = { nullptr, nullptr }, ev_done[2] = { nullptr, nullptr }, ev_start = nullptr;
    size_t host_chunk_bytes = (size_t)64 << 20;         // LORA_PHY_HOST_CHUNK_MB
    // mixed stream batches: one launch per (sf, bandwidth, osr) group, each a small grid of sequential walkers -- the groups
    // run side by side on these streams (one each up to N_SIDE groups, longest walks launched first) and join the
    // context's stream at the end
    static constexpr int N_SIDE = 32;
    cudaStream_t side[N_SIDE] = {};
    cudaEvent_t side_done[N_SIDE] = {}, side_start = nullptr;
    std::recursive_mutex mu;                            // held for a whole entry point (staging buffers, scratch, err)
};

namespace {

int fail(lora_phy_cuda_ctx* c, int code, const char* what, cudaError_t e = cudaSuccess)
{
    if (c) {
        c->err = what;
        if (e != cudaSuccess) { c->err += ": "; c->err += cudaGetErrorString(e); }
    }
    return code;
}

#define CK(call)                                                                  \
    do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail(ctx, -EIO, #call, e_); } while (0)

bool bw_ok(uint32_t hz) { return hz == 125000u || hz == 250000u || hz == 500000u; }

int check_cfg(lora_phy_cuda_ctx* ctx, const lora_phy_cuda_cfg* cfg)
{
    if (!ctx) return -EINVAL;
    if (!cfg) return fail(ctx, -EINVAL, "cfg is NULL");
    if (cfg->sf < 2 || cfg->sf > 12) return fail(ctx, -EINVAL, "sf out of range 2..12");
    if (cfg->osr < 1 || cfg->osr > 64) return fail(ctx, -EINVAL, "osr out of range 1..64");
    if (!bw_ok(cfg->bw_hz)) return fail(ctx, -EINVAL, "bw_hz must be 125000, 250000 or 500000");
    if (cfg->window > 1) return fail(ctx, -EINVAL, "window must be 0 (none) or 1 (Hann)");
    return 0;
}

int get_chirp(lora_phy_cuda_ctx* ctx, int sf, int bw_hz, int osr, cf** out)
{
    auto key = std::make_tuple(sf, bw_hz, osr);
    auto it = ctx->chirps.find(key);
    if (it != ctx->chirps.end()) { *out = it->second; return 0; }
    cf* p = nullptr;
    CK(cudaMalloc(&p, sizeof(cf) * ((size_t)1 << sf) * osr));
    CK(lk::launch_build_chirp(sf, osr, (float)bw_hz / 125000.0f, p, ctx->stream));
    ++ctx->launches;
    ctx->chirps[key] = p;
    *out = p;
    return 0;
}

struct Guard {
    int prev = -1;
    explicit Guard(int dev) { cudaGetDevice(&prev); if (prev != dev) cudaSetDevice(dev); else prev = -1; }
    ~Guard() { if (prev >= 0) cudaSetDevice(prev); }
};

// Every entry point that touches a context holds its (recursive) mutex from the first use of the staging buffers /
// scratch / error string to the last copy-out, so two threads sharing one context (the C++ drop-in wrapper routes
// every workspace through one) serialise instead of racing on d_in / d_out / h_out; the entry points nest
// (a *_host call runs the device call under the same lock).
struct CtxGuard {
    std::lock_guard<std::recursive_mutex> lock;
    Guard dev;
    explicit CtxGuard(lora_phy_cuda_ctx* c) : lock(c->mu), dev(c->device) {}
};

} // namespace

extern "C" {

int lora_phy_cuda_abi_version(void) { return LORA_PHY_CUDA_ABI_VERSION; }

int lora_phy_cuda_create(lora_phy_cuda_ctx** out, int device)
{
    if (!out) return -EINVAL;
    *out = nullptr;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0 || device < 0 || device >= n) return -ENODEV;
    lora_phy_cuda_ctx* ctx = new (std::nothrow) lora_phy_cuda_ctx;
    if (!ctx) return -ENOMEM;
    ctx->device = device;
    Guard g(device);
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) { delete ctx; return -ENODEV; }
    if (prop.major < 10) { delete ctx; return -ENODEV; }     // sm_100a code only: no other architecture, no fallback
    ctx->sm_count = prop.multiProcessorCount;
    if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) { delete ctx; return -EIO; }
    ctx->own_stream = true;
    for (int sf = 2; sf <= 12; ++sf) {
        const size_t N = (size_t)1 << sf;
        SfTables& t = ctx->tab[sf];
        const int cs = lk::compact_size_of(sf);
        if (cudaMalloc(&t.tw, N * sizeof(cf)) != cudaSuccess || cudaMalloc(&t.ct, (size_t)cs * sizeof(cf)) != cudaSuccess ||
            cudaMalloc(&t.hann, N * sizeof(float)) != cudaSuccess ||
            lk::launch_build_tables(sf, t.tw, t.ct, cs, t.hann, ctx->stream) != cudaSuccess) {
            lora_phy_cuda_destroy(ctx);
            return -EIO;
        }
        ctx->launches += 2;
    }
    if (cudaMalloc(&ctx->cert_stats, 8 * sizeof(unsigned long long)) != cudaSuccess ||
        cudaMemsetAsync(ctx->cert_stats, 0, 8 * sizeof(unsigned long long), ctx->stream) != cudaSuccess) {
        lora_phy_cuda_destroy(ctx);
        return -EIO;
    }
    if (cudaStreamSynchronize(ctx->stream) != cudaSuccess) { lora_phy_cuda_destroy(ctx); return -EIO; }
    if (cudaMallocHost(reinterpret_cast<void**>(&ctx->h_flagged), sizeof(unsigned)) != cudaSuccess ||
        cudaEventCreateWithFlags(&ctx->ev_flagged, cudaEventDisableTiming) != cudaSuccess) {
        lora_phy_cuda_destroy(ctx);
        return -ENOMEM;
    }
    *ctx->h_flagged = 0;
    *out = ctx;
    return 0;
}

void lora_phy_cuda_destroy(lora_phy_cuda_ctx* ctx)
{
    if (!ctx) return;
    CtxGuard g(ctx);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    for (int sf = 2; sf <= 12; ++sf) {
        if (ctx->tab[sf].tw) cudaFree(ctx->tab[sf].tw);
        if (ctx->tab[sf].ct) cudaFree(ctx->tab[sf].ct);
        if (ctx->tab[sf].hann) cudaFree(ctx->tab[sf].hann);
    }
    if (ctx->cert_stats) cudaFree(ctx->cert_stats);
    for (auto& kv : ctx->chirps) cudaFree(kv.second);
    for (int k = 0; k < 3; ++k) for (int sf = 0; sf <= 12; ++sf) if (ctx->sync_tab[k][sf]) cudaFree(ctx->sync_tab[k][sf]);
    ctx->lfsr.release();
    for (auto& b : ctx->mc) b.release();
    for (auto& b : ctx->txrx) b.release();
    ctx->state.release(); ctx->phases.release(); ctx->tmp.release(); ctx->flag.release(); ctx->pmax.release();
    ctx->d_in.release(); ctx->d_out.release(); ctx->d_aux.release();
    ctx->h_in.release(); ctx->h_out.release();
    for (int i = 0; i < 2; ++i) {
        if (ctx->ev_h2d[i]) cudaEventDestroy(ctx->ev_h2d[i]);
        if (ctx->ev_done[i]) cudaEventDestroy(ctx->ev_done[i]);
    }
    if (ctx->ev_start) cudaEventDestroy(ctx->ev_start);
    if (ctx->ev_flagged) cudaEventDestroy(ctx->ev_flagged);
    if (ctx->h_flagged) cudaFreeHost(ctx->h_flagged);
    for (int i = 0; i < lora_phy_cuda_ctx::N_SIDE; ++i) {
        if (ctx->side_done[i]) cudaEventDestroy(ctx->side_done[i]);
        if (ctx->side[i]) cudaStreamDestroy(ctx->side[i]);
    }
    if (ctx->side_start) cudaEventDestroy(ctx->side_start);
    if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
    if (ctx->own_stream && ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

const char* lora_phy_cuda_last_error(const lora_phy_cuda_ctx* ctx) { return ctx ? ctx->err.c_str() : "no context"; }
int lora_phy_cuda_device(const lora_phy_cuda_ctx* ctx) { return ctx ? ctx->device : -1; }
uint64_t lora_phy_cuda_launch_count(const lora_phy_cuda_ctx* ctx) { return ctx ? ctx->launches : 0; }

int lora_phy_cuda_set_stream(lora_phy_cuda_ctx* ctx, void* stream)
{
    if (!ctx) return -EINVAL;
    CtxGuard g(ctx);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    if (ctx->own_stream && ctx->stream) cudaStreamDestroy(ctx->stream);
    ctx->stream = static_cast<cudaStream_t>(stream);
    ctx->own_stream = false;
    return 0;
}

int lora_phy_cuda_cert_stats(lora_phy_cuda_ctx* ctx, uint64_t out[4])
{
    if (!ctx || !out) return -EINVAL;
    CtxGuard g(ctx);
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaMemcpy(out, ctx->cert_stats, 4 * sizeof(uint64_t), cudaMemcpyDeviceToHost));
    return 0;
}

int lora_phy_cuda_cert_stats_ex(lora_phy_cuda_ctx* ctx, uint64_t* out, size_t n)
{
    if (!ctx || !out || n > 8) return -EINVAL;
    CtxGuard g(ctx);
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaMemcpy(out, ctx->cert_stats, n * sizeof(uint64_t), cudaMemcpyDeviceToHost));
    return 0;
}

int lora_phy_cuda_synchronize(lora_phy_cuda_ctx* ctx)
{
    if (!ctx) return -EINVAL;
    CtxGuard g(ctx);
    CK(cudaStreamSynchronize(ctx->stream));
    return 0;
}

// ---------------------------------------------------------------------------------------------
static int demod_common(lora_phy_cuda_ctx* ctx, const lora_phy_cuda_cfg* cfg, int mode, const void* d_iq,
                        size_t packet_stride, const uint32_t* d_counts, size_t spp, size_t n_packets,
                        uint16_t* d_symbols, size_t symbol_stride, uint8_t* d_sync, float* d_cfo,
                        float* d_to, int32_t* d_status)
{
    int rc = check_cfg(ctx, cfg);
    if (rc) return rc;
    if (!d_iq && n_packets && spp) return fail(ctx, -EINVAL, "d_iq is NULL");
    if (reinterpret_cast<uintptr_t>(d_iq) & 7u) return fail(ctx, -EINVAL, "d_iq is not 8-byte aligned");
    if (spp > packet_stride) return fail(ctx, -EINVAL, "samples_per_packet exceeds packet_stride");
    if (spp > 0xffffffffull || n_packets > 0xffffffffull) return fail(ctx, -EINVAL, "batch too large");
    const size_t N = (size_t)1 << cfg->sf, step = N * cfg->osr;
    const size_t total_max = spp / step;
    uint32_t dsm = 0, est_uniform = 2;
    if (mode == lk::MODE_EST_ONLY) {
        est_uniform = (uint32_t)total_max;
    } else {
        if (!d_symbols && total_max) return fail(ctx, -EINVAL, "d_symbols is NULL");
        if (mode == lk::MODE_HIGHLEVEL && !d_counts) {           // phy.cpp:190-194
            if (spp % step != 0) return fail(ctx, -EINVAL, "sample count is not a multiple of (1<<sf)*osr");
            if (total_max < 2) return fail(ctx, -ERANGE, "fewer than two symbols");
        }
        dsm = total_max >= 2 ? (uint32_t)(total_max - 2) : (uint32_t)total_max;
        if (d_counts && total_max >= 2 && dsm == 0) dsm = 1;     // a ragged 1-symbol packet still emits one
        if (dsm > symbol_stride) return fail(ctx, -ERANGE, "symbol_stride smaller than the data symbols per packet");
    }
    if (n_packets == 0) return 0;
    CtxGuard g(ctx);
    CK(ctx->state.reserve(n_packets * sizeof(lk::PacketState)));

    lk::DemodArgs a;
    a.iq = static_cast<const cf*>(d_iq);
    a.packet_stride = packet_stride;
    a.sample_counts = d_counts;
    a.samples_per_packet = (uint32_t)spp;
    a.n_packets = (uint32_t)n_packets;
    a.osr = cfg->osr;
    a.mode = mode;
    a.no_scratch = (cfg->flags & LORA_PHY_CUDA_F_NO_SCRATCH) ? 1 : 0;
    a.est_uniform = est_uniform;
    a.data_symbols_max = dsm;
    a.dsm_magic = dsm ? (~0ull / dsm) + 1ull : 0ull;
    a.window = cfg->window ? ctx->tab[cfg->sf].hann : nullptr;
    a.down = nullptr;
    if (mode == lk::MODE_HIGHLEVEL) {
        cf* d = nullptr;
        rc = get_chirp(ctx, (int)cfg->sf, (int)cfg->bw_hz, 1, &d);   // phy.cpp:205-208: osr = 1 table
        if (rc) return rc;
        a.down = d;
    }
    a.tw = ctx->tab[cfg->sf].tw;
    a.ct = ctx->tab[cfg->sf].ct;
    a.power_scale = lk::power_scale_of((int)cfg->sf);
    a.state = static_cast<lk::PacketState*>(ctx->state.p);
    a.symbols = d_symbols;
    a.symbol_stride = symbol_stride;
    a.sync = d_sync;
    a.cfo = d_cfo;
    a.time_offset = d_to;
    a.status = d_status;
    a.force_exact = 0;
    a.no_prune = 0;
    a.unit_scale = 0; a.skip_sync = 0; a.flagged = nullptr; a.only_flagged = nullptr; a.pmax = nullptr; a.n_flagged = nullptr;
    a.cert_stats = ctx->cert_stats;
    unsigned launches = 0;
    // SF10-12, pre-dechirped IQ (the cfg-3 shape): two launches.  (1) demod_lean.cu: the exact two-symbol estimator by
    // teams of warps, then the streaming pruned transform with one warp per packet; it flags every packet it cannot
    // settle; (2) the certified kernel over the flagged packets.
    static const bool lean_off = getenv("LORA_PHY_NO_LEAN") != nullptr;
    bool given_up = false;                                  // the lean route is skipped because it handed most packets over
    if (!lean_off && !(cfg->flags & (LORA_PHY_CUDA_F_ONLY_PREPASS | LORA_PHY_CUDA_F_ONLY_SYMBOLS | LORA_PHY_CUDA_F_TWO_KERNEL | LORA_PHY_CUDA_F_EXACT)) &&
        lk::lean_supported((int)cfg->sf, a) && lk::cert_supported((int)cfg->sf, a) && [&]() {
            // verdict of the previous call, if it has arrived: mostly handed over -> the certified kernel alone for a while
            if (ctx->flagged_pending && cudaEventQuery(ctx->ev_flagged) == cudaSuccess) {
                ctx->flagged_pending = false;
                if (ctx->flagged_of >= 64 && (size_t)*ctx->h_flagged * 3 > ctx->flagged_of) ctx->lean_skip = 7;
            }
            if (ctx->lean_skip > 0) { --ctx->lean_skip; given_up = true; return false; }
            return true;
        }()) {
        CK(ctx->flag.reserve(((n_packets + 15) & ~(size_t)15) + 16));
        CK(ctx->pmax.reserve(n_packets * sizeof(float)));
        CK(cudaMemsetAsync(ctx->flag.p, 0, ((n_packets + 15) & ~(size_t)15) + 16, ctx->stream));     // flags, then (16-byte aligned) the flagged-packet counter
        lk::DemodArgs p = a;
        cudaError_t e;
        p = a;
        p.flagged = static_cast<uint8_t*>(ctx->flag.p);
        p.pmax = static_cast<float*>(ctx->pmax.p);
        p.n_flagged = reinterpret_cast<unsigned*>(static_cast<uint8_t*>(ctx->flag.p) + ((n_packets + 15) & ~(size_t)15));
        e = lk::launch_demod_lean((int)cfg->sf, p, ctx->sm_count, ctx->stream);
        ++launches;
        if (e != cudaSuccess) return fail(ctx, -EIO, "lean demodulator launch", e);
        p = a;
        p.only_flagged = static_cast<const uint8_t*>(ctx->flag.p);
        p.pmax = static_cast<float*>(ctx->pmax.p);
        p.n_flagged = reinterpret_cast<unsigned*>(static_cast<uint8_t*>(ctx->flag.p) + ((n_packets + 15) & ~(size_t)15));
        e = lk::launch_demod_cert((int)cfg->sf, p, ctx->sm_count, ctx->stream);
        ++launches;
        ctx->launches += launches;
        if (e != cudaSuccess) return fail(ctx, -EIO, "certified demodulator launch", e);
        if (!ctx->flagged_pending) {                       // (a still-pending word keeps its own call's count)
            CK(cudaMemcpyAsync(ctx->h_flagged, p.n_flagged, sizeof(unsigned), cudaMemcpyDeviceToHost, ctx->stream));
            CK(cudaEventRecord(ctx->ev_flagged, ctx->stream));
            ctx->flagged_of = n_packets;
            ctx->flagged_pending = true;
        }
        return 0;
    }
    // default: one launch, one pipeline per packet, margin-certified throughput transform with exact replay of the
    // windows that cannot be proven (demod_cert.cu)
    if (!(cfg->flags & (LORA_PHY_CUDA_F_ONLY_PREPASS | LORA_PHY_CUDA_F_ONLY_SYMBOLS | LORA_PHY_CUDA_F_TWO_KERNEL | LORA_PHY_CUDA_F_EXACT)) &&
        lk::cert_supported((int)cfg->sf, a)) {
        // A batch the lean route gave up on is noisy.  For single-warp windows (SF7-9) the copy of the kernel without the
        // pruned transform then runs: their full transform with tier 1 is as fast as the pruned one on clean windows
        // (4.0 TB/s either way at SF7) and 18-27 % faster on noise once the pruned code is out of the walk loop
        // (8 GB batches, SF7: 2 240 -> 2 643 GB/s at -6 dB, 2 235 -> 2 837 at +3 dB).  SF10-12 keep it: a packet fails the
        // lean pass on ONE bad window of 66, while 77 % of SF12 windows at +3 dB still settle on one bin class (measured
        // 2 998 GB/s with, 2 465 without).  Same results either way -- a pruned window is only ever accepted by the tier-1
        // test the full transform applies first.
        static const bool keep_prune = getenv("LORA_PHY_KEEP_PRUNE") != nullptr;       // A/B
        a.no_prune = given_up && !keep_prune && cfg->sf <= 9;
        cudaError_t e = lk::launch_demod_cert((int)cfg->sf, a, ctx->sm_count, ctx->stream);
        ++ctx->launches;
        if (e != cudaSuccess) return fail(ctx, -EIO, "certified demodulator launch", e);
        return 0;
    }
    // F_EXACT (A/B): the certified kernel with every window sent through its operation-for-operation replay
    if (!(cfg->flags & (LORA_PHY_CUDA_F_ONLY_PREPASS | LORA_PHY_CUDA_F_ONLY_SYMBOLS | LORA_PHY_CUDA_F_TWO_KERNEL)) &&
        lk::cert_supported((int)cfg->sf, a)) {
        a.force_exact = 1;
        cudaError_t e = lk::launch_demod_cert((int)cfg->sf, a, ctx->sm_count, ctx->stream);
        ++ctx->launches;
        if (e != cudaSuccess) return fail(ctx, -EIO, "certified demodulator launch (exact replay)", e);
        return 0;
    }
    const bool do_pre = !(cfg->flags & LORA_PHY_CUDA_F_ONLY_SYMBOLS);
    const bool do_sym = mode != lk::MODE_EST_ONLY && !(cfg->flags & LORA_PHY_CUDA_F_ONLY_PREPASS);
    cudaError_t e = lk::launch_demod((int)cfg->sf, a, ctx->sm_count, ctx->stream, do_pre, do_sym, &launches);
    ctx->launches += launches;
    if (e != cudaSuccess) return fail(ctx, -EIO, "demodulator launch", e);
    return 0;
}

int lora_phy_cuda_demodulate_batch(lora_phy_cuda_ctx* ctx, const lora_phy_cuda_cfg* cfg, const void* d_iq,
                                   size_t packet_stride, const uint32_t* d_sample_counts,
                                   size_t samples_per_packet, size_t n_packets, uint16_t* d_symbols,
                                   size_t symbol_stride, uint8_t* d_sync, float* d_cfo, float* d_time_offset,
                                   int32_t* d_status)
{
    const int mode = (cfg && (cfg->flags & LORA_PHY_CUDA_F_DECHIRP)) ? lk::MODE_HIGHLEVEL : lk::MODE_LEGACY;
    return demod_common(ctx, cfg, mode, d_iq, packet_stride, d_sample_counts, samples_per_packet, n_packets,
                        d_symbols, symbol_stride, d_sync, d_cfo, d_time_offset, d_status);
}

int lora_phy_cuda_estimate_offsets_batch(lora_phy_cuda_ctx* ctx, const lora_phy_cuda_cfg* cfg, const void* d_iq,
                                         size_t packet_stride, const uint32_t* d_sample_counts,
                                         size_t samples_per_packet, size_t n_packets, float* d_cfo,
                                         float* d_time_offset)
{
    return demod_common(ctx, cfg, lk::MODE_EST_ONLY, d_iq, packet_stride, d_sample_counts, samples_per_packet,
                        n_packets, nullptr, 0, nullptr, d_cfo, d_time_offset, nullptr);
}

int lora_phy_cuda_demodulate_batch_host(lora_phy_cuda_ctx* ctx, const lora_phy_cuda_cfg* cfg, const void* h_iq,
                                        size_t packet_stride, const uint32_t* h_counts, size_t spp,
                                        size_t n_packets, uint16_t* h_symbols, size_t symbol_stride,
                                        uint8_t* h_sync, float* h_cfo, float* h_to, int32_t* h_status)
{
    int rc = check_cfg(ctx, cfg);
    if (rc) return rc;
    if ((!h_iq && n_packets && spp) || (!h_symbols && n_packets && symbol_stride)) return fail(ctx, -EINVAL, "NULL buffer");
    if (spp > packet_stride) return fail(ctx, -EINVAL, "samples_per_packet exceeds packet_stride");
    if (n_packets == 0) return demod_common(ctx, cfg, (cfg->flags & LORA_PHY_CUDA_F_DECHIRP) ? lk::MODE_HIGHLEVEL : lk::MODE_LEGACY,
                                            h_iq, packet_stride, nullptr, spp, 0, h_symbols, symbol_stride, nullptr, nullptr, nullptr, nullptr);
    CtxGuard g(ctx);
    if (!ctx->copy_stream) {
        CK(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
        for (int i = 0; i < 2; ++i) {
            CK(cudaEventCreateWithFlags(&ctx->ev_h2d[i], cudaEventDisableTiming));
            CK(cudaEventCreateWithFlags(&ctx->ev_done[i], cudaEventDisableTiming));
        }
        CK(cudaEventCreateWithFlags(&ctx->ev_start, cudaEventDisableTiming));
        if (const char* mb = getenv("LORA_PHY_HOST_CHUNK_MB")) {
            const long v = strtol(mb, nullptr, 10);
            if (v > 0 && v <= 4096) ctx->host_chunk_bytes = (size_t)v << 20;
        }
    }
    const size_t pkt_bytes = packet_stride * sizeof(cf);
    const size_t sym_bytes = n_packets * symbol_stride * sizeof(uint16_t);
    // device scratch layout (d_out): symbols | cfo | time_offset | status | sync
    const size_t off_cfo = (sym_bytes + 15) & ~(size_t)15;
    const size_t off_to = off_cfo + n_packets * 4, off_st = off_to + n_packets * 4, off_sync = off_st + n_packets * 4;
    const size_t out_bytes = off_sync + n_packets;
    size_t chunk = pkt_bytes ? ctx->host_chunk_bytes / pkt_bytes : n_packets;
    if (chunk < 1) chunk = 1;
    if (chunk > n_packets) chunk = n_packets;
    const size_t half = (chunk * pkt_bytes + 255) & ~(size_t)255;       // one chunk of IQ; d_in holds two
    CK(ctx->d_in.reserve(2 * half + 16));
    CK(ctx->d_out.reserve(out_bytes));
    CK(ctx->h_out.reserve(out_bytes));
    uint32_t* d_counts = nullptr;
    if (h_counts) {
        for (size_t p = 0; p < n_packets; ++p)
            if (h_counts[p] > spp) return fail(ctx, -EINVAL, "a sample count exceeds samples_per_packet");
        CK(ctx->d_aux.reserve(n_packets * 4));
        d_counts = static_cast<uint32_t*>(ctx->d_aux.p);
        CK(cudaMemcpyAsync(d_counts, h_counts, n_packets * 4, cudaMemcpyHostToDevice, ctx->stream));
    }
    CK(cudaMemsetAsync(ctx->d_out.p, 0, out_bytes, ctx->stream));
    CK(cudaEventRecord(ctx->ev_start, ctx->stream));                     // the copy stream starts after earlier work of the context
    CK(cudaStreamWaitEvent(ctx->copy_stream, ctx->ev_start, 0));
    unsigned char* o = static_cast<unsigned char*>(ctx->d_out.p);
    const unsigned char* src = static_cast<const unsigned char*>(h_iq);
    size_t k = 0;
    for (size_t lo = 0; lo < n_packets; lo += chunk, ++k) {
        const size_t n = std::min(chunk, n_packets - lo);
        const int b = (int)(k & 1);
        unsigned char* d_chunk = static_cast<unsigned char*>(ctx->d_in.p) + (size_t)b * half;
        if (k >= 2) CK(cudaStreamWaitEvent(ctx->copy_stream, ctx->ev_done[b], 0));       // chunk k-2 has left this half
        CK(cudaMemcpyAsync(d_chunk, src + lo * pkt_bytes, n * pkt_bytes, cudaMemcpyHostToDevice, ctx->copy_stream));
        CK(cudaEventRecord(ctx->ev_h2d[b], ctx->copy_stream));
        CK(cudaStreamWaitEvent(ctx->stream, ctx->ev_h2d[b], 0));
        rc = lora_phy_cuda_demodulate_batch(ctx, cfg, d_chunk, packet_stride, d_counts ? d_counts + lo : nullptr, spp, n,
                                            reinterpret_cast<uint16_t*>(o) + lo * symbol_stride, symbol_stride, o + off_sync + lo,
                                            reinterpret_cast<float*>(o + off_cfo) + lo, reinterpret_cast<float*>(o + off_to) + lo,
                                            reinterpret_cast<int32_t*>(o + off_st) + lo);
        if (rc) { cudaStreamSynchronize(ctx->copy_stream); cudaStreamSynchronize(ctx->stream); return rc; }
        CK(cudaEventRecord(ctx->ev_done[b], ctx->stream));
    }
    CK(cudaMemcpyAsync(ctx->h_out.p, ctx->d_out.p, out_bytes, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    const unsigned char* h = static_cast<const unsigned char*>(ctx->h_out.p);
    const int32_t* st = reinterpret_cast<const int32_t*>(h + off_st);
    // only the symbols the reference would have written: rows of equal length leave in one strided copy
    const size_t step = ((size_t)1 << cfg->sf) * cfg->osr;
    const size_t total = step ? spp / step : 0;
    const size_t uniform = total >= 2 ? total - 2 : total;
    bool all_uniform = !h_counts;
    for (size_t p = 0; all_uniform && p < n_packets; ++p) all_uniform = st[p] == (int32_t)uniform;
    if (all_uniform) {
        if (uniform == symbol_stride) memcpy(h_symbols, h, n_packets * symbol_stride * 2);
        else for (size_t p = 0; p < n_packets; ++p) memcpy(h_symbols + p * symbol_stride, h + p * symbol_stride * 2, uniform * 2);
    } else {
        for (size_t p = 0; p < n_packets; ++p) {
            const int32_t cnt = st[p];
            if (cnt > 0) memcpy(h_symbols + p * symbol_stride, h + p * symbol_stride * 2, (size_t)cnt * 2);
        }
    }
    if (h_cfo) memcpy(h_cfo, h + off_cfo, n_packets * 4);
    if (h_to) memcpy(h_to, h + off_to, n_packets * 4);
    if (h_status) memcpy(h_status, h + off_st, n_packets * 4);
    if (h_sync) memcpy(h_sync, h + off_sync, n_packets);
    return 0;
}

int lora_phy_cuda_compensate_offsets_batch(lora_phy_cuda_ctx* ctx, const lora_phy_cuda_cfg* cfg, void* d_iq,
                                           size_t packet_stride, size_t spp, size_t n_packets,
                                           const float* d_cfo, const float* d_to)
{
    int rc = check_cfg(ctx, cfg);
    if (rc) return rc;
    if (!d_iq || !d_cfo || !d_to) return fail(ctx, -EINVAL, "NULL buffer");
    if (spp > packet_stride) return fail(ctx, -EINVAL, "samples_per_packet exceeds packet_stride");
    if (!n_packets || !spp) return 0;
    CtxGuard g(ctx);
    const size_t bytes = n_packets * packet_stride * sizeof(cf);
    CK(ctx->tmp.reserve(bytes));
    CK(lk::launch_compensate(static_cast<const cf*>(d_iq), static_cast<cf*>(ctx->tmp.p), packet_stride, (uint32_t)spp,
                             (uint32_t)n_packets, (int)cfg->sf, cfg->osr, d_cfo, d_to, ctx->stream));
    ++ctx->launches;
    CK(cudaMemcpy2DAsync(d_iq, packet_stride * sizeof(cf), ctx->tmp.p, packet_stride * sizeof(cf), spp * sizeof(cf),
                         n_packets, cudaMemcpyDeviceToDevice, ctx->stream));
    return 0;
}

int lora_phy_cuda_estimate_offsets_batch_host(lora_phy_cuda_ctx* ctx, const lora_phy_cuda_cfg* cfg, const void* h_iq,
                                              size_t packet_stride, size_t spp, size_t n_packets, float* h_cfo,
                                              float* h_to)
{
    int rc = check_cfg(ctx, cfg);
    if (rc) return rc;
    if (!h_iq || !h_cfo || !h_to) return fail(ctx, -EINVAL, "NULL buffer");
    if (spp > packet_stride) return fail(ctx, -EINVAL, "samples_per_packet exceeds packet_stride");
    if (!n_packets) return 0;
    CtxGuard g(ctx);
    const size_t iq_bytes = n_packets * packet_stride * sizeof(cf);
    CK(ctx->d_in.reserve(iq_bytes + 16));
    CK(ctx->d_out.reserve(n_packets * 8));
    CK(cudaMemcpyAsync(ctx->d_in.p, h_iq, iq_bytes, cudaMemcpyHostToDevice, ctx->stream));
    float* d_cfo = static_cast<float*>(ctx->d_out.p);
    float* d_to = d_cfo + n_packets;
    // packets without a whole symbol leave the outputs untouched (phy.cpp:84-90): seed them with the caller's values
    CK(cudaMemcpyAsync(d_cfo, h_cfo, n_packets * 4, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(d_to, h_to, n_packets * 4, cudaMemcpyHostToDevice, ctx->stream));
    rc = lora_phy_cuda_estimate_offsets_batch(ctx, cfg, ctx->d_in.p, packet_stride, nullptr, spp, n_packets, d_cfo, d_to);
    if (rc) return rc;
    CK(cudaMemcpyAsync(h_cfo, d_cfo, n_packets * 4, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaMemcpyAsync(h_to, d_to, n_packets * 4, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return 0;
}

int lora_phy_cuda_compensate_offsets_batch_host(lora_phy_cuda_ctx* ctx, const lora_phy_cuda_cfg* cfg, void* h_iq,
                                                size_t packet_stride, size_t spp, size_t n_packets,
                                                const float* h_cfo, const float* h_to)
{
    int rc = check_cfg(ctx, cfg);
    if (rc) return rc;
    if (!h_iq || !h_cfo || !h_to) return fail(ctx, -EINVAL, "NULL buffer");
    if (spp > packet_stride) return fail(ctx, -EINVAL, "samples_per_packet exceeds packet_stride");
    if (!n_packets || !spp) return 0;
    CtxGuard g(ctx);
    const size_t iq_bytes = n_packets * packet_stride * sizeof(cf);
    CK(ctx->d_in.reserve(iq_bytes + 16));
    CK(ctx->d_aux.reserve(n_packets * 8));
    float* d_cfo = static_cast<float*>(ctx->d_aux.p);
    float* d_to = d_cfo + n_packets;
    CK(cudaMemcpyAsync(ctx->d_in.p, h_iq, iq_bytes, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(d_cfo, h_cfo, n_packets * 4, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(d_to, h_to, n_packets * 4, cudaMemcpyHostToDevice, ctx->stream));
    rc = lora_phy_cuda_compensate_offsets_batch(ctx, cfg, ctx->d_in.p, packet_stride, spp, n_packets, d_cfo, d_to);
    if (rc) return rc;
    CK(cudaMemcpyAsync(h_iq, ctx->d_in.p, iq_bytes, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return 0;
}

int lora_phy_cuda_detect_batch(lora_phy_cuda_ctx* ctx, uint32_t sf, const void* d_windows, size_t n_windows,
                               void* d_spectrum, uint32_t* d_index, float* d_power, float* d_power_avg,
                               float* d_findex)
{
    if (!ctx) return -EINVAL;
    if (sf < 2 || sf > 12) return fail(ctx, -EINVAL, "sf out of range 2..12");
    if (!d_windows && n_windows) return fail(ctx, -EINVAL, "d_windows is NULL");
    if (!n_windows) return 0;
    CtxGuard g(ctx);
    lk::DetectArgs a;
    a.windows = static_cast<const cf*>(d_windows);
    a.n_windows = (uint32_t)n_windows;
    a.tw = ctx->tab[sf].tw; a.ct = ctx->tab[sf].ct; a.power_scale = lk::power_scale_of((int)sf);
    a.spectrum = static_cast<cf*>(d_spectrum); a.index = d_index; a.power = d_power;
    a.power_avg = d_power_avg; a.findex = d_findex;
    CK(lk::launch_detect((int)sf, a, ctx->sm_count, ctx->stream));
    ++ctx->launches;
    return 0;
}

// ---------------------------------------------------------------------------------------------
long lora_phy_cuda_modulate_batch(lora_phy_cuda_ctx* ctx, const lora_phy_cuda_cfg* cfg, const uint16_t* d_symbols,
                                  size_t symbol_stride, size_t spp_sym, size_t n_packets, void* d_iq,
                                  size_t packet_stride)
{
    int rc = check_cfg(ctx, cfg);
    if (rc) return rc;
    if ((!d_symbols && spp_sym && n_packets) || (!d_iq && n_packets)) return fail(ctx, -EINVAL, "NULL buffer");
    if (spp_sym > symbol_stride && n_packets > 1) return fail(ctx, -EINVAL, "symbols_per_packet exceeds symbol_stride");
    const size_t produced = (spp_sym + 2) * ((size_t)1 << cfg->sf) * cfg->osr;
    if (produced > packet_stride) return fail(ctx, -ERANGE, "packet_stride smaller than (symbols+2)*N*osr");
    if (!n_packets) return (long)produced;
    CtxGuard g(ctx);
    CK(ctx->phases.reserve(n_packets * (spp_sym + 2) * sizeof(float)));
    lk::ModArgs a;
    a.symbols = d_symbols; a.symbol_stride = symbol_stride; a.symbols_per_packet = (uint32_t)spp_sym;
    a.n_packets = (uint32_t)n_packets; a.iq = static_cast<cf*>(d_iq); a.packet_stride = packet_stride;
    a.sf = (int)cfg->sf; a.osr = cfg->osr; a.bw_scale = (float)cfg->bw_hz / 125000.0f;
    a.amplitude = cfg->amplitude; a.sync = cfg->sync_word & 0xff;
    a.phase_scratch = static_cast<float*>(ctx->phases.p);
    a.frame = 0; a.padding = 0;
    a.channel = ctx->mc_channel;                    // off unless the Monte-Carlo driver is the caller
    unsigned launches = 0;
    cudaError_t e = lk::launch_modulate(a, ctx->sm_count, ctx->stream, &launches);
    ctx->launches += launches;
    if (e != cudaSuccess) return fail(ctx, -EIO, "modulator launch", e);
    return (long)produced;
}

// Allocation-free variant (no counterpart in the reference; include/lora_phy_cuda.h): the phases live in modcodes.cu's
// static device array, nothing here or below allocates.
long lora_phy_cuda_modulate_batch_static(lora_phy_cuda_ctx* ctx, const lora_phy_cuda_cfg* cfg, const uint16_t* d_symbols,
                                         size_t symbol_stride, size_t spp_sym, size_t n_packets, void* d_iq,
                                         size_t packet_stride)
{
    int rc = check_cfg(ctx, cfg);
    if (rc) return rc;
    if ((!d_symbols && spp_sym && n_packets) || (!d_iq && n_packets)) return fail(ctx, -EINVAL, "NULL buffer");
    if (spp_sym > symbol_stride && n_packets > 1) return fail(ctx, -EINVAL, "symbols_per_packet exceeds symbol_stride");
    const size_t produced = (spp_sym + 2) * ((size_t)1 << cfg->sf) * cfg->osr;
    if (produced > packet_stride) return fail(ctx, -ERANGE, "packet_stride smaller than (symbols+2)*N*osr");
    if (!n_packets) return (long)produced;
    if (n_packets > 0xffffffffull || spp_sym > 0xfffffffdull ||
        n_packets > (size_t)LORA_PHY_CUDA_STATIC_MAX_CHIRPS / (spp_sym + 2))
        return fail(ctx, -E2BIG, "modulate_batch_static: n_packets * (symbols_per_packet + 2) exceeds "
                                 "LORA_PHY_CUDA_STATIC_MAX_CHIRPS (1048576); split the batch or use lora_phy_cuda_modulate_batch");
    CtxGuard g(ctx);
    float* phases = nullptr;
    cudaError_t e = lk::mod_static_scratch(&phases);          // cudaGetSymbolAddress, once per device, this device current
    if (e != cudaSuccess) return fail(ctx, -EIO, "cudaGetSymbolAddress(static modulator scratch)", e);
    lk::ModArgs a;
    a.symbols = d_symbols; a.symbol_stride = symbol_stride; a.symbols_per_packet = (uint32_t)spp_sym;
    a.n_packets = (uint32_t)n_packets; a.iq = static_cast<cf*>(d_iq); a.packet_stride = packet_stride;
    a.sf = (int)cfg->sf; a.osr = cfg->osr; a.bw_scale = (float)cfg->bw_hz / 125000.0f;
    a.amplitude = cfg->amplitude; a.sync = cfg->sync_word & 0xff;
    a.phase_scratch = phases;
    a.frame = 0; a.padding = 0;
    a.channel = lk::McChannel{};
    unsigned launches = 0;
    e = lk::launch_modulate(a, ctx->sm_count, ctx->stream, &launches);
    ctx->launches += launches;
    if (e != cudaSuccess) return fail(ctx, -EIO, "modulator launch (static scratch)", e);
    return (long)produced;
}

long lora_phy_cuda_frame_samples(uint32_t sf, uint32_t osr, size_t n_symbols, uint32_t padding_symbols)
{
    if (sf < 5 || sf > 12 || osr == 0 || osr > 8) return -EINVAL;
    const size_t NN = ((size_t)1 << sf) * osr;
    return (long)((14 + n_symbols + padding_symbols) * NN + NN / 4);
}

long lora_phy_cuda_modulate_frames_batch(lora_phy_cuda_ctx* ctx, const lora_phy_cuda_cfg* cfg, uint32_t padding_symbols,
                                         const uint16_t* d_symbols, size_t symbol_stride, size_t spf,
                                         size_t n_frames, void* d_iq, size_t frame_stride)
{
    int rc = check_cfg(ctx, cfg);
    if (rc) return rc;
    if ((!d_symbols && spf && n_frames) || (!d_iq && n_frames)) return fail(ctx, -EINVAL, "NULL buffer");
    if (spf > symbol_stride && n_frames > 1) return fail(ctx, -EINVAL, "symbols_per_frame exceeds symbol_stride");
    if (cfg->osr & (cfg->osr - 1)) return fail(ctx, -EINVAL, "frame format: oversampling must be a power of two");
    const long produced = lora_phy_cuda_frame_samples(cfg->sf, cfg->osr, spf, padding_symbols);
    if (produced < 0) return fail(ctx, (int)produced, "bad frame geometry");
    if ((size_t)produced > frame_stride) return fail(ctx, -ERANGE, "frame_stride smaller than the frame");
    if (!n_frames) return produced;
    CtxGuard g(ctx);
    CK(ctx->phases.reserve(n_frames * (spf + 15) * sizeof(float)));
    lk::ModArgs a;
    a.channel = lk::McChannel{};
    a.symbols = d_symbols; a.symbol_stride = symbol_stride; a.symbols_per_packet = (uint32_t)spf;
    a.n_packets = (uint32_t)n_frames; a.iq = static_cast<cf*>(d_iq); a.packet_stride = frame_stride;
    a.sf = (int)cfg->sf; a.osr = cfg->osr; a.bw_scale = (float)(cfg->bw_hz / 125000u);   // 1: upstream's frame; 2 / 4: bw_scale semantics
    a.amplitude = cfg->amplitude; a.sync = cfg->sync_word & 0xff;
    a.phase_scratch = static_cast<float*>(ctx->phases.p);
    a.frame = 1; a.padding = padding_symbols;
    unsigned launches = 0;
    cudaError_t e = lk::launch_modulate(a, ctx->sm_count, ctx->stream, &launches);
    ctx->launches += launches;
    if (e != cudaSuccess) return fail(ctx, -EIO, "frame modulator launch", e);
    return produced;
}

long lora_phy_cuda_modulate_batch_host(lora_phy_cuda_ctx* ctx, const lora_phy_cuda_cfg* cfg, const uint16_t* h_symbols,
                                       size_t symbol_stride, size_t spp_sym, size_t n_packets, void* h_iq,
                                       size_t packet_stride)
{
    int rc = check_cfg(ctx, cfg);
    if (rc) return rc;
    if ((!h_symbols && spp_sym && n_packets) || (!h_iq && n_packets)) return fail(ctx, -EINVAL, "NULL buffer");
    const size_t produced = (spp_sym + 2) * ((size_t)1 << cfg->sf) * cfg->osr;
    if (produced > packet_stride) return fail(ctx, -ERANGE, "packet_stride smaller than (symbols+2)*N*osr");
    if (!n_packets) return (long)produced;
    CtxGuard g(ctx);
    const size_t sym_bytes = (n_packets * symbol_stride + 1) * 2, iq_bytes = n_packets * packet_stride * sizeof(cf);
    CK(ctx->d_in.reserve(sym_bytes));
    CK(ctx->d_out.reserve(iq_bytes));
    if (spp_sym) CK(cudaMemcpyAsync(ctx->d_in.p, h_symbols, ((n_packets - 1) * symbol_stride + spp_sym) * 2, cudaMemcpyHostToDevice, ctx->stream));
    long r = lora_phy_cuda_modulate_batch(ctx, cfg, static_cast<const uint16_t*>(ctx->d_in.p), symbol_stride, spp_sym,
                                          n_packets, ctx->d_out.p, packet_stride);
    if (r < 0) return r;
    CK(cudaMemcpy2DAsync(h_iq, packet_stride * sizeof(cf), ctx->d_out.p, packet_stride * sizeof(cf), produced * sizeof(cf),
                         n_packets, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return r;
}

int lora_phy_cuda_dechirp_batch(lora_phy_cuda_ctx* ctx, const lora_phy_cuda_cfg* cfg, void* d_iq, size_t packet_stride,
                                size_t spp, size_t n_packets)
{
    int rc = check_cfg(ctx, cfg);
    if (rc) return rc;
    if (!d_iq && n_packets && spp) return fail(ctx, -EINVAL, "d_iq is NULL");
    if (reinterpret_cast<uintptr_t>(d_iq) & 7u) return fail(ctx, -EINVAL, "d_iq is not 8-byte aligned");
    if (spp > packet_stride) return fail(ctx, -EINVAL, "samples_per_packet exceeds packet_stride");
    if (!n_packets || !spp) return 0;
    CtxGuard g(ctx);
    cf* table = nullptr;
    rc = get_chirp(ctx, (int)cfg->sf, (int)cfg->bw_hz, (int)cfg->osr, &table);
    if (rc) return rc;
    CK(lk::launch_dechirp(static_cast<cf*>(d_iq), packet_stride, (uint32_t)spp, (uint32_t)n_packets, table,
                          (uint32_t)(((size_t)1 << cfg->sf) * cfg->osr), ctx->stream));
    ++ctx->launches;
    return 0;
}

// ---------------------------------------------------------------------------------------------
long lora_phy_cuda_encode_batch(lora_phy_cuda_ctx* ctx, const uint8_t* d_payload, size_t payload_stride,
                                size_t payload_len, size_t n_packets, uint16_t* d_symbols, size_t symbol_stride)
{
    if (!ctx) return -EINVAL;
    if ((!d_payload || !d_symbols) && n_packets && payload_len) return fail(ctx, -EINVAL, "NULL buffer");
    if (payload_len > payload_stride && n_packets > 1) return fail(ctx, -EINVAL, "payload_len exceeds payload_stride");
    if (2 * payload_len > symbol_stride) return fail(ctx, -ERANGE, "symbol_stride smaller than 2*payload_len");
    CtxGuard g(ctx);
    CK(lk::launch_encode(d_payload, payload_stride, (uint32_t)payload_len, (uint32_t)n_packets, d_symbols, symbol_stride, ctx->stream));
    if (n_packets && payload_len) ++ctx->launches;
    return (long)(2 * payload_len);
}

long lora_phy_cuda_decode_batch(lora_phy_cuda_ctx* ctx, const uint16_t* d_symbols, size_t symbol_stride,
                                const int32_t* d_counts, size_t spp, size_t n_packets, uint8_t* d_payload,
                                size_t payload_stride, uint8_t* d_crc_ok)
{
    if (!ctx) return -EINVAL;
    if ((!d_symbols || !d_payload) && n_packets && spp) return fail(ctx, -EINVAL, "NULL buffer");
    if (!d_counts && (spp & 1)) return fail(ctx, -EINVAL, "odd symbol count");      // LoRaDecoder.cpp:10
    if (spp / 2 > payload_stride) return fail(ctx, -ERANGE, "payload_stride smaller than symbols/2");
    CtxGuard g(ctx);
    CK(lk::launch_decode(d_symbols, symbol_stride, d_counts, (uint32_t)spp, (uint32_t)n_packets, d_payload, payload_stride, d_crc_ok, ctx->stream));
    if (n_packets) ctx->launches += (spp / 2 ? 1 : 0) + (d_crc_ok ? 1 : 0);
    return (long)(spp / 2);
}

long lora_phy_cuda_encode_batch_host(lora_phy_cuda_ctx* ctx, const uint8_t* h_payload, size_t payload_stride,
                                     size_t payload_len, size_t n_packets, uint16_t* h_symbols, size_t symbol_stride)
{
    if (!ctx) return -EINVAL;
    if ((!h_payload || !h_symbols) && n_packets && payload_len) return fail(ctx, -EINVAL, "NULL buffer");
    if (2 * payload_len > symbol_stride) return fail(ctx, -ERANGE, "symbol_stride smaller than 2*payload_len");
    if (!n_packets || !payload_len) return (long)(2 * payload_len);
    CtxGuard g(ctx);
    const size_t in_bytes = (n_packets - 1) * payload_stride + payload_len, out_bytes = n_packets * symbol_stride * 2;
    CK(ctx->d_in.reserve(in_bytes));
    CK(ctx->d_out.reserve(out_bytes));
    CK(cudaMemcpyAsync(ctx->d_in.p, h_payload, in_bytes, cudaMemcpyHostToDevice, ctx->stream));
    long r = lora_phy_cuda_encode_batch(ctx, static_cast<const uint8_t*>(ctx->d_in.p), payload_stride, payload_len, n_packets,
                                        static_cast<uint16_t*>(ctx->d_out.p), symbol_stride);
    if (r < 0) return r;
    CK(cudaMemcpy2DAsync(h_symbols, symbol_stride * 2, ctx->d_out.p, symbol_stride * 2, 2 * payload_len * 2, n_packets,
                         cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return r;
}

long lora_phy_cuda_decode_batch_host(lora_phy_cuda_ctx* ctx, const uint16_t* h_symbols, size_t symbol_stride,
                                     const int32_t* h_counts, size_t spp, size_t n_packets, uint8_t* h_payload,
                                     size_t payload_stride, uint8_t* h_crc_ok)
{
    if (!ctx) return -EINVAL;
    if ((!h_symbols || !h_payload) && n_packets && spp) return fail(ctx, -EINVAL, "NULL buffer");
    if (!h_counts && (spp & 1)) return fail(ctx, -EINVAL, "odd symbol count");
    if (spp / 2 > payload_stride) return fail(ctx, -ERANGE, "payload_stride smaller than symbols/2");
    if (!n_packets) return (long)(spp / 2);
    CtxGuard g(ctx);
    const size_t in_bytes = ((n_packets - 1) * symbol_stride + spp) * 2;
    const size_t pay_bytes = n_packets * payload_stride;
    CK(ctx->d_in.reserve(in_bytes + 16));
    CK(ctx->d_out.reserve(pay_bytes + n_packets + 16));
    CK(ctx->d_aux.reserve(n_packets * 4));
    if (in_bytes) CK(cudaMemcpyAsync(ctx->d_in.p, h_symbols, in_bytes, cudaMemcpyHostToDevice, ctx->stream));
    int32_t* d_counts = nullptr;
    if (h_counts) {
        d_counts = static_cast<int32_t*>(ctx->d_aux.p);
        CK(cudaMemcpyAsync(d_counts, h_counts, n_packets * 4, cudaMemcpyHostToDevice, ctx->stream));
    }
    uint8_t* d_pay = static_cast<uint8_t*>(ctx->d_out.p);
    uint8_t* d_crc = d_pay + pay_bytes;
    CK(cudaMemsetAsync(d_pay, 0, pay_bytes + n_packets, ctx->stream));
    long r = lora_phy_cuda_decode_batch(ctx, static_cast<const uint16_t*>(ctx->d_in.p), symbol_stride, d_counts, spp,
                                        n_packets, d_pay, payload_stride, d_crc);
    if (r < 0) return r;
    CK(ctx->h_out.reserve(pay_bytes + n_packets));
    CK(cudaMemcpyAsync(ctx->h_out.p, d_pay, pay_bytes + n_packets, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    const uint8_t* h = static_cast<const uint8_t*>(ctx->h_out.p);
    for (size_t p = 0; p < n_packets; ++p) {
        int32_t cnt = h_counts ? h_counts[p] : (int32_t)spp;
        size_t nb = (cnt < 0 || (cnt & 1)) ? 0 : (size_t)cnt / 2;
        if (nb > spp / 2) nb = spp / 2;
        memcpy(h_payload + p * payload_stride, h + p * payload_stride, nb);
    }
    if (h_crc_ok) memcpy(h_crc_ok, h + pay_bytes, n_packets);
    return r;
}

// ---------------------------------------------------------------------------------------------
// Tables of the upstream receiver (LoRaDemod.cpp:96-117): std::polar(1.0, double phase) narrowed to float, built
// with the host's libm exactly as the reference's constructor does, once per (context, sf, bandwidth scale).
// bw_k = 2 / 4: the reference's bandwidth semantics (ChirpGenerator.hpp:28-30: bw_scale on start frequency and slope).
static int sync_tables(lora_phy_cuda_ctx* ctx, uint32_t sf, uint32_t bw_k, cf** out)
{
    const int slot = bw_k == 1 ? 0 : bw_k == 2 ? 1 : 2;
    if (ctx->sync_tab[slot][sf]) { *out = ctx->sync_tab[slot][sf]; return 0; }
    const size_t N = (size_t)1 << sf, fine_n = N * 128;
    std::vector<cf> host(2 * N + fine_n);
    float phase = (float)(-M_PI * (double)bw_k);
    double acc = 0.0;
    for (size_t i = 0; i < N; ++i) {
        acc += phase;
        const std::complex<double> e = std::polar(1.0, acc);
        host[i] = fx::mk((float)e.real(), (float)-e.imag());          // _upChirpTable = conj(entry)
        host[N + i] = fx::mk((float)e.real(), (float)e.imag());       // _downChirpTable
        phase = (float)((double)phase + (2 * M_PI * (double)bw_k) / (double)N);   // float += double: summed in double, rounded once
    }
    acc = 0.0;
    phase = (float)(2.0 * M_PI / (double)fine_n);
    for (size_t i = 0; i < fine_n; ++i) {
        acc += phase;
        const std::complex<double> e = std::polar(1.0, acc);
        host[2 * N + i] = fx::mk((float)e.real(), (float)e.imag());
    }
    cf* d = nullptr;
    CK(cudaMalloc(&d, host.size() * sizeof(cf)));
    CK(cudaMemcpyAsync(d, host.data(), host.size() * sizeof(cf), cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    ctx->sync_tab[slot][sf] = d;
    *out = d;
    return 0;
}

// one launch per distinct (sf, bandwidth scale, osr) among the streams, rows selected through an id list
static int detect_streams_common(lora_phy_cuda_ctx* ctx, const uint8_t* h_sf, uint32_t sf_all, const uint8_t* h_bw_k,
                                 const uint8_t* h_osr, uint32_t sync_word, float thresh_db, uint32_t mtu, const void* d_iq,
                                 size_t stream_stride, const uint32_t* d_counts, size_t n_samples, size_t n_streams,
                                 size_t max_frames, uint32_t* d_frame_start, int32_t* d_frame_err, int32_t* d_frame_nsym,
                                 int16_t* d_symbols, uint32_t* d_n_frames, uint32_t* d_end_pos, const float* d_cfo,
                                 const float* d_to)
{
    if (!ctx) return -EINVAL;
    if (!mtu || mtu > 32767) return fail(ctx, -EINVAL, "mtu out of range 1..32767");
    if (n_samples > stream_stride && n_streams > 1) return fail(ctx, -EINVAL, "n_samples exceeds stream_stride");
    if (n_samples > 0xf0000000ull || n_streams > 0xffffffffull || max_frames > 0xffffffffull) return fail(ctx, -EINVAL, "too large");
    if (n_streams && (!d_iq || !d_n_frames || (max_frames && (!d_frame_start || !d_frame_err || !d_frame_nsym || !d_symbols))))
        return fail(ctx, -EINVAL, "NULL buffer");
    if ((d_cfo == nullptr) != (d_to == nullptr)) return fail(ctx, -EINVAL, "cfo and time_offset go together");
    if (reinterpret_cast<uintptr_t>(d_iq) & 7u) return fail(ctx, -EINVAL, "d_iq is not 8-byte aligned");
    std::map<std::tuple<int, int, int>, std::vector<uint32_t>> groups;
    for (size_t s = 0; s < n_streams; ++s) {
        const int sf = h_sf ? h_sf[s] : (int)sf_all, k = h_bw_k ? h_bw_k[s] : 1, osr = h_osr ? h_osr[s] : 1;
        if (sf < 7 || sf > 12) return fail(ctx, -EINVAL, "stream sf out of range 7..12");
        if (k != 1 && k != 2 && k != 4) return fail(ctx, -EINVAL, "stream bandwidth scale must be 1, 2 or 4");
        if (osr < 1 || osr > 8) return fail(ctx, -EINVAL, "stream osr out of range 1..8");
        groups[std::make_tuple(sf, k, osr)].push_back((uint32_t)s);
    }
    if (!n_streams) return 0;
    CtxGuard g(ctx);
    uint32_t* d_ids = nullptr;
    const bool one_group = groups.size() == 1;
    if (!one_group) {
        CK(ctx->tmp.reserve(n_streams * sizeof(uint32_t)));
        d_ids = static_cast<uint32_t*>(ctx->tmp.p);
        size_t off = 0;
        for (auto& kv : groups) {                                  // rows of one group are contiguous in the id list
            CK(cudaMemcpyAsync(d_ids + off, kv.second.data(), kv.second.size() * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
            off += kv.second.size();
        }
        CK(cudaStreamSynchronize(ctx->stream));                    // the host vectors go out of scope below
    }
    constexpr int N_SIDE = lora_phy_cuda_ctx::N_SIDE;
    const int n_side = one_group ? 0 : (int)std::min<size_t>(groups.size(), (size_t)N_SIDE);
    if (!one_group) {
        if (!ctx->side_start) CK(cudaEventCreateWithFlags(&ctx->side_start, cudaEventDisableTiming));
        for (int i = 0; i < n_side; ++i) {
            if (ctx->side[i]) continue;
            CK(cudaStreamCreateWithFlags(&ctx->side[i], cudaStreamNonBlocking));
            CK(cudaEventCreateWithFlags(&ctx->side_done[i], cudaEventDisableTiming));
        }
    }
    for (auto& kv : groups) {                                       // tables first: building one synchronises the stream
        cf* tab = nullptr;
        int rc = sync_tables(ctx, (uint32_t)std::get<0>(kv.first), (uint32_t)std::get<1>(kv.first), &tab);
        if (rc) return rc;
    }
    if (!one_group) {
        CK(cudaEventRecord(ctx->side_start, ctx->stream));
        for (int i = 0; i < n_side; ++i) CK(cudaStreamWaitEvent(ctx->side[i], ctx->side_start, 0));
    }
    // A group's time is the walk of its longest row -- windows = samples / (N * osr) -- not its size: launch the longest
    // walks first, every group on a stream of its own while they last.
    struct Launch { const std::tuple<int, int, int>* key; const std::vector<uint32_t>* rows; size_t off; };
    std::vector<Launch> order;
    {
        size_t off = 0;
        for (auto& kv : groups) { order.push_back({ &kv.first, &kv.second, off }); off += kv.second.size(); }
        std::stable_sort(order.begin(), order.end(), [](const Launch& x, const Launch& y) {
            return ((size_t)1 << std::get<0>(*x.key)) * (size_t)std::get<2>(*x.key) < ((size_t)1 << std::get<0>(*y.key)) * (size_t)std::get<2>(*y.key);
        });
    }
    size_t gi = 0;
    for (const Launch& L : order) {
        const auto& key = *L.key;
        const size_t off = L.off, rows = L.rows->size();
        const uint32_t sf = (uint32_t)std::get<0>(key), k = (uint32_t)std::get<1>(key), osr = (uint32_t)std::get<2>(key);
        cf* tab = nullptr;
        int rc = sync_tables(ctx, sf, k, &tab);
        if (rc) return rc;
        const size_t N = (size_t)1 << sf;
        lk::StreamSyncArgs a;
        a.iq = static_cast<const cf*>(d_iq); a.stream_stride = stream_stride; a.counts = d_counts;
        a.n_samples = (uint32_t)n_samples; a.n_streams = (uint32_t)rows;
        a.up = tab; a.down = tab + N; a.fine = tab + 2 * N;
        a.tw = ctx->tab[sf].tw; a.ct = ctx->tab[sf].ct; a.power_scale = lk::power_scale_of((int)sf);
        a.sync = sync_word & 0xff; a.thresh = thresh_db; a.mtu = mtu; a.max_frames = (uint32_t)max_frames;
        a.osr = osr; a.bw_k = k;
        a.sync0 = (uint32_t)((((sync_word >> 4) & 0xf) * k) % (N / 8));     // the modulator puts the nibbles at nibble*8*bw_k bins
        a.sync1 = (uint32_t)(((sync_word & 0xf) * k) % (N / 8));
        a.comp_cfo = d_cfo; a.comp_to = d_to;
        { const char* e = getenv("LORA_PHY_STREAM_SERIAL"); a.force_serial = e && *e == '1'; }
        a.frame_start = d_frame_start; a.frame_err = d_frame_err; a.frame_nsym = d_frame_nsym; a.symbols = d_symbols;
        a.n_frames = d_n_frames; a.end_pos = d_end_pos; a.ids = one_group ? nullptr : d_ids + off;
        CK(lk::launch_stream_sync((int)sf, a, one_group ? ctx->stream : ctx->side[gi % (size_t)n_side]));
        ++ctx->launches;
        ++gi;
    }
    if (!one_group) {
        for (int i = 0; i < n_side; ++i) {
            CK(cudaEventRecord(ctx->side_done[i], ctx->side[i]));
            CK(cudaStreamWaitEvent(ctx->stream, ctx->side_done[i], 0));
        }
    }
    return 0;
}

int lora_phy_cuda_detect_stream(lora_phy_cuda_ctx* ctx, uint32_t sf, uint32_t sync_word, float thresh_db, uint32_t mtu,
                                const void* d_iq, size_t stream_stride, const uint32_t* d_counts, size_t n_samples,
                                size_t n_streams, size_t max_frames, uint32_t* d_frame_start, int32_t* d_frame_err,
                                int32_t* d_frame_nsym, int16_t* d_symbols, uint32_t* d_n_frames, uint32_t* d_end_pos)
{
    if (!ctx) return -EINVAL;
    if (sf < 7 || sf > 12) return fail(ctx, -EINVAL, "sf out of range 7..12");
    return detect_streams_common(ctx, nullptr, sf, nullptr, nullptr, sync_word, thresh_db, mtu, d_iq, stream_stride, d_counts,
                                 n_samples, n_streams, max_frames, d_frame_start, d_frame_err, d_frame_nsym, d_symbols,
                                 d_n_frames, d_end_pos, nullptr, nullptr);
}

int lora_phy_cuda_detect_streams_mixed(lora_phy_cuda_ctx* ctx, const uint8_t* h_sf, uint32_t sync_word, float thresh_db,
                                       uint32_t mtu, const void* d_iq, size_t stream_stride, const uint32_t* d_counts,
                                       size_t n_samples, size_t n_streams, size_t max_frames, uint32_t* d_frame_start,
                                       int32_t* d_frame_err, int32_t* d_frame_nsym, int16_t* d_symbols,
                                       uint32_t* d_n_frames, uint32_t* d_end_pos)
{
    if (!ctx) return -EINVAL;
    if (!h_sf && n_streams) return fail(ctx, -EINVAL, "h_sf is NULL");
    return detect_streams_common(ctx, h_sf, 0, nullptr, nullptr, sync_word, thresh_db, mtu, d_iq, stream_stride, d_counts,
                                 n_samples, n_streams, max_frames, d_frame_start, d_frame_err, d_frame_nsym, d_symbols,
                                 d_n_frames, d_end_pos, nullptr, nullptr);
}

int lora_phy_cuda_detect_streams_ex(lora_phy_cuda_ctx* ctx, const uint8_t* h_sf, const uint8_t* h_bw_scale, const uint8_t* h_osr,
                                    uint32_t sync_word, float thresh_db, uint32_t mtu, const void* d_iq, size_t stream_stride,
                                    const uint32_t* d_counts, size_t n_samples, size_t n_streams, size_t max_frames,
                                    const float* d_cfo, const float* d_time_offset, uint32_t* d_frame_start,
                                    int32_t* d_frame_err, int32_t* d_frame_nsym, int16_t* d_symbols, uint32_t* d_n_frames,
                                    uint32_t* d_end_pos)
{
    if (!ctx) return -EINVAL;
    if (!h_sf && n_streams) return fail(ctx, -EINVAL, "h_sf is NULL");
    return detect_streams_common(ctx, h_sf, 0, h_bw_scale, h_osr, sync_word, thresh_db, mtu, d_iq, stream_stride, d_counts,
                                 n_samples, n_streams, max_frames, d_frame_start, d_frame_err, d_frame_nsym, d_symbols,
                                 d_n_frames, d_end_pos, d_cfo, d_time_offset);
}

// ---------------------------------------------------------------------------------------------
static int chain_args(const lora_phy_cuda_chain_cfg* cc, lk::ChainArgs* a)
{
    if (!cc) return -EINVAL;
    const uint32_t ppm = cc->ppm ? cc->ppm : cc->sf;
    if (cc->sf < 7 || cc->sf > 12 || ppm > cc->sf || ppm < 5 || cc->cr > 4) return -EINVAL;
    a->sf = cc->sf; a->ppm = ppm; a->rdd = cc->cr;
    a->explicit_hdr = cc->explicit_header != 0; a->crc = cc->crc != 0; a->whitening = cc->whitening != 0;
    a->error_check = cc->error_check != 0; a->crc_check = cc->crc_check != 0; a->out_hdr = cc->out_header != 0;
    a->data_length = cc->data_length;
    return 0;
}

long lora_phy_cuda_chain_symbols(const lora_phy_cuda_chain_cfg* cc, size_t payload_len)
{
    lk::ChainArgs a;
    if (chain_args(cc, &a) || payload_len > 255) return -EINVAL;
    const size_t nbytes = payload_len + (a.crc ? 2 : 0);
    const size_t ncw = ((nbytes * 2 + (a.explicit_hdr ? 5 : 0) + a.ppm - 1) / a.ppm) * a.ppm;   // LoRaEncoder.cpp:171
    return (long)(8 + (ncw / a.ppm - 1) * (4 + a.rdd));                                        // :172
}

long lora_phy_cuda_encode_chain_batch(lora_phy_cuda_ctx* ctx, const lora_phy_cuda_chain_cfg* cc,
                                      const uint8_t* d_payload, size_t payload_stride, size_t payload_len,
                                      size_t n_packets, uint16_t* d_symbols, size_t symbol_stride)
{
    if (!ctx) return -EINVAL;
    lk::ChainArgs a;
    if (chain_args(cc, &a)) return fail(ctx, -EINVAL, "bad chain configuration (sf 7..12, 5 <= ppm <= sf, cr 0..4)");
    if (payload_len > 255) return fail(ctx, -EINVAL, "payload longer than 255 bytes");
    if ((!d_payload && payload_len && n_packets) || (!d_symbols && n_packets)) return fail(ctx, -EINVAL, "NULL buffer");
    if (payload_len > payload_stride && n_packets > 1) return fail(ctx, -EINVAL, "payload_len exceeds payload_stride");
    const long nsym = lora_phy_cuda_chain_symbols(cc, payload_len);
    if ((size_t)nsym > symbol_stride) return fail(ctx, -ERANGE, "symbol_stride smaller than the symbols per packet");
    if (n_packets > 0xffffffffull) return fail(ctx, -EINVAL, "batch too large");
    CtxGuard g(ctx);
    CK(lk::launch_chain_encode(a, d_payload, payload_stride, (uint32_t)payload_len, (uint32_t)n_packets, d_symbols,
                               symbol_stride, ctx->sm_count, ctx->stream));
    if (n_packets) ++ctx->launches;
    return nsym;
}

int lora_phy_cuda_decode_chain_batch(lora_phy_cuda_ctx* ctx, const lora_phy_cuda_chain_cfg* cc,
                                     const uint16_t* d_symbols, size_t symbol_stride, const int32_t* d_counts,
                                     size_t n_symbols, size_t n_packets, uint8_t* d_payload, size_t payload_stride,
                                     int32_t* d_len, uint8_t* d_flags)
{
    if (!ctx) return -EINVAL;
    lk::ChainArgs a;
    if (chain_args(cc, &a)) return fail(ctx, -EINVAL, "bad chain configuration (sf 7..12, 5 <= ppm <= sf, cr 0..4)");
    if ((!d_symbols || !d_payload || !d_len) && n_packets) return fail(ctx, -EINVAL, "NULL buffer");
    if (n_symbols > symbol_stride && n_packets > 1) return fail(ctx, -EINVAL, "n_symbols exceeds symbol_stride");
    const size_t rows = 4 + a.rdd, nsym = ((n_symbols + rows - 1) / rows) * rows;
    if (n_symbols > (size_t)lk::chain_max_symbols() || (nsym / rows) * a.ppm + 32 > 2048)
        return fail(ctx, -EINVAL, "too many symbols per packet for the coding chain (a 255-byte packet needs < 700)");
    if (n_packets > 0xffffffffull) return fail(ctx, -EINVAL, "batch too large");
    if (!n_packets) return 0;
    CtxGuard g(ctx);
    if (!ctx->lfsr.p) {
        const size_t n = (size_t)lk::chain_lfsr_table_size();
        std::string host(n, '\0');
        lk::chain_lfsr_table(reinterpret_cast<uint8_t*>(&host[0]));
        CK(ctx->lfsr.reserve(n));
        CK(cudaMemcpyAsync(ctx->lfsr.p, host.data(), n, cudaMemcpyHostToDevice, ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));
    }
    CK(lk::launch_chain_decode(a, d_symbols, symbol_stride, d_counts, (uint32_t)n_symbols, (uint32_t)n_packets,
                               static_cast<const uint8_t*>(ctx->lfsr.p), d_payload, payload_stride, d_len, d_flags,
                               ctx->sm_count, ctx->stream));
    ++ctx->launches;
    return 0;
}

// ---------------------------------------------------------------------------------------------
// README facade pipelines: LoRaPHY::modulate (payload -> IQ) and LoRaPHY::demodulate (IQ -> payload), batched.
long lora_phy_cuda_tx_batch(lora_phy_cuda_ctx* ctx, const lora_phy_cuda_cfg* cfg, const lora_phy_cuda_chain_cfg* cc,
                            const uint8_t* d_payload, size_t payload_stride, size_t payload_len, size_t n_packets,
                            void* d_iq, size_t packet_stride)
{
    int rc = check_cfg(ctx, cfg);
    if (rc) return rc;
    if (!cc || cc->sf != cfg->sf) return fail(ctx, -EINVAL, "chain configuration missing or for another sf");
    const long nsym = lora_phy_cuda_chain_symbols(cc, payload_len);
    if (nsym < 0) return fail(ctx, -EINVAL, "bad chain configuration");
    const size_t produced = ((size_t)nsym + 2) * ((size_t)1 << cfg->sf) * cfg->osr;
    if (produced > packet_stride) return fail(ctx, -ERANGE, "packet_stride smaller than (symbols+2)*N*osr");
    if (!n_packets) return (long)produced;
    {
        CtxGuard g(ctx);
        CK(ctx->txrx[0].reserve(n_packets * (size_t)nsym * 2));
    }
    uint16_t* sym = static_cast<uint16_t*>(ctx->txrx[0].p);
    long r = lora_phy_cuda_encode_chain_batch(ctx, cc, d_payload, payload_stride, payload_len, n_packets, sym, (size_t)nsym);
    if (r < 0) return r;
    return lora_phy_cuda_modulate_batch(ctx, cfg, sym, (size_t)nsym, (size_t)nsym, n_packets, d_iq, packet_stride);
}

int lora_phy_cuda_rx_batch(lora_phy_cuda_ctx* ctx, const lora_phy_cuda_cfg* cfg, const lora_phy_cuda_chain_cfg* cc,
                           void* d_iq, size_t packet_stride, size_t spp, size_t n_packets, uint8_t* d_payload,
                           size_t payload_stride, int32_t* d_len, uint8_t* d_flags, uint8_t* d_sync, float* d_cfo,
                           float* d_time_offset)
{
    int rc = check_cfg(ctx, cfg);
    if (rc) return rc;
    if (!cc || cc->sf != cfg->sf) return fail(ctx, -EINVAL, "chain configuration missing or for another sf");
    if ((!d_iq || !d_payload || !d_len) && n_packets) return fail(ctx, -EINVAL, "NULL buffer");
    const size_t step = ((size_t)1 << cfg->sf) * cfg->osr, total = spp / step;
    if (total < 2) return fail(ctx, -ERANGE, "fewer than two symbols per packet");
    if (!n_packets) return 0;
    const size_t nsym = total - 2;
    {
        CtxGuard g(ctx);
        CK(ctx->txrx[0].reserve(n_packets * (nsym + 1) * 2));
        CK(ctx->txrx[1].reserve(n_packets * 4));
    }
    uint16_t* sym = static_cast<uint16_t*>(ctx->txrx[0].p);
    int32_t* status = static_cast<int32_t*>(ctx->txrx[1].p);
    lora_phy_cuda_cfg c = *cfg;
    c.flags &= ~(uint32_t)LORA_PHY_CUDA_F_DECHIRP;                 // caller-side dechirp + lora_demodulate (e2e_chain_test.cpp:83-101)
    rc = lora_phy_cuda_dechirp_batch(ctx, &c, d_iq, packet_stride, spp, n_packets);
    if (rc < 0) return rc;
    rc = lora_phy_cuda_demodulate_batch(ctx, &c, d_iq, packet_stride, nullptr, spp, n_packets, sym, nsym ? nsym : 1, d_sync,
                                        d_cfo, d_time_offset, status);
    if (rc < 0) return rc;
    return lora_phy_cuda_decode_chain_batch(ctx, cc, sym, nsym ? nsym : 1, status, nsym, n_packets, d_payload, payload_stride,
                                            d_len, d_flags);
}

long lora_phy_cuda_tx_batch_host(lora_phy_cuda_ctx* ctx, const lora_phy_cuda_cfg* cfg, const lora_phy_cuda_chain_cfg* cc,
                                 const uint8_t* h_payload, size_t payload_stride, size_t payload_len, size_t n_packets,
                                 void* h_iq, size_t packet_stride)
{
    int rc = check_cfg(ctx, cfg);
    if (rc) return rc;
    if ((!h_payload && payload_len && n_packets) || (!h_iq && n_packets)) return fail(ctx, -EINVAL, "NULL buffer");
    if (!cc) return fail(ctx, -EINVAL, "chain configuration missing");
    const long nsym = lora_phy_cuda_chain_symbols(cc, payload_len);
    if (nsym < 0) return fail(ctx, -EINVAL, "bad chain configuration");
    const size_t produced = ((size_t)nsym + 2) * ((size_t)1 << cfg->sf) * cfg->osr;
    if (produced > packet_stride) return fail(ctx, -ERANGE, "packet_stride smaller than (symbols+2)*N*osr");
    if (!n_packets) return (long)produced;
    CtxGuard g(ctx);
    const size_t in_bytes = (n_packets - 1) * payload_stride + payload_len, iq_bytes = n_packets * packet_stride * sizeof(cf);
    CK(ctx->d_in.reserve(in_bytes + 16));
    CK(ctx->d_out.reserve(iq_bytes));
    if (in_bytes) CK(cudaMemcpyAsync(ctx->d_in.p, h_payload, in_bytes, cudaMemcpyHostToDevice, ctx->stream));
    long r = lora_phy_cuda_tx_batch(ctx, cfg, cc, static_cast<const uint8_t*>(ctx->d_in.p), payload_stride, payload_len,
                                    n_packets, ctx->d_out.p, packet_stride);
    if (r < 0) return r;
    CK(cudaMemcpy2DAsync(h_iq, packet_stride * sizeof(cf), ctx->d_out.p, packet_stride * sizeof(cf), produced * sizeof(cf),
                         n_packets, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return r;
}

int lora_phy_cuda_rx_batch_host(lora_phy_cuda_ctx* ctx, const lora_phy_cuda_cfg* cfg, const lora_phy_cuda_chain_cfg* cc,
                                const void* h_iq, size_t packet_stride, size_t spp, size_t n_packets, uint8_t* h_payload,
                                size_t payload_stride, int32_t* h_len, uint8_t* h_flags, uint8_t* h_sync, float* h_cfo,
                                float* h_time_offset)
{
    int rc = check_cfg(ctx, cfg);
    if (rc) return rc;
    if ((!h_iq || !h_payload || !h_len) && n_packets) return fail(ctx, -EINVAL, "NULL buffer");
    if (spp > packet_stride) return fail(ctx, -EINVAL, "samples_per_packet exceeds packet_stride");
    if (!n_packets) return 0;
    CtxGuard g(ctx);
    const size_t iq_bytes = ((n_packets - 1) * packet_stride + spp) * sizeof(cf), pay_bytes = n_packets * payload_stride;
    // results: payload | len (int32) | cfo | time_offset | flags | sync
    const size_t o_len = (pay_bytes + 15) & ~(size_t)15, o_cfo = o_len + n_packets * 4, o_to = o_cfo + n_packets * 4,
                 o_fl = o_to + n_packets * 4, o_sy = o_fl + n_packets, res_bytes = o_sy + n_packets;
    CK(ctx->d_in.reserve(iq_bytes));
    CK(ctx->txrx[2].reserve(res_bytes));
    CK(cudaMemcpyAsync(ctx->d_in.p, h_iq, iq_bytes, cudaMemcpyHostToDevice, ctx->stream));
    uint8_t* res = static_cast<uint8_t*>(ctx->txrx[2].p);
    CK(cudaMemsetAsync(res, 0, res_bytes, ctx->stream));
    rc = lora_phy_cuda_rx_batch(ctx, cfg, cc, ctx->d_in.p, packet_stride, spp, n_packets, res, payload_stride,
                                reinterpret_cast<int32_t*>(res + o_len), res + o_fl, res + o_sy,
                                reinterpret_cast<float*>(res + o_cfo), reinterpret_cast<float*>(res + o_to));
    if (rc < 0) return rc;
    CK(ctx->h_out.reserve(res_bytes));
    CK(cudaMemcpyAsync(ctx->h_out.p, res, res_bytes, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    const uint8_t* h = static_cast<const uint8_t*>(ctx->h_out.p);
    memcpy(h_payload, h, pay_bytes);
    memcpy(h_len, h + o_len, n_packets * 4);
    if (h_cfo) memcpy(h_cfo, h + o_cfo, n_packets * 4);
    if (h_time_offset) memcpy(h_time_offset, h + o_to, n_packets * 4);
    if (h_flags) memcpy(h_flags, h + o_fl, n_packets);
    if (h_sync) memcpy(h_sync, h + o_sy, n_packets);
    return 0;
}

// ---------------------------------------------------------------------------------------------
// AWGN Monte-Carlo driver (BASELINE.json configs[4]): per chunk of packets, all on the device,
//   payload draw -> coding chain -> lora_modulate -> channel + dechirp -> demodulate -> decode chain -> score.
int lora_phy_cuda_awgn_roundtrip(lora_phy_cuda_ctx* ctx, const lora_phy_cuda_chain_cfg* cc, float noise_std,
                                 uint64_t seed, uint32_t cell_id, uint64_t first_packet, size_t n_packets,
                                 size_t payload_len, uint64_t* d_counts, uint8_t* d_packet_ok,
                                 uint16_t* d_rx_symbols, size_t rx_symbol_stride, uint8_t* d_rx_payload,
                                 size_t rx_payload_stride, int32_t* d_rx_len)
{
    if (!ctx) return -EINVAL;
    if (!cc || !d_counts) return fail(ctx, -EINVAL, "NULL argument");
    if (payload_len > 255) return fail(ctx, -EINVAL, "payload_len exceeds 255");
    if (!(noise_std >= 0.0f)) return fail(ctx, -EINVAL, "noise_std must be >= 0");
    const long nsym_l = lora_phy_cuda_chain_symbols(cc, payload_len);
    if (nsym_l <= 0) return fail(ctx, -EINVAL, "bad chain configuration");
    const size_t nsym = (size_t)nsym_l, N = (size_t)1 << cc->sf, spp = (nsym + 2) * N;
    if ((d_rx_symbols && rx_symbol_stride < nsym) || (d_rx_payload && rx_payload_stride < payload_len))
        return fail(ctx, -ERANGE, "output stride too small");
    if (!n_packets) return 0;
    std::lock_guard<std::mutex> mc_lock(ctx->mc_mu);
    CtxGuard g(ctx);
    if (const char* e = getenv("LORA_PHY_MC_SCRATCH_MB")) {
        const long mb = atol(e);
        if (mb > 0) ctx->mc_scratch_bytes = (size_t)mb << 20;
    }
    size_t chunk = ctx->mc_scratch_bytes / (spp * sizeof(cf));
    if (chunk < 1) chunk = 1;
    if (chunk > n_packets) chunk = n_packets;
    const size_t dec_stride = 260;
    CK(ctx->mc[0].reserve(chunk * 256));
    CK(ctx->mc[1].reserve(chunk * nsym * 2));
    CK(ctx->mc[2].reserve(chunk * spp * sizeof(cf)));
    CK(ctx->mc[3].reserve(chunk * nsym * 2));
    CK(ctx->mc[4].reserve(chunk * 8));
    CK(ctx->mc[5].reserve(chunk * dec_stride));
    CK(ctx->mc[6].reserve(chunk * 4));
    CK(ctx->mc[7].reserve(chunk));
    lora_phy_cuda_cfg cfg = { cc->sf, 125000u, 1u, 0u, 0u, 0x12u, 1.0f, 0u };
    cf* table = nullptr;
    {
        int rc = get_chirp(ctx, (int)cc->sf, 125000, 1, &table);
        if (rc) return rc;
    }
    uint8_t* pay = static_cast<uint8_t*>(ctx->mc[0].p);
    uint16_t* txs = static_cast<uint16_t*>(ctx->mc[1].p);
    cf* iq = static_cast<cf*>(ctx->mc[2].p);
    uint16_t* rxs = static_cast<uint16_t*>(ctx->mc[3].p);
    int32_t* status = static_cast<int32_t*>(ctx->mc[4].p);
    uint8_t* sync = reinterpret_cast<uint8_t*>(status + chunk);
    uint8_t* dec = static_cast<uint8_t*>(ctx->mc[5].p);
    int32_t* len = static_cast<int32_t*>(ctx->mc[6].p);
    uint8_t* flags = static_cast<uint8_t*>(ctx->mc[7].p);
    for (size_t done = 0; done < n_packets; done += chunk) {
        const size_t n = std::min(chunk, n_packets - done);
        lk::McArgs a;
        a.seed = seed; a.first_packet = first_packet + done; a.cell = cell_id; a.n_packets = (uint32_t)n;
        a.payload = pay; a.payload_stride = 256; a.payload_len = (uint32_t)payload_len;
        a.iq = iq; a.packet_stride = spp; a.spp = (uint32_t)spp; a.chirp = table; a.chirp_len = (uint32_t)N;
        a.noise_std = noise_std;
        a.tx_sym = txs; a.rx_sym = rxs; a.sym_stride = nsym; a.n_sym = (uint32_t)nsym;
        a.rx_payload = dec; a.rx_stride = dec_stride; a.rx_len = len; a.status = status; a.sync = sync;
        a.sync_word = 0x12; a.counts = reinterpret_cast<unsigned long long*>(d_counts);
        a.packet_ok = d_packet_ok ? d_packet_ok + done : nullptr;
        CK(lk::launch_mc_payload(a, ctx->sm_count, ctx->stream));
        ++ctx->launches;
        long r = lora_phy_cuda_encode_chain_batch(ctx, cc, pay, 256, payload_len, n, txs, nsym);
        if (r < 0) return (int)r;
        // synthesis + AWGN + caller-side dechirp in ONE pass: the modulator's store applies the channel, so the frame
        // crosses HBM once on its way to the demodulator (LORA_PHY_MC_UNFUSED=1: the separate channel kernel, A/B)
        static const bool unfused = getenv("LORA_PHY_MC_UNFUSED") != nullptr;
        if (!unfused) {
            ctx->mc_channel.on = 1; ctx->mc_channel.seed = seed; ctx->mc_channel.first_packet = first_packet + done;
            ctx->mc_channel.cell = cell_id; ctx->mc_channel.noise_std = noise_std; ctx->mc_channel.chirp = table;
            ctx->mc_channel.chirp_mask = (uint32_t)N - 1;
        }
        r = lora_phy_cuda_modulate_batch(ctx, &cfg, txs, nsym, nsym, n, iq, spp);
        ctx->mc_channel.on = 0;
        if (r < 0) return (int)r;
        if (unfused) {
            CK(lk::launch_awgn_dechirp(a, ctx->sm_count, ctx->stream));
            ++ctx->launches;
        }
        int rc = lora_phy_cuda_demodulate_batch(ctx, &cfg, iq, spp, nullptr, spp, n, rxs, nsym, sync, nullptr, nullptr, status);
        if (rc < 0) return rc;
        rc = lora_phy_cuda_decode_chain_batch(ctx, cc, rxs, nsym, status, nsym, n, dec, dec_stride, len, flags);
        if (rc < 0) return rc;
        CK(lk::launch_mc_score(a, ctx->stream));
        ++ctx->launches;
        if (d_rx_symbols)
            CK(cudaMemcpy2DAsync(d_rx_symbols + done * rx_symbol_stride, rx_symbol_stride * 2, rxs, nsym * 2, nsym * 2, n,
                                 cudaMemcpyDeviceToDevice, ctx->stream));
        if (d_rx_payload && payload_len)
            CK(cudaMemcpy2DAsync(d_rx_payload + done * rx_payload_stride, rx_payload_stride, dec, dec_stride, payload_len, n,
                                 cudaMemcpyDeviceToDevice, ctx->stream));
        if (d_rx_len) CK(cudaMemcpyAsync(d_rx_len + done, len, n * 4, cudaMemcpyDeviceToDevice, ctx->stream));
    }
    return 0;
}

int lora_phy_cuda_copy_table(lora_phy_cuda_ctx* ctx, uint32_t sf, uint32_t bw_hz, int which, void* h_out)
{
    if (!ctx || !h_out) return -EINVAL;
    if (sf < 2 || sf > 12) return fail(ctx, -EINVAL, "sf out of range 2..12");
    CtxGuard g(ctx);
    const size_t N = (size_t)1 << sf;
    if (which == 0) { CK(cudaMemcpyAsync(h_out, ctx->tab[sf].tw, N * sizeof(cf), cudaMemcpyDeviceToHost, ctx->stream)); }
    else if (which == 1) { CK(cudaMemcpyAsync(h_out, ctx->tab[sf].hann, N * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream)); }
    else if (which == 2) {
        if (!bw_ok(bw_hz)) return fail(ctx, -EINVAL, "bad bw_hz");
        cf* t = nullptr;
        int rc = get_chirp(ctx, (int)sf, (int)bw_hz, 1, &t);
        if (rc) return rc;
        CK(cudaMemcpyAsync(h_out, t, N * sizeof(cf), cudaMemcpyDeviceToHost, ctx->stream));
    } else return fail(ctx, -EINVAL, "unknown table");
    CK(cudaStreamSynchronize(ctx->stream));
    return 0;
}

} // extern "C"
