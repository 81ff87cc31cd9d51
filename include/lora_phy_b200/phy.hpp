// lora_phy_b200/phy.hpp -- the reference's public C++ API (include/lora_phy/phy.hpp:104-224) as thin,
// header-only C++17 wrappers over the C ABI of liblora_phy_cuda.so (include/lora_phy_cuda.h).
//
// Same namespace, names, argument meaning and error precedence as the reference, so its own test
// sources compile against this header unchanged (add -llora_phy_cuda).  Every call is a batch of ONE
// packet through the *_host entry points; the many-packet entry point is the C ABI itself.
// Differences, all deliberate and listed in DESIGN.md:
//   * capacity is checked BEFORE anything is written (the reference reports -ERANGE after writing,
//     phy.cpp:62-65,74-77);
//   * init() rejects sf outside 2..12 with -EINVAL (the reference overruns twiddles[4096]);
//   * there is no CPU path: without an sm_100 device every compute call returns -ENODEV.
#pragma once

#include <cerrno>
#include <complex>
#include <cstddef>
#include <cstdint>
#include <mutex>
#include <stdexcept>
#include <string>
#include <sys/types.h>
#include <vector>

#include "../lora_phy_cuda.h"

// Plan carrier kept for source compatibility (phy.hpp:85-86 touch plan_fwd / plan_inv; phy.cpp:14-18
// deduces the spreading factor from plan_fwd.nfft).  The GPU path holds its plans in the context.
template <typename T>
struct kissfft_plan {
    std::size_t nfft{};
    bool inverse{};
};

namespace lora_phy {

constexpr float PI = 3.14159265358979323846f;

enum class window_type { window_none, window_hann };                       // phy.hpp:31-34
enum class bandwidth : unsigned { bw_125 = 125000, bw_250 = 250000, bw_500 = 500000 };   // phy.hpp:40-44
constexpr float bw_to_hz(bandwidth bw) { return static_cast<float>(static_cast<unsigned>(bw)); }
constexpr float bw_scale(bandwidth bw) { return bw_to_hz(bw) / 125000.0f; }

struct lora_params {                                                         // phy.hpp:53-60
    unsigned sf{};
    bandwidth bw{bandwidth::bw_125};
    unsigned cr{};
    unsigned osr{1};
    window_type window{window_type::window_none};
    uint8_t sync_word{0x12};
};

struct lora_metrics {                                                        // phy.hpp:67-71
    bool crc_ok{};
    float cfo{};
    float time_offset{};
};

struct lora_workspace {                                                      // phy.hpp:79-94 (same fields, same order)
    uint16_t* symbol_buf{};
    std::complex<float>* fft_in{};
    std::complex<float>* fft_out{};
    float* window{};
    window_type window_kind{window_type::window_none};
    kissfft_plan<float> plan_fwd{};
    kissfft_plan<float> plan_inv{};
    lora_metrics metrics{};
    unsigned osr{1};
    bandwidth bw{bandwidth::bw_125};
    uint8_t sync_word{0x12};
};

struct lora_demod_workspace {                                                // phy.hpp:175-192 (host-visible part)
    static const std::size_t MAX_N = 4096;
    std::size_t N{};
    float window[MAX_N];
    window_type window_kind{window_type::window_none};
    lora_metrics metrics{};
    std::complex<float>* scratch{};
    std::size_t scratch_len{};
    unsigned sf{};
};

namespace detail {

// One context per process and device, created on first use (the reference's API has no handle to
// hang it on; buffers stay caller-owned as in phy.hpp:3-7).
inline lora_phy_cuda_ctx* shared_ctx(int* err = nullptr)
{
    static std::mutex mu;
    static lora_phy_cuda_ctx* ctx = nullptr;
    static int rc = 0;
    std::lock_guard<std::mutex> lock(mu);
    if (!ctx && rc == 0) rc = lora_phy_cuda_create(&ctx, 0);
    if (err) *err = rc;
    return ctx;
}

inline unsigned deduce_sf(const lora_workspace* ws)                          // phy.cpp:13-18
{
    unsigned sf = 0;
    while ((std::size_t(1) << sf) < ws->plan_fwd.nfft) ++sf;
    return sf;
}

inline lora_phy_cuda_cfg make_cfg(unsigned sf, bandwidth bw, unsigned osr, window_type w, uint32_t flags,
                                  uint8_t sync = 0x12, float amplitude = 1.0f)
{
    lora_phy_cuda_cfg c{};
    c.sf = sf; c.bw_hz = static_cast<unsigned>(bw); c.osr = osr ? osr : 1u;
    c.window = w == window_type::window_hann ? 1u : 0u;
    c.flags = flags; c.sync_word = sync; c.amplitude = amplitude;
    return c;
}

} // namespace detail

// ---- high-level API (phy.hpp:104-161) ------------------------------------------------------------
inline int init(lora_workspace* ws, const lora_params* cfg)                  // phy.cpp:27-52
{
    if (!ws || !cfg) return -EINVAL;
    if (cfg->sf < 2 || cfg->sf > 12) return -EINVAL;
    ws->plan_fwd.nfft = std::size_t(1) << cfg->sf; ws->plan_fwd.inverse = false;
    ws->plan_inv.nfft = ws->plan_fwd.nfft;         ws->plan_inv.inverse = true;
    ws->metrics = {};
    ws->osr = cfg->osr ? cfg->osr : 1u;
    ws->bw = cfg->bw;
    ws->sync_word = cfg->sync_word;
    ws->window_kind = cfg->window;
    if (ws->window_kind != window_type::window_none && !ws->window) return -ENOMEM;
    if (ws->window) {
        if (ws->window_kind == window_type::window_hann) {
            int rc = 0;
            lora_phy_cuda_ctx* ctx = detail::shared_ctx(&rc);
            if (!ctx) return rc;
            rc = lora_phy_cuda_copy_table(ctx, cfg->sf, 125000, 1, ws->window);   // bit-exact Hann, LoRaDemod.cpp:18-24
            if (rc) return rc;
        } else {
            for (std::size_t i = 0; i < ws->plan_fwd.nfft; ++i) ws->window[i] = 1.0f;
        }
    }
    return 0;
}

inline void reset(lora_workspace* ws) { if (ws) ws->metrics = {}; }         // phy.cpp:54-56

inline const lora_metrics* get_last_metrics(const lora_workspace* ws) { return ws ? &ws->metrics : nullptr; }

inline ssize_t encode(lora_workspace* ws, const uint8_t* payload, std::size_t payload_len, uint16_t* symbols,
                      std::size_t symbol_cap)                                  // phy.cpp:58-66
{
    if (!ws || !payload || !symbols) return -EINVAL;
    if (2 * payload_len > symbol_cap) return -ERANGE;
    int rc = 0;
    lora_phy_cuda_ctx* ctx = detail::shared_ctx(&rc);
    if (!ctx) return rc;
    return lora_phy_cuda_encode_batch_host(ctx, payload, payload_len, payload_len, 1, symbols, 2 * payload_len);
}

inline ssize_t decode(lora_workspace* ws, const uint16_t* symbols, std::size_t symbol_count, uint8_t* payload,
                      std::size_t payload_cap)                                 // phy.cpp:245-261
{
    if (!ws || !symbols || !payload) return -EINVAL;
    if (symbol_count & 1) return -EINVAL;                                    // LoRaDecoder.cpp:10
    if (symbol_count / 2 > payload_cap) return -ERANGE;
    int rc = 0;
    lora_phy_cuda_ctx* ctx = detail::shared_ctx(&rc);
    if (!ctx) return rc;
    uint8_t crc = 0;
    const long r = lora_phy_cuda_decode_batch_host(ctx, symbols, symbol_count, nullptr, symbol_count, 1, payload,
                                                   symbol_count / 2 ? symbol_count / 2 : 1, &crc);
    if (r >= 0) ws->metrics.crc_ok = crc != 0;
    return r;
}

inline ssize_t modulate(lora_workspace* ws, const uint16_t* symbols, std::size_t symbol_count,
                        std::complex<float>* iq, std::size_t iq_cap)           // phy.cpp:68-79
{
    if (!ws || !symbols || !iq) return -EINVAL;
    const unsigned sf = detail::deduce_sf(ws);
    const std::size_t produced = (symbol_count + 2) * (std::size_t(1) << sf) * (ws->osr ? ws->osr : 1u);
    if (produced > iq_cap) return -ERANGE;
    int rc = 0;
    lora_phy_cuda_ctx* ctx = detail::shared_ctx(&rc);
    if (!ctx) return rc;
    const lora_phy_cuda_cfg c = detail::make_cfg(sf, ws->bw, ws->osr, window_type::window_none, 0, ws->sync_word, 1.0f);
    return lora_phy_cuda_modulate_batch_host(ctx, &c, symbols, symbol_count ? symbol_count : 1, symbol_count, 1, iq, produced);
}

inline ssize_t demodulate(lora_workspace* ws, const std::complex<float>* iq, std::size_t sample_count,
                          uint16_t* symbols, std::size_t symbol_cap)           // phy.cpp:182-243
{
    if (!ws || !iq || !symbols) return -EINVAL;
    const unsigned sf = detail::deduce_sf(ws);
    const unsigned osr = ws->osr ? ws->osr : 1u;
    const std::size_t step = (std::size_t(1) << sf) * osr;
    if (sample_count % step != 0) return -EINVAL;
    const std::size_t total = sample_count / step;
    if (total < 2) return -ERANGE;
    if (total - 2 > symbol_cap) return -ERANGE;
    int rc = 0;
    lora_phy_cuda_ctx* ctx = detail::shared_ctx(&rc);
    if (!ctx) return rc;
    const lora_phy_cuda_cfg c = detail::make_cfg(sf, ws->bw, osr, ws->window ? ws->window_kind : window_type::window_none,
                                                 LORA_PHY_CUDA_F_DECHIRP);
    uint8_t sync = 0;
    float cfo = 0.0f, to = 0.0f;
    int32_t status = 0;
    rc = lora_phy_cuda_demodulate_batch_host(ctx, &c, iq, sample_count, nullptr, sample_count, 1, symbols,
                                             total - 2 ? total - 2 : 1, &sync, &cfo, &to, &status);
    if (rc) return rc;
    if (status < 0) return status;
    ws->metrics.cfo = cfo;
    ws->metrics.time_offset = to;
    ws->sync_word = sync;                                                    // phy.cpp:240-241
    return status;
}

inline void estimate_offsets(lora_workspace* ws, const std::complex<float>* samples, std::size_t sample_count)
{                                                                            // phy.cpp:81-148; silent on bad input (:84)
    if (!ws || !samples) return;
    const unsigned sf = detail::deduce_sf(ws);
    const unsigned osr = ws->osr ? ws->osr : 1u;
    if (sample_count / ((std::size_t(1) << sf) * osr) == 0) return;
    lora_phy_cuda_ctx* ctx = detail::shared_ctx();
    if (!ctx) return;
    const lora_phy_cuda_cfg c = detail::make_cfg(sf, ws->bw, osr, ws->window ? ws->window_kind : window_type::window_none, 0);
    float cfo = ws->metrics.cfo, to = ws->metrics.time_offset;
    if (lora_phy_cuda_estimate_offsets_batch_host(ctx, &c, samples, sample_count, sample_count, 1, &cfo, &to) == 0) {
        ws->metrics.cfo = cfo;
        ws->metrics.time_offset = to;
    }
}

inline void compensate_offsets(const lora_workspace* ws, std::complex<float>* samples, std::size_t sample_count)
{                                                                            // phy.cpp:150-180; silent on bad input (:153)
    if (!ws || !samples || !sample_count) return;
    lora_phy_cuda_ctx* ctx = detail::shared_ctx();
    if (!ctx) return;
    const lora_phy_cuda_cfg c = detail::make_cfg(detail::deduce_sf(ws), ws->bw, ws->osr, window_type::window_none, 0);
    const float cfo = ws->metrics.cfo, to = ws->metrics.time_offset;
    lora_phy_cuda_compensate_offsets_batch_host(ctx, &c, samples, sample_count, sample_count, 1, &cfo, &to);
}

// ---- legacy helpers (phy.hpp:195-224) ---------------------------------------------------------------
inline void lora_demod_init(lora_demod_workspace* ws, unsigned sf, window_type win = window_type::window_none,
                            std::complex<float>* scratch = nullptr, std::size_t max_samples = 0)
{                                                                            // LoRaDemod.cpp:11-33
    if (!ws || sf < 2 || sf > 12) return;
    ws->sf = sf;
    ws->N = std::size_t(1) << sf;
    ws->window_kind = win;
    ws->metrics = {};
    ws->scratch = scratch;
    ws->scratch_len = max_samples;
    if (win == window_type::window_hann) {
        if (lora_phy_cuda_ctx* ctx = detail::shared_ctx()) lora_phy_cuda_copy_table(ctx, sf, 125000, 1, ws->window);
    } else {
        for (std::size_t i = 0; i < ws->N; ++i) ws->window[i] = 1.0f;
    }
}

inline void lora_demod_free(lora_demod_workspace* ws) { if (ws) { ws->N = 0; ws->scratch = nullptr; ws->scratch_len = 0; } }

inline std::size_t lora_modulate(const uint16_t* symbols, std::size_t symbol_count, std::complex<float>* out_samples,
                                 unsigned sf, unsigned osr, bandwidth bw, float amplitude = 1.0f, uint8_t sync = 0x12)
{                                                                            // LoRaMod.cpp:8-43
    lora_phy_cuda_ctx* ctx = detail::shared_ctx();
    if (!ctx || !out_samples) return 0;
    const lora_phy_cuda_cfg c = detail::make_cfg(sf, bw, osr, window_type::window_none, 0, sync, amplitude);
    const std::size_t produced = (symbol_count + 2) * (std::size_t(1) << sf) * (osr ? osr : 1u);
    const long r = lora_phy_cuda_modulate_batch_host(ctx, &c, symbols, symbol_count ? symbol_count : 1, symbol_count, 1,
                                                     out_samples, produced);
    return r < 0 ? 0 : static_cast<std::size_t>(r);
}

inline ssize_t lora_demodulate(lora_demod_workspace* ws, const std::complex<float>* samples, std::size_t sample_count,
                               uint16_t* out_symbols, unsigned osr, uint8_t* out_sync = nullptr)
{                                                                            // LoRaDemod.cpp:50-197
    if (!ws || !samples || !out_symbols || !osr || !ws->N) return -EINVAL;   // (the reference does not validate, :50-58)
    int rc = 0;
    lora_phy_cuda_ctx* ctx = detail::shared_ctx(&rc);
    if (!ctx) return rc;
    const bool have_scratch = ws->scratch && ws->scratch_len >= sample_count;   // LoRaDemod.cpp:70
    const lora_phy_cuda_cfg c = detail::make_cfg(ws->sf, bandwidth::bw_125, osr, ws->window_kind,
                                                 have_scratch ? 0u : LORA_PHY_CUDA_F_NO_SCRATCH);
    const std::size_t total = sample_count / (ws->N * osr);
    const std::size_t cap = total >= 2 ? total - 2 : total;
    uint8_t sync = 0;
    float cfo = 0.0f, to = 0.0f;
    int32_t status = 0;
    rc = lora_phy_cuda_demodulate_batch_host(ctx, &c, samples, sample_count, nullptr, sample_count, 1, out_symbols,
                                             cap ? cap : 1, &sync, &cfo, &to, &status);
    if (rc) return rc;
    if (status < 0) return status;
    ws->metrics.cfo = cfo;
    ws->metrics.time_offset = to;
    if (out_sync) *out_sync = sync;
    return status;
}

inline std::size_t lora_encode(const uint8_t* bytes, std::size_t byte_count, uint16_t* out_symbols, unsigned /*sf*/)
{                                                                            // LoRaEncoder.cpp:6-18
    lora_phy_cuda_ctx* ctx = detail::shared_ctx();
    if (!ctx || !bytes || !out_symbols) return 0;
    const long r = lora_phy_cuda_encode_batch_host(ctx, bytes, byte_count, byte_count, 1, out_symbols, 2 * byte_count);
    return r < 0 ? 0 : static_cast<std::size_t>(r);
}

inline ssize_t lora_decode(const uint16_t* symbols, std::size_t symbol_count, uint8_t* out_bytes)
{                                                                            // LoRaDecoder.cpp:7-21
    if (symbol_count & 1) return -EINVAL;
    int rc = 0;
    lora_phy_cuda_ctx* ctx = detail::shared_ctx(&rc);
    if (!ctx) return rc;
    if (!symbols || !out_bytes) return -EINVAL;
    return lora_phy_cuda_decode_batch_host(ctx, symbols, symbol_count, nullptr, symbol_count, 1, out_bytes,
                                           symbol_count / 2 ? symbol_count / 2 : 1, nullptr);
}

// ---- README facade (README.md:40-59 of the reference): LoRaParams / LoRaPHY ----------------------------------
// The reference's README advertises this class but its sources never define it; here it is real, over the
// whole-modem pipelines of the C ABI: modulate = upstream coding chain (CRC, explicit header, FEC at `cr`,
// whitening, interleaver, Gray: external/LoRa-SDR/LoRaEncoder.cpp:161-233) + lora_modulate (LoRaMod.cpp:8-43);
// demodulate = dechirp + lora_demodulate (LoRaDemod.cpp:50-197) + decode chain (LoRaDecoder.cpp:196-397).
// The *_batch members are the many-packet entry points (equal-length payloads / packets, row-major).
struct LoRaParams {
    unsigned sf = 7;
    unsigned bw = 125000;
    unsigned cr = 1;                 // coding rate 4/(4+cr), 1..4 (0 = uncoded)
    bool explicit_header = true;
    bool crc_enabled = true;
    unsigned osr = 1;
    uint8_t sync_word = 0x12;
    unsigned implicit_length = 0;    // payload bytes when explicit_header is false
};

class LoRaPHY {
public:
    explicit LoRaPHY(const LoRaParams& p) : p_(p)
    {
        int rc = 0;
        ctx_ = detail::shared_ctx(&rc);
        if (!ctx_) throw std::runtime_error("lora_phy_cuda_create failed (" + std::to_string(rc) + "): no sm_100 device");
        cfg_ = lora_phy_cuda_cfg{};
        cfg_.sf = p.sf; cfg_.bw_hz = p.bw; cfg_.osr = p.osr ? p.osr : 1u; cfg_.sync_word = p.sync_word; cfg_.amplitude = 1.0f;
        cc_ = lora_phy_cuda_chain_cfg{};
        cc_.sf = p.sf; cc_.cr = p.cr; cc_.explicit_header = p.explicit_header; cc_.crc = p.crc_enabled; cc_.whitening = 1;
        cc_.crc_check = 1; cc_.data_length = p.implicit_length;
        if (lora_phy_cuda_chain_symbols(&cc_, 1) < 0 || (p.bw != 125000 && p.bw != 250000 && p.bw != 500000))
            throw std::invalid_argument("LoRaParams: sf 7..12, bw 125/250/500 kHz, cr 0..4");
    }

    std::size_t samples_for(std::size_t payload_len) const
    {
        return (static_cast<std::size_t>(lora_phy_cuda_chain_symbols(&cc_, payload_len)) + 2) * (std::size_t(1) << p_.sf) * cfg_.osr;
    }

    std::vector<std::complex<float>> modulate(const std::vector<uint8_t>& payload) { return modulate_batch(payload, 1); }

    // n_packets payloads of payload.size() / n_packets bytes each -> n_packets * samples_for(len) samples
    std::vector<std::complex<float>> modulate_batch(const std::vector<uint8_t>& payload, std::size_t n_packets)
    {
        if (!n_packets || payload.size() % n_packets) throw std::invalid_argument("payload size not a multiple of n_packets");
        const std::size_t len = payload.size() / n_packets, spp = samples_for(len);
        std::vector<std::complex<float>> iq(n_packets * spp);
        const long r = lora_phy_cuda_tx_batch_host(ctx_, &cfg_, &cc_, payload.data(), len, len, n_packets, iq.data(), spp);
        if (r < 0) throw std::runtime_error(std::string("modulate: ") + lora_phy_cuda_last_error(ctx_));
        return iq;
    }

    // One packet of whole symbols -> payload bytes; empty when the decoder dropped it (bad header / CRC).
    std::vector<uint8_t> demodulate(const std::vector<std::complex<float>>& iq)
    {
        auto rows = demodulate_batch(iq, 1);
        return rows.empty() ? std::vector<uint8_t>{} : rows[0];
    }

    std::vector<std::vector<uint8_t>> demodulate_batch(const std::vector<std::complex<float>>& iq, std::size_t n_packets)
    {
        if (!n_packets || iq.size() % n_packets) throw std::invalid_argument("sample count not a multiple of n_packets");
        const std::size_t spp = iq.size() / n_packets;
        std::vector<uint8_t> pay(n_packets * 260);
        len_.assign(n_packets, 0); flags_.assign(n_packets, 0); sync_.assign(n_packets, 0);
        cfo_.assign(n_packets, 0.0f); to_.assign(n_packets, 0.0f);
        const int rc = lora_phy_cuda_rx_batch_host(ctx_, &cfg_, &cc_, iq.data(), spp, spp, n_packets, pay.data(), 260,
                                                   len_.data(), flags_.data(), sync_.data(), cfo_.data(), to_.data());
        if (rc < 0) throw std::runtime_error(std::string("demodulate: ") + lora_phy_cuda_last_error(ctx_));
        std::vector<std::vector<uint8_t>> out(n_packets);
        for (std::size_t p = 0; p < n_packets; ++p)
            if (len_[p] > 0) out[p].assign(pay.begin() + p * 260, pay.begin() + p * 260 + len_[p]);
        return out;
    }

    // results of the last demodulate call, per packet
    bool crc_ok(std::size_t p = 0) const { return p < flags_.size() && (flags_[p] & 0x0c) == 0x0c; }
    bool dropped(std::size_t p = 0) const { return p < len_.size() && len_[p] < 0; }
    uint8_t sync_word(std::size_t p = 0) const { return p < sync_.size() ? sync_[p] : 0; }
    lora_metrics metrics(std::size_t p = 0) const
    {
        lora_metrics m{};
        if (p < cfo_.size()) { m.cfo = cfo_[p]; m.time_offset = to_[p]; m.crc_ok = crc_ok(p); }
        return m;
    }
    const LoRaParams& params() const { return p_; }

private:
    LoRaParams p_;
    lora_phy_cuda_ctx* ctx_ = nullptr;
    lora_phy_cuda_cfg cfg_{};
    lora_phy_cuda_chain_cfg cc_{};
    std::vector<int32_t> len_;
    std::vector<uint8_t> flags_, sync_;
    std::vector<float> cfo_, to_;
};

} // namespace lora_phy

// The README uses the facade unqualified.
using lora_phy::LoRaParams;
using lora_phy::LoRaPHY;
