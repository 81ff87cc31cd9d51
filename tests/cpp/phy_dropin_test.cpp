// Drop-in check of include/lora_phy_b200/phy.hpp: the calls and expectations mirror the reference's
// tests/error_code_test.cpp (argument precedence), tests/roundtrip_test.cpp:30-31 (DEADBEEF),
// tests/no_alloc_test.cpp:35 (symbols 0,1,12,34,56 at SF7) and the README "Hello" round trip.
//   phy_dropin_test args   -> argument / errno contract only (runs without a GPU)
//   phy_dropin_test gpu    -> everything (needs a B200)
#include <lora_phy/phy.hpp>
#include <cstdio>
#include <cstring>
#include <vector>

using namespace lora_phy;
static int fails = 0;
#define EXPECT(cond) do { if (!(cond)) { std::printf("FAIL %s:%d %s\n", __FILE__, __LINE__, #cond); ++fails; } } while (0)

int main(int argc, char** argv)
{
    const bool gpu = argc > 1 && std::strcmp(argv[1], "gpu") == 0;
    lora_params cfg{};
    cfg.sf = 7; cfg.bw = bandwidth::bw_125; cfg.cr = 1; cfg.osr = 1; cfg.sync_word = 0x12;

    // ---- init (error_code_test.cpp:19-33) ----
    EXPECT(init(nullptr, &cfg) == -EINVAL);
    lora_workspace tmp{};
    EXPECT(init(&tmp, nullptr) == -EINVAL);
    cfg.window = window_type::window_hann;
    EXPECT(init(&tmp, &cfg) == -ENOMEM);
    cfg.window = window_type::window_none;
    lora_params bad = cfg; bad.sf = 13;
    EXPECT(init(&tmp, &bad) == -EINVAL);                       // tightened

    const size_t N = size_t(1) << cfg.sf;
    std::vector<std::complex<float>> fft_in(N), fft_out(N);
    std::vector<float> window(N);
    lora_workspace ws{};
    ws.fft_in = fft_in.data(); ws.fft_out = fft_out.data(); ws.window = window.data();
    EXPECT(init(&ws, &cfg) == 0);
    EXPECT(window[0] == 1.0f && window[N - 1] == 1.0f);
    EXPECT(get_last_metrics(nullptr) == nullptr && get_last_metrics(&ws) == &ws.metrics);

    // ---- encode / modulate / demodulate / decode argument errors (error_code_test.cpp:47-120) ----
    uint8_t payload[4] = {1, 2, 3, 4};
    uint16_t symbols[64];
    EXPECT(encode(nullptr, payload, 4, symbols, 8) == -EINVAL);
    EXPECT(encode(&ws, nullptr, 4, symbols, 8) == -EINVAL);
    EXPECT(encode(&ws, payload, 4, symbols, 1) == -ERANGE);
    std::complex<float> iq8[8];
    EXPECT(modulate(nullptr, symbols, 1, iq8, 8) == -EINVAL);
    EXPECT(modulate(&ws, symbols, 1, iq8, 1) == -ERANGE);
    std::complex<float> ten[10];
    EXPECT(demodulate(nullptr, ten, 10, symbols, 8) == -EINVAL);
    EXPECT(demodulate(&ws, ten, 10, symbols, 8) == -EINVAL);            // misaligned
    std::vector<std::complex<float>> few(N), many(N * 4);
    EXPECT(demodulate(&ws, few.data(), N, symbols, 8) == -ERANGE);      // fewer than two symbols
    EXPECT(demodulate(&ws, many.data(), N * 4, symbols, 1) == -ERANGE); // capacity
    uint8_t out[64];
    EXPECT(decode(nullptr, symbols, 2, out, 8) == -EINVAL);
    EXPECT(decode(&ws, symbols, 3, out, 8) == -EINVAL);                 // odd count
    EXPECT(decode(&ws, symbols, 8, out, 1) == -ERANGE);
    EXPECT(lora_decode(symbols, 3, out) == -EINVAL);
    estimate_offsets(nullptr, ten, 10);                                 // silent (phy.cpp:84)
    compensate_offsets(nullptr, ten, 10);                               // silent (phy.cpp:153)

    if (!gpu) {
        // no device, no fallback: compute calls report -ENODEV instead of running anywhere else
        int rc = 0;
        if (!detail::shared_ctx(&rc)) {
            EXPECT(rc == -ENODEV);
            EXPECT(encode(&ws, payload, 4, symbols, 8) == -ENODEV);
        }
        std::printf(fails ? "phy_dropin_test args: %d failure(s)\n" : "phy_dropin_test args: ok\n", fails);
        return fails ? 1 : 0;
    }

    // ---- DEADBEEF (roundtrip_test.cpp:30-31) ----
    const uint8_t dead[4] = {0xDE, 0xAD, 0xBE, 0xEF};
    EXPECT(encode(&ws, dead, 4, symbols, 64) == 8);
    EXPECT(decode(&ws, symbols, 8, out, 64) == 4 && std::memcmp(out, dead, 4) == 0);
    EXPECT(lora_encode(dead, 4, symbols, 7) == 8);
    EXPECT(lora_decode(symbols, 8, out) == 4 && std::memcmp(out, dead, 4) == 0);

    // ---- legacy path: symbols 0,1,12,34,56 (no_alloc_test.cpp:35): modulate, caller-side dechirp, demodulate ----
    const uint16_t syms[5] = {0, 1, 12, 34, 56};
    std::vector<std::complex<float>> iq(7 * N), down(N), scratch(7 * N);
    EXPECT(lora_modulate(syms, 5, iq.data(), 7, 1, bandwidth::bw_125, 1.0f, 0x12) == 7 * N);
    EXPECT(lora_phy_cuda_copy_table(detail::shared_ctx(), 7, 125000, 2, down.data()) == 0);
    for (size_t i = 0; i < iq.size(); ++i) iq[i] *= down[i % N];
    lora_demod_workspace dws;
    lora_demod_init(&dws, 7, window_type::window_none, scratch.data(), scratch.size());
    uint16_t got[8] = {0};
    uint8_t sync = 0;
    EXPECT(lora_demodulate(&dws, iq.data(), iq.size(), got, 1, &sync) == 5);
    EXPECT(std::memcmp(got, syms, sizeof syms) == 0 && sync == 0x12);
    lora_demod_workspace nos;
    lora_demod_init(&nos, 7);
    for (auto& x : iq) x *= 2.0f;                                        // above 1.0 without scratch: LoRaDemod.cpp:70-72
    EXPECT(lora_demodulate(&nos, iq.data(), iq.size(), got, 1, &sync) == -ERANGE);
    EXPECT(lora_demodulate(&dws, iq.data(), iq.size(), got, 1, &sync) == 5 && std::memcmp(got, syms, sizeof syms) == 0);
    lora_demod_free(&dws);

    // ---- README "Hello": encode -> modulate -> demodulate -> decode through the workspace API ----
    const uint8_t hello[5] = {0x48, 0x65, 0x6C, 0x6C, 0x6F};
    EXPECT(encode(&ws, hello, 5, symbols, 64) == 10);
    std::vector<std::complex<float>> tx(12 * N);
    EXPECT(modulate(&ws, symbols, 10, tx.data(), tx.size()) == (ssize_t)(12 * N));
    uint16_t rx[16];
    const ssize_t n = demodulate(&ws, tx.data(), tx.size(), rx, 16);
    EXPECT(n == 10);
    const ssize_t nb = decode(&ws, rx, 10, out, 64);
    EXPECT(nb == 5);
    std::printf("hello -> %02x%02x%02x%02x%02x cfo=%g time_offset=%g sync=%02x crc_ok=%d\n", out[0], out[1], out[2], out[3],
                out[4], ws.metrics.cfo, ws.metrics.time_offset, ws.sync_word, (int)ws.metrics.crc_ok);
    // the reference's own high-level path returns 0ca9a0a0a3 here (its estimator is ill-posed on
    // un-dechirped input); parity means reproducing that, which tests/test_gpu_parity.py checks against the oracle
    EXPECT(out[0] == 0x0c && out[1] == 0xa9 && out[2] == 0xa0 && out[3] == 0xa0 && out[4] == 0xa3);

    // ---- README facade, verbatim (reference README.md:42-59): LoRaParams / LoRaPHY ----
    {
        LoRaParams params;
        params.sf = 7; params.bw = 125000; params.cr = 1; params.explicit_header = true; params.crc_enabled = true;
        LoRaPHY phy(params);
        std::vector<uint8_t> payload = {0x48, 0x65, 0x6C, 0x6C, 0x6F};
        auto iq_samples = phy.modulate(payload);
        EXPECT(iq_samples.size() == (18 + 2) * 128);                     // CR4/5 explicit header CRC: 18 data symbols
        auto decoded_payload = phy.demodulate(iq_samples);
        EXPECT(decoded_payload == payload);
        EXPECT(phy.crc_ok() && !phy.dropped() && phy.sync_word() == 0x12);
        iq_samples[700] = -iq_samples[700];                              // a glitch the FEC / FFT gain absorbs
        EXPECT(phy.demodulate(iq_samples) == payload);
        for (std::size_t i = 5 * 128; i < 9 * 128; ++i) iq_samples[i] = 0;   // four symbols wiped: CRC / header must catch it
        EXPECT(phy.demodulate(iq_samples) != payload);
        // many-packet entry point: 6 packets x 9 bytes at SF9 CR4/8
        LoRaParams p9; p9.sf = 9; p9.cr = 4;
        LoRaPHY phy9(p9);
        std::vector<uint8_t> batch(6 * 9);
        for (std::size_t i = 0; i < batch.size(); ++i) batch[i] = (uint8_t)(i * 37 + 11);
        auto iq9 = phy9.modulate_batch(batch, 6);
        EXPECT(iq9.size() == 6 * phy9.samples_for(9));
        auto rows = phy9.demodulate_batch(iq9, 6);
        for (std::size_t p = 0; p < 6; ++p)
            EXPECT(rows[p] == std::vector<uint8_t>(batch.begin() + 9 * p, batch.begin() + 9 * p + 9) && phy9.crc_ok(p));
        bool threw = false;
        try { LoRaParams bad; bad.sf = 13; LoRaPHY x(bad); } catch (const std::invalid_argument&) { threw = true; }
        EXPECT(threw);
    }

    std::printf(fails ? "phy_dropin_test gpu: %d failure(s)\n" : "phy_dropin_test gpu: ok\n", fails);
    return fails ? 1 : 0;
}
