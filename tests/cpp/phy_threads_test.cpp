// Two host threads, one lora_demod_workspace each (how the reference is used in parallel: phy.hpp:74-78, one
// workspace per thread), through the drop-in header -- which routes both through ONE process-wide context.
// Every entry point holds the context's mutex from staging to copy-out, so the threads must serialise instead
// of racing on the staging buffers: each thread's symbols must be its own, every iteration.
//   phy_threads_test   (needs a B200)
#include <lora_phy/phy.hpp>
#include <atomic>
#include <cstdio>
#include <cstring>
#include <thread>
#include <vector>

using namespace lora_phy;
static std::atomic<int> fails{0};
#define EXPECT(cond) do { if (!(cond)) { std::printf("FAIL %s:%d %s\n", __FILE__, __LINE__, #cond); ++fails; } } while (0)

static void worker(unsigned sf, unsigned seed, int iters)
{
    const size_t N = size_t(1) << sf;
    std::vector<std::complex<float>> down(N);
    EXPECT(lora_phy_cuda_copy_table(detail::shared_ctx(), sf, 125000, 2, down.data()) == 0);
    for (int it = 0; it < iters; ++it) {
        const size_t n_sym = 3 + (size_t)((seed + it) % 9);            // different sizes: the staging buffers get re-reserved
        std::vector<uint16_t> syms(n_sym), got(n_sym + 2);
        unsigned x = seed * 2654435761u + (unsigned)it * 40503u;
        for (auto& s : syms) { x = x * 1664525u + 1013904223u; s = (uint16_t)((x >> 8) & (N - 1)); }
        std::vector<std::complex<float>> iq((n_sym + 2) * N), scratch((n_sym + 2) * N);
        EXPECT(lora_modulate(syms.data(), n_sym, iq.data(), sf, 1, bandwidth::bw_125, 1.0f, 0x12) == (n_sym + 2) * N);
        for (size_t i = 0; i < iq.size(); ++i) iq[i] *= down[i % N];
        lora_demod_workspace ws;
        lora_demod_init(&ws, sf, window_type::window_none, scratch.data(), scratch.size());
        uint8_t sync = 0;
        const ssize_t r = lora_demodulate(&ws, iq.data(), iq.size(), got.data(), 1, &sync);
        EXPECT(r == (ssize_t)n_sym);
        EXPECT(sync == 0x12);
        EXPECT(std::memcmp(got.data(), syms.data(), n_sym * sizeof(uint16_t)) == 0);
        lora_demod_free(&ws);
    }
}

int main()
{
    std::thread a(worker, 7u, 1u, 40), b(worker, 9u, 2u, 40), c(worker, 8u, 3u, 40);
    a.join(); b.join(); c.join();
    std::printf(fails ? "phy_threads_test: %d failure(s)\n" : "phy_threads_test: ok\n", fails.load());
    return fails ? 1 : 0;
}
