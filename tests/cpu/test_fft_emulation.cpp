// Thread-loop emulation of csrc/fft_exact.cuh on the CPU: checks, for SF 2..12, that the
// register-pass FFT yields spectra bit-identical to the oracle's kissfft restatement, and
// reports the worst shared-memory bank-conflict degree of every exchange (64-bit accesses,
// 16 lanes per phase) for a 256-thread CTA laid out like the kernels lay it out.
// Usage: test_fft_emulation  -> prints one JSON line per SF, exit 0 iff all spectra match.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <algorithm>
#include "fft_exact.cuh"
#include "../../oracle/lora_oracle.h"

using fx::cf;

static int worst_conflict(const std::vector<int>& addr /* one per CTA thread, float2 units, -1 = inactive */)
{
    int worst = 1;
    for (size_t base = 0; base < addr.size(); base += 16) {
        int cnt[16] = {0};
        std::vector<int> seen[16];
        for (size_t l = base; l < std::min(addr.size(), base + 16); ++l) {
            if (addr[l] < 0) continue;
            int b = addr[l] & 15;
            if (std::find(seen[b].begin(), seen[b].end(), addr[l]) == seen[b].end()) {
                seen[b].push_back(addr[l]);
                ++cnt[b];
            }
        }
        for (int b = 0; b < 16; ++b) worst = std::max(worst, cnt[b]);
    }
    return worst;
}

template <int SF, bool FULL>
static bool run(unsigned seed, int* conflicts_out)
{
    using P = fx::Plan<SF>;
    constexpr int N = P::N, T = P::T, R = P::R;
    constexpr int CTA = 256;
    constexpr int G = CTA / T > 0 ? CTA / T : 1;       // transforms per CTA
    std::vector<ocplx> tw(N);
    oracle_twiddles(SF, tw.data());
    std::vector<cf> ct(P::compact_size());
    for (int j = 0; j < P::compact_size(); ++j) ct[j] = P::compact_entry(j, reinterpret_cast<const cf*>(tw.data()));

    srand(seed);
    bool ok = true;
    int worst = 1;
    std::vector<ocplx> in(G * N), ref(G * N);
    for (auto& z : in) { z.re = (float)rand() / RAND_MAX - 0.5f; z.im = (float)rand() / RAND_MAX - 0.5f; }
    for (int g = 0; g < G; ++g) oracle_fft(SF, in.data() + g * N, ref.data() + g * N);

    std::vector<cf> regs((size_t)G * T * R);
    std::vector<cf> smA((size_t)G * N), smB((size_t)G * N);
    auto V = [&](int g, int tid) -> cf(&)[R] { return *reinterpret_cast<cf(*)[R]>(&regs[((size_t)g * T + tid) * R]); };

    for (int g = 0; g < G; ++g)
        for (int tid = 0; tid < T; ++tid) {
            cf(&v)[R] = V(g, tid);
            for (int c = 0; c < R; ++c) { v[c].x = in[g * N + tid + T * c].re; v[c].y = in[g * N + tid + T * c].im; }
            P::template pass0<FULL>(v, reinterpret_cast<const cf*>(tw.data()));
        }
    cf* cur = smA.data();
    cf* nxt = smB.data();
    for (int p = 1; p < P::NP; ++p) {
        // store results of pass p-1
        for (int r = 0; r < R; ++r) {
            std::vector<int> addr(G * T, -1);
            for (int g = 0; g < G; ++g)
                for (int tid = 0; tid < T; ++tid) {
                    int o = (p == 1) ? P::pos_after_pass0(tid, r) : P::pos_in_pass(p - 1, tid, r);
                    int a = g * N + P::phys(o);
                    addr[g * T + tid] = a;
                    cur[a] = V(g, tid)[r];
                }
            worst = std::max(worst, worst_conflict(addr));
        }
        for (int r = 0; r < R; ++r) {
            std::vector<int> addr(G * T, -1);
            for (int g = 0; g < G; ++g)
                for (int tid = 0; tid < T; ++tid) {
                    int a = g * N + P::phys(P::pos_in_pass(p, tid, r));
                    addr[g * T + tid] = a;
                    V(g, tid)[r] = cur[a];
                }
            worst = std::max(worst, worst_conflict(addr));
        }
        for (int g = 0; g < G; ++g)
            for (int tid = 0; tid < T; ++tid) P::pass(p, V(g, tid), tid, ct.data());
        std::swap(cur, nxt);
    }
    for (int g = 0; g < G; ++g)
        for (int tid = 0; tid < T; ++tid)
            for (int r = 0; r < R; ++r) {
                int bin = P::bin_of_reg(tid, r);
                cf a = V(g, tid)[r];
                ocplx b = ref[g * N + bin];
                if (FULL) { if (memcmp(&a, &b, 8) != 0) ok = false; }
                else if (!(a.x == b.re && a.y == b.im)) ok = false;   // zero signs may differ
            }
    *conflicts_out = worst;
    return ok;
}

template <int SF>
static bool both()
{
    int c1 = 0, c2 = 0;
    bool a = run<SF, true>(1000 + SF, &c1);
    bool b = run<SF, false>(2000 + SF, &c2);
    printf("{\"sf\": %d, \"T\": %d, \"passes\": %d, \"w0\": %d, \"exact_full\": %s, \"exact_fast\": %s, \"worst_bank_conflict\": %d}\n",
           SF, fx::Plan<SF>::T, fx::Plan<SF>::NP, fx::Plan<SF>::W0, a ? "true" : "false", b ? "true" : "false", c1);
    return a && b;
}

int main()
{
    bool ok = true;
    ok &= both<2>(); ok &= both<3>(); ok &= both<4>(); ok &= both<5>(); ok &= both<6>(); ok &= both<7>();
    ok &= both<8>(); ok &= both<9>(); ok &= both<10>(); ok &= both<11>(); ok &= both<12>();
    return ok ? 0 : 1;
}
