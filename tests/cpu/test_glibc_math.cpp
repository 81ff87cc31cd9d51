// Host-side check of csrc/glibc_math.cuh against the real glibc (the libm the reference
// links).  Usage: test_glibc_math [stride]   -- walks the float bit patterns with the
// given stride (1 = exhaustive, ~2^32 arguments per function) and prints mismatch counts.
// Exit code 0 iff every function stays within its allowed mismatch budget:
//   sinf/cosf/log10f : double-precision polynomial, glibc's x86-64 build uses an FMA
//                      ifunc variant whose contraction choices we do not copy -> a
//                      mismatch needs the double result within ~1e-16 of a float rounding
//                      boundary (expected rate ~1e-9); budget 1e-7 of arguments.
//   hypotf/atan2f    : must be exact (0 mismatches).
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>
#include <atomic>
#include "glibc_math.cuh"

static inline bool same(float a, float b)
{
    if (std::isnan(a) && std::isnan(b)) return true;
    return gm::f2u(a) == gm::f2u(b);
}

int main(int argc, char** argv)
{
    const uint64_t stride = argc > 1 ? strtoull(argv[1], nullptr, 10) : 4099;
    const int T = (int)std::thread::hardware_concurrency();
    std::atomic<uint64_t> bad_sin{0}, bad_cos{0}, bad_sc{0}, bad_log{0}, bad_hyp{0}, bad_at2{0}, bad_atan{0}, total{0};
    std::vector<std::thread> pool;
    for (int t = 0; t < T; ++t) pool.emplace_back([&, t] {
        uint64_t n = 0, bs = 0, bc = 0, bsc = 0, bl = 0, bh = 0, ba = 0, bat = 0;
        uint32_t lcg = 12345u + 977u * (uint32_t)t;
        for (uint64_t u = (uint64_t)t * stride; u < (1ULL << 32); u += stride * (uint64_t)T) {
            const float x = gm::u2f((uint32_t)u);
            ++n;
            if (!same(gm::sinf_g(x), sinf(x))) ++bs;
            if (!same(gm::cosf_g(x), cosf(x))) ++bc;
            float s, c; gm::sincosf_g(x, &s, &c);
            if (!same(s, sinf(x)) || !same(c, cosf(x))) ++bsc;
            if (!same(gm::log10f_g(x), log10f(x))) ++bl;
            if (!same(gm::atanf_g(x), atanf(x))) ++bat;
            lcg = lcg * 1664525u + 1013904223u;
            // second operand: same exponent neighbourhood most of the time, arbitrary sometimes
            uint32_t yb = (lcg & 0x100) ? lcg : (((uint32_t)u & 0x7f800000u) + ((int32_t)(lcg >> 27) - 16) * 0x00800000u) | (lcg & 0x807fffffu);
            const float y = gm::u2f(yb);
            if (!same(gm::hypotf_g(x, y), hypotf(x, y))) ++bh;
            if (!same(gm::atan2f_g(y, x), atan2f(y, x))) ++ba;
            if (!same(gm::atan2f_g(x, y), atan2f(x, y))) ++ba;
        }
        total += n; bad_sin += bs; bad_cos += bc; bad_sc += bsc; bad_log += bl; bad_hyp += bh; bad_at2 += ba; bad_atan += bat;
    });
    for (auto& th : pool) th.join();
    const uint64_t n = total.load();
    printf("{\"args\": %llu, \"sinf\": %llu, \"cosf\": %llu, \"sincosf\": %llu, \"log10f\": %llu, \"hypotf\": %llu, \"atan2f\": %llu, \"atanf\": %llu}\n",
           (unsigned long long)n, (unsigned long long)bad_sin.load(), (unsigned long long)bad_cos.load(),
           (unsigned long long)bad_sc.load(), (unsigned long long)bad_log.load(), (unsigned long long)bad_hyp.load(),
           (unsigned long long)bad_at2.load(), (unsigned long long)bad_atan.load());
    const double budget = 1e-7 * (double)n + 1.0;
    bool ok = bad_hyp.load() == 0 && bad_at2.load() == 0 && bad_atan.load() == 0 &&
              (double)bad_sin.load() <= budget && (double)bad_cos.load() <= budget &&
              (double)bad_sc.load() <= 2 * budget && (double)bad_log.load() <= budget;
    return ok ? 0 : 1;
}
