// Thread-loop emulation of csrc/fft_fast.cuh on the CPU (SF 7..12): the fast transform of the certified
// demodulator path against a double-precision DFT, plus the properties the kernel relies on and that no
// GPU tool checks on this pool (compute-sanitizer is closed, profiles/r02_sanitizer_unavailable.txt):
//   * every exchange writes each address of a window exactly once and reads only addresses it wrote
//     (no two threads race on an entry; reads and writes are separated by the kernel's barriers);
//   * the addresses stay inside the window's slot (Plan::pitch());
//   * every 64-bit access phase (16 consecutive lanes of the CTA) is bank-conflict free, for the raw
//     sample loads, the exchange stores and loads, and the middle-pass twiddle loads;
//   * the output bin map is a permutation of 0..N-1.
// Prints one JSON line per SF; exit 0 iff everything holds and the relative spectrum error is < 2e-6.
#include <algorithm>
#include <cmath>
#include <complex>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "fft_fast.cuh"
#include "fft_soa.cuh"
#include "fft_prune.cuh"

using ff::cf;

static int worst_conflict(const std::vector<long>& addr /* per CTA lane, complex entries, -1 = idle */)
{
    int worst = 1;
    for (size_t base = 0; base < addr.size(); base += 16) {
        std::vector<long> seen[16];
        for (size_t l = base; l < std::min(addr.size(), base + 16); ++l) {
            if (addr[l] < 0) continue;
            auto& v = seen[addr[l] & 15];
            if (std::find(v.begin(), v.end(), addr[l]) == v.end()) v.push_back(addr[l]);
        }
        for (auto& v : seen) worst = std::max(worst, (int)v.size());
    }
    return worst;
}

template <int SF>
static bool run(unsigned seed)
{
    using P = ff::Plan<SF>;
    constexpr int N = P::N, T = P::T, NP = P::NP;
    constexpr int CTA = 256;
    constexpr int G = CTA / T;                       // windows a CTA's threads hold at once
    const int PITCH = P::pitch();
    const double PI = 3.14159265358979323846;
    bool ok = true;
    int worst = 1;

    srand(seed);
    std::vector<std::complex<double>> x((size_t)G * N);
    for (auto& z : x) z = { (double)rand() / RAND_MAX - 0.5, (double)rand() / RAND_MAX - 0.5 };
    // one window is a pure tone: the case the kernel cares about
    for (int i = 0; i < N; ++i) x[i] = std::polar(1.0, 2 * PI * 37.0 * i / N);

    std::vector<cf> regs((size_t)G * T * 16);
    auto V = [&](int g, int tid) -> cf(&)[16] { return *reinterpret_cast<cf(*)[16]>(&regs[((size_t)g * T + tid) * 16]); };
    std::vector<cf> sm((size_t)G * PITCH);
    std::vector<int> written((size_t)G * PITCH);

    // raw loads: lane reads win[tid + T*c]
    for (int c = 0; c < 16; ++c) {
        std::vector<long> addr(CTA, -1);
        for (int g = 0; g < G; ++g)
            for (int tid = 0; tid < T; ++tid) addr[g * T + tid] = (long)g * PITCH + tid + T * c;
        worst = std::max(worst, worst_conflict(addr));
    }
    for (int g = 0; g < G; ++g)
        for (int tid = 0; tid < T; ++tid) {
            cf(&v)[16] = V(g, tid);
            for (int c = 0; c < 16; ++c) v[c] = ff::mk((float)x[(size_t)g * N + tid + T * c].real(), (float)x[(size_t)g * N + tid + T * c].imag());
        }

    for (int p = 0; p < NP; ++p) {
        for (int g = 0; g < G; ++g)
            for (int tid = 0; tid < T; ++tid) {
                cf(&v)[16] = V(g, tid);
                P::butterflies(p, v);
                if (p + 1 < NP) {                                   // twiddle W_M^{k * i_rest}
                    const int w = P::width(p), M = 1 << (P::lo(p) + w), ir = P::i_rest(p, tid);
                    for (int e = 0; e < 16; ++e) {
                        const int k = e & ((1 << w) - 1);
                        const double ang = -2 * PI * (double)((k * ir) % M) / M;
                        v[e] = ff::fcmul(v[e], ff::mk((float)cos(ang), (float)sin(ang)));
                    }
                    if (p == 0 && ir != tid) ok = false;
                }
            }
        if (p + 1 == NP) break;
        // twiddle-table loads of the middle pass: entry (k << lo) | i_rest
        if (p == 1)
            for (int k = 0; k < (1 << P::width(1)); ++k) {
                std::vector<long> addr(CTA, -1);
                for (int g = 0; g < G; ++g)
                    for (int tid = 0; tid < T; ++tid) addr[g * T + tid] = (k << P::lo(1)) | P::i_rest(1, tid);
                worst = std::max(worst, worst_conflict(addr));
            }
        std::fill(written.begin(), written.end(), 0);
        for (int e = 0; e < 16; ++e) {
            std::vector<long> addr(CTA, -1);
            for (int g = 0; g < G; ++g)
                for (int tid = 0; tid < T; ++tid) {
                    const int a = P::wr_base(p, tid) + P::wr_off(p, e);
                    if (a < 0 || a >= PITCH) { ok = false; continue; }
                    addr[g * T + tid] = (long)g * PITCH + a;
                    if (written[(size_t)g * PITCH + a]++) ok = false;           // two writers for one entry
                    sm[(size_t)g * PITCH + a] = V(g, tid)[e];
                }
            worst = std::max(worst, worst_conflict(addr));
        }
        for (int e = 0; e < 16; ++e) {
            std::vector<long> addr(CTA, -1);
            for (int g = 0; g < G; ++g)
                for (int tid = 0; tid < T; ++tid) {
                    const int a = P::rd_base(p, tid) + P::rd_off(p, e);
                    if (a < 0 || a >= PITCH || written[(size_t)g * PITCH + a] != 1) { ok = false; continue; }
                    written[(size_t)g * PITCH + a] = 2;                          // read exactly once as well
                    addr[g * T + tid] = (long)g * PITCH + a;
                    V(g, tid)[e] = sm[(size_t)g * PITCH + a];
                }
            worst = std::max(worst, worst_conflict(addr));
        }
    }

    // spectrum against the double DFT, bin map a permutation
    double max_rel = 0.0;
    for (int g = 0; g < G; ++g) {
        std::vector<std::complex<double>> X(N);
        std::vector<int> hit(N, 0);
        double peak = 0.0;
        for (int k = 0; k < N; ++k) {
            std::complex<double> acc = 0;
            for (int i = 0; i < N; ++i) acc += x[(size_t)g * N + i] * std::polar(1.0, -2 * PI * (double)((long)i * k % N) / N);
            X[k] = acc;
            peak = std::max(peak, std::abs(acc));
        }
        for (int tid = 0; tid < T; ++tid)
            for (int e = 0; e < 16; ++e) {
                const int bin = P::bin_base(tid) + P::bin_off(e);
                if (bin < 0 || bin >= N || hit[bin]++) { ok = false; continue; }
                const cf y = V(g, tid)[e];
                max_rel = std::max(max_rel, std::abs(std::complex<double>(y.x, y.y) - X[bin]) / peak);
            }
    }
    if (max_rel > 2e-6) ok = false;
    if (worst != 1) ok = false;
    std::printf("{\"sf\": %d, \"passes\": %d, \"pitch\": %d, \"max_rel_err\": %.3g, \"worst_bank_conflict\": %d, \"match\": %s}\n", SF, NP,
                PITCH, max_rel, worst, ok ? "true" : "false");
    return ok;
}

// The SF7 struct-of-arrays path (fft_soa.cuh): pass 0 on lanes (s_e, d_e), packed twiddle, pair exchange in two planes,
// packed pass 1 -- same checks as above for a CTA of 32 windows x 8 threads with the kernel's window pitch.
static bool run_soa7(unsigned seed)
{
    constexpr int N = 128, T = 8, CTA = 256, G = CTA / T, PITCH = ff::Plan<7>::pitch();
    const double PI = 3.14159265358979323846;
    bool ok = true;
    int worst = 1;
    srand(seed);
    std::vector<std::complex<double>> x((size_t)G * N);
    for (auto& z : x) z = { (double)rand() / RAND_MAX - 0.5, (double)rand() / RAND_MAX - 0.5 };
    for (int i = 0; i < N; ++i) x[i] = std::polar(1.0, 2 * PI * 37.0 * i / N);
    std::vector<fs::P2> R((size_t)G * T * 8), I((size_t)G * T * 8);
    std::vector<fs::P2> sm((size_t)G * PITCH);
    std::vector<int> written((size_t)G * PITCH, 0);
    const fs::P2 S = fs::mk2(0.70710678118654752440f, 0.70710678118654752440f), NS = fs::mk2(-0.70710678118654752440f, -0.70710678118654752440f);
    for (int g = 0; g < G; ++g)
        for (int t = 0; t < T; ++t) {
            cf y[16];
            for (int c = 0; c < 16; ++c) y[c] = ff::mk((float)x[(size_t)g * N + t + T * c].real(), (float)x[(size_t)g * N + t + T * c].imag());
            fs::P2 r[8], im[8];
            fs::pass0_r16(y, r, im, S, NS);
            for (int q = 0; q < 8; ++q) {
                const double a0 = -2 * PI * (double)((t * 2 * q) % N) / N, a1 = -2 * PI * (double)((t * (2 * q + 1)) % N) / N;
                fs::cmul(r[q], im[q], fs::mk2((float)cos(a0), (float)cos(a1)), fs::mk2((float)sin(a0), (float)sin(a1)));
                R[((size_t)g * T + t) * 8 + q] = r[q];
                I[((size_t)g * T + t) * 8 + q] = im[q];
            }
        }
    for (int plane = 0; plane < 2; ++plane)
        for (int q = 0; q < 8; ++q) {
            std::vector<long> addr(CTA, -1);
            for (int g = 0; g < G; ++g)
                for (int t = 0; t < T; ++t) {
                    const int a = plane * fs::SF7_PLANE + fs::sf7_wr(t, q);
                    if (a >= PITCH || written[(size_t)g * PITCH + a]++) ok = false;
                    addr[g * T + t] = (long)g * PITCH + a;
                    sm[(size_t)g * PITCH + a] = (plane ? I : R)[((size_t)g * T + t) * 8 + q];
                }
            worst = std::max(worst, worst_conflict(addr));
        }
    double max_rel = 0.0;
    for (int plane = 0; plane < 2; ++plane)
        for (int r = 0; r < 8; ++r) {
            std::vector<long> addr(CTA, -1);
            for (int g = 0; g < G; ++g)
                for (int t = 0; t < T; ++t) {
                    const int a = plane * fs::SF7_PLANE + fs::sf7_rd(t, r);
                    if (a >= PITCH || written[(size_t)g * PITCH + a] != 1) ok = false;
                    written[(size_t)g * PITCH + a] = 2;
                    addr[g * T + t] = (long)g * PITCH + a;
                    (plane ? I : R)[((size_t)g * T + t) * 8 + r] = sm[(size_t)g * PITCH + a];
                }
            worst = std::max(worst, worst_conflict(addr));
        }
    for (int g = 0; g < G; ++g) {
        std::vector<std::complex<double>> X(N);
        std::vector<int> hit(N, 0);
        double peak = 0.0;
        for (int k = 0; k < N; ++k) {
            std::complex<double> acc = 0;
            for (int i = 0; i < N; ++i) acc += x[(size_t)g * N + i] * std::polar(1.0, -2 * PI * (double)((long)i * k % N) / N);
            X[k] = acc;
            peak = std::max(peak, std::abs(acc));
        }
        for (int t = 0; t < T; ++t) {
            fs::P2 r[8], im[8];
            for (int q = 0; q < 8; ++q) { r[q] = R[((size_t)g * T + t) * 8 + q]; im[q] = I[((size_t)g * T + t) * 8 + q]; }
            fs::dft8(r, im, S, NS);
            for (int q = 0; q < 8; ++q)
                for (int lane = 0; lane < 2; ++lane) {
                    const int bin = fs::sf7_bin(t, q, lane);
                    if (bin < 0 || bin >= N || hit[bin]++) { ok = false; continue; }
                    const std::complex<double> y(lane ? fs::hi(r[q]) : fs::lo(r[q]), lane ? fs::hi(im[q]) : fs::lo(im[q]));
                    max_rel = std::max(max_rel, std::abs(y - X[bin]) / peak);
                }
        }
    }
    if (max_rel > 2e-6 || worst != 1) ok = false;
    std::printf("{\"sf\": 7, \"layout\": \"soa pairs\", \"pitch\": %d, \"max_rel_err\": %.3g, \"worst_bank_conflict\": %d, \"match\": %s}\n", PITCH,
                max_rel, worst, ok ? "true" : "false");
    return ok;
}

// The pruned transform (fft_prune.cuh): for a random bin class k0, the value chain of demod_cert.cu's pruned step --
// pass 0 output k0 (or, geometric variant, sum_c r_c h^c on the raw samples with the rotation folded into h), pass-0
// twiddle, [T > 32: R-point sum across warps + W_T twiddle], radix-2 lane transform with the stage table -- replayed
// lane by lane with shuffles modelled as array reads.  Every bin of the class must match the double DFT of the
// rotated samples and the bin map must cover the class exactly once.
template <int SF>
static bool run_prune(unsigned seed, bool geometric)
{
    using P = ff::Plan<SF>;
    using Q = fp::Prune<SF>;
    constexpr int N = P::N, T = P::T, W = Q::W, R = Q::R, LW = Q::LW;
    const double PI = 3.14159265358979323846;
    bool ok = true;
    srand(seed);
    const double rate = -2 * PI * 0.37 / N;                          // the walk's rotation exp(j rate i)
    std::vector<std::complex<double>> x(N), y(N);
    for (int i = 0; i < N; ++i) {
        x[i] = { (double)rand() / RAND_MAX - 0.5, (double)rand() / RAND_MAX - 0.5 };
        y[i] = x[i] * std::polar(1.0, rate * i);
    }
    std::vector<cf> tws((size_t)(LW > 1 ? LW - 1 : 1) * W);
    for (int s = 0; s + 1 < LW; ++s)
        for (int l = 0; l < W; ++l) {
            const double a = -2 * PI * Q::stage_exp(s, l) / W;
            tws[(size_t)s * W + l] = ff::mk((float)cos(a), (float)sin(a));
        }
    double max_rel = 0.0;
    for (int k0 = 0; k0 < 16; ++k0) {
        std::vector<cf> val(T);
        for (int tid = 0; tid < T; ++tid) {
            cf v[16];
            cf pv;
            if (geometric) {
                // h = G W16^k0, G = exp(j rate T); powers as the kernel builds them (float products of accurate factors)
                cf hp[4];
                for (int m = 0; m < 4; ++m) {
                    const double ga = rate * (double)(T << m), wa = -2 * PI * (double)((k0 << m) & 15) / 16.0;
                    hp[m] = ff::fcmul(ff::mk((float)cos(ga), (float)sin(ga)), ff::mk((float)cos(wa), (float)sin(wa)));
                }
                for (int c = 0; c < 16; ++c) v[c] = ff::mk((float)x[tid + T * c].real(), (float)x[tid + T * c].imag());
                auto cma = [](cf acc, cf a, cf w) {
                    return ff::mk(ff::fma_(-a.y, w.y, ff::fma_(a.x, w.x, acc.x)), ff::fma_(a.y, w.x, ff::fma_(a.x, w.y, acc.y)));
                };
                for (int c = 0; c < 8; ++c) v[c] = cma(v[c], v[c + 8], hp[3]);
                for (int c = 0; c < 4; ++c) v[c] = cma(v[c], v[c + 4], hp[2]);
                for (int c = 0; c < 2; ++c) v[c] = cma(v[c], v[c + 2], hp[1]);
                pv = cma(v[0], v[1], hp[0]);
                pv = ff::fcmul(pv, ff::mk((float)cos(rate * tid), (float)sin(rate * tid)));      // F_tid (scale 1)
            } else {
                for (int c = 0; c < 16; ++c) v[c] = ff::mk((float)y[tid + T * c].real(), (float)y[tid + T * c].imag());
                ff::dft16(v);
                pv = v[k0];
            }
            const double a = -2 * PI * (double)P::tw0_exp(tid, k0) / N;
            val[tid] = ff::fcmul(pv, ff::mk((float)cos(a), (float)sin(a)));
        }
        // cross-warp stage
        std::vector<cf> c(T);
        for (int tid = 0; tid < T; ++tid) {
            if (R == 1) { c[tid] = val[tid]; continue; }
            const int ka = tid >> 5, l = tid & 31;
            cf acc = val[l];
            for (int w = 1; w < R; ++w) {
                const double a = -2 * PI * (double)Q::sum_exp(ka, w) / R;
                acc = ff::add(acc, ff::fcmul(val[l + 32 * w], ff::mk((float)cos(a), (float)sin(a))));
            }
            const double a = -2 * PI * (double)Q::cross_exp(ka, l) / T;
            c[tid] = ff::fcmul(acc, ff::mk((float)cos(a), (float)sin(a)));
        }
        // lane transform: stage s pairs lane l with l ^ h
        for (int s = 0; s < LW; ++s) {
            const int h = Q::half(s);
            std::vector<cf> nx(T);
            for (int tid = 0; tid < T; ++tid) {
                const int l = tid & (W - 1);
                const cf o = c[tid ^ h];
                const bool upper = (l & h) != 0;
                const cf own = upper ? ff::mk(-c[tid].x, -c[tid].y) : c[tid];
                cf r = ff::add(o, own);
                if (s != LW - 1) r = ff::fcmul(r, tws[(size_t)s * W + l]);
                nx[tid] = r;
            }
            c.swap(nx);
        }
        // compare with the double DFT of the rotated samples; the map must hit every bin of the class once
        std::vector<int> hit(N, 0);
        double peak = 0.0;
        std::vector<std::complex<double>> X(N);
        for (int k = k0; k < N; k += 16) {
            std::complex<double> acc = 0;
            for (int i = 0; i < N; ++i) acc += y[i] * std::polar(1.0, -2 * PI * (double)((long)i * k % N) / N);
            X[k] = acc;
        }
        for (int i = 0; i < N; ++i) peak += std::abs(y[i]);
        for (int tid = 0; tid < T; ++tid) {
            const int bin = Q::bin(k0, tid >> 5 & (R - 1), tid & (W - 1));
            if (bin < 0 || bin >= N || (bin & 15) != k0 || hit[bin]++) { ok = false; continue; }
            max_rel = std::max(max_rel, std::abs(std::complex<double>(c[tid].x, c[tid].y) - X[bin]) / peak);
        }
        for (int k = k0; k < N; k += 16) if (hit[k] != 1) ok = false;
    }
    // per-bin error against sum |y_i| (the certificate's unit): the budget for "our side" is 96 * 2^-24 = 5.7e-6
    if (max_rel > 2e-6) ok = false;
    std::printf("{\"sf\": %d, \"layout\": \"pruned %s\", \"lanes\": %d, \"warps\": %d, \"max_err_over_sum_abs\": %.3g, \"worst_bank_conflict\": 1, \"match\": %s}\n",
                SF, geometric ? "geometric" : "pass0", W, R, max_rel, ok ? "true" : "false");
    return ok;
}

// The streaming form of the pruned transform (demod_lean.cu, SF10-12): a window passes one warp in R = N / 512 chunks of
// 512 samples; lane l keeps R accumulators (thread t = l + 32 j of the layout above), every chunk adds 16 / R terms to each
// by halving levels and joins with h^(c0) built from the table's powers of two; then the twiddles exp(j t theta), an
// R-point DFT in registers, the W_T cross twiddles and R lane transforms.  Same check: every bin of every class against
// the double DFT of the rotated samples, each class covered exactly once.
template <int SF>
static bool run_lean(unsigned seed)
{
    using Q = fp::Prune<SF>;
    constexpr int N = 1 << SF, T = N / 16, R = N / 512, CP = 16 / R;
    const double PI = 3.14159265358979323846;
    bool ok = true;
    srand(seed);
    const double rate = -2 * PI * 0.31 / N;
    std::vector<std::complex<double>> x(N), y(N);
    for (int i = 0; i < N; ++i) {
        x[i] = { (double)rand() / RAND_MAX - 0.5, (double)rand() / RAND_MAX - 0.5 };
        y[i] = x[i] * std::polar(1.0, rate * i);
    }
    auto fc = [](double a) { return ff::mk((float)cos(a), (float)sin(a)); };
    auto cma = [](cf acc, cf a, cf w) {
        return ff::mk(ff::fma_(-a.y, w.y, ff::fma_(a.x, w.x, acc.x)), ff::fma_(a.y, w.x, ff::fma_(a.x, w.y, acc.y)));
    };
    std::vector<cf> tws(4 * 32);
    for (int s = 0; s < 4; ++s) for (int l = 0; l < 32; ++l) tws[s * 32 + l] = fc(-2 * PI * Q::stage_exp(s, l) / 32.0);
    double max_rel = 0.0, sum_abs = 0.0;
    for (int i = 0; i < N; ++i) sum_abs += std::abs(y[i]);
    for (int k0 = 0; k0 < 16; ++k0) {
        cf hp[4];
        for (int m = 0; m < 4; ++m) hp[m] = ff::fcmul(fc(rate * (double)(T << m)), fc(-2 * PI * (double)((k0 << m) & 15) / 16.0));
        std::vector<cf> lanes((size_t)32 * R);                     // [l][ka] after the R-point DFT and the cross twiddle
        for (int l = 0; l < 32; ++l) {
            cf acc[8];
            for (int q = 0; q < R; ++q) {
                cf v[16];
                for (int m = 0; m < 16; ++m) { const auto z = x[512 * q + l + 32 * m]; v[m] = ff::mk((float)z.real(), (float)z.imag()); }
                if (CP >= 8) for (int i = 0; i < 4 * R; ++i) v[i] = cma(v[i], v[i + 4 * R], hp[2]);
                if (CP >= 4) for (int i = 0; i < 2 * R; ++i) v[i] = cma(v[i], v[i + 2 * R], hp[1]);
                for (int i = 0; i < R; ++i) v[i] = cma(v[i], v[i + R], hp[0]);
                if (q == 0) { for (int j = 0; j < R; ++j) acc[j] = v[j]; continue; }
                const int e = q * CP;
                cf hq = ff::mk(1.0f, 0.0f); bool first = true;
                if (e & 2) { hq = hp[1]; first = false; }
                if (e & 4) { hq = first ? hp[2] : ff::fcmul(hq, hp[2]); first = false; }
                if (e & 8) { hq = first ? hp[3] : ff::fcmul(hq, hp[3]); first = false; }
                for (int j = 0; j < R; ++j) acc[j] = cma(acc[j], v[j], hq);
            }
            const cf el = ff::fcmul(fc(rate * l), fc(-2 * PI * (double)(l * k0) / N));
            for (int j = 0; j < R; ++j) {
                cf u = ff::fcmul(acc[j], el);
                if (j > 0) u = ff::fcmul(u, ff::fcmul(fc(rate * 32.0 * j), fc(-2 * PI * (double)((32 * j * k0) % N) / N)));
                acc[j] = u;
            }
            if (R == 2) ff::dft2(acc[0], acc[1]);
            if (R == 4) ff::dft4(acc[0], acc[1], acc[2], acc[3]);
            if (R == 8) ff::dft8(acc[0], acc[1], acc[2], acc[3], acc[4], acc[5], acc[6], acc[7]);
            for (int ka = 0; ka < R; ++ka)
                lanes[(size_t)l * R + ka] = ka ? ff::fcmul(acc[ka], fc(-2 * PI * (double)Q::cross_exp(ka, l) / T)) : acc[ka];
        }
        for (int s = 0; s < 5; ++s) {
            const int h = 16 >> s;
            std::vector<cf> nx(lanes.size());
            for (int l = 0; l < 32; ++l)
                for (int ka = 0; ka < R; ++ka) {
                    const cf own0 = lanes[(size_t)l * R + ka], o = lanes[(size_t)(l ^ h) * R + ka];
                    const cf own = (l & h) ? ff::mk(-own0.x, -own0.y) : own0;
                    cf r = ff::add(o, own);
                    if (s != 4) r = ff::fcmul(r, tws[s * 32 + l]);
                    nx[(size_t)l * R + ka] = r;
                }
            lanes.swap(nx);
        }
        std::vector<int> hit(N, 0);
        for (int l = 0; l < 32; ++l)
            for (int ka = 0; ka < R; ++ka) {
                const int bin = Q::bin(k0, ka, l);
                if (bin < 0 || bin >= N || (bin & 15) != k0 || hit[bin]++) { ok = false; continue; }
                std::complex<double> acc = 0;
                for (int i = 0; i < N; ++i) acc += y[i] * std::polar(1.0, -2 * PI * (double)((long)i * bin % N) / N);
                const cf g = lanes[(size_t)l * R + ka];
                max_rel = std::max(max_rel, std::abs(std::complex<double>(g.x, g.y) - acc) / sum_abs);
            }
        for (int k = k0; k < N; k += 16) if (hit[k] != 1) ok = false;
    }
    if (max_rel > 2e-6) ok = false;
    std::printf("{\"sf\": %d, \"layout\": \"pruned streaming\", \"lanes\": 32, \"accumulators\": %d, \"max_err_over_sum_abs\": %.3g, \"worst_bank_conflict\": 1, \"match\": %s}\n",
                SF, R, max_rel, ok ? "true" : "false");
    return ok;
}

int main()
{
    bool ok = true;
    ok &= run_lean<10>(21);
    ok &= run_lean<11>(22);
    ok &= run_lean<12>(23);
    ok &= run_prune<7>(11, true);
    ok &= run_prune<8>(12, true);
    ok &= run_prune<9>(13, true);
    ok &= run_prune<7>(14, false);
    ok &= run_prune<10>(15, false);
    ok &= run_prune<11>(16, false);
    ok &= run_prune<12>(17, false);
    ok &= run_prune<12>(18, true);
    ok &= run_soa7(7);
    ok &= run<7>(1);
    ok &= run<8>(2);
    ok &= run<9>(3);
    ok &= run<10>(4);
    ok &= run<11>(5);
    ok &= run<12>(6);
    return ok ? 0 : 1;
}
