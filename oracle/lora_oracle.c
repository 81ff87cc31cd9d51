/* oracle/lora_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Plain-C CPU restatement of the reference's LoRa PHY hot path (see lora_oracle.h
 * for scope and parity status: PINNED against oracle/_ref and the reference's
 * known-answer vectors by tests/test_oracle.py).
 *
 * Build: gcc -std=c11 -O2 -ffp-contract=off (the reference binary contains no
 * FMA; every float op below rounds once, like the reference's SSE code).
 * libm calls are the same glibc entry points the reference resolves to:
 * std::cos(float)->cosf, std::abs(complex)->hypotf, std::arg->atan2f,
 * std::exp(complex)->cexpf, std::polar -> (rho*cosf, rho*sinf).
 */
#define _GNU_SOURCE
#include "lora_oracle.h"

#include <complex.h>
#include <errno.h>
#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

static const float O_PI = 3.14159265358979323846f; /* phy.hpp:20 */

/* ---- complex helpers: std::complex<float> semantics, one rounding per op ---- */
static inline ocplx c_mk(float re, float im) { ocplx z = { re, im }; return z; }
static inline ocplx c_add(ocplx a, ocplx b) { return c_mk(a.re + b.re, a.im + b.im); }
static inline ocplx c_sub(ocplx a, ocplx b) { return c_mk(a.re - b.re, a.im - b.im); }
static inline ocplx c_mul(ocplx a, ocplx b)
{
    /* (ac - bd, ad + bc): what g++ -O2 emits for finite operands */
    return c_mk(a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re);
}
static inline ocplx c_scale(ocplx a, float s) { return c_mk(a.re * s, a.im * s); }

/* ---- FFT ------------------------------------------------------------------ */

/* kissfft.hh:24-29: phinc is a float, i*phinc is rounded to float, exp() is cexpf */
void oracle_twiddles(unsigned sf, ocplx* tw)
{
    const int nfft = 1 << sf;
    const float phinc = (float)(-2.0 * acos(-1.0) / nfft);
    for (int i = 0; i < nfft; ++i) {
        float _Complex e = cexpf(CMPLXF(0.0f, (float)i * phinc));
        tw[i] = c_mk(crealf(e), cimagf(e));
    }
}

/* kissfft.hh:81-97: factor 4 while divisible, then 2.  Returned outermost first. */
static int plan_radices(unsigned sf, int* radix, int* rem)
{
    int n = 1 << sf, stages = 0;
    while (n > 1) {
        int p = (n % 4 == 0) ? 4 : 2;
        n /= p;
        radix[stages] = p;
        rem[stages] = n;
        ++stages;
    }
    return stages;
}

/* kissfft.hh:106-185 restated iteratively: digit-reversed gather (what the m==1
 * leaf copy of the recursion amounts to), then butterflies from the innermost
 * stage outwards.  Per-element arithmetic and operand order follow kf_bfly2
 * (:155-162) and kf_bfly4 (:164-185, forward transform). */
void oracle_fft(unsigned sf, const ocplx* in, ocplx* out)
{
    const int N = 1 << sf;
    int radix[16], rem[16];
    const int stages = plan_radices(sf, radix, rem);
    static _Thread_local ocplx tw[ORACLE_MAX_N];
    static _Thread_local unsigned tw_sf = 0xffffffffu;
    if (tw_sf != sf) { oracle_twiddles(sf, tw); tw_sf = sf; }

    if (N == 1) { out[0] = in[0]; return; }

    for (int o = 0; o < N; ++o) {
        /* position o = sum q_j * rem_j  <->  input index sum q_j * prod_{l<j} radix_l */
        int r = o, idx = 0, stride = 1;
        for (int j = 0; j < stages; ++j) {
            int q = r / rem[j];
            r -= q * rem[j];
            idx += q * stride;
            stride *= radix[j];
        }
        out[o] = in[idx];
    }

    for (int j = stages - 1; j >= 0; --j) {
        const int p = radix[j], m = rem[j];
        int fstride = 1;
        for (int l = 0; l < j; ++l) fstride *= radix[l];
        for (int base = 0; base < N; base += p * m) {
            ocplx* F = out + base;
            if (p == 2) {
                for (int k = 0; k < m; ++k) {
                    ocplx t = c_mul(F[m + k], tw[k * fstride]);
                    F[m + k] = c_sub(F[k], t);
                    F[k] = c_add(F[k], t);
                }
            } else {
                for (int k = 0; k < m; ++k) {
                    ocplx s0 = c_mul(F[k + m], tw[k * fstride]);
                    ocplx s1 = c_mul(F[k + 2 * m], tw[k * fstride * 2]);
                    ocplx s2 = c_mul(F[k + 3 * m], tw[k * fstride * 3]);
                    ocplx s5 = c_sub(F[k], s1);
                    F[k] = c_add(F[k], s1);
                    ocplx s3 = c_add(s0, s2);
                    ocplx s4 = c_sub(s0, s2);
                    s4 = c_mk(s4.im, -s4.re); /* forward: multiply by -j */
                    F[k + 2 * m] = c_sub(F[k], s3);
                    F[k] = c_add(F[k], s3);
                    F[k + m] = c_add(s5, s4);
                    F[k + 3 * m] = c_sub(s5, s4);
                }
            }
        }
    }
}

/* LoRaDetector.hpp:39-74 */
size_t oracle_detect(unsigned sf, const ocplx* in, ocplx* spectrum,
                     float* power, float* power_avg, float* findex)
{
    const size_t N = (size_t)1 << sf;
    oracle_fft(sf, in, spectrum);
    size_t max_index = 0;
    float max_value = 0.0f;
    double total = 0.0;
    for (size_t i = 0; i < N; ++i) {
        float re = spectrum[i].re, im = spectrum[i].im;
        float mag2 = re * re + im * im;
        total += mag2;
        if (mag2 > max_value) { max_index = i; max_value = mag2; }
    }
    const float power_scale = (float)(20 * log10((double)N)); /* :29 */
    const float noise = sqrtf((float)(total - max_value));
    const float fundamental = sqrtf(max_value);
    if (power_avg) *power_avg = 20 * log10f(noise) - power_scale;
    if (power) *power = 20 * log10f(fundamental) - power_scale;
    ocplx l = spectrum[max_index > 0 ? max_index - 1 : N - 1];
    ocplx r = spectrum[max_index < N - 1 ? max_index + 1 : 0];
    float left = hypotf(l.re, l.im), right = hypotf(r.re, r.im);
    const double demon = (2.0 * fundamental) - right - left;
    if (findex) *findex = (demon == 0.0) ? 0.0f : (float)(0.5 * (right - left) / demon);
    return max_index;
}

/* LoRaDemod.cpp:18-27 / phy.cpp:39-50 */
void oracle_window(unsigned sf, int hann, float* w)
{
    const size_t N = (size_t)1 << sf;
    for (size_t i = 0; i < N; ++i)
        w[i] = hann ? 0.5f - 0.5f * cosf(2.0f * O_PI * (float)i / ((float)N - 1.0f)) : 1.0f;
}

/* Shared by the legacy (LoRaDemod.cpp:80-136) and high-level (phy.cpp:93-147)
 * estimators.  index_tiebreak selects LoRaDemod.cpp:102 vs phy.cpp:116. */
static void estimate_core(unsigned sf, unsigned osr, int hann, const float* w,
                          const ocplx* x, size_t n_est, int index_tiebreak,
                          float* cfo, float* time_offset)
{
    const size_t N = (size_t)1 << sf, step = N * osr;
    static _Thread_local ocplx fin[ORACLE_MAX_N], fout[ORACLE_MAX_N];
    float sum_index = 0.0f, phase_diff = 0.0f, prev_phase = 0.0f;
    int have_prev = 0;
    unsigned sum_t = 0;
    for (size_t s = 0; s < n_est; ++s) {
        const ocplx* sym = x + s * step;
        float best_p = -1e30f, best_fi = 0.0f;
        size_t best_idx = 0;
        unsigned best_t = 0;
        ocplx best_bin = c_mk(0.0f, 0.0f);
        for (unsigned t = 0; t < osr; ++t) {
            for (size_t i = 0; i < N; ++i) {
                ocplx v = sym[t + i * osr];
                if (hann) v = c_scale(v, w[i]);
                fin[i] = v;
            }
            float p, pav, fi;
            size_t idx = oracle_detect(sf, fin, fout, &p, &pav, &fi);
            if (p > best_p || (index_tiebreak && p == best_p && idx < best_idx)) {
                best_p = p; best_idx = idx; best_fi = fi; best_t = t; best_bin = fout[idx];
            }
        }
        sum_t += best_t;
        sum_index += (float)best_idx + best_fi;
        float phase = atan2f(best_bin.im, best_bin.re);
        if (have_prev) {
            float d = phase - prev_phase;
            while (d > O_PI) d -= 2.0f * O_PI;
            while (d < -O_PI) d += 2.0f * O_PI;
            phase_diff += d;
        }
        prev_phase = phase;
        have_prev = 1;
    }
    float avg_index = sum_index / (float)n_est;
    float cfo_coarse = avg_index / (float)N;
    float cfo_fine = 0.0f;
    if (n_est > 1)
        cfo_fine = (phase_diff / (float)(n_est - 1)) / (2.0f * O_PI * (float)N);
    *cfo = cfo_coarse + cfo_fine;
    float frac = avg_index - floorf(avg_index + 0.5f);
    float avg_t = (float)sum_t / (float)n_est;
    *time_offset = avg_t - frac * (float)N * (float)osr;
}

/* Data loop shared by LoRaDemod.cpp:138-176 and phy.cpp:201-238.  `down` != NULL
 * applies the high-level path's downchirp multiply first (phy.cpp:224-225). */
static void demod_loop(unsigned sf, unsigned osr, int hann, const float* w, const ocplx* down,
                       const ocplx* x, size_t sample_count, float cfo, float time_offset,
                       uint16_t* idx_out /* one per symbol, incl. sync */)
{
    const size_t N = (size_t)1 << sf, step = N * osr, total = sample_count / step;
    static _Thread_local ocplx fin[ORACLE_MAX_N], fout[ORACLE_MAX_N];
    int t_off = (int)roundf(time_offset);
    float rate = -2.0f * O_PI * cfo / (float)N;
    for (size_t s = 0; s < total; ++s) {
        size_t base = s * step;
        if (t_off > 0) {
            if (base + (size_t)t_off + step <= sample_count) base += (size_t)t_off;
        } else if (t_off < 0) {
            size_t off = (size_t)(-t_off);
            if (off <= base) base -= off;
        }
        const ocplx* sym = x + base;
        float start = rate * ((float)(s * N) + (float)t_off / (float)osr);
        for (size_t i = 0; i < N; ++i) {
            float ph = start + rate * (float)i;
            ocplx rot = c_mk(cosf(ph), sinf(ph));
            ocplx v = sym[i * osr];
            if (down) v = c_mul(v, down[i]);
            v = c_mul(v, rot);
            if (hann) v = c_scale(v, w[i]);
            fin[i] = v;
        }
        float p, pav, fi;
        idx_out[s] = (uint16_t)oracle_detect(sf, fin, fout, &p, &pav, &fi);
    }
}

static uint8_t sync_from(unsigned sf, uint16_t sw0, uint16_t sw1)
{
    unsigned shift = sf > 4 ? sf - 4 : 0; /* LoRaDemod.cpp:180-189 */
    uint8_t hi = (uint8_t)(sw0 >> shift) & 0x0f, lo = (uint8_t)(sw1 >> shift) & 0x0f;
    return (uint8_t)((hi << 4) | lo);
}

/* LoRaDemod.cpp:50-197 */
long oracle_lora_demodulate(unsigned sf, int window, int with_scratch,
                            const ocplx* samples, size_t sample_count, unsigned osr,
                            uint16_t* out_symbols, uint8_t* out_sync,
                            float* out_cfo, float* out_time_offset)
{
    const size_t N = (size_t)1 << sf, step = N * osr, total = sample_count / step;
    const int have_sync = total >= 2;
    float* w = (float*)malloc(N * sizeof(float));
    oracle_window(sf, window, w);

    /* :61-78 max-abs scan and optional normalisation into scratch */
    const ocplx* x = samples;
    ocplx* scratch = NULL;
    float max_amp = 0.0f;
    for (size_t i = 0; i < sample_count; ++i) {
        float r = fabsf(samples[i].re), im = fabsf(samples[i].im);
        float m = (r < im) ? im : r; /* std::max(r, im) */
        if (m > max_amp) max_amp = m;
    }
    if (max_amp > 1.0f) {
        if (!with_scratch) { free(w); return -ERANGE; }
        scratch = (ocplx*)malloc((sample_count ? sample_count : 1) * sizeof(ocplx));
        float scale = 1.0f / max_amp;
        for (size_t i = 0; i < sample_count; ++i) scratch[i] = c_scale(samples[i], scale);
        x = scratch;
    }

    float cfo, time_offset;
    estimate_core(sf, osr, window, w, x, total < 2 ? total : 2, 1, &cfo, &time_offset);
    if (out_cfo) *out_cfo = cfo;
    if (out_time_offset) *out_time_offset = time_offset;

    uint16_t* idx = (uint16_t*)malloc((total ? total : 1) * sizeof(uint16_t));
    demod_loop(sf, osr, window, w, NULL, x, sample_count, cfo, time_offset, idx);
    size_t produced = 0;
    for (size_t s = 0; s < total; ++s)
        if (!have_sync || s >= 2) out_symbols[produced++] = idx[s];
    if (out_sync) *out_sync = have_sync ? sync_from(sf, idx[0], idx[1]) : 0;
    free(idx); free(scratch); free(w);
    return have_sync ? (long)produced : (long)total;
}

/* ---- modulator -------------------------------------------------------------- */

/* ChirpGenerator.hpp:24-51.  Note :49 calls the C `floor` (double) on a float
 * quotient, so the wrap is evaluated in double and rounded once on assignment. */
int oracle_gen_chirp(ocplx* out, int N, int osr, int NN, float f0, int down, float ampl,
                     float* phase, float bw_scale)
{
    const float fMin = -O_PI * bw_scale / (float)osr;
    const float fMax = O_PI * bw_scale / (float)osr;
    const float fStep = (2 * O_PI * bw_scale) / (float)(N * osr * osr);
    float f = fMin + f0;
    float acc = *phase;
    int i;
    for (i = 0; i < NN; ++i) {
        f += fStep;
        if (f > fMax) f -= (fMax - fMin);
        if (down) acc -= f; else acc += f;
        out[i] = c_mk(ampl * cosf(acc), ampl * sinf(acc));
    }
    acc = (float)((double)acc - floor((double)(acc / (2 * O_PI))) * 2 * (double)O_PI);
    *phase = acc;
    return i;
}

static float bw_scale_of(unsigned bw_hz) { return (float)bw_hz / 125000.0f; } /* phy.hpp:49-51 */

/* LoRaMod.cpp:8-43 */
size_t oracle_lora_modulate(const uint16_t* symbols, size_t n, ocplx* out, unsigned sf,
                            unsigned osr, unsigned bw_hz, float amplitude, uint8_t sync)
{
    const size_t N = (size_t)1 << sf, step = N * osr;
    float phase = 0.0f;
    const float bws = bw_scale_of(bw_hz);
    amplitude = fmaxf(-1.0f, fminf(1.0f, amplitude));
    unsigned shift = sf > 4 ? sf - 4 : 0;
    const uint16_t sw[2] = { (uint16_t)((sync >> 4) << shift), (uint16_t)((sync & 0x0f) << shift) };
    for (size_t s = 0; s < n + 2; ++s) {
        uint16_t v = s < 2 ? sw[s] : symbols[s - 2];
        float f0 = (2.0f * O_PI * (float)v * bws) / ((float)N * (float)osr);
        oracle_gen_chirp(out + s * step, (int)N, (int)osr, (int)step, f0, 0, amplitude, &phase, bws);
    }
    return (n + 2) * step;
}

/* ---- shipped nibble coder ----------------------------------------------------- */
size_t oracle_lora_encode(const uint8_t* bytes, size_t n, uint16_t* out)
{
    for (size_t i = 0; i < n; ++i) {
        out[2 * i] = oracle_enc_h84(bytes[i] >> 4);
        out[2 * i + 1] = oracle_enc_h84(bytes[i] & 0x0f);
    }
    return 2 * n;
}

long oracle_lora_decode(const uint16_t* symbols, size_t n, uint8_t* out)
{
    if (n % 2) return -EINVAL;
    for (size_t i = 0; i + 1 < n; i += 2) {
        int e, b;
        uint8_t hi = oracle_dec_h84((uint8_t)symbols[i], &e, &b) & 0x0f;
        uint8_t lo = oracle_dec_h84((uint8_t)symbols[i + 1], &e, &b) & 0x0f;
        out[i / 2] = (uint8_t)((hi << 4) | lo);
    }
    return (long)(n / 2);
}

/* ---- high-level workspace API ------------------------------------------------- */
void oracle_estimate_offsets(unsigned sf, unsigned osr, int window, const ocplx* samples,
                             size_t sample_count, float* cfo, float* time_offset)
{
    const size_t N = (size_t)1 << sf, step = N * osr, symbols = sample_count / step;
    if (!samples || sample_count == 0 || symbols == 0) return; /* phy.cpp:84,90 */
    float* w = (float*)malloc(N * sizeof(float));
    oracle_window(sf, window, w);
    estimate_core(sf, osr, window, w, samples, symbols, 0, cfo, time_offset);
    free(w);
}

void oracle_compensate_offsets(unsigned sf, unsigned osr, float cfo, float time_offset,
                               ocplx* samples, size_t sample_count)
{
    if (!samples || sample_count == 0) return;
    const size_t N = (size_t)1 << sf;
    float rate = -2.0f * O_PI * cfo / ((float)N * (float)osr);
    for (size_t n = 0; n < sample_count; ++n) {
        float ph = rate * (float)n;
        samples[n] = c_mul(samples[n], c_mk(cosf(ph), sinf(ph)));
    }
    int offset = (int)roundf(time_offset);
    if (offset > 0 && (size_t)offset < sample_count) {
        memmove(samples + offset, samples, (sample_count - (size_t)offset) * sizeof(ocplx));
        memset(samples, 0, (size_t)offset * sizeof(ocplx));
    } else if (offset < 0 && (size_t)(-offset) < sample_count) {
        size_t off = (size_t)(-offset);
        memmove(samples, samples + off, (sample_count - off) * sizeof(ocplx));
        memset(samples + (sample_count - off), 0, off * sizeof(ocplx));
    }
}

long oracle_demodulate(unsigned sf, unsigned bw_hz, unsigned osr, int window,
                       const ocplx* iq, size_t sample_count, uint16_t* symbols,
                       size_t symbol_cap, uint8_t* sync_out, float* cfo_out, float* to_out)
{
    if (!iq || !symbols) return -EINVAL;
    const size_t N = (size_t)1 << sf, step = N * osr;
    if (sample_count % step != 0) return -EINVAL;
    const size_t total = sample_count / step;
    if (total < 2) return -ERANGE;
    if (total - 2 > symbol_cap) return -ERANGE;
    float* w = (float*)malloc(N * sizeof(float));
    oracle_window(sf, window, w);
    float cfo = 0.0f, to = 0.0f;
    estimate_core(sf, osr, window, w, iq, 2, 0, &cfo, &to); /* phy.cpp:196-197 */
    ocplx* down = (ocplx*)malloc(N * sizeof(ocplx));
    float ph = 0.0f;
    oracle_gen_chirp(down, (int)N, 1, (int)N, 0.0f, 1, 1.0f, &ph, bw_scale_of(bw_hz)); /* :205-208 */
    uint16_t* idx = (uint16_t*)malloc(total * sizeof(uint16_t));
    demod_loop(sf, osr, window, w, down, iq, sample_count, cfo, to, idx);
    for (size_t s = 2; s < total; ++s) symbols[s - 2] = idx[s];
    if (sync_out) *sync_out = sync_from(sf, idx[0], idx[1]);
    if (cfo_out) *cfo_out = cfo;
    if (to_out) *to_out = to;
    free(idx); free(down); free(w);
    return (long)(total - 2);
}

long oracle_decode(const uint16_t* symbols, size_t n, uint8_t* payload, size_t cap, int* crc_ok)
{
    if (!symbols || !payload) return -EINVAL;
    long produced = oracle_lora_decode(symbols, n, payload); /* writes before the cap check, like phy.cpp:249-251 */
    if (produced < 0) return produced;
    if ((size_t)produced > cap) return -ERANGE;
    if (produced >= 4) {
        uint16_t provided = (uint16_t)(payload[produced - 2] | (payload[produced - 1] << 8));
        uint16_t calc = oracle_data_checksum(payload + 2, (int)(produced - 4));
        *crc_ok = provided == calc;
    } else {
        *crc_ok = 0;
    }
    return produced;
}

/* ---- LoRaCodes.hpp helpers ------------------------------------------------------ */
static inline unsigned bit(unsigned v, unsigned n) { return (v >> n) & 1u; }

uint8_t oracle_header_checksum(const uint8_t* h)
{
    unsigned a[4], b[4], c[4];
    for (unsigned i = 0; i < 4; ++i) { a[i] = bit(h[0], 4 + i); b[i] = bit(h[0], i); c[i] = bit(h[1], i); }
    unsigned r = (a[0] ^ a[1] ^ a[2] ^ a[3]) << 4;
    r |= (a[3] ^ b[1] ^ b[2] ^ b[3] ^ c[0]) << 3;
    r |= (a[2] ^ b[0] ^ b[3] ^ c[1] ^ c[3]) << 2;
    r |= (a[1] ^ b[0] ^ b[2] ^ c[0] ^ c[1] ^ c[2]) << 1;
    r |= a[0] ^ b[1] ^ c[0] ^ c[1] ^ c[2] ^ c[3];
    return (uint8_t)r;
}

static uint16_t crc16_step8(uint16_t crc, uint16_t poly)
{
    for (int i = 0; i < 8; ++i) crc = (crc & 0x8000) ? (uint16_t)((crc << 1) ^ poly) : (uint16_t)(crc << 1);
    return crc;
}
static uint8_t parity8(uint8_t t) { t ^= t >> 4; t ^= t >> 2; t ^= t >> 1; return t & 1; }

uint16_t oracle_data_checksum(const uint8_t* d, int n)
{
    uint16_t res = 0, crc = 0;
    uint8_t v = 0xff;
    for (int i = 0; i < n; ++i) {
        crc = crc16_step8(res, 0x1021);
        v = (uint8_t)(parity8(v & 0xB8) | (v << 1));
        res = crc ^ d[i];
    }
    res ^= v;
    v = (uint8_t)(parity8(v & 0xB8) | (v << 1));
    res ^= (uint16_t)(v << 8);
    return res;
}

void oracle_whiten_sx1232(uint8_t* b, uint16_t n)
{
    uint8_t msb = 0x01, lsb = 0xFF;
    for (uint16_t j = 0; j < n; ++j) {
        b[j] ^= lsb;
        for (int i = 0; i < 8; ++i) {
            uint8_t prev = msb;
            msb = (lsb & 1) ^ ((lsb >> 5) & 1);
            lsb = (uint8_t)(((lsb >> 1) & 0xFF) | ((prev << 7) & 0x80));
        }
    }
}

void oracle_whiten_table(uint8_t* b, uint16_t n, int ofs_bits, int rdd)
{
    static const int o0[8] = { 6, 4, 2, 0, -112, -114, -302, -34 };
    static const int o1[5] = { 6, 4, 2, 0, -360 };
    static const uint64_t seq[8] = {
        0x0102291EA751AAFFULL, 0xD24B050A8D643A17ULL, 0x5B279B671120B8F4ULL, 0x032B37B9F6FB55A2ULL,
        0x994E0F87E95E2D16ULL, 0x7CBCFC7631984C26ULL, 0x281C8E4F0DAEF7F9ULL, 0x1741886EB7733B15ULL };
    const int len = 510;
    const int* o = rdd == 1 ? o1 : o0;
    for (int j = 0; j < n; ++j) {
        uint8_t x = 0;
        for (int i = 0; i < 4 + rdd; ++i) {
            int t = (o[i] + j + ofs_bits + len) % len;
            if (seq[t >> 6] & ((uint64_t)1 << (t & 0x3F))) x |= (uint8_t)(1 << i);
        }
        b[j] ^= x;
    }
}

static inline uint64_t lfsr_1d(uint64_t r)
{
    return (r >> 8) | (((r >> 32) ^ (r >> 24) ^ (r >> 16) ^ r) << 56);
}

void oracle_whiten_lfsr(uint8_t* b, uint16_t n, int ofs_bits, size_t rdd)
{
    uint64_t r[2];
    if (rdd == 1) { r[0] = 0x05121100F8ECFEEFULL; r[1] = 0xF8ECFEEFEFEFEFEFULL; }
    else          { r[0] = 0x6572D100E85C2EFFULL; r[1] = 0xE85C2EFFFFFFFFFFULL; }
    const uint8_t m = (uint8_t)(0xff >> (4 - rdd));
    int i = 0;
    for (; i < ofs_bits; ++i) r[i & 1] = lfsr_1d(r[i & 1]);
    for (int j = 0; j < n; ++j, ++i) {
        b[j] ^= (uint8_t)(r[i & 1] & m);
        r[i & 1] = lfsr_1d(r[i & 1]);
    }
}

uint16_t oracle_binary_to_gray16(uint16_t v) { return (uint16_t)(v ^ (v >> 1)); }
uint16_t oracle_gray_to_binary16(uint16_t v)
{
    v ^= v >> 8; v ^= v >> 4; v ^= v >> 2; v ^= v >> 1;
    return v;
}

uint8_t oracle_enc_h84(uint8_t x)
{
    unsigned d0 = bit(x, 0), d1 = bit(x, 1), d2 = bit(x, 2), d3 = bit(x, 3);
    unsigned b = x & 0xf;
    b |= (d0 ^ d1 ^ d2) << 4;
    b |= (d1 ^ d2 ^ d3) << 5;
    b |= (d0 ^ d1 ^ d3) << 6;
    b |= (d0 ^ d2 ^ d3) << 7;
    return (uint8_t)b;
}

uint8_t oracle_dec_h84(uint8_t b, int* err, int* bad)
{
    unsigned p0 = bit(b, 0) ^ bit(b, 1) ^ bit(b, 2) ^ bit(b, 4);
    unsigned p1 = bit(b, 1) ^ bit(b, 2) ^ bit(b, 3) ^ bit(b, 5);
    unsigned p2 = bit(b, 0) ^ bit(b, 1) ^ bit(b, 3) ^ bit(b, 6);
    unsigned p3 = bit(b, 0) ^ bit(b, 2) ^ bit(b, 3) ^ bit(b, 7);
    unsigned syn = p0 | (p1 << 1) | (p2 << 2) | (p3 << 3);
    *err = syn != 0;
    *bad = 0;
    switch (syn) {
    case 0xD: return (b ^ 1) & 0xf;
    case 0x7: return (b ^ 2) & 0xf;
    case 0xB: return (b ^ 4) & 0xf;
    case 0xE: return (b ^ 8) & 0xf;
    case 0x0: case 0x1: case 0x2: case 0x4: case 0x8: return b & 0xf;
    default: *bad = 1; return b & 0xf;
    }
}

uint8_t oracle_enc_h74(uint8_t x) { return oracle_enc_h84(x) & 0x7f; }

uint8_t oracle_dec_h74(uint8_t b, int* err)
{
    unsigned p0 = bit(b, 0) ^ bit(b, 1) ^ bit(b, 2) ^ bit(b, 4);
    unsigned p1 = bit(b, 1) ^ bit(b, 2) ^ bit(b, 3) ^ bit(b, 5);
    unsigned p2 = bit(b, 0) ^ bit(b, 1) ^ bit(b, 3) ^ bit(b, 6);
    unsigned syn = p0 | (p1 << 1) | (p2 << 2);
    *err = syn != 0;
    switch (syn) {
    case 0x5: return (b ^ 1) & 0xf;
    case 0x7: return (b ^ 2) & 0xf;
    case 0x3: return (b ^ 4) & 0xf;
    case 0x6: return (b ^ 8) & 0xf;
    default: return b & 0xf;
    }
}

uint8_t oracle_chk_p54(uint8_t b, int* err)
{
    int x = b ^ (b >> 2);
    x = x ^ (x >> 1) ^ (b >> 4);
    *err = x & 1;
    return b & 0xf;
}
uint8_t oracle_enc_p54(uint8_t b)
{
    int x = b ^ (b >> 2);
    x = x ^ (x >> 1);
    return (uint8_t)((b & 0xf) | ((x << 4) & 0x10));
}
uint8_t oracle_chk_p64(uint8_t b, int* err)
{
    int x = b ^ (b >> 1) ^ (b >> 2);
    int y = x ^ b ^ (b >> 3);
    x ^= b >> 4;
    y ^= b >> 5;
    *err = (x | y) & 1;
    return b & 0xf;
}
uint8_t oracle_enc_p64(uint8_t b)
{
    int x = b ^ (b >> 1) ^ (b >> 2);
    int y = x ^ b ^ (b >> 3);
    return (uint8_t)(((x & 1) << 4) | ((y & 1) << 5) | (b & 0xf));
}

void oracle_interleave(const uint8_t* cw, size_t ncw, uint16_t* sym, size_t ppm, size_t rdd)
{
    const size_t rows = 4 + rdd;
    for (size_t blk = 0; blk < ncw / ppm; ++blk)
        for (size_t row = 0; row < rows; ++row) {
            uint16_t s = 0;
            for (size_t col = 0; col < ppm; ++col)
                s |= (uint16_t)(((cw[blk * ppm + (col + row) % ppm] >> row) & 1u) << col);
            sym[blk * rows + row] = s;
        }
}

void oracle_deinterleave(const uint16_t* sym, size_t nsym, uint8_t* cw, size_t ppm, size_t rdd)
{
    const size_t rows = 4 + rdd;
    for (size_t blk = 0; blk < nsym / rows; ++blk)
        for (size_t row = 0; row < rows; ++row)
            for (size_t col = 0; col < ppm; ++col)
                cw[blk * ppm + (col + row) % ppm] |= (uint8_t)(((sym[blk * rows + row] >> col) & 1u) << row);
}

/* ---- timed CPU arm ---------------------------------------------------------------- */
typedef struct {
    unsigned sf, osr; int window;
    const ocplx* iq; size_t lo, hi, spp; uint16_t* out; size_t stride; uint8_t* sync;
} mt_job;

static void* mt_worker(void* arg)
{
    mt_job* j = (mt_job*)arg;
    for (size_t p = j->lo; p < j->hi; ++p) {
        uint8_t sync = 0;
        oracle_lora_demodulate(j->sf, j->window, 1, j->iq + p * j->spp, j->spp, j->osr,
                               j->out + p * j->stride, &sync, NULL, NULL);
        if (j->sync) j->sync[p] = sync;
    }
    return NULL;
}

double oracle_demod_batch_mt(unsigned sf, int window, unsigned osr, const ocplx* iq,
                             size_t n_packets, size_t samples_per_packet,
                             uint16_t* out_symbols, size_t symbol_stride, uint8_t* out_sync,
                             int n_threads)
{
    if (n_threads < 1) n_threads = 1;
    pthread_t* th = (pthread_t*)malloc((size_t)n_threads * sizeof(pthread_t));
    mt_job* jobs = (mt_job*)malloc((size_t)n_threads * sizeof(mt_job));
    struct timespec t0, t1;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    for (int t = 0; t < n_threads; ++t) {
        mt_job j = { sf, osr, window, iq, n_packets * (size_t)t / (size_t)n_threads,
                     n_packets * (size_t)(t + 1) / (size_t)n_threads, samples_per_packet,
                     out_symbols, symbol_stride, out_sync };
        jobs[t] = j;
        pthread_create(&th[t], NULL, mt_worker, &jobs[t]);
    }
    for (int t = 0; t < n_threads; ++t) pthread_join(th[t], NULL);
    clock_gettime(CLOCK_MONOTONIC, &t1);
    free(th); free(jobs);
    return (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
}
