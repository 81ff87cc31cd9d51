/* oracle/lora_stream_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Plain-C restatement of the upstream streaming receiver the reference vendors under external/LoRa-SDR
 * (LoRaDemod.cpp:68-117 tables, :145-327 work loop): continuous IQ is consumed one window at a time;
 * each window is dechirped with the current chirp table and a 128-step fine-tune table, run through the
 * detector (LoRaDetector.hpp:39-74, restated as oracle_detect) and fed to the frame-sync state machine
 *   FRAMESYNC (preamble tracking, sync word0 / word1 match) -> DOWNCHIRP0 -> DOWNCHIRP1 (signed bin error
 *   averaged) -> QUARTERCHIRP (consume N/4 + err/2) -> DATASYMBOLS (emit until mtu or squelch) -> FRAMESYNC.
 * PARITY STATUS: pinned only by this restatement (the Pothos block is not compilable here and no in-tree
 * test runs it); the detector and FFT underneath are pinned against the compiled reference.
 * Upstream leaves _prevValue / _freqError / _finefreqError uninitialised before the first window; 0 here.
 * The FFT is the in-tree kissfft restatement (oracle_fft), as everywhere else in this oracle.
 */
#define _USE_MATH_DEFINES
#define _DEFAULT_SOURCE
#include "lora_oracle.h"
#include <math.h>
#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif
#include <stdlib.h>
#include <string.h>

#define FINE_STEPS 128

/* LoRaDemod.cpp:96-117: std::polar(1.0, double phaseAccum) narrowed to float.  up = conj(entry). */
void oracle_stream_tables_bw(unsigned sf, unsigned bw_k, ocplx* up, ocplx* down, ocplx* fine /* N * 128 */);
void oracle_stream_tables(unsigned sf, ocplx* up, ocplx* down, ocplx* fine /* N * 128 */)
{
    oracle_stream_tables_bw(sf, 1, up, down, fine);
}

/* bw_k = 1: upstream's tables.  bw_k = 2, 4 (BW250 / 500): the reference's bandwidth semantics applied to them --
 * bw_scale multiplies the chirp's start frequency AND its slope (ChirpGenerator.hpp:28-30), so the ramp runs from
 * -pi*k in steps of 2*pi*k/N (frequencies are taken modulo 2*pi by the sampling itself).  This repository's
 * extension of the upstream receiver to cfg 4 of BASELINE.json; with k = 1 every expression is upstream's. */
void oracle_stream_tables_bw(unsigned sf, unsigned bw_k, ocplx* up, ocplx* down, ocplx* fine /* N * 128 */)
{
    const size_t N = (size_t)1 << sf;
    float phase = (float)(-M_PI * (double)bw_k);
    double acc = 0.0;
    for (size_t i = 0; i < N; ++i) {
        acc += phase;
        const double re = cos(acc), im = sin(acc);
        up[i].re = (float)re;   up[i].im = (float)(-im);
        down[i].re = (float)re; down[i].im = (float)im;
        phase = (float)((double)phase + (2 * M_PI * (double)bw_k) / (double)N);   /* float += double: summed in double, rounded once */
    }
    acc = 0.0;
    phase = (float)(2.0 * M_PI / (double)(N * FINE_STEPS));
    for (size_t i = 0; i < N * FINE_STEPS; ++i) {
        acc += phase;
        fine[i].re = (float)cos(acc);
        fine[i].im = (float)sin(acc);
    }
}

static inline ocplx cmulf(ocplx a, ocplx b)
{
    ocplx r;
    r.re = a.re * b.re - a.im * b.im;
    r.im = a.re * b.im + a.im * b.re;
    return r;
}

/* one window: LoRaDemod.cpp:157-166 (fine-tune index recurrence: int -= float * size_t, truncated) */
static void dechirp_window(const ocplx* in, size_t osr, const ocplx* chirp, const ocplx* fine, size_t N, int* ft, float fine_err,
                           ocplx* out)
{
    const int wrap = (int)(N * FINE_STEPS);
    int idx = *ft;
    for (size_t i = 0; i < N; ++i) {
        out[i] = cmulf(cmulf(in[i * osr], chirp[i]), fine[idx]);      /* osr > 1: sample phase 0 of every chip, as LoRaDemod.cpp:94-99 */
        idx = (int)((float)idx - fine_err * (float)FINE_STEPS);
        if (idx < 0) idx += wrap;
        else if (idx >= wrap) idx -= wrap;
    }
    *ft = idx;
}

/* Runs the receiver over iq[0, n).  Frames are reported when their DATASYMBOLS state ends (mtu reached or
 * squelched) -- a frame still open at the end of the buffer is not reported, as upstream would not have
 * posted it yet.  Per frame: start = sample index of the window that matched sync word0, err = the averaged
 * down-chirp bin error (_freqError), nsym symbols (int16) at symbols[f * mtu].  Returns the frame count
 * (frames beyond max_frames are counted but not stored). */
long oracle_stream_demod_ex(unsigned sf, unsigned osr, unsigned bw_k, unsigned sync, float thresh, unsigned mtu, const ocplx* iq,
                            size_t n, size_t max_frames, uint32_t* frame_start, int32_t* frame_err, int32_t* frame_nsym,
                            int16_t* symbols, uint32_t* end_pos);
long oracle_stream_demod(unsigned sf, unsigned sync, float thresh, unsigned mtu, const ocplx* iq, size_t n,
                         size_t max_frames, uint32_t* frame_start, int32_t* frame_err, int32_t* frame_nsym,
                         int16_t* symbols, uint32_t* end_pos)
{
    return oracle_stream_demod_ex(sf, 1, 1, sync, thresh, mtu, iq, n, max_frames, frame_start, frame_err, frame_nsym, symbols, end_pos);
}

/* The receiver on an oversampled stream (osr samples per chip: windows are N*osr samples, every slide is in chips) with
 * the bandwidth-scaled tables above; the sync-word targets follow the modulator (nibble*8*bw_k bins, modulo N) and the
 * two places where the state machine converts a bin offset into a time slide divide by bw_k (a delay of d chips moves
 * the dechirped tone by bw_k*d bins).  With osr = bw_k = 1 every expression is upstream's. */
long oracle_stream_demod_ex(unsigned sf, unsigned osr, unsigned bw_k, unsigned sync, float thresh, unsigned mtu, const ocplx* iq,
                            size_t n, size_t max_frames, uint32_t* frame_start, int32_t* frame_err, int32_t* frame_nsym,
                            int16_t* symbols, uint32_t* end_pos)
{
    const size_t N = (size_t)1 << sf;
    const size_t sync0 = (((size_t)(sync >> 4) & 0xf) * bw_k) % (N / 8), sync1 = (((size_t)sync & 0xf) * bw_k) % (N / 8);
    ocplx* up = (ocplx*)malloc(sizeof(ocplx) * N);
    ocplx* down = (ocplx*)malloc(sizeof(ocplx) * N);
    ocplx* fine = (ocplx*)malloc(sizeof(ocplx) * N * FINE_STEPS);
    ocplx* dec = (ocplx*)malloc(sizeof(ocplx) * N);
    ocplx* spec = (ocplx*)malloc(sizeof(ocplx) * N);
    int16_t* cur = (int16_t*)malloc(sizeof(int16_t) * (mtu ? mtu : 1));
    oracle_stream_tables_bw(sf, bw_k, up, down, fine);
    enum { FRAMESYNC, DOWNCHIRP0, DOWNCHIRP1, QUARTERCHIRP, DATASYMBOLS } state = FRAMESYNC;
    const ocplx* chirp = up;
    short prev_value = 0;
    int freq_error = 0, fine_idx = 0;
    float fine_err = 0.0f;
    size_t sym_count = 0, pos = 0, frames = 0;
    uint32_t cur_start = 0;
    while (pos + 2 * N * osr <= n) {                                     /* :148 */
        dechirp_window(iq + pos, osr, chirp, fine, N, &fine_idx, fine_err, dec);
        float power = 0, power_avg = 0, findex = 0;
        const size_t value = oracle_detect(sf, dec, spec, &power, &power_avg, &findex);
        const float snr = power - power_avg;
        const int squelched = snr < thresh;
        size_t total = 0;
        switch (state) {
        case FRAMESYNC: {                                                /* :179-237 */
            const int syncd = !squelched && (prev_value + 4) / 8 == 0;
            const int match0 = (value + 4) / 8 == sync0;
            int match1 = 0;
            if (syncd && match0) {
                int ft = fine_idx;
                dechirp_window(iq + pos + N * osr, osr, chirp, fine, N, &ft, fine_err, dec);
                float p1, pa1, f1;
                const size_t value1 = oracle_detect(sf, dec, spec, &p1, &pa1, &f1);
                findex = f1; power = p1; power_avg = pa1;              /* detect() overwrites the same variables */
                match1 = (value1 + 4) / 8 == sync1;
            }
            if (syncd && match0 && match1) {
                total = 2 * N;
                state = DOWNCHIRP0;
                chirp = down;
                cur_start = (uint32_t)pos;
            } else if (!squelched) {
                total = N - value / bw_k;                                 /* a delay of d chips moves the tone by bw_k*d bins */
                fine_err += findex;
            } else {
                total = N;
                fine_err = 0;
                fine_idx = 0;
            }
        } break;
        case DOWNCHIRP0: {                                               /* :240-250 */
            state = DOWNCHIRP1;
            total = N;
            int error = (int)value;
            if (value > N / 2) error -= (int)N;
            freq_error = error;
        } break;
        case DOWNCHIRP1: {                                               /* :253-270 */
            state = QUARTERCHIRP;
            total = N;
            chirp = up;
            int error = (int)value;
            if (value > N / 2) error -= (int)N;
            freq_error = (freq_error + error) / 2;
        } break;
        case QUARTERCHIRP: {                                             /* :273-283 */
            state = DATASYMBOLS;
            total = (size_t)((long)(N / 4) + (long)((freq_error / 2) / (int)bw_k));
            fine_err += (float)(freq_error / 2);
            sym_count = 0;
        } break;
        case DATASYMBOLS: {                                              /* :286-310 */
            total = N;
            cur[sym_count++] = (int16_t)value;
            if (sym_count >= mtu || squelched) {
                if (frames < max_frames) {
                    frame_start[frames] = cur_start;
                    frame_err[frames] = freq_error;
                    frame_nsym[frames] = (int32_t)sym_count;
                    memcpy(symbols + frames * mtu, cur, sizeof(int16_t) * sym_count);
                }
                ++frames;
                fine_err = 0;
                state = FRAMESYNC;
            }
        } break;
        }
        pos += total * osr;
        prev_value = (short)value;
    }
    if (end_pos) *end_pos = (uint32_t)pos;
    free(up); free(down); free(fine); free(dec); free(spec); free(cur);
    return (long)frames;
}

/* ---- upstream transmitter: frame format (external/LoRa-SDR/LoRaMod.cpp:109-238) ------------------------
 * genChirp as vendored upstream (external/LoRa-SDR/ChirpGenerator.hpp:22-47): the constants are double
 * expressions narrowed to float, the ramp and the phase are float recurrences, the sample is
 * std::polar(ampl, phase) = (ampl*cosf, ampl*sinf), and the phase is wrapped in double after each chirp. */
static size_t upstream_gen_chirp(ocplx* out, int N, int ovs, int NN, float f0, int down, float ampl, float* phase, int bw_k)
{
    const float fMin = (float)(-M_PI * bw_k / ovs);                 /* bw_k = 1: upstream's constants */
    const float fMax = (float)(M_PI * bw_k / ovs);
    const float fStep = (float)((2 * M_PI * bw_k) / (N * ovs * ovs));
    float f = fMin + f0;
    float acc = *phase;
    int i;
    for (i = 0; i < NN; ++i) {
        f += fStep;
        if (f > fMax) f -= (fMax - fMin);
        if (down) acc -= f; else acc += f;
        out[i].re = ampl * cosf(acc);
        out[i].im = ampl * sinf(acc);
    }
    acc = (float)((double)acc - floor((double)acc / (2 * M_PI)) * 2 * M_PI);
    *phase = acc;
    return (size_t)i;
}

size_t oracle_frame_samples(unsigned sf, unsigned ovs, size_t n_symbols, unsigned padding)
{
    const size_t NN = ((size_t)1 << sf) * ovs;
    return (14 + n_symbols + padding) * NN + NN / 4;
}

/* One frame: 10 up-chirps (:141-146), sync word0 / word1 at (nibble*8) bins (:151-170), two down-chirps and
 * a quarter down-chirp (:175-202), one up-chirp per symbol (:207-221), `padding` symbols of zeros (:226-236). */
size_t oracle_frame_modulate_bw(const uint16_t* symbols, size_t n, ocplx* out, unsigned sf, unsigned ovs, unsigned sync,
                                float ampl, unsigned padding, unsigned bw_k);
size_t oracle_frame_modulate(const uint16_t* symbols, size_t n, ocplx* out, unsigned sf, unsigned ovs, unsigned sync,
                             float ampl, unsigned padding)
{
    return oracle_frame_modulate_bw(symbols, n, out, sf, ovs, sync, ampl, padding, 1);
}

/* the same frame with the reference's bandwidth semantics (LoRaMod.cpp:24-25,36-37: bw_scale on slope and offset) */
size_t oracle_frame_modulate_bw(const uint16_t* symbols, size_t n, ocplx* out, unsigned sf, unsigned ovs, unsigned sync,
                                float ampl, unsigned padding, unsigned bw_k)
{
    const int N = 1 << sf, NN = N * (int)ovs, K = (int)bw_k;
    float phase = 0.0f;
    size_t o = 0;
    for (int k = 0; k < 10; ++k) o += upstream_gen_chirp(out + o, N, (int)ovs, NN, 0.0f, 0, ampl, &phase, K);
    const int sw0 = (int)((sync >> 4) & 0xf) * 8, sw1 = (int)(sync & 0xf) * 8;
    o += upstream_gen_chirp(out + o, N, (int)ovs, NN, (float)((2 * M_PI * sw0 * K) / NN), 0, ampl, &phase, K);
    o += upstream_gen_chirp(out + o, N, (int)ovs, NN, (float)((2 * M_PI * sw1 * K) / NN), 0, ampl, &phase, K);
    o += upstream_gen_chirp(out + o, N, (int)ovs, NN, 0.0f, 1, ampl, &phase, K);
    o += upstream_gen_chirp(out + o, N, (int)ovs, NN, 0.0f, 1, ampl, &phase, K);
    o += upstream_gen_chirp(out + o, N, (int)ovs, NN / 4, 0.0f, 1, ampl, &phase, K);
    for (size_t s = 0; s < n; ++s)
        o += upstream_gen_chirp(out + o, N, (int)ovs, NN, (float)((2 * M_PI * (int)symbols[s] * K) / NN), 0, ampl, &phase, K);
    memset(out + o, 0, sizeof(ocplx) * (size_t)padding * (size_t)NN);
    return o + (size_t)padding * (size_t)NN;
}
