// oracle/ref_shim.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// C entry points around the UNMODIFIED reference sources, which are compiled
// where they lie under /root/reference (see oracle/Makefile, target _ref).
// Nothing from the reference is copied into this repository: this file only
// #includes the reference's public headers and forwards calls.  The resulting
// oracle/_ref/libref_lora_phy.so is used (a) to pin the plain-C restatement in
// oracle/lora_oracle.c, (b) as the checker in tests/, and (c) as the timed CPU
// arm of bench.py (`--impl reference` / cpu_baseline.kind == "reference").
#include <lora_phy/phy.hpp>
#include <lora_phy/LoRaCodes.hpp>
#include <lora_phy/ChirpGenerator.hpp>

#include <atomic>
#include <chrono>
#include <cstring>
#include <thread>
#include <vector>

using cf = std::complex<float>;
using namespace lora_phy;

static bandwidth to_bw(unsigned hz) { return static_cast<bandwidth>(hz); }

extern "C" {

// src/phy/LoRaDemod.cpp:50 via lora_demod_init (:11).  Returns what the reference returns.
long ref_lora_demodulate(unsigned sf, int window, int with_scratch,
                         const float* iq, size_t sample_count, unsigned osr,
                         uint16_t* out_symbols, uint8_t* out_sync,
                         float* out_cfo, float* out_time_offset)
{
    static thread_local lora_demod_workspace ws;
    std::vector<cf> scratch(with_scratch ? sample_count : 0);
    lora_demod_init(&ws, sf, window ? window_type::window_hann : window_type::window_none,
                    with_scratch ? scratch.data() : nullptr, scratch.size());
    ssize_t r = lora_demodulate(&ws, reinterpret_cast<const cf*>(iq), sample_count,
                                out_symbols, osr, out_sync);
    if (out_cfo) *out_cfo = ws.metrics.cfo;
    if (out_time_offset) *out_time_offset = ws.metrics.time_offset;
    lora_demod_free(&ws);
    return static_cast<long>(r);
}

// Spectrum of one N-point window exactly as LoRaDetector::detect sees it (LoRaDetector.hpp:39-74).
long ref_detect(unsigned sf, const float* in_iq, float* out_spectrum,
                float* power, float* power_avg, float* findex)
{
    static thread_local lora_demod_workspace ws;
    lora_demod_init(&ws, sf);
    const size_t N = ws.N;
    for (size_t i = 0; i < N; ++i) ws.detector->feed(i, cf(in_iq[2 * i], in_iq[2 * i + 1]));
    float p, pav, fi;
    size_t idx = ws.detector->detect(p, pav, fi);
    if (out_spectrum) std::memcpy(out_spectrum, ws.fft_out, N * sizeof(cf));
    if (power) *power = p;
    if (power_avg) *power_avg = pav;
    if (findex) *findex = fi;
    lora_demod_free(&ws);
    return static_cast<long>(idx);
}

// src/phy/LoRaMod.cpp:8
size_t ref_lora_modulate(const uint16_t* symbols, size_t n, float* out_iq, unsigned sf,
                         unsigned osr, unsigned bw_hz, float amplitude, uint8_t sync)
{
    return lora_modulate(symbols, n, reinterpret_cast<cf*>(out_iq), sf, osr, to_bw(bw_hz),
                         amplitude, sync);
}

// include/lora_phy/ChirpGenerator.hpp:24
int ref_gen_chirp(float* out_iq, int N, int osr, int NN, float f0, int down, float ampl,
                  float* phase, float bw_scale)
{
    return genChirp(reinterpret_cast<cf*>(out_iq), N, osr, NN, f0, down != 0, ampl, *phase, bw_scale);
}

size_t ref_lora_encode(const uint8_t* bytes, size_t n, uint16_t* out, unsigned sf)
{ return lora_encode(bytes, n, out, sf); }

long ref_lora_decode(const uint16_t* symbols, size_t n, uint8_t* out)
{ return static_cast<long>(lora_decode(symbols, n, out)); }

// ---- high-level workspace API (src/phy/phy.cpp) -----------------------------------------
struct ref_ws {
    lora_workspace ws{};
    std::vector<uint16_t> symbol_buf;
    std::vector<cf> fft_in, fft_out;
    std::vector<float> window;
};

void* ref_ws_create(unsigned sf, unsigned bw_hz, unsigned cr, unsigned osr, int window,
                    uint8_t sync, int* rc)
{
    auto* h = new ref_ws;
    const size_t N = size_t(1) << sf;
    h->symbol_buf.resize(N);
    h->fft_in.resize(N);
    h->fft_out.resize(N * (osr ? osr : 1));
    h->window.resize(N);
    h->ws.symbol_buf = h->symbol_buf.data();
    h->ws.fft_in = h->fft_in.data();
    h->ws.fft_out = h->fft_out.data();
    h->ws.window = h->window.data();
    lora_params p;
    p.sf = sf; p.bw = to_bw(bw_hz); p.cr = cr; p.osr = osr;
    p.window = window ? window_type::window_hann : window_type::window_none;
    p.sync_word = sync;
    int r = init(&h->ws, &p);
    if (rc) *rc = r;
    return h;
}
void ref_ws_destroy(void* h) { delete static_cast<ref_ws*>(h); }
long ref_ws_encode(void* h, const uint8_t* p, size_t n, uint16_t* s, size_t cap)
{ return encode(&static_cast<ref_ws*>(h)->ws, p, n, s, cap); }
long ref_ws_decode(void* h, const uint16_t* s, size_t n, uint8_t* p, size_t cap)
{ return decode(&static_cast<ref_ws*>(h)->ws, s, n, p, cap); }
long ref_ws_modulate(void* h, const uint16_t* s, size_t n, float* iq, size_t cap)
{ return modulate(&static_cast<ref_ws*>(h)->ws, s, n, reinterpret_cast<cf*>(iq), cap); }
long ref_ws_demodulate(void* h, const float* iq, size_t n, uint16_t* s, size_t cap)
{ return demodulate(&static_cast<ref_ws*>(h)->ws, reinterpret_cast<const cf*>(iq), n, s, cap); }
void ref_ws_estimate_offsets(void* h, const float* iq, size_t n)
{ estimate_offsets(&static_cast<ref_ws*>(h)->ws, reinterpret_cast<const cf*>(iq), n); }
void ref_ws_compensate_offsets(void* h, float* iq, size_t n)
{ compensate_offsets(&static_cast<ref_ws*>(h)->ws, reinterpret_cast<cf*>(iq), n); }
void ref_ws_set_metrics(void* h, float cfo, float time_offset)
{ auto& m = static_cast<ref_ws*>(h)->ws.metrics; m.cfo = cfo; m.time_offset = time_offset; }
void ref_ws_get_metrics(void* h, int* crc_ok, float* cfo, float* time_offset, uint8_t* sync)
{
    auto* w = &static_cast<ref_ws*>(h)->ws;
    const lora_metrics* m = get_last_metrics(w);
    if (crc_ok) *crc_ok = m->crc_ok;
    if (cfo) *cfo = m->cfo;
    if (time_offset) *time_offset = m->time_offset;
    if (sync) *sync = w->sync_word;
}
void ref_ws_get_window(void* h, float* out, size_t n)
{ std::memcpy(out, static_cast<ref_ws*>(h)->window.data(), n * sizeof(float)); }
void ref_twiddles(unsigned sf, float* out)   // kissfft.hh:24-29 via kissfft::init
{
    static thread_local kissfft_plan<float> plan;
    kissfft<float>::init(plan, 1 << sf, false);
    std::memcpy(out, plan.twiddles, (size_t(1) << sf) * sizeof(cf));
}

// ---- include/lora_phy/LoRaCodes.hpp helpers ---------------------------------------------
uint8_t  ref_header_checksum(const uint8_t* h) { return headerChecksum(h); }
uint16_t ref_data_checksum(const uint8_t* d, int n) { return sx1272DataChecksum(d, n); }
uint8_t  ref_checksum8(const uint8_t* d, size_t n) { return checksum8(d, n); }
void ref_whiten_sx1232(uint8_t* b, uint16_t n) { SX1232RadioComputeWhitening(b, n); }
void ref_whiten_table(uint8_t* b, uint16_t n, int ofs, int rdd) { Sx1272ComputeWhitening(b, n, ofs, rdd); }
void ref_whiten_lfsr(uint8_t* b, uint16_t n, int ofs, size_t rdd) { Sx1272ComputeWhiteningLfsr(b, n, ofs, rdd); }
uint16_t ref_binary_to_gray16(uint16_t v) { return binaryToGray16(v); }
uint16_t ref_gray_to_binary16(uint16_t v) { return grayToBinary16(v); }
uint8_t ref_enc_h84(uint8_t v) { return encodeHamming84sx(v); }
uint8_t ref_dec_h84(uint8_t v, int* err, int* bad)
{ bool e = false, b = false; uint8_t r = decodeHamming84sx(v, e, b); *err = e; *bad = b; return r; }
uint8_t ref_enc_h74(uint8_t v) { return encodeHamming74sx(v); }
uint8_t ref_dec_h74(uint8_t v, int* err)
{ bool e = false; uint8_t r = decodeHamming74sx(v, e); *err = e; return r; }
uint8_t ref_enc_p54(uint8_t v) { return encodeParity54(v); }
uint8_t ref_chk_p54(uint8_t v, int* err) { bool e = false; uint8_t r = checkParity54(v, e); *err = e; return r; }
uint8_t ref_enc_p64(uint8_t v) { return encodeParity64(v); }
uint8_t ref_chk_p64(uint8_t v, int* err) { bool e = false; uint8_t r = checkParity64(v, e); *err = e; return r; }
void ref_interleave(const uint8_t* cw, size_t ncw, uint16_t* sym, size_t ppm, size_t rdd)
{ diagonalInterleaveSx(cw, ncw, sym, ppm, rdd); }
void ref_deinterleave(const uint16_t* sym, size_t nsym, uint8_t* cw, size_t ppm, size_t rdd)
{ diagonalDeterleaveSx(sym, nsym, cw, ppm, rdd); }
void ref_deinterleave2(const uint16_t* sym, size_t nsym, uint8_t* cw, size_t ppm, size_t rdd)
{ diagonalDeterleaveSx2(sym, nsym, cw, ppm, rdd); }

// ---- timed CPU arm: one packet per thread, private workspace per thread -----------------
// Mirrors the shape of tests/performance_test.cpp:112-125 (demod-only).  Returns seconds.
double ref_demod_batch_mt(unsigned sf, int window, unsigned osr, const float* iq,
                          size_t n_packets, size_t samples_per_packet,
                          uint16_t* out_symbols, size_t symbol_stride, uint8_t* out_sync,
                          int n_threads)
{
    if (n_threads < 1) n_threads = 1;
    std::vector<std::thread> pool;
    std::atomic<int> ready{0};
    std::atomic<bool> go{false};
    auto worker = [&](int tid) {
        auto* ws = new lora_demod_workspace;
        std::vector<cf> scratch(samples_per_packet);
        lora_demod_init(ws, sf, window ? window_type::window_hann : window_type::window_none,
                        scratch.data(), scratch.size());
        ready.fetch_add(1);
        while (!go.load()) std::this_thread::yield();
        const size_t lo = n_packets * size_t(tid) / size_t(n_threads);
        const size_t hi = n_packets * size_t(tid + 1) / size_t(n_threads);
        for (size_t p = lo; p < hi; ++p) {
            uint8_t sync = 0;
            lora_demodulate(ws, reinterpret_cast<const cf*>(iq) + p * samples_per_packet,
                            samples_per_packet, out_symbols + p * symbol_stride, osr, &sync);
            if (out_sync) out_sync[p] = sync;
        }
        lora_demod_free(ws);
        delete ws;
    };
    for (int t = 0; t < n_threads; ++t) pool.emplace_back(worker, t);
    while (ready.load() < n_threads) std::this_thread::yield();
    auto t0 = std::chrono::steady_clock::now();
    go.store(true);
    for (auto& th : pool) th.join();
    auto t1 = std::chrono::steady_clock::now();
    return std::chrono::duration<double>(t1 - t0).count();
}

} // extern "C"
