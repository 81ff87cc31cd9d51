// modcodes.cu -- K3 chirp synthesis, caller-side dechirp, K4 shipped nibble coder,
// compensate_offsets.  Reference loops restated:
//   lora_modulate        src/phy/LoRaMod.cpp:8-43
//   genChirp             include/lora_phy/ChirpGenerator.hpp:24-51
//   lora_encode/decode   src/phy/LoRaEncoder.cpp:6-18, src/phy/LoRaDecoder.cpp:7-21
//   decode CRC check     src/phy/phy.cpp:252-259, LoRaCodes.hpp:69-105
//   compensate_offsets   src/phy/phy.cpp:150-180
#include <mutex>
#include "internal.h"
#include "glibc_math.cuh"
#include "mc_noise.cuh"

namespace lk {

// ------------------------------------------------------------------------------------------
// K3.  The modulator's phase is ONE serial float32 recurrence per packet: genChirp accumulates
// the frequency ramp and the phase sample by sample and carries the (wrapped) phase from chirp
// to chirp (LoRaMod.cpp:14,27,32,40).  Reproducing the reference IQ to 1e-5 requires replaying
// exactly that sequence of float adds, so the work is split in two kernels:
//   mod_phase_kernel   one thread per packet walks all chirps with adds only (no sincos) and
//                      records the start phase of every chirp;
//   mod_emit_kernel    one thread per chirp replays its N*osr adds from the recorded start and
//                      emits (amp*cosf, amp*sinf); a warp stages 32 chirps x 32 samples in
//                      shared memory so that global stores are 256-byte row segments.
// ------------------------------------------------------------------------------------------
struct ChirpConst { float fMin, fMax, fStep, span; };

__device__ __forceinline__ ChirpConst chirp_const(int N, unsigned osr, float bw_scale)
{
    const float PI_F = 3.14159265358979323846f;
    ChirpConst c;
    c.fMin = __fdiv_rn(fx::fmul(-PI_F, bw_scale), (float)osr);
    c.fMax = __fdiv_rn(fx::fmul(PI_F, bw_scale), (float)osr);
    c.fStep = __fdiv_rn(fx::fmul(2 * PI_F, bw_scale), (float)(int)(N * osr * osr));
    c.span = fx::fsub(c.fMax, c.fMin);
    return c;
}

__device__ __forceinline__ float symbol_f0(unsigned v, int N, unsigned osr, float bw_scale)
{
    const float PI_F = 3.14159265358979323846f;   // LoRaMod.cpp:24-25,36-37
    return __fdiv_rn(fx::fmul(fx::fmul(2.0f * PI_F, (float)v), bw_scale), fx::fmul((float)N, (float)osr));
}

__device__ __forceinline__ unsigned chirp_value(const ModArgs& a, uint32_t pid, uint32_t s)
{
    const unsigned shift = a.sf > 4 ? a.sf - 4 : 0;
    if (s == 0) return (uint16_t)(((a.sync & 0xff) >> 4) << shift);
    if (s == 1) return (uint16_t)((a.sync & 0x0f) << shift);
    return a.symbols[(size_t)pid * a.symbol_stride + (s - 2)];
}

// ---- upstream frame format (external/LoRa-SDR/LoRaMod.cpp:109-238, ChirpGenerator.hpp:22-47) -------------
// Constants there are double expressions narrowed to float, f0 = float((2*M_PI*v)/NN), the wrap is all double.
constexpr double PI_D = 3.14159265358979323846;

// bw_k = 1: upstream's constants; 2 / 4: the reference's bandwidth semantics on the same frame (LoRaMod.cpp:24-25,36-37)
__device__ __forceinline__ ChirpConst frame_const(int N, unsigned ovs, int bw_k)
{
    ChirpConst c;
    c.fMin = (float)(-PI_D * (double)bw_k / (double)(int)ovs);
    c.fMax = (float)(PI_D * (double)bw_k / (double)(int)ovs);
    c.fStep = (float)((2 * PI_D * (double)bw_k) / (double)(int)(N * ovs * ovs));
    c.span = fx::fsub(c.fMax, c.fMin);
    return c;
}

struct ChirpDesc { float f0; uint32_t len; uint32_t offset; bool down; };

// chirp s of packet pid: start frequency, direction, length and sample offset inside the packet
__device__ __forceinline__ ChirpDesc chirp_desc(const ModArgs& a, uint32_t pid, uint32_t s, int N, uint32_t step)
{
    ChirpDesc d;
    if (!a.frame) {
        d.f0 = symbol_f0(chirp_value(a, pid, s), N, a.osr, a.bw_scale);
        d.len = step; d.offset = s * step; d.down = false;
        return d;
    }
    int v = 0;
    d.down = s >= 12 && s < 15;
    d.len = s == 14 ? step / 4 : step;
    d.offset = s <= 14 ? s * step : 14 * step + step / 4 + (s - 15) * step;
    if (s == 10) v = (int)((a.sync >> 4) & 0xf) * 8;
    else if (s == 11) v = (int)(a.sync & 0xf) * 8;
    else if (s >= 15) v = (int)a.symbols[(size_t)pid * a.symbol_stride + (s - 15)];
    d.f0 = (float)((2 * PI_D * (double)v * (double)(int)a.bw_scale) / (double)step);
    return d;
}

__device__ __forceinline__ float wrap_phase_frame(float acc)
{
    return (float)((double)acc - floor((double)acc / (2 * PI_D)) * 2 * PI_D);
}

__device__ __forceinline__ float wrap_phase(float acc)
{
    // ChirpGenerator.hpp:49: floor() is the C double function, the update rounds once
    const float PI_F = 3.14159265358979323846f;
    const double q = floor((double)__fdiv_rn(acc, 2 * PI_F));
    return (float)((double)acc - q * 2 * (double)PI_F);
}

__global__ void __launch_bounds__(128) mod_phase_kernel(const ModArgs a)
{
    const uint32_t pid = blockIdx.x * blockDim.x + threadIdx.x;
    if (pid >= a.n_packets) return;
    const int N = 1 << a.sf;
    const uint32_t step = (uint32_t)N * a.osr;
    const ChirpConst c = a.frame ? frame_const(N, a.osr, (int)a.bw_scale) : chirp_const(N, a.osr, a.bw_scale);
    const uint32_t chirps = mod_chirps(a);
    float phase = 0.0f;
    float* out = a.phase_scratch + (size_t)pid * chirps;
    for (uint32_t s = 0; s < chirps; ++s) {
        out[s] = phase;
        const ChirpDesc d = chirp_desc(a, pid, s, N, step);
        float f = fx::fadd(c.fMin, d.f0);
        for (uint32_t i = 0; i < d.len; ++i) {
            f = fx::fadd(f, c.fStep);
            if (f > c.fMax) f = fx::fsub(f, c.span);
            phase = d.down ? fx::fsub(phase, f) : fx::fadd(phase, f);
        }
        phase = a.frame ? wrap_phase_frame(phase) : wrap_phase(phase);
    }
}

constexpr int EMIT_WARPS = 4;

__global__ void __launch_bounds__(EMIT_WARPS * 32) mod_emit_kernel(const ModArgs a)
{
    __shared__ cf tile[EMIT_WARPS][32][33];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int N = 1 << a.sf;
    const uint32_t step = (uint32_t)N * a.osr;
    const ChirpConst c = a.frame ? frame_const(N, a.osr, (int)a.bw_scale) : chirp_const(N, a.osr, a.bw_scale);
    const uint32_t chirps = mod_chirps(a);
    const unsigned long long n_chirps = (unsigned long long)a.n_packets * chirps;
    const float amp = a.frame ? a.amplitude : fmaxf(-1.0f, fminf(1.0f, a.amplitude));      // LoRaMod.cpp:18 (upstream: as given)
    const unsigned long long warps_total = (unsigned long long)gridDim.x * EMIT_WARPS;
    for (unsigned long long w = (unsigned long long)blockIdx.x * EMIT_WARPS + warp; w * 32 < n_chirps; w += warps_total) {
        const unsigned long long chirp = w * 32 + lane;
        const bool active = chirp < n_chirps;
        const uint32_t pid = active ? (uint32_t)(chirp / chirps) : 0;
        const uint32_t s = active ? (uint32_t)(chirp % chirps) : 0;
        float phase = active ? a.phase_scratch[chirp] : 0.0f;
        ChirpDesc d = chirp_desc(a, pid, active ? s : 0, N, step);
        if (!active) d.len = 0;
        float f = fx::fadd(c.fMin, d.f0);
        for (uint32_t j0 = 0; j0 < step; j0 += 32) {
            const uint32_t jn = min(32u, step - j0);
            const uint32_t jm = j0 < d.len ? min(jn, d.len - j0) : 0;
            for (uint32_t j = 0; j < jm; ++j) {
                f = fx::fadd(f, c.fStep);
                if (f > c.fMax) f = fx::fsub(f, c.span);
                phase = d.down ? fx::fsub(phase, f) : fx::fadd(phase, f);
                float sn, cs;
                gm::sincosf_g(phase, &sn, &cs);
                tile[warp][lane][j] = fx::mk(fx::fmul(amp, cs), fx::fmul(amp, sn));
            }
            __syncwarp();
#pragma unroll 4
            for (int r = 0; r < 32; ++r) {
                const unsigned long long ch = w * 32 + r;
                if (ch < n_chirps && (uint32_t)lane < jn) {
                    const uint32_t rp = (uint32_t)(ch / chirps), rs = (uint32_t)(ch % chirps);
                    uint32_t off = rs * step, len = step;
                    if (a.frame) {
                        len = rs == 14 ? step / 4 : step;
                        off = rs <= 14 ? rs * step : 14 * step + step / 4 + (rs - 15) * step;
                    }
                    if (j0 + lane < len) {
                        // Monte-Carlo driver: noise and the caller-side dechirp are applied HERE, so the clean frame never
                        // exists in HBM and the channel costs no pass of its own (same arithmetic as awgn_dechirp_kernel)
                        cf v = tile[warp][r][lane];
                        if (a.channel.on) v = mc_channel_sample(a.channel, a.channel.first_packet + rp, off + j0 + lane, v);
                        a.iq[(size_t)rp * a.packet_stride + off + j0 + lane] = v;
                    }
                }
            }
            __syncwarp();
        }
    }
    if (a.frame && a.padding) {                                     // LoRaMod.cpp:226-236: padding symbols of zeros
        const uint32_t pad = a.padding * step, body = 14 * step + step / 4 + a.symbols_per_packet * step;
        const unsigned long long total = (unsigned long long)a.n_packets * pad;
        for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
             i += (unsigned long long)gridDim.x * blockDim.x)
            a.iq[(size_t)(i / pad) * a.packet_stride + body + (uint32_t)(i % pad)] = fx::mk(0.0f, 0.0f);
    }
}

cudaError_t launch_modulate(const ModArgs& a, int sm_count, cudaStream_t stream, unsigned* launches)
{
    if (a.n_packets == 0) return cudaSuccess;
    mod_phase_kernel<<<(a.n_packets + 127) / 128, 128, 0, stream>>>(a);
    ++*launches;
    const unsigned long long n_chirps = (unsigned long long)a.n_packets * mod_chirps(a);
    unsigned long long need = (n_chirps + EMIT_WARPS * 32 - 1) / (EMIT_WARPS * 32);
    unsigned max_ctas = (unsigned)sm_count * 64u;
    mod_emit_kernel<<<need < max_ctas ? (unsigned)need : max_ctas, EMIT_WARPS * 32, 0, stream>>>(a);
    ++*launches;
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------
// Allocation-free modulator (lora_phy_cuda_modulate_batch_static): the per-chirp start phases go to a file-scope device
// array instead of the context's growable buffer.  Every GPU that loads the library holds the array (4 MiB) until exit.
// In host code `g_mod_phase_static` is only a host-side stand-in for the symbol, NOT a device address: the address the
// kernels get comes from cudaGetSymbolAddress, asked once per device with that device current.
// ------------------------------------------------------------------------------------------
static __device__ __align__(256) float g_mod_phase_static[MOD_STATIC_MAX_CHIRPS];

cudaError_t mod_static_scratch(float** out)
{
    static std::mutex mu;
    static float* addr[64] = {};
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(mu);
    if (!addr[dev]) {
        float* p = nullptr;
        e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_mod_phase_static);
        if (e != cudaSuccess) return e;
        if (!p) return cudaErrorInvalidSymbol;
        addr[dev] = p;
    }
    *out = addr[dev];
    return cudaSuccess;
}

// ------------------------------------------------------------------------------------------
// caller-side dechirp: iq[s*L + i] *= table[i]  (tests/e2e_chain_test.cpp:88-92)
// ------------------------------------------------------------------------------------------
__global__ void dechirp_kernel(cf* iq, size_t packet_stride, uint32_t spp, uint32_t n_packets,
                               const cf* __restrict__ table, uint32_t L)
{
    const unsigned long long total = (unsigned long long)n_packets * spp;
    for (unsigned long long k = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; k < total;
         k += (unsigned long long)gridDim.x * blockDim.x) {
        const uint32_t pid = (uint32_t)(k / spp), j = (uint32_t)(k % spp);
        if (j >= (spp / L) * L) continue;                 // trailing partial symbol untouched
        cf* p = iq + (size_t)pid * packet_stride + j;
        const float2 t = __ldg(reinterpret_cast<const float2*>(table + (j % L)));
        *p = fx::cmul(*p, fx::mk(t.x, t.y));
    }
}

cudaError_t launch_dechirp(cf* iq, size_t packet_stride, uint32_t spp, uint32_t n_packets,
                           const cf* table, uint32_t L, cudaStream_t stream)
{
    const unsigned long long total = (unsigned long long)n_packets * spp;
    if (!total) return cudaSuccess;
    unsigned long long need = (total + 255) / 256;
    dechirp_kernel<<<need < 148u * 32u ? (unsigned)need : 148u * 32u, 256, 0, stream>>>(iq, packet_stride, spp, n_packets, table, L);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------
// K4 shipped coder: Hamming(8,4) nibble <-> symbol (LoRaCodes.hpp:229-281)
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned enc_h84(unsigned x)
{
    const unsigned d0 = x & 1, d1 = (x >> 1) & 1, d2 = (x >> 2) & 1, d3 = (x >> 3) & 1;
    return (x & 0xf) | ((d0 ^ d1 ^ d2) << 4) | ((d1 ^ d2 ^ d3) << 5) | ((d0 ^ d1 ^ d3) << 6) | ((d0 ^ d2 ^ d3) << 7);
}

__device__ __forceinline__ unsigned dec_h84(unsigned b)
{
    const unsigned b0 = b & 1, b1 = (b >> 1) & 1, b2 = (b >> 2) & 1, b3 = (b >> 3) & 1;
    const unsigned p0 = b0 ^ b1 ^ b2 ^ ((b >> 4) & 1);
    const unsigned p1 = b1 ^ b2 ^ b3 ^ ((b >> 5) & 1);
    const unsigned p2 = b0 ^ b1 ^ b3 ^ ((b >> 6) & 1);
    const unsigned p3 = b0 ^ b2 ^ b3 ^ ((b >> 7) & 1);
    const unsigned syn = p0 | (p1 << 1) | (p2 << 2) | (p3 << 3);
    // syndrome -> data-bit flip mask (0xD,0x7,0xB,0xE correct d0..d3; everything else leaves the nibble)
    const unsigned flip = syn == 0xD ? 1u : syn == 0x7 ? 2u : syn == 0xB ? 4u : syn == 0xE ? 8u : 0u;
    return (b ^ flip) & 0xf;
}

__global__ void encode_kernel(const uint8_t* payload, size_t payload_stride, uint32_t len, uint32_t n_packets,
                              uint16_t* symbols, size_t symbol_stride)
{
    const unsigned long long total = (unsigned long long)n_packets * len;
    for (unsigned long long k = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; k < total;
         k += (unsigned long long)gridDim.x * blockDim.x) {
        const uint32_t pid = (uint32_t)(k / len), i = (uint32_t)(k % len);
        const unsigned byte = payload[(size_t)pid * payload_stride + i];
        uint16_t* o = symbols + (size_t)pid * symbol_stride + 2 * i;
        o[0] = (uint16_t)enc_h84(byte >> 4);
        o[1] = (uint16_t)enc_h84(byte & 0xf);
    }
}

// sx1272DataChecksum (LoRaCodes.hpp:92-105) -- one thread per packet, bytes only
__device__ unsigned data_checksum(const uint8_t* d, int n)
{
    unsigned res = 0, crc = 0, v = 0xff;
    for (int i = 0; i < n; ++i) {
        crc = res;
        for (int k = 0; k < 8; ++k) crc = (crc & 0x8000) ? ((crc << 1) ^ 0x1021) & 0xffff : (crc << 1) & 0xffff;
        unsigned t = v & 0xB8; t ^= t >> 4; t ^= t >> 2; t ^= t >> 1;
        v = ((t & 1) | (v << 1)) & 0xff;
        res = crc ^ d[i];
    }
    res ^= v;
    unsigned t = v & 0xB8; t ^= t >> 4; t ^= t >> 2; t ^= t >> 1;
    v = ((t & 1) | (v << 1)) & 0xff;
    res ^= v << 8;
    return res & 0xffff;
}

__global__ void decode_kernel(const uint16_t* symbols, size_t symbol_stride, const int32_t* counts,
                              uint32_t spp, uint32_t n_packets, uint8_t* payload, size_t payload_stride)
{
    const uint32_t half = spp / 2;
    const unsigned long long total = (unsigned long long)n_packets * half;
    for (unsigned long long k = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; k < total;
         k += (unsigned long long)gridDim.x * blockDim.x) {
        const uint32_t pid = (uint32_t)(k / half), i = (uint32_t)(k % half);
        const int32_t cnt = counts ? counts[pid] : (int32_t)spp;
        if (cnt < 0 || (cnt & 1) || 2 * i + 1 >= (uint32_t)cnt) continue;
        const uint16_t* s = symbols + (size_t)pid * symbol_stride + 2 * i;
        payload[(size_t)pid * payload_stride + i] = (uint8_t)((dec_h84(s[0] & 0xff) << 4) | dec_h84(s[1] & 0xff));
    }
}

__global__ void crc_kernel(const int32_t* counts, uint32_t spp, uint32_t n_packets, const uint8_t* payload,
                           size_t payload_stride, uint8_t* crc_ok)
{
    const uint32_t pid = blockIdx.x * blockDim.x + threadIdx.x;
    if (pid >= n_packets) return;
    const int32_t cnt = counts ? counts[pid] : (int32_t)spp;
    int produced = (cnt < 0 || (cnt & 1)) ? 0 : cnt / 2;
    if ((uint32_t)produced > spp / 2) produced = (int)(spp / 2);
    const uint8_t* p = payload + (size_t)pid * payload_stride;
    uint8_t ok = 0;
    if (produced >= 4) {                                   // phy.cpp:252-259
        const unsigned provided = p[produced - 2] | (p[produced - 1] << 8);
        ok = provided == data_checksum(p + 2, produced - 4);
    }
    crc_ok[pid] = ok;
}

cudaError_t launch_encode(const uint8_t* payload, size_t payload_stride, uint32_t len, uint32_t n_packets,
                          uint16_t* symbols, size_t symbol_stride, cudaStream_t stream)
{
    const unsigned long long total = (unsigned long long)n_packets * len;
    if (!total) return cudaSuccess;
    unsigned long long need = (total + 255) / 256;
    encode_kernel<<<need < 148u * 32u ? (unsigned)need : 148u * 32u, 256, 0, stream>>>(payload, payload_stride, len, n_packets, symbols, symbol_stride);
    return cudaGetLastError();
}

cudaError_t launch_decode(const uint16_t* symbols, size_t symbol_stride, const int32_t* counts, uint32_t spp,
                          uint32_t n_packets, uint8_t* payload, size_t payload_stride, uint8_t* crc_ok,
                          cudaStream_t stream)
{
    if (!n_packets) return cudaSuccess;
    const unsigned long long total = (unsigned long long)n_packets * (spp / 2);
    if (total) {
        unsigned long long need = (total + 255) / 256;
        decode_kernel<<<need < 148u * 32u ? (unsigned)need : 148u * 32u, 256, 0, stream>>>(symbols, symbol_stride, counts, spp, n_packets, payload, payload_stride);
    }
    if (crc_ok) crc_kernel<<<(n_packets + 127) / 128, 128, 0, stream>>>(counts, spp, n_packets, payload, payload_stride, crc_ok);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------
// compensate_offsets (phy.cpp:150-180): out[n] = rot(x[n - offset]) with zero fill; `in` and
// `out` are distinct buffers (the ABI layer copies back for the in-place contract).
// ------------------------------------------------------------------------------------------
__global__ void compensate_kernel(const cf* in, cf* out, size_t packet_stride, uint32_t spp, uint32_t n_packets,
                                  int sf, uint32_t osr, const float* cfo, const float* time_offset)
{
    const float PI_F = 3.14159265358979323846f;
    const unsigned long long total = (unsigned long long)n_packets * spp;
    for (unsigned long long k = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; k < total;
         k += (unsigned long long)gridDim.x * blockDim.x) {
        const uint32_t pid = (uint32_t)(k / spp), n = (uint32_t)(k % spp);
        const float rate = __fdiv_rn(fx::fmul(-2.0f * PI_F, cfo[pid]), fx::fmul((float)(1 << sf), (float)osr));
        int offset = (int)roundf(time_offset[pid]);
        if (offset > 0 && (uint32_t)offset >= spp) offset = 0;        // phy.cpp:165,170 guards
        if (offset < 0 && (uint32_t)(-offset) >= spp) offset = 0;
        const long long src = (long long)n - offset;
        cf r = fx::mk(0.0f, 0.0f);
        if (src >= 0 && src < (long long)spp) {
            const cf x = in[(size_t)pid * packet_stride + (size_t)src];
            const float ph = fx::fmul(rate, (float)(uint32_t)src);
            float sn, cs;
            gm::sincosf_g(ph, &sn, &cs);
            r = fx::cmul(x, fx::mk(cs, sn));
        }
        out[(size_t)pid * packet_stride + n] = r;
    }
}

cudaError_t launch_compensate(const cf* in, cf* out, size_t packet_stride, uint32_t spp, uint32_t n_packets,
                              int sf, uint32_t osr, const float* cfo, const float* time_offset, cudaStream_t stream)
{
    const unsigned long long total = (unsigned long long)n_packets * spp;
    if (!total) return cudaSuccess;
    unsigned long long need = (total + 255) / 256;
    compensate_kernel<<<need < 148u * 32u ? (unsigned)need : 148u * 32u, 256, 0, stream>>>(in, out, packet_stride, spp, n_packets, sf, osr, cfo, time_offset);
    return cudaGetLastError();
}

} // namespace lk
