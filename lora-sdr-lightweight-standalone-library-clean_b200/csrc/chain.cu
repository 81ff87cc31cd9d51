// chain.cu -- K4, full LoRa coding chain as integer warp kernels (one warp per packet, everything in
// shared memory between the payload bytes and the symbols):
//   TX  external/LoRa-SDR/LoRaEncoder.cpp:161-233  CRC-16 -> 3-byte explicit header as 5 H(8,4) codewords ->
//       first interleaver block always at RDD 4 -> remaining codewords at RDD (4/4 parity-free, 4/5, 4/6
//       parities, H(7,4), H(8,4)) -> table whitening (header codewords skipped, bit offset continues) ->
//       diagonal interleave -> gray-to-binary -> << (SF - PPM)
//   RX  external/LoRa-SDR/LoRaDecoder.cpp:196-397  + half LSB, >> (SF - PPM), binary-to-gray -> diagonal
//       deinterleave -> LFSR dewhitening -> header decode (length, CRC flag, RDD) -> FEC decode ->
//       CRC check / drop
// built on the bit helpers of include/lora_phy/LoRaCodes.hpp (:43-67 header checksum, :69-105 CRC,
// :147-189 whitening, :201-222 gray, :229-371 Hamming / parity, :376-412 diagonal interleaver).
// Lanes stride over codewords / symbols / bytes; the only serial pieces are the CRC (lane 0) and the
// header parse.  The work is byte-sized and tiny next to the demodulator (0.2 % of its bytes).
#include "internal.h"

namespace lk {

namespace {

constexpr int HEADER_RDD = 4, N_HEADER_SYMBOLS = 8, N_HEADER_CODEWORDS = 5;
constexpr int CH_WARPS = 4;
constexpr int MAX_CW = 2048, MAX_SYM = 1216, MAX_BYTES = 1032;      // decoder capacity (n_symbols <= 1200)

struct ChainSmem {
    uint16_t sym[MAX_SYM];
    uint8_t cw[MAX_CW];
    uint8_t bytes[MAX_BYTES];
    int hdr[8];
};

__device__ __forceinline__ unsigned bitof(unsigned v, unsigned n) { return (v >> n) & 1u; }
__device__ __forceinline__ unsigned round_up(unsigned n, unsigned f) { return ((n + f - 1) / f) * f; }

__device__ __forceinline__ unsigned enc_h84(unsigned x)                // LoRaCodes.hpp:229-242
{
    const unsigned d0 = x & 1, d1 = (x >> 1) & 1, d2 = (x >> 2) & 1, d3 = (x >> 3) & 1;
    return (x & 0xf) | ((d0 ^ d1 ^ d2) << 4) | ((d1 ^ d2 ^ d3) << 5) | ((d0 ^ d1 ^ d3) << 6) | ((d0 ^ d2 ^ d3) << 7);
}
__device__ __forceinline__ unsigned enc_p54(unsigned b) { unsigned x = b ^ (b >> 2); x ^= x >> 1; return (b & 0xf) | ((x << 4) & 0x10); }
__device__ __forceinline__ unsigned enc_p64(unsigned b)
{
    const unsigned x = b ^ (b >> 1) ^ (b >> 2), y = x ^ b ^ (b >> 3);
    return ((x & 1) << 4) | ((y & 1) << 5) | (b & 0xf);
}
__device__ __forceinline__ unsigned enc_fec(unsigned rdd, unsigned nib)   // LoRaEncoder.cpp:132-158
{
    nib &= 0xf;
    switch (rdd) {
    case 0: return nib;
    case 1: return enc_p54(nib);
    case 2: return enc_p64(nib);
    case 3: return enc_h84(nib) & 0x7f;
    default: return enc_h84(nib);
    }
}

// decoders return the data nibble (H codes) / the low nibble (parities) and OR into err / bad
__device__ __forceinline__ unsigned dec_h84(unsigned b, unsigned& err, unsigned& bad)   // LoRaCodes.hpp:250-281
{
    const unsigned b0 = b & 1, b1 = (b >> 1) & 1, b2 = (b >> 2) & 1, b3 = (b >> 3) & 1;
    const unsigned syn = (b0 ^ b1 ^ b2 ^ bitof(b, 4)) | ((b1 ^ b2 ^ b3 ^ bitof(b, 5)) << 1) |
                         ((b0 ^ b1 ^ b3 ^ bitof(b, 6)) << 2) | ((b0 ^ b2 ^ b3 ^ bitof(b, 7)) << 3);
    err |= syn != 0;
    const unsigned flip = syn == 0xD ? 1u : syn == 0x7 ? 2u : syn == 0xB ? 4u : syn == 0xE ? 8u : 0u;
    // single parity-bit errors (1, 2, 4, 8) and 0 leave the nibble; every other syndrome is a double error
    bad |= (syn != 0 && flip == 0 && syn != 1 && syn != 2 && syn != 4 && syn != 8);
    return (b ^ flip) & 0xf;
}
__device__ __forceinline__ unsigned dec_h74(unsigned b, unsigned& err)                  // LoRaCodes.hpp:306-334
{
    const unsigned b0 = b & 1, b1 = (b >> 1) & 1, b2 = (b >> 2) & 1, b3 = (b >> 3) & 1;
    const unsigned syn = (b0 ^ b1 ^ b2 ^ bitof(b, 4)) | ((b1 ^ b2 ^ b3 ^ bitof(b, 5)) << 1) | ((b0 ^ b1 ^ b3 ^ bitof(b, 6)) << 2);
    err |= syn != 0;
    const unsigned flip = syn == 0x5 ? 1u : syn == 0x7 ? 2u : syn == 0x3 ? 4u : syn == 0x6 ? 8u : 0u;
    return (b ^ flip) & 0xf;
}
__device__ __forceinline__ unsigned chk_p54(unsigned b, unsigned& err)                  // LoRaCodes.hpp:340-345
{
    unsigned x = b ^ (b >> 2);
    x = x ^ (x >> 1) ^ (b >> 4);
    err |= x & 1;
    return b & 0xf;
}
__device__ __forceinline__ unsigned chk_p64(unsigned b, unsigned& err)                  // LoRaCodes.hpp:357-365
{
    unsigned x = b ^ (b >> 1) ^ (b >> 2);
    unsigned y = x ^ b ^ (b >> 3);
    x ^= b >> 4;
    y ^= b >> 5;
    err |= (x | y) & 1;
    return b & 0xf;
}
__device__ __forceinline__ unsigned dec_fec(unsigned rdd, unsigned cw, unsigned& err, unsigned& bad)
{
    switch (rdd) {
    case 0: return cw;                      // LoRaDecoder.cpp:349-351: used raw
    case 1: return chk_p54(cw, err);
    case 2: return chk_p64(cw, err);
    case 3: return dec_h74(cw, err);
    default: return dec_h84(cw, err, bad);
    }
}

__device__ __forceinline__ unsigned header_checksum(unsigned h0, unsigned h1)           // LoRaCodes.hpp:43-67
{
    unsigned a[4], b[4], c[4];
#pragma unroll
    for (unsigned i = 0; i < 4; ++i) { a[i] = bitof(h0, 4 + i); b[i] = bitof(h0, i); c[i] = bitof(h1, i); }
    return ((a[0] ^ a[1] ^ a[2] ^ a[3]) << 4) | ((a[3] ^ b[1] ^ b[2] ^ b[3] ^ c[0]) << 3) |
           ((a[2] ^ b[0] ^ b[3] ^ c[1] ^ c[3]) << 2) | ((a[1] ^ b[0] ^ b[2] ^ c[0] ^ c[1] ^ c[2]) << 1) |
           (a[0] ^ b[1] ^ c[0] ^ c[1] ^ c[2] ^ c[3]);
}

__device__ unsigned data_checksum_s(const uint8_t* d, int n)                           // LoRaCodes.hpp:69-105
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

// Sx1272ComputeWhitening (LoRaCodes.hpp:147-167): codeword j of a run starting at bit offset ofs
__device__ __forceinline__ unsigned whiten_table(int j, int ofs, int rdd)
{
    const int o0[8] = { 6, 4, 2, 0, -112, -114, -302, -34 };
    const int o1[5] = { 6, 4, 2, 0, -360 };
    const unsigned long long seq[8] = {
        0x0102291EA751AAFFULL, 0xD24B050A8D643A17ULL, 0x5B279B671120B8F4ULL, 0x032B37B9F6FB55A2ULL,
        0x994E0F87E95E2D16ULL, 0x7CBCFC7631984C26ULL, 0x281C8E4F0DAEF7F9ULL, 0x1741886EB7733B15ULL };
    unsigned x = 0;
    for (int i = 0; i < 4 + rdd; ++i) {
        const int t = ((rdd == 1 ? o1[i] : o0[i]) + j + ofs + 510) % 510;
        if (seq[t >> 6] & (1ULL << (t & 0x3F))) x |= 1u << i;
    }
    return x;
}

__device__ __forceinline__ unsigned gray_to_binary16(unsigned v)                      // LoRaCodes.hpp:212-222
{
    v ^= v >> 8; v ^= v >> 4; v ^= v >> 2; v ^= v >> 1;
    return v & 0xffff;
}

// diagonalInterleaveSx (LoRaCodes.hpp:376-393): symbol `row` of one block
__device__ __forceinline__ unsigned interleave_row(const uint8_t* cw, unsigned ppm, unsigned row)
{
    unsigned s = 0;
    for (unsigned col = 0; col < ppm; ++col) s |= ((cw[(col + row) % ppm] >> row) & 1u) << col;
    return s;
}
// diagonalDeterleaveSx (LoRaCodes.hpp:396-412): codeword k of one block of `rows` symbols
__device__ __forceinline__ unsigned deinterleave_cw(const uint16_t* sym, unsigned ppm, unsigned rows, unsigned k)
{
    unsigned c = 0;
    for (unsigned row = 0; row < rows; ++row) {
        const unsigned col = (k + ppm - row % ppm) % ppm;          // (col + row) % ppm == k
        c |= ((sym[row] >> col) & 1u) << row;
    }
    return c;
}

} // namespace

// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(CH_WARPS * 32) chain_encode_kernel(const ChainArgs a, const uint8_t* payload,
                                                                     size_t payload_stride, uint32_t len,
                                                                     uint32_t n_packets, uint16_t* symbols,
                                                                     size_t symbol_stride)
{
    __shared__ ChainSmem sm[CH_WARPS];
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    ChainSmem& s = sm[w];
    const unsigned ppm = a.ppm, rdd = a.rdd, sf = a.sf;
    const unsigned nbytes = len + (a.crc ? 2u : 0u);
    const unsigned hdrn = a.explicit_hdr ? N_HEADER_CODEWORDS : 0;
    const unsigned ncw = round_up(nbytes * 2 + hdrn, ppm);
    const unsigned nsym = N_HEADER_SYMBOLS + (ncw / ppm - 1) * (4 + rdd);
    for (uint32_t pid = blockIdx.x * CH_WARPS + w; pid < n_packets; pid += gridDim.x * CH_WARPS) {
        const uint8_t* p = payload + (size_t)pid * payload_stride;
        for (unsigned i = lane; i < len; i += 32) s.bytes[i] = p[i];
        __syncwarp();
        if (lane == 0 && a.crc) {                                        // LoRaEncoder.cpp:176-180
            const unsigned crc = data_checksum_s(s.bytes, (int)len);
            s.bytes[len] = (uint8_t)(crc & 0xff);
            s.bytes[len + 1] = (uint8_t)(crc >> 8);
        }
        __syncwarp();
        for (unsigned c = lane; c < ncw; c += 32) {
            unsigned v;
            if (c < hdrn) {                                              // :182-194
                const unsigned h0 = len & 0xff, h1 = (a.crc ? 1u : 0u) | (rdd << 1), h2 = header_checksum(h0, h1);
                const unsigned nib = c == 0 ? h0 >> 4 : c == 1 ? h0 & 0xf : c == 2 ? h1 & 0xf : c == 3 ? h2 >> 4 : h2 & 0xf;
                v = enc_h84(nib);
            } else {
                const unsigned d = c - hdrn;
                const unsigned b = (d >> 1) < nbytes ? s.bytes[d >> 1] : 0u;   // padding nibbles: 0 (see oracle header)
                const unsigned nib = (d & 1) ? b >> 4 : b & 0xf;
                if (c < ppm) {                                           // first block: RDD 4, whitening offset 0 (:196-199)
                    v = enc_fec(4, nib);
                    if (a.whitening) v ^= whiten_table((int)(c - hdrn), 0, HEADER_RDD);
                } else {                                                 // :201-207
                    v = enc_fec(rdd, nib);
                    if (a.whitening) v ^= whiten_table((int)(c - ppm), (int)(ppm - hdrn), (int)rdd);
                }
            }
            s.cw[c] = (uint8_t)v;
        }
        __syncwarp();
        uint16_t* out = symbols + (size_t)pid * symbol_stride;
        for (unsigned i = lane; i < nsym; i += 32) {                     // :210-220
            unsigned v;
            if (i < N_HEADER_SYMBOLS) v = interleave_row(s.cw, ppm, i);
            else {
                const unsigned j = i - N_HEADER_SYMBOLS, rows = 4 + rdd;
                v = interleave_row(s.cw + ppm + (j / rows) * ppm, ppm, j % rows);
            }
            out[i] = (uint16_t)(gray_to_binary16(v) << (sf - ppm));
        }
        __syncwarp();
    }
}

// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(CH_WARPS * 32) chain_decode_kernel(const ChainArgs a, const uint16_t* symbols,
                                                                     size_t symbol_stride, const int32_t* counts,
                                                                     uint32_t n_symbols, uint32_t n_packets,
                                                                     const uint8_t* lfsr, uint8_t* payload,
                                                                     size_t payload_stride, int32_t* out_len,
                                                                     uint8_t* out_flags)
{
    __shared__ ChainSmem sm[CH_WARPS];
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    ChainSmem& s = sm[w];
    const unsigned ppm = a.ppm, sf = a.sf, cfg_rdd = a.rdd;
    const uint8_t* lf4 = lfsr;                 // whitening bytes of the RDD != 1 register pair (mask applied here)
    const uint8_t* lf1 = lfsr + MAX_CW;        // RDD == 1 register pair
    for (uint32_t pid = blockIdx.x * CH_WARPS + w; pid < n_packets; pid += gridDim.x * CH_WARPS) {
        const int32_t cnt = counts ? counts[pid] : (int32_t)n_symbols;
        int result = -1;
        unsigned flags = 0;
        if (cnt >= N_HEADER_SYMBOLS && (uint32_t)cnt <= n_symbols) {                       // LoRaDecoder.cpp:208
            const unsigned n_in = (unsigned)cnt;
            const unsigned nsym = round_up(n_in, 4 + cfg_rdd);
            const unsigned ncw = (nsym / (4 + cfg_rdd)) * ppm;
            const uint16_t* in = symbols + (size_t)pid * symbol_stride;
            for (unsigned i = lane; i < nsym; i += 32) {                                   // :218-222, uint16 arithmetic
                unsigned v = i < n_in ? in[i] : 0u;
                v = (v + ((1u << (sf - ppm)) / 2)) & 0xffff;
                v >>= (sf - ppm);
                s.sym[i] = (uint16_t)(v ^ (v >> 1));
            }
            for (unsigned i = lane; i < ncw + 32 && i < MAX_CW; i += 32) s.cw[i] = 0;   // upstream may read a little past ncw
            __syncwarp();
            const unsigned hdrn = a.explicit_hdr ? N_HEADER_CODEWORDS : 0;
            for (unsigned c = lane; c < ncw; c += 32) {                                    // :226-256
                unsigned v = 0;
                if (cfg_rdd != HEADER_RDD) {
                    if (c < ppm) {
                        v = deinterleave_cw(s.sym, ppm, N_HEADER_SYMBOLS, c);
                        if (c >= hdrn) v ^= lf4[c - hdrn] & 0xff;
                    } else if (nsym > N_HEADER_SYMBOLS) {                                  // :238 (nothing at all otherwise)
                        const unsigned rows = 4 + cfg_rdd, blk = (c - ppm) / ppm, k = (c - ppm) % ppm;
                        if ((blk + 1) * rows <= nsym - N_HEADER_SYMBOLS)
                            v = deinterleave_cw(s.sym + N_HEADER_SYMBOLS + blk * rows, ppm, rows, k);
                        const unsigned wi = (a.explicit_hdr ? ppm - N_HEADER_CODEWORDS : ppm) + (c - ppm);
                        v ^= (cfg_rdd == 1 ? lf1[wi] : lf4[wi]) & (0xffu >> (4 - cfg_rdd));
                    }
                } else {
                    v = deinterleave_cw(s.sym + (c / ppm) * 8, ppm, 8, c % ppm);
                    if (c >= hdrn) v ^= lf4[c - hdrn] & 0xff;
                }
                s.cw[c] = (uint8_t)v;
            }
            for (unsigned i = lane; i < (ncw + 1) / 2 + 8 && i < MAX_BYTES; i += 32) s.bytes[i] = 0;
            __syncwarp();

            // ---- header (lane 0; every lane reads the outcome) : :284-316 ----
            if (lane == 0) {
                unsigned err = 0, bad = 0, rdd = cfg_rdd, dofs = 0, cofs = 0, packet_len, data_len;
                int check_crc = a.crc_check, ok = 1;
                if (a.explicit_hdr) {
                    unsigned b0 = dec_h84(s.cw[1], err, bad) & 0xf;
                    b0 |= (dec_h84(s.cw[0], err, bad) << 4) & 0xff;
                    const unsigned b1 = dec_h84(s.cw[2], err, bad) & 0xf;
                    unsigned b2 = dec_h84(s.cw[4], err, bad) & 0xf;
                    b2 |= (dec_h84(s.cw[3], err, bad) << 4) & 0xff;
                    b2 ^= header_checksum(b0, b1);
                    s.bytes[0] = (uint8_t)b0; s.bytes[1] = (uint8_t)b1; s.bytes[2] = (uint8_t)b2;
                    if (err && a.error_check) ok = 0;
                    if (!(b1 & 1)) check_crc = 0;
                    rdd = (b1 >> 1) & 7;
                    if (rdd > 4) ok = 0;
                    packet_len = b0;
                    data_len = packet_len + ((b1 & 1) ? 5 : 3);
                    cofs = N_HEADER_CODEWORDS;
                    dofs = 6;
                } else {
                    packet_len = a.data_length;
                    data_len = a.crc_check ? packet_len + 2 : packet_len;
                }
                if (data_len > (ncw + 1) / 2) ok = 0;
                if (ok) {                                                                  // :318-344 rest of the first block
                    for (; cofs < ppm; ++cofs, ++dofs) {
                        if (dofs & 1) s.bytes[dofs >> 1] |= (uint8_t)(dec_h84(s.cw[cofs], err, bad) << 4);
                        else s.bytes[dofs >> 1] = (uint8_t)(dec_h84(s.cw[cofs], err, bad) & 0xf);
                    }
                    if (dofs & 1) {
                        s.bytes[dofs >> 1] |= (uint8_t)(dec_fec(rdd, s.cw[cofs++], err, bad) << 4);
                        dofs++;
                    }
                    dofs >>= 1;
                    if (err && a.error_check) ok = 0;
                }
                s.hdr[0] = ok; s.hdr[1] = (int)rdd; s.hdr[2] = (int)dofs; s.hdr[3] = (int)cofs;
                s.hdr[4] = (int)packet_len; s.hdr[5] = (int)data_len; s.hdr[6] = check_crc; s.hdr[7] = (int)(err | (bad << 1));
            }
            __syncwarp();
            int ok = s.hdr[0];
            const unsigned rdd = (unsigned)s.hdr[1], dofs = (unsigned)s.hdr[2], cofs = (unsigned)s.hdr[3];
            const unsigned packet_len = (unsigned)s.hdr[4];
            unsigned data_len = (unsigned)s.hdr[5];
            unsigned err = (unsigned)s.hdr[7] & 1, bad = ((unsigned)s.hdr[7] >> 1) & 1;
            if (ok) {                                                                      // :348-364, one byte per lane
                for (unsigned i = dofs + lane; i < data_len; i += 32) {
                    const unsigned c0 = cofs + 2 * (i - dofs);
                    unsigned v;
                    if (rdd == 0) v = (s.cw[c0] & 0xf) | ((s.cw[c0 + 1] << 4) & 0xff);
                    else if (rdd >= 3) v = (dec_fec(rdd, s.cw[c0], err, bad) & 0xf) | ((dec_fec(rdd, s.cw[c0 + 1], err, bad) << 4) & 0xff);
                    else v = dec_fec(rdd, s.cw[c0], err, bad) | ((dec_fec(rdd, s.cw[c0 + 1], err, bad) << 4) & 0xff);
                    s.bytes[i] = (uint8_t)v;
                }
                err = __any_sync(0xffffffffu, err != 0);
                bad = __any_sync(0xffffffffu, bad != 0);
                if (err && a.error_check) ok = 0;
            }
            __syncwarp();
            unsigned out_ofs = 0, crc_present = 0, crc_match = 0;
            if (ok) {                                                                      // :366-391
                if (lane == 0) {
                    int keep = 1;
                    if (a.explicit_hdr) {
                        if (s.bytes[1] & 1) {
                            const unsigned crc = data_checksum_s(s.bytes + 3, (int)packet_len);
                            const unsigned pkt_crc = s.bytes[3 + packet_len] | (s.bytes[4 + packet_len] << 8);
                            crc_present = 1;
                            crc_match = crc == pkt_crc;
                            if (crc != pkt_crc && s.hdr[6]) keep = 0;
                            s.bytes[3 + packet_len] ^= (uint8_t)crc;
                            s.bytes[4 + packet_len] ^= (uint8_t)(crc >> 8);
                        }
                    } else if (s.hdr[6]) {
                        const unsigned crc = data_checksum_s(s.bytes, (int)a.data_length);
                        const unsigned pkt_crc = s.bytes[a.data_length] | (s.bytes[a.data_length + 1] << 8);
                        crc_present = 1;
                        crc_match = crc == pkt_crc;
                        if (crc != pkt_crc) keep = 0;
                        s.bytes[a.data_length] ^= (uint8_t)crc;
                        s.bytes[a.data_length + 1] ^= (uint8_t)(crc >> 8);
                    }
                    s.hdr[0] = keep; s.hdr[1] = (int)(crc_present | (crc_match << 1));
                }
                __syncwarp();
                ok = s.hdr[0];
                crc_present = (unsigned)s.hdr[1] & 1; crc_match = ((unsigned)s.hdr[1] >> 1) & 1;
                if (a.explicit_hdr && !a.out_hdr) { out_ofs = 3; data_len -= 5; }           // unsigned wrap -> rejected below
            }
            flags = err | (bad << 1);
            if (ok) {
                flags |= (crc_present << 2) | (crc_match << 3);
                if (data_len > payload_stride || data_len > MAX_BYTES) result = -34;
                else {
                    uint8_t* o = payload + (size_t)pid * payload_stride;
                    for (unsigned i = lane; i < data_len; i += 32) o[i] = s.bytes[out_ofs + i];
                    result = (int)data_len;
                }
            }
        }
        if (lane == 0) {
            if (out_len) out_len[pid] = result;
            if (out_flags) out_flags[pid] = (uint8_t)flags;
        }
        __syncwarp();
    }
}

cudaError_t launch_chain_encode(const ChainArgs& a, const uint8_t* payload, size_t payload_stride, uint32_t len,
                                uint32_t n_packets, uint16_t* symbols, size_t symbol_stride, int sm_count,
                                cudaStream_t stream)
{
    if (!n_packets) return cudaSuccess;
    unsigned need = (n_packets + CH_WARPS - 1) / CH_WARPS, cap = (unsigned)sm_count * 16u;
    chain_encode_kernel<<<need < cap ? need : cap, CH_WARPS * 32, 0, stream>>>(a, payload, payload_stride, len, n_packets,
                                                                              symbols, symbol_stride);
    return cudaGetLastError();
}

cudaError_t launch_chain_decode(const ChainArgs& a, const uint16_t* symbols, size_t symbol_stride,
                                const int32_t* counts, uint32_t n_symbols, uint32_t n_packets, const uint8_t* lfsr,
                                uint8_t* payload, size_t payload_stride, int32_t* out_len, uint8_t* out_flags,
                                int sm_count, cudaStream_t stream)
{
    if (!n_packets) return cudaSuccess;
    unsigned need = (n_packets + CH_WARPS - 1) / CH_WARPS, cap = (unsigned)sm_count * 16u;
    chain_decode_kernel<<<need < cap ? need : cap, CH_WARPS * 32, 0, stream>>>(a, symbols, symbol_stride, counts, n_symbols,
                                                                              n_packets, lfsr, payload, payload_stride,
                                                                              out_len, out_flags);
    return cudaGetLastError();
}

// Sx1272ComputeWhiteningLfsr (LoRaCodes.hpp:176-189): the byte XORed into codeword i of a run that starts at
// bit offset 0, for both register pairs (RDD == 1 / others), unmasked.  Host side, once per context.
void chain_lfsr_table(uint8_t* out /* 2 * MAX_CW */)
{
    for (int t = 0; t < 2; ++t) {
        unsigned long long r[2];
        if (t == 1) { r[0] = 0x05121100F8ECFEEFULL; r[1] = 0xF8ECFEEFEFEFEFEFULL; }
        else        { r[0] = 0x6572D100E85C2EFFULL; r[1] = 0xE85C2EFFFFFFFFFFULL; }
        for (int i = 0; i < MAX_CW; ++i) {
            out[t * MAX_CW + i] = (uint8_t)(r[i & 1] & 0xff);
            const unsigned long long x = r[i & 1];
            r[i & 1] = (x >> 8) | (((x >> 32) ^ (x >> 24) ^ (x >> 16) ^ x) << 56);
        }
    }
}
int chain_lfsr_table_size() { return 2 * MAX_CW; }
int chain_max_symbols() { return 1200; }

} // namespace lk
