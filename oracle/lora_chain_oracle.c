/* oracle/lora_chain_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Plain-C restatement of the upstream LoRa coding chain the reference vendors under
 * external/LoRa-SDR (pinned commit c545c51; not compilable here: #include <Pothos/Framework.hpp>):
 *   TX  LoRaEncoder.cpp:161-233  CRC -> 3-byte header as 5 H(8,4) codewords -> first block at RDD 4 ->
 *       rest at RDD -> table whitening (header codewords skipped) -> diagonal interleave -> gray -> shift
 *   RX  LoRaDecoder.cpp:196-397  round/shift -> gray -> deinterleave -> LFSR dewhitening -> header decode ->
 *       FEC decode per RDD -> CRC check / drop
 * built on the LoRaCodes.hpp helpers restated in lora_oracle.c (each pinned against the compiled
 * reference by tests/test_oracle.py).  PARITY STATUS OF THE CHAIN ORDER: pinned only by this
 * restatement -- no reference test exercises the Pothos blocks in-tree.
 * One deliberate choice: LoRaEncoder.cpp:132-158 reads bytes[dOfs >> 1] past the end of its vector for
 * the codewords that pad the last interleaver block (undefined upstream); here those nibbles are 0.
 */
#include "lora_oracle.h"
#include <string.h>

#define HEADER_RDD 4
#define N_HEADER_SYMBOLS 8
#define N_HEADER_CODEWORDS 5

static unsigned round_up(unsigned num, unsigned factor) { return ((num + factor - 1) / factor) * factor; }   /* LoRaCodes.hpp:12 */

static uint8_t enc_fec(unsigned rdd, uint8_t nib)                       /* LoRaEncoder.cpp:132-158 */
{
    switch (rdd) {
    case 0: return nib & 0xf;
    case 1: return oracle_enc_p54(nib);
    case 2: return oracle_enc_p64(nib);
    case 3: return oracle_enc_h74(nib);
    default: return oracle_enc_h84(nib);
    }
}

static uint8_t dec_fec(unsigned rdd, uint8_t cw, int* error, int* bad)  /* LoRaDecoder.cpp:312-350 */
{
    int e = 0, b = 0;
    uint8_t r;
    switch (rdd) {
    case 0: return cw;                                   /* used raw upstream (masked / shifted by the caller) */
    case 1: r = oracle_chk_p54(cw, &e); break;
    case 2: r = oracle_chk_p64(cw, &e); break;
    case 3: r = oracle_dec_h74(cw, &e); break;
    default: r = oracle_dec_h84(cw, &e, &b); break;
    }
    if (e) *error = 1;
    if (b) *bad = 1;
    return r;
}

size_t oracle_chain_num_symbols(const oracle_chain_cfg* c, size_t payload_len)
{
    const unsigned ppm = c->ppm ? c->ppm : c->sf;
    const unsigned nbytes = (unsigned)payload_len + (c->crc ? 2u : 0u);
    const unsigned ncw = round_up(nbytes * 2 + (c->explicit_hdr ? N_HEADER_CODEWORDS : 0), ppm);
    return N_HEADER_SYMBOLS + (ncw / ppm - 1) * (4 + c->rdd);
}

long oracle_encode_chain(const oracle_chain_cfg* c, const uint8_t* payload, size_t payload_len, uint16_t* symbols)
{
    const unsigned sf = c->sf, ppm = c->ppm ? c->ppm : c->sf, rdd = c->rdd;
    if (ppm > sf || ppm < 5 || rdd > 4 || payload_len > 255) return -22;
    uint8_t bytes[260];
    uint8_t cw[640];
    const unsigned nbytes = (unsigned)payload_len + (c->crc ? 2u : 0u);
    memset(bytes, 0, sizeof bytes);
    memset(cw, 0, sizeof cw);
    memcpy(bytes, payload, payload_len);
    const unsigned ncw = round_up(nbytes * 2 + (c->explicit_hdr ? N_HEADER_CODEWORDS : 0), ppm);
    const unsigned nsym = N_HEADER_SYMBOLS + (ncw / ppm - 1) * (4 + rdd);
    unsigned cofs = 0, dofs = 0;
    if (c->crc) {                                                        /* LoRaEncoder.cpp:176-180 */
        const uint16_t crc = oracle_data_checksum(bytes, (int)payload_len);
        bytes[payload_len] = (uint8_t)(crc & 0xff);
        bytes[payload_len + 1] = (uint8_t)(crc >> 8);
    }
    if (c->explicit_hdr) {                                               /* :182-194 */
        uint8_t hdr[3];
        hdr[0] = (uint8_t)payload_len;
        hdr[1] = (uint8_t)((c->crc ? 1 : 0) | (rdd << 1));
        hdr[2] = oracle_header_checksum(hdr);
        cw[cofs++] = oracle_enc_h84(hdr[0] >> 4);
        cw[cofs++] = oracle_enc_h84(hdr[0] & 0xf);
        cw[cofs++] = oracle_enc_h84(hdr[1] & 0xf);
        cw[cofs++] = oracle_enc_h84(hdr[2] >> 4);
        cw[cofs++] = oracle_enc_h84(hdr[2] & 0xf);
    }
    const unsigned cofs1 = cofs;
    for (; cofs < ppm; ++cofs, ++dofs) {                                 /* first block: always RDD 4 (:196) */
        const uint8_t b = (dofs >> 1) < nbytes ? bytes[dofs >> 1] : 0;
        cw[cofs] = enc_fec(4, (dofs & 1) ? (uint8_t)(b >> 4) : (uint8_t)(b & 0xf));
    }
    if (c->whitening) oracle_whiten_table(cw + cofs1, (uint16_t)(ppm - cofs1), 0, HEADER_RDD);
    if (ncw > ppm) {                                                     /* :201-207 */
        for (; cofs < ncw; ++cofs, ++dofs) {
            const uint8_t b = (dofs >> 1) < nbytes ? bytes[dofs >> 1] : 0;
            cw[cofs] = enc_fec(rdd, (dofs & 1) ? (uint8_t)(b >> 4) : (uint8_t)(b & 0xf));
        }
        if (c->whitening) oracle_whiten_table(cw + ppm, (uint16_t)(ncw - ppm), (int)(ppm - cofs1), (int)rdd);
    }
    oracle_interleave(cw, ppm, symbols, ppm, HEADER_RDD);                /* :210-214 */
    if (ncw > ppm) oracle_interleave(cw + ppm, ncw - ppm, symbols + N_HEADER_SYMBOLS, ppm, rdd);
    for (unsigned i = 0; i < nsym; ++i)                                  /* :217-220 */
        symbols[i] = (uint16_t)(oracle_gray_to_binary16(symbols[i]) << (sf - ppm));
    return (long)nsym;
}

/* Returns the number of bytes written to out (LoRaDecoder.cpp:388-392), or -1 when upstream drops the
 * packet (drop()), -22 on a bad configuration.  *flags: bit0 some codeword had a non-zero syndrome,
 * bit1 an uncorrectable H(8,4) word, bit2 the header announces a CRC, bit3 the CRC matched. */
long oracle_decode_chain(const oracle_chain_cfg* c, const uint16_t* in_symbols, size_t n_in, uint8_t* out, size_t cap,
                         int* flags)
{
    const unsigned sf = c->sf, ppm = c->ppm ? c->ppm : c->sf;
    int rdd = (int)c->rdd;
    if (flags) *flags = 0;
    if (ppm > sf || ppm < 5 || c->rdd > 4 || n_in > 1200) return -22;
    if (n_in < N_HEADER_SYMBOLS) return -1;                              /* :208 */
    const unsigned nsym = round_up((unsigned)n_in, 4 + c->rdd);
    const unsigned ncw = (nsym / (4 + c->rdd)) * ppm;
    if (ncw + 32 > 2048) return -22;
    uint16_t sym[1216];
    uint8_t cw[2048];
    uint8_t bytes[1032];
    memset(sym, 0, sizeof sym);
    memset(cw, 0, sizeof cw);
    memset(bytes, 0, sizeof bytes);
    for (unsigned i = 0; i < n_in; ++i) sym[i] = in_symbols[i];
    for (unsigned i = 0; i < nsym; ++i) {                                /* :218-222 (uint16 arithmetic) */
        uint16_t s = (uint16_t)(sym[i] + (uint16_t)((1u << (sf - ppm)) / 2));
        s = (uint16_t)(s >> (sf - ppm));
        sym[i] = oracle_binary_to_gray16(s);
    }
    if (rdd != HEADER_RDD) {                                             /* :228-246 */
        oracle_deinterleave(sym, N_HEADER_SYMBOLS, cw, ppm, HEADER_RDD);
        if (c->explicit_hdr) oracle_whiten_lfsr(cw + N_HEADER_CODEWORDS, (uint16_t)(ppm - N_HEADER_CODEWORDS), 0, HEADER_RDD);
        else oracle_whiten_lfsr(cw, (uint16_t)ppm, 0, HEADER_RDD);
        if (nsym > N_HEADER_SYMBOLS) {
            oracle_deinterleave(sym + N_HEADER_SYMBOLS, nsym - N_HEADER_SYMBOLS, cw + ppm, ppm, (size_t)rdd);
            if (c->explicit_hdr) oracle_whiten_lfsr(cw + ppm, (uint16_t)(ncw - ppm), (int)(ppm - N_HEADER_CODEWORDS), (size_t)rdd);
            else oracle_whiten_lfsr(cw + ppm, (uint16_t)(ncw - ppm), (int)ppm, (size_t)rdd);
        }
    } else {                                                             /* :247-255 */
        oracle_deinterleave(sym, nsym, cw, ppm, (size_t)rdd);
        if (c->explicit_hdr) oracle_whiten_lfsr(cw + N_HEADER_CODEWORDS, (uint16_t)(ncw - N_HEADER_CODEWORDS), 0, (size_t)rdd);
        else oracle_whiten_lfsr(cw, (uint16_t)ncw, 0, (size_t)rdd);
    }
    int error = 0, bad = 0;
    const unsigned nbytes_cap = (ncw + 1) / 2;                           /* bytes.size(), :276 */
    unsigned dofs = 0, cofs = 0, packet_len = 0, data_len = 0;
    int check_crc = c->crc_check;
    if (c->explicit_hdr) {                                               /* :284-306 */
        bytes[0] = (uint8_t)(dec_fec(4, cw[1], &error, &bad) & 0xf);
        bytes[0] |= (uint8_t)(dec_fec(4, cw[0], &error, &bad) << 4);
        bytes[1] = (uint8_t)(dec_fec(4, cw[2], &error, &bad) & 0xf);
        bytes[2] = (uint8_t)(dec_fec(4, cw[4], &error, &bad) & 0xf);
        bytes[2] |= (uint8_t)(dec_fec(4, cw[3], &error, &bad) << 4);
        bytes[2] ^= oracle_header_checksum(bytes);
        if (error && c->error_check) goto drop;
        if (0 == (bytes[1] & 1)) check_crc = 0;
        rdd = (bytes[1] >> 1) & 0x7;
        if (rdd > 4) goto drop;
        packet_len = bytes[0];
        data_len = packet_len + ((bytes[1] & 1) ? 5 : 3);
        cofs = N_HEADER_CODEWORDS;
        dofs = 6;
    } else {                                                             /* :307-314 */
        packet_len = c->data_length;
        data_len = c->crc_check ? packet_len + 2 : packet_len;
    }
    if (data_len > nbytes_cap) goto drop;                                /* :316 */
    for (; cofs < ppm; ++cofs, ++dofs) {                                 /* :318-323 */
        if (dofs & 1) bytes[dofs >> 1] |= (uint8_t)(dec_fec(4, cw[cofs], &error, &bad) << 4);
        else bytes[dofs >> 1] = (uint8_t)(dec_fec(4, cw[cofs], &error, &bad) & 0xf);
    }
    if (dofs & 1) {                                                      /* :325-342 */
        bytes[dofs >> 1] |= (uint8_t)(dec_fec((unsigned)rdd, cw[cofs++], &error, &bad) << 4);
        dofs++;
    }
    dofs >>= 1;
    if (error && c->error_check) goto drop;
    for (unsigned i = dofs; i < data_len; ++i) {                         /* :348-364 */
        if (rdd == 0) {
            bytes[i] = cw[cofs++] & 0xf;
            bytes[i] |= (uint8_t)(cw[cofs++] << 4);
        } else if (rdd == 3 || rdd == 4) {
            bytes[i] = (uint8_t)(dec_fec((unsigned)rdd, cw[cofs++], &error, &bad) & 0xf);
            bytes[i] |= (uint8_t)(dec_fec((unsigned)rdd, cw[cofs++], &error, &bad) << 4);
        } else {
            bytes[i] = dec_fec((unsigned)rdd, cw[cofs++], &error, &bad);
            bytes[i] |= (uint8_t)(dec_fec((unsigned)rdd, cw[cofs++], &error, &bad) << 4);
        }
    }
    if (error && c->error_check) goto drop;
    dofs = 0;
    int crc_present = 0, crc_match = 0;
    if (c->explicit_hdr) {                                               /* :370-381 */
        if (bytes[1] & 1) {
            const uint16_t crc = oracle_data_checksum(bytes + 3, (int)packet_len);
            const uint16_t pkt_crc = (uint16_t)(bytes[3 + packet_len] | (bytes[4 + packet_len] << 8));
            crc_present = 1;
            crc_match = crc == pkt_crc;
            if (crc != pkt_crc && check_crc) goto drop;
            bytes[3 + packet_len] ^= (uint8_t)crc;
            bytes[4 + packet_len] ^= (uint8_t)(crc >> 8);
        }
        if (!c->out_hdr) { dofs = 3; data_len -= 5; }
    } else if (check_crc) {                                              /* :383-391 */
        const uint16_t crc = oracle_data_checksum(bytes, (int)c->data_length);
        const uint16_t pkt_crc = (uint16_t)(bytes[c->data_length] | (bytes[c->data_length + 1] << 8));
        crc_present = 1;
        crc_match = crc == pkt_crc;
        if (crc != pkt_crc) goto drop;
        bytes[c->data_length] ^= (uint8_t)crc;
        bytes[c->data_length + 1] ^= (uint8_t)(crc >> 8);
    }
    if (flags) *flags = (error ? 1 : 0) | (bad ? 2 : 0) | (crc_present ? 4 : 0) | (crc_match ? 8 : 0);
    if (data_len > cap) return -34;       /* also catches the unsigned wrap of data_len - 5 upstream would crash on */
    memcpy(out, bytes + dofs, data_len);
    return (long)data_len;
drop:
    if (flags) *flags = (error ? 1 : 0) | (bad ? 2 : 0);
    return -1;
}
