"""Regenerated golden set for the vector classes the reference names but does not ship usable
data for (modulation / awgn absent; hamming / interleaver / crc placeholders), plus the upstream-chain, frame
and stream rows.  Produced from the ORACLE (oracle/liblora_oracle.so), which is itself pinned against the
compiled reference and the reference's known answers (tests/test_oracle.py); this file freezes its outputs so
that any later change to the oracle or the CUDA path shows up as a diff.  Run:  python tests/golden/make_regenerated.py
Layout of regenerated.npz (little-endian numpy arrays; `manifest` lists every case):
  chain_{i}_{payload,symbols}            upstream coding chain, parameters in chain_{i}_cfg = [sf, cr, explicit, crc, ppm]
  hamming_{84,74}_table, parity_{64,54}_table   encoder tables for nibbles 0..15
  crc_{i}_{data,value}                   sx1272DataChecksum
  interleave_{sf}_{rdd}_{cw,sym}         diagonal interleaver, one block
  modulation_{i}_{symbols,iq}            lora_modulate, parameters in modulation_{i}_cfg = [sf, osr, bw_hz, sync]
  frame_{sf}_{symbols,iq}                upstream frame format (amplitude 0.3, padding 1)
  stream_{sf}_{iq,start,err,nsym,symbols,end}  streaming receiver on a noisy 2-frame stream (thresh -8 dB, mtu 24)
  awgn_{i}_{cfg,rx_sym,rx_payload,length}      Monte-Carlo packets, cfg = [sf, cr, cell, packet, payload_len], std in awgn_{i}_std
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import cpu_libs  # noqa: E402


def build(o):
    rng = np.random.default_rng(0x601DE2)
    out, manifest = {}, []

    def put(name, arr):
        out[name] = np.ascontiguousarray(arr)
        manifest.append(name)

    i = 0
    for sf in (7, 9, 12):
        for cr, expl, crc, ppm in ((1, 1, 1, 0), (4, 1, 1, 0), (2, 0, 1, 0), (3, 1, 0, sf - 2)):
            L = int(rng.integers(1, 40))
            pay = rng.integers(0, 256, L, dtype=np.uint8)
            cfg = o.chain_cfg(sf, cr=cr, explicit_header=expl, crc=crc, ppm=ppm, data_length=L)
            put(f"chain_{i}_cfg", np.array([sf, cr, expl, crc, ppm], np.uint32))
            put(f"chain_{i}_payload", pay)
            put(f"chain_{i}_symbols", o.encode_chain(cfg, pay))
            i += 1
    put("n_chain", np.array([i], np.uint32))
    import ctypes as C
    for name, key in (("enc_h84", "hamming_84_table"), ("enc_h74", "hamming_74_table"), ("enc_p64", "parity_64_table"),
                      ("enc_p54", "parity_54_table")):
        f = o.fn(name, C.c_uint8, [C.c_uint8])
        put(key, np.array([f(v) for v in range(16)], np.uint8))
    for k, data in enumerate((b"Hello", b"World", b"Test", b"LoRa", bytes(range(1, 9)), bytes(8), b"\xff" * 8)):
        put(f"crc_{k}_data", np.frombuffer(data, np.uint8))
        put(f"crc_{k}_value", np.array([o.data_checksum(np.frombuffer(data, np.uint8))], np.uint16))
    put("n_crc", np.array([7], np.uint32))
    i = 0
    for sf, osr, bw, sync in ((7, 1, 125000, 0x12), (8, 2, 250000, 0x34), (10, 1, 500000, 0xab), (12, 1, 125000, 0x12)):
        syms = rng.integers(0, 1 << sf, 5).astype(np.uint16)
        put(f"modulation_{i}_cfg", np.array([sf, osr, bw, sync], np.uint32))
        put(f"modulation_{i}_symbols", syms)
        put(f"modulation_{i}_iq", o.lora_modulate(syms, sf, osr, bw, 0.9, sync))
        i += 1
    put("n_modulation", np.array([i], np.uint32))
    for sf in (7, 10):
        syms = rng.integers(0, 1 << sf, 6).astype(np.uint16)
        put(f"frame_{sf}_symbols", syms)
        put(f"frame_{sf}_iq", o.frame_modulate(syms, sf, ampl=0.3, padding=1))
        N = 1 << sf
        a = o.frame_modulate(rng.integers(0, N, 20).astype(np.uint16), sf, ampl=1.0, padding=2)
        b = o.frame_modulate(rng.integers(0, N, 9).astype(np.uint16), sf, ampl=1.0, padding=2)
        x = np.concatenate([np.zeros(N + 37, np.complex64), a, np.zeros(211, np.complex64), b, np.zeros(3 * N, np.complex64)])
        x = x * np.exp(2j * np.pi * 0.2 * np.arange(x.size) / N)
        x = (x + 0.1 * (rng.standard_normal(x.size) + 1j * rng.standard_normal(x.size))).astype(np.complex64)
        n, start, err, nsym, sym, end = o.stream_demod(x, sf, thresh=-8.0, mtu=24, max_frames=4)
        assert n == 2
        put(f"stream_{sf}_iq", x)
        put(f"stream_{sf}_start", start[:n]); put(f"stream_{sf}_err", err[:n]); put(f"stream_{sf}_nsym", nsym[:n])
        put(f"stream_{sf}_symbols", sym[:n]); put(f"stream_{sf}_end", np.array([end], np.uint32))
    i = 0
    for sf, cr, snr in ((7, 4, -9.0), (7, 1, -11.0), (9, 4, -15.0), (11, 2, -19.0)):
        std = np.float32(np.sqrt(10.0 ** (-snr / 10.0) / 2.0))
        for pkt in (0, 1, (1 << 32) + 7):
            r = o.awgn_roundtrip(o.chain_cfg(sf, cr=cr), float(std), 0xA76E, sf << 8 | 1, pkt, 9)
            put(f"awgn_{i}_cfg", np.array([sf, cr, sf << 8 | 1, pkt, 9], np.uint64))
            put(f"awgn_{i}_std", np.array([std], np.float32))
            put(f"awgn_{i}_rx_sym", r["rx_sym"]); put(f"awgn_{i}_rx_payload", r["rx_payload"])
            put(f"awgn_{i}_tx_payload", r["tx_payload"])
            put(f"awgn_{i}_length", np.array([r["length"]], np.int32))
            i += 1
    put("n_awgn", np.array([i], np.uint32))
    out["manifest"] = np.array(manifest)
    return out


if __name__ == "__main__":
    o = cpu_libs.load_oracle()
    z = build(o)
    np.savez_compressed(os.path.join(HERE, "regenerated.npz"), **z)
    print("regenerated.npz:", len(z["manifest"]), "arrays,", os.path.getsize(os.path.join(HERE, "regenerated.npz")), "bytes")
