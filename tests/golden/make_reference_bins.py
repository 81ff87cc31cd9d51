"""Golden vectors in the REFERENCE'S OWN .bin layouts, regenerated from the compiled
reference (oracle/_ref/libref_lora_phy.so) so that they finally agree with the reference's code:

  modulation_tests.bin   the file tests/bit_exact_test.cpp:58-104 loads and the reference never shipped
                         (.MISSING_LARGE_BLOBS): u32 count; per record u8 type, u32 sf, u32 bw_khz, u32 cr_idx,
                         u32 flags, u32 len, payload[len], u32 nsamp, f64 iq[2*nsamp] -- one record per profile of
                         tests/profiles.yaml.  IQ = lora_modulate(lora_encode(source payload)); the payload field holds
                         what the reference's own chain (dechirp -> lora_demodulate -> lora_decode) returns for that IQ,
                         which is the source payload at BW125 and the (bw_scale*s mod N) image of it at BW250 / 500
                          -- so bit_exact_test passes on the reference build by construction and passes
                         on the B200 drop-in build iff that build is bit-exact.
  awgn_tests.bin         same layout, the same frames plus seeded complex Gaussian noise (also absent upstream).
  detection_tests.bin    u32 count; u8 type, u32 sf, u32 bw_khz, u32 cr_idx, u32 nsym, u32 sym[nsym], u32 nsamp,
                         f64 iq[2*nsamp]: real chirps now (upstream: uniform noise), sym = the reference's demodulator output.
  hamming_tests.bin      u32 32; 16 x {u8 0, data, enc, dec, err, bad} H(8,4); 16 x {u8 1, data, enc, dec, err} H(7,4),
                         decoded from a codeword with one flipped bit (upstream: enc == data).
  interleaver_tests.bin  u32 count; {u8 0, u32 sf, u32 n_cw, cw[n_cw], u32 n_sym, u16 sym[n_sym], out[n_cw], u8 ok}
                         n_cw = PPM = sf codewords of 4+RDD bits, n_sym = 4+RDD (upstream: identity placeholder).
  crc_tests.bin          u32 5; {u8 0, u32 len, payload, u16 crc, u8 1} with crc = sx1272DataChecksum (upstream: byte sum).
  manifest.json          sha256 and meaning of every file, the source payloads, the generator's seed.

The big files are NOT committed (3 MB of float64 IQ): this script rebuilds them bit for bit wherever oracle/_ref exists,
and tests/golden/reference_bins_manifest.json pins their hashes.  Usage:
    python tests/golden/make_reference_bins.py OUT_DIR        (writes OUT_DIR/*.bin + manifest.json)
"""
import hashlib
import json
import os
import struct
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import cpu_libs  # noqa: E402

SEED = 0xC0FFEE
PROFILES = [("sf7_bw125_cr45", 7, 125000, 1), ("sf7_bw125_cr47", 7, 125000, 3), ("sf8_bw125_cr45", 8, 125000, 1),
            ("sf9_bw250_cr48", 9, 250000, 4), ("sf10_bw250_cr47", 10, 250000, 3), ("sf11_bw500_cr45", 11, 500000, 1),
            ("sf12_bw500_cr45", 12, 500000, 1)]
PAYLOAD_LEN = 8


def ref_chain(ref, iq, sf, bw):
    """tests/bit_exact_test.cpp:131-166: caller-side dechirp, lora_demodulate, lora_decode."""
    N = 1 << sf
    down = ref.downchirp(sf, bw)
    d = (iq.reshape(-1, N) * down[None, :]).astype(np.complex64).ravel()
    r, sym, sync, cfo, to = ref.lora_demodulate(d, sf)
    n, pay = ref.lora_decode(sym[:r])
    return sym[:r].copy(), pay


def frame_record(sf, bw, cr, payload, iq):
    b = struct.pack("<BIIIII", 0, sf, bw // 1000, cr, 0, len(payload)) + bytes(payload)
    x = np.empty(2 * iq.size, np.float64)
    x[0::2], x[1::2] = iq.real, iq.imag
    return b + struct.pack("<I", iq.size) + x.tobytes()


def build(out_dir):
    ref = cpu_libs.load_ref()
    if ref is None:
        raise SystemExit("oracle/_ref/libref_lora_phy.so is missing: run `make -C oracle ref` where /root/reference exists")
    os.makedirs(out_dir, exist_ok=True)
    rng = np.random.default_rng(SEED)
    notes = {}
    files = {}

    mod, awgn, sources = [], [], {}
    for name, sf, bw, cr in PROFILES:
        src = rng.integers(0, 256, PAYLOAD_LEN, dtype=np.uint8)
        iq = ref.lora_modulate(ref.lora_encode(src, sf), sf, 1, bw, 1.0, 0x12)
        _, expect = ref_chain(ref, iq, sf, bw)
        mod.append(frame_record(sf, bw, cr, expect, iq))
        sigma = 0.35
        noisy = (iq + sigma * (rng.standard_normal(iq.size) + 1j * rng.standard_normal(iq.size))).astype(np.complex64)
        _, expect_n = ref_chain(ref, noisy, sf, bw)
        awgn.append(frame_record(sf, bw, cr, expect_n, noisy))
        sources[name] = {"source_payload": bytes(src).hex(), "reference_decodes_clean": bytes(expect).hex(),
                         "reference_decodes_noisy": bytes(expect_n).hex(), "noise_sigma_per_component": sigma}
    files["modulation_tests.bin"] = struct.pack("<I", len(mod)) + b"".join(mod)
    files["awgn_tests.bin"] = struct.pack("<I", len(awgn)) + b"".join(awgn)
    notes["frames"] = sources

    det = []
    for sf, cr in ((7, 1), (7, 4), (9, 1), (9, 4), (12, 1), (12, 4)):
        N = 1 << sf
        nsym = {7: 8, 9: 10, 12: 13}[sf]
        tx = rng.integers(0, min(N, 256), nsym).astype(np.uint16)
        iq = ref.lora_modulate(tx, sf, 1, 125000, 1.0, 0x12)
        iq = (iq + 0.5 * (rng.standard_normal(iq.size) + 1j * rng.standard_normal(iq.size))).astype(np.complex64)
        sym, _ = ref_chain(ref, iq, sf, 125000)
        x = np.empty(2 * iq.size, np.float64)
        x[0::2], x[1::2] = iq.real, iq.imag
        det.append(struct.pack("<BIIII", 0, sf, 125, cr, sym.size) + sym.astype("<u4").tobytes() + struct.pack("<I", iq.size) + x.tobytes())
    files["detection_tests.bin"] = struct.pack("<I", len(det)) + b"".join(det)

    ham = []
    for v in range(16):
        enc = ref.u8fn("enc_h84", v)
        dec, err, bad = ref.u8fn_err("dec_h84", enc ^ (1 << (v % 8)), two=True)
        ham.append(struct.pack("<BBBBBB", 0, v, enc, dec, err, bad))
    for v in range(16):
        enc = ref.u8fn("enc_h74", v)
        dec, err = ref.u8fn_err("dec_h74", enc ^ (1 << (v % 7)))
        ham.append(struct.pack("<BBBBB", 1, v, enc, dec, err))
    files["hamming_tests.bin"] = struct.pack("<I", 32) + b"".join(ham)

    inter = []
    for sf in (7, 9, 12):
        for rdd in range(5):
            cw = rng.integers(0, 1 << (4 + rdd), sf, dtype=np.uint8)
            sym = ref.interleave(cw, sf, rdd)
            back = ref.deinterleave(sym, sf, rdd)
            inter.append(struct.pack("<BII", 0, sf, cw.size) + cw.tobytes() + struct.pack("<I", sym.size) + sym.astype("<u2").tobytes() +
                         back.tobytes() + struct.pack("<B", int(np.array_equal(back, cw))))
    files["interleaver_tests.bin"] = struct.pack("<I", len(inter)) + b"".join(inter)

    crc = []
    for data in (b"Hello", b"World", bytes(range(1, 9)), bytes(8), b"\xff" * 8):
        c = ref.data_checksum(np.frombuffer(data, np.uint8))
        crc.append(struct.pack("<BI", 0, len(data)) + data + struct.pack("<HB", c, 1))
    files["crc_tests.bin"] = struct.pack("<I", len(crc)) + b"".join(crc)

    manifest = {"generator": "tests/golden/make_reference_bins.py", "seed": hex(SEED), "from": "oracle/_ref/libref_lora_phy.so",
                "files": {}, "notes": notes}
    for name, blob in files.items():
        with open(os.path.join(out_dir, name), "wb") as f:
            f.write(blob)
        manifest["files"][name] = {"bytes": len(blob), "sha256": hashlib.sha256(blob).hexdigest()}
    with open(os.path.join(out_dir, "manifest.json"), "w") as f:
        json.dump(manifest, f, indent=1, sort_keys=True)
    return manifest


if __name__ == "__main__":
    m = build(sys.argv[1] if len(sys.argv) > 1 else os.path.join(HERE, "_reference_bins"))
    print(json.dumps(m["files"], indent=1))
