"""Generates tests/golden/* from the reference tree (run in the build container only; the GPU box
has no /root/reference).  Vectors are DATA copied/derived from vectors/golden, never code.

  equal_power_iq.b64          verbatim copy (4 complex64, tie-break fixture of equal_power_bin_test.cpp)
  sync_word_iq.b64            verbatim copy of the shipped (damaged) fixture
  sync_word_iq_repaired.npy   the same base64 stream with the 3 dropped characters re-inserted
                              (offsets 383 '/', 640 '/', 2559 '+'): 256 complex64 =
                              lora_modulate(no symbols, SF7, osr1, BW125, amp 1, sync 0xAB)
  detection_inputs.npz        the IQ inputs + symbol lists of detection_tests.bin, first 16 symbols per
                              case (the stored IQ is uniform noise: adversarial argmax input)
  hamming/interleaver/crc     small placeholder files, verbatim (inputs only are used)
  ref_outputs.npz             outputs of the compiled reference (oracle/_ref) on seeded inputs, so the
                              oracle stays pinned on the GPU box even if oracle/_ref were absent
"""
import base64
import os
import shutil
import struct
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import cpu_libs  # noqa: E402

REF = "/root/reference/vectors/golden"


def main():
    for f in ("equal_power_iq.b64", "sync_word_iq.b64", "hamming_tests.bin", "interleaver_tests.bin",
              "crc_tests.bin", "sync_word_tests.bin"):
        shutil.copyfile(os.path.join(REF, f), os.path.join(HERE, f))
        os.chmod(os.path.join(HERE, f), 0o644)
    ref = cpu_libs.load_ref()
    want = ref.lora_modulate(np.zeros(0, np.uint16), 7, 1, 125000, 1.0, 0xAB)
    s = "".join(open(os.path.join(REF, "sync_word_iq.b64")).read().split())
    for off, ch in ((383, "/"), (640, "/"), (2559, "+")):
        s = s[:off] + ch + s[off:]
    raw = base64.b64decode(s + "=" * (-len(s) % 4))
    iq = np.frombuffer(raw[:256 * 8], dtype=np.complex64)
    assert np.array_equal(iq.view(np.uint32), want.view(np.uint32)), "repaired fixture != reference modulate"
    np.save(os.path.join(HERE, "sync_word_iq_repaired.npy"), iq)

    # detection_tests.bin: u32 count; {u8 0,u32 sf,u32 bw,u32 cr,u32 nsym,u32 sym[nsym],u32 nsamp,f64[2*nsamp]}
    b = open(os.path.join(REF, "detection_tests.bin"), "rb").read()
    pos = 0
    (cnt,) = struct.unpack_from("<I", b, pos); pos += 4
    cases = {}
    for c in range(cnt):
        pos += 1
        sf, bw, cr, nsym = struct.unpack_from("<4I", b, pos); pos += 16
        syms = np.frombuffer(b, dtype="<u4", count=nsym, offset=pos); pos += 4 * nsym
        (nsamp,) = struct.unpack_from("<I", b, pos); pos += 4
        iqd = np.frombuffer(b, dtype="<f8", count=2 * nsamp, offset=pos); pos += 16 * nsamp
        keep = min(nsym, 16) * (1 << sf)
        cases[f"case{c}_meta"] = np.array([sf, bw, cr, nsym], np.uint32)
        cases[f"case{c}_symbols"] = syms.astype(np.uint32)
        cases[f"case{c}_iq"] = (iqd[0:2 * keep:2] + 1j * iqd[1:2 * keep:2]).astype(np.complex64)
    assert pos == len(b)
    np.savez_compressed(os.path.join(HERE, "detection_inputs.npz"), **cases)

    # reference outputs on seeded inputs (pins the oracle without oracle/_ref)
    rng = np.random.default_rng(0xC0FFEE)
    out = {}
    for sf in (2, 5, 7, 8, 9, 12):
        N = 1 << sf
        x = (rng.standard_normal(N) + 1j * rng.standard_normal(N)).astype(np.complex64)
        idx, spec, p, pav, fi = ref.detect(x, sf)
        out[f"detect_sf{sf}_in"] = x
        out[f"detect_sf{sf}_spec"] = spec
        out[f"detect_sf{sf}_scalars"] = np.array([idx, p, pav, fi], np.float64)
        syms = rng.integers(0, N, 6).astype(np.uint16)
        for osr in (1, 2):
            iq = ref.lora_modulate(syms, sf, osr, 125000, 1.0, 0x12)
            down = ref.gen_chirp(N, osr, N * osr, 0.0, True, 1.0, 0.0, 1.0)[0]
            d = (iq.reshape(-1, N * osr) * down[None, :]).astype(np.complex64).ravel()
            d = (d + 0.4 * (rng.standard_normal(d.size) + 1j * rng.standard_normal(d.size))).astype(np.complex64)
            r, s_, sync, cfo, to = ref.lora_demodulate(d, sf, osr, 0, 1)
            out[f"demod_sf{sf}_osr{osr}_in"] = d
            out[f"demod_sf{sf}_osr{osr}_mod"] = iq
            out[f"demod_sf{sf}_osr{osr}_syms_in"] = syms
            out[f"demod_sf{sf}_osr{osr}_out"] = s_
            out[f"demod_sf{sf}_osr{osr}_scalars"] = np.array([r, sync], np.int64)
            out[f"demod_sf{sf}_osr{osr}_metrics"] = np.array([cfo, to], np.float32)
    np.savez_compressed(os.path.join(HERE, "ref_outputs.npz"), **out)
    print("golden fixtures written")


if __name__ == "__main__":
    main()
