"""Golden vectors in the reference's own .bin layouts.

tests/golden/make_reference_bins.py regenerates, from the compiled reference, the files the reference names but does
not ship usable data for -- modulation_tests.bin / awgn_tests.bin (absent upstream), detection / hamming / interleaver /
crc (placeholders upstream) -- and tests/golden/reference_bins_manifest.json pins their SHA-256.  With them the
reference's own tests/bit_exact_test.cpp, which has never run anywhere, runs twice: compiled against the reference's
sources (oracle/_ref/refbin, CPU) and, unchanged, against the B200 drop-in header (oracle/_ref/reftests, GPU).
"""
import json
import os
import shutil
import struct
import subprocess
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REFBIN = os.path.join(ROOT, "oracle", "_ref", "refbin")
REFTESTS = os.path.join(ROOT, "oracle", "_ref", "reftests")
sys.path.insert(0, os.path.join(HERE, "golden"))


@pytest.fixture(scope="module")
def bins(tmp_path_factory, ref):
    import make_reference_bins
    d = tmp_path_factory.mktemp("refbins")
    manifest = make_reference_bins.build(str(d))
    return str(d), manifest


def sandbox(tmp_path, bins_dir, which="modulation_tests.bin"):
    """The working directory bit_exact_test expects: tests/profiles.yaml and vectors/golden/modulation_tests.bin."""
    (tmp_path / "tests").mkdir(exist_ok=True)
    shutil.copy(os.path.join(HERE, "golden", "profiles.yaml"), tmp_path / "tests" / "profiles.yaml")
    g = tmp_path / "vectors" / "golden"
    g.mkdir(parents=True, exist_ok=True)
    shutil.copy(os.path.join(bins_dir, which), g / "modulation_tests.bin")
    return str(tmp_path)


def test_generator_is_pinned(bins):
    d, manifest = bins
    want = json.load(open(os.path.join(HERE, "golden", "reference_bins_manifest.json")))
    assert {k: v["sha256"] for k, v in manifest["files"].items()} == {k: v["sha256"] for k, v in want["files"].items()}
    for name in ("crc_tests.bin", "hamming_tests.bin", "interleaver_tests.bin"):          # the small ones are committed
        assert open(os.path.join(d, name), "rb").read() == open(os.path.join(HERE, "golden", "reference_layout", name), "rb").read()


def test_corrected_small_files_agree_with_the_oracle_port(bins, oracle):
    d, _ = bins
    b = open(os.path.join(d, "crc_tests.bin"), "rb").read()
    n, = struct.unpack_from("<I", b, 0)
    o = 4
    for _ in range(n):
        t, ln = struct.unpack_from("<BI", b, o)
        data = b[o + 5:o + 5 + ln]
        crc, ok = struct.unpack_from("<HB", b, o + 5 + ln)
        o += 5 + ln + 3
        assert ok == 1 and crc == oracle.data_checksum(np.frombuffer(data, np.uint8))
    assert o == len(b)
    b = open(os.path.join(d, "hamming_tests.bin"), "rb").read()
    assert struct.unpack_from("<I", b, 0)[0] == 32
    o = 4
    for v in range(16):
        t, data, enc, dec, err, bad = struct.unpack_from("<BBBBBB", b, o)
        o += 6
        assert (t, data, enc) == (0, v, oracle.u8fn("enc_h84", v)) and (dec, err, bad) == (v, 1, 0)
        assert enc != data or v == 0                      # upstream's placeholder had enc == data
    for v in range(16):
        t, data, enc, dec, err = struct.unpack_from("<BBBBB", b, o)
        o += 5
        assert (t, data, enc, dec, err) == (1, v, oracle.u8fn("enc_h74", v), v, 1)
    b = open(os.path.join(d, "interleaver_tests.bin"), "rb").read()
    n, = struct.unpack_from("<I", b, 0)
    o = 4
    for _ in range(n):
        t, sf, ncw = struct.unpack_from("<BII", b, o)
        cw = np.frombuffer(b, np.uint8, ncw, o + 9)
        nsym, = struct.unpack_from("<I", b, o + 9 + ncw)
        sym = np.frombuffer(b, "<u2", nsym, o + 13 + ncw)
        back = np.frombuffer(b, np.uint8, ncw, o + 13 + ncw + 2 * nsym)
        ok = b[o + 13 + 2 * ncw + 2 * nsym]
        o += 14 + 2 * ncw + 2 * nsym
        assert ncw == sf and ok == 1 and np.array_equal(back, cw)
        assert np.array_equal(oracle.interleave(cw, sf, nsym - 4), sym)
    assert o == len(b)


@pytest.mark.parametrize("which", ["modulation_tests.bin", "awgn_tests.bin"])
def test_reference_bit_exact_test_on_the_reference_build(bins, tmp_path, which):
    exe = os.path.join(REFBIN, "bit_exact_test")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/refbin/bit_exact_test not built (needs /root/reference at build time)")
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300, cwd=sandbox(tmp_path, bins[0], which))
    assert r.returncode == 0 and r.stdout.count("passed.") == 7, r.stdout + r.stderr


@pytest.mark.gpu
@pytest.mark.parametrize("which", ["modulation_tests.bin", "awgn_tests.bin"])
def test_reference_bit_exact_test_on_the_b200_dropin_build(bins, tmp_path, which):
    """The one reference test that had never run anywhere, compiled unchanged against include/lora_phy_b200/phy.hpp."""
    exe, cpu = os.path.join(REFTESTS, "bit_exact_test"), os.path.join(REFBIN, "bit_exact_test")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/reftests/bit_exact_test not built (needs /root/reference at build time)")
    cwd = sandbox(tmp_path, bins[0], which)
    r = subprocess.run([exe], capture_output=True, text=True, timeout=600, cwd=cwd)
    assert r.returncode == 0 and r.stdout.count("passed.") == 7, r.stdout + r.stderr
    if os.path.exists(cpu):
        assert subprocess.run([cpu], capture_output=True, text=True, timeout=300, cwd=cwd).stdout == r.stdout


@pytest.mark.gpu
def test_detection_bin_symbols_on_gpu(bins, oracle):
    """detection_tests.bin in the upstream layout, with real chirps: the stored symbols (the reference's output) come
    back from the CUDA demodulator on every path."""
    import lora_phy_b200 as lp
    import torch
    phy = lp.LoraPhyCuda(0)
    b = open(os.path.join(bins[0], "detection_tests.bin"), "rb").read()
    n, = struct.unpack_from("<I", b, 0)
    o = 4
    for _ in range(n):
        t, sf, bw, cr, nsym = struct.unpack_from("<BIIII", b, o)
        sym = np.frombuffer(b, "<u4", nsym, o + 17)
        nsamp, = struct.unpack_from("<I", b, o + 17 + 4 * nsym)
        x = np.frombuffer(b, np.float64, 2 * nsamp, o + 21 + 4 * nsym)
        o += 21 + 4 * nsym + 16 * nsamp
        iq = (x[0::2] + 1j * x[1::2]).astype(np.complex64)
        N = 1 << sf
        d = (iq.reshape(-1, N) * oracle.downchirp(sf)[None, :]).astype(np.complex64).ravel()
        for flags in (0, lp.F_EXACT, lp.F_TWO_KERNEL):
            out = phy.demodulate_batch(lp.make_cfg(sf, flags=flags), torch.from_numpy(d[None, :].copy()).to(phy.device))
            phy.synchronize()
            assert int(out["status"][0]) == nsym and np.array_equal(out["symbols"][0, :nsym].cpu().numpy().astype(np.uint32), sym)
    assert o == len(b)
    phy.close()
