"""The C++ drop-in boundary (include/lora_phy_b200/phy.hpp over the C ABI).

CPU: tests/cpp/phy_dropin_test.cpp compiles and its argument / errno contract (the precedence pinned by
the reference's tests/error_code_test.cpp) holds without a device.
GPU: the same program end to end, and the reference's OWN test sources (roundtrip, no_alloc,
equal_power_bin, odd_symbol_count, scratch_buffer_error) compiled against the drop-in header by
oracle/Makefile `reftests` in the build container -- they must pass unchanged on the B200 path."""
import os
import subprocess

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
PKG = os.path.join(ROOT, "lora-sdr-lightweight-standalone-library-clean_b200")
REFTESTS = os.path.join(ROOT, "oracle", "_ref", "reftests")


def build_dropin(name="phy_dropin_test"):
    out_dir = os.path.join(HERE, "cpp", "build")
    os.makedirs(out_dir, exist_ok=True)
    exe = os.path.join(out_dir, name)
    if not os.path.exists(os.path.join(PKG, "liblora_phy_cuda.so")):
        import lora_phy_b200
        lora_phy_b200.build()
    subprocess.run(["g++", "-std=c++17", "-O1", "-Wall", "-I", os.path.join(ROOT, "include"),
                    os.path.join(HERE, "cpp", name + ".cpp"), "-o", exe, "-L", PKG, "-llora_phy_cuda", "-pthread",
                    "-Wl,-rpath," + PKG], check=True)
    return exe


def test_dropin_header_argument_contract():
    import torch
    if torch.cuda.is_available():
        pytest.skip("covered by the gpu run")
    r = subprocess.run([build_dropin(), "args"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr


@pytest.mark.gpu
def test_dropin_header_end_to_end():
    r = subprocess.run([build_dropin(), "gpu"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr


def test_threads_program_compiles():
    build_dropin("phy_threads_test")


@pytest.mark.gpu
def test_two_threads_one_workspace_each():
    """the drop-in wrapper funnels every workspace through one context; its entry points must serialise
    (context mutex held from staging to copy-out) so reference-style one-workspace-per-thread use stays correct."""
    r = subprocess.run([build_dropin("phy_threads_test")], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["roundtrip_test", "no_alloc_test", "equal_power_bin_test", "odd_symbol_count_test",
                                  "scratch_buffer_error_test"])
def test_reference_own_tests_pass_on_b200(name, tmp_path):
    exe = os.path.join(REFTESTS, name)
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/reftests not built (needs /root/reference at build time)")
    (tmp_path / "vectors").mkdir()
    os.symlink(os.path.join(HERE, "golden"), tmp_path / "vectors" / "golden")
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300, cwd=tmp_path)
    assert r.returncode == 0, r.stdout + r.stderr


REFBIN = os.path.join(ROOT, "oracle", "_ref", "refbin")


def _run(exe, args, cwd):
    r = subprocess.run([exe] + args, capture_output=True, text=True, timeout=300, cwd=cwd)
    return r.returncode, r.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("sf,bw,payload", [(7, 125000, "48656c6c6f"), (9, 250000, "00ff10ef5aa5c33c"), (12, 125000, "deadbeef"),
                                           (8, 500000, "0102030405060708090a0b0c0d0e0f10")])
def test_reference_cli_runners_on_b200(sf, bw, payload, tmp_path):
    """The reference's own tx_runner / rx_runner (raw interleaved float32 IQ on disk, payload / CRC / CFO / time-offset
    report), compiled unchanged against the drop-in header, against the same sources linked with the reference's own
    library: identical IQ files and identical reports."""
    for d in (REFBIN, REFTESTS):
        if not os.path.exists(os.path.join(d, "tx_runner")):
            pytest.skip("oracle/_ref runners not built (needs /root/reference at build time)")
    args = [f"--payload={payload}", f"--sf={sf}", f"--bw={bw}"]
    assert _run(os.path.join(REFBIN, "tx_runner"), args + ["--out=cpu.bin"], tmp_path)[0] == 0
    assert _run(os.path.join(REFTESTS, "tx_runner"), args + ["--out=gpu.bin"], tmp_path)[0] == 0
    cpu, gpu = (tmp_path / "cpu.bin").read_bytes(), (tmp_path / "gpu.bin").read_bytes()
    assert len(cpu) == (len(payload) + 2) * (1 << sf) * 8 and cpu == gpu
    rargs = ["--in=cpu.bin", f"--sf={sf}", f"--bw={bw}", "--report-offsets"]
    want = _run(os.path.join(REFBIN, "rx_runner"), rargs, tmp_path)
    got = _run(os.path.join(REFTESTS, "rx_runner"), rargs, tmp_path)
    assert want[0] == 0 and got == want, (want, got)
    assert _run(os.path.join(REFTESTS, "rx_runner"), ["--in=missing.bin"], tmp_path)[0] == 1


@pytest.mark.gpu
@pytest.mark.parametrize("extra", [[], ["--osr=2"], ["--sf=9", "--bytes=12", "--window=hann"], ["--sf=8", "--bw=250000"]])
def test_reference_vector_dump_on_b200(extra, tmp_path):
    """lora_phy_vector_dump (per-stage CSV / bin dumps, runners/lora_phy_vector_dump.cpp:164-203): the directory the
    drop-in build writes equals the reference build's, file by file."""
    for d in (REFBIN, REFTESTS):
        if not os.path.exists(os.path.join(d, "lora_phy_vector_dump")):
            pytest.skip("oracle/_ref runners not built (needs /root/reference at build time)")
    (tmp_path / "cpu").mkdir()
    (tmp_path / "gpu").mkdir()
    a = _run(os.path.join(REFBIN, "lora_phy_vector_dump"), ["--out=cpu", "--seed=11"] + extra, tmp_path)
    b = _run(os.path.join(REFTESTS, "lora_phy_vector_dump"), ["--out=gpu", "--seed=11"] + extra, tmp_path)
    assert a[0] == b[0] == 0, (a, b)
    names = sorted(os.listdir(tmp_path / "cpu"))
    assert names and names == sorted(os.listdir(tmp_path / "gpu"))
    for n in names:
        assert (tmp_path / "cpu" / n).read_bytes() == (tmp_path / "gpu" / n).read_bytes(), n
