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


def build_dropin():
    out_dir = os.path.join(HERE, "cpp", "build")
    os.makedirs(out_dir, exist_ok=True)
    exe = os.path.join(out_dir, "phy_dropin_test")
    if not os.path.exists(os.path.join(PKG, "liblora_phy_cuda.so")):
        import lora_phy_b200
        lora_phy_b200.build()
    subprocess.run(["g++", "-std=c++17", "-O1", "-Wall", "-I", os.path.join(ROOT, "include"),
                    os.path.join(HERE, "cpp", "phy_dropin_test.cpp"), "-o", exe, "-L", PKG, "-llora_phy_cuda",
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
